set -x
timeout 600 python -m epfl_megatron_b200.ops.selfcheck attn_bwd_small attn_bwd_mha attn_bwd_gqa attn_bwd_mqa_long attn_bwd_full attn_bwd_ragged > gpurun_out/selfcheck_attn_bwd.log 2>&1; echo "selfcheck rc=$?"; grep -v SUMMARY gpurun_out/selfcheck_attn_bwd.log | cut -c1-400
timeout 300 python tools/kernel_bench.py attn_bwd > gpurun_out/kb_attn_bwd3.log 2>&1; cat gpurun_out/kb_attn_bwd3.log | cut -c1-330
