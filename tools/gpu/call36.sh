set -x
timeout 150 python -m epfl_megatron_b200.ops.selfcheck attn_bwd2_small attn_bwd2_mha attn_bwd2_gqa attn_bwd2_mqa_long attn_bwd2_full attn_bwd2_ragged > gpurun_out/selfcheck_bwd2.log 2>&1; echo "selfcheck rc=$?"; grep -v SUMMARY gpurun_out/selfcheck_bwd2.log | cut -c1-300
timeout 90 python tools/kernel_bench.py attn_bwd > gpurun_out/kb_attn_bwd_v2.log 2>&1; cut -c1-420 gpurun_out/kb_attn_bwd_v2.log
