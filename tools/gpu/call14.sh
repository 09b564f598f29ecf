set -x
timeout 600 python -m epfl_megatron_b200.ops.selfcheck attn_fwd_one_tile attn_fwd attn_fwd_mqa_long attn_fwd_full attn_fwd_ragged attn_fwd_long attn_fwd_gen1 > gpurun_out/selfcheck_attn2.log 2>&1; echo "selfcheck rc=$?"; grep -v SUMMARY gpurun_out/selfcheck_attn2.log | cut -c1-600
timeout 300 python tools/kernel_bench.py attn > gpurun_out/kb_attn2.log 2>&1; cat gpurun_out/kb_attn2.log | cut -c1-600
