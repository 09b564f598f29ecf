set -x
timeout 120 python -m epfl_megatron_b200.ops.selfcheck attn_fwd attn_fwd_mqa_long attn_fwd_full attn_fwd_ragged attn_fwd_one_tile attn_fwd_long attn_bwd2_gqa > gpurun_out/selfcheck_fwd2.log 2>&1; echo "selfcheck rc=$?"; grep -v SUMMARY gpurun_out/selfcheck_fwd2.log | cut -c1-300
B200_BENCH_ATTN_BWD_V2=0 timeout 60 python tools/kernel_bench.py attn > gpurun_out/kb_attn_fwd_lean.log 2>&1; cut -c1-330 gpurun_out/kb_attn_fwd_lean.log
