set -x
timeout 60 python -m epfl_megatron_b200.ops.selfcheck attn_bwd2_small > gpurun_out/selfcheck_bwd2_small.log 2>&1; rc=$?; echo "small rc=$rc"; cut -c1-500 gpurun_out/selfcheck_bwd2_small.log | tail -3
if grep -q "attn_bwd2_small OK" gpurun_out/selfcheck_bwd2_small.log; then
  timeout 150 python -m epfl_megatron_b200.ops.selfcheck attn_bwd2_mha attn_bwd2_gqa attn_bwd2_mqa_long attn_bwd2_full attn_bwd2_ragged attn_bwd_small > gpurun_out/selfcheck_bwd2.log 2>&1; echo "rest rc=$?"; grep -v SUMMARY gpurun_out/selfcheck_bwd2.log | cut -c1-300
  timeout 90 python tools/kernel_bench.py attn_bwd > gpurun_out/kb_attn_bwd_v2.log 2>&1; cut -c1-420 gpurun_out/kb_attn_bwd_v2.log
fi
