# Round 2 (1 GPU): FP8 cast + GEMM (tcgen05 kind::f8f6f4): numerics and throughput.
set -x
for n in fp8_quantize_e4m3 fp8_quantize_e5m2_ragged gemm_fp8 gemm_fp8_hybrid_big gemm_fp8_ragged_f32 gemm_fp8_wgrad_acc gemm_nt_pair gemm_tn_pair_acc; do timeout 120 python -m epfl_megatron_b200.ops.selfcheck --one $n 2>/dev/null | tail -1 | cut -c1-300; done
timeout 200 python tools/kernel_bench.py fp8 2>&1 | grep kernel | cut -c1-330
