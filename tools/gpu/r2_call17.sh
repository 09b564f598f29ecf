# Round 2 (1 GPU): LayerNorm on the bulk-copy ring: numerics, A/B vs the block kernels, model-level tests.
set -x
for n in layernorm layernorm_h4096 layernorm_h1024 layernorm_h5120 rmsnorm rmsnorm_h8192_cta_path; do timeout 120 python -m epfl_megatron_b200.ops.selfcheck --one $n 2>/dev/null | tail -1 | cut -c1-260; done
timeout 120 python tools/kernel_bench.py layernorm 2>&1 | grep kernel | cut -c1-260
B200_NORM_IMPL=block timeout 120 python tools/kernel_bench.py layernorm 2>&1 | grep kernel | cut -c1-260
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
