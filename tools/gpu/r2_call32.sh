# Round 2 (1 GPU): RMSNorm forward with the packed bf16 multiply: exactness vs the fp32 restatement, timing, kernel tests.
set -x
for n in rmsnorm rmsnorm_h1024 rmsnorm_h5120 rmsnorm_h8192_cta_path; do timeout 120 python -m epfl_megatron_b200.ops.selfcheck --one $n 2>/dev/null | tail -1 | cut -c1-160; done
timeout 120 python tools/kernel_bench.py norm 2>&1 | grep '"rmsnorm_fwd"' | grep '"M": 4096, "H": 4096' | cut -c1-200
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
