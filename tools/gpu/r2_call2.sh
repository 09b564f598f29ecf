# Round 2, call 2 (1 GPU): bulk-copy-ring RMSNorm: numerics, A/B against the register-resident kernels, step time.
set -x
for n in rmsnorm rmsnorm_h1024 rmsnorm_h5120 rmsnorm_h8192_cta_path rmsnorm_h1000_cta_path; do timeout 120 python -m epfl_megatron_b200.ops.selfcheck --one $n | tail -1 | cut -c1-400; done
timeout 200 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
timeout 120 python tools/kernel_bench.py norm > gpurun_out/kb_norm_tma.jsonl 2>&1; cut -c1-220 gpurun_out/kb_norm_tma.jsonl
B200_NORM_IMPL=warp timeout 120 python tools/kernel_bench.py norm > gpurun_out/kb_norm_warp.jsonl 2>&1; cut -c1-220 gpurun_out/kb_norm_warp.jsonl
timeout 240 python bench.py --steps 6 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n1.log | grep -o '"ms_per_step": [0-9.]*'
