# Round 2, call 3 (1 GPU): grouped-consumer RMSNorm ring: numerics + kernel bench + step time.
set -x
for n in rmsnorm rmsnorm_h1024 rmsnorm_h5120 rmsnorm_h8192_cta_path rmsnorm_h1000_cta_path; do timeout 120 python -m epfl_megatron_b200.ops.selfcheck --one $n 2>/dev/null | tail -1 | cut -c1-200; done
timeout 120 python tools/kernel_bench.py norm > gpurun_out/kb_norm_tma2.jsonl 2>&1; cut -c1-220 gpurun_out/kb_norm_tma2.jsonl
timeout 200 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
timeout 240 python bench.py --steps 6 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n1.log | grep -o '"ms_per_step": [0-9.]*'
