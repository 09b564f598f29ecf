set -x
timeout 300 python -m epfl_megatron_b200.ops.selfcheck rope glu > gpurun_out/selfcheck_ew.log 2>&1; echo "selfcheck rc=$?"; cut -c1-300 gpurun_out/selfcheck_ew.log
timeout 300 python tools/kernel_bench.py elementwise > gpurun_out/kb_ew.log 2>&1; cut -c1-330 gpurun_out/kb_ew.log
timeout 600 python -m pytest tests/test_kernels_gpu.py -q -x -k "rope or glu or swiglu" > gpurun_out/pytest_ew.log 2>&1; tail -2 gpurun_out/pytest_ew.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n1.log | grep -o '"ms_per_step": [0-9.]*'
