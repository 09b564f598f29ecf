set -x
timeout 600 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 400 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_n1_final.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n1_final.log | grep -o '"ms_per_step": [0-9.]*'
timeout 200 python tools/kernel_bench.py elementwise > gpurun_out/kb_ew2.log 2>&1; grep -E "swiglu" gpurun_out/kb_ew2.log | cut -c1-200
timeout 400 python tools/sanitize.py --tool memcheck --timeout 360 > gpurun_out/sanitize_memcheck.log 2>&1; echo "sanitize rc=$?"; tail -16 gpurun_out/sanitize_memcheck.log | cut -c1-300
