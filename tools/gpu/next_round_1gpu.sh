# First 1-GPU call of the next round: the single-GPU tier the driver runs at round end, after the CPU-only
# changes listed in DESIGN.md section 5.6 item 7.
#   /usr/local/graft/bin/gpurun --timeout 900 -- 'bash tools/gpu/next_round_1gpu.sh'
set -x
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 240 python bench.py --steps 6 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n1.log | grep -o '"ms_per_step": [0-9.]*'
