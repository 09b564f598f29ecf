# Round 2, call 1 (1 GPU): GPU test tier, smoke, headline both arms, norm kernel baseline.
set -x
timeout 700 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 300 python bench.py --impl reference --steps 6 --warmup 3 > gpurun_out/bench_ref_n1.log 2>&1; echo "ref rc=$?"; tail -1 gpurun_out/bench_ref_n1.log | cut -c1-300
timeout 240 python bench.py --steps 6 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n1.log | grep -o '"ms_per_step": [0-9.]*'
timeout 120 python tools/kernel_bench.py norm > gpurun_out/kb_norm.jsonl 2>&1; cat gpurun_out/kb_norm.jsonl | cut -c1-250
