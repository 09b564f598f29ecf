set -x
timeout 400 python bench.py --steps 3 --warmup 3 --torch_profile > gpurun_out/bench_n1_prof.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n1_prof.log | grep -o '"ms_per_step": [0-9.]*'
