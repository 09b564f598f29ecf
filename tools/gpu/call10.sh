set -x
timeout 900 python bench.py --steps 3 --warmup 3 --torch_profile > gpurun_out/bench_ours.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_ours.log | cut -c1-300
head -50 gpurun_out/torch_profile_n1.md | cut -c1-200
