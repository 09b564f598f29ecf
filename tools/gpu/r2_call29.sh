# Round 2 (1 GPU): FP8 loss lag on the 7B synthetic benchmark: delayed scaling with head-room vs current scaling.
set -x
B200_BENCH_FP8_EXTRA="--fp8_margin 2" timeout 240 python bench.py --steps 6 --warmup 4 --fp8 > gpurun_out/bench_n1_fp8_margin2.log 2>&1; tail -1 gpurun_out/bench_n1_fp8_margin2.log | grep -o '"ms_per_step": [0-9.]*\|"losses": [^]]*]'
B200_FP8_CURRENT_SCALING=1 timeout 240 python bench.py --steps 6 --warmup 4 --fp8 > gpurun_out/bench_n1_fp8_current.log 2>&1; tail -1 gpurun_out/bench_n1_fp8_current.log | grep -o '"ms_per_step": [0-9.]*\|"losses": [^]]*]'
