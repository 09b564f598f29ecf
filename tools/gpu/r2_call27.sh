# Round 2 (1 GPU): where does the FP8 loss lag come from? wgrad in bf16 (--no_fp8_wgrad) vs all-FP8.
set -x
B200_BENCH_FP8_EXTRA="--no_fp8_wgrad" timeout 240 python bench.py --steps 6 --warmup 4 --fp8 > gpurun_out/bench_n1_fp8_nowgrad.log 2>&1; tail -1 gpurun_out/bench_n1_fp8_nowgrad.log | grep -o '"ms_per_step": [0-9.]*\|"losses": [^]]*]'
