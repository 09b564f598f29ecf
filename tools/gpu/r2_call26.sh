# Round 2 (1 GPU): 7B step, bf16 vs FP8 hybrid linears with identical warm-up / seeds: loss curves and step time.
set -x
timeout 240 python bench.py --steps 6 --warmup 4 > gpurun_out/bench_n1_bf16_w4.log 2>&1; tail -1 gpurun_out/bench_n1_bf16_w4.log | grep -o '"ms_per_step": [0-9.]*\|"losses": [^]]*]'
timeout 240 python bench.py --steps 6 --warmup 4 --fp8 > gpurun_out/bench_n1_fp8.log 2>&1; tail -1 gpurun_out/bench_n1_fp8.log | grep -o '"ms_per_step": [0-9.]*\|"losses": [^]]*]'
timeout 200 python -m pytest tests/test_model_gpu.py -x -q -k fp8 2>&1 | tail -2
