# Round 2 (1 GPU): FP8 hybrid linears end to end: tiny-model parity test, then the 7B step bf16 vs fp8 (A/B, not the headline).
set -x
timeout 300 python -m pytest tests/test_model_gpu.py -x -q -k fp8 2>&1 | tail -12 | cut -c1-400
timeout 240 python bench.py --steps 4 --warmup 4 --fp8 > gpurun_out/bench_n1_fp8.log 2>&1; echo "fp8 bench rc=$?"; tail -1 gpurun_out/bench_n1_fp8.log | grep -o '"ms_per_step": [0-9.]*\|"losses": [^]]*]' ; grep -E "Error|error" gpurun_out/bench_n1_fp8.log | tail -3
