set -x
python -m pytest tests -m "gpu and not multigpu" -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_ours.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_ours.log
python bench.py --profile --debug_layers 2 --steps 1 --warmup 1 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python bench.py --profile --debug_layers 2 --steps 1 --warmup 1 > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python tools/kernel_bench.py gemm > gpurun_out/kb_gemm.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_kernel -s 10 -c 2 -o gpurun_out/prof_gemm python tools/kernel_bench.py gemm > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
python tools/kernel_bench.py attn > gpurun_out/kb_attn.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:attn -s 6 -c 1 -o gpurun_out/prof_attn python tools/kernel_bench.py attn > gpurun_out/ncu_attn.log 2>&1; echo "ncu attn rc=$?"
