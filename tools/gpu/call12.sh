set -x
timeout 300 python tools/kernel_bench.py gemm > gpurun_out/kb_gemm3.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_pair_kernel -s 10 -c 2 -o gpurun_out/prof_gemm_pair python tools/kernel_bench.py gemm > gpurun_out/ncu_gemm_pair.log 2>&1; echo "ncu gemm rc=$?"
