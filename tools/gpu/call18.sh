set -x
timeout 300 python tools/kernel_bench.py attn_bwd > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_kernel -s 12 -c 2 -o gpurun_out/prof_attn_bwd2 python tools/kernel_bench.py attn_bwd > gpurun_out/ncu_attn_bwd2.log 2>&1; echo "ncu rc=$?"
