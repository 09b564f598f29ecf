set -x
timeout 200 python tools/gpu/ncu_targets.py > gpurun_out/ncu_targets_plain.log 2>&1 && timeout 500 ncu --set full --clock-control none --import-source on -k regex:"gemm_bf16_pair_kernel|attn_fwd2_kernel|glu_bwd_kernel|rope_kernel" -s 5 -c 5 -o gpurun_out/prof_top5 -f python tools/gpu/ncu_targets.py > gpurun_out/ncu_top5.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_top5.log
