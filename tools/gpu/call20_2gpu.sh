set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $B --master-port 29511 bench.py --gpus 2 --tp 1 --steps 3 --warmup 3 --fused_dp on --torch_profile > gpurun_out/bench_dp2_fused_b24.log 2>&1; echo "fused b24 rc=$?"; tail -1 gpurun_out/bench_dp2_fused_b24.log | grep -o '"ms_per_step": [0-9.]*'
cp gpurun_out/torch_profile_n2.md gpurun_out/torch_profile_n2_dpfused.md
B200_DP_BLOCKS=96 timeout 600 $B --master-port 29512 bench.py --gpus 2 --tp 1 --steps 3 --warmup 3 --fused_dp on > gpurun_out/bench_dp2_fused_b96.log 2>&1; echo "fused b96 rc=$?"; tail -1 gpurun_out/bench_dp2_fused_b96.log | grep -o '"ms_per_step": [0-9.]*'
B200_DP_NO_OVERLAP=1 timeout 600 $B --master-port 29513 bench.py --gpus 2 --tp 1 --steps 3 --warmup 3 --fused_dp on > gpurun_out/bench_dp2_fused_noov.log 2>&1; echo "fused no-overlap rc=$?"; tail -1 gpurun_out/bench_dp2_fused_noov.log | grep -o '"ms_per_step": [0-9.]*'
grep -E "dp_reduce|adam_push|adam_kernel|signal_kernel|dp_wait|nccl" gpurun_out/torch_profile_n2_dpfused.md | cut -c1-160
