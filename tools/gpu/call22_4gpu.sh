set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 600 $B --master-port 29511 bench.py --gpus 4 --steps 6 --warmup 3 --fused_dp on > gpurun_out/bench_n4_fused_ce96.log 2>&1; echo "n4 fused ce96 rc=$?"; tail -1 gpurun_out/bench_n4_fused_ce96.log | grep -o '"ms_per_step": [0-9.]*'
timeout 600 $B --master-port 29514 bench.py --gpus 4 --steps 6 --warmup 3 --fused_dp off > gpurun_out/bench_n4_nccl.log 2>&1; echo "n4 nccl rc=$?"; tail -1 gpurun_out/bench_n4_nccl.log | grep -o '"ms_per_step": [0-9.]*'
B200_DP_BLOCKS=48 timeout 600 $B --master-port 29512 bench.py --gpus 4 --steps 6 --warmup 3 --fused_dp on > gpurun_out/bench_n4_fused_ce48.log 2>&1; echo "n4 fused ce48 rc=$?"; tail -1 gpurun_out/bench_n4_fused_ce48.log | grep -o '"ms_per_step": [0-9.]*'
B200_DP_NO_OVERLAP=1 timeout 600 $B --master-port 29513 bench.py --gpus 4 --steps 6 --warmup 3 --fused_dp on > gpurun_out/bench_n4_fused_noov.log 2>&1; echo "n4 fused no-overlap rc=$?"; tail -1 gpurun_out/bench_n4_fused_noov.log | grep -o '"ms_per_step": [0-9.]*'
