set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_multigpu.py -q -x -k "dp2_fused" > gpurun_out/pytest_dp2.log 2>&1; echo "pytest dp2 rc=$?"; grep -E "KernelError|RuntimeError|passed|failed|Assertion|Error" gpurun_out/pytest_dp2.log | cut -c1-1200 | tail -12
timeout 600 $B --master-port 29511 bench.py --gpus 4 --tp 1 --steps 5 --warmup 3 > gpurun_out/bench_dp4_fused_lanes.log 2>&1; echo "dp4 fused lanes rc=$?"; tail -1 gpurun_out/bench_dp4_fused_lanes.log | grep -o '"ms_per_step": [0-9.]*'
B200_DP_LANES=0 timeout 600 $B --master-port 29512 bench.py --gpus 4 --tp 1 --steps 5 --warmup 3 > gpurun_out/bench_dp4_fused_nolanes.log 2>&1; echo "dp4 fused nolanes rc=$?"; tail -1 gpurun_out/bench_dp4_fused_nolanes.log | grep -o '"ms_per_step": [0-9.]*'
timeout 600 $B --master-port 29513 bench.py --gpus 4 --tp 1 --steps 5 --warmup 3 --fused_dp off > gpurun_out/bench_dp4_nccl.log 2>&1; echo "dp4 nccl rc=$?"; tail -1 gpurun_out/bench_dp4_nccl.log | grep -o '"ms_per_step": [0-9.]*'
timeout 600 $B --master-port 29514 bench.py --gpus 4 --tp 1 --steps 5 --warmup 3 --bucket_mb 64 > gpurun_out/bench_dp4_fused_lanes_b64.log 2>&1; echo "dp4 fused lanes 64MB rc=$?"; tail -1 gpurun_out/bench_dp4_fused_lanes_b64.log | grep -o '"ms_per_step": [0-9.]*'
