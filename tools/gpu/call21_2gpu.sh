set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_multigpu.py -q -x -k "dp2_fused" > gpurun_out/pytest_dp2.log 2>&1; echo "pytest dp2 rc=$?"; grep -E "KernelError|RuntimeError|passed|failed|Assertion|Error" gpurun_out/pytest_dp2.log | cut -c1-1200 | tail -12
timeout 600 $B --master-port 29511 bench.py --gpus 2 --tp 1 --steps 4 --warmup 3 --fused_dp on > gpurun_out/bench_dp2_fused_ce.log 2>&1; echo "fused ce rc=$?"; tail -1 gpurun_out/bench_dp2_fused_ce.log | grep -o '"ms_per_step": [0-9.]*'
B200_DP_BLOCKS=48 timeout 600 $B --master-port 29512 bench.py --gpus 2 --tp 1 --steps 4 --warmup 3 --fused_dp on > gpurun_out/bench_dp2_fused_ce48.log 2>&1; echo "fused ce b48 rc=$?"; tail -1 gpurun_out/bench_dp2_fused_ce48.log | grep -o '"ms_per_step": [0-9.]*'
B200_DP_NO_OVERLAP=1 timeout 600 $B --master-port 29513 bench.py --gpus 2 --tp 1 --steps 4 --warmup 3 --fused_dp on > gpurun_out/bench_dp2_fused_noov.log 2>&1; echo "fused no-overlap rc=$?"; tail -1 gpurun_out/bench_dp2_fused_noov.log | grep -o '"ms_per_step": [0-9.]*'
timeout 600 $B --master-port 29514 bench.py --gpus 2 --tp 1 --steps 4 --warmup 3 --fused_dp off > gpurun_out/bench_dp2_nccl.log 2>&1; echo "nccl rc=$?"; tail -1 gpurun_out/bench_dp2_nccl.log | grep -o '"ms_per_step": [0-9.]*'
