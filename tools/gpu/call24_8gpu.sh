set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_multigpu.py -q -x -k "dp4_fused" > gpurun_out/pytest_dp4.log 2>&1; echo "pytest dp4 rc=$?"; grep -E "KernelError|RuntimeError|passed|failed|Assertion|Error" gpurun_out/pytest_dp4.log | cut -c1-1200 | tail -12
timeout 600 $B --master-port 29511 bench.py --gpus 8 --steps 6 --warmup 3 > gpurun_out/bench_n8_fused.log 2>&1; echo "n8 fused rc=$?"; tail -1 gpurun_out/bench_n8_fused.log | grep -o '"ms_per_step": [0-9.]*'
timeout 600 $B --master-port 29514 bench.py --gpus 8 --steps 6 --warmup 3 --fused_dp off > gpurun_out/bench_n8_nccl.log 2>&1; echo "n8 nccl rc=$?"; tail -1 gpurun_out/bench_n8_nccl.log | grep -o '"ms_per_step": [0-9.]*'
