# Round 2 (4 GPUs): the 4-GPU-only parity tests and the headline TP2 x DP2 on the frozen multi-GPU code
# (host-side tensor-parallel group with two TP groups, interleaved tile order).
set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 500 python -m pytest tests/test_multigpu.py -x -q -k "pp2_dp2 or fused_collective_gemms_match_nccl[4]" > gpurun_out/pytest_mg4.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_mg4.log
timeout 200 $B --master-port 29514 bench.py --gpus 4 --steps 6 --warmup 3 > gpurun_out/bench_n4.log 2>&1; echo "n4 rc=$?"; tail -1 gpurun_out/bench_n4.log | grep -o '"ms_per_step": [0-9.]*'
