# First 8-GPU call of the next round: confirms what was changed after the GPU budget of round 1 ran out.
#   /usr/local/graft/bin/gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu/next_round_8gpu.sh'
set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
# 1. every multi-GPU parity test (fused DP on DP2 / TP2xDP4 / PP2xDP2, TP2+SP, 1F1B, interleaved, TP2xPP2xDP2)
timeout 600 python -m pytest tests/test_multigpu.py -q > gpurun_out/pytest_multigpu8.log 2>&1; echo "pytest multigpu rc=$?"; tail -3 gpurun_out/pytest_multigpu8.log
# 2. Falcon-40B TP8: auto policy (NCCL all-gather + own GEMM above tp 4, in-kernel reduce-scatter) vs everything in-kernel
timeout 170 $B --master-port 29512 bench.py --gpus 8 --model falcon-40b --tp 8 --seq 2048 --steps 3 --warmup 3 > gpurun_out/bench_falcon40b_tp8_policy.log 2>&1; echo "falcon policy rc=$?"; tail -1 gpurun_out/bench_falcon40b_tp8_policy.log | grep -o '"ms_per_step": [0-9.]*'
B200_FUSED_AG_MAX_TP=8 timeout 170 $B --master-port 29513 bench.py --gpus 8 --model falcon-40b --tp 8 --seq 2048 --steps 3 --warmup 3 > gpurun_out/bench_falcon40b_tp8_allfused.log 2>&1; echo "falcon all fused rc=$?"; tail -1 gpurun_out/bench_falcon40b_tp8_allfused.log | grep -o '"ms_per_step": [0-9.]*'
# 3. headline config once more
timeout 300 $B --master-port 29514 bench.py --gpus 8 --steps 6 --warmup 3 > gpurun_out/bench_n8.log 2>&1; echo "n8 rc=$?"; tail -1 gpurun_out/bench_n8.log | grep -o '"ms_per_step": [0-9.]*'
