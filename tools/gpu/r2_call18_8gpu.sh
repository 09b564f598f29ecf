# Round 2 (8 GPUs), validation of the frozen multi-GPU code: tp 8 fused collectives (batched partial loads), Falcon-40B TP8
# under both reduce-scatter policies, the 8-GPU parity tests, headline TP2 x DP4.
set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 $B --master-port 29531 tests/helpers/fused_comm_worker.py --bench > gpurun_out/fused_comm_n8.log 2>&1; echo "fused comm n8 rc=$?"; grep -E "Error|timed out" gpurun_out/fused_comm_n8.log | tail -3
timeout 170 $B --master-port 29513 bench.py --gpus 8 --model falcon-40b --tp 8 --seq 2048 --steps 3 --warmup 3 > gpurun_out/bench_falcon40b_tp8_rsfused.log 2>&1; echo "falcon RS fused rc=$?"; tail -1 gpurun_out/bench_falcon40b_tp8_rsfused.log | grep -o '"ms_per_step": [0-9.]*'
B200_FUSED_RS_MAX_TP=4 timeout 170 $B --master-port 29512 bench.py --gpus 8 --model falcon-40b --tp 8 --seq 2048 --steps 3 --warmup 3 > gpurun_out/bench_falcon40b_tp8_nccl.log 2>&1; echo "falcon NCCL collectives rc=$?"; tail -1 gpurun_out/bench_falcon40b_tp8_nccl.log | grep -o '"ms_per_step": [0-9.]*'
timeout 200 $B --master-port 29514 bench.py --gpus 8 --steps 6 --warmup 3 > gpurun_out/bench_n8.log 2>&1; echo "n8 rc=$?"; tail -1 gpurun_out/bench_n8.log | grep -o '"ms_per_step": [0-9.]*'
timeout 300 python -m pytest tests/test_multigpu.py -x -q -k "dp4_fused_with_tp2 or tp2_pp2_dp2" > gpurun_out/pytest_mg8.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_mg8.log
