# Round 2 (4 GPUs): fused collective GEMMs at tp 4 (batched partial loads), 4-GPU parity tests, headline TP2xDP2 + reference arm.
set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 200 $B --master-port 29531 tests/helpers/fused_comm_worker.py --bench > gpurun_out/fused_comm_n4.log 2>&1; echo "fused comm n4 rc=$?"; grep -E "Error|timed out" gpurun_out/fused_comm_n4.log | tail -3
timeout 500 python -m pytest tests/test_multigpu.py -x -q -k "pp2_dp2 or fused_collective_gemms_match_nccl[4] or interleaved" > gpurun_out/pytest_mg4.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_mg4.log
timeout 200 $B --master-port 29514 bench.py --gpus 4 --steps 6 --warmup 3 > gpurun_out/bench_n4.log 2>&1; echo "n4 rc=$?"; tail -1 gpurun_out/bench_n4.log | grep -o '"ms_per_step": [0-9.]*'
timeout 300 $B --master-port 29515 bench.py --impl reference --gpus 4 --steps 4 --warmup 3 > gpurun_out/bench_ref_n4.log 2>&1; echo "ref n4 rc=$?"; tail -1 gpurun_out/bench_ref_n4.log | cut -c1-400
