# Round 2 (2 GPUs): fused collective GEMMs (epilogue-fused reduce-scatter, all-reduce mode) vs NCCL, then TP2 tests and the headline.
set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $B --master-port 29531 tests/helpers/fused_comm_worker.py --bench > gpurun_out/fused_comm_n2.log 2>&1; echo "fused comm rc=$?"; grep -E "RESULT|Error|error|assert|timed out" gpurun_out/fused_comm_n2.log | cut -c1-4000 | tail -6
timeout 400 python -m pytest tests/test_multigpu.py -x -q -k "tp2 or tied or fused_collective" > gpurun_out/pytest_mg2.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_mg2.log
timeout 300 $B --master-port 29532 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_n2.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n2.log | grep -o '"ms_per_step": [0-9.]*'
