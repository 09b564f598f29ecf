# Round 2 (2 GPUs): peer-memory embedding + cross-entropy exchange: TP2 parity tests (SP, no SP, tied), verify mode, headline.
set -x
timeout 700 python -m pytest tests/test_multigpu.py -x -q -k "tp2 or tied" > gpurun_out/pytest_mg2c.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_mg2c.log | cut -c1-300
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $B --master-port 29552 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_n2.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n2.log | grep -o '"ms_per_step": [0-9.]*'; tail -1 gpurun_out/bench_n2.log | grep -o '"losses": [^]]*]'
