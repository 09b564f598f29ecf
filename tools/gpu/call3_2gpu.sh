set -x
nvidia-smi -L
timeout 900 python -m pytest tests/test_multigpu.py -x -q > gpurun_out/pytest_multigpu.log 2>&1; echo "pytest multigpu rc=$?"; tail -15 gpurun_out/pytest_multigpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_ours_n2.log 2>&1; echo "bench ours n2 rc=$?"; tail -3 gpurun_out/bench_ours_n2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_ref_n2.log 2>&1; echo "bench ref n2 rc=$?"; tail -3 gpurun_out/bench_ref_n2.log
