# Round 2 (2 GPUs): TP2 without sequence parallelism on the fused all-reduce GEMMs; fused-path evidence via B200_COMM_VERIFY.
set -x
timeout 500 python -m pytest tests/test_multigpu.py -x -q -k "without_sequence_parallel or tp2_sequence" > gpurun_out/pytest_mg2b.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_mg2b.log
B200_COMM_VERIFY=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tests/helpers/run_tiny.py --tp=2 --no_sp --steps=2 --num_microbatches=2 2>&1 | tail -3
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $B --master-port 29542 bench.py --gpus 2 --steps 4 --warmup 3 --no_sp > gpurun_out/bench_n2_nosp_fused.log 2>&1; tail -1 gpurun_out/bench_n2_nosp_fused.log | grep -o '"ms_per_step": [0-9.]*'
timeout 300 $B --master-port 29543 bench.py --gpus 2 --steps 4 --warmup 3 --no_sp --fused_comm off > gpurun_out/bench_n2_nosp_nccl.log 2>&1; tail -1 gpurun_out/bench_n2_nosp_nccl.log | grep -o '"ms_per_step": [0-9.]*'
