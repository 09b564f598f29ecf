set -x
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tests/helpers/fused_comm_worker.py --bench > gpurun_out/fused_comm_n2.log 2>&1; echo "fused comm rc=$?"; grep -E "RESULT|Error|error|assert" gpurun_out/fused_comm_n2.log | cut -c1-3000 | tail -12
timeout 900 python -m pytest tests/test_multigpu.py -q -k "pp2 or tp2 or dp2" > gpurun_out/pytest_multigpu.log 2>&1; echo "pytest multigpu rc=$?"; grep -E "KernelError|passed|failed" gpurun_out/pytest_multigpu.log | cut -c1-800 | tail -8
