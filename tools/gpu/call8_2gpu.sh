set -x
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tests/helpers/fused_comm_worker.py --bench > gpurun_out/fused_comm_n2.log 2>&1; echo "fused comm rc=$?"; grep -E "RESULT|Error|error|assert" gpurun_out/fused_comm_n2.log | cut -c1-3000 | tail -12
