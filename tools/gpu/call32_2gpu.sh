set -x
B200_COMM_POISON=1 B200_COMM_VERIFY=1 timeout 400 python -m pytest tests/test_multigpu.py -q -x -k "tp2_sequence" > gpurun_out/pytest_tp2_verify.log 2>&1; echo "pytest tp2 poison+verify rc=$?"; grep -E "KernelError|RuntimeError|passed|failed|Assertion|Error" gpurun_out/pytest_tp2_verify.log | cut -c1-600 | tail -8
