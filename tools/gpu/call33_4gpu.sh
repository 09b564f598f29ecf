set -x
timeout 400 python -m pytest tests/test_multigpu.py -q -x -k "pp2_dp2_fused" > gpurun_out/pytest_pp2dp2.log 2>&1; echo "pytest pp2dp2 rc=$?"; grep -E "KernelError|RuntimeError|passed|failed|Assertion|Error" gpurun_out/pytest_pp2dp2.log | cut -c1-900 | tail -8
