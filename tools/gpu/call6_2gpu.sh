set -x
timeout 900 python -m pytest tests/test_multigpu.py -q -k "pp2 or tp2" > gpurun_out/pytest_multigpu.log 2>&1; echo "pytest multigpu rc=$?"; grep -E "KernelError|passed|failed" gpurun_out/pytest_multigpu.log | cut -c1-800 | tail -8
