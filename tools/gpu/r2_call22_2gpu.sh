# Round 2 (2 GPUs): every multi-GPU test that fits two devices, on the frozen multi-GPU code.
set -x
timeout 900 python -m pytest tests/test_multigpu.py -q > gpurun_out/pytest_mg2_full.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_mg2_full.log
