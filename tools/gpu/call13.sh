set -x
timeout 600 python -m epfl_megatron_b200.ops.selfcheck rmsnorm rmsnorm_h1024 rmsnorm_h5120 rmsnorm_h8192_cta_path rmsnorm_h1000_cta_path > gpurun_out/selfcheck_norm.log 2>&1; echo "selfcheck rc=$?"; grep -v SUMMARY gpurun_out/selfcheck_norm.log | cut -c1-500
timeout 300 python tools/kernel_bench.py norm > gpurun_out/kb_norm.log 2>&1; cat gpurun_out/kb_norm.log | cut -c1-400
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu 2>&1 | tail -3
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_ours.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_ours.log | cut -c1-330
