set -x
timeout 300 python tools/sanitize.py --tool memcheck --timeout 280 rope glu gemm_nt_pair_small attn_fwd_one_tile > gpurun_out/sanitize_memcheck.log 2>&1; echo "sanitize rc=$?"; tail -6 gpurun_out/sanitize_memcheck.log | cut -c1-300; head -c 1500 gpurun_out/sanitize_memcheck.raw.log
