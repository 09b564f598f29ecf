set -x
timeout 600 python -m epfl_megatron_b200.ops.selfcheck gemm_nt_pair_small gemm_nt_pair gemm_nt_pair_ragged gemm_nn_pair gemm_tn_pair gemm_tt_pair gemm_tn_pair_acc gemm_nt_pair_f32 gemm_nt_big gemm_tn_acc > gpurun_out/selfcheck_gemm.log 2>&1; echo "selfcheck rc=$?"; grep -v SUMMARY gpurun_out/selfcheck_gemm.log | cut -c1-300
timeout 300 python tools/kernel_bench.py gemm > gpurun_out/kb_gemm4.log 2>&1; cat gpurun_out/kb_gemm4.log | cut -c1-420
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_ours.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_ours.log | cut -c1-330
