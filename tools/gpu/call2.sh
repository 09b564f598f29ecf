set -x
timeout 600 python -m epfl_megatron_b200.ops.selfcheck gemm_nt_pair_small gemm_nt_pair gemm_nt_pair_ragged gemm_nn_pair gemm_tn_pair gemm_tt_pair gemm_tn_pair_acc gemm_nt_pair_f32 gemm_nt_big gemm_nt_1cta > gpurun_out/selfcheck_gemm.log 2>&1; echo "selfcheck rc=$?"
grep -v SUMMARY gpurun_out/selfcheck_gemm.log | tail -20
timeout 300 python tools/kernel_bench.py gemm > gpurun_out/kb_gemm2.log 2>&1; echo "kb rc=$?"; cat gpurun_out/kb_gemm2.log | cut -c1-400
