"""Per-kernel timing on one B200: own kernels vs the library op they replace.

CUDA-event timing, 5 warm-ups, L2 flushed between iterations (a 256 MiB memset).
Writes JSON lines to stdout and to gpurun_out/kernel_bench.jsonl.
Usage: python tools/kernel_bench.py [gemm] [attn] [norm] [elementwise] [adam]
"""
from __future__ import annotations

import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from epfl_megatron_b200.ops import lib as L
from epfl_megatron_b200.ops import selfcheck as SC

PEAKS = {"hbm_gbs": 6448.7, "bf16_tflops": 1658.4}
try:
    with open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) as f:
        PEAKS.update(json.load(f))
except OSError:
    pass

_flush = None


def timeit(fn, iters=10, warmup=5):
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        _flush.zero_()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


OUT = []


def report(**kw):
    OUT.append(kw)
    print(json.dumps(kw), flush=True)


def bench_gemm():
    shapes = [
        ("qkv_7b", 4096, 12288, 4096), ("dense_7b", 4096, 4096, 4096),
        ("h4h_7b", 4096, 22016, 4096), ("4hh_7b", 4096, 4096, 11008),
        ("head_7b", 4096, 32000, 4096), ("sq8k", 8192, 8192, 8192),
    ]
    for name, M, N, K in shapes:
        A = torch.randn(M, K, device="cuda").bfloat16()
        B = torch.randn(N, K, device="cuda").bfloat16()
        C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * M * N * K
        t_own = timeit(lambda: SC.gemm(A, B, M, N, K, 0, 0, 0, c=C, kernel=2))
        t_1cta = timeit(lambda: SC.gemm(A, B, M, N, K, 0, 0, 0, c=C, kernel=1))
        t_lib = timeit(lambda: torch.matmul(A, B.t(), out=C))
        report(kernel="gemm_nt", shape=name, M=M, N=N, K=K, ms_own=t_own, ms_own_1cta=t_1cta,
               ms_cublas=t_lib, tflops_own=fl / t_own / 1e9, tflops_own_1cta=fl / t_1cta / 1e9,
               tflops_cublas=fl / t_lib / 1e9,
               frac_of_measured_peak=fl / t_own / 1e9 / PEAKS["bf16_tflops"])
    # dgrad / wgrad forms on the MLP shape
    M, N, K = 4096, 11008, 4096
    dY = torch.randn(M, N, device="cuda").bfloat16()
    W = torch.randn(N, K, device="cuda").bfloat16()
    X = torch.randn(M, K, device="cuda").bfloat16()
    dX = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
    t_own = timeit(lambda: SC.gemm(dY, W, M, K, N, 0, 1, 0, c=dX))
    t_lib = timeit(lambda: torch.matmul(dY, W, out=dX))
    report(kernel="gemm_dgrad", M=M, N=K, K=N, ms_own=t_own, ms_cublas=t_lib,
           tflops_own=2.0 * M * N * K / t_own / 1e9, tflops_cublas=2.0 * M * N * K / t_lib / 1e9)
    G = torch.zeros(N, K, device="cuda")
    t_own = timeit(lambda: SC.gemm(dY, X, N, K, M, 1, 1, 2, c=G))
    t_1cta = timeit(lambda: SC.gemm(dY, X, N, K, M, 1, 1, 2, c=G, kernel=1))

    def lib_wgrad():
        G.add_(torch.matmul(dY.t(), X))
    t_lib = timeit(lib_wgrad)
    report(kernel="gemm_wgrad_accum_fp32", M=N, N=K, K=M, ms_own=t_own, ms_own_1cta=t_1cta,
           ms_cublas_plus_add=t_lib,
           tflops_own=2.0 * M * N * K / t_own / 1e9, tflops_lib=2.0 * M * N * K / t_lib / 1e9)


def bench_attn():
    try:
        from flash_attn import flash_attn_func
    except Exception as e:  # noqa: BLE001
        flash_attn_func = None
        print("flash_attn unavailable:", e)
    for name, S, B, G, QPG in [("7b_mha", 4096, 1, 32, 1), ("70b_gqa_tp4", 4096, 1, 2, 8),
                               ("falcon40b_tp8", 2048, 1, 1, 16)]:
        HN = 128
        qkv = torch.randn(S, B, G, QPG + 2, HN, device="cuda").bfloat16()
        scale = 1 / math.sqrt(HN)
        fl = 4.0 * S * S * G * QPG * HN * B / 2
        t_own = timeit(lambda: SC.attn_fwd_fused_qkv(qkv, G, QPG, HN, scale, True))
        t_gen1 = timeit(lambda: SC.attn_fwd_fused_qkv(qkv, G, QPG, HN, scale, True, entry="b200_attn_fwd"))
        rec = dict(kernel="attn_fwd", shape=name, S=S, heads=G * QPG, kv_heads=G, ms_own=t_own,
                   tflops_own=fl / t_own / 1e9, ms_own_gen1=t_gen1, tflops_own_gen1=fl / t_gen1 / 1e9)
        if flash_attn_func is not None:
            def ref_path():
                # what the reference does: expand K/V to all heads, [s,b]->[b,s] copies
                q = qkv[:, :, :, :QPG].reshape(S, B, G * QPG, HN)
                k = qkv[:, :, :, QPG:QPG + 1].expand(S, B, G, QPG, HN).reshape(S, B, G * QPG, HN)
                v = qkv[:, :, :, QPG + 1:].expand(S, B, G, QPG, HN).reshape(S, B, G * QPG, HN)
                q, k, v = (t.transpose(0, 1).contiguous() for t in (q, k, v))
                o = flash_attn_func(q, k, v, causal=True)
                return o.transpose(0, 1).contiguous()
            qb = torch.randn(B, S, G * QPG, HN, device="cuda").bfloat16()
            kb = torch.randn(B, S, G, HN, device="cuda").bfloat16()
            t_fa = timeit(lambda: flash_attn_func(qb, kb, kb, causal=True))
            t_ref = timeit(ref_path)
            rec.update(ms_fa2_kernel_only=t_fa, ms_reference_path=t_ref, tflops_fa2=fl / t_fa / 1e9)
        report(**rec)


def bench_attn_bwd():
    from epfl_megatron_b200.ops import functional as Fn
    for name, S, B, G, QPG in [("7b_mha", 4096, 1, 32, 1), ("70b_gqa_tp4", 4096, 1, 2, 8),
                               ("falcon40b_tp8", 2048, 1, 1, 16)]:
        HN = 128
        qkv = torch.randn(S, B, G, QPG + 2, HN, device="cuda").bfloat16()
        dout = torch.randn(S, B, G * QPG * HN, device="cuda").bfloat16()
        scale = 1 / math.sqrt(HN)
        out, lse = SC.attn_fwd_fused_qkv(qkv, G, QPG, HN, scale, True)
        dqkv = torch.empty_like(qkv)
        fl = 10.0 * S * S * G * QPG * HN * B / 2     # 5 GEMMs of the single-pass formulation
        t_own = timeit(lambda: Fn.attn_bwd_fused_qkv(qkv, out, dout, lse, dqkv, scale, True,
                                                     entry="b200_attn_bwd"))
        rec = dict(kernel="attn_bwd", shape=name, S=S, heads=G * QPG, kv_heads=G, ms_own=t_own,
                   tflops_own_5gemm_equiv=fl / t_own / 1e9)
        if os.environ.get("B200_BENCH_ATTN_BWD_V2", "1") == "1":
            try:
                t_v2 = timeit(lambda: Fn.attn_bwd_fused_qkv(qkv, out, dout, lse, dqkv, scale, True,
                                                            entry="b200_attn_bwd_v2"))
                rec.update(ms_own_v2=t_v2, tflops_own_v2_5gemm_equiv=fl / t_v2 / 1e9)
            except Exception as e:  # noqa: BLE001
                rec.update(v2_error=str(e)[:200])
        try:
            k = qkv[:, :, :, QPG].permute(1, 0, 2, 3)
            v = qkv[:, :, :, QPG + 1].permute(1, 0, 2, 3)
            q = qkv[:, :, :, :QPG].reshape(S, B, G * QPG, HN).permute(1, 0, 2, 3)
            o4 = out.view(S, B, G * QPG, HN).permute(1, 0, 2, 3)
            do4 = dout.view(S, B, G * QPG, HN).permute(1, 0, 2, 3)
            dq = torch.empty(B, S, G * QPG, HN, device="cuda", dtype=torch.bfloat16)
            dk = torch.empty(B, S, G, HN, device="cuda", dtype=torch.bfloat16)
            dv = torch.empty_like(dk)
            t_fa = timeit(lambda: Fn._fa2_backward(do4, q, k, v, o4, lse, dq, dk, dv, scale, True))
            rec.update(ms_fa2=t_fa, tflops_fa2=fl / t_fa / 1e9)
        except Exception as e:  # noqa: BLE001
            rec.update(fa2_error=str(e)[:200])
        report(**rec)


def bench_norm():
    """RMSNorm forward / backward, with and without the fused residual add (B200_NORM_IMPL=warp|block
    in the environment selects the older register-resident kernels for an A/B)."""
    impl = os.environ.get("B200_NORM_IMPL", "tma")
    for M, H in ((4096, 4096), (2048, 4096), (4096, 5120), (2048, 8192)):
        x = torch.randn(M, H, device="cuda").bfloat16()
        res = torch.randn(M, H, device="cuda").bfloat16()
        w = torch.ones(H, device="cuda").bfloat16()
        y, h = torch.empty_like(x), torch.empty_like(x)
        rstd = torch.empty(M, device="cuda")
        dy = torch.randn_like(x)
        dx = torch.empty_like(x)
        dw = torch.zeros(H, device="cuda")
        scratch = torch.empty(L.lib().b200_norm_bwd_grid(M), H, device="cuda")
        for use_res in (False, True):
            t = timeit(lambda: L.call("b200_rmsnorm_fwd", x.data_ptr(), L.ptr(res if use_res else None),
                                      w.data_ptr(), y.data_ptr(), L.ptr(h if use_res else None),
                                      rstd.data_ptr(), M, H, 1e-5, L.stream_ptr()))
            by = (4.0 if use_res else 2.0) * M * H * 2
            rec = dict(kernel="rmsnorm_fwd", impl=impl, residual=use_res, M=M, H=H, ms_own=t,
                       gbs_own=by / t / 1e6, frac_of_measured_hbm=by / t / 1e6 / PEAKS["hbm_gbs"])
            if not use_res and (M, H) == (4096, 4096):
                def eager():
                    xf = x.float()
                    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5)).type_as(x) * w
                rec["ms_eager"] = timeit(eager)
            report(**rec)
            t = timeit(lambda: L.call("b200_rmsnorm_bwd", dy.data_ptr(), x.data_ptr(), w.data_ptr(),
                                      rstd.data_ptr(), L.ptr(res if use_res else None), dx.data_ptr(),
                                      dw.data_ptr(), scratch.data_ptr(), M, H, 0, L.stream_ptr()))
            by = (4.0 if use_res else 3.0) * M * H * 2
            report(kernel="rmsnorm_bwd(+dw)", impl=impl, dres=use_res, M=M, H=H, ms_own=t,
                   gbs_own=by / t / 1e6, frac_of_measured_hbm=by / t / 1e6 / PEAKS["hbm_gbs"])


def bench_fp8():
    """FP8 GEMM (tcgen05 kind::f8f6f4, CTA pairs) vs the bf16 kernel and cuBLAS bf16 on the 7B shapes."""
    for (M, N, K) in ((4096, 12288, 4096), (4096, 22016, 4096), (4096, 4096, 11008), (8192, 8192, 8192)):
        A8 = torch.randn(M, K, device="cuda").to(torch.float8_e4m3fn)
        B8 = torch.randn(N, K, device="cuda").to(torch.float8_e4m3fn)
        A = torch.randn(M, K, device="cuda").bfloat16()
        B = torch.randn(N, K, device="cuda").bfloat16()
        C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        one = torch.ones(1, device="cuda")
        t8 = timeit(lambda: L.call("b200_gemm_fp8", A8.data_ptr(), B8.data_ptr(), C.data_ptr(), M, N, K, 0, 0, K, K, N,
                                   0, one.data_ptr(), one.data_ptr(), L.stream_ptr()))
        t16 = timeit(lambda: SC.gemm(A, B, M, N, K, 0, 0, 0, c=C, kernel=2))
        tcb = timeit(lambda: torch.matmul(A, B.t(), out=C))
        fl = 2.0 * M * N * K
        report(kernel="gemm_fp8", M=M, N=N, K=K, ms_fp8=t8, tflops_fp8=fl / t8 / 1e9, ms_bf16_own=t16,
               tflops_bf16_own=fl / t16 / 1e9, ms_bf16_cublas=tcb, tflops_bf16_cublas=fl / tcb / 1e9)
    x = torch.randn(4096, 11008, device="cuda").bfloat16()
    y = torch.empty(4096, 11008, device="cuda", dtype=torch.uint8)
    yT = torch.empty(11008, 4096, device="cuda", dtype=torch.uint8)
    sc, am = torch.ones(1, device="cuda"), torch.zeros(1, device="cuda")
    t = timeit(lambda: L.call("b200_fp8_quantize", x.data_ptr(), y.data_ptr(), yT.data_ptr(), 4096, 11008, 11008,
                              sc.data_ptr(), am.data_ptr(), 0, L.stream_ptr()))
    by = 4096 * 11008 * 4.0
    report(kernel="fp8_quantize(+transpose,+amax)", rows=4096, cols=11008, ms_own=t, gbs_own=by / t / 1e6,
           frac_of_measured_hbm=by / t / 1e6 / PEAKS["hbm_gbs"])


def bench_layernorm():
    impl = os.environ.get("B200_NORM_IMPL", "tma")
    for M, H in ((4096, 4096), (2048, 8192)):
        x = torch.randn(M, H, device="cuda").bfloat16()
        w = torch.ones(H, device="cuda").bfloat16()
        b = torch.zeros(H, device="cuda").bfloat16()
        y, dx = torch.empty_like(x), torch.empty_like(x)
        mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
        dy = torch.randn_like(x)
        dg, db = torch.zeros(H, device="cuda"), torch.zeros(H, device="cuda")
        scratch = torch.empty(L.lib().b200_norm_bwd_grid(M), 2, H, device="cuda")
        t = timeit(lambda: L.call("b200_layernorm_fwd", x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(),
                                  mean.data_ptr(), rstd.data_ptr(), M, H, 1e-5, L.stream_ptr()))
        by = 2.0 * M * H * 2
        report(kernel="layernorm_fwd", impl=impl, M=M, H=H, ms_own=t, gbs_own=by / t / 1e6,
               frac_of_measured_hbm=by / t / 1e6 / PEAKS["hbm_gbs"],
               ms_torch=timeit(lambda: torch.nn.functional.layer_norm(x, (H,), w, b)))
        t = timeit(lambda: L.call("b200_layernorm_bwd", dy.data_ptr(), x.data_ptr(), w.data_ptr(), mean.data_ptr(),
                                  rstd.data_ptr(), dx.data_ptr(), dg.data_ptr(), db.data_ptr(), scratch.data_ptr(),
                                  M, H, 0, L.stream_ptr()))
        by = 3.0 * M * H * 2
        report(kernel="layernorm_bwd(+dgamma,dbeta)", impl=impl, M=M, H=H, ms_own=t, gbs_own=by / t / 1e6,
               frac_of_measured_hbm=by / t / 1e6 / PEAKS["hbm_gbs"])


def bench_elementwise():
    M, F = 4096, 11008
    x = torch.randn(M, 2 * F, device="cuda").bfloat16()
    y = torch.empty(M, F, device="cuda", dtype=torch.bfloat16)
    t_own = timeit(lambda: L.call("b200_glu_fwd", x.data_ptr(), y.data_ptr(), M, F, 3, L.stream_ptr()))

    def eager():
        a, g = x.chunk(2, -1)
        return a * torch.nn.functional.silu(g)
    t_lib = timeit(eager)
    by = 3.0 * M * F * 2
    report(kernel="swiglu_fwd", M=M, F=F, ms_own=t_own, ms_eager=t_lib, gbs_own=by / t_own / 1e6,
           frac_of_measured_hbm=by / t_own / 1e6 / PEAKS["hbm_gbs"])
    dy = torch.randn(M, F, device="cuda").bfloat16()
    dx = torch.empty_like(x)
    t_own = timeit(lambda: L.call("b200_glu_bwd", dy.data_ptr(), x.data_ptr(), dx.data_ptr(), M, F, 3,
                                  L.stream_ptr()))
    by = 5.0 * M * F * 2
    report(kernel="swiglu_bwd", M=M, F=F, ms_own=t_own, gbs_own=by / t_own / 1e6,
           frac_of_measured_hbm=by / t_own / 1e6 / PEAKS["hbm_gbs"])
    S, G, NS, HN = 4096, 32, 3, 128
    qkv = torch.randn(S, 1, G, NS, HN, device="cuda").bfloat16()
    inv = 1.0 / (10000 ** (torch.arange(0, HN, 2, device="cuda").float() / HN))
    ang = torch.outer(torch.arange(S, device="cuda").float(), inv)
    table = torch.stack([ang.cos(), ang.sin()], -1).contiguous()
    t_own = timeit(lambda: L.call("b200_rope", qkv.data_ptr(), table.data_ptr(), None, S, 1, G, NS, NS - 1, HN,
                                  0, S, 0, L.stream_ptr()))
    by = 2.0 * S * G * (NS - 1) * HN * 2
    report(kernel="rope_fused_qkv", S=S, G=G, HN=HN, ms_own=t_own, gbs_own=by / t_own / 1e6,
           frac_of_measured_hbm=by / t_own / 1e6 / PEAKS["hbm_gbs"])
    T, V = 4096, 32000
    logits = torch.randn(T, V, device="cuda").bfloat16()
    labels = torch.randint(0, V, (T,), device="cuda")
    stats = torch.empty(T, 3, device="cuda")
    t_own = timeit(lambda: L.call("b200_xent_stats", logits.data_ptr(), labels.data_ptr(),
                                  stats.data_ptr(), T, V, 0, L.stream_ptr()))
    t_lib = timeit(lambda: torch.nn.functional.cross_entropy(logits.float(), labels, reduction="none"))
    by = 1.0 * T * V * 2
    report(kernel="xent_stats", T=T, V=V, ms_own=t_own, ms_eager_fp32=t_lib, gbs_own=by / t_own / 1e6,
           frac_of_measured_hbm=by / t_own / 1e6 / PEAKS["hbm_gbs"])


def bench_adam():
    n = 1 << 28
    p, g, m, v = (torch.zeros(n, device="cuda") for _ in range(4))
    p16 = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    t_own = timeit(lambda: L.call("b200_adam_step", p.data_ptr(), g.data_ptr(), m.data_ptr(),
                                  v.data_ptr(), p16.data_ptr(), n, 1e-4, 0.9, 0.95, 1e-8, 0.1, 1, 1,
                                  1, None, None, L.stream_ptr()), iters=5, warmup=2)
    by = n * (4 * 4 + 3 * 4 + 2.0)
    report(kernel="adam_flat", n=n, ms_own=t_own, gbs_own=by / t_own / 1e6,
           frac_of_measured_hbm=by / t_own / 1e6 / PEAKS["hbm_gbs"])


if __name__ == "__main__":
    which = sys.argv[1:] or ["gemm", "attn", "attn_bwd", "norm", "elementwise", "adam"]
    for w in which:
        try:
            globals()["bench_" + w]()
        except Exception as e:  # noqa: BLE001
            report(kernel=w, error=f"{type(e).__name__}: {e}"[:500])
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/kernel_bench.jsonl", "a") as f:
        for r in OUT:
            f.write(json.dumps(r) + "\n")
