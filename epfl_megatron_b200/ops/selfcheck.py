"""Numerical self-checks of every CUDA kernel against plain PyTorch fp32 math.

Each ``check_*`` function raises AssertionError on mismatch and returns a small dict of
error statistics. They are used by ``tests/test_kernels_gpu.py`` (pytest, ``-m gpu``) and
by ``python -m epfl_megatron_b200.ops.selfcheck`` which runs every check in its own
subprocess under a timeout, so one faulting kernel cannot take the others down.
"""
from __future__ import annotations

import json
import math
import subprocess
import sys
import time

import torch

from . import lib as L


def _dev():
    return torch.device("cuda", torch.cuda.current_device())


def _close(a, b, atol, rtol, what):
    a = a.detach().float()
    b = b.detach().float()
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = (err > tol)
    stats = {"max_abs_err": float(err.max()), "ref_max": float(b.abs().max())}
    assert not bool(bad.any()), f"{what}: {int(bad.sum())} / {bad.numel()} mismatches, {stats}"
    assert torch.isfinite(a).all(), f"{what}: non-finite output"
    return stats


def check_rmsnorm(M=1024, H=4096):
    dev = _dev()
    torch.manual_seed(0)
    x = torch.randn(M, H, device=dev).bfloat16()
    res = torch.randn(M, H, device=dev).bfloat16()
    w = (1 + 0.1 * torch.randn(H, device=dev)).bfloat16()
    dy = torch.randn(M, H, device=dev).bfloat16()
    out = {}
    for use_res in (False, True):
        y = torch.empty_like(x)
        h = torch.empty_like(x) if use_res else None
        rstd = torch.empty(M, device=dev, dtype=torch.float32)
        L.call("b200_rmsnorm_fwd", x.data_ptr(), L.ptr(res if use_res else None), w.data_ptr(),
               y.data_ptr(), L.ptr(h), rstd.data_ptr(), M, H, 1e-5, L.stream_ptr())
        hin = (x.float() + res.float()).bfloat16() if use_res else x
        hf = hin.float().requires_grad_(True)
        wf = w.float().requires_grad_(True)
        yr = (hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + 1e-5)) * wf
        out[f"fwd_res{int(use_res)}"] = _close(y, yr, 2e-2, 2e-2, "rmsnorm fwd")
        if use_res:
            _close(h, hin, 0, 0, "rmsnorm residual out")
        yr.backward(dy.float())
        grid = L.lib().b200_norm_bwd_grid(M)
        scratch = torch.empty(grid, H, device=dev, dtype=torch.float32)
        dx = torch.empty_like(x)
        dw = torch.zeros(H, device=dev, dtype=torch.float32)
        L.call("b200_rmsnorm_bwd", dy.data_ptr(), hin.data_ptr(), w.data_ptr(), rstd.data_ptr(),
               None, dx.data_ptr(), dw.data_ptr(), scratch.data_ptr(), M, H, 0, L.stream_ptr(),
               launches=2)
        out[f"dx_res{int(use_res)}"] = _close(dx, hf.grad, 2e-2, 2e-2, "rmsnorm dx")
        out[f"dw_res{int(use_res)}"] = _close(dw, wf.grad, 0.5, 2e-2, "rmsnorm dw")
        # the kernel multiplies dy by the bf16-ROUNDED xhat (what the forward fed to the weight):
        # against that restatement only the fp32 summation order differs, so a dropped row shows
        dw_exact = (dy.float() * (hin.float() * rstd[:, None]).bfloat16().float()).sum(0)
        _close(dw, dw_exact, 1e-3 * float(dw_exact.abs().max()), 1e-3, "rmsnorm dw (rounded xhat)")
        # gradient arriving through the residual branch + accumulation into dw
        dres = torch.randn(M, H, device=dev).bfloat16()
        dx2 = torch.empty_like(x)
        L.call("b200_rmsnorm_bwd", dy.data_ptr(), hin.data_ptr(), w.data_ptr(), rstd.data_ptr(),
               dres.data_ptr(), dx2.data_ptr(), dw.data_ptr(), scratch.data_ptr(), M, H, 1, L.stream_ptr(),
               launches=2)
        _close(dx2, hf.grad + dres.float(), 3e-2, 2e-2, "rmsnorm dx + dres")
        _close(dw, 2 * wf.grad, 1.0, 2e-2, "rmsnorm dw accumulate")
    return out


def check_layernorm(M=512, H=8192):
    dev = _dev()
    torch.manual_seed(1)
    x = torch.randn(M, H, device=dev).bfloat16()
    w = (1 + 0.1 * torch.randn(H, device=dev)).bfloat16()
    b = (0.1 * torch.randn(H, device=dev)).bfloat16()
    dy = torch.randn(M, H, device=dev).bfloat16()
    y = torch.empty_like(x)
    mean = torch.empty(M, device=dev)
    rstd = torch.empty(M, device=dev)
    L.call("b200_layernorm_fwd", x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(),
           mean.data_ptr(), rstd.data_ptr(), M, H, 1e-5, L.stream_ptr())
    xf, wf, bf = (t.float().requires_grad_(True) for t in (x, w, b))
    yr = torch.nn.functional.layer_norm(xf, (H,), wf, bf, 1e-5)
    out = {"fwd": _close(y, yr, 2e-2, 2e-2, "layernorm fwd")}
    yr.backward(dy.float())
    grid = L.lib().b200_norm_bwd_grid(M)
    scratch = torch.empty(grid, 2, H, device=dev)
    dx = torch.empty_like(x)
    dg = torch.zeros(H, device=dev)
    db = torch.zeros(H, device=dev)
    L.call("b200_layernorm_bwd", dy.data_ptr(), x.data_ptr(), w.data_ptr(), mean.data_ptr(),
           rstd.data_ptr(), dx.data_ptr(), dg.data_ptr(), db.data_ptr(), scratch.data_ptr(), M, H,
           0, L.stream_ptr(), launches=2)
    out["dx"] = _close(dx, xf.grad, 2e-2, 2e-2, "layernorm dx")
    out["dg"] = _close(dg, wf.grad, 0.5, 2e-2, "layernorm dgamma")
    out["db"] = _close(db, bf.grad, 0.5, 2e-2, "layernorm dbeta")
    return out


def _rope_ref(x, table, pos):
    # x [..., hn] fp32, interleaved pairs; table [P, hn/2, 2]
    xc = torch.view_as_complex(x.float().reshape(*x.shape[:-1], -1, 2))
    f = torch.view_as_complex(table[pos].contiguous())
    return torch.view_as_real(xc * f).flatten(-2)


def check_rope(S=512, B=2, G=4, QPG=4, HN=128):
    dev = _dev()
    torch.manual_seed(2)
    nslot = QPG + 2
    x = torch.randn(S, B, G, nslot, HN, device=dev).bfloat16()
    inv = 1.0 / (10000 ** (torch.arange(0, HN, 2, device=dev).float() / HN))
    ang = torch.outer(torch.arange(S, device=dev).float(), inv)
    table = torch.stack([ang.cos(), ang.sin()], -1).contiguous()
    ref = x.clone().float()
    pos = torch.arange(S, device=dev)
    f = table[pos][:, None, None, None]  # [S,1,1,1,hn/2,2]
    xr = ref[..., :QPG + 1, :].reshape(S, B, G, QPG + 1, HN // 2, 2)
    rot = torch.stack([xr[..., 0] * f[..., 0] - xr[..., 1] * f[..., 1],
                       xr[..., 0] * f[..., 1] + xr[..., 1] * f[..., 0]], -1).flatten(-2)
    ref[..., :QPG + 1, :] = rot
    y = x.clone()
    L.call("b200_rope", y.data_ptr(), table.data_ptr(), None, S, B, G, nslot, QPG + 1, HN, 0, S, 0,
           L.stream_ptr())
    out = {"fwd": _close(y, ref, 2e-2, 1e-2, "rope fwd")}
    assert torch.equal(y[..., QPG + 1, :], x[..., QPG + 1, :]), "rope touched V"
    # inverse rotation restores the input (up to bf16 rounding)
    L.call("b200_rope", y.data_ptr(), table.data_ptr(), None, S, B, G, nslot, QPG + 1, HN, 0, S, 1,
           L.stream_ptr())
    out["roundtrip"] = _close(y, x, 5e-2, 2e-2, "rope inverse")
    return out


def check_glu(M=2048, F=5504):
    dev = _dev()
    torch.manual_seed(3)
    acts = {0: lambda g: g, 1: torch.nn.functional.gelu, 2: torch.relu,
            3: torch.nn.functional.silu}
    out = {}
    for kind, act in acts.items():
        x = torch.randn(M, 2 * F, device=dev).bfloat16()
        dy = torch.randn(M, F, device=dev).bfloat16()
        y = torch.empty(M, F, device=dev, dtype=torch.bfloat16)
        L.call("b200_glu_fwd", x.data_ptr(), y.data_ptr(), M, F, kind, L.stream_ptr())
        xf = x.float().requires_grad_(True)
        a, g = xf.chunk(2, -1)
        yr = a * act(g)
        out[f"fwd{kind}"] = _close(y, yr, 2e-2, 1e-2, f"glu{kind} fwd")
        yr.backward(dy.float())
        dx = torch.empty_like(x)
        L.call("b200_glu_bwd", dy.data_ptr(), x.data_ptr(), dx.data_ptr(), M, F, kind,
               L.stream_ptr())
        out[f"bwd{kind}"] = _close(dx, xf.grad, 3e-2, 1e-2, f"glu{kind} bwd")
    return out


def check_bias_gelu(M=1024, H=4096):
    dev = _dev()
    torch.manual_seed(4)
    out = {}
    for tanh in (0, 1):
        for use_bias in (False, True):
            x = torch.randn(M, H, device=dev).bfloat16()
            b = torch.randn(H, device=dev).bfloat16() if use_bias else None
            dy = torch.randn(M, H, device=dev).bfloat16()
            y = torch.empty_like(x)
            L.call("b200_bias_gelu_fwd", x.data_ptr(), L.ptr(b), y.data_ptr(), M, H, tanh,
                   L.stream_ptr())
            xf = x.float().requires_grad_(True)
            t = xf + (b.float() if use_bias else 0)
            yr = torch.nn.functional.gelu(t, approximate="tanh" if tanh else "none")
            out[f"fwd{tanh}{int(use_bias)}"] = _close(y, yr, 2e-2, 1e-2, "gelu fwd")
            yr.backward(dy.float())
            dx = torch.empty_like(x)
            L.call("b200_bias_gelu_bwd", dy.data_ptr(), x.data_ptr(), L.ptr(b), dx.data_ptr(), M,
                   H, tanh, L.stream_ptr())
            out[f"bwd{tanh}{int(use_bias)}"] = _close(dx, xf.grad, 3e-2, 1e-2, "gelu bwd")
    return out


def check_dropout_add(M=1024, H=4096, p=0.25):
    dev = _dev()
    torch.manual_seed(5)
    x = (torch.randn(M, H, device=dev).abs() + 0.5).bfloat16()
    zero = torch.zeros(M, H, device=dev, dtype=torch.bfloat16)
    r = torch.randn(M, H, device=dev).bfloat16()
    y = torch.empty_like(x)
    # residual = 0 makes the mask exactly observable
    L.call("b200_dropout_add_fwd", x.data_ptr(), None, zero.data_ptr(), y.data_ptr(), M, H, p, 1234,
           7, L.stream_ptr())
    kept = y.float() != 0
    frac = float(kept.float().mean())
    assert abs(frac - (1 - p)) < 0.01, f"keep fraction {frac}"
    _close(y.float()[kept], (x.float() / (1 - p))[kept], 5e-2, 2e-2, "dropout scale")
    dy = torch.ones_like(x)
    dx = torch.empty_like(x)
    L.call("b200_dropout_bwd", dy.data_ptr(), dx.data_ptr(), M, H, p, 1234, 7, L.stream_ptr())
    assert bool(((dx.float() > 0) == kept).all()), "dropout bwd mask differs from fwd mask"
    # a different offset must give a different mask
    L.call("b200_dropout_bwd", dy.data_ptr(), dx.data_ptr(), M, H, p, 1234, 8, L.stream_ptr())
    assert not bool(((dx.float() > 0) == kept).all()), "dropout mask ignores the offset"
    # p = 0 must be an exact add
    L.call("b200_dropout_add_fwd", x.data_ptr(), None, r.data_ptr(), y.data_ptr(), M, H, 0.0, 1, 1,
           L.stream_ptr())
    _close(y, (x.float() + r.float()), 2e-2, 1e-2, "dropout p=0")
    return {"keep_frac": frac}


def check_softmax(b=2, np_=4, sq=256, sk=256):
    dev = _dev()
    torch.manual_seed(6)
    out = {}
    scale = 0.3
    x = torch.randn(b, np_, sq, sk, device=dev).bfloat16() * 4
    dy = torch.randn(b, np_, sq, sk, device=dev).bfloat16()
    for mode in (0, 1, 2):
        mask = None
        xf = x.float().requires_grad_(True)
        s = xf * scale
        if mode == 1:
            tri = torch.ones(sq, sk, device=dev, dtype=torch.bool).triu(1)
            yr = torch.softmax(s.masked_fill(tri, float("-inf")), -1)
        elif mode == 2:
            mask = (torch.rand(b, 1, sq, sk, device=dev) < 0.3)
            mask[:, :, 3, :] = True  # a fully masked row must come out as zeros
            yr = torch.softmax(s.masked_fill(mask, -10000.0), -1)
            yr = yr * (~mask.all(-1, keepdim=True)).float()
        else:
            yr = torch.softmax(s, -1)
        y = torch.empty_like(x)
        m8 = mask.to(torch.uint8).contiguous() if mask is not None else None
        L.call("b200_softmax_fwd", x.data_ptr(), L.ptr(m8), y.data_ptr(), b, np_, sq, sk,
               b if mask is not None else 1, scale, mode, L.stream_ptr())
        out[f"fwd{mode}"] = _close(y, yr, 1e-3 * 4, 2e-2, f"softmax{mode} fwd")
        if mode == 2:
            assert float(y[:, :, 3, :].abs().max()) == 0.0, "fully masked row not zero"
        if mode == 1:
            assert float(y[0, 0].triu(1).abs().max()) == 0.0, "upper triangle not zero"
        yr.backward(dy.float())
        dx = torch.empty_like(x)
        L.call("b200_softmax_bwd", dy.data_ptr(), y.data_ptr(), dx.data_ptr(), b * np_ * sq, sk,
               scale, L.stream_ptr())
        out[f"bwd{mode}"] = _close(dx, xf.grad, 4e-3, 3e-2, f"softmax{mode} bwd")
    return out


def check_xent(T=1024, V=32000, tp=2):
    dev = _dev()
    torch.manual_seed(7)
    logits = (torch.randn(T, V, device=dev) * 3).bfloat16()
    labels = torch.randint(0, V, (T,), device=dev)
    Vp = V // tp
    stats = torch.empty(tp, T, 3, device=dev)
    for r in range(tp):
        shard = logits[:, r * Vp:(r + 1) * Vp].contiguous()
        L.call("b200_xent_stats", shard.data_ptr(), labels.data_ptr(), stats[r].data_ptr(), T, Vp,
               r * Vp, L.stream_ptr())
    loss = torch.empty(T, device=dev)
    lse = torch.empty(T, device=dev)
    L.call("b200_xent_combine", stats.data_ptr(), loss.data_ptr(), lse.data_ptr(), T, tp,
           L.stream_ptr())
    lf = logits.float().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lf, labels, reduction="none")
    out = {"loss": _close(loss, ref, 2e-3, 1e-3, "xent loss")}
    g = torch.rand(T, device=dev)
    ref.backward(g)
    for r in range(tp):
        shard = logits[:, r * Vp:(r + 1) * Vp].contiguous()
        L.call("b200_xent_bwd", shard.data_ptr(), labels.data_ptr(), lse.data_ptr(), g.data_ptr(),
               T, Vp, r * Vp, L.stream_ptr())
        out[f"bwd{r}"] = _close(shard, lf.grad[:, r * Vp:(r + 1) * Vp], 2e-3, 2e-2, "xent bwd")
    return out


def check_adam(n=1 << 22):
    dev = _dev()
    torch.manual_seed(8)
    p = torch.randn(n, device=dev)
    g = torch.randn(n, device=dev) * 0.1
    m = torch.zeros(n, device=dev)
    v = torch.zeros(n, device=dev)
    p16 = torch.empty(n, device=dev, dtype=torch.bfloat16)
    pr = p.clone().requires_grad_(True)
    opt = torch.optim.AdamW([pr], lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    coef = torch.tensor([0.5], device=dev)
    skip = torch.zeros(1, device=dev)
    for step in (1, 2, 3):
        pr.grad = g * 0.5
        opt.step()
        L.call("b200_adam_step", p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(),
               p16.data_ptr(), n, 1e-3, 0.9, 0.95, 1e-8, 0.1, step, 1, 1, coef.data_ptr(),
               skip.data_ptr(), L.stream_ptr())
    out = {"p": _close(p, pr.detach(), 1e-5, 1e-4, "adam p"),
           "p16": _close(p16, pr.detach(), 2e-2, 1e-2, "adam p16")}
    skip.fill_(1.0)
    before = p.clone()
    L.call("b200_adam_step", p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(),
           p16.data_ptr(), n, 1e-3, 0.9, 0.95, 1e-8, 0.1, 4, 1, 1, coef.data_ptr(), skip.data_ptr(),
           L.stream_ptr())
    assert torch.equal(before, p), "adam did not honour skip flag"
    # sumsq / scale
    res = torch.zeros(2, device=dev)
    scratch = torch.empty(L.lib().b200_sumsq_scratch_elems(), device=dev)
    L.call("b200_sumsq", g.data_ptr(), n, res.data_ptr(), scratch.data_ptr(), 0, L.stream_ptr(),
           launches=2)
    out["sumsq"] = _close(res[0], (g.double() ** 2).sum().float(), 0, 1e-4, "sumsq")
    gg = g.clone()
    L.call("b200_scale", gg.data_ptr(), n, coef.data_ptr(), 2.0, L.stream_ptr())
    out["scale"] = _close(gg, g, 0, 1e-6, "scale")
    return out


def check_casts(n=1 << 20):
    dev = _dev()
    src = torch.randn(n, device=dev).bfloat16()
    dst = torch.ones(n, device=dev)
    L.call("b200_accum_bf16_to_f32", src.data_ptr(), dst.data_ptr(), n, 0.5, 1, L.stream_ptr())
    _close(dst, 1 + 0.5 * src.float(), 1e-6, 1e-6, "accum")
    o = torch.empty(n, device=dev, dtype=torch.bfloat16)
    L.call("b200_cast_f32_to_bf16", dst.data_ptr(), o.data_ptr(), n, L.stream_ptr())
    _close(o, dst.bfloat16(), 0, 0, "cast")
    return {}


def gemm(a, b, M, N, K, a_mn, b_mn, out_mode, c=None, max_ctas=0, kernel=0):
    """Thin wrapper used by checks and by ops.linear. kernel: 0 auto, 1 single CTA, 2 CTA pair."""
    if c is None:
        c = torch.empty(M, N, device=a.device,
                        dtype=torch.bfloat16 if out_mode == 0 else torch.float32)
    L.call("b200_gemm_bf16_variant", a.data_ptr(), b.data_ptr(), c.data_ptr(), M, N, K, a_mn, b_mn,
           a.stride(0), b.stride(0), c.stride(0), out_mode, max_ctas, kernel, L.stream_ptr())
    return c


def check_gemm(variant="nt", M=512, N=768, K=512, out_mode=0, kernel=0):
    """variant: nt (fwd), nn (dgrad: B MN-major), tn (wgrad: both MN-major), tt."""
    dev = _dev()
    torch.manual_seed(9)
    a_mn = variant[0] == "t"
    b_mn = variant[1] == "n"
    A = torch.randn(M, K, device=dev).bfloat16()
    Bm = torch.randn(N, K, device=dev).bfloat16()
    ref = A.float() @ Bm.float().t()
    a_st = A.t().contiguous() if a_mn else A
    b_st = Bm.t().contiguous() if b_mn else Bm
    if out_mode == 2:
        c = torch.randn(M, N, device=dev)
        ref = ref + c
        out = gemm(a_st, b_st, M, N, K, int(a_mn), int(b_mn), 2, c=c, kernel=kernel)
    else:
        out = gemm(a_st, b_st, M, N, K, int(a_mn), int(b_mn), out_mode, kernel=kernel)
    tol = 0.02 * math.sqrt(K) if out_mode == 0 else 1e-2 * math.sqrt(K) * 0.1
    return _close(out, ref, tol, 1e-2, f"gemm {variant} {M}x{N}x{K} mode{out_mode}")


def check_fp8_quantize(rows=1024, cols=2048, e5m2=False):
    """One-pass cast (+ transpose, + amax) against torch's float8 conversion of the same scaled values."""
    dev = _dev()
    torch.manual_seed(12)
    x = (torch.randn(rows, cols, device=dev) * 3).bfloat16()
    scale = torch.tensor([7.5], device=dev)
    amax = torch.zeros(1, device=dev)
    y = torch.empty(rows, cols, device=dev, dtype=torch.uint8)
    yT = torch.empty(cols, rows, device=dev, dtype=torch.uint8)
    L.call("b200_fp8_quantize", x.data_ptr(), y.data_ptr(), yT.data_ptr(), rows, cols, cols, scale.data_ptr(),
           amax.data_ptr(), int(e5m2), L.stream_ptr())
    dt, lim = (torch.float8_e5m2, 57344.0) if e5m2 else (torch.float8_e4m3fn, 448.0)
    ref = (x.float() * 7.5).clamp(-lim, lim).to(dt).view(torch.uint8)
    assert torch.equal(y, ref), f"fp8 cast: {(y != ref).sum().item()} mismatching bytes"
    assert torch.equal(yT, ref.t().contiguous()), "fp8 transposed cast differs"
    assert float(amax) == float(x.float().abs().max()), (float(amax), float(x.float().abs().max()))
    return {"bytes": rows * cols}


def check_gemm_fp8(M=512, N=768, K=1024, a_e5m2=False, b_e5m2=False, out_mode=0):
    dev = _dev()
    torch.manual_seed(13)
    ta = torch.float8_e5m2 if a_e5m2 else torch.float8_e4m3fn
    tb = torch.float8_e5m2 if b_e5m2 else torch.float8_e4m3fn
    A = torch.randn(M, K, device=dev).to(ta)
    B = torch.randn(N, K, device=dev).to(tb)
    sa, sb = torch.tensor([0.5], device=dev), torch.tensor([0.25], device=dev)
    ref = (A.float() @ B.float().t()) * 0.125
    if out_mode == 2:
        C = torch.randn(M, N, device=dev)
        ref = ref + C
    else:
        C = torch.empty(M, N, device=dev, dtype=torch.bfloat16 if out_mode == 0 else torch.float32)
    L.call("b200_gemm_fp8", A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, K, int(a_e5m2), int(b_e5m2), K, K, N,
           out_mode, sa.data_ptr(), sb.data_ptr(), L.stream_ptr())
    tol = (2e-2 if out_mode == 0 else 1e-3) * math.sqrt(K) * 0.125
    return _close(C, ref, tol, 1e-2 if out_mode == 0 else 1e-4, f"gemm fp8 {M}x{N}x{K} mode{out_mode}")


def attn_ref(q, k, v, scale, causal):
    # q [s,b,np,hn]  k,v [sk,b,g,hn] -> out [s,b,np,hn], lse [b,np,s]
    s, b, np_, hn = q.shape
    sk, g = k.shape[0], k.shape[2]
    rep = np_ // g
    qf = q.float().permute(1, 2, 0, 3)
    kf = k.float().permute(1, 2, 0, 3).repeat_interleave(rep, 1)
    vf = v.float().permute(1, 2, 0, 3).repeat_interleave(rep, 1)
    sc = torch.matmul(qf, kf.transpose(-1, -2)) * scale
    if causal:
        mask = torch.ones(s, sk, device=q.device, dtype=torch.bool).triu(1 + sk - s)
        sc = sc.masked_fill(mask, float("-inf"))
    lse = torch.logsumexp(sc, -1)
    o = torch.matmul(torch.softmax(sc, -1), vf)
    return o.permute(2, 0, 1, 3).contiguous(), lse


ATTN_FWD_ENTRY = "b200_attn_fwd_v2"    # two-tile ping-pong kernel; "b200_attn_fwd" = first generation


def attn_fwd_fused_qkv(qkv, G, qpg, hn, scale, causal=True, entry=None):
    """qkv: [s, b, G, qpg+2, hn] bf16 (the fused projection output). Returns out, lse."""
    s, b = qkv.shape[0], qkv.shape[1]
    np_ = G * qpg
    out = torch.empty(s, b, np_ * hn, device=qkv.device, dtype=torch.bfloat16)
    lse = torch.empty(b, np_, s, device=qkv.device, dtype=torch.float32)
    ss, sb, sg, sj = qkv.stride(0), qkv.stride(1), qkv.stride(2), qkv.stride(3)
    base = qkv.data_ptr()
    esz = 2
    L.call(entry or ATTN_FWD_ENTRY, base, base + qpg * sj * esz, base + (qpg + 1) * sj * esz,
           out.data_ptr(), lse.data_ptr(), s, s, b, G, qpg, hn, ss, sb, sg, sj, ss, sb, sg,
           float(scale), int(causal), L.stream_ptr())
    return out, lse


def check_attn_fwd(S=512, B=2, G=2, QPG=4, HN=128, causal=True, entry=None):
    dev = _dev()
    torch.manual_seed(10)
    qkv = torch.randn(S, B, G, QPG + 2, HN, device=dev).bfloat16()
    scale = 1.0 / math.sqrt(HN)
    out, lse = attn_fwd_fused_qkv(qkv, G, QPG, HN, scale, causal, entry=entry)
    q = qkv[:, :, :, :QPG].reshape(S, B, G * QPG, HN)
    k = qkv[:, :, :, QPG]
    v = qkv[:, :, :, QPG + 1]
    ro, rl = attn_ref(q, k, v, scale, causal)
    st = {"out": _close(out.view(S, B, G * QPG, HN), ro, 2e-2, 2e-2, "attn out"),
          "lse": _close(lse, rl, 2e-3, 1e-3, "attn lse")}
    return st


def check_attn_bwd(S=512, B=2, G=2, QPG=4, HN=128, causal=True, entry=None):
    """Own backward vs autograd through the fp32 restatement of the same attention."""
    from . import functional as Fn
    dev = _dev()
    torch.manual_seed(11)
    qkv = torch.randn(S, B, G, QPG + 2, HN, device=dev).bfloat16()
    dout = (torch.randn(S, B, G * QPG * HN, device=dev) * 0.5).bfloat16()
    scale = 1.0 / math.sqrt(HN)
    out, lse = attn_fwd_fused_qkv(qkv, G, QPG, HN, scale, causal)
    dqkv = torch.full_like(qkv, float("nan"))
    Fn.attn_bwd_fused_qkv(qkv, out, dout, lse, dqkv, scale, causal, entry=entry)
    ref_in = qkv.float().requires_grad_()
    q = ref_in[:, :, :, :QPG].reshape(S, B, G * QPG, HN)
    ro, _ = attn_ref(q, ref_in[:, :, :, QPG], ref_in[:, :, :, QPG + 1], scale, causal)
    ro.backward(dout.float().view(S, B, G * QPG, HN))
    rg = ref_in.grad
    return {"dq": _close(dqkv[:, :, :, :QPG], rg[:, :, :, :QPG], 3e-2, 3e-2, "attn dq"),
            "dk": _close(dqkv[:, :, :, QPG], rg[:, :, :, QPG], 3e-2 * QPG ** 0.5, 3e-2, "attn dk"),
            "dv": _close(dqkv[:, :, :, QPG + 1], rg[:, :, :, QPG + 1], 3e-2 * QPG ** 0.5, 3e-2, "attn dv")}


CHECKS = {
    "rmsnorm": check_rmsnorm,
    "rmsnorm_h1024": lambda: check_rmsnorm(777, 1024),
    "rmsnorm_h5120": lambda: check_rmsnorm(300, 5120),
    "rmsnorm_h8192_cta_path": lambda: check_rmsnorm(256, 8192),
    "rmsnorm_h1000_cta_path": lambda: check_rmsnorm(128, 1000),
    "layernorm": check_layernorm,
    "layernorm_h4096": lambda: check_layernorm(1001, 4096),
    "layernorm_h1024": lambda: check_layernorm(333, 1024),
    "layernorm_h5120": lambda: check_layernorm(64, 5120),
    "rope": check_rope,
    "glu": check_glu,
    "bias_gelu": check_bias_gelu,
    "dropout_add": check_dropout_add,
    "softmax": check_softmax,
    "xent": check_xent,
    "adam": check_adam,
    "casts": check_casts,
    "gemm_nt": lambda: check_gemm("nt"),
    "gemm_nt_1cta": lambda: check_gemm("nt", kernel=1),
    "gemm_nt_1cta_big": lambda: check_gemm("nt", 2048, 2048, 1024, kernel=1),
    "gemm_nn_1cta": lambda: check_gemm("nn", kernel=1),
    "gemm_tn_1cta_acc": lambda: check_gemm("tn", 768, 1024, 2048, out_mode=2, kernel=1),
    "gemm_nt_pair": lambda: check_gemm("nt", kernel=2),
    "gemm_nt_pair_small": lambda: check_gemm("nt", 128, 128, 64, kernel=2),
    "gemm_nt_pair_ragged": lambda: check_gemm("nt", 1000, 968, 520, kernel=2),
    "gemm_nn_pair": lambda: check_gemm("nn", 1024, 1280, 768, kernel=2),
    "gemm_tn_pair": lambda: check_gemm("tn", 1024, 1280, 768, kernel=2),
    "gemm_tt_pair": lambda: check_gemm("tt", 512, 768, 512, kernel=2),
    "gemm_tn_pair_acc": lambda: check_gemm("tn", 768, 1024, 2048, out_mode=2, kernel=2),
    "gemm_nt_pair_f32": lambda: check_gemm("nt", 512, 512, 256, out_mode=1, kernel=2),
    "gemm_nt_big": lambda: check_gemm("nt", 4096, 4096, 4096),
    "gemm_nt_ragged": lambda: check_gemm("nt", 1000, 968, 520),
    "gemm_nn": lambda: check_gemm("nn"),
    "gemm_tn": lambda: check_gemm("tn"),
    "gemm_tt": lambda: check_gemm("tt"),
    "gemm_tn_acc": lambda: check_gemm("tn", 768, 1024, 2048, out_mode=2),
    "gemm_nt_f32": lambda: check_gemm("nt", 256, 128, 256, out_mode=1),
    "fp8_quantize_e4m3": check_fp8_quantize,
    "fp8_quantize_e5m2_ragged": lambda: check_fp8_quantize(208, 1040, True),
    "gemm_fp8": check_gemm_fp8,
    "gemm_fp8_hybrid_big": lambda: check_gemm_fp8(4096, 4096, 4096, True, False),
    "gemm_fp8_ragged_f32": lambda: check_gemm_fp8(1000 // 16 * 16, 968 // 16 * 16, 528, False, False, 1),
    "gemm_fp8_wgrad_acc": lambda: check_gemm_fp8(768, 1024, 2048, True, False, 2),
    "attn_fwd": check_attn_fwd,
    "attn_fwd_mqa_long": lambda: check_attn_fwd(1024, 1, 1, 8, 128, True),
    "attn_fwd_full": lambda: check_attn_fwd(384, 1, 2, 2, 128, False),
    "attn_fwd_ragged": lambda: check_attn_fwd(328, 1, 1, 2, 128, True),
    "attn_fwd_one_tile": lambda: check_attn_fwd(128, 1, 1, 1, 128, True),
    "attn_fwd_long": lambda: check_attn_fwd(2048, 1, 2, 1, 128, True),
    "attn_fwd_gen1": lambda: check_attn_fwd(entry="b200_attn_fwd"),
    "attn_fwd_gen1_mqa_long": lambda: check_attn_fwd(1024, 1, 1, 8, 128, True, entry="b200_attn_fwd"),
    "attn_bwd_small": lambda: check_attn_bwd(128, 1, 1, 1, 128, True),
    "attn_bwd_mha": lambda: check_attn_bwd(512, 2, 4, 1, 128, True),
    "attn_bwd_gqa": check_attn_bwd,
    "attn_bwd_mqa_long": lambda: check_attn_bwd(1024, 1, 1, 8, 128, True),
    "attn_bwd_full": lambda: check_attn_bwd(384, 1, 2, 2, 128, False),
    "attn_bwd_ragged": lambda: check_attn_bwd(328, 1, 1, 2, 128, True),
    "attn_bwd2_small": lambda: check_attn_bwd(128, 1, 1, 1, 128, True, entry="b200_attn_bwd_v2"),
    "attn_bwd2_mha": lambda: check_attn_bwd(512, 2, 4, 1, 128, True, entry="b200_attn_bwd_v2"),
    "attn_bwd2_gqa": lambda: check_attn_bwd(entry="b200_attn_bwd_v2"),
    "attn_bwd2_mqa_long": lambda: check_attn_bwd(1024, 1, 1, 8, 128, True, entry="b200_attn_bwd_v2"),
    "attn_bwd2_full": lambda: check_attn_bwd(384, 1, 2, 2, 128, False, entry="b200_attn_bwd_v2"),
    "attn_bwd2_ragged": lambda: check_attn_bwd(328, 1, 1, 2, 128, True, entry="b200_attn_bwd_v2"),
}


# Model-sized shapes (seconds each, several GB of fp32 restatement): not part of the default list / the
# pytest tier; run by name, e.g. ``python -m epfl_megatron_b200.ops.selfcheck --one gemm_fp8_7b_mlp_up``.
# They exist to narrow down the open FP8 loss gap at 7B.
EXTRA_CHECKS = {
    "gemm_fp8_7b_qkv": lambda: check_gemm_fp8(4096, 12288, 4096),
    "gemm_fp8_7b_mlp_up": lambda: check_gemm_fp8(4096, 22016, 4096),
    "gemm_fp8_7b_mlp_down": lambda: check_gemm_fp8(4096, 4096, 11008),
    "gemm_fp8_7b_dgrad_up": lambda: check_gemm_fp8(4096, 4096, 22016, True, False),
    "gemm_fp8_7b_wgrad_up": lambda: check_gemm_fp8(22016, 4096, 4096, True, False, 2),
    "fp8_quantize_7b_weight": lambda: check_fp8_quantize(22016, 4096),
    "fp8_quantize_7b_act_e5m2": lambda: check_fp8_quantize(4096, 22016, True),
}


def _run_one(name):
    t0 = time.time()
    res = (CHECKS.get(name) or EXTRA_CHECKS[name])()
    torch.cuda.synchronize()
    return {"ok": True, "secs": round(time.time() - t0, 2), "stats": res}


def main(argv):
    if len(argv) >= 2 and argv[0] == "--one":
        try:
            print("RESULT " + json.dumps(_run_one(argv[1])))
        except Exception as e:  # noqa: BLE001
            print("RESULT " + json.dumps({"ok": False, "error": f"{type(e).__name__}: {e}"[:2000]}))
            return 1
        return 0
    names = argv or list(CHECKS)
    summary = {}
    for n in names:
        try:
            r = subprocess.run([sys.executable, "-m", "epfl_megatron_b200.ops.selfcheck", "--one", n],
                               capture_output=True, text=True, timeout=180)
            line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            summary[n] = json.loads(line[-1][7:]) if line else {
                "ok": False, "error": f"exit {r.returncode}: {r.stderr[-1500:]}"}
        except subprocess.TimeoutExpired:
            summary[n] = {"ok": False, "error": "timeout (180 s)"}
        print(n, "OK" if summary[n].get("ok") else "FAIL " + summary[n].get("error", ""), flush=True)
    print("SUMMARY " + json.dumps(summary))
    return 0 if all(v.get("ok") for v in summary.values()) else 1


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
