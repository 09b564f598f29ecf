"""Autograd-aware fused ops. bf16 CUDA tensors run the sm_100a kernels of
``libb200kernels.so``; everything else (CPU, fp32, fp16) runs the PyTorch restatement in
``ops.ref``. ``set_kernel_mode('off')`` forces the restatement everywhere (A/B testing,
the ``--b200_kernels`` flag); it is a test switch, not a second backend.
"""
from __future__ import annotations

import math
import os

import torch

from . import lib as L
from . import ref

_MODE = "auto"       # auto | on | off
_GEMM_MODE = "auto"  # auto | own | cublas


def set_kernel_mode(mode: str):
    global _MODE
    assert mode in ("auto", "on", "off")
    _MODE = mode


def set_gemm_mode(mode: str):
    global _GEMM_MODE
    assert mode in ("auto", "own", "cublas")
    _GEMM_MODE = mode


_HOST_FUSED_WGRAD = os.environ.get("B200_HOST_FUSED_WGRAD", "0") == "1"


def kernels_enabled() -> bool:
    return _MODE != "off" and L.available()


def _k(t: torch.Tensor) -> bool:
    """Should tensor ``t`` take the CUDA-kernel path?"""
    return t.is_cuda and t.dtype == torch.bfloat16 and _MODE != "off" and L.available()


def _notify_grad_ready(param):
    cb = getattr(param, "_b200_grad_ready", None)
    if cb is not None:
        cb()


# ----------------------------------------------------------------------------- norms
class _RMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, eps, residual):
        shape = x.shape
        H = shape[-1]
        x2 = x.reshape(-1, H)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        y = torch.empty_like(x2)
        rstd = torch.empty(M, device=x.device, dtype=torch.float32)
        h = None
        if residual is not None:
            r2 = residual.reshape(-1, H)
            if not r2.is_contiguous():
                r2 = r2.contiguous()
            h = torch.empty_like(x2)
            ok = L.try_call("b200_rmsnorm_fwd", x2.data_ptr(), r2.data_ptr(), weight.data_ptr(),
                            y.data_ptr(), h.data_ptr(), rstd.data_ptr(), M, H, eps, L.stream_ptr())
        else:
            ok = L.try_call("b200_rmsnorm_fwd", x2.data_ptr(), None, weight.data_ptr(),
                            y.data_ptr(), None, rstd.data_ptr(), M, H, eps, L.stream_ptr())
        if not ok:
            raise L.KernelError(f"rmsnorm: unsupported hidden size {H}")
        ctx.save_for_backward(h if h is not None else x2, weight, rstd)
        ctx.has_res = residual is not None
        ctx.shape = shape
        ctx.weight_ref = weight
        if residual is not None:
            return y.view(shape), h.view(shape)
        return y.view(shape)

    @staticmethod
    def backward(ctx, dy, dh=None):
        xin, weight, rstd = ctx.saved_tensors
        M, H = xin.shape
        dy2 = dy.reshape(M, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dres = None
        if ctx.has_res and dh is not None:
            dres = dh.reshape(M, H)
            if not dres.is_contiguous():
                dres = dres.contiguous()
        grid = L.lib().b200_norm_bwd_grid(M)
        scratch = torch.empty(grid, H, device=dy.device, dtype=torch.float32)
        dx = torch.empty_like(xin)
        w = ctx.weight_ref
        main = getattr(w, "main_grad", None)
        fuse = main is not None and main.dtype == torch.float32 and getattr(w, "_b200_fuse_wgrad", False)
        dw = main if fuse else torch.empty(H, device=dy.device, dtype=torch.float32)
        # the grad buffer is only "logically" zeroed at the start of a step: the first gradient
        # of the step must overwrite the slice, later ones (more micro-batches) accumulate
        first = fuse and getattr(w, "_b200_main_grad_empty", False)
        L.call("b200_rmsnorm_bwd", dy2.data_ptr(), xin.data_ptr(), weight.data_ptr(),
               rstd.data_ptr(), L.ptr(dres), dx.data_ptr(), dw.data_ptr(), scratch.data_ptr(), M, H,
               1 if (fuse and not first) else 0, L.stream_ptr(), launches=2)
        dxv = dx.view(ctx.shape)
        if fuse:
            w._b200_main_grad_empty = False
            _notify_grad_ready(w)
            gw = None
        else:
            gw = dw.to(weight.dtype)
        if ctx.has_res:
            return dxv, gw, None, dxv
        return dxv, gw, None, None


def rmsnorm(x, weight, eps=1e-5, residual=None):
    """y = rmsnorm(x [+ residual]) * weight. With ``residual`` returns (y, x + residual)."""
    if _k(x) and x.shape[-1] % 8 == 0 and x.shape[-1] <= 8192:
        return _RMSNormFn.apply(x, weight, eps, residual)
    if residual is not None:
        h = x + residual
        return ref.rmsnorm(h, weight, eps), h
    return ref.rmsnorm(x, weight, eps)


class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        shape = x.shape
        H = shape[-1]
        x2 = x.reshape(-1, H)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(M, device=x.device, dtype=torch.float32)
        rstd = torch.empty(M, device=x.device, dtype=torch.float32)
        L.call("b200_layernorm_fwd", x2.data_ptr(), weight.data_ptr(), L.ptr(bias), y.data_ptr(),
               mean.data_ptr(), rstd.data_ptr(), M, H, eps, L.stream_ptr())
        ctx.save_for_backward(x2, weight, mean, rstd)
        ctx.shape = shape
        ctx.has_bias = bias is not None
        ctx.bias_dtype = bias.dtype if bias is not None else None
        return y.view(shape)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, mean, rstd = ctx.saved_tensors
        M, H = x2.shape
        dy2 = dy.reshape(M, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        grid = L.lib().b200_norm_bwd_grid(M)
        scratch = torch.empty(grid, 2, H, device=dy.device, dtype=torch.float32)
        dx = torch.empty_like(x2)
        dg = torch.empty(H, device=dy.device, dtype=torch.float32)
        db = torch.empty(H, device=dy.device, dtype=torch.float32) if ctx.has_bias else None
        L.call("b200_layernorm_bwd", dy2.data_ptr(), x2.data_ptr(), weight.data_ptr(),
               mean.data_ptr(), rstd.data_ptr(), dx.data_ptr(), dg.data_ptr(), L.ptr(db),
               scratch.data_ptr(), M, H, 0, L.stream_ptr(), launches=2)
        return (dx.view(ctx.shape), dg.to(weight.dtype),
                db.to(ctx.bias_dtype) if db is not None else None, None)


def layernorm(x, weight, bias, eps=1e-5):
    if _k(x) and x.shape[-1] % 8 == 0 and x.shape[-1] <= 8192:
        return _LayerNormFn.apply(x, weight, bias, eps)
    return ref.layernorm(x, weight, bias, eps)


# ----------------------------------------------------------------------------- RoPE
def _fresh_alias(t):
    """A new tensor object over the same memory that autograd does not track as a view."""
    out = torch.empty(0, dtype=t.dtype, device=t.device)
    out.set_(t.untyped_storage(), t.storage_offset(), t.shape, t.stride())
    return out


class _RopeQKVFn(torch.autograd.Function):
    """Rotation of the q and k slots of a fused [s,b,g,nslot,hn] tensor, physically in place:
    the projection output is overwritten (its producer never re-reads it) and handed back
    as a fresh tensor so autograd sees an ordinary out-of-place op."""

    @staticmethod
    def forward(ctx, qkv, table, position_ids, nrot, pos_offset):
        s, b, g, nslot, hn = qkv.shape
        assert qkv.is_contiguous()
        L.call("b200_rope", qkv.data_ptr(), table.data_ptr(), L.ptr(position_ids), s, b, g, nslot,
               nrot, hn, pos_offset, table.shape[0], 0, L.stream_ptr())
        ctx.save_for_backward(table, position_ids)
        ctx.meta = (nrot, pos_offset)
        return _fresh_alias(qkv)

    @staticmethod
    def backward(ctx, dqkv):
        table, position_ids = ctx.saved_tensors
        nrot, pos_offset = ctx.meta
        if not dqkv.is_contiguous():
            dqkv = dqkv.contiguous()
        s, b, g, nslot, hn = dqkv.shape
        # the incoming gradient is produced by the attention backward for this op only
        L.call("b200_rope", dqkv.data_ptr(), table.data_ptr(), L.ptr(position_ids), s, b, g, nslot,
               nrot, hn, pos_offset, table.shape[0], 1, L.stream_ptr())
        return _fresh_alias(dqkv), None, None, None, None


def rope_fused_qkv(qkv, table, position_ids=None, pos_offset=0):
    """qkv: [s, b, g, q_per_g + 2, hn]; rotates all q slots and the k slot, leaves v."""
    nslot = qkv.shape[3]
    if _k(qkv) and qkv.is_contiguous() and qkv.shape[-1] % 8 == 0:
        pid = position_ids.contiguous() if position_ids is not None else None
        return _RopeQKVFn.apply(qkv, table, pid, nslot - 1, pos_offset)
    rot = ref.apply_rope(qkv[:, :, :, :nslot - 1], table, position_ids, pos_offset)
    return torch.cat([rot, qkv[:, :, :, nslot - 1:]], dim=3)


def rope(x, table, position_ids=None, pos_offset=0):
    """Generic RoPE on [s, b, heads, hn] (used by the inference / KV-cache path)."""
    return ref.apply_rope(x, table, position_ids, pos_offset)


# ----------------------------------------------------------------------------- activations
class _GluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kind):
        F2 = x.shape[-1]
        Fh = F2 // 2
        x2 = x.reshape(-1, F2)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        y = torch.empty(M, Fh, device=x.device, dtype=x.dtype)
        L.call("b200_glu_fwd", x2.data_ptr(), y.data_ptr(), M, Fh, kind, L.stream_ptr())
        ctx.save_for_backward(x2)
        ctx.kind = kind
        ctx.shape = x.shape
        return y.view(*x.shape[:-1], Fh)

    @staticmethod
    def backward(ctx, dy):
        (x2,) = ctx.saved_tensors
        M, F2 = x2.shape
        dy2 = dy.reshape(M, F2 // 2)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = torch.empty_like(x2)
        L.call("b200_glu_bwd", dy2.data_ptr(), x2.data_ptr(), dx.data_ptr(), M, F2 // 2, ctx.kind,
               L.stream_ptr())
        return dx.view(ctx.shape), None


def glu(x, kind: str):
    if _k(x) and (x.shape[-1] // 2) % 8 == 0:
        return _GluFn.apply(x, ref.GLU_KIND[kind])
    return ref.glu(x, kind)


class _BiasGeluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bias, tanh_approx):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        y = torch.empty_like(x2)
        L.call("b200_bias_gelu_fwd", x2.data_ptr(), L.ptr(bias), y.data_ptr(), x2.shape[0], H,
               int(tanh_approx), L.stream_ptr())
        ctx.save_for_backward(x2, bias)
        ctx.tanh = int(tanh_approx)
        ctx.shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, bias = ctx.saved_tensors
        M, H = x2.shape
        dy2 = dy.reshape(M, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = torch.empty_like(x2)
        L.call("b200_bias_gelu_bwd", dy2.data_ptr(), x2.data_ptr(), L.ptr(bias), dx.data_ptr(), M,
               H, ctx.tanh, L.stream_ptr())
        db = dx.float().sum(0).to(bias.dtype) if bias is not None else None
        return dx.view(ctx.shape), db, None


def bias_gelu(x, bias=None, tanh_approx=True):
    if _k(x) and x.shape[-1] % 8 == 0:
        return _BiasGeluFn.apply(x, bias, tanh_approx)
    return ref.bias_gelu(x, bias, tanh_approx)


class _DropoutAddFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bias, residual, p, seed, offset):
        H = x.shape[-1]
        x2 = x.reshape(-1, H).contiguous()
        r2 = residual.reshape(-1, H).contiguous()
        out = torch.empty_like(x2)
        L.call("b200_dropout_add_fwd", x2.data_ptr(), L.ptr(bias), r2.data_ptr(), out.data_ptr(),
               x2.shape[0], H, p, seed, offset, L.stream_ptr())
        ctx.meta = (p, seed, offset, x.shape, bias is not None, bias.dtype if bias is not None else None)
        return out.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        p, seed, offset, shape, has_bias, bdt = ctx.meta
        H = shape[-1]
        dy2 = dy.reshape(-1, H).contiguous()
        if p == 0.0:
            dx = dy2
        else:
            dx = torch.empty_like(dy2)
            L.call("b200_dropout_bwd", dy2.data_ptr(), dx.data_ptr(), dy2.shape[0], H, p, seed,
                   offset, L.stream_ptr())
        db = dx.float().sum(0).to(bdt) if has_bias else None
        return dx.view(shape), db, dy, None, None, None


def bias_dropout_add(x, bias, residual, p: float, training: bool):
    """residual + dropout(x + bias) (model/transformer.py:532-578). The Philox seed is drawn
    from the active torch CUDA generator, so the TP RNG tracker semantics carry over."""
    if not training:
        p = 0.0
    if p == 0.0 and bias is None:
        return residual + x
    if _k(x) and x.shape[-1] % 8 == 0 and x.shape == residual.shape:
        seed, off = _philox_state(x.device) if p > 0 else (0, 0)
        return _DropoutAddFn.apply(x, bias, residual, float(p), seed, off)
    t = x if bias is None else x + bias
    return residual + torch.nn.functional.dropout(t, p=p, training=training and p > 0)


def _philox_state(device):
    """(seed, offset) of the active CUDA generator; consumes 4 counters, no device sync.
    Under ``get_cuda_rng_tracker().fork()`` this is the tensor-parallel RNG stream."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    gen = torch.cuda.default_generators[idx]
    off = gen.get_offset()
    gen.set_offset(off + 4)
    return int(gen.initial_seed()) & ((1 << 63) - 1), int(off)


# ----------------------------------------------------------------------------- softmax
class _SoftmaxFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, mask, scale, mode):
        b, np_, sq, sk = x.shape
        xc = x.contiguous()
        y = torch.empty_like(xc)
        m8 = None
        mb = 1
        if mode == 2:
            m8 = mask.to(torch.uint8).contiguous()
            mb = m8.shape[0]
        L.call("b200_softmax_fwd", xc.data_ptr(), L.ptr(m8), y.data_ptr(), b, np_, sq, sk, mb,
               float(scale), mode, L.stream_ptr())
        ctx.save_for_backward(y)
        ctx.scale = float(scale)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dyc = dy.contiguous()
        dx = torch.empty_like(y)
        L.call("b200_softmax_bwd", dyc.data_ptr(), y.data_ptr(), dx.data_ptr(),
               y.numel() // y.shape[-1], y.shape[-1], ctx.scale, L.stream_ptr())
        return dx, None, None, None


def scaled_softmax(x, scale=1.0, mask=None, causal=False):
    """x [b, np, sq, sk]. mask: bool [b or 1, 1, sq, sk], True = masked out."""
    if _k(x) and x.dim() == 4 and x.shape[-1] % 8 == 0 and not (causal and mask is not None):
        mode = 1 if causal else (2 if mask is not None else 0)
        return _SoftmaxFn.apply(x, mask, scale, mode)
    return ref.scaled_softmax(x, scale, mask, causal)


# ----------------------------------------------------------------------------- attention
def _fa2_backward(dout, q, k, v, out, lse, dq, dk, dv, scale, causal):
    from flash_attn.flash_attn_interface import _flash_attn_backward
    _flash_attn_backward(dout, q, k, v, out, lse, dq, dk, dv, 0.0, scale, causal, -1, -1, 0.0,
                         None, False, None)


class _AttnFusedQKVFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, scale, causal):
        s, b, g, nslot, hn = qkv.shape
        qpg = nslot - 2
        np_ = g * qpg
        out = torch.empty(s, b, np_ * hn, device=qkv.device, dtype=qkv.dtype)
        lse = torch.empty(b, np_, s, device=qkv.device, dtype=torch.float32)
        ss, sb, sg, sj = qkv.stride(0), qkv.stride(1), qkv.stride(2), qkv.stride(3)
        base = qkv.data_ptr()
        L.call(_ATTN_FWD_ENTRY, base, base + qpg * sj * 2, base + (qpg + 1) * sj * 2,
               out.data_ptr(), lse.data_ptr(), s, s, b, g, qpg, hn, ss, sb, sg, sj, ss, sb, sg,
               float(scale), int(causal), L.stream_ptr())
        ctx.save_for_backward(qkv, out, lse)
        ctx.meta = (float(scale), bool(causal))
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse = ctx.saved_tensors
        scale, causal = ctx.meta
        s, b, g, nslot, hn = qkv.shape
        qpg = nslot - 2
        np_ = g * qpg
        dqkv = torch.empty_like(qkv)
        if not dout.is_contiguous():
            dout = dout.contiguous()
        if _ATTN_BWD == "own":
            attn_bwd_fused_qkv(qkv, out, dout, lse, dqkv, scale, causal)
            return dqkv, None, None
        # library backward (FlashAttention-2), kept for A/B runs: --b200_attn_bwd fa2
        # [b, s, heads, hn] strided views (no copies when qpg == 1)
        k = qkv[:, :, :, qpg].permute(1, 0, 2, 3)
        v = qkv[:, :, :, qpg + 1].permute(1, 0, 2, 3)
        dk = dqkv[:, :, :, qpg].permute(1, 0, 2, 3)
        dv = dqkv[:, :, :, qpg + 1].permute(1, 0, 2, 3)
        o4 = out.view(s, b, np_, hn).permute(1, 0, 2, 3)
        do4 = dout.reshape(s, b, np_, hn).permute(1, 0, 2, 3)
        if qpg == 1:
            q = qkv[:, :, :, 0].permute(1, 0, 2, 3)
            dq = dqkv[:, :, :, 0].permute(1, 0, 2, 3)
            _fa2_backward(do4, q, k, v, o4, lse, dq, dk, dv, scale, causal)
        else:
            q = qkv[:, :, :, :qpg].reshape(s, b, np_, hn).permute(1, 0, 2, 3)
            dq = torch.empty(b, s, np_, hn, device=qkv.device, dtype=qkv.dtype)
            _fa2_backward(do4, q, k, v, o4, lse, dq, dk, dv, scale, causal)
            dqkv[:, :, :, :qpg].copy_(dq.permute(1, 0, 2, 3).reshape(s, b, g, qpg, hn))
        return dqkv, None, None


_ATTN_BWD = "own"   # own | fa2
_ATTN_FWD_ENTRY = "b200_attn_fwd_v2"   # two query tiles per CTA; "b200_attn_fwd" = first generation
# "b200_attn_bwd" = first generation (one S / dP buffer, 128-row streamed tiles);
# "b200_attn_bwd_v2" = 64-row streamed blocks with S / dP double-buffered in TMEM
_ATTN_BWD_ENTRY = os.environ.get("B200_ATTN_BWD_ENTRY", "b200_attn_bwd")


def set_attn_bwd_mode(mode: str):
    global _ATTN_BWD
    assert mode in ("own", "fa2")
    _ATTN_BWD = mode


def attn_bwd_fused_qkv(qkv, out, dout, lse, dqkv, scale, causal=True, entry=None):
    """dqkv (same layout as qkv: [s, b, g, q_per_g+2, hn]) <- attention backward, own tcgen05
    kernels (csrc/attn_bwd_tcgen05.cu). out/dout: [s, b, np*hn] contiguous, lse [b, np, s]."""
    s, b, g, nslot, hn = qkv.shape
    qpg = nslot - 2
    n = L.lib().b200_attn_bwd_scratch_elems(s, s, b, g, qpg)
    scratch = torch.empty(n, device=qkv.device, dtype=torch.float32)
    ss, sb, sg, sj = qkv.stride(0), qkv.stride(1), qkv.stride(2), qkv.stride(3)
    ds, db, dg, dj = dqkv.stride(0), dqkv.stride(1), dqkv.stride(2), dqkv.stride(3)
    base, dbase = qkv.data_ptr(), dqkv.data_ptr()
    L.call(entry or _ATTN_BWD_ENTRY, base, base + qpg * sj * 2, base + (qpg + 1) * sj * 2, out.data_ptr(),
           dout.data_ptr(), lse.data_ptr(), dbase, dbase + qpg * dj * 2, dbase + (qpg + 1) * dj * 2,
           scratch.data_ptr(), s, s, b, g, qpg, hn, ss, sb, sg, sj, ss, sb, sg, ds, db, dg, dj,
           ds, db, dg, float(scale), int(causal), L.stream_ptr(), launches=3)
    return dqkv


def attention_fused_qkv(qkv, scale=None, causal=True, dropout_p=0.0, attn_mask=None):
    """Self-attention straight from the fused projection output.

    qkv: [s, b, g, q_per_g + 2, hn] with slots (q_0 .. q_{n/g-1}, k, v) per KV group --
    the checkpoint layout of ``query_key_value``. Returns
    [s, b, np*hn]. GQA/MQA never expands K/V.
    """
    s, b, g, nslot, hn = qkv.shape
    qpg = nslot - 2
    if scale is None:
        scale = 1.0 / math.sqrt(hn)
    if _k(qkv) and hn == 128 and dropout_p == 0.0 and attn_mask is None:
        return _AttnFusedQKVFn.apply(qkv, scale, causal)
    q = qkv[:, :, :, :qpg].reshape(s, b, g * qpg, hn)
    return ref.attention(q, qkv[:, :, :, qpg], qkv[:, :, :, qpg + 1], scale, causal, attn_mask,
                         dropout_p)


# ----------------------------------------------------------------------------- GEMM
def _own_gemm_ok(*ts):
    return all(t.is_cuda and t.dtype == torch.bfloat16 and t.stride(-1) == 1
               and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0 for t in ts) and L.available() \
        and _MODE != "off"


def _use_own(M, N, K, force_own):
    """`auto` takes the tcgen05 CTA-pair kernel for every GEMM with at least one full 256 x 256
    tile; tiny problems stay on the library (launch-bound either way)."""
    if force_own or _GEMM_MODE == "own":
        return True
    return _GEMM_MODE == "auto" and M >= 256 and N >= 256 and K >= 64


def gemm_nt(x2d, w, force_own=False):
    """x2d [M,K] @ w[N,K]^T -> [M,N]."""
    if _use_own(x2d.shape[0], w.shape[0], x2d.shape[1], force_own) and _own_gemm_ok(x2d, w):
        M, K = x2d.shape
        N = w.shape[0]
        out = torch.empty(M, N, device=x2d.device, dtype=torch.bfloat16)
        L.call("b200_gemm_bf16", x2d.data_ptr(), w.data_ptr(), out.data_ptr(), M, N, K, 0, 0,
               x2d.stride(0), w.stride(0), N, 0, 0, L.stream_ptr())
        return out
    return torch.matmul(x2d, w.t())


def gemm_nn(dy2d, w, force_own=False):
    """dy2d [M,N] @ w[N,K] -> [M,K] (dgrad)."""
    if _use_own(dy2d.shape[0], w.shape[1], dy2d.shape[1], force_own) and _own_gemm_ok(dy2d, w):
        M, N = dy2d.shape
        K = w.shape[1]
        out = torch.empty(M, K, device=dy2d.device, dtype=torch.bfloat16)
        L.call("b200_gemm_bf16", dy2d.data_ptr(), w.data_ptr(), out.data_ptr(), M, K, N, 0, 1,
               dy2d.stride(0), w.stride(0), K, 0, 0, L.stream_ptr())
        return out
    return torch.matmul(dy2d, w)


def wgrad_accumulate(dy2d, x2d, main_grad, accumulate=True):
    """main_grad[N,K] (fp32) (+)= dy2d[M,N]^T @ x2d[M,K] in ONE kernel (N5 equivalent).
    Returns False when the fused kernel cannot be used (caller falls back).
    B200_HOST_FUSED_WGRAD=1 restates the epilogue with torch on tensors the kernel does not take
    (CPU / fp32), so the arrival order it creates -- gradient written outside autograd, then
    ``_b200_grad_ready`` -- can be exercised on Gloo."""
    if _HOST_FUSED_WGRAD and not _k(dy2d) and main_grad.dtype == torch.float32:
        if accumulate:
            main_grad.addmm_(dy2d.t().float(), x2d.float())
        else:
            torch.mm(dy2d.t().float(), x2d.float(), out=main_grad)
        return True
    if _GEMM_MODE == "cublas" or not _own_gemm_ok(dy2d, x2d) or main_grad.dtype != torch.float32 \
            or not main_grad.is_contiguous():
        return False
    M, N = dy2d.shape
    K = x2d.shape[1]
    L.call("b200_gemm_bf16", dy2d.data_ptr(), x2d.data_ptr(), main_grad.data_ptr(), N, K, M, 1, 1,
           dy2d.stride(0), x2d.stride(0), K, 2 if accumulate else 1, 0, L.stream_ptr())
    return True
