"""FP8 linear layers with delayed scaling on the library's own 8-bit GEMM.

The reference offers FP8 only by swapping its transformer layers for Transformer Engine's
(megatron/model/transformer.py:931-950, 985-1012; flags ``--fp8_e4m3 --fp8_hybrid --fp8_margin
--fp8_interval --fp8_amax_history_len --fp8_amax_compute_algo --no_fp8_wgrad``,
megatron/arguments.py:369-386). Transformer Engine is not part of this build; the same recipe runs
here on ``b200_gemm_fp8`` (tcgen05 ``kind::f8f6f4``) and ``b200_fp8_quantize``:

* forward   y  = fp8(x) fp8(W)^T                     E4M3 x E4M3
* dgrad     dx = fp8(dy) fp8(W^T)^T                  E5M2 (``--fp8_hybrid``) or E4M3 x E4M3
* wgrad     dW += fp8(dy^T) fp8(x^T)^T  (fp32)       unless ``--no_fp8_wgrad``

Every operand is K-major, so each cast also writes the transposed copy the other GEMM needs (one pass
over the bf16 tensor, which also produces its amax). Scaling is DELAYED as in Transformer Engine: the
cast of step t uses a scale derived from the amax history of earlier steps
(``scale = fp8_max / amax / 2**margin``, amax = most recent or the max over the history), everything
stays on the device: scales and their inverses live in one pool tensor that the GEMM epilogue reads.

The scaling state is NOT written to checkpoints (Transformer Engine stores it as ``_extra_state``): the
first step of a process, fresh or resumed, scales every cast by the tensor's own amax (one extra
reduction per cast) and seeds the history from there, so a resumed run needs no state. FP8 layers
keep NCCL collectives under tensor parallelism (the in-kernel collectives are bf16 kernels).
"""
from __future__ import annotations

import os

import torch

from . import lib as L

E4M3_MAX, E5M2_MAX = 448.0, 57344.0
X, W, G = 0, 1, 2

# B200_FP8_CURRENT_SCALING=1: scale every cast by the tensor's own amax (one extra reduction per cast)
# instead of the delayed scale; useful when statistics move fast (the first steps of a run, tiny
# synthetic benchmarks), not part of the reference's Transformer Engine recipe.
_CURRENT_SCALING = os.environ.get("B200_FP8_CURRENT_SCALING", "0") == "1"
# B200_FP8_EMULATE=1: the casts and GEMMs below are restated with torch's float8 dtypes (any device), so
# the whole recipe -- scale bookkeeping, amax histories, weight-cast caching, the three GEMMs of a layer --
# runs on CPU / Gloo through the real model code (tests/test_fp8_recipe_cpu.py).
_EMULATE = os.environ.get("B200_FP8_EMULATE", "0") == "1"
_CAP = 4096
_POOL = None            # dict of pool tensors, created on first use
_N = 0                  # metas registered
_STEP = 0               # optimizer steps seen (weights are re-cast once per step)
_RECIPE = None


class Recipe:
    def __init__(self, hybrid=False, margin=0, interval=1, history_len=1, algo="most_recent", wgrad=True):
        self.hybrid, self.margin, self.interval = hybrid, margin, max(1, interval)
        self.history_len, self.algo, self.wgrad = max(1, history_len), algo, wgrad


def recipe_from_args(args):
    """None when the run does not ask for FP8."""
    if not (getattr(args, "fp8_e4m3", False) or getattr(args, "fp8_hybrid", False)):
        return None
    return Recipe(hybrid=args.fp8_hybrid, margin=args.fp8_margin, interval=args.fp8_interval,
                  history_len=args.fp8_amax_history_len, algo=args.fp8_amax_compute_algo,
                  wgrad=getattr(args, "fp8_wgrad", True))


def _pool(device, recipe):
    global _POOL, _RECIPE
    if _POOL is None:
        _RECIPE = recipe
        fmax = torch.tensor([E4M3_MAX, E4M3_MAX, E5M2_MAX if recipe.hybrid else E4M3_MAX], device=device)
        _POOL = {"hist": torch.zeros(_CAP, 3, recipe.history_len, device=device),
                 "scale": torch.ones(_CAP, 3, device=device), "scale_inv": torch.ones(_CAP, 3, device=device),
                 "fmax": fmax}
    return _POOL


class Meta:
    """Scaling state of one linear layer: rows of the pool tensors + the cached FP8 weight."""

    def __init__(self, device, recipe: Recipe):
        global _N
        assert _N < _CAP, "too many FP8 linear layers"
        self.recipe = recipe
        self.pool = _pool(device, recipe)
        self.i = _N
        _N += 1
        self.w8 = self.wT8 = None
        self.w_step = -1

    def scale_ptr(self, role):
        return self.pool["scale"][self.i, role].data_ptr()

    def inv_ptr(self, role):
        return self.pool["scale_inv"][self.i, role].data_ptr()

    def amax_ptr(self, role):
        return self.pool["hist"][self.i, role, 0].data_ptr()


def attach(weight, args):
    """Give ``weight`` an FP8 meta when the arguments ask for FP8 and the library can run it."""
    recipe = recipe_from_args(args)
    if recipe is None:
        return None
    if not _EMULATE and (not weight.is_cuda or weight.dtype != torch.bfloat16 or not L.available()):
        return None
    if weight.shape[0] % 16 or weight.shape[1] % 16:
        return None
    weight._b200_fp8 = Meta(weight.device, recipe)
    return weight._b200_fp8


def advance_step():
    """Once per optimizer step: turn the amax values collected during the step into the scales of the
    next one, roll the histories, invalidate the cached FP8 weights. A handful of small device ops for
    the whole model; no host synchronisation."""
    global _STEP
    _STEP += 1
    if _POOL is None or _N == 0:
        return
    r = _RECIPE
    hist, scale, inv = _POOL["hist"][:_N], _POOL["scale"][:_N], _POOL["scale_inv"][:_N]
    if _STEP % r.interval == 0:
        amax = hist.amax(dim=-1) if r.algo == "max" else hist[..., 0]
        new = _POOL["fmax"] / amax / (2.0 ** r.margin)
        ok = (amax > 0) & torch.isfinite(amax) & torch.isfinite(new) & (new > 0)
        scale.copy_(torch.where(ok, new, scale))
        inv.copy_(1.0 / scale)
    if r.history_len > 1:
        hist.copy_(torch.roll(hist, 1, dims=-1))
    hist[..., 0] = 0


def quantize(x2d, meta: Meta, role, transpose, track_amax=True):
    """bf16 [rows, cols] -> (fp8 [rows, cols], fp8 [cols, rows] or None) with the role's current scale."""
    rows, cols = x2d.shape
    e5m2 = int(role == G and meta.recipe.hybrid)
    if _EMULATE:
        return _quantize_emulated(x2d, meta, role, transpose, track_amax, e5m2)
    y = torch.empty(rows, cols, device=x2d.device, dtype=torch.uint8)
    yT = torch.empty(cols, rows, device=x2d.device, dtype=torch.uint8) if transpose else None
    if (_STEP == 0 or _CURRENT_SCALING) and track_amax:
        # No history yet: the very first step scales by the tensor's own amax (two extra small device ops)
        # instead of the neutral scale 1, which would flush bf16-sized gradients to zero in E5M2.
        am = x2d.detach().abs().amax().float()
        sc = meta.pool["fmax"][role] / am / (2.0 ** meta.recipe.margin)
        sc = torch.where(torch.isfinite(sc) & (sc > 0), sc, torch.ones_like(sc))
        meta.pool["scale"][meta.i, role] = sc
        meta.pool["scale_inv"][meta.i, role] = 1.0 / sc
    L.call("b200_fp8_quantize", x2d.data_ptr(), y.data_ptr(), L.ptr(yT), rows, cols, x2d.stride(0),
           meta.scale_ptr(role), meta.amax_ptr(role) if track_amax else None, e5m2, L.stream_ptr())
    return y, yT


def _quantize_emulated(x2d, meta, role, transpose, track_amax, e5m2):
    """torch restatement of b200_fp8_quantize: same scale, same saturation, same amax update."""
    xf = x2d.detach().float()
    am = xf.abs().amax()
    if (_STEP == 0 or _CURRENT_SCALING) and track_amax:
        sc = meta.pool["fmax"][role] / am / (2.0 ** meta.recipe.margin)
        sc = torch.where(torch.isfinite(sc) & (sc > 0), sc, torch.ones_like(sc))
        meta.pool["scale"][meta.i, role] = sc
        meta.pool["scale_inv"][meta.i, role] = 1.0 / sc
    if track_amax:
        h = meta.pool["hist"]
        h[meta.i, role, 0] = torch.maximum(h[meta.i, role, 0], am)
    lim, dt = (E5M2_MAX, torch.float8_e5m2) if e5m2 else (E4M3_MAX, torch.float8_e4m3fn)
    y = (xf * meta.pool["scale"][meta.i, role]).clamp(-lim, lim).to(dt)
    return y, (y.t().contiguous() if transpose else None)


def usable(x2d, weight) -> bool:
    m = getattr(weight, "_b200_fp8", None)
    if _EMULATE:
        return m is not None and x2d.dim() == 2 and x2d.is_floating_point()
    return (m is not None and x2d.is_cuda and x2d.dtype == torch.bfloat16 and x2d.dim() == 2
            and x2d.shape[0] % 16 == 0 and x2d.shape[1] % 16 == 0 and x2d.stride(1) == 1 and x2d.stride(0) % 8 == 0)


def _weights(weight, meta: Meta):
    if meta.w_step != _STEP or meta.w8 is None:
        meta.w8, meta.wT8 = quantize(weight.detach(), meta, W, True)
        meta.w_step = _STEP
    return meta.w8, meta.wT8


def _gemm(a8, b8, out, M, N, K, a_e5m2, b_e5m2, out_mode, meta, role_a, role_b):
    """out (bf16 store: 0, fp32 store: 1, fp32 accumulate: 2) = A B^T / (scale_a * scale_b)."""
    if _EMULATE:
        inv = meta.pool["scale_inv"]
        res = (a8.float() @ b8.float().t()) * (inv[meta.i, role_a] * inv[meta.i, role_b])
        if out_mode == 2:
            out.add_(res.to(out.dtype))
        else:
            out.copy_(res.to(out.dtype))
        return out
    L.call("b200_gemm_fp8", a8.data_ptr(), b8.data_ptr(), out.data_ptr(), M, N, K, int(a_e5m2), int(b_e5m2),
           a8.stride(0), b8.stride(0), out.stride(0), out_mode, meta.inv_ptr(role_a), meta.inv_ptr(role_b),
           L.stream_ptr())
    return out


def forward_gemm(x2d, weight):
    """y [M, N] = x2d [M, K] weight[N, K]^T in FP8 (E4M3 x E4M3)."""
    meta = weight._b200_fp8
    x8, _ = quantize(x2d, meta, X, False)
    w8, _ = _weights(weight, meta)
    M, K = x2d.shape
    N = weight.shape[0]
    out = torch.empty(M, N, device=x2d.device, dtype=x2d.dtype)
    return _gemm(x8, w8, out, M, N, K, 0, 0, 0, meta, X, W)


def dgrad_gemm(dy2d, weight):
    """(dx [M, K], dy^T in FP8 for the weight gradient)."""
    meta = weight._b200_fp8
    dy8, dyT8 = quantize(dy2d, meta, G, meta.recipe.wgrad)
    _, wT8 = _weights(weight, meta)
    M, N = dy2d.shape
    K = weight.shape[1]
    dx = torch.empty(M, K, device=dy2d.device, dtype=dy2d.dtype)
    _gemm(dy8, wT8, dx, M, K, N, meta.recipe.hybrid, 0, 0, meta, G, W)
    return dx, dyT8


def wgrad_gemm(dyT8, x2d, weight, main_grad, accumulate):
    """main_grad [N, K] (fp32) (+)= dy^T x in FP8; x is cast again with the forward's scale (its amax was
    recorded by the forward cast)."""
    meta = weight._b200_fp8
    _, xT8 = quantize(x2d, meta, X, True, track_amax=False)
    N, M = dyT8.shape
    K = x2d.shape[1]
    return _gemm(dyT8, xT8, main_grad, N, K, M, meta.recipe.hybrid, 0, 2 if accumulate else 1, meta, G, X)


def reset():
    """Tests: forget every registered layer."""
    global _POOL, _N, _STEP, _RECIPE
    _POOL, _N, _STEP, _RECIPE = None, 0, 0, None
