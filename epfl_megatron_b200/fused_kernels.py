"""``megatron.fused_kernels`` surface on top of the sm_100a library.

The reference JIT-builds five torch extensions in ``fused_kernels.load(args)``
(megatron/fused_kernels/__init__.py:19-95) and callers reach them by module name:
``scaled_upper_triang_masked_softmax_cuda``, ``scaled_masked_softmax_cuda``, ``scaled_softmax_cuda``,
``fused_mix_prec_layer_norm_cuda`` and ``fused_dense_cuda``. Here there is ONE in-tree library
(``ops/build.py`` -> ``libb200kernels.so``); ``load`` builds it if needed, and the five names below are
adapters with the reference extensions' call signatures over the kernels of that library (or the
PyTorch restatement on CPU), so code and tests written against the extensions keep working:

    from megatron import fused_kernels
    fused_kernels.load(args)
    y = fused_kernels.scaled_masked_softmax_cuda.forward(x, mask, scale)

Unlike the reference's kernels there is no limit on the key length (theirs: 16 < sk <= 4096, 2048 for
the causal one) and no batch-per-block divisibility rule.
"""
from __future__ import annotations

import torch

from .ops import functional as Fn
from .ops import lib as L
from .ops import ref


def load(args=None):
    """Build (if stale) and load the native libraries; a no-op on a machine without nvcc when the
    libraries are already in the tree. ``args`` is accepted for signature parity and not used."""
    from .ops import build
    try:
        build.build_all()
    except Exception:  # noqa: BLE001 - no compiler here: use what is in the tree
        pass
    return L.available()


def _softmax_fwd(x, mask, scale, mode):
    if Fn._k(x) and x.dim() == 4 and x.shape[-1] % 8 == 0:
        return Fn._SoftmaxFn.apply(x, mask, float(scale), mode)
    return ref.scaled_softmax(x, float(scale), mask, mode == 1)


def _softmax_bwd(grad, probs, scale):
    """scale * (g * y - y * sum(g * y)); the reference overwrites ``grad`` in place and returns it."""
    if Fn._k(grad) and grad.shape[-1] % 8 == 0:
        g, y = grad.contiguous(), probs.contiguous()
        dx = torch.empty_like(y)
        L.call("b200_softmax_bwd", g.data_ptr(), y.data_ptr(), dx.data_ptr(), y.numel() // y.shape[-1],
               y.shape[-1], float(scale), L.stream_ptr())
        return dx
    gy = grad.float() * probs.float()
    return (float(scale) * (gy - probs.float() * gy.sum(-1, keepdim=True))).to(grad.dtype)


class scaled_upper_triang_masked_softmax_cuda:      # noqa: N801 - the reference's extension name
    """Causal softmax of ``scale * x`` over [attn_batches, sq, sk] (scaled_upper_triang_masked_softmax.h)."""

    @staticmethod
    def forward(inputs, scale):
        x4 = inputs.unsqueeze(0)
        return _softmax_fwd(x4, None, scale, 1).squeeze(0)

    @staticmethod
    def backward(output_grads, softmax_results, scale):
        return _softmax_bwd(output_grads, softmax_results, scale)


class scaled_masked_softmax_cuda:                   # noqa: N801
    """Softmax of ``scale * x`` over [b, np, sq, sk] with a uint8 / bool mask [b or 1, 1, sq, sk]
    (1 = masked; a fully masked row gives zeros), scaled_masked_softmax.h."""

    @staticmethod
    def forward(inputs, mask, scale):
        return _softmax_fwd(inputs, mask.bool() if mask is not None else None, scale, 2 if mask is not None else 0)

    @staticmethod
    def backward(output_grads, softmax_results, scale):
        return _softmax_bwd(output_grads, softmax_results, scale)

    @staticmethod
    def get_batch_per_block(query_seq_len, key_seq_len, batches, attn_heads):
        """The reference sizes its thread blocks with this; the kernels here have no such constraint.
        Returns the reference's value for the common configurations so divisibility checks written
        against it keep passing (scaled_masked_softmax.h:get_batch_per_block)."""
        log2 = max(int(key_seq_len - 1).bit_length(), 0)
        pow2 = 1 << log2
        warp = min(pow2, 32)
        batches_per_warp = 2 if pow2 <= 128 else 1
        return (128 // warp) * batches_per_warp


class scaled_softmax_cuda:                          # noqa: N801
    """Unmasked softmax of ``scale * x`` over [b, np, sq, sk] (scaled_softmax.cpp)."""

    @staticmethod
    def forward(inputs, scale):
        return _softmax_fwd(inputs, None, scale, 0)

    @staticmethod
    def backward(output_grads, softmax_results, scale):
        return _softmax_bwd(output_grads, softmax_results, scale)


class fused_mix_prec_layer_norm_cuda:               # noqa: N801
    """Mixed-precision LayerNorm (layer_norm_cuda.cpp): output in the parameter dtype, fp32 mean and
    inverse standard deviation per row."""

    @staticmethod
    def forward_affine(inputs, normalized_shape, weight, bias, eps):
        H = int(torch.Size(normalized_shape).numel())
        x2 = inputs.reshape(-1, H).contiguous()
        M = x2.shape[0]
        if Fn._k(x2) and weight.dtype == torch.bfloat16:
            y = torch.empty_like(x2)
            mean = torch.empty(M, device=x2.device, dtype=torch.float32)
            rstd = torch.empty(M, device=x2.device, dtype=torch.float32)
            L.call("b200_layernorm_fwd", x2.data_ptr(), weight.data_ptr(), L.ptr(bias), y.data_ptr(),
                   mean.data_ptr(), rstd.data_ptr(), M, H, float(eps), L.stream_ptr())
            return y.view_as(inputs), mean, rstd
        xf = x2.float()
        mean = xf.mean(-1)
        rstd = torch.rsqrt(xf.var(-1, unbiased=False) + eps)
        y = (xf - mean[:, None]) * rstd[:, None] * weight.float()
        if bias is not None:
            y = y + bias.float()
        return y.to(weight.dtype).view_as(inputs), mean, rstd

    @staticmethod
    def backward_affine(grad_output, mean, invvar, inputs, normalized_shape, weight, bias, eps):
        H = int(torch.Size(normalized_shape).numel())
        g2 = grad_output.reshape(-1, H).contiguous()
        x2 = inputs.reshape(-1, H).contiguous()
        M = x2.shape[0]
        if Fn._k(x2) and Fn._k(g2) and weight.dtype == torch.bfloat16:
            dx = torch.empty_like(x2)
            dg = torch.empty(H, device=x2.device, dtype=torch.float32)
            db = torch.empty(H, device=x2.device, dtype=torch.float32)
            scratch = torch.empty(L.lib().b200_norm_bwd_grid(M), 2, H, device=x2.device, dtype=torch.float32)
            L.call("b200_layernorm_bwd", g2.data_ptr(), x2.data_ptr(), weight.data_ptr(), mean.data_ptr(),
                   invvar.data_ptr(), dx.data_ptr(), dg.data_ptr(), db.data_ptr(), scratch.data_ptr(), M, H, 0,
                   L.stream_ptr(), launches=2)
            return dx.view_as(inputs), dg.to(weight.dtype), db.to(weight.dtype)
        xh = (x2.float() - mean[:, None]) * invvar[:, None]
        gf = g2.float()
        gw = gf * weight.float()
        dx = invvar[:, None] * (gw - gw.mean(-1, keepdim=True) - xh * (gw * xh).mean(-1, keepdim=True))
        return dx.to(inputs.dtype).view_as(inputs), (gf * xh).sum(0).to(weight.dtype), gf.sum(0).to(weight.dtype)


class fused_dense_cuda:                             # noqa: N801
    """``main_grad += grad_output^T @ input`` in one GEMM (fused_weight_gradient_dense.cu:128-152)."""

    @staticmethod
    def wgrad_gemm_accum_fp32(input, grad_output, main_grad):      # noqa: A002 - reference argument name
        x2 = input.reshape(-1, input.shape[-1])
        g2 = grad_output.reshape(-1, grad_output.shape[-1])
        if x2.is_cuda and Fn.wgrad_accumulate(g2.contiguous(), x2.contiguous(), main_grad, accumulate=True):
            return
        main_grad.addmm_(g2.t().float(), x2.float())


# the reference's apex-era alias of the same extension
fused_weight_gradient_mlp_cuda = fused_dense_cuda
