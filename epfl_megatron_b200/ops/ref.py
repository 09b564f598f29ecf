"""Pure-PyTorch restatement of every fused op.

Serves two purposes: (a) the CPU / Gloo execution path (BASELINE config 1) and any dtype
the CUDA kernels do not cover, (b) the numerical oracle the kernel tests compare
against. Semantics follow the reference modules cited in each docstring.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def _norm_out_dtype(x, weight):
    """An fp32 residual stream (--fp32_residual_connection) normalised with half-precision weights feeds
    half-precision GEMMs: the output takes the weight's dtype, as the reference's mixed-dtype layer norm
    does (model/fused_layer_norm.py:29-41, 106-122). Otherwise the input's dtype."""
    return weight.dtype if x.dtype == torch.float32 and weight.dtype != torch.float32 else x.dtype


def rmsnorm(x, weight, eps):
    """fp32 x*rsqrt(mean(x^2)+eps), cast back, THEN scale (model/fused_layer_norm.py:125-139)."""
    xf = x.float()
    y = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    out = y.type_as(x) * weight
    return out.to(weight.dtype) if _norm_out_dtype(x, weight) != x.dtype else out


def layernorm(x, weight, bias, eps):
    return F.layer_norm(x.float(), (x.shape[-1],), weight.float(),
                        None if bias is None else bias.float(), eps).to(_norm_out_dtype(x, weight))


def rope_table(head_dim: int, max_pos: int, theta: float = 10000.0, scaling_factor: float = 1.0,
               device=None):
    """(cos, sin) table [max_pos, head_dim/2, 2] fp32; positions divided by the linear
    scaling factor (model/positional_embeddings.py:7-21)."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, device=device)[: head_dim // 2].float()
                           / head_dim))
    t = torch.arange(max_pos, device=device).float() / scaling_factor
    ang = torch.outer(t, inv)
    return torch.stack([ang.cos(), ang.sin()], dim=-1).contiguous()


def apply_rope(x, table, position_ids=None, pos_offset: int = 0, inverse: bool = False):
    """Rotate adjacent pairs (Meta 'interleaved' convention). x: [s, b, ..., hn];
    position_ids: [b, s] or None (model/positional_embeddings.py:24-51)."""
    s, b = x.shape[0], x.shape[1]
    if position_ids is None:
        f = table[pos_offset:pos_offset + s][:, None]            # [s,1,hn/2,2]
    else:
        f = table[position_ids.clamp(max=table.shape[0] - 1)].transpose(0, 1)  # [s,b,hn/2,2]
    extra = x.dim() - 3
    f = f.reshape(f.shape[0], f.shape[1], *([1] * extra), f.shape[-2], 2)
    xf = x.float().reshape(*x.shape[:-1], -1, 2)
    cos, sin = f[..., 0], f[..., 1]
    if inverse:
        sin = -sin
    out = torch.stack([xf[..., 0] * cos - xf[..., 1] * sin,
                       xf[..., 0] * sin + xf[..., 1] * cos], dim=-1)
    return out.flatten(-2).type_as(x)


GLU_ACTS = {
    "liglu": lambda g: g,
    "geglu": F.gelu,
    "reglu": F.relu,
    "swiglu": F.silu,
}
GLU_KIND = {"liglu": 0, "geglu": 1, "reglu": 2, "swiglu": 3}


def glu(x, kind: str):
    """x1 * act(x2): the SECOND half is the gate (model/glu_activations.py:11-54)."""
    a, g = x.chunk(2, dim=-1)
    return a * GLU_ACTS[kind](g)


def bias_gelu(x, bias, tanh_approx: bool):
    t = x if bias is None else x + bias
    return F.gelu(t, approximate="tanh" if tanh_approx else "none")


def scaled_softmax(x, scale, mask=None, causal=False):
    """The three fused-softmax variants of model/fused_softmax.py. mask True = masked."""
    s = x.float() * scale
    if causal:
        sq, sk = s.shape[-2], s.shape[-1]
        tri = torch.ones(sq, sk, device=x.device, dtype=torch.bool).triu(1 + sk - sq)
        s = s.masked_fill(tri, float("-inf"))
    if mask is not None:
        s = s.masked_fill(mask, -10000.0)
    y = torch.softmax(s, dim=-1)
    if mask is not None:
        y = y * (~mask.all(dim=-1, keepdim=True)).to(y.dtype)
    return y.type_as(x)


def attention(q, k, v, scale, causal=True, attn_mask=None, dropout_p=0.0):
    """q [sq,b,np,hn], k/v [sk,b,g,hn] -> [sq,b,np*hn]. GQA by head repetition."""
    sq, b, np_, hn = q.shape
    g = k.shape[2]
    qf = q.permute(1, 2, 0, 3)
    kf = k.permute(1, 2, 0, 3)
    vf = v.permute(1, 2, 0, 3)
    if g != np_:
        kf = kf.repeat_interleave(np_ // g, dim=1)
        vf = vf.repeat_interleave(np_ // g, dim=1)
    sk = kf.shape[2]
    scores = torch.matmul(qf.float(), kf.float().transpose(-1, -2)) * scale
    if causal:
        tri = torch.ones(sq, sk, device=q.device, dtype=torch.bool).triu(1 + sk - sq)
        scores = scores.masked_fill(tri, float("-inf"))
    if attn_mask is not None:
        scores = scores.masked_fill(attn_mask, -10000.0)
    p = torch.softmax(scores, dim=-1)
    if dropout_p > 0:
        p = F.dropout(p, dropout_p)
    o = torch.matmul(p, vf.float()).to(q.dtype)
    return o.permute(2, 0, 1, 3).reshape(sq, b, np_ * hn)


def adam_update(p, g, m, v, lr, beta1, beta2, eps, wd, step, adamw=True, bias_correction=True):
    """In-place AdamW math used by the fused kernel (apex FusedAdam semantics)."""
    if not adamw and wd != 0:
        g = g + wd * p
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1 = 1 - beta1 ** step if bias_correction else 1.0
    bc2 = 1 - beta2 ** step if bias_correction else 1.0
    upd = (m / bc1) / ((v / bc2).sqrt() + eps)
    if adamw and wd != 0:
        upd = upd + wd * p
    p.add_(upd, alpha=-lr)


def default_softmax_scale(head_dim):
    return 1.0 / math.sqrt(head_dim)
