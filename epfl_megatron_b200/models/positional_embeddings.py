"""RoPE helpers (megatron/model/positional_embeddings.py:7-51 surface). The rotation
table is a real (cos, sin) tensor kept on the compute device (the reference keeps a
complex64 CPU tensor and copies it to the GPU on every call)."""
import torch

from ..ops import ref


def precompute_freqs_cis(dim: int, end: int, theta: float = 10000.0, scaling_factor: float = 1.0,
                         device=None) -> torch.Tensor:
    """[end, dim/2, 2] fp32 (cos, sin); positions are divided by ``scaling_factor``."""
    return ref.rope_table(dim, end, theta, scaling_factor, device)


def apply_rotary_emb(xq, xk, freqs_cis, position_ids=None):
    """xq/xk: [s, b, heads, hn]; position_ids: [b, s] or None."""
    return (ref.apply_rope(xq, freqs_cis, position_ids), ref.apply_rope(xk, freqs_cis, position_ids))
