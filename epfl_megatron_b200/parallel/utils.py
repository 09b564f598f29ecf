"""Small helpers shared by the tensor-parallel code (megatron/core/utils.py:10-124 and
megatron/core/tensor_parallel/utils.py:9-108 behaviour)."""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import state


def ensure_divisibility(numerator, denominator):
    assert numerator % denominator == 0, f"{numerator} is not divisible by {denominator}"


def divide(numerator, denominator):
    ensure_divisibility(numerator, denominator)
    return numerator // denominator


def split_tensor_along_last_dim(tensor, num_partitions, contiguous_split_chunks=False):
    size = divide(tensor.size(-1), num_partitions)
    parts = torch.split(tensor, size, dim=-1)
    if contiguous_split_chunks:
        return tuple(p.contiguous() for p in parts)
    return parts


def split_tensor_into_1d_equal_chunks(tensor, new_buffer=False):
    """This rank's 1/tp slice of the flattened tensor."""
    n = tensor.numel() // state.get_tensor_model_parallel_world_size()
    lo = n * state.get_tensor_model_parallel_rank()
    flat = tensor.reshape(-1)
    if new_buffer:
        out = torch.empty(n, dtype=tensor.dtype, device=tensor.device)
        out.copy_(flat[lo:lo + n])
        return out
    return flat[lo:lo + n]


def gather_split_1d_tensor(tensor):
    """Inverse of split_tensor_into_1d_equal_chunks."""
    ws = state.get_tensor_model_parallel_world_size()
    out = torch.empty(tensor.numel() * ws, dtype=tensor.dtype, device=tensor.device)
    all_gather_into(out, tensor.contiguous(), state.get_tensor_model_parallel_group())
    return out


class VocabUtility:
    """[first, last) vocabulary range owned by a rank."""

    @staticmethod
    def vocab_range_from_per_partition_vocab_size(per_partition_vocab_size, rank, world_size):
        lo = rank * per_partition_vocab_size
        return lo, lo + per_partition_vocab_size

    @staticmethod
    def vocab_range_from_global_vocab_size(global_vocab_size, rank, world_size):
        return VocabUtility.vocab_range_from_per_partition_vocab_size(
            divide(global_vocab_size, world_size), rank, world_size)


# -- viewless tensors (pipeline "free the output .data" trick, core/utils.py:48-124) --
def _kernel_make_viewless_tensor(inp, requires_grad):
    out = torch.empty((1,), dtype=inp.dtype, device=inp.device, requires_grad=requires_grad)
    out.data = inp.data
    return out


class MakeViewlessTensor(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp, requires_grad):
        return _kernel_make_viewless_tensor(inp, requires_grad)

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output, None


def make_viewless_tensor(inp, requires_grad, keep_graph):
    if inp._base is None:
        return inp
    if keep_graph:
        return MakeViewlessTensor.apply(inp, requires_grad)
    return _kernel_make_viewless_tensor(inp, requires_grad)


def assert_viewless_tensor(tensor, extra_msg=None):
    if isinstance(tensor, list):
        [assert_viewless_tensor(t) for t in tensor]
        return tensor
    if not isinstance(tensor, torch.Tensor):
        return tensor
    assert tensor._base is None, (
        f"Ensure tensor._base is None before setting tensor.data or storing tensor to memory "
        f"buffer. Otherwise, a memory leak will occur (and likely accumulate over iterations). {extra_msg}")
    return tensor


def safely_set_viewless_tensor_data(tensor, new_data_tensor):
    assert_viewless_tensor(tensor, extra_msg="FYI, tensor._base has shape %s, and new_data_tensor has shape %s." % (
        "--" if tensor._base is None else tensor._base.shape, new_data_tensor.shape))
    tensor.data = new_data_tensor


# -- collectives that also work on Gloo (no reduce_scatter_tensor / all_gather_into_tensor) --
def _is_gloo(group):
    try:
        return dist.get_backend(group) == "gloo"
    except Exception:  # noqa: BLE001
        return False


def all_gather_into(out, inp, group, async_op=False):
    """out (ws*n contiguous along dim 0) <- concat of every rank's inp."""
    if _is_gloo(group):
        ws = dist.get_world_size(group)
        chunks = list(out.view(ws, -1).unbind(0)) if out.dim() == 1 else list(out.chunk(ws, dim=0))
        tmp = [torch.empty_like(inp) for _ in range(ws)]
        work = dist.all_gather(tmp, inp.contiguous(), group=group)
        for c, t in zip(chunks, tmp):
            c.copy_(t.view_as(c))
        return work if async_op else None
    return dist.all_gather_into_tensor(out, inp.contiguous(), group=group, async_op=async_op)


def reduce_scatter_into(out, inp, group, async_op=False):
    """out <- this rank's dim-0 chunk of sum over ranks of inp."""
    if _is_gloo(group):
        ws, r = dist.get_world_size(group), dist.get_rank(group)
        tmp = inp.clone()
        dist.all_reduce(tmp, group=group)
        chunk = tmp.view(ws, -1)[r] if tmp.dim() == 1 else tmp.chunk(ws, dim=0)[r]
        out.copy_(chunk.view_as(out))
        return None
    return dist.reduce_scatter_tensor(out, inp.contiguous(), group=group, async_op=async_op)
