"""The seven autograd mappings between tensor-parallel regions
(megatron/core/tensor_parallel/mappings.py:13-278 behaviour):

    copy <-> reduce, scatter(last dim) <-> gather(last dim),
    SP scatter(first dim) <-> SP gather(first dim),
    SP gather <-> reduce-scatter, reduce-scatter <-> gather.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import state
from .utils import all_gather_into, reduce_scatter_into, split_tensor_along_last_dim


def _tp():
    return state.get_tensor_model_parallel_world_size()


def _reduce(x):
    if _tp() == 1:
        return x
    dist.all_reduce(x, group=state.get_tensor_model_parallel_group())
    return x


def _split_along_last_dim(x):
    if _tp() == 1:
        return x
    parts = split_tensor_along_last_dim(x, _tp())
    return parts[state.get_tensor_model_parallel_rank()].contiguous()


def _split_along_first_dim(x):
    ws = _tp()
    if ws == 1:
        return x
    assert x.shape[0] % ws == 0, "first dimension not divisible by tensor parallel size"
    n = x.shape[0] // ws
    r = state.get_tensor_model_parallel_rank()
    return x[r * n:(r + 1) * n].contiguous()


def _gather_along_last_dim(x):
    ws = _tp()
    if ws == 1:
        return x
    parts = [torch.empty_like(x) for _ in range(ws)]
    parts[state.get_tensor_model_parallel_rank()] = x
    dist.all_gather(parts, x.contiguous(), group=state.get_tensor_model_parallel_group())
    return torch.cat(parts, dim=-1).contiguous()


def _gather_along_first_dim(x):
    ws = _tp()
    if ws == 1:
        return x
    shape = list(x.shape)
    shape[0] *= ws
    out = torch.empty(shape, dtype=x.dtype, device=x.device)
    all_gather_into(out, x.contiguous(), state.get_tensor_model_parallel_group())
    return out


def _reduce_scatter_along_first_dim(x):
    ws = _tp()
    if ws == 1:
        return x
    shape = list(x.shape)
    assert shape[0] % ws == 0, "first dimension not divisible by tensor parallel size"
    shape[0] //= ws
    out = torch.empty(shape, dtype=x.dtype, device=x.device)
    reduce_scatter_into(out, x.contiguous(), state.get_tensor_model_parallel_group())
    return out


def _mk(name, fwd, bwd):
    def forward(ctx, x):
        return fwd(x)

    def backward(ctx, g):
        return bwd(g)

    def symbolic(graph, x):
        return fwd(x)
    return type(name, (torch.autograd.Function,), {
        "forward": staticmethod(forward), "backward": staticmethod(backward),
        "symbolic": staticmethod(symbolic)})


_ident = lambda t: t  # noqa: E731

_CopyToModelParallelRegion = _mk("_CopyToModelParallelRegion", _ident, lambda g: _reduce(g.contiguous()))
_ReduceFromModelParallelRegion = _mk("_ReduceFromModelParallelRegion", lambda x: _reduce(x.contiguous()), _ident)
_ScatterToModelParallelRegion = _mk("_ScatterToModelParallelRegion", _split_along_last_dim, _gather_along_last_dim)
_GatherFromModelParallelRegion = _mk("_GatherFromModelParallelRegion", _gather_along_last_dim, _split_along_last_dim)
_ScatterToSequenceParallelRegion = _mk("_ScatterToSequenceParallelRegion", _split_along_first_dim, _gather_along_first_dim)
_ReduceScatterToSequenceParallelRegion = _mk("_ReduceScatterToSequenceParallelRegion",
                                             _reduce_scatter_along_first_dim, _gather_along_first_dim)


class _GatherFromSequenceParallelRegion(torch.autograd.Function):
    """Backward is a reduce-scatter when the consumer computes in tensor-parallel mode
    (every rank holds a partial gradient), a plain split otherwise."""

    @staticmethod
    def forward(ctx, x, tensor_parallel_output_grad=True):
        ctx.tp_grad = tensor_parallel_output_grad
        return _gather_along_first_dim(x)

    @staticmethod
    def backward(ctx, g):
        if ctx.tp_grad:
            return _reduce_scatter_along_first_dim(g), None
        return _split_along_first_dim(g), None


def copy_to_tensor_model_parallel_region(x):
    return _CopyToModelParallelRegion.apply(x)


def reduce_from_tensor_model_parallel_region(x):
    return _ReduceFromModelParallelRegion.apply(x)


def scatter_to_tensor_model_parallel_region(x):
    return _ScatterToModelParallelRegion.apply(x)


def gather_from_tensor_model_parallel_region(x):
    return _GatherFromModelParallelRegion.apply(x)


def scatter_to_sequence_parallel_region(x):
    return _ScatterToSequenceParallelRegion.apply(x)


def gather_from_sequence_parallel_region(x, tensor_parallel_output_grad=True):
    return _GatherFromSequenceParallelRegion.apply(x, tensor_parallel_output_grad)


def reduce_scatter_to_sequence_parallel_region(x):
    return _ReduceScatterToSequenceParallelRegion.apply(x)
