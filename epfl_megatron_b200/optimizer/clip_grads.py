"""Global-norm gradient clipping over lists of tensors
(megatron/optimizer/clip_grads.py:16-136 surface). The flat-buffer optimizers use the
fused, sync-free variant in ``optimizer.py``; these functions serve the generic path
(tasks, tests, user code) and return device tensors rather than Python floats."""
import math

import torch
import torch.distributed as dist

from ..parallel import state as mpu
from ..parallel.layers import param_is_not_tensor_parallel_duplicate
from ..models.module import param_is_not_shared


def _group_or_none(g):
    return g if dist.is_initialized() else None


def clip_grad_norm_fp32(parameters, grads_for_norm, max_norm, norm_type=2, model_parallel_group=None):
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    if isinstance(grads_for_norm, torch.Tensor):
        grads_for_norm = [grads_for_norm]
    grads = [p.grad.detach() for p in parameters if p.grad is not None]
    max_norm, norm_type = float(max_norm), float(norm_type)
    dev = grads[0].device if grads else (grads_for_norm[0].device if grads_for_norm else "cpu")
    if norm_type == math.inf:
        total = torch.zeros(1, device=dev)
        for g in grads_for_norm:
            total = torch.maximum(total, g.abs().max().float().view(1))
        if _group_or_none(model_parallel_group) is not None:
            dist.all_reduce(total, op=dist.ReduceOp.MAX, group=model_parallel_group)
        total_norm = total[0]
    else:
        if grads_for_norm:
            norms = torch._foreach_norm(grads_for_norm, norm_type)
            total = torch.stack([n.float() ** norm_type for n in norms]).sum().view(1)
        else:
            total = torch.zeros(1, device=dev)
        if _group_or_none(model_parallel_group) is not None:
            dist.all_reduce(total, op=dist.ReduceOp.SUM, group=model_parallel_group)
        total_norm = total[0] ** (1.0 / norm_type)
    clip_coeff = torch.clamp(max_norm / (total_norm + 1.0e-6), max=1.0)
    if grads:
        torch._foreach_mul_(grads, clip_coeff)
    return total_norm


def count_zeros_fp32(parameters, model_parallel_group=None):
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    total = None
    for p in parameters:
        g = getattr(p, "grad", None)
        if g is None:
            continue
        if param_is_not_shared(p) and param_is_not_tensor_parallel_duplicate(p):
            z = (g.numel() - torch.count_nonzero(g.detach())).float().view(1)
            total = z if total is None else total + z
    if total is None:
        total = torch.zeros(1)
    if _group_or_none(model_parallel_group) is not None:
        dist.all_reduce(total, op=dist.ReduceOp.SUM, group=model_parallel_group)
    return total[0]
