"""Small collectives used by generation under pipeline parallelism
(megatron/text_generation/communication.py:13-185 surface).

Everything is a no-op / identity when the pipeline has a single stage. Tensors live on the
device of the process group's backend (CUDA for NCCL, CPU for Gloo).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..parallel import state as mpu


def _dev():
    if dist.is_initialized() and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def recv_from_prev_pipeline_rank_(recv_buffer=None):
    """In-place receive from the previous stage (no-op on the first stage)."""
    if not mpu.is_pipeline_first_stage():
        assert recv_buffer is not None
        dist.recv(recv_buffer, mpu.get_pipeline_model_parallel_prev_rank())


def send_to_next_pipeline_rank(tensor=None):
    if not mpu.is_pipeline_last_stage():
        assert tensor is not None
        dist.send(tensor.contiguous(), mpu.get_pipeline_model_parallel_next_rank())


def broadcast_from_last_pipeline_stage(size, dtype, tensor=None):
    """Every stage gets the last stage's tensor."""
    if mpu.get_pipeline_model_parallel_world_size() == 1:
        return tensor
    if mpu.is_pipeline_last_stage():
        assert tensor is not None
        tensor = tensor.contiguous()
    else:
        tensor = torch.empty(size, dtype=dtype, device=_dev())
    dist.broadcast(tensor, mpu.get_pipeline_model_parallel_last_rank(),
                   group=mpu.get_pipeline_model_parallel_group())
    return tensor


def broadcast_from_last_to_first_pipeline_stage(size, dtype, tensor=None):
    """Only the first stage receives (over the embedding group, which holds first + last)."""
    first, last = mpu.is_pipeline_first_stage(), mpu.is_pipeline_last_stage()
    if first and last:
        return tensor
    if not (first or last):
        return None
    if last:
        tensor = tensor.contiguous()
    else:
        tensor = torch.empty(size, dtype=dtype, device=_dev())
    dist.broadcast(tensor, mpu.get_pipeline_model_parallel_last_rank(), group=mpu.get_embedding_group())
    return tensor


def copy_from_last_to_first_pipeline_stage(size, dtype, tensor=None):
    """tensor (a slice that may be non-contiguous) on the first stage <- last stage's values."""
    first, last = mpu.is_pipeline_first_stage(), mpu.is_pipeline_last_stage()
    if (first and last) or not (first or last):
        return
    buf = tensor.contiguous() if last else torch.empty(size, dtype=dtype, device=_dev())
    dist.broadcast(buf, mpu.get_pipeline_model_parallel_last_rank(), group=mpu.get_embedding_group())
    if first:
        tensor.copy_(buf.view_as(tensor))


def broadcast_tensor(size, dtype, tensor=None, rank=0):
    """World broadcast from `rank`; other ranks allocate."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return tensor
    if dist.get_rank() == rank:
        tensor = tensor.to(_dev()).contiguous()
    else:
        tensor = torch.empty(size, dtype=dtype, device=_dev())
    dist.broadcast(tensor, rank)
    return tensor


def broadcast_list(size, dtype, list_values=None, rank=0):
    t = None
    if not dist.is_initialized() or dist.get_rank() == rank:
        t = torch.tensor(list_values, dtype=dtype, device=_dev())
    return broadcast_tensor(size, dtype, tensor=t, rank=rank)


def broadcast_int_list(size, int_list=None, rank=0):
    return broadcast_list(size, torch.int64, list_values=int_list, rank=rank)


def broadcast_float_list(size, float_list=None, rank=0):
    return broadcast_list(size, torch.float32, list_values=float_list, rank=rank)
