"""Broadcast a dict of integer tensors from TP rank 0 to its group with two collectives
(megatron/core/tensor_parallel/data.py:22-105 behaviour): one for the packed shapes, one
for the flattened payload."""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import state

_MAX_DATA_DIM = 5


def _device():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() \
        and dist.get_backend() != "gloo" else torch.device("cpu")


def _check_data_types(keys, data, target_dtype):
    for key in keys:
        assert data[key].dtype == target_dtype, \
            f"{key} has data type {data[key].dtype} which is different than {target_dtype}"


def _build_key_size_numel_dictionaries(keys, data):
    sizes = [0] * (_MAX_DATA_DIM * len(keys))
    if state.get_tensor_model_parallel_rank() == 0:
        for i, key in enumerate(keys):
            assert data[key].dim() < _MAX_DATA_DIM, "you should increase MAX_DATA_DIM"
            for j, s in enumerate(data[key].size()):
                sizes[i * _MAX_DATA_DIM + j] = s
    t = torch.tensor(sizes, dtype=torch.long, device=_device())
    dist.broadcast(t, state.get_tensor_model_parallel_src_rank(),
                   group=state.get_tensor_model_parallel_group())
    flat = t.cpu().tolist()
    key_size, key_numel, total = {}, {}, 0
    for i, key in enumerate(keys):
        shape, n = [], 1
        for j in range(_MAX_DATA_DIM):
            s = flat[i * _MAX_DATA_DIM + j]
            if s == 0:
                break
            shape.append(s)
            n *= s
        key_size[key], key_numel[key] = shape, n
        total += n
    return key_size, key_numel, total


def broadcast_data(keys, data, datatype):
    if state.get_tensor_model_parallel_world_size() == 1 and data is not None:
        dev = _device()
        return {k: data[k].to(dev, non_blocking=True) for k in keys}
    key_size, key_numel, total = _build_key_size_numel_dictionaries(keys, data)
    dev = _device()
    if state.get_tensor_model_parallel_rank() == 0:
        _check_data_types(keys, data, datatype)
        flat = torch.cat([data[k].contiguous().view(-1) for k in keys], dim=0).to(dev)
    else:
        flat = torch.empty(total, device=dev, dtype=datatype)
    dist.broadcast(flat, state.get_tensor_model_parallel_src_rank(),
                   group=state.get_tensor_model_parallel_group())
    out, off = {}, 0
    for key in keys:
        n = key_numel[key]
        out[key] = flat.narrow(0, off, n).view(key_size[key])
        off += n
    return out
