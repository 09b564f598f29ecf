"""Broadcast a dict of integer tensors from TP rank 0 to its group with two collectives
(megatron/core/tensor_parallel/data.py:22-105 behaviour): one for the packed shapes, one
for the flattened payload.

The reference broadcasts the shapes on the GPU and reads them back (``.cpu()``), which drains the
stream once per micro-batch: the host cannot enqueue the next micro-batch while the previous one
runs. Here the shapes go through a host-side (Gloo) group of the same ranks when one exists
(parallel/state.py), and the payload is staged in pinned memory and copied asynchronously, so
the launch queue stays ahead of the GPU."""
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
    cpu_group = state.get_tensor_model_parallel_cpu_group()
    if cpu_group is not None:
        t = torch.tensor(sizes, dtype=torch.long)
        dist.broadcast(t, state.get_tensor_model_parallel_src_rank(), group=cpu_group)
        flat = t.tolist()
    else:
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


_PINNED, _PINNED_EVENTS, _PINNED_NEXT = [], [], 0


def _stage_payload(parts, dev):
    """Concatenate host tensors into one device tensor without a blocking pageable copy: a ring of
    four pinned staging buffers, each reused only after the copy that last read it has finished."""
    global _PINNED_NEXT
    if dev.type != "cuda":
        return torch.cat(parts, dim=0).to(dev)
    if len(parts) == 1 and parts[0].is_cuda:
        return parts[0]
    if len(parts) == 1 and parts[0].is_pinned():
        return parts[0].to(dev, non_blocking=True)
    n = sum(p.numel() for p in parts)
    i = _PINNED_NEXT % 4
    _PINNED_NEXT += 1
    while len(_PINNED) <= i:
        _PINNED.append(None)
        _PINNED_EVENTS.append(None)
    buf = _PINNED[i]
    if buf is None or buf.numel() < n or buf.dtype != parts[0].dtype:
        buf = _PINNED[i] = torch.empty(max(n, 1), dtype=parts[0].dtype).pin_memory()
    elif _PINNED_EVENTS[i] is not None:
        _PINNED_EVENTS[i].synchronize()
    torch.cat([p.cpu() for p in parts], dim=0, out=buf[:n])
    out = buf[:n].to(dev, non_blocking=True)
    ev = torch.cuda.Event()
    ev.record()
    _PINNED_EVENTS[i] = ev
    return out


def broadcast_data(keys, data, datatype):
    if state.get_tensor_model_parallel_world_size() == 1 and data is not None:
        dev = _device()
        return {k: data[k].to(dev, non_blocking=True) for k in keys}
    key_size, key_numel, total = _build_key_size_numel_dictionaries(keys, data)
    dev = _device()
    if state.get_tensor_model_parallel_rank() == 0:
        _check_data_types(keys, data, datatype)
        flat = _stage_payload([data[k].contiguous().view(-1) for k in keys], dev)
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
