"""SIGTERM -> flag, agreed on by all ranks through an all-gather, so every rank
checkpoints and exits together (megatron/dist_signal_handler.py:6-81 behaviour)."""
from __future__ import annotations

import signal

import torch
import torch.distributed as dist


def _device():
    if dist.is_initialized() and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def all_gather_item(item, dtype, group=None, async_op=False, local_rank=None):
    if not dist.is_available() or not dist.is_initialized():
        return [item]
    dev = _device()
    t = torch.tensor([item], device=dev, dtype=dtype)
    out = [torch.zeros(1, dtype=dtype, device=dev) for _ in range(dist.get_world_size(group))]
    dist.all_gather(out, t, group, async_op)
    return [e.item() for e in out]


class DistributedSignalHandler:
    def __init__(self, sig=signal.SIGTERM):
        self.sig = sig
        self.released = True

    def signals_received(self):
        return all_gather_item(self._signal_received, dtype=torch.int32)

    def __enter__(self):
        self._signal_received = False
        self.released = False
        self.original_handler = signal.getsignal(self.sig)

        def handler(signum, frame):
            self._signal_received = True

        signal.signal(self.sig, handler)
        return self

    def __exit__(self, type, value, tb):
        self.release()

    def release(self):
        if self.released:
            return False
        signal.signal(self.sig, self.original_handler)
        self.released = True
        return True
