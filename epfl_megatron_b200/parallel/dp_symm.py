"""Peer-mapped gradient / parameter buffers of one data-parallel group and the two fused
collectives of the distributed optimizer on them (csrc/dp_comm.cu):

* ``reduce_bucket``: reduce-scatter of one bucket of the fp32 gradient buffer. Every rank tells its
  peers "my copy of bucket b is final" (one flag store per peer); the rank that owns a slice of the
  bucket pulls that slice out of every peer's buffer over NVLink, averages and overwrites its own
  copy. Runs on a side stream under the rest of the backward pass.
* ``adam_push`` (called by the optimizer): the Adam kernel writes the new bf16 parameters of the
  owned shard into the parameter buffer of every rank, so the post-step all-gather is its epilogue;
  ``params_barrier`` then waits until all peers' shards have landed.

Replaces megatron/optimizer/distrib_optimizer.py:553-567 (scale + ``reduce_scatter_tensor``) and
:592-608 (``_all_gather_base`` + per-parameter copies). Only used when all ranks of the group sit on
one NVSwitch domain (same host); otherwise the NCCL path of models/distributed.py stays in charge.

Why no extra barrier protects the buffers: the optimizer all-reduces the gradient-norm statistics
over ALL ranks between the last pull of a step and the first parameter push, so (a) nobody pushes
parameters into a rank that is still running backward, and (b) nobody starts overwriting a gradient
bucket of step e+1 before every owner has pulled step e (the next forward waits for all pushes).
"""
from __future__ import annotations

import ctypes as C
import os
import socket

import torch
import torch.distributed as dist

from ..ops import lib as L
from .symm import _wrap

_KEEP = []      # symmetric allocations live for the life of the process


def symm_tensor(numel: int, dtype: torch.dtype, device) -> torch.Tensor:
    """Zero-filled tensor in its own cudaMalloc allocation (exportable with CUDA IPC)."""
    nbytes = numel * torch.empty((), dtype=dtype).element_size()
    base = C.c_void_p()
    L.call("b200_symm_alloc", max(nbytes, 16), C.byref(base), launches=0)
    t = _wrap(base.value, max(nbytes, 16), device)[:nbytes].view(dtype)
    t._b200_symm_base = base.value
    _KEEP.append(t)
    return t


def group_on_one_host(group) -> bool:
    names = [None] * dist.get_world_size(group)
    dist.all_gather_object(names, socket.gethostname(), group=group)
    return len(set(names)) == 1


def _exchange(group, make_ptrs):
    """Export local allocations and import the peers' ones: for each allocation the base address per
    rank of the group. ``make_ptrs()`` returns the local device pointers (it may allocate). The one
    collective in here is reached by every rank even when a local step failed, so that a failure on
    one rank (CUDA IPC not permitted, out of memory) surfaces as an exception on ALL ranks instead of
    a hang."""
    rank, n = dist.get_rank(group), dist.get_world_size(group)
    ptrs, mine, err = None, None, None
    try:
        ptrs = make_ptrs()
        mine = []
        for ptr in ptrs:
            handle = (C.c_ubyte * 64)()
            L.call("b200_symm_export", ptr, handle, launches=0)
            mine.append(bytes(handle))
    except Exception as e:  # noqa: BLE001
        err, mine = e, None
    everyone = [None] * n
    dist.all_gather_object(everyone, mine, group=group)
    if err is not None:
        raise err
    if any(h is None for h in everyone):
        raise RuntimeError("a peer could not export its buffers")
    out = []
    for i, ptr in enumerate(ptrs):
        bases = []
        for r in range(n):
            if r == rank:
                bases.append(ptr)
                continue
            p = C.c_void_p()
            L.call("b200_symm_import", (C.c_ubyte * 64).from_buffer_copy(everyone[r][i]), C.byref(p), launches=0)
            bases.append(p.value)
        out.append(bases)
    return out


class DPWorkspace:
    """grad: fp32 [N] and param: bf16 [N], both from ``symm_tensor``; shard r = [r*N/dp, (r+1)*N/dp)."""

    def __init__(self, group, grad: torch.Tensor, param: torch.Tensor, n_buckets: int, max_bucket: int = 0):
        self.group = group
        self.dp = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        assert 2 <= self.dp <= 8 and grad.dtype == torch.float32 and param.dtype == torch.bfloat16
        self.N = grad.numel()
        assert self.N % (8 * self.dp) == 0 and param.numel() == self.N
        self.shard = self.N // self.dp
        self.shard_lo, self.shard_hi = self.rank * self.shard, (self.rank + 1) * self.shard
        self.device = grad.device
        self.n_buckets = n_buckets
        def local_buffers():
            self.flags = symm_tensor((n_buckets + 1) * self.dp + 8, torch.int32, self.device)
            return [grad.data_ptr(), param.data_ptr(), self.flags.data_ptr()]

        # the caller follows up with a collective agreement (optimizer._Chunk), which is also the
        # barrier that orders "everyone has mapped everyone" before the first kernel
        self.peer_grad, self.peer_param, self.peer_flags = _exchange(group, local_buffers)
        self.epoch = 0          # one per training step (bucket flags)
        self.param_epoch = 0    # one per successful optimizer step
        self.side = torch.cuda.Stream(device=self.device)
        # "my bucket is final" stores go out on their own stream, which never waits for a peer, so
        # the order in which ranks get to their waits cannot matter
        self.sig = torch.cuda.Stream(device=self.device)
        self.blocks_overlap = int(os.environ.get("B200_DP_BLOCKS", "96"))
        self.blocks_exposed = 148 * 4
        self.allow_overlap = os.environ.get("B200_DP_NO_OVERLAP", "0") != "1"
        # how an overlapped bucket travels: "ce" = copy engines into local staging + local reduce
        # kernel (no SMs taken from the GEMMs), "pull" = peer loads inside the reduce kernel
        self.overlap_mode = os.environ.get("B200_DP_OVERLAP_MODE", "ce")
        self.max_bucket = min(max_bucket, self.shard) if max_bucket else self.shard
        self._stage = None
        self._pending = False

    # ---- gradients -------------------------------------------------------------------------
    def new_step(self):
        self.epoch += 1

    def reduce_bucket(self, index: int, start: int, end: int, overlapped: bool):
        """Announce bucket [start, end) and reduce the part of it this rank owns."""
        main = torch.cuda.current_stream()
        overlapped = overlapped and self.allow_overlap
        if overlapped:
            self.sig.wait_stream(main)
            self.side.wait_stream(main)
            stream, sig = self.side, self.sig
            self._pending = True
        else:
            if self._pending:
                main.wait_stream(self.side)
                main.wait_stream(self.sig)
                self._pending = False
            stream = sig = main
        st = stream.cuda_stream
        dp, me = self.dp, self.rank
        word = index * dp
        addrs = (C.c_void_p * 8)()
        n = 0
        for r in range(dp):
            if r != me:
                addrs[n] = self.peer_flags[r] + 4 * (word + me)
                n += 1
        L.call("b200_symm_signal", addrs, n, self.epoch & 0xFFFFFFFF, sig.cuda_stream)
        lo, hi = max(start, self.shard_lo), min(end, self.shard_hi)
        if lo < hi:
            grads = (C.c_void_p * 8)()
            for r in range(dp):
                grads[r] = self.peer_grad[r] + 4 * lo
            if overlapped and self.overlap_mode == "ce":
                if self._stage is None:
                    self._stage = torch.empty(dp - 1, self.max_bucket, dtype=torch.float32, device=self.device)
                assert hi - lo <= self.max_bucket
                L.call("b200_dp_wait", self.flags.data_ptr() + 4 * word, dp, me, self.epoch & 0xFFFFFFFF, 1, st)
                i = 0
                for r in range(dp):
                    if r != me:
                        L.call("b200_copy_async", self._stage[i].data_ptr(), grads[r], 4 * (hi - lo), st, launches=0)
                        grads[r] = self._stage[i].data_ptr()
                        i += 1
            L.call("b200_dp_reduce", grads, dp, me, hi - lo, 1.0 / dp, self.flags.data_ptr() + 4 * word,
                   self.epoch & 0xFFFFFFFF, self.blocks_overlap if overlapped else self.blocks_exposed, st)

    def finish(self):
        if self._pending:
            torch.cuda.current_stream().wait_stream(self.side)
            torch.cuda.current_stream().wait_stream(self.sig)
            self._pending = False

    # ---- parameters ------------------------------------------------------------------------
    def param_dsts(self, lo: int):
        """Addresses of element ``lo`` of the bf16 parameter buffer on every rank (own first)."""
        d = (C.c_void_p * 8)()
        order = [self.rank] + [r for r in range(self.dp) if r != self.rank]
        for i, r in enumerate(order):
            d[i] = self.peer_param[r] + 2 * lo
        return d

    def params_barrier(self):
        """After the Adam launches of a step: publish 'my shard is everywhere', wait for the peers'."""
        self.param_epoch += 1
        st = L.stream_ptr()
        word = self.n_buckets * self.dp
        addrs = (C.c_void_p * 8)()
        n = 0
        for r in range(self.dp):
            if r != self.rank:
                addrs[n] = self.peer_flags[r] + 4 * (word + self.rank)
                n += 1
        L.call("b200_symm_signal", addrs, n, self.param_epoch & 0xFFFFFFFF, st)
        L.call("b200_dp_wait", self.flags.data_ptr() + 4 * word, self.dp, self.rank,
               self.param_epoch & 0xFFFFFFFF, 2, st)

    def check(self):
        """Synchronise and raise if a cross-GPU wait timed out (debug / tests)."""
        rec = (C.c_uint * 8)()
        L.call("b200_dp_debug_read", rec, launches=0)
        if rec[0]:
            site = {1: "bucket-ready flag of a peer", 2: "parameter shard of a peer"}.get(rec[0], str(rec[0]))
            raise RuntimeError(f"data-parallel fused collective timed out waiting for the {site}: "
                               f"saw {rec[1]}, wanted {rec[2]}, peer {rec[3]}, block {rec[4]}")
