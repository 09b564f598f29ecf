"""Local data-parallel wrapper with one contiguous gradient buffer per dtype
(megatron/model/distributed.py:15-232 surface: ``main_grad`` views, ``zero_grad_buffer``,
``allreduce_gradients``).

Differences that matter on B200:
* zeroing is LOGICAL: ``zero_grad_buffer`` marks every parameter's slice as empty and the
  first gradient of the step is stored (beta = 0) instead of accumulated, so the 4 B/param
  memset and the first read-modify-write disappear (the reference memsets the whole
  buffer every step, training.py:397-399);
* weight gradients of the linear layers are accumulated into ``main_grad`` by the GEMM
  epilogue itself (ops.functional.wgrad_accumulate);
* the buffer is split into buckets in the order gradients are produced; with
  ``overlap_grad_reduce`` a bucket's data-parallel average is launched (asynchronously, on
  NCCL's stream) the moment every parameter of the bucket has received its last
  gradient of the step, i.e. under the rest of the backward pass. The reference reduces
  once, after backward, with nothing to hide it behind (distributed.py:202-232).
"""
from __future__ import annotations

import math
from abc import ABC, abstractmethod

import torch
import torch.distributed as dist

from ..parallel import state as mpu
from .module import MegatronModule


def default_param_group_key(name, param):
    """(wd_mult, lr_mult): biases and 1-D parameters get no weight decay
    (megatron/optimizer/__init__.py:13-61 rule)."""
    no_wd = name.endswith(".bias") or param.dim() == 1
    return (0.0 if no_wd else 1.0, 1.0)


class MemoryBuffer:
    def __init__(self, numel, numel_padded, dtype, device, symmetric=False):
        self.numel, self.numel_padded, self.dtype = numel, numel_padded, dtype
        if symmetric:
            # own cudaMalloc allocation that the data-parallel peers map (parallel/dp_symm.py)
            from ..parallel.dp_symm import symm_tensor
            self.data = symm_tensor(numel_padded, dtype, device)
        else:
            self.data = torch.zeros(numel_padded, dtype=dtype, device=device, requires_grad=False)

    def zero(self):
        self.data.zero_()

    def get(self, shape, start_index):
        end = start_index + shape.numel()
        assert end <= self.numel, "requested tensor is out of the buffer range."
        return self.data[start_index:end].view(shape)


class DistributedDataParallelBase(MegatronModule, ABC):
    def __init__(self, module):
        super().__init__()
        self.module = module

    @abstractmethod
    def allreduce_gradients(self):
        pass

    def forward(self, *inputs, **kwargs):
        return self.module(*inputs, **kwargs)

    def state_dict(self, prefix="", keep_vars=False):
        return self.module.state_dict(prefix=prefix, keep_vars=keep_vars)

    def state_dict_for_save_checkpoint(self, prefix="", keep_vars=False):
        return self.module.state_dict_for_save_checkpoint(prefix=prefix, keep_vars=keep_vars)

    def load_state_dict(self, state_dict, strict=True):
        self.module.load_state_dict(state_dict, strict=strict)


class _Bucket:
    __slots__ = ("dtype", "start", "end", "params", "pending", "early_ok", "launched", "index")

    def __init__(self, dtype, start, index=0):
        self.dtype, self.start, self.end, self.index = dtype, start, start, index
        self.params, self.pending, self.early_ok, self.launched = [], 0, True, False


class DistributedDataParallel(DistributedDataParallelBase):
    def __init__(self, module, accumulate_allreduce_grads_in_fp32, use_contiguous_buffers,
                 fuse_wgrad: bool = True, group_key_fn=None, overlap_grad_reduce: bool = False,
                 bucket_mb: float = 256.0, fused_dp: bool = False):
        super().__init__(module)
        self.overlap_grad_reduce = overlap_grad_reduce
        # fp32 buffers live in peer-mappable memory; the distributed optimizer attaches a
        # DPWorkspace per buffer and the reduction becomes the in-kernel reduce-scatter
        self.fused_dp = fused_dp
        self.fused_dp_required = False        # set by the caller for --b200_fused_dp on
        self._dp_ws = {}
        self._reduced = False
        self.bucket_bytes = int(bucket_mb * (1 << 20))
        self._buckets, self._bucket_of, self._arrivals, self._expected = [], {}, {}, 0
        # gradient contributions a parameter receives per micro-batch (1 for most; 2 for a word
        # embedding that is also the LM head: lookup gradient through the autograd hook + the
        # head's weight-gradient epilogue; 3 for T5's shared table). Measured on the first step.
        self._uses, self._calibrating = None, False
        self._handles = []
        self.accumulate_allreduce_grads_in_fp32 = accumulate_allreduce_grads_in_fp32
        self.use_contiguous_buffers = use_contiguous_buffers
        if accumulate_allreduce_grads_in_fp32:
            assert use_contiguous_buffers
        self._grad_buffers = None
        self._grad_buffer_param_index_map = None
        self._params = [p for p in self.module.parameters() if p.requires_grad]
        if not self.use_contiguous_buffers:
            return
        self._grad_buffers = {}
        self._grad_buffer_param_index_map = {}
        dp = mpu.get_data_parallel_world_size()

        def buf_dtype(p):
            return torch.float if accumulate_allreduce_grads_in_fp32 else p.dtype

        sizes = {}
        for p in self._params:
            sizes[buf_dtype(p)] = sizes.get(buf_dtype(p), 0) + p.data.nelement()
        dev = self._params[0].device if self._params else torch.device("cpu")
        # Layout: parameters are grouped by optimizer group key (weight-decay / lr multipliers)
        # so every optimizer group is ONE contiguous index range of the buffer, and inside a
        # group they are laid out in reverse module order (== order gradients are produced).
        names = {p: n for n, p in self.module.named_parameters()}
        self.group_key_fn = group_key_fn or default_param_group_key
        keyed = [(self.group_key_fn(names.get(p, ""), p), i, p) for i, p in enumerate(self._params)]
        keys = sorted({k for k, _, _ in keyed})
        self.layout = {}        # dtype -> [(param, start, end, key)] in buffer order
        self.group_ranges = {}  # dtype -> {key: (start, end)}
        cursor = {dt: 0 for dt in sizes}
        self._bucket_cuts = {dt: set() for dt in sizes}
        import os
        lanes_on = fused_dp and os.environ.get("B200_DP_LANES", "1") != "0"
        for key in keys:
            members = [(i, p) for k, i, p in keyed if k == key]
            for i, p, cut in self._buffer_order(list(reversed(members)), buf_dtype, dp if lanes_on else 1):
                dt = buf_dtype(p)
                if cut:
                    self._bucket_cuts[dt].add((cursor[dt] + 7) // 8 * 8)
                start = cursor[dt]
                # keep every parameter 16-byte aligned in the bf16 mirror of this index space
                start = (start + 7) // 8 * 8
                end = start + p.data.nelement()
                cursor[dt] = end
                self.layout.setdefault(dt, []).append((p, start, end, key))
                lo, hi = self.group_ranges.setdefault(dt, {}).get(key, (start, start))
                self.group_ranges[dt][key] = (lo, end)
        for dt in list(sizes):
            n = cursor[dt]
            unit = dp * 8
            self._grad_buffers[dt] = MemoryBuffer(n, unit * int(math.ceil(n / unit)), dt, dev,
                                                  symmetric=fused_dp and dt == torch.float and dp > 1)
            # a group range is extended to the start of the next group so ranges tile the buffer
            ordered = sorted(self.group_ranges[dt].items(), key=lambda kv: kv[1][0])
            for j, (key, (lo, hi)) in enumerate(ordered):
                nxt = ordered[j + 1][1][0] if j + 1 < len(ordered) else self._grad_buffers[dt].numel_padded
                self.group_ranges[dt][key] = (lo, nxt)
            for p, start, end, key in self.layout[dt]:
                p.main_grad = self._grad_buffers[dt].get(p.data.shape, start)
                p._b200_fuse_wgrad = bool(fuse_wgrad and dt == torch.float)
                p._b200_main_grad_empty = False
                self._grad_buffer_param_index_map.setdefault(dt, {})[p] = (start, end)
        self._build_buckets()
        self.grad_accs = []
        for p in self._params:
            tmp = p.expand_as(p)
            acc = tmp.grad_fn.next_functions[0][0]
            acc.register_hook(self._make_param_hook(p))
            self.grad_accs.append(acc)

    def _buffer_order(self, produced, buf_dtype, lanes):
        """(index, param, starts_a_bucket) in buffer order. ``produced`` is the order in which
        backward finishes the gradients. With ``lanes`` > 1 (fused data-parallel reduction, where a
        rank reduces the buckets that fall into ITS contiguous 1/dp of the buffer) bucket-sized
        runs of parameters are dealt round-robin to ``lanes`` regions laid out one after the
        other, so consecutive buckets of the backward pass belong to different owners and no
        rank ends up with all of its reduction work at the very end of the step."""
        if lanes <= 1:
            return [(i, p, False) for i, p in produced]
        runs, cur, size = [], [], 0
        for i, p in produced:
            cur.append((i, p))
            size += p.data.nelement() * torch.empty((), dtype=buf_dtype(p)).element_size()
            if size >= self.bucket_bytes:
                runs.append(cur)
                cur, size = [], 0
        if cur:
            runs.append(cur)
        out = []
        for lane in range(lanes):
            for run in runs[lane::lanes]:
                out += [(i, p, j == 0) for j, (i, p) in enumerate(run)]
        return out

    def _make_param_hook(self, param):
        def hook(*unused):
            if param.grad is not None:
                if getattr(param, "_b200_main_grad_empty", False):
                    param.main_grad.copy_(param.grad.data)
                    param._b200_main_grad_empty = False
                else:
                    param.main_grad.add_(param.grad.data)
                param.grad = None
                cb = getattr(param, "_b200_grad_ready", None)
                if cb is not None:
                    cb()
        return hook

    # ---- bucketed, overlapped data-parallel reduction ----------------------------------------
    def _build_buckets(self):
        """Contiguous index ranges of each grad buffer, cut at parameter boundaries once a range
        holds ``bucket_bytes``. Buffer order == order in which backward produces the gradients."""
        for dt, items in self.layout.items():
            esz = self._grad_buffers[dt].data.element_size()
            cur = None
            first = len(self._buckets)
            cuts = self._bucket_cuts.get(dt, ())
            for p, start, end, _ in items:
                if start in cuts:
                    cur = None
                if cur is None:
                    cur = _Bucket(dt, start, len(self._buckets) - first)
                    self._buckets.append(cur)
                cur.params.append(p)
                cur.end = end
                if getattr(p, "shared", False):
                    cur.early_ok = False       # tied weights receive several gradients per micro-batch
                self._bucket_of[p] = cur
                if (cur.end - cur.start) * esz >= self.bucket_bytes:
                    cur = None
            if cur is not None or items:
                # the padding at the end of the buffer travels with the last bucket
                self._buckets[-1].end = self._grad_buffers[dt].numel_padded
        for p in self._params:
            p._b200_grad_ready = (lambda q=p: self._on_grad_ready(q))

    def _on_grad_ready(self, p):
        """Called (stream-ordered) after a gradient contribution of ``p`` was written."""
        if not self._expected:
            return
        n = self._arrivals.get(p, 0) + 1
        self._arrivals[p] = n
        if self._calibrating:
            return
        b = self._bucket_of[p]
        target = self._uses.get(p, 0) * self._expected
        if b.launched or (target and n > target):
            raise RuntimeError(
                "a gradient contribution arrived after its bucket's data-parallel reduction was launched "
                f"(arrival {n}, expected {target}): the parameter is used a different number of times "
                "per micro-batch than on the first step; run with --no_overlap_grad_reduce")
        if n != target:
            return
        b.pending -= 1
        if b.pending == 0 and b.early_ok and not b.launched and self.overlap_grad_reduce:
            self._launch(b, async_op=True)

    def buckets_of(self, dtype):
        return [b for b in self._buckets if b.dtype == dtype]

    def attach_dp_workspace(self, dtype, ws):
        """Called by the distributed optimizer: buckets of this buffer are reduce-scattered by
        parallel/dp_symm.DPWorkspace (only the owner's shard ends up averaged)."""
        self._dp_ws[dtype] = ws

    def _launch(self, b, async_op):
        b.launched = True
        ws = self._dp_ws.get(b.dtype)
        if ws is not None:
            ws.reduce_bucket(b.index, b.start, b.end, overlapped=async_op)
            return
        view = self._grad_buffers[b.dtype].data[b.start:b.end]
        grp = mpu.get_data_parallel_group()
        avg = dist.ReduceOp.AVG if view.is_cuda else dist.ReduceOp.SUM
        if not view.is_cuda:
            view.mul_(1.0 / mpu.get_data_parallel_world_size())
        h = dist.all_reduce(view, op=avg, group=grp, async_op=async_op)
        if async_op:
            self._handles.append(h)

    def finish_grad_reduce(self) -> bool:
        """Reduce whatever was not launched during backward and wait for everything. Returns True
        when this wrapper averaged its whole buffer over the data-parallel group (the optimizer
        then skips its own reduction), False when overlap is off."""
        if not self._expected:
            return self._reduced
        if self._calibrating:
            # first step: nothing was launched early; learn how often each parameter is written
            self._uses = {}
            for p, n in self._arrivals.items():
                if n % self._expected == 0:
                    self._uses[p] = n // self._expected
                else:
                    self._bucket_of[p].early_ok = False
            self._calibrating = False
        for b in self._buckets:
            if not b.launched:
                self._launch(b, async_op=False)
        for h in self._handles:
            h.wait()
        self._handles = []
        for ws in self._dp_ws.values():
            ws.finish()
        self._expected = 0
        self._reduced = True
        return True

    def zero_grad_buffer(self):
        """Logical zero: the next gradient written to each slice overwrites it."""
        assert self._grad_buffers is not None, "buffers are not initialized."
        for p in self._params:
            p._b200_main_grad_empty = True
        self._expected = 0
        self._reduced = False
        if (self.overlap_grad_reduce or self._dp_ws) and self._buckets and dist.is_initialized() \
                and mpu.get_data_parallel_world_size() > 1:
            from .. import global_vars
            self._expected = global_vars.get_num_microbatches()
            self._arrivals = {}
            self._calibrating = self._uses is None
            for b in self._buckets:
                b.pending, b.launched = len(b.params), False
            for ws in self._dp_ws.values():
                ws.new_step()

    def finalize_grad_buffer(self):
        """Physically zero the slices that received no gradient this step."""
        for p in self._params:
            if getattr(p, "_b200_main_grad_empty", False):
                p.main_grad.zero_()
                p._b200_main_grad_empty = False

    def broadcast_params(self):
        for p in self.module.parameters():
            dist.broadcast(p.data, src=mpu.get_data_parallel_src_rank(), group=mpu.get_data_parallel_group())

    def allreduce_gradients(self):
        """Average gradients over the data-parallel group."""
        dp = mpu.get_data_parallel_world_size()
        if self._grad_buffers is not None:
            self.finalize_grad_buffer()
            for _, buf in self._grad_buffers.items():
                if dp > 1:
                    buf.data /= dp
                    dist.all_reduce(buf.data, group=mpu.get_data_parallel_group())
            return
        if dp == 1:
            return
        buckets = {}
        for p in self.module.parameters():
            if p.requires_grad and p.grad is not None:
                buckets.setdefault(p.data.type(), []).append(p)
        for tp, bucket in buckets.items():
            grads = [p.grad.data for p in bucket]
            flat = torch._utils._flatten_dense_tensors(grads)
            flat /= dp
            dist.all_reduce(flat, group=mpu.get_data_parallel_group())
            for g, s in zip(grads, torch._utils._unflatten_dense_tensors(flat, grads)):
                g.copy_(s)
