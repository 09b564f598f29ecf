"""Optimizers on flat buffers.

Public surface of megatron/optimizer/optimizer.py:58-783 and
megatron/optimizer/distrib_optimizer.py:331-700 (``reduce_model_grads``, ``step``,
``gather_model_params``, ``zero_grad``, ``state_dict``, ``param_groups``,
``get_loss_scale``, ``scale_loss``, ``reload_model_params``), re-designed for one index
space shared by four flat buffers per model chunk:

    grad   fp32 [N]   (owned by the local DDP wrapper; GEMM epilogues accumulate into it)
    param  bf16 [N]   (the model's parameters are VIEWS into it)
    main   fp32 [n]   master weights      }  n = N for the plain optimizer,
    m, v   fp32 [n]   Adam moments        }  n = N / dp for the distributed optimizer

Every optimizer group (weight-decay / lr multipliers) is one contiguous range of the
index space, so a step is one fused Adam launch per (group x buffer) that reads the
gradient, applies unscale * clip from a DEVICE scalar, updates main/m/v and writes the
bf16 parameters in place. The global gradient norm, the clip coefficient and the
found-inf flag stay on the device: a bf16 step performs no host synchronisation (the
reference syncs on ``.item()`` in clip_grads.py:96 and optimizer.py:397-432).

Distributed optimizer (ZeRO-1): reduce-scatter of the grad buffer over DP, Adam on the
owned shard, all-gather of the bf16 param buffer -- the parameters alias that buffer, so
the per-parameter copy-back of the reference (distrib_optimizer.py:592-610) disappears.
"""
from __future__ import annotations

import math
from abc import ABC, abstractmethod

import torch
import torch.distributed as dist

from ..models.module import param_is_not_shared
from ..ops import lib as L
from ..ops import ref as ops_ref
from ..parallel import state as mpu
from ..parallel.layers import param_is_not_tensor_parallel_duplicate
from ..parallel.utils import all_gather_into, reduce_scatter_into


def _unwrap(model, classes):
    while isinstance(model, classes):
        model = model.module
    return model


class _Segment:
    """One optimizer group restricted to the locally owned part of one flat buffer."""
    __slots__ = ("chunk", "group", "lo", "hi")

    def __init__(self, chunk, group, lo, hi):
        self.chunk, self.group, self.lo, self.hi = chunk, group, lo, hi


class _Chunk:
    """Flat state attached to one grad buffer of one model chunk."""

    def __init__(self, ddp, dtype, distributed, keep_main):
        self.ddp = ddp
        self.grad = ddp._grad_buffers[dtype].data
        self.N = self.grad.numel()
        self.layout = ddp.layout[dtype]
        self.group_ranges = ddp.group_ranges[dtype]
        dev = self.grad.device
        dp = mpu.get_data_parallel_world_size() if distributed else 1
        r = mpu.get_data_parallel_rank() if distributed else 0
        assert self.N % dp == 0
        self.shard_lo, self.shard_hi = r * (self.N // dp), (r + 1) * (self.N // dp)
        n = self.shard_hi - self.shard_lo
        pdtype = self.layout[0][0].dtype
        # fused data-parallel path (parallel/dp_symm.py): gradient and parameter buffers are mapped
        # by the peers, which pull gradient slices and push parameter shards from inside kernels
        self.ws = None
        fused = (distributed and dp > 1 and getattr(ddp, "fused_dp", False) and dtype == torch.float
                 and pdtype == torch.bfloat16 and hasattr(self.grad, "_b200_symm_base"))
        # parameter buffer: the model parameters become views into it
        if fused:
            from ..parallel.dp_symm import DPWorkspace, symm_tensor
            self.param = symm_tensor(self.N, pdtype, dev)
        else:
            self.param = torch.zeros(self.N, dtype=pdtype, device=dev)
        for p, s, e, _ in self.layout:
            self.param[s:e].copy_(p.data.reshape(-1))
            p.data = self.param[s:e].view(p.data.shape)
        self.keep_main = keep_main and pdtype != torch.float32
        if self.keep_main:
            self.main = self.param[self.shard_lo:self.shard_hi].float()
        else:
            self.main = self.param[self.shard_lo:self.shard_hi]
        self.m = torch.zeros(n, dtype=torch.float32, device=dev)
        self.v = torch.zeros(n, dtype=torch.float32, device=dev)
        if fused:
            bks = ddp.buckets_of(dtype)
            grp = mpu.get_data_parallel_group()
            ws, err = None, ""
            try:
                ws = DPWorkspace(grp, self.grad, self.param, len(bks), max(b.end - b.start for b in bks))
            except Exception as e:  # noqa: BLE001 - e.g. CUDA IPC not permitted in this container
                err = f"{type(e).__name__}: {e}"
            # every rank of the group must take the same path
            ok = torch.tensor([1 if ws is not None else 0], device=dev, dtype=torch.int32)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=grp)
            if int(ok.item()) == 1:
                self.ws = ws
                ddp.attach_dp_workspace(dtype, self.ws)
            else:
                if getattr(ddp, "fused_dp_required", False):
                    raise RuntimeError(f"--b200_fused_dp on: peer-mapped workspace could not be created ({err})")
                if dist.get_rank() == 0:
                    print(f"WARNING: fused data-parallel path disabled, falling back to NCCL ({err or 'a peer failed'})",
                          flush=True)

    def local(self, lo, hi):
        """Intersection of [lo, hi) with the owned shard, in shard-local coordinates."""
        a, b = max(lo, self.shard_lo), min(hi, self.shard_hi)
        if a >= b:
            return None
        return a - self.shard_lo, b - self.shard_lo


class MegatronOptimizer(ABC):
    """Base: bookkeeping shared by all optimizers."""

    def __init__(self, models, cfg, grad_scaler=None):
        self.models = models
        self.cfg = cfg
        self.grad_scaler = grad_scaler
        self.clip_grad = cfg["clip_grad"]
        self.log_num_zeros_in_grad = cfg["log_num_zeros_in_grad"]

    @abstractmethod
    def step(self, args, timers):
        pass

    def get_loss_scale(self):
        if self.grad_scaler is None:
            return self._one
        return self.grad_scaler.scale

    def scale_loss(self, loss):
        if self.grad_scaler is None:
            return loss
        return self.get_loss_scale() * loss

    def gather_model_params(self, args, timers):
        pass


class FlatOptimizer(MegatronOptimizer):
    """Adam/SGD over flat buffers; ``distributed=True`` shards state over the DP group."""

    def __init__(self, models, cfg, grad_scaler=None, distributed=False):
        super().__init__(models, cfg, grad_scaler)
        self.distributed = distributed
        self.kind = cfg["optimizer"]
        self.step_count = 0
        self.chunks = []
        keys = set()
        for model in models:
            ddp = self._find_ddp(model)
            for dtype in ddp._grad_buffers:
                ch = _Chunk(ddp, dtype, distributed, keep_main=True)
                self.chunks.append(ch)
                keys.update(ch.group_ranges)
        self.param_groups = []
        self._group_index = {}
        for key in sorted(keys, reverse=True):   # (1,1) wd group first, like the reference
            wd_mult, lr_mult = key
            self._group_index[key] = len(self.param_groups)
            self.param_groups.append({"lr": cfg["lr"], "weight_decay": cfg["weight_decay"] * wd_mult,
                                      "wd_mult": wd_mult, "lr_mult": lr_mult,
                                      "betas": (cfg["adam_beta1"], cfg["adam_beta2"]),
                                      "eps": cfg["adam_eps"], "momentum": cfg["sgd_momentum"],
                                      "params": []})
        self.segments = []
        for ch in self.chunks:
            for key, (lo, hi) in ch.group_ranges.items():
                loc = ch.local(lo, hi)
                if loc is not None:
                    self.segments.append(_Segment(ch, self._group_index[key], loc[0], loc[1]))
        dev = self.chunks[0].grad.device if self.chunks else torch.device("cpu")
        self._dev = dev
        self._one = torch.ones(1, dtype=torch.float32, device=dev)
        self._coef = torch.ones(1, dtype=torch.float32, device=dev)
        self._skip = torch.zeros(1, dtype=torch.float32, device=dev)
        self._norm_out = torch.zeros(2, dtype=torch.float32, device=dev)
        self._use_kernels = dev.type == "cuda" and L.available()
        # Adam writes the bf16 shard into every rank's parameter buffer (fused all-gather)
        self._push_params = self.kind == "adam" and any(ch.ws is not None for ch in self.chunks)
        if self._use_kernels:
            self._sumsq_scratch = torch.empty(L.lib().b200_sumsq_scratch_elems(), dtype=torch.float32,
                                              device=dev)
        self._dup_views = self._collect_excluded_views()
        self._seqpar_params, self._all_params = [], []
        for ch in self.chunks:
            for p, s, e, _ in ch.layout:
                self._all_params.append(p)
                if getattr(p, "sequence_parallel", False):
                    self._seqpar_params.append(p)

    # ------------------------------------------------------------------ helpers
    @staticmethod
    def _find_ddp(model):
        m = model
        while not hasattr(m, "_grad_buffers"):
            assert hasattr(m, "module"), "FlatOptimizer needs the local DDP wrapper (--DDP_impl local)"
            m = m.module
        assert m._grad_buffers is not None, "contiguous grad buffers are required"
        return m

    def _collect_excluded_views(self):
        """Grad slices that must NOT count in the global norm on this rank: parameters
        replicated across TP (counted on tp rank 0 only) and shared (tied) parameters."""
        views = []
        for ch in self.chunks:
            for p, s, e, _ in ch.layout:
                if param_is_not_shared(p) and param_is_not_tensor_parallel_duplicate(p):
                    continue
                loc = ch.local(s, e)
                if loc is not None:
                    views.append(ch.grad[ch.shard_lo + loc[0]:ch.shard_lo + loc[1]])
        return views

    def get_parameters(self):
        return list(self._all_params)

    def zero_grad(self, set_to_none=True):
        for p in self._all_params:
            p.grad = None

    def reload_model_params(self):
        """Master weights <- model weights (after loading a checkpoint without optimizer state)."""
        for ch in self.chunks:
            if ch.keep_main:
                ch.main.copy_(ch.param[ch.shard_lo:ch.shard_hi])

    # ------------------------------------------------------------------ grad reduction
    def allreduce_layernorm_grads(self, args):
        if mpu.get_tensor_model_parallel_world_size() > 1 and args.sequence_parallel and self._seqpar_params:
            grads = [p.main_grad for p in self._seqpar_params]
            flat = torch._utils._flatten_dense_tensors(grads)
            dist.all_reduce(flat, group=mpu.get_tensor_model_parallel_group())
            for g, s in zip(grads, torch._utils._unflatten_dense_tensors(flat, grads)):
                g.copy_(s)

    def allreduce_word_embedding_grads(self, args):
        if not dist.is_initialized():
            return
        if mpu.is_rank_in_embedding_group(ignore_virtual=True) and \
                mpu.get_pipeline_model_parallel_world_size() > 1:
            from ..models.distributed import DistributedDataParallelBase
            from ..models.module import Float16Module
            if mpu.is_pipeline_first_stage(ignore_virtual=True):
                unwrapped = self.models[0]
            elif mpu.is_pipeline_last_stage(ignore_virtual=True):
                unwrapped = self.models[-1]
            else:
                unwrapped = self.models[0]
            unwrapped = _unwrap(unwrapped, (DistributedDataParallelBase, Float16Module))
            if getattr(unwrapped, "share_word_embeddings", False):
                w = unwrapped.word_embeddings_weight()
                g = w.main_grad if hasattr(w, "main_grad") else w.grad
                dist.all_reduce(g, group=mpu.get_embedding_group())

    def allreduce_position_embedding_grads(self, args):
        if dist.is_initialized() and mpu.is_rank_in_position_embedding_group() and \
                mpu.get_pipeline_model_parallel_world_size() > 1 and \
                args.pipeline_model_parallel_split_rank is not None:
            from ..models.distributed import DistributedDataParallelBase
            from ..models.module import Float16Module
            m = _unwrap(self.models[0], (DistributedDataParallelBase, Float16Module))
            w = m.language_model.embedding.position_embeddings.weight
            dist.all_reduce(w.main_grad, group=mpu.get_position_embedding_group())

    def reduce_model_grads(self, args, timers):
        # Order: (1) tied-embedding all-reduces over the pipeline ends, on gradients no bucket has
        # touched yet (buckets holding shared parameters are never launched early) -- they must see
        # whole gradients, and a reduce-scatter leaves only the owner's shard valid; (2) the
        # data-parallel reduction, which also waits for the buckets launched under backward;
        # (3) the tensor-parallel all-reduce of the sequence-parallel norm gradients: all ranks of
        # a TP group own the same shard range, and doing it after (2) keeps it from racing with
        # bucket reductions still in flight. Sums commute, the result equals the reference's
        # layernorm -> embedding -> data-parallel order (optimizer.py:280-301).
        for ch in self.chunks:
            ch.ddp.finalize_grad_buffer()
        timers("embedding-grads-all-reduce", log_level=1).start(barrier=args.barrier_with_L1_time)
        self.allreduce_word_embedding_grads(args)
        self.allreduce_position_embedding_grads(args)
        timers("embedding-grads-all-reduce").stop()
        timers("grads-reduce", log_level=1).start(barrier=args.barrier_with_L1_time)
        dp = mpu.get_data_parallel_world_size()
        if dp > 1:
            grp = mpu.get_data_parallel_group()
            for ch in self.chunks:
                if ch.ddp.finish_grad_reduce():
                    continue     # averaged bucket by bucket under the backward pass
                if self._use_kernels:
                    L.call("b200_scale", ch.grad.data_ptr(), ch.N, None, 1.0 / dp, L.stream_ptr())
                else:
                    ch.grad.mul_(1.0 / dp)
                if self.distributed:
                    reduce_scatter_into(ch.grad[ch.shard_lo:ch.shard_hi], ch.grad, grp)
                else:
                    dist.all_reduce(ch.grad, group=grp)
        timers("grads-reduce").stop()
        timers("layernorm-grads-all-reduce", log_level=1).start(barrier=args.barrier_with_L1_time)
        self.allreduce_layernorm_grads(args)
        timers("layernorm-grads-all-reduce").stop()

    def gather_model_params(self, args, timers):
        if not self.distributed or mpu.get_data_parallel_world_size() == 1:
            return
        timers("params-all-gather", log_level=1).start(barrier=args.barrier_with_L1_time)
        grp = mpu.get_data_parallel_group()
        for ch in self.chunks:
            if ch.ws is not None and self._push_params:
                ch.ws.params_barrier()     # the Adam kernel already wrote every rank's copy
            else:
                all_gather_into(ch.param, ch.param[ch.shard_lo:ch.shard_hi], grp)
        timers("params-all-gather").stop()

    def check_dp_consistency(self):
        """Debug aid for the in-kernel collectives: after a step every
        data-parallel replica must hold bit-identical parameters. Compares a 64-bit checksum of each
        parameter buffer across the data-parallel group and raises on the first mismatch. Syncs."""
        if not dist.is_initialized() or mpu.get_data_parallel_world_size() == 1:
            return
        grp = mpu.get_data_parallel_group()
        for i, ch in enumerate(self.chunks):
            words = ch.param.view(torch.int16 if ch.param.element_size() == 2 else torch.int32)
            # position-weighted so that swapped shards do not cancel
            n = words.numel()
            w = (torch.arange(n, device=words.device, dtype=torch.int64) % 8191) + 1
            cs = (words.to(torch.int64) * w).sum().view(1)
            lo, hi = cs.clone(), cs.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=grp)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=grp)
            if int(lo) != int(hi):
                raise RuntimeError(f"data-parallel replicas diverged: parameter buffer {i} checksum {int(cs)} on "
                                   f"dp rank {mpu.get_data_parallel_rank()}, group min {int(lo)} max {int(hi)}")
            if ch.ws is not None:
                ch.ws.check()

    # ------------------------------------------------------------------ step
    def _grad_sumsq(self):
        """out[0] = sum of squares of the locally owned, non-duplicated gradients; out[1] =
        non-finite flag. Stays on the device."""
        out = self._norm_out
        if self._use_kernels:
            first = True
            for ch in self.chunks:
                g = ch.grad[ch.shard_lo:ch.shard_hi]
                L.call("b200_sumsq", g.data_ptr(), g.numel(), out.data_ptr(), self._sumsq_scratch.data_ptr(),
                       0 if first else 1, L.stream_ptr(), launches=2)
                first = False
            if first:
                out.zero_()
        else:
            tot = torch.zeros((), dtype=torch.float32, device=self._dev)
            for ch in self.chunks:
                g = ch.grad[ch.shard_lo:ch.shard_hi]
                tot = tot + (g.double() ** 2).sum().float()
            out[0] = tot
            out[1] = (~torch.isfinite(tot)).float()
        if self._dup_views:
            norms = torch._foreach_norm(self._dup_views)
            out[0] -= torch.stack([n.float() ** 2 for n in norms]).sum()
        return out

    def step(self, args, timers):
        timers("optimizer-clip-main-grad", log_level=1).start(barrier=args.barrier_with_L1_time)
        stats = self._grad_sumsq().clone()
        if dist.is_initialized():
            # every rank holds a different piece (TP/PP shards; DP shards for the distributed
            # optimizer): the plain optimizer sums over the model-parallel group, the
            # distributed one over all ranks
            grp = None if self.distributed else mpu.get_model_parallel_group()
            if (dist.get_world_size(grp) if grp is not None else dist.get_world_size()) > 1:
                dist.all_reduce(stats, group=grp)
        total_sq = stats[0].clamp_min(0.0)
        found_inf_t = (stats[1] > 0) | ~torch.isfinite(total_sq)
        inv_scale = self.grad_scaler.inv_scale[0] if self.grad_scaler is not None else None
        grad_norm = total_sq.sqrt()
        if inv_scale is not None:
            grad_norm = grad_norm * inv_scale
        coef = torch.ones((), dtype=torch.float32, device=self._dev)
        if self.clip_grad > 0.0:
            coef = torch.clamp(self.clip_grad / (grad_norm + 1.0e-6), max=1.0)
        if inv_scale is not None:
            coef = coef * inv_scale
        self._coef.copy_(coef.view(1))
        self._skip.copy_(found_inf_t.float().view(1))
        timers("optimizer-clip-main-grad").stop()

        found_inf = False
        if self.grad_scaler is not None:
            # only loss-scaled (fp16) training pays for a host sync, as it must decide whether
            # to advance the schedule
            found_inf = bool(found_inf_t.item())
            self.grad_scaler.update(found_inf)
            if found_inf:
                return False, None, None
        num_zeros = self._count_zeros() if self.log_num_zeros_in_grad else None

        timers("optimizer-inner-step", log_level=1).start(barrier=args.barrier_with_L1_time)
        self.step_count += 1
        for seg in self.segments:
            self._update_segment(seg)
        timers("optimizer-inner-step").stop()
        return True, grad_norm, num_zeros

    def _update_segment(self, seg):
        ch, g = seg.chunk, self.param_groups[seg.group]
        lo, hi = seg.lo, seg.hi
        n = hi - lo
        grad = ch.grad[ch.shard_lo + lo:ch.shard_lo + hi]
        main, m, v = ch.main[lo:hi], ch.m[lo:hi], ch.v[lo:hi]
        p16 = ch.param[ch.shard_lo + lo:ch.shard_lo + hi] if ch.keep_main else None
        lr, wd = g["lr"], g["weight_decay"]
        b1, b2 = g["betas"]
        if self._use_kernels and n % 4 == 0:
            if self.kind == "adam" and ch.ws is not None and self._push_params:
                assert n % 8 == 0 and (ch.shard_lo + lo) % 8 == 0
                L.call("b200_adam_step_push", main.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(),
                       ch.ws.param_dsts(ch.shard_lo + lo), ch.ws.dp, n, lr, b1, b2, g["eps"], wd,
                       self.step_count, 1, 1, self._coef.data_ptr(), self._skip.data_ptr(), L.stream_ptr())
            elif self.kind == "adam":
                L.call("b200_adam_step", main.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(),
                       L.ptr(p16), n, lr, b1, b2, g["eps"], wd, self.step_count, 1, 1,
                       self._coef.data_ptr(), self._skip.data_ptr(), L.stream_ptr())
            else:
                L.call("b200_sgd_step", main.data_ptr(), grad.data_ptr(), m.data_ptr(), L.ptr(p16), n, lr,
                       g["momentum"], wd, 1 if self.step_count == 1 else 0, self._coef.data_ptr(),
                       self._skip.data_ptr(), L.stream_ptr())
            return
        # PyTorch path (CPU runs, kernels switched off): same contract as the kernels, which leave
        # the segment untouched when the gradients of this step were not finite
        if bool(self._skip.item()):
            return
        gr = grad * self._coef
        if self.kind == "adam":
            ops_ref.adam_update(main, gr, m, v, lr, b1, b2, g["eps"], wd, self.step_count)
        else:
            gr = gr + wd * main
            if self.step_count == 1:
                m.copy_(gr)
            else:
                m.mul_(g["momentum"]).add_(gr)
            main.add_(m if g["momentum"] != 0 else gr, alpha=-lr)
        if p16 is not None:
            p16.copy_(main)

    def _count_zeros(self):
        tot = torch.zeros((), dtype=torch.float32, device=self._dev)
        for ch in self.chunks:
            g = ch.grad[ch.shard_lo:ch.shard_hi]
            tot = tot + (g.numel() - torch.count_nonzero(g)).float()
        if dist.is_initialized() and dist.get_world_size() > 1:
            t = tot.view(1).clone()
            dist.all_reduce(t, group=None if self.distributed else mpu.get_model_parallel_group())
            tot = t[0]
        return tot

    # ------------------------------------------------------------------ checkpoint
    def _per_param(self, ch, flat, p_start, p_end):
        loc = ch.local(p_start, p_end)
        if loc is None or (loc[1] - loc[0]) != (p_end - p_start):
            return None
        return flat[loc[0]:loc[1]]

    @staticmethod
    def _layout_signature(ch):
        """Order-sensitive fingerprint of (start, end) of every parameter in the flat buffer."""
        h = 1469598103934665603
        for _, s, e, _ in ch.layout:
            for x in (s, e):
                h = ((h ^ x) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
        return h

    def state_dict(self):
        sd = {"step": self.step_count,
              "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}
        if self.grad_scaler:
            sd["grad_scaler"] = self.grad_scaler.state_dict()
        if self.distributed:
            # raw shards: only loadable with the same DP size, like the reference
            sd["shard_fp32_from_float16_groups"] = [ch.main.detach().cpu().clone() for ch in self.chunks]
            sd["optimizer"] = {"state": {i: {"exp_avg": ch.m.cpu().clone(), "exp_avg_sq": ch.v.cpu().clone()}
                                         for i, ch in enumerate(self.chunks)},
                               "param_groups": sd["param_groups"]}
            sd["shard_ranges"] = [(ch.shard_lo, ch.shard_hi, ch.N) for ch in self.chunks]
            # the raw shards are only meaningful for the buffer order they were cut from (it changes
            # with --b200_fused_dp, which rotates bucket owners): store a fingerprint of it
            sd["layout_signature"] = [self._layout_signature(ch) for ch in self.chunks]
            return sd
        state, mains, idx = {}, [], 0
        for ch in self.chunks:
            for p, s, e, _ in ch.layout:
                state[idx] = {"exp_avg": ch.m[s:e].view(p.shape).cpu().clone(),
                              "exp_avg_sq": ch.v[s:e].view(p.shape).cpu().clone(), "step": self.step_count}
                mains.append(ch.main[s:e].view(p.shape).float().cpu().clone())
                idx += 1
        sd["optimizer"] = {"state": state, "param_groups": sd["param_groups"]}
        sd["fp32_from_fp16_params"] = [mains]
        return sd

    def load_state_dict(self, sd):
        self.step_count = int(sd.get("step", 0))
        for g, saved in zip(self.param_groups, sd.get("param_groups", [])):
            for k in ("lr", "weight_decay", "betas", "eps", "momentum"):
                if k in saved:
                    g[k] = saved[k]
        if self.grad_scaler and "grad_scaler" in sd:
            self.grad_scaler.load_state_dict(sd["grad_scaler"])
        inner = sd["optimizer"]["state"]
        if self.distributed:
            for i, ch in enumerate(self.chunks):
                lo, hi, n = sd["shard_ranges"][i]
                assert (lo, hi, n) == (ch.shard_lo, ch.shard_hi, ch.N), \
                    "distributed-optimizer checkpoints can only be loaded with the same data-parallel size"
                sig = sd.get("layout_signature")
                if sig is not None and sig[i] != self._layout_signature(ch):
                    raise RuntimeError(
                        "distributed-optimizer checkpoint was written with a different parameter order in "
                        "the flat buffers (different --b200_fused_dp / --ddp_bucket_mb setting or model "
                        "structure); load it with the settings it was saved with, or load the model only "
                        "(--no_load_optim)")
                ch.main.copy_(sd["shard_fp32_from_float16_groups"][i])
                ch.m.copy_(inner[i]["exp_avg"])
                ch.v.copy_(inner[i]["exp_avg_sq"])
            return
        mains = sd["fp32_from_fp16_params"][0]
        idx = 0
        for ch in self.chunks:
            for p, s, e, _ in ch.layout:
                ch.m[s:e].copy_(inner[idx]["exp_avg"].reshape(-1))
                ch.v[s:e].copy_(inner[idx]["exp_avg_sq"].reshape(-1))
                ch.main[s:e].copy_(mains[idx].reshape(-1))
                idx += 1


# Names kept for users of the reference API.
class Float16OptimizerWithFloat16Params(FlatOptimizer):
    def __init__(self, models, cfg, grad_scaler=None):
        super().__init__(models, cfg, grad_scaler, distributed=False)


class FP32Optimizer(FlatOptimizer):
    def __init__(self, models, cfg):
        super().__init__(models, cfg, None, distributed=False)


class DistributedOptimizer(FlatOptimizer):
    def __init__(self, models, cfg, grad_scaler=None):
        super().__init__(models, cfg, grad_scaler, distributed=True)
