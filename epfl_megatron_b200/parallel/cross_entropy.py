"""Vocab-parallel cross-entropy (megatron/core/tensor_parallel/cross_entropy.py:16-126).

CUDA bf16 path: one fused pass per rank over its [T, V/tp] bf16 shard produces
(max, sum-exp, target-logit); the triples are exchanged with ONE all-gather (the
reference issues three all-reduces on an fp32 copy of the logits), folded by a tiny
kernel, and the backward rewrites the logits buffer in place.
PyTorch path (CPU / fp32 / label smoothing): restatement of the reference math.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from ..ops import lib as L
from ..ops import functional as Fn
from . import state
from .utils import VocabUtility, all_gather_into


class _VocabParallelCrossEntropyRef(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vocab_parallel_logits, target, label_smoothing=0.0):
        grp = state.get_tensor_model_parallel_group() if state.get_tensor_model_parallel_world_size() > 1 else None
        logits = vocab_parallel_logits.float() if vocab_parallel_logits.dtype != torch.float32 \
            else vocab_parallel_logits.clone()
        mx = logits.max(dim=-1)[0]
        if grp is not None:
            dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=grp)
        logits = logits - mx.unsqueeze(-1)
        part = logits.size(-1)
        rank, ws = state.get_tensor_model_parallel_rank(), state.get_tensor_model_parallel_world_size()
        lo, hi = VocabUtility.vocab_range_from_per_partition_vocab_size(part, rank, ws)
        tmask = (target < lo) | (target >= hi)
        mt = (target - lo).masked_fill(tmask, 0)
        l2 = logits.view(-1, part)
        ar = torch.arange(l2.size(0), device=l2.device)
        pred = l2[ar, mt.view(-1)].clone().contiguous().view_as(target)
        pred = pred.masked_fill(tmask, 0.0)
        if grp is not None:
            dist.all_reduce(pred, group=grp)
        ex = torch.exp(logits)
        se = ex.sum(dim=-1)
        if grp is not None:
            dist.all_reduce(se, group=grp)
        loss = torch.log(se) - pred
        ex = ex / se.unsqueeze(-1)
        vocab_size = ex.size(-1)
        if label_smoothing > 0:
            assert 1.0 > label_smoothing > 0.0
            smoothing = label_smoothing * vocab_size / (vocab_size - 1)
            log_probs = torch.log(ex)
            mean_log_probs = log_probs.mean(dim=-1)
            loss = (1.0 - smoothing) * loss - smoothing * mean_log_probs
        ctx.label_smoothing, ctx.vocab_size = label_smoothing, vocab_size
        ctx.in_dtype = vocab_parallel_logits.dtype
        ctx.save_for_backward(ex, tmask, mt.view(-1))
        return loss

    @staticmethod
    def backward(ctx, grad_output):
        softmax, tmask, mt = ctx.saved_tensors
        g = softmax
        part = softmax.size(-1)
        g2 = g.view(-1, part)
        ar = torch.arange(g2.size(0), device=g2.device)
        upd = 1.0 - tmask.view(-1).float()
        if ctx.label_smoothing > 0:
            smoothing = ctx.label_smoothing * ctx.vocab_size / (ctx.vocab_size - 1)
            g2[ar, mt] -= (1.0 - smoothing) * upd
            g2[:, :] -= smoothing / ctx.vocab_size
        else:
            g2[ar, mt] -= upd
        g = g * grad_output.unsqueeze(-1)
        return g.to(ctx.in_dtype), None, None


def _peer_workspace(T):
    """The tensor-parallel group's peer-memory workspace when the fused collectives are in use
    (``--b200_fused_comm`` not off, group of 2..8 GPUs on one host), else None -> NCCL all-gather."""
    from . import layers
    tp = state.get_tensor_model_parallel_world_size()
    if layers._FUSED_COMM == "off" or not (2 <= tp <= 8) or os.environ.get("B200_FUSED_XENT", "1") == "0" \
            or not layers.tp_group_on_one_host():
        return None
    ws = layers._workspace(T, 256, 256)
    return ws if T <= ws.max_rows else None


class _VocabParallelCrossEntropyFused(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        V = logits.shape[-1]
        l2 = logits.reshape(-1, V)
        if not l2.is_contiguous():
            l2 = l2.contiguous()
        T = l2.shape[0]
        tgt = target.reshape(-1).contiguous()
        ws = state.get_tensor_model_parallel_world_size()
        rank = state.get_tensor_model_parallel_rank()
        lo = rank * V
        tpws = _peer_workspace(T) if ws > 1 else None
        if tpws is not None:
            # the statistics kernel stores its triples into every rank's table over NVLink
            loss, lse = tpws.xent_exchange(l2, tgt, lo)
        else:
            stats = torch.empty(T, 3, device=logits.device, dtype=torch.float32)
            L.call("b200_xent_stats", l2.data_ptr(), tgt.data_ptr(), stats.data_ptr(), T, V, lo,
                   L.stream_ptr())
            if ws > 1:
                allst = torch.empty(ws, T, 3, device=logits.device, dtype=torch.float32)
                all_gather_into(allst.view(ws * T, 3), stats, state.get_tensor_model_parallel_group())
            else:
                allst = stats
            loss = torch.empty(T, device=logits.device, dtype=torch.float32)
            lse = torch.empty(T, device=logits.device, dtype=torch.float32)
            L.call("b200_xent_combine", allst.data_ptr(), loss.data_ptr(), lse.data_ptr(), T, ws,
                   L.stream_ptr())
        ctx.save_for_backward(l2, tgt, lse)
        ctx.meta = (logits.shape, lo)
        return loss.view(target.shape)

    @staticmethod
    def backward(ctx, grad_output):
        l2, tgt, lse = ctx.saved_tensors
        shape, lo = ctx.meta
        T, V = l2.shape
        g = grad_output.reshape(-1).float().contiguous()
        # in place on the saved logits buffer: the forward result is not needed again
        L.call("b200_xent_bwd", l2.data_ptr(), tgt.data_ptr(), lse.data_ptr(), g.data_ptr(), T, V,
               lo, L.stream_ptr())
        return l2.view(shape), None


def vocab_parallel_cross_entropy(vocab_parallel_logits, target, label_smoothing=0.0):
    """logits [..., V/tp], target [...] (global ids) -> per-token loss [...] (fp32)."""
    x = vocab_parallel_logits
    if label_smoothing == 0.0 and Fn._k(x) and x.shape[-1] % 8 == 0:
        return _VocabParallelCrossEntropyFused.apply(x, target)
    return _VocabParallelCrossEntropyRef.apply(x, target, label_smoothing)
