"""Tensor-parallel layers: VocabParallelEmbedding, ColumnParallelLinear, RowParallelLinear
and the autograd linear that carries the sequence-parallel collectives.

Public surface of megatron/core/tensor_parallel/layers.py:41-701. What differs inside:

* the weight gradient is accumulated straight into the fp32 ``weight.main_grad`` by the
  tcgen05 GEMM epilogue (reference: apex ``fused_weight_gradient_mlp_cuda``, layers.py:298-305);
* collectives that overlap with GEMMs run on a dedicated communication stream ordered
  with CUDA events (reference: async_op + CUDA_DEVICE_MAX_CONNECTIONS=1, layers.py:255-296);
* the backward of a sequence-parallel ColumnParallelLinear re-gathers its input while the
  dgrad GEMM runs and reduce-scatters dgrad while the wgrad GEMM runs;
* with sequence parallelism the embedding is reduce-scattered along the sequence
  instead of all-reduced and then discarded (reference: layers.py:209 +
  language_model.py:255-256).
"""
from __future__ import annotations

import math
import ctypes
import os
from typing import Callable, Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.nn.parameter import Parameter

from ..ops import fp8
from ..ops import functional as Fn
from . import state
from .mappings import (copy_to_tensor_model_parallel_region, gather_from_tensor_model_parallel_region,
                       reduce_from_tensor_model_parallel_region,
                       reduce_scatter_to_sequence_parallel_region,
                       scatter_to_tensor_model_parallel_region)
from .random import get_cuda_rng_tracker
from .utils import VocabUtility, all_gather_into, divide, reduce_scatter_into

_MODEL_PARALLEL_ATTRIBUTE_DEFAULTS = {"tensor_model_parallel": False, "partition_dim": -1,
                                      "partition_stride": 1}


def param_is_not_tensor_parallel_duplicate(param):
    return (hasattr(param, "tensor_model_parallel") and param.tensor_model_parallel) or \
        state.get_tensor_model_parallel_rank() == 0


def set_tensor_model_parallel_attributes(tensor, is_parallel, dim, stride):
    for a in _MODEL_PARALLEL_ATTRIBUTE_DEFAULTS:
        assert not hasattr(tensor, a)
    tensor.tensor_model_parallel = is_parallel
    tensor.partition_dim = dim
    tensor.partition_stride = stride


def set_defaults_if_not_set_tensor_model_parallel_attributes(tensor):
    for a, v in _MODEL_PARALLEL_ATTRIBUTE_DEFAULTS.items():
        if not hasattr(tensor, a):
            setattr(tensor, a, v)


def copy_tensor_model_parallel_attributes(dst, src):
    for a in _MODEL_PARALLEL_ATTRIBUTE_DEFAULTS:
        if hasattr(src, a):
            setattr(dst, a, getattr(src, a))


def _initialize_affine_weight_gpu(weight, init_method, partition_dim, stride=1):
    """Each TP rank draws its shard from the tensor-parallel RNG stream."""
    set_tensor_model_parallel_attributes(weight, True, partition_dim, stride)
    with get_cuda_rng_tracker().fork():
        init_method(weight)


def _initialize_affine_weight_cpu(weight, output_size, input_size, per_partition_size,
                                  partition_dim, init_method, stride=1, return_master_weight=False,
                                  params_dtype=torch.float32):
    """Build the full weight on the host, keep this rank's (strided) shard: results do
    not depend on the TP degree."""
    set_tensor_model_parallel_attributes(weight, True, partition_dim, stride)
    master = torch.empty(output_size, input_size, dtype=torch.float, requires_grad=False)
    init_method(master)
    master = master.to(dtype=params_dtype)
    per_stride = divide(per_partition_size, stride)
    pieces = torch.split(master, per_stride, dim=partition_dim)
    rank, ws = state.get_tensor_model_parallel_rank(), state.get_tensor_model_parallel_world_size()
    mine = pieces[rank::ws]
    with torch.no_grad():
        torch.cat(mine, dim=partition_dim, out=weight)
    return master if return_master_weight else None


class VocabParallelEmbedding(torch.nn.Module):
    """Embedding table sharded along the vocabulary."""

    def __init__(self, num_embeddings, embedding_dim, *, init_method=torch.nn.init.xavier_normal_,
                 params_dtype=torch.float32, use_cpu_initialization=False,
                 perform_initialization=True):
        super().__init__()
        self.num_embeddings, self.embedding_dim = num_embeddings, embedding_dim
        self.padding_idx = None
        self.tensor_model_parallel_size = state.get_tensor_model_parallel_world_size()
        self.vocab_start_index, self.vocab_end_index = VocabUtility.vocab_range_from_global_vocab_size(
            num_embeddings, state.get_tensor_model_parallel_rank(), self.tensor_model_parallel_size)
        self.num_embeddings_per_partition = self.vocab_end_index - self.vocab_start_index
        dev = "cpu" if use_cpu_initialization or not torch.cuda.is_available() else torch.cuda.current_device()
        self.weight = Parameter(torch.empty(self.num_embeddings_per_partition, embedding_dim,
                                            device=dev, dtype=params_dtype))
        if perform_initialization:
            if use_cpu_initialization or not torch.cuda.is_available():
                _initialize_affine_weight_cpu(self.weight, num_embeddings, embedding_dim,
                                              self.num_embeddings_per_partition, 0, init_method,
                                              params_dtype=params_dtype)
            else:
                _initialize_affine_weight_gpu(self.weight, init_method, partition_dim=0, stride=1)
        else:
            set_tensor_model_parallel_attributes(self.weight, True, 0, 1)

    def local_lookup(self, input_):
        """Partial embedding: rows owned by other ranks come out as zeros."""
        if self.tensor_model_parallel_size > 1:
            mask = (input_ < self.vocab_start_index) | (input_ >= self.vocab_end_index)
            idx = (input_ - self.vocab_start_index).masked_fill(mask, 0)
            out = F.embedding(idx, self.weight)
            return out.masked_fill(mask.unsqueeze(-1), 0.0)
        return F.embedding(input_, self.weight)

    def forward(self, input_):
        return reduce_from_tensor_model_parallel_region(self.local_lookup(input_))


_COMM_STREAM = None


def comm_stream():
    """Side stream for collectives that overlap with GEMMs."""
    global _COMM_STREAM
    if _COMM_STREAM is None:
        _COMM_STREAM = torch.cuda.Stream()
    return _COMM_STREAM


class _Overlap:
    """Run a collective on the comm stream after the current stream's work so far; the
    caller later ``wait()``s before consuming the result. On CPU it runs inline."""

    def __init__(self, fn, *tensors):
        self.cuda = tensors[0].is_cuda
        if self.cuda:
            cs = comm_stream()
            cs.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(cs):
                fn()
            for t in tensors:
                t.record_stream(cs)
            self.event = cs.record_event()
        else:
            fn()

    def wait(self):
        if self.cuda:
            torch.cuda.current_stream().wait_event(self.event)


class _FusedVocabEmbedding(torch.autograd.Function):
    """Vocab-parallel word embedding without a reduction (csrc/embed.cu): the forward copies every
    output row from the table shard of the rank that owns the token (peer loads over NVLink), directly
    in [s, b, h] order and, under sequence parallelism, only for this rank's part of the sequence; the
    backward adds the output-gradient rows of the tokens this rank owns into the fp32 ``main_grad``
    (rows of other sequence parts are read from the owners' staging buffers). Replaces the masked
    lookup + all-reduce of megatron/core/tensor_parallel/layers.py:187-210 and the all-gather +
    embedding_dense_backward + main_grad.add_ of its backward."""

    @staticmethod
    def forward(ctx, input_ids, weight, sequence_parallel):
        tp = state.get_tensor_model_parallel_world_size()
        rank = state.get_tensor_model_parallel_rank()
        B, S = input_ids.shape
        Vp, H = weight.shape
        ids = input_ids.contiguous()
        sequence_parallel = sequence_parallel and tp > 1
        s_out = S // tp if sequence_parallel else S
        s_off = rank * s_out if sequence_parallel else 0
        out = torch.empty(s_out, B, H, device=weight.device, dtype=weight.dtype)
        if tp > 1:
            ws = _workspace(S * B, H, H)
            tabs = ws.peer_tables(weight)
            ws.tables_barrier()    # the peers' optimizer steps that wrote their shards are complete
        else:
            tabs = [weight.data_ptr()]
        arr = (ctypes.c_void_p * 8)(*tabs)
        Fn.L.call("b200_embed_fwd", ids.data_ptr(), arr, out.data_ptr(), S, B, s_off, s_out, H, Vp, tp,
                  Fn.L.stream_ptr())
        if tp > 1 and ws.verify:
            # B200_COMM_VERIFY=1: masked local lookup + NCCL all-reduce, the reference's formulation
            lo = rank * Vp
            mask = (ids < lo) | (ids >= lo + Vp)
            part = F.embedding((ids - lo).masked_fill(mask, 0), weight).masked_fill(mask.unsqueeze(-1), 0.0)
            dist.all_reduce(part, group=state.get_tensor_model_parallel_group())
            want = part.transpose(0, 1)[s_off:s_off + s_out]
            if not torch.equal(out, want):
                raise RuntimeError(f"peer-memory embedding (rank {rank}): {(out != want).sum().item()} elements "
                                   "differ from lookup + all-reduce")
        ctx.save_for_backward(ids)
        ctx.weight_ref = weight
        ctx.sp = sequence_parallel
        return out

    @staticmethod
    def backward(ctx, dout):
        (ids,) = ctx.saved_tensors
        w = ctx.weight_ref
        tp = state.get_tensor_model_parallel_world_size()
        rank = state.get_tensor_model_parallel_rank()
        B, S = ids.shape
        Vp, H = w.shape
        g = dout.reshape(-1, H)
        if not g.is_contiguous():
            g = g.contiguous()
        main = w.main_grad
        if getattr(w, "_b200_main_grad_empty", False):
            main.zero_()
            w._b200_main_grad_empty = False
        st = Fn.L.stream_ptr()
        if ctx.sp:
            ws = _workspace(S * B, H, H)
            e, srcs = ws.gather_begin(g)
            arr = (ctypes.c_void_p * 8)(*srcs)
            Fn.L.call("b200_embed_bwd", ids.data_ptr(), arr, tp, main.data_ptr(), S, B, H, rank * Vp, Vp, st)
            ws.gather_end(e)
        else:
            arr = (ctypes.c_void_p * 8)(g.data_ptr())
            Fn.L.call("b200_embed_bwd", ids.data_ptr(), arr, 1, main.data_ptr(), S, B, H, rank * Vp, Vp, st)
        Fn._notify_grad_ready(w)
        return None, None, None


def fused_vocab_embedding_ok(input_ids, weight, sequence_parallel) -> bool:
    """Own embedding kernels: bf16 table with an fp32 main_grad (local DDP buffers); with tensor
    parallelism a group of 2..8 GPUs on one host with equal vocabulary shards (peer memory). On one GPU
    the same kernels copy the rows in [s, b, h] order and scatter-add the gradient into main_grad."""
    tp = state.get_tensor_model_parallel_world_size()
    if tp > 8 or (tp > 1 and _FUSED_COMM == "off") or not Fn.kernels_enabled() \
            or os.environ.get("B200_FUSED_EMBED", "1") == "0":      # A/B switch
        return False
    main = getattr(weight, "main_grad", None)
    ok = (weight.is_cuda and weight.dtype == torch.bfloat16 and weight.shape[1] % 8 == 0 and input_ids.dim() == 2
          and input_ids.dtype == torch.int64 and main is not None and main.dtype == torch.float32
          and getattr(weight, "_b200_fuse_wgrad", False) and weight.is_contiguous() and main.is_contiguous()
          and (not sequence_parallel or input_ids.shape[1] % tp == 0))
    if ok and sequence_parallel and tp > 1:
        B, S = input_ids.shape
        ok = (S // tp) * B * weight.shape[1] * 2 <= _workspace(S * B, weight.shape[1], weight.shape[1]).stage_bytes
    if not ok:
        return False
    if tp == 1:
        return True
    if not tp_group_on_one_host():
        return False
    # the tables must be mappable into each other (CUDA IPC): collective on first use, cached afterwards
    B, S = input_ids.shape
    return _workspace(S * B, weight.shape[1], weight.shape[1]).peer_tables(weight) is not None


class LinearWithGradAccumulationAndAsyncCommunication(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input, weight, bias, gradient_accumulation_fusion, async_grad_allreduce,
                sequence_parallel):
        ctx.save_for_backward(input, weight)
        ctx.use_bias = bias is not None
        ctx.gradient_accumulation_fusion = gradient_accumulation_fusion
        ctx.async_grad_allreduce = async_grad_allreduce
        ctx.sequence_parallel = sequence_parallel
        ctx.weight_ref = weight
        if sequence_parallel:
            ws = state.get_tensor_model_parallel_world_size()
            shape = list(input.shape)
            shape[0] *= ws
            total = state.get_global_memory_buffer().get_tensor(shape, input.dtype, "mpu")
            all_gather_into(total, input.contiguous(), state.get_tensor_model_parallel_group())
        else:
            total = input
        K = total.shape[-1]
        t2 = total.reshape(-1, K)
        if fp8.usable(t2, weight):
            out = fp8.forward_gemm(t2, weight)            # E4M3 x E4M3, delayed scaling (ops/fp8.py)
        else:
            out = Fn.gemm_nt(t2, weight)
        out = out.view(*total.shape[:-1], weight.shape[0])
        if bias is not None:
            out = out + bias
        return out

    @staticmethod
    def backward(ctx, grad_output):
        input, weight = ctx.saved_tensors
        w = ctx.weight_ref
        grp = state.get_tensor_model_parallel_group() if (ctx.sequence_parallel or ctx.async_grad_allreduce) else None
        gather = None
        if ctx.sequence_parallel:
            ws = state.get_tensor_model_parallel_world_size()
            shape = list(input.shape)
            shape[0] *= ws
            total = state.get_global_memory_buffer().get_tensor(shape, input.dtype, "mpu")
            inp_c = input.contiguous()
            gather = _Overlap(lambda: all_gather_into(total, inp_c, grp), total, inp_c)
        else:
            total = input
        N = grad_output.shape[-1]
        go2 = grad_output.reshape(-1, N)
        if not go2.is_contiguous():
            go2 = go2.contiguous()
        dyT8 = None
        if fp8.usable(go2, weight):
            grad_input, dyT8 = fp8.dgrad_gemm(go2, weight)
            grad_input = grad_input.view(*grad_output.shape[:-1], weight.shape[1])
        else:
            grad_input = Fn.gemm_nn(go2, weight).view(*grad_output.shape[:-1], weight.shape[1])
        if gather is not None:
            gather.wait()
        red = None
        if ctx.async_grad_allreduce:
            red = _Overlap(lambda: dist.all_reduce(grad_input, group=grp), grad_input)
        sub_grad_input = None
        if ctx.sequence_parallel:
            assert not ctx.async_grad_allreduce
            sub_grad_input = torch.empty(input.shape, dtype=input.dtype, device=input.device)
            red = _Overlap(lambda: reduce_scatter_into(sub_grad_input, grad_input, grp),
                           sub_grad_input, grad_input)
        t2 = total.reshape(-1, total.shape[-1])
        main = getattr(w, "main_grad", None)
        grad_weight = None
        fused = False
        if ctx.gradient_accumulation_fusion and main is not None:
            first = getattr(w, "_b200_main_grad_empty", False)
            if dyT8 is not None and main.dtype == torch.float32 and main.is_contiguous() and fp8.usable(t2, weight):
                fp8.wgrad_gemm(dyT8, t2, weight, main, accumulate=not first)
                fused = True
            else:
                fused = Fn.wgrad_accumulate(go2, t2, main, accumulate=not first)
            if fused:
                w._b200_main_grad_empty = False
                Fn._notify_grad_ready(w)
        if not fused:
            grad_weight = torch.matmul(go2.t(), t2)
        grad_bias = go2.sum(dim=0) if ctx.use_bias else None
        if red is not None:
            red.wait()
        if ctx.sequence_parallel:
            return sub_grad_input, grad_weight, grad_bias, None, None, None
        return grad_input, grad_weight, grad_bias, None, None, None


# ---------------------------------------------------------------------------------------------
# Sequence-parallel linears with the collective inside the GEMM kernel (parallel/symm.py,
# csrc/gemm_comm_tcgen05.cu). Same math as the class above + the mappings around it.
# ---------------------------------------------------------------------------------------------
_FUSED_COMM = "auto"    # auto | on | off  (--b200_fused_comm)


# Per-op policy of "auto": the in-kernel all-gather is used up to this tensor-parallel size. At tp 8
# the transfer (7/8 of the activation over NVLink) cannot hide behind 1/8 of the GEMM and NCCL
# all-gather + GEMM is faster (0.119 vs 0.19 ms on the Falcon-40B shapes).
# "--b200_fused_comm on" forces every fused path.
_FUSED_AG_MAX_TP = int(os.environ.get("B200_FUSED_AG_MAX_TP", "4"))
# GEMM -> all-reduce in one kernel (no sequence parallelism) wins at tp 2 (0.127 vs 0.135 ms on
# 4096 x 4096 x 2048) and loses to cuBLAS + NCCL's NVLS all-reduce at tp 4 / 8 on the 7B-sized shards
# (profiles/fused_comm_n4_r2_call15.log, fused_comm_n8_r2_call14.log).
_FUSED_AR_MAX_TP = int(os.environ.get("B200_FUSED_AR_MAX_TP", "2"))
# GEMM -> reduce-scatter: in-kernel up to tp 4; at tp 8 the 2048-row activations of Falcon-40B leave each
# rank 1/8 of an already small GEMM to hide 7/8 of the output behind, and cuBLAS + NCCL is faster (0.128 vs
# 0.204 ms; Falcon-40B TP8 step 1035 vs 1133 ms, profiles/bench_r2_call18_falcon40b_tp8_*.log).
_FUSED_RS_MAX_TP = int(os.environ.get("B200_FUSED_RS_MAX_TP", "4"))


def _fused_all_gather(tp: int) -> bool:
    return _FUSED_COMM == "on" or tp <= _FUSED_AG_MAX_TP


def _fused_reduce_scatter(tp: int) -> bool:
    return _FUSED_COMM == "on" or tp <= _FUSED_RS_MAX_TP


def set_fused_comm_mode(mode: str):
    global _FUSED_COMM
    assert mode in ("auto", "on", "off")
    _FUSED_COMM = mode


_TP_ONE_HOST = None


def tp_group_on_one_host() -> bool:
    """Peer memory (CUDA IPC) only reaches the GPUs of this host. Collective over the tensor-parallel
    group on first use (every rank gets here in the same forward pass), cached afterwards."""
    global _TP_ONE_HOST
    if _TP_ONE_HOST is None:
        from .dp_symm import group_on_one_host
        _TP_ONE_HOST = group_on_one_host(state.get_tensor_model_parallel_group())
    return _TP_ONE_HOST


def _fused_comm_ok(rows_local: int, k_gather: int, n_scatter: int, *tensors) -> bool:
    """rows_local: rows of the sequence shard; k_gather: width of the gathered operand;
    n_scatter: width of the reduce-scattered result."""
    if _FUSED_COMM == "off" or not Fn.kernels_enabled():
        return False
    ok = (all(t.is_cuda and t.dtype == torch.bfloat16 for t in tensors)
          and not any(hasattr(t, "_b200_fp8") for t in tensors)      # FP8 layers: NCCL collectives + 8-bit GEMMs
          and rows_local % 256 == 0 and k_gather % 256 == 0 and n_scatter % 8 == 0
          and 2 <= state.get_tensor_model_parallel_world_size() <= 8)
    ok = ok and tp_group_on_one_host()
    if not ok and _FUSED_COMM == "on":
        raise RuntimeError(f"--b200_fused_comm on: unsupported shape rows/rank {rows_local}, "
                           f"gathered width {k_gather}, scattered width {n_scatter}")
    return ok


def _workspace(rows_total: int, k_gather: int, n_scatter: int, allreduce: bool = False):
    """The group's workspace, sized once for the whole model when the global args are known
    (every gathered / scattered activation of a decoder is [s*b, hidden])."""
    from .symm import get_workspace
    try:
        from .. import global_vars
        a = global_vars.get_args()
        rows_total = max(rows_total, a.seq_length * a.micro_batch_size)
        k_gather = n_scatter = max(k_gather, n_scatter, a.hidden_size)
        # tensor parallelism without sequence parallelism: every reduction is an all-reduce
        allreduce = allreduce or (a.tensor_model_parallel_size > 1 and not a.sequence_parallel)
    except Exception:  # noqa: BLE001 - used outside a configured run (tests, tools)
        pass
    return get_workspace(state.get_tensor_model_parallel_group(), rows_total, k_gather, n_scatter, allreduce)


def _accumulate_wgrad(w, go2, t2, gradient_accumulation_fusion):
    """-> grad_weight or None when it went straight into w.main_grad."""
    main = getattr(w, "main_grad", None)
    if gradient_accumulation_fusion and main is not None:
        first = getattr(w, "_b200_main_grad_empty", False)
        if Fn.wgrad_accumulate(go2, t2, main, accumulate=not first):
            w._b200_main_grad_empty = False
            Fn._notify_grad_ready(w)
            return None
    return torch.matmul(go2.t(), t2)


class _FusedColumnSP(torch.autograd.Function):
    """[s/tp, b, h] -> all-gather along s (in kernel) -> x W^T. Backward: dgrad GEMM with the
    reduce-scatter in its epilogue; wgrad from the gathered input kept from forward (the
    reference gathers it a second time, layers.py:255-270, to save the memory)."""

    @staticmethod
    def forward(ctx, input, weight, gradient_accumulation_fusion):
        s_local, b, K = input.shape
        x2 = input.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        tp = state.get_tensor_model_parallel_world_size()
        ws = _workspace(x2.shape[0] * tp, K, K)
        if _fused_all_gather(tp):
            out, gathered = ws.allgather_gemm(x2, weight)
        else:
            gathered = torch.empty(x2.shape[0] * tp, K, dtype=x2.dtype, device=x2.device)
            all_gather_into(gathered, x2, state.get_tensor_model_parallel_group())
            out = Fn.gemm_nt(gathered, weight)
        ctx.save_for_backward(gathered, weight)
        ctx.weight_ref = weight
        ctx.gaf = gradient_accumulation_fusion
        ctx.in_shape = input.shape
        return out.view(s_local * tp, b, weight.shape[0])

    @staticmethod
    def backward(ctx, grad_output):
        gathered, weight = ctx.saved_tensors
        N = grad_output.shape[-1]
        go2 = grad_output.reshape(-1, N)
        if not go2.is_contiguous():
            go2 = go2.contiguous()
        K = weight.shape[1]
        ws = _workspace(go2.shape[0], K, K)
        tp = state.get_tensor_model_parallel_world_size()
        if _fused_reduce_scatter(tp):
            dx = ws.gemm_reduce_scatter(go2, weight, b_mn_major=True)    # [M/tp, K]
        else:
            full = Fn.gemm_nn(go2, weight)
            dx = torch.empty(go2.shape[0] // tp, K, dtype=go2.dtype, device=go2.device)
            reduce_scatter_into(dx, full, state.get_tensor_model_parallel_group())
        gw = _accumulate_wgrad(ctx.weight_ref, go2, gathered, ctx.gaf)
        return dx.view(ctx.in_shape), gw, None


class _FusedRowSP(torch.autograd.Function):
    """[s, b, k/tp] -> x W^T with the reduce-scatter along s in the GEMM epilogue -> [s/tp, b, h].
    Backward: the all-gather of dY feeds the dgrad GEMM from inside the kernel and the gathered
    dY is reused for the wgrad."""

    @staticmethod
    def forward(ctx, input, weight, gradient_accumulation_fusion):
        s, b, K = input.shape
        x2 = input.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        tp = state.get_tensor_model_parallel_world_size()
        N = weight.shape[0]
        ws = _workspace(x2.shape[0], N, N)
        if _fused_reduce_scatter(tp):
            out = ws.gemm_reduce_scatter(x2, weight)                      # [M/tp, N]
        else:
            full = Fn.gemm_nt(x2, weight)
            out = torch.empty(x2.shape[0] // tp, N, dtype=x2.dtype, device=x2.device)
            reduce_scatter_into(out, full, state.get_tensor_model_parallel_group())
        ctx.save_for_backward(x2, weight)
        ctx.weight_ref = weight
        ctx.gaf = gradient_accumulation_fusion
        ctx.in_shape = input.shape
        return out.view(s // tp, b, N)

    @staticmethod
    def backward(ctx, grad_output):
        x2, weight = ctx.saved_tensors
        N = grad_output.shape[-1]
        dy2 = grad_output.reshape(-1, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        ws = _workspace(x2.shape[0], N, N)
        if _fused_all_gather(state.get_tensor_model_parallel_world_size()):
            dx, gathered_dy = ws.allgather_gemm(dy2, weight, b_mn_major=True)  # [M, K/tp], [M, N]
        else:
            gathered_dy = torch.empty(x2.shape[0], N, dtype=dy2.dtype, device=dy2.device)
            all_gather_into(gathered_dy, dy2, state.get_tensor_model_parallel_group())
            dx = Fn.gemm_nn(gathered_dy, weight)
        gw = _accumulate_wgrad(ctx.weight_ref, gathered_dy, x2, ctx.gaf)
        return dx.view(ctx.in_shape), gw, None


class _FusedRowAR(torch.autograd.Function):
    """Tensor parallelism WITHOUT sequence parallelism: [s, b, k/tp] -> x W^T with the all-reduce in the
    GEMM kernel (the owner of a row chunk reduces it in its epilogue and stores the final tiles into
    every rank's output) -> [s, b, h], the same bits on all ranks. Replaces matmul + NCCL all-reduce
    (megatron/core/tensor_parallel/layers.py:694, mappings.py:13-23). Backward is local: dY is
    replicated."""

    @staticmethod
    def forward(ctx, input, weight, gradient_accumulation_fusion):
        K = input.shape[-1]
        x2 = input.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        N = weight.shape[0]
        ws = _workspace(x2.shape[0], N, N, allreduce=True)
        out = ws.gemm_allreduce(x2, weight)                               # [M, N]
        ctx.save_for_backward(x2, weight)
        ctx.weight_ref = weight
        ctx.gaf = gradient_accumulation_fusion
        ctx.in_shape = input.shape
        return out.view(*input.shape[:-1], N)

    @staticmethod
    def backward(ctx, grad_output):
        x2, weight = ctx.saved_tensors
        N = grad_output.shape[-1]
        dy2 = grad_output.reshape(-1, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = Fn.gemm_nn(dy2, weight)
        gw = _accumulate_wgrad(ctx.weight_ref, dy2, x2, ctx.gaf)
        return dx.view(ctx.in_shape), gw, None


class _FusedColumnAR(torch.autograd.Function):
    """ColumnParallelLinear without sequence parallelism: the forward is a local GEMM on the
    replicated input; the backward's dgrad needs the sum over ranks (layers.py:278-283, or the
    backward of copy_to_tensor_model_parallel_region, mappings.py:139-140): dgrad GEMM + all-reduce in
    one kernel."""

    @staticmethod
    def forward(ctx, input, weight, gradient_accumulation_fusion):
        K = input.shape[-1]
        x2 = input.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        out = Fn.gemm_nt(x2, weight)
        ctx.save_for_backward(x2, weight)
        ctx.weight_ref = weight
        ctx.gaf = gradient_accumulation_fusion
        ctx.in_shape = input.shape
        return out.view(*input.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, grad_output):
        x2, weight = ctx.saved_tensors
        N = grad_output.shape[-1]
        go2 = grad_output.reshape(-1, N)
        if not go2.is_contiguous():
            go2 = go2.contiguous()
        K = weight.shape[1]
        ws = _workspace(go2.shape[0], K, K, allreduce=True)
        dx = ws.gemm_allreduce(go2, weight, b_mn_major=True)               # [M, K], summed over ranks
        gw = _accumulate_wgrad(ctx.weight_ref, go2, x2, ctx.gaf)
        return dx.view(ctx.in_shape), gw, None


def _fused_allreduce_ok(x, weight, n_out) -> bool:
    """Plain tensor parallelism (no sequence parallelism): rows must split into 256-row chunks per rank."""
    if x.dim() < 2 or state.get_tensor_model_parallel_world_size() < 2:
        return False
    rows = x.numel() // x.shape[-1]
    tp = state.get_tensor_model_parallel_world_size()
    if _FUSED_COMM != "on" and tp > _FUSED_AR_MAX_TP:
        return False
    return rows % tp == 0 and _fused_comm_ok(rows // tp, 256, n_out, x, weight)


def _maybe_attach_fp8(weight):
    """First forward of a Column / RowParallelLinear: give its weight an FP8 scaling state when the run
    asks for FP8 (``--fp8_e4m3`` / ``--fp8_hybrid``); by now the parameter is bf16 on the GPU."""
    if hasattr(weight, "_b200_fp8") or getattr(weight, "_b200_fp8_checked", False):
        return
    weight._b200_fp8_checked = True
    try:
        from .. import global_vars
        args = global_vars.get_args()
    except Exception:  # noqa: BLE001 - layers used outside a configured run
        return
    fp8.attach(weight, args)


def linear_with_grad_accumulation_and_async_allreduce(input, weight, bias, gradient_accumulation_fusion,
                                                      async_grad_allreduce,
                                                      sequence_parallel_enabled):
    if sequence_parallel_enabled and input.dim() == 3 and \
            _fused_comm_ok(input.shape[0] * input.shape[1], input.shape[2], input.shape[2], input, weight):
        out = _FusedColumnSP.apply(input, weight, gradient_accumulation_fusion)
        return out + bias if bias is not None else out
    return LinearWithGradAccumulationAndAsyncCommunication.apply(
        input, weight, bias, gradient_accumulation_fusion, async_grad_allreduce,
        sequence_parallel_enabled)


class ColumnParallelLinear(torch.nn.Module):
    """Y = X A^T + b with A sharded along its output rows: A = [A_1; ...; A_p]."""

    def __init__(self, input_size, output_size, *, bias=True, gather_output=True,
                 init_method=torch.nn.init.xavier_normal_, stride=1, keep_master_weight_for_test=False,
                 skip_bias_add=False, async_tensor_model_parallel_allreduce=True,
                 params_dtype=torch.float32, use_cpu_initialization=False, perform_initialization=True,
                 gradient_accumulation_fusion=False, sequence_parallel_enabled: bool = False,
                 world_size: Optional[int] = None):
        super().__init__()
        self.input_size, self.output_size = input_size, output_size
        self.gather_output, self.skip_bias_add = gather_output, skip_bias_add
        ws = world_size if world_size is not None else state.get_tensor_model_parallel_world_size()
        self.world_size = ws
        self.output_size_per_partition = divide(output_size, ws)
        cpu = use_cpu_initialization or not torch.cuda.is_available()
        dev = "cpu" if cpu else torch.cuda.current_device()
        self.weight = Parameter(torch.empty(self.output_size_per_partition, input_size, device=dev,
                                            dtype=params_dtype))
        self.master_weight = None
        if perform_initialization:
            if cpu:
                self.master_weight = _initialize_affine_weight_cpu(
                    self.weight, output_size, input_size, self.output_size_per_partition, 0,
                    init_method, stride=stride, return_master_weight=keep_master_weight_for_test,
                    params_dtype=params_dtype)
            else:
                _initialize_affine_weight_gpu(self.weight, init_method, partition_dim=0, stride=stride)
        else:
            set_tensor_model_parallel_attributes(self.weight, True, 0, stride)
        if bias:
            self.bias = Parameter(torch.zeros(self.output_size_per_partition, device=dev, dtype=params_dtype))
            set_tensor_model_parallel_attributes(self.bias, True, 0, stride)
        else:
            self.register_parameter("bias", None)
        self.async_tensor_model_parallel_allreduce = async_tensor_model_parallel_allreduce and ws > 1
        self.sequence_parallel_enabled = sequence_parallel_enabled and ws > 1
        if self.async_tensor_model_parallel_allreduce and self.sequence_parallel_enabled:
            raise RuntimeError("`async_tensor_model_parallel_allreduce` and `sequence_parallel_enabled` "
                               "cannot be enabled at the same time.")
        self.gradient_accumulation_fusion = gradient_accumulation_fusion

    def forward(self, input_):
        bias = self.bias if not self.skip_bias_add else None
        _maybe_attach_fp8(self.weight)
        if self.world_size > 1 and not self.sequence_parallel_enabled and torch.is_grad_enabled() \
                and input_.requires_grad and _fused_allreduce_ok(input_, self.weight, self.input_size):
            # dgrad GEMM + all-reduce in one kernel in the backward pass
            out = _FusedColumnAR.apply(input_, self.weight, self.gradient_accumulation_fusion)
            if bias is not None:
                out = out + bias
            if self.gather_output:
                out = gather_from_tensor_model_parallel_region(out)
            return out, (self.bias if self.skip_bias_add else None)
        if self.async_tensor_model_parallel_allreduce or self.sequence_parallel_enabled or self.world_size == 1:
            x = input_
        else:
            x = copy_to_tensor_model_parallel_region(input_)
        out = linear_with_grad_accumulation_and_async_allreduce(
            x, self.weight, bias, self.gradient_accumulation_fusion,
            self.async_tensor_model_parallel_allreduce, self.sequence_parallel_enabled)
        if self.gather_output and self.world_size > 1:
            assert not self.sequence_parallel_enabled
            out = gather_from_tensor_model_parallel_region(out)
        return out, (self.bias if self.skip_bias_add else None)


class RowParallelLinear(torch.nn.Module):
    """Y = X A^T + b with A sharded along its input columns: A = [A_1, ..., A_p]."""

    def __init__(self, input_size, output_size, *, bias=True, input_is_parallel=False,
                 init_method=torch.nn.init.xavier_normal_, stride=1, keep_master_weight_for_test=False,
                 skip_bias_add=False, params_dtype=torch.float32, use_cpu_initialization=False,
                 perform_initialization=True, gradient_accumulation_fusion=False,
                 sequence_parallel_enabled: bool = False, world_size: Optional[int] = None):
        super().__init__()
        self.input_size, self.output_size = input_size, output_size
        self.input_is_parallel, self.skip_bias_add = input_is_parallel, skip_bias_add
        ws = world_size if world_size is not None else state.get_tensor_model_parallel_world_size()
        self.world_size = ws
        self.input_size_per_partition = divide(input_size, ws)
        self.gradient_accumulation_fusion = gradient_accumulation_fusion
        self.sequence_parallel_enabled = sequence_parallel_enabled and ws > 1
        if sequence_parallel_enabled and not input_is_parallel:
            raise RuntimeError("To enable `sequence_parallel_enabled`, `input_is_parallel` must be `True`")
        cpu = use_cpu_initialization or not torch.cuda.is_available()
        dev = "cpu" if cpu else torch.cuda.current_device()
        self.weight = Parameter(torch.empty(output_size, self.input_size_per_partition, device=dev,
                                            dtype=params_dtype))
        self.master_weight = None
        if perform_initialization:
            if cpu:
                self.master_weight = _initialize_affine_weight_cpu(
                    self.weight, output_size, input_size, self.input_size_per_partition, 1,
                    init_method, stride=stride, return_master_weight=keep_master_weight_for_test,
                    params_dtype=params_dtype)
            else:
                _initialize_affine_weight_gpu(self.weight, init_method, partition_dim=1, stride=stride)
        else:
            set_tensor_model_parallel_attributes(self.weight, True, 1, stride)
        if bias:
            self.bias = Parameter(torch.zeros(output_size, device=dev, dtype=params_dtype))
            setattr(self.bias, "sequence_parallel", self.sequence_parallel_enabled)
        else:
            self.register_parameter("bias", None)

    def forward(self, input_):
        _maybe_attach_fp8(self.weight)
        if self.input_is_parallel or self.world_size == 1:
            x = input_
        else:
            assert not self.sequence_parallel_enabled
            x = scatter_to_tensor_model_parallel_region(input_)
        if self.sequence_parallel_enabled and x.dim() == 3 and x.shape[0] % self.world_size == 0 and \
                _fused_comm_ok(x.shape[0] * x.shape[1] // self.world_size, self.output_size, self.output_size,
                               x, self.weight):
            out_ = _FusedRowSP.apply(x, self.weight, self.gradient_accumulation_fusion)
            if not self.skip_bias_add:
                return (out_ + self.bias if self.bias is not None else out_), None
            return out_, self.bias
        if self.world_size > 1 and not self.sequence_parallel_enabled and \
                _fused_allreduce_ok(x, self.weight, self.output_size):
            out_ = _FusedRowAR.apply(x, self.weight, self.gradient_accumulation_fusion)
            if not self.skip_bias_add:
                return (out_ + self.bias if self.bias is not None else out_), None
            return out_, self.bias
        out_parallel = linear_with_grad_accumulation_and_async_allreduce(
            x, self.weight, None, self.gradient_accumulation_fusion, False, False)
        if self.sequence_parallel_enabled:
            out_ = reduce_scatter_to_sequence_parallel_region(out_parallel)
        elif self.world_size > 1:
            out_ = reduce_from_tensor_model_parallel_region(out_parallel)
        else:
            out_ = out_parallel
        if not self.skip_bias_add:
            return (out_ + self.bias if self.bias is not None else out_), None
        return out_, self.bias
