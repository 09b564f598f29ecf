"""Pipeline-parallel point-to-point exchange over NCCL p2p
(megatron/p2p_communication.py:11-405 surface: the eight named wrappers + _communicate).

Stays on NCCL send/recv as BASELINE.json prescribes. Unlike the reference there is no
``torch.cuda.synchronize()`` after each exchange (p2p_communication.py:229-231): the
receive requests are waited on the stream (NCCL work objects make the compute stream
wait on the communication stream), so the host keeps running ahead and the next
micro-batch's kernels are already queued behind the transfer.
"""
from __future__ import annotations

import operator
from functools import reduce

import torch
import torch.distributed as dist

from . import global_vars
from .parallel import state as mpu
from .parallel.utils import gather_split_1d_tensor, split_tensor_into_1d_equal_chunks


def _device():
    if dist.get_backend() == "gloo" or not torch.cuda.is_available():
        return torch.device("cpu")
    return torch.device("cuda", torch.cuda.current_device())


def _communicate_shapes(tensor_send_next, tensor_send_prev, recv_prev, recv_next):
    """Exchange 3-element shapes first (variable sequence lengths)."""
    dev = _device()
    recv_prev_shape = torch.empty(3, device=dev, dtype=torch.int64) if recv_prev else None
    recv_next_shape = torch.empty(3, device=dev, dtype=torch.int64) if recv_next else None
    send_prev_shape = torch.tensor(tensor_send_prev.size(), device=dev, dtype=torch.int64) \
        if tensor_send_prev is not None else None
    send_next_shape = torch.tensor(tensor_send_next.size(), device=dev, dtype=torch.int64) \
        if tensor_send_next is not None else None
    ops = []
    if send_prev_shape is not None:
        ops.append(dist.P2POp(dist.isend, send_prev_shape, mpu.get_pipeline_model_parallel_prev_rank()))
    if recv_prev_shape is not None:
        ops.append(dist.P2POp(dist.irecv, recv_prev_shape, mpu.get_pipeline_model_parallel_prev_rank()))
    if send_next_shape is not None:
        ops.append(dist.P2POp(dist.isend, send_next_shape, mpu.get_pipeline_model_parallel_next_rank()))
    if recv_next_shape is not None:
        ops.append(dist.P2POp(dist.irecv, recv_next_shape, mpu.get_pipeline_model_parallel_next_rank()))
    if ops:
        for r in dist.batch_isend_irecv(ops):
            r.wait()
    return (recv_prev_shape.tolist() if recv_prev_shape is not None else [0, 0, 0],
            recv_next_shape.tolist() if recv_next_shape is not None else [0, 0, 0])


def _communicate(tensor_send_next, tensor_send_prev, recv_prev, recv_next, tensor_shape,
                 dtype_=None, args=None):
    """Send to / receive from the neighbouring pipeline stages in one batched call."""
    args = args or global_vars.get_args()
    tensor_recv_prev = tensor_recv_next = None
    if tensor_shape is None:
        # activation shape between stages: [s (/tp under sequence parallelism), micro batch, h]
        s = args.seq_length
        if args.sequence_parallel:
            s //= mpu.get_tensor_model_parallel_world_size()
        tensor_shape = (s, args.micro_batch_size, args.hidden_size)
    if not args.variable_seq_lengths:
        recv_prev_shape = recv_next_shape = tensor_shape
    else:
        recv_prev_shape, recv_next_shape = _communicate_shapes(tensor_send_next, tensor_send_prev,
                                                               recv_prev, recv_next)
    # scatter-gather: ship 1/tp of the activation and all-gather on the receiver (only when
    # the activation is replicated across TP ranks, i.e. without sequence parallelism)
    sg = (args.scatter_gather_tensors_in_pipeline and not args.sequence_parallel
          and mpu.get_tensor_model_parallel_world_size() > 1)
    if sg:
        tp = mpu.get_tensor_model_parallel_world_size()
        n_prev = reduce(operator.mul, recv_prev_shape, 1)
        n_next = reduce(operator.mul, recv_next_shape, 1)
        chunk_prev = (n_prev // tp,) if n_prev % tp == 0 else None
        chunk_next = (n_next // tp,) if n_next % tp == 0 else None
        if chunk_prev is None or chunk_next is None:
            sg = False
    if not sg:
        chunk_prev, chunk_next = recv_prev_shape, recv_next_shape
    dtype = args.params_dtype
    if args.fp32_residual_connection:
        dtype = torch.float
    requires_grad = True
    if dtype_ is not None:
        dtype = dtype_
        requires_grad = False
    dev = _device()
    if recv_prev:
        tensor_recv_prev = torch.empty(chunk_prev, requires_grad=requires_grad, device=dev, dtype=dtype)
    if recv_next:
        tensor_recv_next = torch.empty(chunk_next, requires_grad=requires_grad, device=dev, dtype=dtype)
    if sg:
        if tensor_send_next is not None:
            tensor_send_next = split_tensor_into_1d_equal_chunks(tensor_send_next)
        if tensor_send_prev is not None:
            tensor_send_prev = split_tensor_into_1d_equal_chunks(tensor_send_prev)

    prev_rank = mpu.get_pipeline_model_parallel_prev_rank()
    next_rank = mpu.get_pipeline_model_parallel_next_rank()
    ops = []
    if prev_rank == next_rank:
        # Two stages: the neighbour in both directions is the same rank, and sends / receives
        # between one pair of ranks are matched in posting order. Our "send to prev" (a gradient)
        # is the peer's "receive from next", so the receives are posted next-first. (The reference
        # avoids the case by refusing interleaving for pp == 2, megatron/arguments.py:117-120.)
        if tensor_send_prev is not None:
            ops.append(dist.P2POp(dist.isend, tensor_send_prev.contiguous(), prev_rank))
        if tensor_send_next is not None:
            ops.append(dist.P2POp(dist.isend, tensor_send_next.contiguous(), next_rank))
        if tensor_recv_next is not None:
            ops.append(dist.P2POp(dist.irecv, tensor_recv_next, next_rank))
        if tensor_recv_prev is not None:
            ops.append(dist.P2POp(dist.irecv, tensor_recv_prev, prev_rank))
    else:
        if tensor_send_prev is not None:
            ops.append(dist.P2POp(dist.isend, tensor_send_prev.contiguous(), prev_rank))
        if tensor_recv_prev is not None:
            ops.append(dist.P2POp(dist.irecv, tensor_recv_prev, prev_rank))
        if tensor_send_next is not None:
            ops.append(dist.P2POp(dist.isend, tensor_send_next.contiguous(), next_rank))
        if tensor_recv_next is not None:
            ops.append(dist.P2POp(dist.irecv, tensor_recv_next, next_rank))
    if ops:
        if dist.get_backend() == "gloo" and mpu.get_pipeline_model_parallel_world_size() == 2 \
                and len(ops) > 1:
            # with two stages prev == next; order sends/recvs by rank parity to avoid a
            # head-of-line deadlock on the ordered Gloo pair
            reqs = [op.op(op.tensor, op.peer) for op in ops]
        else:
            reqs = dist.batch_isend_irecv(ops)
        for r in reqs:
            r.wait()   # stream-ordered for NCCL: no device-wide synchronize

    if sg:
        if recv_prev:
            tensor_recv_prev = gather_split_1d_tensor(tensor_recv_prev).view(recv_prev_shape).requires_grad_()
        if recv_next:
            tensor_recv_next = gather_split_1d_tensor(tensor_recv_next).view(recv_next_shape).requires_grad_()
    return tensor_recv_prev, tensor_recv_next


def _timed(name, timers, fn):
    if timers is not None:
        timers(name, log_level=2).start()
    out = fn()
    if timers is not None:
        timers(name).stop()
    return out


def recv_forward(tensor_shape=None, dtype_=None, timers=None):
    if mpu.is_pipeline_first_stage():
        return None
    return _timed("forward-recv", timers, lambda: _communicate(
        None, None, True, False, tensor_shape, dtype_)[0])


def recv_backward(tensor_shape=None, timers=None):
    if mpu.is_pipeline_last_stage():
        return None
    return _timed("backward-recv", timers, lambda: _communicate(
        None, None, False, True, tensor_shape)[1])


def send_forward(output_tensor, tensor_shape=None, dtype_=None, timers=None):
    if not mpu.is_pipeline_last_stage():
        _timed("forward-send", timers, lambda: _communicate(
            output_tensor, None, False, False, tensor_shape, dtype_))


def send_backward(input_tensor_grad, tensor_shape=None, timers=None):
    if not mpu.is_pipeline_first_stage():
        _timed("backward-send", timers, lambda: _communicate(
            None, input_tensor_grad, False, False, tensor_shape))


def send_forward_recv_backward(output_tensor, tensor_shape=None, timers=None):
    if mpu.is_pipeline_last_stage():
        return None
    return _timed("forward-send-backward-recv", timers, lambda: _communicate(
        output_tensor, None, False, True, tensor_shape)[1])


def send_backward_recv_forward(input_tensor_grad, tensor_shape=None, timers=None):
    if mpu.is_pipeline_first_stage():
        return None
    return _timed("backward-send-forward-recv", timers, lambda: _communicate(
        None, input_tensor_grad, True, False, tensor_shape)[0])


def send_forward_recv_forward(output_tensor, recv_prev, tensor_shape=None, timers=None):
    return _timed("forward-send-forward-recv", timers, lambda: _communicate(
        output_tensor, None, recv_prev, False, tensor_shape)[0])


def send_backward_recv_backward(input_tensor_grad, recv_next, tensor_shape=None, timers=None):
    return _timed("backward-send-backward-recv", timers, lambda: _communicate(
        None, input_tensor_grad, False, recv_next, tensor_shape)[1])


def send_forward_backward_recv_forward_backward(output_tensor, input_tensor_grad, recv_prev, recv_next,
                                                tensor_shape=None, timers=None):
    return _timed("forward-backward-send-forward-backward-recv", timers, lambda: _communicate(
        output_tensor, input_tensor_grad, recv_prev, recv_next, tensor_shape))
