"""Forward/backward schedules: no pipelining (gradient accumulation), 1F1B and interleaved
1F1B with virtual stages (megatron/schedules.py:18-722 surface).

All three share ``forward_step`` / ``backward_step``. After a stage has sent its output
downstream, the output's storage is released (only its autograd graph is needed) and the
backward is started through the engine directly, which skips the shape check -- the
memory trick of schedules.py:36-88. Interleaving is allowed for pp >= 2 (the reference
refuses pp == 2).
"""
from __future__ import annotations

from contextlib import contextmanager

import torch
from torch.autograd.variable import Variable
from torch.nn.parallel.distributed import DistributedDataParallel as torchDDP

from . import global_vars, p2p_communication
from .models.distributed import DistributedDataParallelBase as LocalDDP
from .models.enums import ModelType
from .models.module import Float16Module
from .parallel import state as mpu


def unwrap_model(model, module_instances=(torchDDP,)):
    as_list = isinstance(model, list)
    models = model if as_list else [model]
    out = []
    for m in models:
        while isinstance(m, module_instances):
            m = m.module
        out.append(m)
    return out if as_list else out[0]


def get_forward_backward_func(args=None):
    args = args or global_vars.get_args()
    if mpu.get_pipeline_model_parallel_world_size() > 1:
        if args.virtual_pipeline_model_parallel_size is not None:
            assert global_vars.get_num_microbatches() % args.pipeline_model_parallel_size == 0, \
                "number of microbatches is not divisible by pipeline-parallel size when using interleaved schedule"
            return forward_backward_pipelining_with_interleaving
        return forward_backward_pipelining_without_interleaving
    return forward_backward_no_pipelining


def deallocate_output_tensor(out):
    """Keep the graph, drop the storage (the data has already been sent downstream)."""
    if out is None:
        return
    assert isinstance(out, torch.Tensor), f"expected Tensor, found {type(out).__name__}."
    assert out._base is None, "counter-productive to free a view of another tensor."
    out.data = torch.empty((1,), device=out.device, dtype=out.dtype)


def custom_backward(output, grad_output):
    """torch.autograd.backward without the output/grad shape check."""
    assert output.numel() == 1, "output should be pseudo-'freed' in schedule, to optimize memory"
    assert isinstance(output, torch.Tensor), f"output == '{type(output).__name__}'."
    assert isinstance(grad_output, (torch.Tensor, type(None))), f"grad_output == '{type(grad_output).__name__}'."
    if grad_output is None:
        assert output.numel() == 1, "implicit grad requires scalar output."
        grad_output = torch.ones_like(output, memory_format=torch.preserve_format)
    Variable._execution_engine.run_backward(
        tensors=(output,), grad_tensors=(grad_output,), keep_graph=False, create_graph=False,
        inputs=tuple(), allow_unreachable=True, accumulate_grad=True)


def forward_step(forward_step_func, data_iterator, model, input_tensor, forward_data_store, timers,
                 collect_non_loss_data=False, args=None):
    args = args or global_vars.get_args()
    if timers is not None:
        timers("forward-compute", log_level=2).start()
    unwrapped = unwrap_model(model, (torchDDP, LocalDDP, Float16Module))
    unwrap_output = False
    if not isinstance(input_tensor, list):
        input_tensor = [input_tensor]
        unwrap_output = True
    unwrapped.set_input_tensor(input_tensor)
    output_tensor, loss_func = forward_step_func(data_iterator, model)
    if mpu.is_pipeline_last_stage():
        if not collect_non_loss_data:
            output_tensor = loss_func(output_tensor)
            loss, loss_reduced = output_tensor
            output_tensor = loss / global_vars.get_num_microbatches()
            forward_data_store.append(loss_reduced)
        else:
            forward_data_store.append(loss_func(output_tensor, non_loss_data=True))
    if timers is not None:
        timers("forward-compute").stop()
    if mpu.is_pipeline_stage_after_split() and args.model_type == ModelType.encoder_and_decoder:
        return [output_tensor, input_tensor[-1]]
    return output_tensor if unwrap_output else [output_tensor]


def backward_step(optimizer, input_tensor, output_tensor, output_tensor_grad, timers, args=None):
    args = args or global_vars.get_args()
    if timers is not None:
        timers("backward-compute", log_level=2).start()
    unwrap_grad = False
    if not isinstance(input_tensor, list):
        input_tensor = [input_tensor]
        unwrap_grad = True
    for t in input_tensor:
        if t is not None:
            t.retain_grad()
    if not isinstance(output_tensor, list):
        output_tensor = [output_tensor]
    if not isinstance(output_tensor_grad, list):
        output_tensor_grad = [output_tensor_grad]
    if output_tensor_grad[0] is None and optimizer is not None:
        output_tensor = [optimizer.scale_loss(output_tensor[0])] + output_tensor[1:]
    if output_tensor_grad[0] is None:
        torch.autograd.backward(output_tensor[0], grad_tensors=None)
    else:
        custom_backward(output_tensor[0], output_tensor_grad[0])
    grads = [None if t is None else t.grad for t in input_tensor]
    if mpu.get_pipeline_model_parallel_world_size() > 1 and mpu.is_pipeline_stage_after_split() and \
            args.model_type == ModelType.encoder_and_decoder:
        if output_tensor_grad[1] is not None:
            grads[-1].add_(output_tensor_grad[1])
    if timers is not None:
        timers("backward-compute").stop()
    return grads[0] if unwrap_grad else grads


@contextmanager
def dummy_handler():
    yield


def forward_backward_no_pipelining(forward_step_func, data_iterator, model, optimizer, timers, forward_only,
                                   collect_non_loss_data=False):
    assert len(model) == 1
    model = model[0]
    ctx = model.no_sync if isinstance(model, torchDDP) else dummy_handler
    store = []
    n = global_vars.get_num_microbatches()
    with ctx():
        for _ in range(n - 1):
            out = forward_step(forward_step_func, data_iterator, model, None, store, timers, collect_non_loss_data)
            if not forward_only:
                backward_step(optimizer, None, out, None, timers)
    out = forward_step(forward_step_func, data_iterator, model, None, store, timers, collect_non_loss_data)
    if not forward_only:
        backward_step(optimizer, None, out, None, timers)
    return store


def get_tensor_shapes(rank, model_type, args=None):
    """Shapes exchanged between stages: (s or s/tp, micro batch, h); T5 decoder stages also
    carry the encoder output."""
    args = args or global_vars.get_args()
    tp = mpu.get_tensor_model_parallel_world_size()
    seq = args.seq_length // tp if args.sequence_parallel else args.seq_length
    shapes = []
    if model_type == ModelType.encoder_and_decoder:
        dec = args.decoder_seq_length // tp if args.sequence_parallel else args.decoder_seq_length
        if mpu.is_pipeline_stage_before_split(rank):
            shapes.append((seq, args.micro_batch_size, args.hidden_size))
        else:
            shapes.append((dec, args.micro_batch_size, args.hidden_size))
            shapes.append((seq, args.micro_batch_size, args.hidden_size))
    else:
        shapes.append((seq, args.micro_batch_size, args.hidden_size))
    return shapes


def recv_forward(tensor_shapes, timers):
    return [None if s is None else p2p_communication.recv_forward(s, timers=timers) for s in tensor_shapes]


def recv_backward(tensor_shapes, timers):
    return [None if s is None else p2p_communication.recv_backward(s, timers=timers) for s in tensor_shapes]


def send_forward(output_tensors, tensor_shapes, timers):
    if not isinstance(output_tensors, list):
        output_tensors = [output_tensors]
    for t, s in zip(output_tensors, tensor_shapes):
        if s is not None:
            p2p_communication.send_forward(t, s, timers=timers)


def send_backward(input_tensor_grads, tensor_shapes, timers):
    if not isinstance(input_tensor_grads, list):
        input_tensor_grads = [input_tensor_grads]
    for t, s in zip(input_tensor_grads, tensor_shapes):
        if s is not None:
            p2p_communication.send_backward(t, s, timers=timers)


def send_forward_recv_backward(output_tensors, tensor_shapes, timers):
    if not isinstance(output_tensors, list):
        output_tensors = [output_tensors]
    return [None if s is None else p2p_communication.send_forward_recv_backward(t, s, timers=timers)
            for t, s in zip(output_tensors, tensor_shapes)]


def send_backward_recv_forward(input_tensor_grads, tensor_shapes, timers):
    if not isinstance(input_tensor_grads, list):
        input_tensor_grads = [input_tensor_grads]
    return [None if s is None else p2p_communication.send_backward_recv_forward(t, s, timers=timers)
            for t, s in zip(input_tensor_grads, tensor_shapes)]


def forward_backward_pipelining_without_interleaving(forward_step_func, data_iterator, model, optimizer, timers,
                                                     forward_only, collect_non_loss_data=False):
    """1F1B: (pp - rank - 1) warm-up forwards, steady one-forward-one-backward, cool-down."""
    args = global_vars.get_args()
    assert len(model) == 1
    model = model[0]
    n_mb = global_vars.get_num_microbatches()
    pp, rank = mpu.get_pipeline_model_parallel_world_size(), mpu.get_pipeline_model_parallel_rank()
    n_warm = min(pp - rank - 1, n_mb)
    n_steady = n_mb - n_warm
    model_type = args.model_type
    recv_shapes = get_tensor_shapes(rank - 1, model_type)
    send_shapes = get_tensor_shapes(rank, model_type)
    store, inputs, outputs = [], [], []

    for _ in range(n_warm):
        inp = recv_forward(recv_shapes, timers)
        out = forward_step(forward_step_func, data_iterator, model, inp, store, timers, collect_non_loss_data)
        send_forward(out, send_shapes, timers)
        if not forward_only:
            inputs.append(inp)
            outputs.append(out)
            deallocate_output_tensor(out[0])
    if n_steady > 0:
        inp = recv_forward(recv_shapes, timers)
    for i in range(n_steady):
        last = i == n_steady - 1
        out = forward_step(forward_step_func, data_iterator, model, inp, store, timers, collect_non_loss_data)
        if forward_only:
            send_forward(out, send_shapes, timers)
            if not last:
                inp = recv_forward(recv_shapes, timers)
            continue
        out_grad = send_forward_recv_backward(out, send_shapes, timers)
        inputs.append(inp)
        outputs.append(out)
        deallocate_output_tensor(out[0])
        inp_b, out_b = inputs.pop(0), outputs.pop(0)
        inp_grad = backward_step(optimizer, inp_b, out_b, out_grad, timers)
        if last:
            inp = None
            send_backward(inp_grad, recv_shapes, timers)
        else:
            inp = send_backward_recv_forward(inp_grad, recv_shapes, timers)
    if not forward_only:
        for _ in range(n_warm):
            inp_b, out_b = inputs.pop(0), outputs.pop(0)
            out_grad = recv_backward(send_shapes, timers)
            inp_grad = backward_step(optimizer, inp_b, out_b, out_grad, timers)
            send_backward(inp_grad, recv_shapes, timers)
    return store


def forward_backward_pipelining_with_interleaving(forward_step_func, data_iterator, model, optimizer, timers,
                                                  forward_only, collect_non_loss_data=False):
    """Interleaved 1F1B: each rank owns ``len(model)`` model chunks (virtual stages); the
    unit of work is (micro-batch, chunk). Virtual micro-batch k maps to chunk
    (k mod (pp * chunks)) // pp, reversed for the backward direction."""
    args = global_vars.get_args()
    n_chunks = len(model)
    in_tensors = [[] for _ in range(n_chunks)]
    out_tensors = [[] for _ in range(n_chunks)]
    store = []
    out_grads = [[] for _ in range(n_chunks)] if not forward_only else None
    pp, rank = mpu.get_pipeline_model_parallel_world_size(), mpu.get_pipeline_model_parallel_rank()
    seq = args.seq_length // mpu.get_tensor_model_parallel_world_size() if args.sequence_parallel \
        else args.seq_length
    shape = (seq, args.micro_batch_size, args.hidden_size)
    n_mb = global_vars.get_num_microbatches() * n_chunks
    all_warm = False
    if forward_only:
        n_warm = n_mb
    elif global_vars.get_num_microbatches() == pp:
        n_warm, all_warm = n_mb, True
    else:
        n_warm = min((pp - rank - 1) * 2 + (n_chunks - 1) * pp, n_mb)
    n_steady = n_mb - n_warm

    def chunk_of(k, forward):
        c = (k % (pp * n_chunks)) // pp
        return c if forward else n_chunks - c - 1

    def fwd_helper(k):
        c = chunk_of(k, True)
        mpu.set_virtual_pipeline_model_parallel_rank(c)
        if mpu.is_pipeline_first_stage() and len(in_tensors[c]) == len(out_tensors[c]):
            in_tensors[c].append(None)
        out = forward_step(forward_step_func, data_iterator[c], model[c], in_tensors[c][-1], store, timers,
                           collect_non_loss_data)
        out_tensors[c].append(out)
        if forward_only:
            in_tensors[c].pop()
            out_tensors[c].pop()
        return out

    def bwd_helper(k):
        c = chunk_of(k, False)
        mpu.set_virtual_pipeline_model_parallel_rank(c)
        if mpu.is_pipeline_last_stage() and len(out_grads[c]) == 0:
            out_grads[c].append(None)
        return backward_step(optimizer, in_tensors[c].pop(0), out_tensors[c].pop(0), out_grads[c].pop(0), timers)

    mpu.set_virtual_pipeline_model_parallel_rank(0)
    in_tensors[0].append(p2p_communication.recv_forward(shape, timers=timers))
    for k in range(n_warm):
        out = fwd_helper(k)
        nxt = chunk_of(k + 1, True)
        recv_prev = True
        if mpu.is_pipeline_first_stage(ignore_virtual=True) and nxt == 0:
            recv_prev = False
        if k == n_mb - 1:
            recv_prev = False
        if mpu.is_pipeline_last_stage():
            out = None
        if k == n_warm - 1 and not forward_only and not all_warm:
            recv_next = not mpu.is_pipeline_last_stage(ignore_virtual=True)
            inp, og = p2p_communication.send_forward_backward_recv_forward_backward(
                out, None, recv_prev=recv_prev, recv_next=recv_next, tensor_shape=shape, timers=timers)
            out_grads[n_chunks - 1].append(og)
        else:
            inp = p2p_communication.send_forward_recv_forward(out, recv_prev=recv_prev, tensor_shape=shape,
                                                              timers=timers)
        in_tensors[nxt].append(inp)
        deallocate_output_tensor(out)

    for i in range(n_steady):
        fk = i + n_warm
        out = fwd_helper(fk)
        inp_grad = bwd_helper(i)
        fc = chunk_of(fk, True)
        mpu.set_virtual_pipeline_model_parallel_rank(fc)
        if mpu.is_pipeline_last_stage():
            out = None
        bc = chunk_of(i, False)
        mpu.set_virtual_pipeline_model_parallel_rank(bc)
        if mpu.is_pipeline_first_stage():
            inp_grad = None
        recv_prev = True
        if mpu.is_pipeline_first_stage(ignore_virtual=True):
            nfc = chunk_of(fk - (pp - 1), True)
            if nfc == n_chunks - 1:
                recv_prev = False
            nfc += 1
        else:
            nfc = chunk_of(fk + 1, True)
        recv_next = True
        if mpu.is_pipeline_last_stage(ignore_virtual=True):
            nbc = chunk_of(i - (pp - 1), False)
            if nbc == 0:
                recv_next = False
            nbc -= 1
        else:
            nbc = chunk_of(i + 1, False)
        if i == n_steady - 1:
            recv_prev = False
        inp, og = p2p_communication.send_forward_backward_recv_forward_backward(
            out, inp_grad, recv_prev=recv_prev, recv_next=recv_next, tensor_shape=shape, timers=timers)
        deallocate_output_tensor(out)
        if recv_prev:
            in_tensors[nfc].append(inp)
        if recv_next:
            out_grads[nbc].append(og)

    if not forward_only:
        if all_warm:
            out_grads[n_chunks - 1].append(p2p_communication.recv_backward(shape, timers=timers))
        for k in range(n_steady, n_mb):
            inp_grad = bwd_helper(k)
            nbc = chunk_of(k + 1, False)
            recv_next = True
            if mpu.is_pipeline_last_stage(ignore_virtual=True) and nbc == n_chunks - 1:
                recv_next = False
            if k == n_mb - 1:
                recv_next = False
            out_grads[nbc].append(p2p_communication.send_backward_recv_backward(
                inp_grad, recv_next=recv_next, tensor_shape=shape, timers=timers))
    return store
