"""Named RNG streams and activation recomputation
(megatron/core/tensor_parallel/random.py:31-252 behaviour).

* ``CudaRNGStatesTracker``: named generator states; ``fork(name)`` swaps the state in for
  the duration of a ``with`` block. The default state is seeded with ``seed`` (same on
  all TP ranks: dropout outside TP regions), the ``model-parallel-rng`` state with
  ``seed + 2718 + tp_rank`` (different per TP rank: dropout inside TP regions and weight
  init of sharded parameters).
* ``checkpoint(fn, distribute_saved_activations, *args)``: recompute ``fn`` in backward
  while restoring CPU / device / tracker RNG states; optionally keep only this rank's
  1/tp slice of the saved input.

Works without CUDA too (CPU generator states), which the Gloo test path needs.
"""
from __future__ import annotations

import contextlib

import torch
from torch.utils.checkpoint import detach_variable

from . import state
from .utils import gather_split_1d_tensor, safely_set_viewless_tensor_data, split_tensor_into_1d_equal_chunks

_MODEL_PARALLEL_RNG_TRACKER_NAME = "model-parallel-rng"


def _cuda():
    return torch.cuda.is_available()


def _get_state():
    return torch.cuda.get_rng_state() if _cuda() else torch.get_rng_state()


def _set_state(s):
    if _cuda():
        torch.cuda.set_rng_state(s)
    else:
        torch.set_rng_state(s)


def _seed(seed):
    if _cuda():
        torch.cuda.manual_seed(seed)
    else:
        torch.manual_seed(seed)


class CudaRNGStatesTracker:
    def __init__(self):
        self.states_ = {}
        self.seeds_ = set()

    def reset(self):
        self.states_ = {}
        self.seeds_ = set()

    def get_states(self):
        return dict(self.states_)

    def set_states(self, states):
        self.states_ = states

    def add(self, name, seed):
        if seed in self.seeds_:
            raise Exception(f"seed {seed} already exists")
        if name in self.states_:
            raise Exception(f"cuda rng state {name} already exists")
        self.seeds_.add(seed)
        orig = _get_state()
        _seed(seed)
        self.states_[name] = _get_state()
        _set_state(orig)

    @contextlib.contextmanager
    def fork(self, name=_MODEL_PARALLEL_RNG_TRACKER_NAME):
        if name not in self.states_:
            raise Exception(f"cuda rng state {name} is not added")
        orig = _get_state()
        _set_state(self.states_[name])
        try:
            yield
        finally:
            self.states_[name] = _get_state()
            _set_state(orig)


_TRACKER = CudaRNGStatesTracker()


def get_cuda_rng_tracker():
    return _TRACKER


def model_parallel_cuda_manual_seed(seed):
    offset = seed + 2718
    tp_seed = offset + state.get_tensor_model_parallel_rank()
    _TRACKER.reset()
    _seed(seed)
    _TRACKER.add(_MODEL_PARALLEL_RNG_TRACKER_NAME, tp_seed)


class CheckpointFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, run_function, distribute_saved_activations, *args):
        ctx.run_function = run_function
        ctx.distribute = distribute_saved_activations
        ctx.cpu_state = torch.get_rng_state()
        ctx.dev_state = torch.cuda.get_rng_state() if _cuda() else None
        ctx.tracker_states = get_cuda_rng_tracker().get_states()
        with torch.no_grad():
            outputs = run_function(*args)
        if distribute_saved_activations:
            ctx.input_0_shape = args[0].data.shape
            safely_set_viewless_tensor_data(
                args[0], split_tensor_into_1d_equal_chunks(args[0].data, new_buffer=True))
        ctx.save_for_backward(*args)
        return outputs

    @staticmethod
    def backward(ctx, *grads):
        if not torch.autograd._is_checkpoint_valid():
            raise RuntimeError("Checkpointing is not compatible with .grad(), use .backward()")
        inputs = ctx.saved_tensors
        if ctx.distribute:
            safely_set_viewless_tensor_data(
                inputs[0], gather_split_1d_tensor(inputs[0].data).view(ctx.input_0_shape))
        cpu_now = torch.get_rng_state()
        dev_now = torch.cuda.get_rng_state() if _cuda() else None
        tracker_now = get_cuda_rng_tracker().get_states()
        torch.set_rng_state(ctx.cpu_state)
        if ctx.dev_state is not None:
            torch.cuda.set_rng_state(ctx.dev_state)
        get_cuda_rng_tracker().set_states(ctx.tracker_states)
        detached = detach_variable(inputs)
        with torch.enable_grad():
            outputs = ctx.run_function(*detached)
        torch.set_rng_state(cpu_now)
        if dev_now is not None:
            torch.cuda.set_rng_state(dev_now)
        get_cuda_rng_tracker().set_states(tracker_now)
        if isinstance(outputs, torch.Tensor):
            outputs = (outputs,)
        pairs = [(o, g) for o, g in zip(outputs, grads) if isinstance(o, torch.Tensor) and o.requires_grad]
        torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
        return (None, None) + tuple(i.grad if isinstance(i, torch.Tensor) else i for i in detached)


def checkpoint(function, distribute_saved_activations, *args):
    return CheckpointFunction.apply(function, distribute_saved_activations, *args)
