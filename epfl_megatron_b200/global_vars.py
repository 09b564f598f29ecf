"""Process-wide singletons: args, micro-batch calculator, tokenizer, TB/WandB writer,
timers, signal handler (megatron/global_vars.py:22-192 surface)."""
from __future__ import annotations

import os
import sys

import torch

from . import dist_signal_handler
from .microbatches import build_num_microbatches_calculator
from .timers import Timers

_ARGS = None
_NUM_MICROBATCHES_CALCULATOR = None
_TOKENIZER = None
_TENSORBOARD_WRITER = None
_ADLR_AUTORESUME = None
_TIMERS = None
_SIGNAL_HANDLER = None
_COUNTERS = {}


def _ensure(var, name):
    assert var is not None, f"{name} is not initialized."
    return var


def get_args():
    return _ensure(_ARGS, "args")


def get_num_microbatches():
    return _NUM_MICROBATCHES_CALCULATOR.get()


def get_current_global_batch_size():
    return _NUM_MICROBATCHES_CALCULATOR.get_current_global_batch_size()


def update_num_microbatches(consumed_samples, consistency_check=True):
    _NUM_MICROBATCHES_CALCULATOR.update(consumed_samples, consistency_check)


def get_tokenizer():
    return _ensure(_TOKENIZER, "tokenizer")


def get_tensorboard_writer():
    return _TENSORBOARD_WRITER  # may legitimately be None


def get_adlr_autoresume():
    return _ADLR_AUTORESUME


def get_timers():
    return _ensure(_TIMERS, "timers")


def get_signal_handler():
    return _ensure(_SIGNAL_HANDLER, "signal handler")


def get_counters():
    return _COUNTERS


def _set_signal_handler():
    global _SIGNAL_HANDLER
    _SIGNAL_HANDLER = dist_signal_handler.DistributedSignalHandler().__enter__()


def set_args(args):
    global _ARGS
    _ARGS = args


def set_global_variables(args, build_tokenizer_flag=True):
    global _NUM_MICROBATCHES_CALCULATOR, _TIMERS
    assert args is not None
    set_args(args)
    _NUM_MICROBATCHES_CALCULATOR = build_num_microbatches_calculator(args)
    if build_tokenizer_flag:
        _build_tokenizer(args)
    _set_tensorboard_writer(args)
    _set_adlr_autoresume(args)
    if getattr(args, "nvtx_ranges", False):
        from .timers import set_nvtx_ranges
        set_nvtx_ranges(True)
        args.timing_log_level = 2       # every phase gets a (real) timer, hence a range
    _TIMERS = Timers(args.timing_log_level, args.timing_log_option)
    if args.exit_signal_handler:
        _set_signal_handler()


def _build_tokenizer(args):
    global _TOKENIZER
    from .tokenizer import build_tokenizer
    _TOKENIZER = build_tokenizer(args)
    return _TOKENIZER


def rebuild_tokenizer(args):
    global _TOKENIZER
    _TOKENIZER = None
    return _build_tokenizer(args)


def _set_tensorboard_writer(args):
    """The LAST global rank owns the writer (it is on the last pipeline stage, the only
    one that sees the loss), as in global_vars.py:119-153."""
    global _TENSORBOARD_WRITER
    if args.rank != args.world_size - 1:
        return
    if getattr(args, "wandb_logger", False):
        from .wandb_logger import WandBConfig, WandbTBShim
        cfg = WandBConfig.from_args(args)
        _TENSORBOARD_WRITER = WandbTBShim(cfg)
        print("> setting wandb ...", flush=True)
    elif getattr(args, "tensorboard_dir", None):
        try:
            from torch.utils.tensorboard import SummaryWriter
            print("> setting tensorboard ...", flush=True)
            _TENSORBOARD_WRITER = SummaryWriter(log_dir=args.tensorboard_dir,
                                                max_queue=args.tensorboard_queue_size)
        except Exception:  # noqa: BLE001
            from .utils.jsonl_writer import JsonlScalarWriter
            print("WARNING: TensorBoard unavailable, writing scalars as JSONL instead.", flush=True)
            os.makedirs(args.tensorboard_dir, exist_ok=True)
            _TENSORBOARD_WRITER = JsonlScalarWriter(os.path.join(args.tensorboard_dir, "scalars.jsonl"))


def _set_adlr_autoresume(args):
    global _ADLR_AUTORESUME
    if getattr(args, "adlr_autoresume", False):
        if args.rank == 0:
            print("enabling autoresume ...", flush=True)
        sys.path.append(os.environ.get("SUBMIT_SCRIPTS", "."))
        try:
            from userlib.auto_resume import AutoResume
        except ImportError:
            print("ADLR autoresume is not available, exiting ...")
            sys.exit()
        _ADLR_AUTORESUME = AutoResume


def destroy_global_vars():
    global _ARGS, _NUM_MICROBATCHES_CALCULATOR, _TOKENIZER, _TENSORBOARD_WRITER
    global _ADLR_AUTORESUME, _TIMERS, _SIGNAL_HANDLER
    _ARGS = _NUM_MICROBATCHES_CALCULATOR = _TOKENIZER = _TENSORBOARD_WRITER = None
    _ADLR_AUTORESUME = _TIMERS = _SIGNAL_HANDLER = None
