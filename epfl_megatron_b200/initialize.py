"""Process initialisation (megatron/initialize.py:26-275 surface): parse and validate
arguments, set globals, create the process groups, seed the RNG streams.

Unlike the reference (initialize.py:36) a GPU is not mandatory: with
``--distributed_backend gloo`` everything runs on CPU tensors (BASELINE config 1).
Native code is built ahead of time by ``__graft_entry__.build`` / ``ops.build`` and loaded
lazily by ``ops.lib``; nothing is JIT-compiled here.
"""
from __future__ import annotations

import os
import random
import time
from datetime import timedelta

import numpy as np
import torch
import torch.distributed as dist

from . import global_vars
from .arguments import parse_args, validate_args
from .checkpointing import load_args_from_checkpoint
from .ops import functional as Fn
from .parallel import random as tp_random
from .parallel import state as mpu


def initialize_megatron(extra_args_provider=None, args_defaults=None, ignore_unknown_args=False,
                        allow_no_cuda=True, args_list=None):
    args_defaults = args_defaults or {}
    args = parse_args(extra_args_provider, ignore_unknown_args, args_list)
    if args.use_checkpoint_args or args_defaults.get("use_checkpoint_args", False):
        assert args.load is not None, "--use_checkpoint_args requires --load argument"
        load_args_from_checkpoint(args)
    if not torch.cuda.is_available():
        if not allow_no_cuda:
            raise AssertionError("this entry point requires CUDA")
        if args.distributed_backend == "nccl":
            args.distributed_backend = "gloo"
        args.use_cpu_initialization = True
    validate_args(args, args_defaults)
    global_vars.set_global_variables(args)
    Fn.set_kernel_mode({"auto": "auto", "on": "on", "off": "off"}[args.b200_kernels])
    Fn.set_gemm_mode(args.b200_gemm)
    from .parallel import layers as _tp_layers
    _tp_layers.set_fused_comm_mode(args.b200_fused_comm)
    _initialize_distributed(args)
    if args.rank == 0:
        print(f"> setting random seeds to {args.seed} ...", flush=True)
    _set_random_seed(args.seed, args.data_parallel_random_init)
    args = global_vars.get_args()
    _init_autoresume()
    return None


def _initialize_distributed(args):
    n_dev = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if dist.is_initialized():
        if args.rank == 0:
            print("torch distributed is already initialized, skipping initialization ...", flush=True)
        args.rank = dist.get_rank()
        args.world_size = dist.get_world_size()
    else:
        if args.rank == 0:
            print("> initializing torch distributed ...", flush=True)
        if n_dev > 0 and args.distributed_backend == "nccl":
            device = args.rank % n_dev
            if args.local_rank is not None:
                assert args.local_rank == device, "expected local-rank to be the same as rank % device-count."
            else:
                args.local_rank = device
            torch.cuda.set_device(device)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group(backend=args.distributed_backend, world_size=args.world_size, rank=args.rank,
                                timeout=timedelta(minutes=10))
    if n_dev > 0 or args.distributed_backend == "gloo":
        if mpu.model_parallel_is_initialized():
            print("model parallel is already initialized")
        else:
            mpu.initialize_model_parallel(args.tensor_model_parallel_size, args.pipeline_model_parallel_size,
                                          args.virtual_pipeline_model_parallel_size,
                                          args.pipeline_model_parallel_split_rank)
            if args.rank == 0:
                print(f"> initialized tensor model parallel with size {mpu.get_tensor_model_parallel_world_size()}")
                print(f"> initialized pipeline model parallel with size {mpu.get_pipeline_model_parallel_world_size()}")


def _init_autoresume():
    autoresume = global_vars.get_adlr_autoresume()
    if autoresume:
        dist.barrier()
        autoresume.init()
        dist.barrier()


def _set_random_seed(seed_, data_parallel_random_init=False):
    """seed + 100 * pp_rank (+ 10 * dp_rank) for python / numpy / torch; the TP tracker adds
    2718 + tp_rank for the tensor-parallel stream."""
    if seed_ is None or seed_ <= 0:
        raise ValueError(f"Seed ({seed_}) should be a positive integer.")
    seed = seed_ + 100 * mpu.get_pipeline_model_parallel_rank()
    if data_parallel_random_init:
        seed = seed + 10 * mpu.get_data_parallel_rank()
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    tp_random.model_parallel_cuda_manual_seed(seed)


def write_args_to_tensorboard():
    args = global_vars.get_args()
    writer = global_vars.get_tensorboard_writer()
    if writer:
        for arg in vars(args):
            writer.add_text(arg, str(getattr(args, arg)), global_step=args.iteration)


def set_jit_fusion_options():
    """Kept for API parity; the fused ops are ahead-of-time compiled CUDA kernels, there is
    no JIT fuser to configure or warm up."""
    return None
