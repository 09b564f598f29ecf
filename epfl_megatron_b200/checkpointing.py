"""Checkpoint save / load in the Megatron layout (megatron/checkpointing.py:35-730 surface).

    <save>/latest_checkpointed_iteration.txt            an integer or the word ``release``
    <save>/iter_0001234/mp_rank_TT[_PPP]/model_optim_rng.pt
    with --use_distributed_optimizer:
    <save>/iter_0001234/mp_rank_TT[_PPP]/model_rng.pt        (DP rank 0 writes the model)
    <save>/iter_0001234/mp_rank_TT[_PPP]_DDD/optim.pt        (every DP rank writes its shard)

The top-level dict carries ``args`` (the whole Namespace), ``checkpoint_version`` 3.0,
``iteration``, ``model`` (or ``model0..modelK`` per virtual-pipeline chunk), ``optimizer``,
``opt_param_scheduler`` and ``rng_state``. Both the modern key spellings and the legacy
aliases written by the converters load (handled in models/language_model.py).
"""
from __future__ import annotations

import os
import random
import sys

import numpy as np
import torch
import torch.distributed as dist

from . import global_vars
from .parallel import random as tp_random
from .parallel import state as mpu
from .utils import print_rank_0, unwrap_model

_CHECKPOINT_VERSION = None


def set_checkpoint_version(value):
    global _CHECKPOINT_VERSION
    if _CHECKPOINT_VERSION is not None:
        assert _CHECKPOINT_VERSION == value, "checkpoint versions do not match"
    _CHECKPOINT_VERSION = value


def get_checkpoint_version():
    return _CHECKPOINT_VERSION


def check_checkpoint_args(checkpoint_args, args=None):
    """The architecture and model-parallel layout of a resumed run must match the checkpoint."""
    args = args or global_vars.get_args()

    def same(name, old_name=None):
        ck = getattr(checkpoint_args, old_name or name, None) if old_name else getattr(checkpoint_args, name, None)
        if ck is None:
            return
        cur = getattr(args, name)
        assert ck == cur, f"{name} value from checkpoint ({ck}) is not equal to the input argument value ({cur})."

    same("num_layers")
    same("hidden_size")
    same("num_attention_heads")
    if getattr(args, "vocab_file", None):
        same("max_position_embeddings")
        same("make_vocab_size_divisible_by")
        same("padded_vocab_size")
        same("tokenizer_type")
    if getattr(args, "data_parallel_random_init", False):
        same("data_parallel_random_init")
    if get_checkpoint_version() is None or get_checkpoint_version() < 3.0:
        same("tensor_model_parallel_size", old_name="model_parallel_size")
    else:
        same("tensor_model_parallel_size")
        same("pipeline_model_parallel_size")


def ensure_directory_exists(filename):
    os.makedirs(os.path.dirname(filename), exist_ok=True)


def get_checkpoint_names(checkpoints_path, iteration, use_distributed_optimizer, release=False,
                         pipeline_parallel=None, tensor_rank=None, pipeline_rank=None):
    directory = "release" if release else "iter_{:07d}".format(iteration)
    if pipeline_parallel is None:
        pipeline_parallel = mpu.get_pipeline_model_parallel_world_size() > 1
    if tensor_rank is None:
        tensor_rank = mpu.get_tensor_model_parallel_rank()
    if pipeline_rank is None:
        pipeline_rank = mpu.get_pipeline_model_parallel_rank()
    if not pipeline_parallel:
        common = os.path.join(checkpoints_path, directory, f"mp_rank_{tensor_rank:02d}")
    else:
        common = os.path.join(checkpoints_path, directory, f"mp_rank_{tensor_rank:02d}_{pipeline_rank:03d}")
    if use_distributed_optimizer:
        model_name = os.path.join(common, "model_rng.pt")
        optim_name = os.path.join(common + f"_{mpu.get_data_parallel_rank():03d}", "optim.pt")
    else:
        model_name = optim_name = os.path.join(common, "model_optim_rng.pt")
    return model_name, optim_name


def find_checkpoint_rank_0(checkpoints_path, iteration, use_distributed_optimizer, release=False):
    """A checkpoint file of rank 0 regardless of whether it was saved with pp > 1."""
    for pp in (False, True):
        names = get_checkpoint_names(checkpoints_path, iteration, use_distributed_optimizer, release,
                                     pipeline_parallel=pp, tensor_rank=0, pipeline_rank=0)
        if os.path.isfile(names[0]):
            return names
    return None, None


def get_checkpoint_tracker_filename(checkpoints_path):
    return os.path.join(checkpoints_path, "latest_checkpointed_iteration.txt")


def read_metadata(tracker_filename):
    iteration, release = 0, False
    with open(tracker_filename, "r") as f:
        meta = f.read().strip()
        try:
            iteration = int(meta)
        except ValueError:
            release = meta == "release"
            if not release:
                print_rank_0(f"ERROR: Invalid metadata file {tracker_filename}. Exiting")
                sys.exit()
    assert iteration > 0 or release, f"error parsing metadata file {tracker_filename}"
    if dist.is_initialized():
        dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
        t = torch.tensor([iteration], dtype=torch.long, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        mx = int(t[0].item())
        if iteration != mx:
            print(f"WARNING: on rank {dist.get_rank()} found iteration {iteration} in the metadata while "
                  f"max iteration across the ranks is {mx}, replacing it with max iteration.", flush=True)
    else:
        mx = iteration
    return mx, release


def get_rng_state(args=None):
    args = args or global_vars.get_args()
    state = {"random_rng_state": random.getstate(), "np_rng_state": np.random.get_state(),
             "torch_rng_state": torch.get_rng_state(),
             "cuda_rng_state": torch.cuda.get_rng_state() if torch.cuda.is_available() else None,
             "rng_tracker_states": tp_random.get_cuda_rng_tracker().get_states()}
    if dist.is_initialized() and mpu.get_data_parallel_world_size() > 1 and args.data_parallel_random_init:
        out = [None] * mpu.get_data_parallel_world_size()
        dist.all_gather_object(out, state, group=mpu.get_data_parallel_group())
        return out
    return [state]


def save_checkpoint(iteration, model, optimizer, opt_param_scheduler, args=None):
    args = args or global_vars.get_args()
    model = unwrap_model(model)
    print_rank_0(f"saving checkpoint at iteration {iteration:7d} to {args.save}")
    rng_state = get_rng_state(args)
    model_name, optim_name = get_checkpoint_names(args.save, iteration, args.use_distributed_optimizer)
    dp_rank = mpu.get_data_parallel_rank()
    if args.use_distributed_optimizer or dp_rank == 0 or not dist.is_initialized():
        model_sd, optim_sd = {}, {}
        if dp_rank == 0 or not dist.is_initialized():
            model_sd["args"] = args
            model_sd["checkpoint_version"] = 3.0
            model_sd["iteration"] = iteration
            if len(model) == 1:
                model_sd["model"] = model[0].state_dict_for_save_checkpoint()
            else:
                for i in range(len(model)):
                    mpu.set_virtual_pipeline_model_parallel_rank(i)
                    model_sd[f"model{i}"] = model[i].state_dict_for_save_checkpoint()
            if not args.no_save_rng:
                model_sd["rng_state"] = rng_state
        if not args.no_save_optim:
            if optimizer is not None:
                optim_sd["optimizer"] = optimizer.state_dict()
            if opt_param_scheduler is not None:
                optim_sd["opt_param_scheduler"] = opt_param_scheduler.state_dict()
        if args.use_distributed_optimizer:
            if model_sd:
                ensure_directory_exists(model_name)
                torch.save(model_sd, model_name)
            if optim_sd:
                ensure_directory_exists(optim_name)
                torch.save(optim_sd, optim_name)
        else:
            sd = {**model_sd, **optim_sd}
            if sd:
                ensure_directory_exists(model_name)
                torch.save(sd, model_name)
    if dist.is_initialized():
        dist.barrier()
    print_rank_0(f"  successfully saved checkpoint at iteration {iteration:7d} to {args.save}")
    if not dist.is_initialized() or dist.get_rank() == 0:
        with open(get_checkpoint_tracker_filename(args.save), "w") as f:
            f.write(str(iteration))
    if dist.is_initialized():
        dist.barrier()


def _transpose_first_dim(t, num_splits, num_splits_first, model, args):
    """Re-order a fused QKV parameter saved with checkpoint version 0 / 1.0."""
    shape = t.size()
    hn = args.hidden_size // args.num_attention_heads
    np_ = args.num_attention_heads // mpu.get_tensor_model_parallel_world_size()
    if num_splits_first:
        t = t.view(num_splits, np_, hn, *shape[1:]).transpose(0, 1).contiguous()
    else:
        t = t.view(np_, hn, num_splits, *shape[1:]).transpose(1, 2).contiguous()
    return t.view(*shape)


def fix_query_key_value_ordering(model, checkpoint_version, args=None):
    args = args or global_vars.get_args()
    if checkpoint_version >= 2.0:
        return
    if isinstance(model, list):
        assert len(model) == 1
        model = model[0]
    for name, param in model.named_parameters():
        if name.endswith((".query_key_value.weight", ".query_key_value.bias")):
            param.data.copy_(_transpose_first_dim(param.data, 3, checkpoint_version == 0, model, args))
        if name.endswith((".key_value.weight", ".key_value.bias")):
            param.data.copy_(_transpose_first_dim(param.data, 2, checkpoint_version == 0, model, args))
    print_rank_0(f" succesfully fixed query-key-values ordering for checkpoint version {checkpoint_version}")


def _install_unpickle_aliases():
    """Reference checkpoints pickle ``megatron.*`` classes (enums inside args, legacy loss
    scalers); the ``megatron`` facade package at the repo root provides them."""
    try:
        import megatron  # noqa: F401
    except ImportError:
        pass


def _torch_load(path):
    _install_unpickle_aliases()
    return torch.load(path, map_location="cpu", weights_only=False)


def _load_base_checkpoint(load_dir, use_distributed_optimizer, rank0=False):
    tracker = get_checkpoint_tracker_filename(load_dir)
    if not os.path.isfile(tracker):
        if not rank0:
            print_rank_0(f"WARNING: could not find the metadata file {tracker} ")
            print_rank_0("    will not load any checkpoints and will start from random")
        return None, None, False
    iteration, release = read_metadata(tracker)
    if rank0:
        names = find_checkpoint_rank_0(load_dir, iteration, use_distributed_optimizer, release)
    else:
        names = get_checkpoint_names(load_dir, iteration, use_distributed_optimizer, release)
        if release:
            print_rank_0(f" loading release checkpoint from {load_dir}")
        else:
            print_rank_0(f" loading checkpoint from {load_dir} at iteration {iteration}")
    model_name, optim_name = names
    if model_name is None:
        return None, None, release
    try:
        model_sd = _torch_load(model_name)
        optim_sd = _torch_load(optim_name) if use_distributed_optimizer and os.path.isfile(optim_name) \
            else model_sd
    except BaseException as e:  # noqa: BLE001
        print_rank_0(f"could not load the checkpoint: {e}")
        sys.exit()
    return model_sd, optim_sd, release


_ARGS_FROM_CHECKPOINT = [
    ("num_layers", False), ("hidden_size", False), ("ffn_hidden_size", False), ("seq_length", False),
    ("num_attention_heads", False), ("num_attention_heads_kv", False), ("kv_channels", False),
    ("max_position_embeddings", False), ("tokenizer_type", False), ("padded_vocab_size", False),
    ("position_embedding_type", False), ("glu_activation", False), ("use_rms_norm", False),
    ("use_bias", False), ("parallel_attn", False), ("parallel_layernorm", False),
    ("tie_embed_logits", False), ("layernorm_epsilon", False), ("use_post_ln", False),
    ("rope_scaling_factor", False), ("make_vocab_size_divisible_by", False),
]


def load_args_from_checkpoint(args, load_arg="load"):
    """Pull architecture flags out of the checkpoint before validation (--use_checkpoint_args)."""
    load_dir = getattr(args, load_arg)
    if load_dir is None:
        return args
    model_sd, _, _ = _load_base_checkpoint(load_dir, use_distributed_optimizer=args.use_distributed_optimizer,
                                           rank0=True)
    if model_sd is None and args.use_distributed_optimizer:
        model_sd, _, _ = _load_base_checkpoint(load_dir, use_distributed_optimizer=False, rank0=True)
    if not model_sd or "args" not in model_sd:
        print("Checkpoint not found or provides no arguments; using provided arguments.")
        return args
    ck = model_sd["args"]
    args.iteration = model_sd.get("iteration", 0)
    for name, force in _ARGS_FROM_CHECKPOINT:
        if not force and getattr(args, name, None) not in (None, False) and name not in (
                "tie_embed_logits", "make_vocab_size_divisible_by", "layernorm_epsilon",
                "rope_scaling_factor", "position_embedding_type"):
            continue
        val = getattr(ck, name, None)
        if val is not None:
            setattr(args, name, val)
    if model_sd.get("checkpoint_version", 0) < 3.0:
        if getattr(ck, "model_parallel_size", None) is not None:
            args.tensor_model_parallel_size = ck.model_parallel_size
    return args


def load_checkpoint(model, optimizer, opt_param_scheduler, load_arg="load", strict=True, args=None):
    args = args or global_vars.get_args()
    load_dir = getattr(args, load_arg)
    model = unwrap_model(model)
    model_sd, optim_sd, release = _load_base_checkpoint(load_dir, args.use_distributed_optimizer)
    if model_sd is None:
        return 0
    set_checkpoint_version(model_sd.get("checkpoint_version", 0))
    if args.finetune or release:
        iteration = 0
    else:
        try:
            iteration = model_sd["iteration"]
        except KeyError:
            iteration = model_sd.get("total_iters")
            if iteration is None:
                print_rank_0("A metadata file exists but unable to load iteration from checkpoint, exiting")
                sys.exit()
    assert args.consumed_train_samples == 0 and args.consumed_valid_samples == 0
    if "args" in model_sd and not args.finetune:
        ck = model_sd["args"]
        check_checkpoint_args(ck, args)
        args.consumed_train_samples = getattr(ck, "consumed_train_samples", 0)
        global_vars.update_num_microbatches(consumed_samples=args.consumed_train_samples)
        args.consumed_valid_samples = getattr(ck, "consumed_valid_samples", 0)
    else:
        print_rank_0("could not find arguments in the checkpoint ...")
    if len(model) == 1:
        model[0].load_state_dict(model_sd["model"], strict=strict)
    else:
        for i in range(len(model)):
            mpu.set_virtual_pipeline_model_parallel_rank(i)
            model[i].load_state_dict(model_sd[f"model{i}"], strict=strict)
    print_rank_0(f" checkpoint version {get_checkpoint_version()}")
    fix_query_key_value_ordering(model, get_checkpoint_version(), args)
    if not release and not args.finetune and not args.no_load_optim:
        try:
            if optimizer is not None:
                optimizer.load_state_dict(optim_sd["optimizer"])
            if opt_param_scheduler is not None:
                if "lr_scheduler" in optim_sd:
                    opt_param_scheduler.load_state_dict(optim_sd["lr_scheduler"])
                else:
                    opt_param_scheduler.load_state_dict(optim_sd["opt_param_scheduler"])
        except KeyError:
            print_rank_0("Unable to load optimizer from checkpoint. Specify --no_load_optim or --finetune "
                         "to prevent attempting to load the optimizer state, exiting ...")
            sys.exit()
    elif optimizer is not None and (args.fp16 or args.bf16):
        optimizer.reload_model_params()
    if not release and not args.finetune and not args.no_load_rng:
        try:
            if "rng_state" in model_sd:
                states = model_sd["rng_state"]
                rs = states[mpu.get_data_parallel_rank()] if args.data_parallel_random_init and len(states) > 1 \
                    else states[0]
                random.setstate(rs["random_rng_state"])
                np.random.set_state(rs["np_rng_state"])
                torch.set_rng_state(rs["torch_rng_state"])
                if rs.get("cuda_rng_state") is not None and torch.cuda.is_available():
                    torch.cuda.set_rng_state(rs["cuda_rng_state"])
                if not rs["rng_tracker_states"]:
                    raise KeyError
                tp_random.get_cuda_rng_tracker().set_states(rs["rng_tracker_states"])
            else:
                random.setstate(model_sd["random_rng_state"])
                np.random.set_state(model_sd["np_rng_state"])
                torch.set_rng_state(model_sd["torch_rng_state"])
                if torch.cuda.is_available():
                    torch.cuda.set_rng_state(model_sd["cuda_rng_state"])
                if not model_sd["rng_tracker_states"]:
                    raise KeyError
                tp_random.get_cuda_rng_tracker().set_states(model_sd["rng_tracker_states"])
        except KeyError:
            print_rank_0("Unable to load rng state from checkpoint. Specify --no_load_rng or --finetune "
                         "to prevent attempting to load the rng state, exiting ...")
            sys.exit()
    if dist.is_initialized():
        dist.barrier()
    print_rank_0(f"  successfully loaded checkpoint from {load_dir} at iteration {iteration}")
    return iteration


def load_biencoder_checkpoint(model, only_query_model=False, only_context_model=False, custom_load_path=None):
    """Load the query / context towers of an ICT biencoder checkpoint."""
    args = global_vars.get_args()
    model = unwrap_model(model)
    load_path = custom_load_path if custom_load_path is not None else args.load
    with open(get_checkpoint_tracker_filename(load_path), "r") as f:
        iteration = int(f.read().strip())
    name, _ = get_checkpoint_names(load_path, iteration, args.use_distributed_optimizer, release=False)
    if mpu.get_data_parallel_rank() == 0:
        print(f"global rank {dist.get_rank() if dist.is_initialized() else 0} is loading checkpoint {name}")
    sd = _torch_load(name)
    ret = sd["model"]
    if only_query_model:
        ret.pop("context_model", None)
    if only_context_model:
        ret.pop("query_model", None)
    assert len(model) == 1
    model[0].load_state_dict(ret)
    if dist.is_initialized():
        dist.barrier()
    if mpu.get_data_parallel_rank() == 0:
        print(f" successfully loaded {name}")
    return model
