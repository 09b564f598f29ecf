"""Training driver (megatron/training.py:54-939 surface): ``pretrain``, ``get_model``,
``setup_model_and_optimizer``, ``train_step``, ``training_log``, ``evaluate``,
``build_train_valid_test_data_iterators``.

Additions: tokens/s, model TFLOP/s per GPU and MFU on the log line (the reference never
reports throughput), a JSONL metrics sink, and a step that performs no host
synchronisation in bf16 (loss / grad-norm values are read only at log time).
"""
from __future__ import annotations

import json
import math
import os
import sys
import time
from datetime import datetime

import torch
import torch.distributed as dist
from torch.nn.parallel.distributed import DistributedDataParallel as torchDDP

from . import global_vars
from .checkpointing import load_checkpoint, save_checkpoint
from .data.data_samplers import build_pretraining_data_loader
from .global_vars import (get_args, get_current_global_batch_size, get_num_microbatches,
                          get_signal_handler, get_tensorboard_writer, get_timers, update_num_microbatches)
from .initialize import initialize_megatron, set_jit_fusion_options, write_args_to_tensorboard
from .models.distributed import DistributedDataParallel as LocalDDP
from .models.enums import ModelType
from .models.module import Float16Module
from .optimizer import get_megatron_optimizer, make_group_key_fn
from .optimizer_param_scheduler import OptimizerParamScheduler
from .parallel import layers as tp_layers
from .parallel import state as mpu
from .schedules import get_forward_backward_func
from .utils import (calc_params_l2_norm, check_adlr_autoresume_termination, is_last_rank, print_rank_0,
                    print_rank_last, report_memory, unwrap_model)
from .utils.fault import maybe_inject

_TRAIN_START_TIME = time.time()


def print_datetime(string):
    if dist.is_initialized():
        dist.barrier()
    print_rank_0("[" + string + "] datetime: {} ".format(datetime.now().strftime("%Y-%m-%d %H:%M:%S")))


def _stat_device():
    if dist.is_initialized() and dist.get_backend() == "gloo" or not torch.cuda.is_available():
        return torch.device("cpu")
    return torch.device("cuda", torch.cuda.current_device())


def pretrain(args, train_valid_test_dataset_provider, model_provider_func, model_type, forward_step_func,
             process_non_loss_data_func=None, collate_fn=None):
    """initialize (done by the caller) -> model/optimizer -> data -> train -> eval -> save."""
    set_jit_fusion_options()
    global _TRAIN_START_TIME
    t = torch.tensor([_TRAIN_START_TIME], dtype=torch.double, device=_stat_device())
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
    _TRAIN_START_TIME = t.item()
    print_rank_0("time to initialize megatron (seconds): {:.3f}".format(time.time() - _TRAIN_START_TIME))
    print_datetime("after megatron is initialized")
    args = get_args()
    timers = get_timers()

    timers("model-and-optimizer-setup", log_level=0).start(barrier=True)
    model, optimizer, opt_param_scheduler = setup_model_and_optimizer(model_provider_func, model_type, args=args)
    timers("model-and-optimizer-setup").stop()
    print_datetime("after model, optimizer, and learning rate scheduler are built")

    timers("train/valid/test-data-iterators-setup", log_level=0).start(barrier=True)
    if args.virtual_pipeline_model_parallel_size is not None:
        its = [build_train_valid_test_data_iterators(train_valid_test_dataset_provider, args, collate_fn)
               for _ in range(len(model))]
        train_it, valid_it, test_it = ([x[i] for x in its] for i in range(3))
    else:
        train_it, valid_it, test_it = build_train_valid_test_data_iterators(
            train_valid_test_dataset_provider, args, collate_fn)
    timers("train/valid/test-data-iterators-setup").stop()
    print_datetime("after dataloaders are built")
    print_rank_0("done with setup ...")
    timers.log(["model-and-optimizer-setup", "train/valid/test-data-iterators-setup"], barrier=True)
    print_rank_0("training ...")

    iteration = 0
    if args.do_train and (args.train_iters or 0) > 0:
        iteration = _train(args, forward_step_func, model, optimizer, opt_param_scheduler, train_it, valid_it,
                           process_non_loss_data_func)
    print_datetime("after training is done")
    if args.do_valid:
        evaluate_and_print_results("the end of training for val data", forward_step_func, valid_it, model,
                                   iteration, process_non_loss_data_func, verbose=False, args=args)
    if args.save and iteration != 0:
        save_checkpoint(iteration, model, optimizer, opt_param_scheduler)
    if args.do_test:
        evaluate_and_print_results("the end of training for test data", forward_step_func, test_it, model, 0,
                                   process_non_loss_data_func, verbose=True, args=args)
    return model, optimizer


def update_train_iters(args):
    if args.train_iters:
        return
    if args.rampup_batch_size is None:
        args.train_iters = args.train_samples // args.global_batch_size
    else:
        iterations, consumed = 0, 0
        while consumed <= int(args.rampup_batch_size[2]):
            update_num_microbatches(consumed, consistency_check=False)
            consumed += get_current_global_batch_size()
            iterations += 1
        update_num_microbatches(0, consistency_check=False)
        iterations += (args.train_samples - consumed) // args.global_batch_size
        args.train_iters = iterations
    print_rank_0(f"setting training iterations to {args.train_iters}")


def _device_for_model():
    return torch.cuda.current_device() if torch.cuda.is_available() else None


def get_model(model_provider_func, model_type=ModelType.encoder_or_decoder, wrap_with_ddp=True, args=None):
    args = args or get_args()
    args.model_type = model_type
    pp = mpu.get_pipeline_model_parallel_world_size()
    if pp > 1 and args.virtual_pipeline_model_parallel_size is not None:
        assert model_type != ModelType.encoder_and_decoder, \
            "Interleaved schedule not supported for model with both encoder and decoder"
        model = []
        for i in range(args.virtual_pipeline_model_parallel_size):
            mpu.set_virtual_pipeline_model_parallel_rank(i)
            m = model_provider_func(pre_process=mpu.is_pipeline_first_stage(),
                                    post_process=mpu.is_pipeline_last_stage())
            m.model_type = model_type
            model.append(m)
    else:
        pre, post = mpu.is_pipeline_first_stage(), mpu.is_pipeline_last_stage()
        if model_type == ModelType.encoder_and_decoder:
            add_enc = add_dec = True
            if pp > 1:
                assert args.pipeline_model_parallel_split_rank is not None, \
                    "Split rank needs to be specified for model with both encoder and decoder"
                rank, split = mpu.get_pipeline_model_parallel_rank(), args.pipeline_model_parallel_split_rank
                pre = rank == 0 or rank == split
                post = rank == split - 1 or rank == pp - 1
                add_enc, add_dec = mpu.is_pipeline_stage_before_split(), mpu.is_pipeline_stage_after_split()
            model = model_provider_func(pre_process=pre, post_process=post, add_encoder=add_enc,
                                        add_decoder=add_dec)
        else:
            model = model_provider_func(pre_process=pre, post_process=post)
        model.model_type = model_type
    if not isinstance(model, list):
        model = [model]
    for m in model:
        for p in m.parameters():
            tp_layers.set_defaults_if_not_set_tensor_model_parallel_attributes(p)
    if mpu.get_data_parallel_rank() == 0:
        n = sum(sum(p.nelement() for p in m.parameters()) for m in model)
        print(" > number of parameters on (tensor, pipeline) model parallel rank ({}, {}): {}".format(
            mpu.get_tensor_model_parallel_rank(), mpu.get_pipeline_model_parallel_rank(), n), flush=True)
    dev = _device_for_model()
    if dev is not None and args.distributed_backend != "gloo":
        for m in model:
            m.cuda(dev)
    if args.fp16 or args.bf16:
        model = [Float16Module(m, args) for m in model]
    if wrap_with_ddp:
        if args.DDP_impl == "torch":
            # torch-DDP semantics on the one gradient engine this framework has (the flat optimizers
            # read the contiguous gradient buffers): gradients in the parameter dtype, averaged over
            # the data-parallel group in 25 MiB buckets while the backward pass is still running
            # (reference: training.py:284-288 wraps in torch.nn.parallel.DistributedDataParallel)
            model = [LocalDDP(m, False, True, fuse_wgrad=False, overlap_grad_reduce=True, bucket_mb=25)
                     for m in model]
            for m in model:
                m.fused_dp_required = False
        elif args.DDP_impl == "local":
            key_fn = make_group_key_fn(getattr(args, "_no_wd_cond", None), getattr(args, "_scale_lr_cond", None),
                                       getattr(args, "_lr_mult", 1.0))
            fused_dp = _fused_dp_enabled(args, dev)
            model = [LocalDDP(m, args.accumulate_allreduce_grads_in_fp32,
                              args.use_contiguous_buffers_in_local_ddp,
                              fuse_wgrad=args.gradient_accumulation_fusion, group_key_fn=key_fn,
                              overlap_grad_reduce=getattr(args, "overlap_grad_reduce", True),
                              bucket_mb=getattr(args, "ddp_bucket_mb", 256), fused_dp=fused_dp) for m in model]
            for m in model:
                m.fused_dp_required = fused_dp and getattr(args, "b200_fused_dp", "auto") == "on"
            if args.data_parallel_random_init:
                for m in model:
                    m.broadcast_params()
        else:
            raise NotImplementedError(f"Unknown DDP implementation specified: {args.DDP_impl}. Exiting.")
    return model


def _fused_dp_enabled(args, dev) -> bool:
    """--b200_fused_dp: in-kernel reduce-scatter / all-gather of the distributed optimizer over peer
    memory (parallel/dp_symm.py). Needs bf16 parameters with fp32 gradient accumulation, Adam, and a
    data-parallel group of 2..8 GPUs on one host. Collective over the data-parallel group."""
    mode = getattr(args, "b200_fused_dp", "auto")
    if mode == "off" or dev is None or args.distributed_backend == "gloo":
        return False
    from .ops import lib as _L
    dp = mpu.get_data_parallel_world_size()
    ok = (_L.available() and args.use_distributed_optimizer and 2 <= dp <= 8 and args.bf16
          and args.accumulate_allreduce_grads_in_fp32 and args.optimizer == "adam")
    if ok:
        from .parallel.dp_symm import group_on_one_host
        ok = group_on_one_host(mpu.get_data_parallel_group())
    if mode == "on" and not ok:
        raise RuntimeError("--b200_fused_dp on needs: CUDA kernels, --use_distributed_optimizer, --bf16 with fp32 "
                           "gradient accumulation, Adam, 2..8 data-parallel ranks on one host")
    return ok


def _get_optimizer_param_scheduler(optimizer, args):
    if args.train_iters or (args.train_iters == 0 and not args.train_samples):
        # train_iters == 0: evaluation-only runs of the task harness (--epochs 0) still build an
        # optimizer and a schedule; give the schedule a nominal length of one step
        if args.lr_decay_iters is None:
            args.lr_decay_iters = max(args.train_iters, 1)
        lr_decay_steps = args.lr_decay_iters * args.global_batch_size
        wd_incr_steps = args.train_iters * args.global_batch_size
        warm = (args.lr_warmup_fraction * lr_decay_steps if args.lr_warmup_fraction is not None
                else args.lr_warmup_iters * args.global_batch_size)
    elif args.train_samples:
        update_train_iters(args)
        if args.lr_decay_samples is None:
            args.lr_decay_samples = args.train_samples
        lr_decay_steps = args.lr_decay_samples
        wd_incr_steps = args.train_samples
        warm = (args.lr_warmup_fraction * lr_decay_steps if args.lr_warmup_fraction is not None
                else args.lr_warmup_samples)
    else:
        raise Exception("either train_iters or train_samples should be provided.")
    return OptimizerParamScheduler(
        optimizer, max_lr=args.lr, min_lr=args.min_lr, lr_warmup_steps=warm, lr_decay_steps=lr_decay_steps,
        lr_decay_style=args.lr_decay_style, start_wd=args.start_weight_decay, end_wd=args.end_weight_decay,
        wd_incr_steps=wd_incr_steps, wd_incr_style=args.weight_decay_incr_style,
        use_checkpoint_opt_param_scheduler=args.use_checkpoint_opt_param_scheduler,
        override_opt_param_scheduler=args.override_opt_param_scheduler)


def setup_model_and_optimizer(model_provider_func, model_type, no_wd_decay_cond=None, scale_lr_cond=None,
                              lr_mult=1.0, args=None):
    args = args or get_args()
    args._no_wd_cond, args._scale_lr_cond, args._lr_mult = no_wd_decay_cond, scale_lr_cond, lr_mult
    model = get_model(model_provider_func, model_type, args=args)
    unwrapped = unwrap_model(model, (torchDDP, LocalDDP, Float16Module))
    optimizer = get_megatron_optimizer(model, no_wd_decay_cond, scale_lr_cond, lr_mult, args=args)
    for k in ("_no_wd_cond", "_scale_lr_cond", "_lr_mult"):   # keep args picklable
        if hasattr(args, k):
            delattr(args, k)
    sched = _get_optimizer_param_scheduler(optimizer, args)
    if args.load is not None:
        timers = get_timers()
        timers("load-checkpoint", log_level=0).start(barrier=True)
        args.iteration = load_checkpoint(model, optimizer, sched, args=args)
        timers("load-checkpoint").stop(barrier=True)
        timers.log(["load-checkpoint"])
        if args.fp16 or args.bf16:
            optimizer.reload_model_params() if args.finetune or args.no_load_optim else None
    else:
        args.iteration = 0
    assert args.DDP_impl == "local" or (len(model) == 1 and mpu.get_pipeline_model_parallel_world_size() == 1)
    if args.iteration == 0 and len(unwrapped) == 1 and hasattr(unwrapped[0], "init_state_dict_from_bert"):
        print_rank_0("Initializing ICT from pretrained BERT model")
        unwrapped[0].init_state_dict_from_bert()
        if args.fp16:
            optimizer.reload_model_params()
    return model, optimizer, sched


def train_step(forward_step_func, data_iterator, model, optimizer, opt_param_scheduler, args=None):
    args = args or get_args()
    timers = get_timers()
    if args.use_contiguous_buffers_in_local_ddp or args.DDP_impl == "torch":
        for part in model:
            part.zero_grad_buffer()
    optimizer.zero_grad()
    timers("forward-backward", log_level=1).start(barrier=args.barrier_with_L1_time)
    fb = get_forward_backward_func(args)
    fwd_bwd_timers = timers if args.timing_log_level > 1 else None
    losses_reduced = fb(forward_step_func, data_iterator, model, optimizer, fwd_bwd_timers, forward_only=False)
    timers("forward-backward").stop()
    if args.empty_unused_memory_level >= 1 and torch.cuda.is_available():
        torch.cuda.empty_cache()
    maybe_inject(args.fault_inject, getattr(args, "curr_iteration", 0), model)
    optimizer.reduce_model_grads(args, timers)
    timers("optimizer", log_level=1).start(barrier=args.barrier_with_L1_time)
    update_successful, grad_norm, num_zeros = optimizer.step(args, timers)
    timers("optimizer").stop()
    from .ops import fp8 as _fp8
    _fp8.advance_step()            # FP8 layers: amax of this step -> scales of the next (no-op without FP8)
    if update_successful:
        optimizer.gather_model_params(args, timers)
        every = getattr(args, "check_dp_consistency_interval", 0)
        if every and hasattr(optimizer, "check_dp_consistency") and \
                (getattr(args, "curr_iteration", 0) + 1) % every == 0:
            optimizer.check_dp_consistency()
        increment = get_num_microbatches() * args.micro_batch_size * args.data_parallel_size
        opt_param_scheduler.step(increment=increment)
        skipped = 0
    else:
        skipped = 1
    if args.empty_unused_memory_level >= 2 and torch.cuda.is_available():
        torch.cuda.empty_cache()
    if mpu.is_pipeline_last_stage(ignore_virtual=True):
        reduced = {}
        for key in losses_reduced[0]:
            vals = [x[key] for x in losses_reduced]
            reduced[key] = sum(vals) / len(vals)
        return reduced, skipped, grad_norm, num_zeros
    return {}, skipped, grad_norm, num_zeros


def model_flops_per_token(args):
    """Training FLOPs per token: 6 * (non-embedding parameters incl. LM head) plus the causal
    attention term 6 * L * s * h (12*L*h*s halved by causality)."""
    h, L, V, s, f = args.hidden_size, args.num_layers, args.padded_vocab_size, args.seq_length, args.ffn_hidden_size
    hn = args.kv_channels
    attn_params = h * (args.num_attention_heads * hn + 2 * args.num_attention_heads_kv * hn) \
        + args.num_attention_heads * hn * h
    mlp_params = h * f * (3 if args.glu_activation else 2)
    n = L * (attn_params + mlp_params) + V * h
    return 6.0 * n + 6.0 * L * s * h


def _peak_tflops():
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            p = json.load(f)
        return float(p.get("bf16_tflops_sustained", 1400.0)), float(p.get("bf16_tflops", 1590.0))
    except OSError:
        return 1400.0, 1590.0


def training_log(loss_dict, total_loss_dict, learning_rate, iteration, loss_scale, report_memory_flag, skipped_iter,
                 grad_norm, params_norm, num_zeros_in_grad, args=None):
    args = args or get_args()
    timers = get_timers()
    writer = get_tensorboard_writer()
    adv, skip_k, nan_k = "advanced iterations", "skipped iterations", "nan iterations"
    total_loss_dict[adv] = total_loss_dict.get(adv, 0) + (0 if skipped_iter else 1)
    total_loss_dict[skip_k] = total_loss_dict.get(skip_k, 0) + skipped_iter
    got_nan = False
    for key in loss_dict:
        if not skipped_iter:
            total_loss_dict[key] = total_loss_dict.get(key, torch.zeros(1, device=loss_dict[key].device)) + loss_dict[key]
    total_loss_dict[nan_k] = total_loss_dict.get(nan_k, 0) + int(got_nan)

    timers_to_log = [
        "forward-backward", "forward-compute", "backward-compute", "batch-generator", "forward-recv",
        "forward-send", "backward-recv", "backward-send", "forward-send-forward-recv",
        "forward-send-backward-recv", "backward-send-forward-recv", "backward-send-backward-recv",
        "forward-backward-send-forward-backward-recv", "layernorm-grads-all-reduce",
        "embedding-grads-all-reduce", "grads-all-reduce", "grads-reduce", "grads-reduce-scatter",
        "params-all-gather", "optimizer-copy-to-main-grad", "optimizer-unscale-and-check-inf",
        "optimizer-clip-main-grad", "optimizer-count-zeros", "optimizer-inner-step",
        "optimizer-copy-main-to-model-params", "optimizer"]
    batch_size = args.micro_batch_size * args.data_parallel_size * get_num_microbatches()
    total_iterations = total_loss_dict[adv] + total_loss_dict[skip_k]

    if writer and iteration % args.tensorboard_log_interval == 0 and is_last_rank():
        if getattr(args, "log_learning_rate_to_tensorboard", True):
            writer.add_scalar("learning-rate", learning_rate, iteration)
            writer.add_scalar("learning-rate vs samples", learning_rate, args.consumed_train_samples)
        if args.log_batch_size_to_tensorboard:
            writer.add_scalar("batch-size", batch_size, iteration)
            writer.add_scalar("batch-size vs samples", batch_size, args.consumed_train_samples)
        for key in loss_dict:
            writer.add_scalar(key, loss_dict[key], iteration)
            writer.add_scalar(key + " vs samples", loss_dict[key], args.consumed_train_samples)
        if getattr(args, "log_loss_scale_to_tensorboard", True):
            writer.add_scalar("loss-scale", loss_scale, iteration)
            writer.add_scalar("loss-scale vs samples", loss_scale, args.consumed_train_samples)
        if args.log_world_size_to_tensorboard:
            writer.add_scalar("world-size", args.world_size, iteration)
            writer.add_scalar("world-size vs samples", args.world_size, args.consumed_train_samples)
        if grad_norm is not None:
            writer.add_scalar("grad-norm", float(grad_norm), iteration)
            writer.add_scalar("grad-norm vs samples", float(grad_norm), args.consumed_train_samples)
        if num_zeros_in_grad is not None:
            writer.add_scalar("num-zeros", float(num_zeros_in_grad), iteration)
            writer.add_scalar("num-zeros vs samples", float(num_zeros_in_grad), args.consumed_train_samples)
        if params_norm is not None:
            writer.add_scalar("params-norm", params_norm, iteration)
            writer.add_scalar("params-norm vs samples", params_norm, args.consumed_train_samples)
        if args.log_memory_to_tensorboard and torch.cuda.is_available():
            ms = torch.cuda.memory_stats()
            writer.add_scalar("mem-reserved-bytes", ms["reserved_bytes.all.current"], iteration)
            writer.add_scalar("mem-allocated-bytes", ms["allocated_bytes.all.current"], iteration)
            writer.add_scalar("mem-allocated-count", ms["allocation.all.current"], iteration)

    if iteration % args.log_interval == 0:
        elapsed = timers("interval-time").elapsed(barrier=True)
        per_iter = elapsed / max(total_iterations, 1)
        tokens = batch_size * args.seq_length
        tok_s = tokens / max(per_iter, 1e-9)
        tflops_gpu = tok_s * model_flops_per_token(args) / 1e12 / args.world_size
        sustained, burst = _peak_tflops()
        if writer and is_last_rank() and args.log_timers_to_tensorboard:
            writer.add_scalar("iteration-time", per_iter, iteration)
        log = " iteration {:8d}/{:8d} |".format(iteration, args.train_iters)
        log += " consumed samples: {:12d} |".format(args.consumed_train_samples)
        log += " elapsed time per iteration (ms): {:.1f} |".format(per_iter * 1000.0)
        log += " learning rate: {:.3E} |".format(learning_rate)
        log += " global batch size: {:5d} |".format(batch_size)
        record = {"iteration": iteration, "ms_per_iteration": per_iter * 1000.0, "tokens_per_s": tok_s,
                  "tflops_per_gpu": tflops_gpu, "mfu_sustained": tflops_gpu / sustained,
                  "mfu_burst": tflops_gpu / burst, "lr": learning_rate}
        for key in total_loss_dict:
            if key in (adv, skip_k, nan_k):
                continue
            avg = total_loss_dict[key].item() / float(max(1, total_loss_dict[adv]))
            if avg > 0.0:
                log += " {}: {:.6E} |".format(key, avg)
            record[key] = avg
            if not math.isfinite(avg):
                total_loss_dict[nan_k] += 1
            total_loss_dict[key] = torch.zeros(1, device=total_loss_dict[key].device)
        log += " loss scale: {:.1f} |".format(float(loss_scale))
        if grad_norm is not None:
            log += " grad norm: {:.3f} |".format(float(grad_norm))
            record["grad_norm"] = float(grad_norm)
        if num_zeros_in_grad is not None:
            log += " num zeros: {:.1f} |".format(float(num_zeros_in_grad))
        if params_norm is not None:
            log += " params norm: {:.3f} |".format(params_norm)
        log += " number of skipped iterations: {:3d} |".format(total_loss_dict[skip_k])
        log += " number of nan iterations: {:3d} |".format(total_loss_dict[nan_k])
        log += " tokens/s: {:.1f} | TFLOP/s/GPU: {:.1f} | MFU: {:.1%} (of measured sustained) |".format(
            tok_s, tflops_gpu, tflops_gpu / sustained)
        total_loss_dict[adv] = total_loss_dict[skip_k] = total_loss_dict[nan_k] = 0
        print_rank_last(log)
        if args.metrics_jsonl and is_last_rank():
            with open(args.metrics_jsonl, "a") as f:
                f.write(json.dumps(record) + "\n")
        if report_memory_flag and learning_rate > 0.0:
            report_memory("(after {} iterations)".format(iteration))
            report_memory_flag = False
        timers.log(timers_to_log, normalizer=args.log_interval)
    return report_memory_flag


def save_checkpoint_and_time(iteration, model, optimizer, opt_param_scheduler):
    timers = get_timers()
    timers("save-checkpoint", log_level=0).start(barrier=True)
    save_checkpoint(iteration, model, optimizer, opt_param_scheduler)
    timers("save-checkpoint").stop(barrier=True)
    timers.log(["save-checkpoint"])


def _train(args, forward_step_func, model, optimizer, opt_param_scheduler, train_data_iterator,
           valid_data_iterator, process_non_loss_data_func):
    timers = get_timers()
    write_args_to_tensorboard()
    for m in model:
        m.train()
    total_loss_dict = {}
    iteration = args.iteration
    timers("interval-time", log_level=0).start(barrier=True)
    print_datetime("before the start of training step")
    report_memory_flag = True
    while iteration < args.train_iters:
        update_num_microbatches(args.consumed_train_samples)
        args.curr_iteration = iteration
        loss_dict, skipped, grad_norm, num_zeros = train_step(forward_step_func, train_data_iterator, model,
                                                              optimizer, opt_param_scheduler, args)
        iteration += 1
        args.consumed_train_samples += mpu.get_data_parallel_world_size() * args.micro_batch_size * \
            get_num_microbatches()
        logging_now = iteration % args.log_interval == 0
        loss_scale = optimizer.get_loss_scale().item() if (logging_now or optimizer.grad_scaler) else 1.0
        params_norm = calc_params_l2_norm(model) if args.log_params_norm else None
        if not logging_now:
            grad_norm_f = None
        else:
            grad_norm_f = float(grad_norm) if grad_norm is not None else None
        report_memory_flag = training_log(loss_dict, total_loss_dict, optimizer.param_groups[0]["lr"], iteration,
                                          loss_scale, report_memory_flag, skipped, grad_norm_f, params_norm,
                                          num_zeros if logging_now else None, args)
        if args.adlr_autoresume and iteration % args.adlr_autoresume_interval == 0:
            check_adlr_autoresume_termination(iteration, model, optimizer, opt_param_scheduler)
        if args.eval_interval and iteration % args.eval_interval == 0 and args.do_valid:
            evaluate_and_print_results("iteration {}".format(iteration), forward_step_func, valid_data_iterator,
                                       model, iteration, process_non_loss_data_func, verbose=False, args=args)
        if hasattr(get_tensorboard_writer(), "flush_all"):
            get_tensorboard_writer().flush_all()
        saved = False
        if args.exit_signal_handler:
            if any(get_signal_handler().signals_received()):
                save_checkpoint_and_time(iteration, model, optimizer, opt_param_scheduler)
                print_datetime("exiting program after receiving SIGTERM.")
                sys.exit()
        if args.save and args.save_interval and iteration % args.save_interval == 0:
            save_checkpoint_and_time(iteration, model, optimizer, opt_param_scheduler)
            saved = True
        if args.exit_duration_in_mins:
            mins = (time.time() - _TRAIN_START_TIME) / 60.0
            flag = torch.tensor([mins > args.exit_duration_in_mins], dtype=torch.int, device=_stat_device())
            if dist.is_initialized():
                dist.all_reduce(flag, op=dist.ReduceOp.MAX)
            if flag.item():
                if args.save and not saved:
                    save_checkpoint_and_time(iteration, model, optimizer, opt_param_scheduler)
                print_datetime("exiting program after {} minutes".format(mins))
                sys.exit()
        if args.exit_interval and iteration % args.exit_interval == 0:
            if args.save and not saved:
                save_checkpoint_and_time(iteration, model, optimizer, opt_param_scheduler)
            if dist.is_initialized():
                dist.barrier()
            print_datetime("exiting program at iteration {}".format(iteration))
            sys.exit()
    return iteration


def evaluate(forward_step_func, data_iterator, model, process_non_loss_data_func, verbose=False, args=None):
    args = args or get_args()
    for m in model:
        m.eval()
    total = {}
    collected = None
    with torch.no_grad():
        it = 0
        while it < args.eval_iters:
            it += 1
            if verbose and it % args.log_interval == 0:
                print_rank_0("Evaluating iter {}/{}".format(it, args.eval_iters))
            fb = get_forward_backward_func(args)
            loss_dicts = fb(forward_step_func, data_iterator, model, optimizer=None, timers=None, forward_only=True)
            if args.empty_unused_memory_level >= 1 and torch.cuda.is_available():
                torch.cuda.empty_cache()
            if mpu.is_pipeline_last_stage(ignore_virtual=True):
                for ld in loss_dicts:
                    for key in ld:
                        total[key] = total.get(key, torch.zeros(1, device=ld[key].device)) + ld[key]
            args.consumed_valid_samples += mpu.get_data_parallel_world_size() * args.micro_batch_size * \
                get_num_microbatches()
        if process_non_loss_data_func is not None and is_last_rank():
            fb = get_forward_backward_func(args)
            collected = fb(forward_step_func, data_iterator, model, optimizer=None, timers=None, forward_only=True,
                           collect_non_loss_data=True)
    for m in model:
        m.train()
    for key in total:
        total[key] /= args.eval_iters * get_num_microbatches()
    return total, collected


def evaluate_and_print_results(prefix, forward_step_func, data_iterator, model, iteration,
                               process_non_loss_data_func, verbose=False, args=None):
    args = args or get_args()
    writer = get_tensorboard_writer()
    total, collected = evaluate(forward_step_func, data_iterator, model, process_non_loss_data_func, verbose, args)
    string = " validation loss at {} | ".format(prefix)
    for key in total:
        v = total[key].item()
        string += "{} value: {:.6E} | ".format(key, v)
        ppl = math.exp(min(20, v))
        string += "{} PPL: {:.6E} | ".format(key, ppl)
        if writer and is_last_rank():
            writer.add_scalar("{} validation".format(key), v, iteration)
            writer.add_scalar("{} validation vs samples".format(key), v, args.consumed_train_samples)
            if args.log_validation_ppl_to_tensorboard:
                writer.add_scalar("{} validation ppl".format(key), ppl, iteration)
                writer.add_scalar("{} validation ppl vs samples".format(key), ppl, args.consumed_train_samples)
    if process_non_loss_data_func is not None and writer and is_last_rank():
        process_non_loss_data_func(collected, iteration, writer)
    length = len(string) + 1
    print_rank_last("-" * length)
    print_rank_last(string)
    print_rank_last("-" * length)


def cyclic_iter(iterable):
    while True:
        for x in iterable:
            yield x


def build_train_valid_test_data_iterators(build_train_valid_test_datasets_provider, args=None, collate_fn=None):
    """Only TP rank 0 builds datasets/loaders; the do_train/valid/test flags are broadcast
    over the TP group (training.py:855-939 behaviour)."""
    args = args or get_args()
    train_dl = valid_dl = test_dl = None
    print_rank_0("> building train, validation, and test datasets ...")
    if args.iteration > 0 and args.consumed_train_samples == 0:
        assert args.train_samples is None, "only backward compatiblity support for iteration-based training"
        args.consumed_train_samples = args.iteration * args.global_batch_size
    if args.iteration > 0 and args.consumed_valid_samples == 0 and args.train_samples is None:
        args.consumed_valid_samples = (args.iteration // args.eval_interval) * args.eval_iters * \
            args.global_batch_size
    if mpu.get_tensor_model_parallel_rank() == 0:
        train_samples = args.train_samples if args.train_samples else (args.train_iters or 0) * args.global_batch_size
        eval_iters = (((args.train_iters or 0) // args.eval_interval + 1) * args.eval_iters) if args.eval_interval else 0
        nums = [train_samples, eval_iters * args.global_batch_size, args.eval_iters * args.global_batch_size]
        print_rank_0(" > datasets target sizes (minimum size):")
        print_rank_0("    train:      {}".format(nums[0]))
        print_rank_0("    validation: {}".format(nums[1]))
        print_rank_0("    test:       {}".format(nums[2]))
        train_ds, valid_ds, test_ds = build_train_valid_test_datasets_provider(nums)
        train_dl = build_pretraining_data_loader(train_ds, args.consumed_train_samples, args)
        valid_dl = build_pretraining_data_loader(valid_ds, args.consumed_valid_samples, args)
        test_dl = build_pretraining_data_loader(test_ds, 0, args)
        flags = torch.tensor([int(train_dl is not None and (args.train_iters or 0) > 0),
                              int(valid_dl is not None and args.eval_iters > 0),
                              int(test_dl is not None and args.eval_iters > 0)], dtype=torch.long,
                             device=_stat_device())
    else:
        flags = torch.zeros(3, dtype=torch.long, device=_stat_device())
    if dist.is_initialized() and mpu.get_tensor_model_parallel_world_size() > 1:
        dist.broadcast(flags, mpu.get_tensor_model_parallel_src_rank(),
                       group=mpu.get_tensor_model_parallel_group())
    args.do_train, args.do_valid, args.do_test = (bool(flags[i].item()) for i in range(3))
    dl_type = args.dataloader_type
    assert dl_type in ("single", "cyclic")

    def mk(dl):
        if dl is None:
            return None
        if dl_type != "single":
            return iter(cyclic_iter(dl))
        # a DataLoader iterator draws its base seed from the default generator when it is created; keep
        # the model's random stream (CPU dropout) independent of it, so that a resumed run, which builds
        # its iterators after the checkpoint's generator state was restored, replays the same numbers
        state = torch.get_rng_state()
        it = iter(dl)
        torch.set_rng_state(state)
        return it

    return mk(train_dl), mk(valid_dl), mk(test_dl)
