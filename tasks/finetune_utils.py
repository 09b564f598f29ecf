"""Epoch-based finetuning loop shared by the tasks (tasks/finetune_utils.py:23-309 surface)."""
import sys
from functools import partial

import torch

from epfl_megatron_b200 import get_args, get_num_microbatches, get_timers, print_rank_0
from epfl_megatron_b200 import training as mtraining
from epfl_megatron_b200.checkpointing import load_checkpoint, save_checkpoint
from epfl_megatron_b200.models import ModelType
from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.utils import average_losses_across_data_parallel_group, calc_params_l2_norm


def _dev():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


def process_batch(batch, is_fp16: bool):
    d = _dev()
    tokens = batch["text"].long().to(d).contiguous()
    types = batch["types"].long().to(d).contiguous()
    labels = batch["label"].long().to(d).contiguous()
    mask = batch["padding_mask"].float().to(d).contiguous()
    if is_fp16:
        mask = mask.half()
    return tokens, types, labels, mask


def cross_entropy_loss_func(labels, output_tensor):
    loss = torch.nn.functional.cross_entropy(output_tensor.contiguous().float(), labels)
    avg = average_losses_across_data_parallel_group([loss])
    return loss, {"lm loss": avg[0]}


def _cross_entropy_forward_step(batch, model):
    timers = get_timers()
    args = get_args()
    timers("batch-generator", log_level=2).start()
    try:
        batch_ = next(batch)
    except TypeError:
        batch_ = batch
    tokens, types, labels, mask = process_batch(batch_, args.fp16)
    timers("batch-generator").stop()
    return model(tokens, mask, tokentype_ids=types), partial(cross_entropy_loss_func, labels)


def build_data_loader(dataset, micro_batch_size, num_workers, drop_last, task_collate_fn=None):
    """Per-rank loader: a DistributedSampler over the data-parallel group."""
    sampler = torch.utils.data.distributed.DistributedSampler(
        dataset, num_replicas=mpu.get_data_parallel_world_size(), rank=mpu.get_data_parallel_rank())
    return torch.utils.data.DataLoader(dataset, batch_size=micro_batch_size, sampler=sampler, shuffle=False,
                                       num_workers=num_workers, drop_last=drop_last, pin_memory=True,
                                       collate_fn=task_collate_fn)


def _build_infinite_size_dataloader(dataloader):
    while True:
        yield from iter(dataloader)


def _build_train_valid_dataloaders(train_dataset, valid_dataset, task_collate_fn=None):
    args = get_args()
    print_rank_0("building train and validation dataloaders ...")
    train = build_data_loader(train_dataset, args.micro_batch_size, args.num_workers, not args.keep_last,
                              task_collate_fn)
    args.train_iters_per_epoch = len(train)
    args.train_iters = args.epochs * args.train_iters_per_epoch
    valid = _build_infinite_size_dataloader(
        build_data_loader(valid_dataset, args.micro_batch_size, args.num_workers, not args.keep_last,
                          task_collate_fn))
    # multiple-choice datasets fold several candidates into one sample
    args.orig_micro_batch_size, args.orig_global_batch_size = args.micro_batch_size, args.global_batch_size
    if hasattr(train_dataset, "sample_multiplier"):
        args.micro_batch_size *= train_dataset.sample_multiplier
        args.global_batch_size *= train_dataset.sample_multiplier
    return train, valid


def _train(model, optimizer, opt_param_scheduler, forward_step, train_dataloader, valid_dataloader,
           end_of_epoch_callback, args):
    timers = get_timers()
    assert get_num_microbatches() == 1, "finetuning with gradient accumulation is not supported"
    for m in model:
        m.train()
    losses_sum = {}
    start_epoch = args.iteration // args.train_iters_per_epoch
    start_iteration = args.iteration % args.train_iters_per_epoch
    iteration = args.iteration
    report_memory = True
    timers("interval-time", log_level=0).start(barrier=True)
    for epoch in range(start_epoch, args.epochs):
        print_rank_0(f"working on epoch {epoch + 1} ...")
        train_dataloader.sampler.set_epoch(args.seed + epoch)
        for it_, batch in enumerate(train_dataloader):
            if it_ < start_iteration:
                continue
            start_iteration = 0
            losses, skipped, grad_norm, num_zeros = mtraining.train_step(
                forward_step, batch, model, optimizer, opt_param_scheduler, args)
            iteration += 1
            params_norm = calc_params_l2_norm(model) if args.log_params_norm else None
            report_memory = mtraining.training_log(
                losses, losses_sum, optimizer.param_groups[0]["lr"], iteration,
                float(optimizer.get_loss_scale()), report_memory, skipped, grad_norm, params_norm, num_zeros)
            saved = False
            if args.save and args.save_interval and iteration % args.save_interval == 0:
                save_checkpoint(iteration, model, optimizer, opt_param_scheduler)
                saved = True
            if args.eval_interval and iteration % args.eval_interval == 0:
                mtraining.evaluate_and_print_results(f"iteration {iteration}", forward_step, valid_dataloader,
                                                     model, iteration, None, False)
            if args.exit_interval and iteration % args.exit_interval == 0:
                if not saved:
                    save_checkpoint(iteration, model, optimizer, opt_param_scheduler)
                if torch.distributed.is_initialized():
                    torch.distributed.barrier()
                print_rank_0(f"exiting program at iteration {iteration}")
                sys.exit()
        if args.save:
            save_checkpoint(iteration, model, optimizer, opt_param_scheduler)
        if end_of_epoch_callback is not None:
            end_of_epoch_callback(model, epoch)


def finetune(train_valid_datasets_provider, model_provider, model_type=ModelType.encoder_or_decoder,
             forward_step=_cross_entropy_forward_step, end_of_epoch_callback_provider=None,
             task_collate_fn=None):
    """Dataloaders -> model + optimizer -> optional pretrained checkpoint -> epochs (or, with
    --epochs 0, one evaluation pass through the end-of-epoch callback)."""
    args = get_args()
    timers = get_timers()
    assert args.rampup_batch_size is None, "batch size scaling is not supported for finetuning"
    timers("train/valid/test dataset/dataloder", log_level=0).start()
    if args.epochs > 0:
        train_ds, valid_ds = train_valid_datasets_provider()
        train_dl, valid_dl = _build_train_valid_dataloaders(train_ds, valid_ds, task_collate_fn)
    else:
        # evaluation only (--epochs 0): the metric callbacks size their loaders from the batch sizes the
        # user gave, which the training branch records while it builds its loaders
        args.train_iters = 0
        args.orig_micro_batch_size, args.orig_global_batch_size = args.micro_batch_size, args.global_batch_size
        train_dl = valid_dl = None
    timers("train/valid/test dataset/dataloder").stop()
    callback = end_of_epoch_callback_provider() if end_of_epoch_callback_provider is not None else None
    model, optimizer, sched = mtraining.setup_model_and_optimizer(model_provider, model_type)
    if args.iteration == 0 and args.pretrained_checkpoint is not None:
        keep_load, keep_rng = args.load, args.no_load_rng
        args.load, args.no_load_rng = args.pretrained_checkpoint, True
        load_checkpoint(model, None, None)
        args.load, args.no_load_rng = keep_load, keep_rng
        optimizer.reload_model_params()
    print_rank_0("done with setups ...\ntraining ...")
    if args.epochs > 0:
        _train(model, optimizer, sched, forward_step, train_dl, valid_dl, callback, args)
    elif callback is not None:
        print_rank_0("evaluation only mode, setting epoch to -1")
        callback(model, epoch=-1, output_predictions=True)
    print_rank_0("done :-)")
