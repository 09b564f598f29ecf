"""Optimizer factory (megatron/optimizer/__init__.py:13-144 surface)."""
from __future__ import annotations

from .. import global_vars
from .grad_scaler import ConstantGradScaler, DynamicGradScaler
from .optimizer import (DistributedOptimizer, FlatOptimizer, Float16OptimizerWithFloat16Params,
                        FP32Optimizer, MegatronOptimizer)


def make_group_key_fn(no_weight_decay_cond=None, scale_lr_cond=None, lr_mult=1.0):
    """(wd_mult, lr_mult) per parameter. Default rule: biases and 1-D parameters are not
    decayed; ``scale_lr_cond`` selects parameters whose LR is multiplied by ``lr_mult``."""
    def key(name, param):
        if no_weight_decay_cond is not None:
            no_wd = no_weight_decay_cond(name, param)
        else:
            no_wd = name.endswith(".bias") or len(param.shape) == 1
        scale = scale_lr_cond(name, param) if scale_lr_cond is not None else False
        return (0.0 if no_wd else 1.0, lr_mult if scale else 1.0)
    return key


def get_param_groups(modules, no_weight_decay_cond=None, scale_lr_cond=None, lr_mult=1.0):
    """torch.optim-style param groups (for users that bring their own optimizer)."""
    key_fn = make_group_key_fn(no_weight_decay_cond, scale_lr_cond, lr_mult)
    buckets = {}
    for module in modules:
        for name, p in module.named_parameters():
            if p.requires_grad:
                buckets.setdefault(key_fn(name, p), []).append(p)
    return [{"params": ps, "wd_mult": k[0], "lr_mult": k[1]} for k, ps in sorted(buckets.items(), reverse=True)]


def get_megatron_optimizer(model, no_weight_decay_cond=None, scale_lr_cond=None, lr_mult=1.0, args=None):
    args = args or global_vars.get_args()
    cfg = dict(optimizer=args.optimizer, lr=args.lr or 0.0, weight_decay=args.weight_decay,
               adam_beta1=args.adam_beta1, adam_beta2=args.adam_beta2, adam_eps=args.adam_eps,
               sgd_momentum=args.sgd_momentum, clip_grad=args.clip_grad,
               log_num_zeros_in_grad=args.log_num_zeros_in_grad)
    if args.optimizer not in ("adam", "sgd"):
        raise Exception(f"{args.optimizer} optimizer is not supported.")
    grad_scaler = None
    if args.fp16 or args.bf16 or args.use_distributed_optimizer:
        if args.loss_scale:
            grad_scaler = ConstantGradScaler(args.loss_scale)
        elif args.fp16:
            grad_scaler = DynamicGradScaler(initial_scale=args.initial_loss_scale,
                                            min_scale=args.min_loss_scale, growth_factor=2.0,
                                            backoff_factor=0.5, growth_interval=args.loss_scale_window,
                                            hysteresis=args.hysteresis)
    if args.use_distributed_optimizer:
        return DistributedOptimizer(model, cfg, grad_scaler)
    if args.fp16 or args.bf16:
        return Float16OptimizerWithFloat16Params(model, cfg, grad_scaler)
    return FP32Optimizer(model, cfg)
