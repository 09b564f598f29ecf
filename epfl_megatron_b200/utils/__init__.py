"""General utilities (megatron/utils.py:23-229 surface)."""
from __future__ import annotations

import os
import sys

import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel as torchDDP

from .. import global_vars
from ..parallel import state as mpu
from ..parallel.layers import param_is_not_tensor_parallel_duplicate


def unwrap_model(model, module_instances=(torchDDP,)):
    from ..schedules import unwrap_model as _u
    return _u(model, module_instances)


def unwrap_to(model, attribute):
    """Peel wrappers (DDP, Float16Module, ...) until the module that has ``attribute``."""
    while not hasattr(model, attribute):
        model = model.module
    return model


def calc_params_l2_norm(model):
    """L2 norm of all (non-duplicated) parameters across the model-parallel group."""
    from ..models.module import param_is_not_shared
    args = global_vars.get_args()
    if not isinstance(model, list):
        model = [model]
    params = []
    for m in model:
        for p in m.parameters():
            if param_is_not_shared(p) and param_is_not_tensor_parallel_duplicate(p):
                params.append(p.data.float() if args.bf16 else p.data)
    if params:
        norms = torch._foreach_norm(params)
        sq = torch.stack([n.float() ** 2 for n in norms]).sum().view(1)
    else:
        sq = torch.zeros(1)
    if dist.is_initialized():
        dist.all_reduce(sq, group=mpu.get_model_parallel_group())
    return sq.item() ** 0.5


def average_losses_across_data_parallel_group(losses):
    """Mean over the DP group of a list of scalar losses (logging only)."""
    avg = torch.cat([l.clone().detach().view(1) for l in losses])
    if dist.is_initialized() and mpu.get_data_parallel_world_size() > 1:
        dist.all_reduce(avg, group=mpu.get_data_parallel_group())
        avg = avg / dist.get_world_size(group=mpu.get_data_parallel_group())
    return avg


def report_memory(name):
    if not torch.cuda.is_available():
        return
    mb = 1024.0 * 1024.0
    s = name + " memory (MB)"
    s += " | allocated: {}".format(torch.cuda.memory_allocated() / mb)
    s += " | max allocated: {}".format(torch.cuda.max_memory_allocated() / mb)
    s += " | reserved: {}".format(torch.cuda.memory_reserved() / mb)
    s += " | max reserved: {}".format(torch.cuda.max_memory_reserved() / mb)
    if mpu.get_data_parallel_rank() == 0:
        print("[Rank {}] {}".format(dist.get_rank() if dist.is_initialized() else 0, s), flush=True)


def print_params_min_max_norm(optimizer, iteration):
    rank = dist.get_rank() if dist.is_initialized() else 0
    out = "iteration, rank, index, tensor-model-parallel, min, max, norm\n"
    for i, p in enumerate(optimizer.get_parameters()):
        out += "{:7d}, {:4d}, {:4d}, {:2d}, {:.6E}, {:.6E}, {:.6E}\n".format(
            iteration, rank, i, int(getattr(p, "tensor_model_parallel", False)),
            p.data.min().item(), p.data.max().item(), torch.linalg.norm(p.data.float()).item())
    print(out, flush=True)


def check_adlr_autoresume_termination(iteration, model, optimizer, opt_param_scheduler):
    from ..checkpointing import save_checkpoint
    args = global_vars.get_args()
    autoresume = global_vars.get_adlr_autoresume()
    if dist.is_initialized():
        dist.barrier()
    if autoresume is not None and autoresume.termination_requested():
        if args.save:
            save_checkpoint(iteration, model, optimizer, opt_param_scheduler)
        print_rank_0(">>> autoresume termination request found!")
        if not dist.is_initialized() or dist.get_rank() == 0:
            autoresume.request_resume()
        print_rank_0(">>> training terminated. Returning")
        sys.exit(0)


def get_ltor_masks_and_position_ids(data, eod_token, reset_position_ids, reset_attention_mask, eod_mask_loss,
                                    build_attention_mask=True):
    """Left-to-right masks and positions for a [b, s] token batch, with optional resets at
    end-of-document tokens. ``build_attention_mask=False`` skips the [b,1,s,s] boolean mask
    (the fused causal attention never reads it; the reference materialises it on every
    micro-batch, utils.py:152-154)."""
    b, s = data.size()
    attention_mask = None
    if build_attention_mask:
        att_b = b if reset_attention_mask else 1
        attention_mask = torch.tril(torch.ones((att_b, s, s), device=data.device)).view(att_b, 1, s, s)
    loss_mask = torch.ones(data.size(), dtype=torch.float, device=data.device)
    if eod_mask_loss:
        loss_mask[data == eod_token] = 0.0
    position_ids = torch.arange(s, dtype=torch.long, device=data.device).unsqueeze(0).expand_as(data)
    if reset_position_ids:
        position_ids = position_ids.clone()
    if reset_position_ids or reset_attention_mask:
        for bi in range(b):
            eod_index = position_ids[bi, data[bi] == eod_token]
            if reset_position_ids:
                eod_index = eod_index.clone()
            prev = 0
            for j in range(eod_index.size()[0]):
                i = eod_index[j]
                if reset_attention_mask and attention_mask is not None:
                    attention_mask[bi, 0, (i + 1):, :(i + 1)] = 0
                if reset_position_ids:
                    position_ids[bi, (i + 1):] -= (i + 1 - prev)
                    prev = i + 1
    if attention_mask is not None:
        attention_mask = attention_mask < 0.5
    return attention_mask, loss_mask, position_ids


def print_rank_0(message):
    if dist.is_initialized():
        if dist.get_rank() == 0:
            print(message, flush=True)
    else:
        print(message, flush=True)


def is_last_rank():
    return not dist.is_initialized() or dist.get_rank() == dist.get_world_size() - 1


def print_rank_last(message):
    if is_last_rank():
        print(message, flush=True)


def print_all_nodes(message):
    if dist.is_initialized():
        local = int(os.environ.get("LOCAL_WORLD_SIZE", "1"))
        if (dist.get_rank() + 1) % local == 0:
            print(message, flush=True)
    else:
        print(message, flush=True)
