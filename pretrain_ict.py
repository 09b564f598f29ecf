"""Pretrain the biencoder retriever with the Inverse Cloze Task: in-batch softmax over
query x context scores, contexts gathered from every data-parallel rank
(pretrain_ict.py:24-176 surface)."""
import math
from functools import partial

import torch
import torch.distributed as dist
import torch.nn.functional as F

from epfl_megatron_b200 import get_args, get_timers, print_rank_0
from epfl_megatron_b200.data.biencoder_dataset_utils import get_ict_batch
from epfl_megatron_b200.initialize import initialize_megatron
from epfl_megatron_b200.models import ModelType
from epfl_megatron_b200.models.biencoder_model import biencoder_model_provider
from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.training import pretrain
from epfl_megatron_b200.utils import average_losses_across_data_parallel_group


def pretrain_ict_model_provider(pre_process=True, post_process=True):
    args = get_args()
    return biencoder_model_provider(
        only_context_model=False, only_query_model=False,
        biencoder_shared_query_context_model=args.biencoder_shared_query_context_model,
        pre_process=pre_process, post_process=post_process)


def get_group_world_size_rank():
    group = mpu.get_data_parallel_group()
    return group, dist.get_world_size(group=group), dist.get_rank(group=group)


class AllgatherFromDataParallelRegion(torch.autograd.Function):
    """Concatenate along dim 0 over the data-parallel group; the backward keeps this rank's rows."""

    @staticmethod
    def forward(ctx, input_):
        assert input_.dim() == 2
        group, world, rank = get_group_world_size_rank()
        parts = [torch.empty_like(input_) for _ in range(world)]
        parts[rank] = input_
        dist.all_gather(parts, input_.contiguous(), group=group)
        return torch.cat(parts, dim=0).contiguous()

    @staticmethod
    def backward(ctx, grad_output):
        group, world, rank = get_group_world_size_rank()
        assert grad_output.shape[0] % world == 0
        return torch.split(grad_output, grad_output.shape[0] // world, dim=0)[rank].contiguous()


def loss_func(output_tensor):
    args = get_args()
    query_logits, context_logits = output_tensor
    micro_batch_size = query_logits.shape[0]
    global_batch_size = dist.get_world_size(group=mpu.get_data_parallel_group()) * micro_batch_size
    all_q = AllgatherFromDataParallelRegion.apply(query_logits)
    all_c = AllgatherFromDataParallelRegion.apply(context_logits)
    scores = torch.matmul(all_q, all_c.t())
    if args.retriever_score_scaling:
        scores = scores / math.sqrt(args.hidden_size)
    softmax_scores = F.log_softmax(scores, dim=1)
    sorted_indices = torch.argsort(softmax_scores, dim=1, descending=True)
    target = torch.arange(global_batch_size, device=scores.device)

    def topk_accuracy(k):
        hits = sum(int(target[i] in sorted_indices[i, :k]) for i in range(global_batch_size))
        return torch.tensor([hits / global_batch_size], dtype=torch.float32, device=scores.device)

    topk_accs = [topk_accuracy(int(k)) for k in args.retriever_report_topk_accuracies]
    loss = F.nll_loss(softmax_scores, target, reduction="mean")
    reduced = average_losses_across_data_parallel_group([loss, *topk_accs])
    # the gradient hook averages over the DP group, the gathered loss already spans it
    loss = loss.float() * mpu.get_data_parallel_world_size()
    stats = {f"top{k}_acc": v * 100 for k, v in zip(args.retriever_report_topk_accuracies, reduced[1:])}
    stats["loss"] = reduced[0]
    return loss, stats


def forward_step(data_iterator, model):
    timers = get_timers()
    timers("batch-generator", log_level=2).start()
    q_tokens, q_mask, c_tokens, c_mask, _ = get_ict_batch(data_iterator)
    timers("batch-generator").stop()
    q_types = torch.zeros_like(q_tokens)
    c_types = torch.zeros_like(c_tokens)
    out = model(q_tokens, q_mask, q_types, c_tokens, c_mask, c_types)
    return out, partial(loss_func)


def train_valid_test_datasets_provider(train_val_test_num_samples):
    from epfl_megatron_b200.data.dataset_utils import build_train_valid_test_datasets
    args = get_args()
    print_rank_0("> building train, validation, and test datasets for BERT ICT...")
    ds = build_train_valid_test_datasets(
        data_prefix=args.data_path, data_impl=args.data_impl, splits_string=args.split,
        train_valid_test_num_samples=train_val_test_num_samples, max_seq_length=args.seq_length,
        masked_lm_prob=args.mask_prob, short_seq_prob=args.short_seq_prob, seed=args.seed,
        skip_warmup=(not args.mmap_warmup), binary_head=False, dataset_type="ict")
    print_rank_0("> finished creating BERT ICT datasets ...")
    return ds


if __name__ == "__main__":
    initialize_megatron(args_defaults={"tokenizer_type": "BertWordPieceLowerCase"})
    pretrain(get_args(), train_valid_test_datasets_provider, pretrain_ict_model_provider,
             ModelType.encoder_or_decoder, forward_step)
