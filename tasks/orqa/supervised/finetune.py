"""Supervised finetuning of the biencoder retriever on Natural Questions
(tasks/orqa/supervised/finetune.py:26-243 surface): in-batch negatives gathered over the
data-parallel group plus optional hard negatives."""
import math
from functools import partial

import torch
import torch.distributed as dist
import torch.nn.functional as F

from epfl_megatron_b200 import get_args, get_timers, get_tokenizer, print_rank_0
from epfl_megatron_b200.models import ModelType
from epfl_megatron_b200.models.biencoder_model import biencoder_model_provider
from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.utils import average_losses_across_data_parallel_group
from finetune_utils import finetune
from orqa.supervised.eval_utils import accuracy_func_provider, process_batch, task_collate_fn


def get_group_world_size_rank():
    """-> (data-parallel group, this rank's index in it, its size)."""
    group = mpu.get_data_parallel_group()
    return group, dist.get_rank(group=group), dist.get_world_size(group=group)


def check_and_append_tensor_for_gather(group, rank, world_size, input_):
    """Pad dim 0 to the longest over the group so all_gather sees equal shapes."""
    n = torch.tensor([[input_.size(0)]], device=input_.device)
    sizes = [torch.empty_like(n) for _ in range(world_size)]
    dist.all_gather(sizes, n, group=group)
    longest = int(torch.cat(sizes).max())
    if longest > input_.size(0):
        pad = (0,) * (input_.dim() * 2 - 1) + (longest - input_.size(0),)
        input_ = F.pad(input_, pad)
    return input_


def orqa(Dataset):
    def cross_entropy_forward_step(batch, model):
        timers = get_timers()
        timers("batch generator", log_level=2).start()
        try:
            batch_ = next(batch)
        except TypeError:
            batch_ = batch
        group, rank, world = get_group_world_size_rank()
        (q_tok, q_mask, q_types, q_pad, c_tok, c_mask, c_types, c_pad, neg, _ref) = process_batch(batch_)
        timers("batch generator").stop()
        local_bs = q_tok.shape[0]
        if neg is not None:
            if world > 1:
                neg = tuple(check_and_append_tensor_for_gather(group, rank, world, t) for t in neg)
            c_tok, c_mask, c_types = (torch.cat([a, b]) for a, b in zip((c_tok, c_mask, c_types), neg))
        out = model(q_tok, q_mask, q_types, c_tok, c_mask, c_types)
        return out, partial(cross_entropy_loss_func, q_tok, c_tok)

    def cross_entropy_loss_func(query_tokens, context_tokens, output_tensor):
        args = get_args()
        q_logits, c_logits = output_tensor
        group, rank, world = get_group_world_size_rank()
        local_q, local_c = q_logits.shape[0], c_logits.shape[0]
        if world > 1:
            qs = [torch.empty_like(q_logits) for _ in range(world)]
            cs = [torch.empty_like(c_logits) for _ in range(world)]
            dist.all_gather(qs, q_logits.detach().contiguous(), group=group)
            dist.all_gather(cs, c_logits.detach().contiguous(), group=group)
            qs[rank], cs[rank] = q_logits, c_logits          # keep this rank's rows differentiable
            all_q, all_c = torch.cat(qs), torch.cat(cs)
        else:
            all_q, all_c = q_logits, c_logits
        scores = all_q @ all_c.t()
        if args.retriever_score_scaling:
            scores = scores / math.sqrt(args.hidden_size)
        lp = F.log_softmax(scores, dim=1)
        # the positive of query i of rank r sits at r * local_c + i
        target = torch.cat([torch.arange(local_q, device=scores.device) + r * local_c for r in range(world)])
        order = torch.argsort(lp, dim=1, descending=True)
        pos = (order == target[:, None]).float().argmax(dim=1)
        accs = [((pos < k).float().mean() * 100).view(1) for k in args.retriever_report_topk_accuracies]
        loss = F.nll_loss(lp, target, reduction="mean")
        reduced = average_losses_across_data_parallel_group([loss, *accs])
        stats = {f"top{k}_acc": v for k, v in zip(args.retriever_report_topk_accuracies, reduced[1:])}
        stats["loss"] = reduced[0]
        return loss * world, stats

    def train_valid_datasets_provider():
        args, tok = get_args(), get_tokenizer()
        return (Dataset("training", args.train_data, tok, args.retriever_seq_length, evaluate=False),
                Dataset("validation", args.valid_data, tok, args.retriever_seq_length, evaluate=True))

    def model_provider(pre_process=True, post_process=True):
        args = get_args()
        print_rank_0(f"building retriever model for {args.task} ...")
        return biencoder_model_provider(
            only_context_model=False, only_query_model=False,
            biencoder_shared_query_context_model=args.biencoder_shared_query_context_model,
            pre_process=pre_process, post_process=post_process)

    def single_dataset_provider(datapath):
        args, tok = get_args(), get_tokenizer()
        name = datapath[0].split("/")[-1].split(".")[0] if isinstance(datapath, list) else \
            datapath.split("/")[-1].split(".")[0]
        paths = datapath if isinstance(datapath, list) else [datapath]
        return Dataset(name, paths, tok, args.retriever_seq_length, evaluate=True)

    def metrics_func_provider():
        return accuracy_func_provider(single_dataset_provider)

    finetune(train_valid_datasets_provider, model_provider, model_type=ModelType.encoder_or_decoder,
             forward_step=cross_entropy_forward_step, end_of_epoch_callback_provider=metrics_func_provider,
             task_collate_fn=task_collate_fn)


def main():
    args = get_args()
    if args.task == "RET-FINETUNE-NQ":
        from orqa.supervised.data import NQSupervisedDataset as Dataset
    else:
        raise NotImplementedError(f"ORQA task {args.task} is not implemented.")
    orqa(Dataset)
