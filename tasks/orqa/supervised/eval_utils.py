"""Average-rank / top-k validation of the retriever over gathered in-batch + negative contexts
(tasks/orqa/supervised/eval_utils.py:17-192 surface)."""
import math
from collections import OrderedDict

import numpy as np
import torch
import torch.distributed as dist
import torch.nn.functional as F

from epfl_megatron_b200 import get_args, print_rank_0
from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.utils import unwrap_model
from finetune_utils import build_data_loader


def _dev():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


def task_collate_fn(batch_data):
    """Stack per-key; negatives of all samples are concatenated; references stay lists."""
    out = OrderedDict()
    for k in batch_data[0]:
        vals = [d[k] for d in batch_data]
        if k == "reference":
            out[k] = vals
        elif k.startswith("neg_context"):
            out[k] = torch.from_numpy(np.concatenate(vals, axis=0))
        else:
            out[k] = torch.from_numpy(np.array(vals))
    return out


def process_batch(batch):
    d = _dev()
    q = (batch["query"].long().to(d), (batch["query_mask"] < 0.5).to(d), batch["query_types"].long().to(d),
         batch["query_pad_mask"].long().to(d))
    c = (batch["context"].long().to(d), (batch["context_mask"] < 0.5).to(d), batch["context_types"].long().to(d),
         batch["context_pad_mask"].long().to(d))
    neg = None
    if "neg_context" in batch:
        neg = (batch["neg_context"].long().to(d), (batch["neg_context_mask"] < 0.5).to(d),
               batch["neg_context_types"].long().to(d))
    return q + c + (neg, batch["reference"])


def accuracy_func_provider(single_dataset_provider, rank0sampler=False):
    args = get_args()
    loaders = []
    for path in args.valid_data:
        ds = single_dataset_provider(path)
        dl = build_data_loader(ds, args.eval_micro_batch_size or args.micro_batch_size,
                               num_workers=args.num_workers, drop_last=False, task_collate_fn=task_collate_fn)
        loaders.append((ds.dataset_name, dl))

    def metrics_func(model, epoch, output_predictions=False):
        print_rank_0("calculating metrics by accuracy func in ORQA...")
        for name, dl in loaders:
            stats = retrieval_loss(model, dl)
            print_rank_0(f"epoch:{epoch} name:{name} " + " ".join(f"{k}: {v:.4f}" for k, v in stats.items()))
    return metrics_func


def retrieval_loss(model, dataloader):
    args = get_args()
    total = 0
    topk = {k: 0.0 for k in args.retriever_report_topk_accuracies}
    rank_sum = 0.0
    for m in model:
        m.eval()
    um = unwrap_model(model[0])
    with torch.no_grad():
        for batch in dataloader:
            (q_tok, q_mask, q_types, _, c_tok, c_mask, c_types, _, neg, _ref) = process_batch(batch)
            if neg is not None:
                c_tok = torch.cat([c_tok, neg[0]])
                c_mask = torch.cat([c_mask, neg[1]])
                c_types = torch.cat([c_types, neg[2]])
            q_logits, c_logits = um(q_tok, q_mask, q_types, c_tok, c_mask, c_types)
            scores = q_logits @ c_logits.t()
            if args.retriever_score_scaling:
                scores = scores / math.sqrt(args.hidden_size)
            n = q_logits.shape[0]
            target = torch.arange(n, device=scores.device)
            order = torch.argsort(F.log_softmax(scores, dim=1), dim=1, descending=True)
            pos = (order == target[:, None]).float().argmax(dim=1)
            rank_sum += float((pos + 1).sum())
            for k in topk:
                topk[k] += float((pos < k).sum())
            total += n
    for m in model:
        m.train()
    t = torch.tensor([rank_sum, total] + [topk[k] for k in topk], dtype=torch.float64, device=_dev())
    if dist.is_initialized():
        dist.all_reduce(t, group=mpu.get_data_parallel_group())
    total = max(float(t[1]), 1.0)
    out = {"rank": float(t[0]) / total}
    for i, k in enumerate(topk):
        out[f"top{k}_acc"] = 100.0 * float(t[2 + i]) / total
    return out
