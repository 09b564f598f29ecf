"""Zero-shot GPT evaluation: WikiText-103 perplexity and LAMBADA last-word accuracy
(tasks/zeroshot_gpt/evaluate.py:28-211 surface). Pipeline aware: activations are passed between
stages with the p2p wrappers and the last stage reduces the metric over the data-parallel group."""
import math

import torch

from epfl_megatron_b200 import get_args, get_tokenizer, is_last_rank, print_rank_0
from epfl_megatron_b200.checkpointing import load_checkpoint
from epfl_megatron_b200.models import GPTModel, LlamaModel, FalconModel
from epfl_megatron_b200.p2p_communication import recv_forward, send_forward
from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.training import get_model
from epfl_megatron_b200.utils import get_ltor_masks_and_position_ids, unwrap_model
from finetune_utils import build_data_loader
from .datasets import build_dataset


def _dev():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


def get_model_provider(eval_metric):
    def model_provider(pre_process=True, post_process=True):
        args = get_args()
        if eval_metric == "loss":
            parallel_output = True
        elif eval_metric == "accuracy":
            parallel_output = False
        else:
            raise NotImplementedError(f"output type for {eval_metric} evaluation metric is not supported.")
        print_rank_0("building GPT model ...")
        name = getattr(args, "model_name", "gpt")
        cls = {"falcon": FalconModel, "llama": LlamaModel, "llama2": LlamaModel, "codellama": LlamaModel}.get(
            name, GPTModel)
        return cls(num_tokentypes=0, parallel_output=parallel_output, pre_process=pre_process,
                   post_process=post_process)
    return model_provider


def process_batch(batch):
    args, tok = get_args(), get_tokenizer()
    loss_mask = batch["pad_mask"].long().to(_dev()).contiguous().byte()
    tokens_ = batch["text"].long().to(_dev()).contiguous()
    labels, tokens = tokens_[:, 1:].contiguous(), tokens_[:, :-1].contiguous()
    mask, _, position_ids = get_ltor_masks_and_position_ids(tokens, tok.eod, args.reset_position_ids,
                                                            args.reset_attention_mask, args.eod_mask_loss)
    return tokens, labels, mask, position_ids, loss_mask


def forward_step(batch, model, eval_metric):
    """-> metric contribution of this batch on the last stage, else None."""
    tokens, labels, mask, position_ids, loss_mask = process_batch(batch)
    args = get_args()
    args.micro_batch_size = len(labels)
    input_tensor = recv_forward()
    um = unwrap_model(model)
    um.set_input_tensor(input_tensor)
    output = model(tokens, position_ids, mask)
    send_forward(output)
    if not mpu.is_pipeline_last_stage():
        return None
    if eval_metric == "loss":
        from epfl_megatron_b200.parallel.cross_entropy import vocab_parallel_cross_entropy
        losses = vocab_parallel_cross_entropy(output.contiguous().float(), labels.contiguous())
        return torch.sum(losses.view(-1) * loss_mask.contiguous().view(-1).float())
    if eval_metric == "accuracy":
        pred = torch.argmax(output, -1)
        correct = (pred == labels).float()
        correct[(1 - loss_mask).bool()] = 1       # only the masked-in (target) positions can be wrong
        return correct.prod(-1).sum()
    raise NotImplementedError(f"forward method for evaluation metric {eval_metric} is not implemented.")


def evaluate(data_loader, model, eval_metric):
    args = get_args()
    model.eval()
    total = torch.zeros((), device=_dev())
    with torch.no_grad():
        for it, batch in enumerate(data_loader):
            if it % args.log_interval == 0:
                print_rank_0(f"> working on iteration: {it}")
            out = forward_step(batch, model, eval_metric)
            if mpu.is_pipeline_last_stage():
                if torch.distributed.is_initialized():
                    torch.distributed.all_reduce(out, group=mpu.get_data_parallel_group())
                total += out
    return total


def evaluate_and_print_results(task, data_loader, model, eval_metric):
    output = evaluate(data_loader, model, eval_metric)
    string = f" validation results on {task} | "
    if is_last_rank():
        if eval_metric == "loss":
            ds = data_loader.dataset
            n_tok, n_orig = ds.num_tokenized_tokens, ds.num_original_tokens
            val_loss = float(output) / (n_tok - 1)
            ppl = math.exp(min(20, val_loss))
            ratio = (n_tok - 1) / (n_orig - 1)
            adj_ppl = math.exp(min(20, val_loss * ratio))
            string += (f"avg loss: {val_loss:.4E} | ppl: {ppl:.4E} | adjusted ppl: {adj_ppl:.4E} | "
                       f"token ratio: {ratio} |")
        elif eval_metric == "accuracy":
            n = len(data_loader.dataset)
            string += f"number correct: {float(output):.4E} | total examples: {n:.4E} | avg accuracy: {float(output) / n:.4E}"
        else:
            raise NotImplementedError(f"evaluation method for {eval_metric} metric is not implemented yet.")
        print("-" * (len(string) + 1))
        print(string)
        print("-" * (len(string) + 1))
    return float(output)


def main():
    args = get_args()
    if args.num_layers_per_virtual_pipeline_stage is not None:
        print("Interleaved pipeline schedule is not yet supported for text generation.")
        return
    if args.task == "LAMBADA":
        eval_metric = "accuracy"
    elif args.task == "WIKITEXT103":
        eval_metric = "loss"
    else:
        raise NotImplementedError(f"{args.task} task is not implemented.")
    model = get_model(get_model_provider(eval_metric), wrap_with_ddp=False)
    if args.load is not None:
        load_checkpoint(model, None, None)
    assert len(model) == 1
    model = model[0]
    dataset = build_dataset(args.task)
    loader = build_data_loader(dataset, args.micro_batch_size, args.num_workers, drop_last=False)
    evaluate_and_print_results(args.task, loader, model, eval_metric)
    print_rank_0("done :-)")
