"""Accuracy evaluation for the classification tasks, pipeline aware
(tasks/eval_utils.py:18-181 surface)."""
import os
import time
from functools import partial

import torch

from epfl_megatron_b200 import get_args, print_rank_last
from epfl_megatron_b200 import is_last_rank
from epfl_megatron_b200.parallel import state as mpu
from epfl_megatron_b200.schedules import get_forward_backward_func
from finetune_utils import build_data_loader, process_batch


def accuracy_func_provider(single_dataset_provider):
    """-> metrics(model, epoch, output_predictions=False) printing per-dataset and overall accuracy."""
    args = get_args()
    loaders = []
    for path in args.valid_data:
        ds = single_dataset_provider(path)
        dl = build_data_loader(ds, args.orig_micro_batch_size, num_workers=args.num_workers,
                               drop_last=(mpu.get_data_parallel_world_size() > 1))
        loaders.append((ds.dataset_name, dl))

    def metrics_func(model, epoch, output_predictions=False):
        print_rank_last("calculating metrics ...")
        correct = total = 0
        named = {} if output_predictions else None
        if output_predictions:
            assert mpu.get_data_parallel_world_size() == 1
        for name, dl in loaders:
            out = calculate_correct_answers(name, model, dl, epoch, output_predictions)
            if output_predictions:
                c, t, preds = out
                named[name] = preds
            else:
                c, t = out
            correct, total = correct + c, total + t
        if is_last_rank():
            print(" >> |epoch: {}| overall: correct / total = {} / {} = {:.4f} %".format(
                epoch, correct, total, 100.0 * correct / max(total, 1)))
        if output_predictions and is_last_rank():
            # next to the evaluated checkpoint (the reference insists on --load; an evaluation-only run
            # usually names its weights with --pretrained_checkpoint)
            where = args.load or args.pretrained_checkpoint
            assert where is not None, "predictions need --load or --pretrained_checkpoint to be written to"
            torch.save(named, os.path.join(where, "predictions.pt"))
    return metrics_func


def calculate_correct_answers(name, model, dataloader, epoch, output_predictions):
    """Forward-only schedule over the loader; the last pipeline stage counts correct argmaxes."""
    args = get_args()
    fb = get_forward_backward_func()
    t0 = time.time()
    for m in model:
        m.eval()
    saved = (args.micro_batch_size, args.global_batch_size)
    ds = dataloader.dataset
    mult = getattr(ds, "sample_multiplier", 1)
    micro = args.orig_micro_batch_size * mult
    args.micro_batch_size, args.global_batch_size = micro, micro * mpu.get_data_parallel_world_size()
    softmaxes, labels_all, ids_all = [], [], []

    def loss_func(output_predictions, labels, output_tensor):
        logits = output_tensor
        d = {}
        if output_predictions:
            d["softmaxes"] = torch.softmax(logits.float(), dim=-1).cpu().numpy().tolist()
            d["labels"] = labels.cpu().numpy().tolist()
        pred = torch.argmax(logits, dim=-1)
        d["total"], d["correct"] = labels.size(0), int((pred == labels).sum().item())
        return 0, d

    def forward_step(batch, model):
        try:
            b = next(batch)
        except TypeError:
            b = batch
        tokens, types, labels, mask = process_batch(b, args.fp16)
        return model(tokens, mask, tokentype_ids=types), partial(loss_func, output_predictions, labels)

    correct = total = 0
    with torch.no_grad():
        for batch in dataloader:
            actual = len(batch["label"])
            args.micro_batch_size = actual * mult
            args.global_batch_size = args.micro_batch_size * mpu.get_data_parallel_world_size()
            dicts = fb(forward_step, batch, model, optimizer=None, timers=None, forward_only=True)
            for d in dicts:
                if output_predictions:
                    softmaxes.extend(d["softmaxes"])
                    labels_all.extend(d["labels"])
                    ids_all.extend(batch["uid"].cpu().numpy().tolist())
                total += d["total"]
                correct += d["correct"]
    for m in model:
        m.train()
    args.micro_batch_size, args.global_batch_size = saved
    if mpu.is_pipeline_last_stage():
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu"
        t = torch.tensor([correct, total], dtype=torch.int64, device=dev)
        if torch.distributed.is_initialized():
            torch.distributed.all_reduce(t, group=mpu.get_data_parallel_group())
        correct, total = t.tolist()
        print_rank_last(" > |epoch: {}| metrics for {}: correct / total = {} / {} = {:.4f} %, "
                        "elapsed time (sec): {:.3f}".format(epoch, name, correct, total,
                                                            100.0 * correct / max(total, 1), time.time() - t0))
        if output_predictions:
            return correct, total, (softmaxes, labels_all, ids_all)
        return correct, total
    if output_predictions:
        return 0, 0, ()
    return 0, 0
