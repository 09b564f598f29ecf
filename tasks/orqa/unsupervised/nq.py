"""Natural Questions (DPR csv: question \\t answers) for zero-shot retrieval evaluation
(tasks/orqa/unsupervised/nq.py:24-215 surface)."""
import csv
from abc import ABC

import numpy as np
import torch
from torch.utils.data import BatchSampler, DataLoader, Dataset

from epfl_megatron_b200 import get_args, get_tokenizer, print_rank_0
from epfl_megatron_b200.data.biencoder_dataset_utils import make_attention_mask
from epfl_megatron_b200.parallel import state as mpu


def get_nq_dataset(qa_data, split):
    args, tok = get_args(), get_tokenizer()
    return NQDataset(f"Google NQ {split} Split", "Google Natural Questions", qa_data, tok,
                     args.retriever_seq_length)


def process_nq_batch(batch):
    return (batch["token_ids"].long(), (batch["token_mask"] < 0.5), batch["token_types"].long(),
            batch["seq_len"].long(), batch["reference"])


class CustomDataLoader(DataLoader):
    """Collates references (lists of strings) without stacking them into tensors."""

    def __init__(self, dataset, eval=False, **kwargs):
        if kwargs.get("collate_fn") is None:
            kwargs["collate_fn"] = self._collate_fn
        self.eval = eval
        super().__init__(dataset, **kwargs)

    @staticmethod
    def _collate_fn(batch):
        out = {}
        for k in batch[0]:
            vals = [d[k] for d in batch]
            out[k] = vals if k == "reference" else torch.from_numpy(np.array(vals))
        return out


def get_one_epoch_nq_dataloader(dataset, micro_batch_size=None):
    args = get_args()
    if micro_batch_size is None:
        micro_batch_size = args.micro_batch_size
    sampler = torch.utils.data.distributed.DistributedSampler(
        dataset, num_replicas=mpu.get_data_parallel_world_size(), rank=mpu.get_data_parallel_rank(),
        shuffle=False)
    return CustomDataLoader(dataset, batch_sampler=BatchSampler(sampler, micro_batch_size, drop_last=False),
                            num_workers=args.num_workers, pin_memory=True)


def build_tokens_types_paddings_from_text(src_text, tokenizer, max_seq_length):
    return build_tokens_types_paddings_from_ids(tokenizer.tokenize(src_text), max_seq_length, tokenizer.cls,
                                                tokenizer.sep, tokenizer.pad)


def build_tokens_types_paddings_from_ids(src_ids, max_seq_length, cls_id, sep_id, pad_id):
    ids = [cls_id] + list(src_ids)
    if len(ids) > max_seq_length - 1:
        ids = ids[:max_seq_length - 1]
    ids.append(sep_id)
    n = len(ids)
    pad = max_seq_length - n
    return ids + [pad_id] * pad, [0] * n + [pad_id] * pad, n


def build_sample(token_ids, token_types, num_tokens, reference):
    ids = np.array(token_ids, dtype=np.int64)
    return {"token_ids": ids, "token_mask": make_attention_mask(ids, ids),
            "token_types": np.array(token_types, dtype=np.int64), "seq_len": num_tokens,
            "reference": reference}


class NQDataset(ABC, Dataset):
    def __init__(self, task_name, dataset_name, datapath, tokenizer, max_seq_length):
        self.task_name, self.dataset_name = task_name, dataset_name
        self.tokenizer, self.max_seq_length = tokenizer, max_seq_length
        print_rank_0(f" > building {task_name} dataset for {dataset_name}:")
        print_rank_0(f"   {datapath}")
        self.samples = self.process_samples_from_single_path(datapath)
        print_rank_0(f"  >> total number of samples: {len(self.samples)}")

    def __len__(self):
        return len(self.samples)

    def __getitem__(self, idx):
        raw = self.samples[idx]
        ids, types, n = build_tokens_types_paddings_from_text(raw["question"], self.tokenizer,
                                                              self.max_seq_length)
        return build_sample(ids, types, n, raw["answers"])

    @staticmethod
    def process_samples_from_single_path(filename):
        print_rank_0(f" > Processing {filename} ...")
        samples = []
        with open(filename, encoding="utf-8") as f:
            for row in csv.reader(f, delimiter="\t"):
                samples.append({"question": row[0], "answers": eval(row[1])})   # noqa: S307 - DPR format
        print_rank_0(f" >> processed {len(samples)} samples.")
        return samples
