"""Wikipedia evidence passages (DPR-style TSV: id, text, title) for open-retrieval QA
(megatron/data/orqa_wiki_dataset.py:17-184 surface)."""
from __future__ import annotations

import csv
import random
from abc import ABC

import numpy as np
import torch

from .. import global_vars
from ..utils import print_rank_0
from .biencoder_dataset_utils import make_attention_mask


def get_open_retrieval_wiki_dataset():
    args = global_vars.get_args()
    tok = global_vars.get_tokenizer()
    return OpenRetrievalEvidenceDataset("2018 Wikipedia from DPR codebase", "evidence", args.evidence_data_path,
                                        tok, args.retriever_seq_length)


def get_open_retrieval_batch(data_iterator):
    from ..parallel.data import broadcast_data
    keys = ["row_id", "context", "context_mask", "context_types", "context_pad_mask"]
    data = None if data_iterator is None else next(data_iterator)
    d = broadcast_data(keys, data, torch.int64)
    return (d["row_id"].long(), d["context"].long(), d["context_mask"] < 0.5, d["context_types"].long(),
            d["context_pad_mask"].long())


def build_tokens_types_paddings_from_text(row, tokenizer, max_seq_length):
    title_ids = tokenizer.tokenize(row["title"])
    ctx_ids = tokenizer.tokenize(row["text"])
    return build_tokens_types_paddings_from_ids(title_ids + [tokenizer.sep_id] + ctx_ids, max_seq_length,
                                                tokenizer.cls, tokenizer.sep, tokenizer.pad)


def build_tokens_types_paddings_from_ids(text_ids, max_seq_length, cls_id, sep_id, pad_id):
    """[CLS] text [SEP], truncated and padded. -> (ids, token types, pad mask)."""
    ids = [cls_id] + list(text_ids)
    if len(ids) > max_seq_length - 1:
        ids = ids[:max_seq_length - 1]
    ids.append(sep_id)
    n = len(ids)
    pad = max_seq_length - n
    ids += [pad_id] * pad
    return ids, [0] * n + [pad_id] * pad, np.array([1] * n + [0] * pad, dtype=np.int64)


def build_sample(row_id, context_ids, context_types, context_pad_mask):
    ids = np.array(context_ids, dtype=np.int64)
    return {"row_id": row_id, "context": ids, "context_mask": make_attention_mask(ids, ids),
            "context_types": np.array(context_types, dtype=np.int64), "context_pad_mask": context_pad_mask}


class OpenRetrievalEvidenceDataset(ABC, torch.utils.data.Dataset):
    def __init__(self, task_name, dataset_name, datapath, tokenizer, max_seq_length):
        self.task_name, self.dataset_name = task_name, dataset_name
        self.tokenizer, self.max_seq_length = tokenizer, max_seq_length
        print_rank_0(f" > building {task_name} dataset for {dataset_name}:")
        print_rank_0(f"   {datapath}")
        self.samples, self.id2text = self.process_samples_from_single_path(datapath)
        args = global_vars.get_args()
        if getattr(args, "sample_rate", 1.0) < 1:
            k = int(len(self.samples) * args.sample_rate)
            self.samples = random.sample(self.samples, k)
        print_rank_0(f"  >> total number of samples: {len(self.samples)}")

    def __len__(self):
        return len(self.samples)

    def __getitem__(self, idx):
        row = self.samples[idx]
        ids, types, pad_mask = build_tokens_types_paddings_from_text(row, self.tokenizer, self.max_seq_length)
        return build_sample(row["doc_id"], ids, types, pad_mask)

    @staticmethod
    def process_samples_from_single_path(filename):
        print_rank_0(f" > Processing {filename} ...")
        rows, id2text = [], {}
        with open(filename, encoding="utf-8") as f:
            reader = csv.reader(f, delimiter="\t")
            next(reader, None)   # header
            for total, row in enumerate(reader, start=1):
                doc_id, text, title = int(row[0]), row[1], row[2]
                rows.append({"doc_id": doc_id, "text": text, "title": title})
                assert doc_id not in id2text
                id2text[doc_id] = (text, title)
                if total % 100000 == 0:
                    print_rank_0(f"  > processed {total} rows so far ...")
        print_rank_0(f" >> processed {len(rows)} samples.")
        return rows, id2text
