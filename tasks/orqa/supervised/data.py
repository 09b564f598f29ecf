"""Supervised retriever data (DPR json: question, positive_ctxs, hard_negative_ctxs, negative_ctxs)
(tasks/orqa/supervised/data.py:16-287 surface)."""
import json
import random
from abc import ABC, abstractmethod

import numpy as np
from torch.utils.data import Dataset

from epfl_megatron_b200 import get_args, print_rank_0
from epfl_megatron_b200.data.biencoder_dataset_utils import make_attention_mask


def build_tokens_types_paddings_from_ids(text_ids, max_seq_length, cls_id, sep_id, pad_id):
    ids = [cls_id] + list(text_ids)
    if len(ids) > max_seq_length - 1:
        ids = ids[:max_seq_length - 1]
    ids.append(sep_id)
    n = len(ids)
    pad = max_seq_length - n
    return ids + [pad_id] * pad, [0] * n + [pad_id] * pad, np.array([1] * n + [0] * pad, dtype=np.int64)


def _ctx_ids(ctx, tokenizer):
    return tokenizer.tokenize(ctx["title"]) + [tokenizer.sep] + tokenizer.tokenize(ctx["text"])


def build_token_types_from_context_list(ctx_list, tokenizer, max_seq_length):
    ids, types, masks = [], [], []
    for ctx in ctx_list:
        i, t, m = build_tokens_types_paddings_from_ids(_ctx_ids(ctx, tokenizer), max_seq_length, tokenizer.cls,
                                                       tokenizer.sep, tokenizer.pad)
        ids.append(i)
        types.append(t)
        masks.append(m)
    return ids, types, masks


def build_tokens_types_paddings_from_text(query, context, tokenizer, max_seq_length):
    q = build_tokens_types_paddings_from_ids(tokenizer.tokenize(query), max_seq_length, tokenizer.cls,
                                             tokenizer.sep, tokenizer.pad)
    c = build_tokens_types_paddings_from_ids(_ctx_ids(context, tokenizer), max_seq_length, tokenizer.cls,
                                             tokenizer.sep, tokenizer.pad)
    return q + c


def build_sample(query_ids, query_types, query_pad_mask, ctx_ids, ctx_types, ctx_pad_mask, answers,
                 neg_ctx_id_list=None, neg_ctx_types_list=None, include_neg=False):
    q, c = np.array(query_ids, dtype=np.int64), np.array(ctx_ids, dtype=np.int64)
    s = {"query": q, "query_mask": make_attention_mask(q, q), "query_types": np.array(query_types, dtype=np.int64),
         "query_pad_mask": query_pad_mask, "context": c, "context_mask": make_attention_mask(c, c),
         "context_types": np.array(ctx_types, dtype=np.int64), "context_pad_mask": ctx_pad_mask,
         "reference": answers}
    if include_neg:
        neg = np.array(neg_ctx_id_list, dtype=np.int64)
        s["neg_context"] = neg
        s["neg_context_types"] = np.array(neg_ctx_types_list, dtype=np.int64)
        s["neg_context_mask"] = np.array([make_attention_mask(i, i) for i in neg], dtype=np.int64)
    return s


def normalize_question(question):
    return question[:-1] if question.endswith("?") else question


class OpenRetrievalAbstractDataset(ABC, Dataset):
    def __init__(self, task_name, dataset_name, datapaths, tokenizer, max_seq_length, evaluate=False):
        args = get_args()
        self.evaluate = evaluate
        self.val_av_rank_hard_neg, self.val_av_rank_other_neg = args.val_av_rank_hard_neg, args.val_av_rank_other_neg
        self.train_with_neg, self.train_hard_neg = args.train_with_neg, args.train_hard_neg
        self.task_name, self.dataset_name = task_name, dataset_name
        self.tokenizer, self.max_seq_length = tokenizer, max_seq_length
        print_rank_0(f" > building {task_name} dataset for {dataset_name}:")
        self.samples = []
        for p in datapaths:
            self.samples.extend(self.process_samples_from_single_path(p))
        if getattr(args, "sample_rate", 1.0) < 1:
            self.samples = random.sample(self.samples, int(len(self.samples) * args.sample_rate))
        print_rank_0(f"  >> total number of samples: {len(self.samples)}")

    def __len__(self):
        return len(self.samples)

    def __getitem__(self, idx):
        raw = self.samples[idx]
        q = build_tokens_types_paddings_from_text(raw["question"], raw["pos_context"], self.tokenizer,
                                                  self.max_seq_length)
        neg_ids = neg_types = None
        include_neg = False
        if self.evaluate:
            negs = raw["negative_context"][:self.val_av_rank_other_neg] + \
                raw["hard_negative_context"][:self.val_av_rank_hard_neg]
            include_neg = True
        elif self.train_with_neg:
            hard = list(raw["hard_negative_context"])
            random.shuffle(hard)
            negs = hard[:self.train_hard_neg]
            if len(negs) < self.train_hard_neg:
                simple = list(raw["negative_context"])
                random.shuffle(simple)
                negs += simple[:self.train_hard_neg - len(negs)]
            include_neg = True
        if include_neg:
            neg_ids, neg_types, _ = build_token_types_from_context_list(negs, self.tokenizer, self.max_seq_length)
        return build_sample(q[0], q[1], q[2], q[3], q[4], q[5], raw["answers"], neg_ids, neg_types, include_neg)

    @staticmethod
    @abstractmethod
    def process_samples_from_single_path(filename):
        """-> list of {question, pos_context, hard_negative_context, negative_context, answers}"""


class NQSupervisedDataset(OpenRetrievalAbstractDataset):
    def __init__(self, name, datapaths, tokenizer, max_seq_length, evaluate=False):
        super().__init__("natural_questions_ret", name, datapaths, tokenizer, max_seq_length, evaluate=evaluate)

    @staticmethod
    def process_samples_from_single_path(filename):
        print_rank_0(f" > Processing {filename} ...")
        with open(filename, encoding="utf-8") as f:
            data = json.load(f)
        samples = []
        for row in data:
            if not row.get("positive_ctxs"):
                continue
            samples.append({"question": normalize_question(row["question"]),
                            "pos_context": row["positive_ctxs"][0],
                            "hard_negative_context": row.get("hard_negative_ctxs", []),
                            "negative_context": row.get("negative_ctxs", []), "answers": row["answers"]})
        print_rank_0(f" >> processed {len(samples)} samples.")
        return samples
