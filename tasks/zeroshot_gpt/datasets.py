"""Zero-shot evaluation datasets: sliding-window language modelling (WikiText-103) and last-word
prediction (LAMBADA) (tasks/zeroshot_gpt/datasets.py:18-147 surface)."""
import json
import math

import numpy as np
import torch

from epfl_megatron_b200 import get_args, get_tokenizer, print_rank_0
from .detokenizer import get_detokenizer


def build_dataset(task):
    if task == "LAMBADA":
        return _build_lambada_dataset()
    if task == "WIKITEXT103":
        return _build_wikitext103_dataset()
    raise NotImplementedError(f"dataset for {task} task is not implemented.")


class _LMDataset(torch.utils.data.Dataset):
    """Windows of seq_len tokens advancing by `overlapping_eval`; only the new tokens of each
    window are scored."""

    def __init__(self, tokens, seq_len, pad_idx, num_original_tokens, num_tokenized_tokens,
                 overlapping_eval=None):
        self.tokens, self.seq_len, self.pad_idx = tokens, seq_len, pad_idx
        self.overlapping_eval = max(1, overlapping_eval or seq_len)
        self.num_original_tokens, self.num_tokenized_tokens = num_original_tokens, num_tokenized_tokens
        targets = max(len(tokens) - 1 - self.overlapping_eval, 0)
        self.total_sequences = max(math.ceil(targets / self.overlapping_eval) + 1, 1)

    def __len__(self):
        return self.total_sequences

    def __getitem__(self, idx):
        start = idx * self.overlapping_eval
        tokens = list(self.tokens[start:start + self.seq_len + 1])
        n = len(tokens)
        pad_mask = [1] * n
        if n < self.seq_len + 1:
            pad_mask += [0] * (self.seq_len + 1 - n)
            tokens += [self.pad_idx] * (self.seq_len + 1 - n)
        pad_mask = np.array(pad_mask[1:])
        if self.overlapping_eval != self.seq_len and idx != 0:
            pad_mask[:-self.overlapping_eval] = 0
        return {"text": np.array(tokens), "pad_mask": pad_mask}


class _LambadaDataset(torch.utils.data.Dataset):
    def __init__(self, path, pad_idx, tokenizer, seq_len, strict=False):
        print_rank_0(f"> building lambada dataset from {path} ...")
        self.seq_len, self.pad_idx, self.tokenizer, self.strict = seq_len, pad_idx, tokenizer, strict
        self.tokens, self.labels = [], []
        with open(path, encoding="utf-8") as f:
            for line in f:
                t, l = self.get_tokens(json.loads(line)["text"])
                self.tokens.append(t)
                self.labels.append(l)

    def get_tokens(self, text):
        if not self.strict:
            ids = self.tokenizer.tokenize(text)
            return ids[:-1], [ids[-1]]
        last = text.split()[-1]
        start = text.rfind(last)
        return self.tokenizer.tokenize(text[:start].strip()), self.tokenizer.tokenize(" " + last)

    def __len__(self):
        return len(self.tokens)

    def __getitem__(self, idx):
        tokens, labels = list(self.tokens[idx]), self.labels[idx]
        pad_mask = [0] * len(tokens) + [1] * len(labels)
        tokens = tokens + labels
        if len(tokens) > self.seq_len + 1:
            # longer than the model's window: keep the end of the passage (the target word is last)
            # instead of failing to collate, which is all the reference does with such a sample
            tokens, pad_mask = tokens[-(self.seq_len + 1):], pad_mask[-(self.seq_len + 1):]
        n = len(tokens)
        if n < self.seq_len + 1:
            pad_mask += [0] * (self.seq_len + 1 - n)
            tokens += [self.pad_idx] * (self.seq_len + 1 - n)
        return {"text": np.array(tokens), "pad_mask": np.array(pad_mask[1:])}


def _build_lambada_dataset():
    args, tok = get_args(), get_tokenizer()
    assert len(args.valid_data) == 1
    ds = _LambadaDataset(args.valid_data[0], tok.eod, tok, args.seq_length, args.strict_lambada)
    print_rank_0(f" > found {len(ds)} samples.")
    return ds


def _build_wikitext103_dataset():
    args, tok = get_args(), get_tokenizer()
    assert len(args.valid_data) == 1
    with open(args.valid_data[0], "rb") as f:
        raw = f.read().decode("utf-8")
    n_orig = len(raw.strip().split(" "))
    detok = get_detokenizer(args.valid_data[0])
    ids = tok.tokenize(detok(raw) if detok else raw)
    ds = _LMDataset(ids, args.seq_length, tok.eod, n_orig, len(ids), args.overlapping_eval)
    print_rank_0(f" > number of original tokens: {n_orig}, number of detokenized tokens: {len(ids)}")
    return ds
