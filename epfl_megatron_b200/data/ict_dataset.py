"""Inverse Cloze Task: a sentence is the pseudo-query, the block it came from the positive
context (with probability 1 - query_in_block_prob the sentence is removed from its block)
(megatron/data/ict_dataset.py:13-157 surface)."""
from __future__ import annotations

import itertools
import random

import numpy as np
import torch

from .. import global_vars
from .dataset_utils import get_indexed_dataset_
from .realm_dataset_utils import get_block_samples_mapping


def make_attention_mask(source_block, target_block):
    return ((target_block[None, :] >= 1) * (source_block[:, None] >= 1)).astype(np.int64)


def get_ict_dataset(use_titles=True, query_in_block_prob=1):
    """One epoch over every block of --data_path (index building / evaluation)."""
    args = global_vars.get_args()
    block_dataset = get_indexed_dataset_(args.data_path, "mmap", True)
    titles_dataset = get_indexed_dataset_(args.titles_data_path, "mmap", True)
    return ICTDataset(name="full", block_dataset=block_dataset, title_dataset=titles_dataset,
                      data_prefix=args.data_path, num_epochs=1, max_num_samples=None,
                      max_seq_length=args.seq_length, seed=1, query_in_block_prob=query_in_block_prob,
                      use_titles=use_titles, use_one_sent_docs=args.use_one_sent_docs)


class ICTDataset(torch.utils.data.Dataset):
    def __init__(self, name, block_dataset, title_dataset, data_prefix, num_epochs, max_num_samples,
                 max_seq_length, query_in_block_prob, seed, use_titles=True, use_one_sent_docs=False,
                 binary_head=False):
        self.name, self.seed, self.max_seq_length = name, seed, max_seq_length
        self.query_in_block_prob = query_in_block_prob
        self.block_dataset, self.title_dataset = block_dataset, title_dataset
        self.rng = random.Random(self.seed)
        self.use_titles, self.use_one_sent_docs = use_titles, use_one_sent_docs
        self.samples_mapping = get_block_samples_mapping(block_dataset, title_dataset, data_prefix, num_epochs,
                                                         max_num_samples, max_seq_length, seed, name,
                                                         use_one_sent_docs)
        tok = global_vars.get_tokenizer()
        self.vocab_id_list = list(tok.inv_vocab.keys())
        self.vocab_id_to_token_list = tok.inv_vocab
        self.cls_id, self.sep_id, self.mask_id, self.pad_id = tok.cls, tok.sep, tok.mask, tok.pad

    def __len__(self):
        return len(self.samples_mapping)

    def __getitem__(self, idx):
        start, end, doc, block_idx = self.samples_mapping[idx].as_tuple()
        if self.use_titles:
            title = self.title_dataset[int(doc)]
            title_pad = 3 + len(title)
        else:
            title, title_pad = None, 2
        block = [self.block_dataset[i] for i in range(start, end)]
        assert len(block) > 1 or self.use_one_sent_docs or self.query_in_block_prob == 1
        pick = self.rng.randint(0, len(block) - 1)
        if self.rng.random() < self.query_in_block_prob:
            query = block[pick].copy()
        else:
            query = block.pop(pick)
        query = query[:self.max_seq_length - 2]
        block = list(itertools.chain(*block))[:self.max_seq_length - title_pad]
        q_tokens, q_pad = self.concat_and_pad_tokens(query)
        c_tokens, c_pad = self.concat_and_pad_tokens(block, title)
        return {"query_tokens": q_tokens, "query_mask": make_attention_mask(q_tokens, q_tokens),
                "query_pad_mask": q_pad, "context_tokens": c_tokens,
                "context_mask": make_attention_mask(c_tokens, c_tokens), "context_pad_mask": c_pad,
                "block_data": np.array([start, end, doc, block_idx]).astype(np.int64)}

    def get_block(self, start_idx, end_idx, doc_idx):
        block = [self.block_dataset[i] for i in range(start_idx, end_idx)]
        title = self.title_dataset[int(doc_idx)]
        block = list(itertools.chain(*block))[:self.max_seq_length - (3 + len(title))]
        return self.concat_and_pad_tokens(block, title)

    def get_null_block(self):
        return self.concat_and_pad_tokens([], [])

    def concat_and_pad_tokens(self, tokens, title=None):
        """[CLS] (title [SEP]) tokens [SEP] padded to max_seq_length; returns (ids, pad mask)."""
        tokens = list(tokens)
        if title is None:
            tokens = [self.cls_id] + tokens + [self.sep_id]
        else:
            tokens = [self.cls_id] + list(title) + [self.sep_id] + tokens + [self.sep_id]
        assert len(tokens) <= self.max_seq_length
        pad = self.max_seq_length - len(tokens)
        return (np.array(tokens + [self.pad_id] * pad), np.array([1] * len(tokens) + [0] * pad))
