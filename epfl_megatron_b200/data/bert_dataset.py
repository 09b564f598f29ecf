"""BERT pre-training samples: sentence-pair windows -> [CLS] A [SEP] B [SEP] with whole-word
n-gram masking and the sentence-order label (megatron/data/bert_dataset.py:20-183 surface)."""
from __future__ import annotations

import numpy as np
import torch

from .. import global_vars
from .dataset_utils import (create_masked_lm_predictions, create_tokens_and_tokentypes, get_a_and_b_segments,
                            get_samples_mapping, pad_and_convert_to_numpy, truncate_segments)


class BertDataset(torch.utils.data.Dataset):
    def __init__(self, name, indexed_dataset, data_prefix, num_epochs, max_num_samples, masked_lm_prob,
                 max_seq_length, short_seq_prob, seed, binary_head):
        self.name, self.seed = name, seed
        self.masked_lm_prob, self.max_seq_length, self.binary_head = masked_lm_prob, max_seq_length, binary_head
        self.indexed_dataset = indexed_dataset
        # three special tokens ([CLS], [SEP], [SEP]) come out of the budget
        self.samples_mapping = get_samples_mapping(indexed_dataset, data_prefix, num_epochs, max_num_samples,
                                                   max_seq_length - 3, short_seq_prob, seed, name, binary_head)
        tok = global_vars.get_tokenizer()
        self.vocab_id_list = list(tok.inv_vocab.keys())
        self.vocab_id_to_token_dict = tok.inv_vocab
        self.cls_id, self.sep_id, self.mask_id, self.pad_id = tok.cls, tok.sep, tok.mask, tok.pad

    def __len__(self):
        return self.samples_mapping.shape[0]

    def __getitem__(self, idx):
        start, end, seq_length = (int(x) for x in self.samples_mapping[idx])
        sample = [self.indexed_dataset[i] for i in range(start, end)]
        # numpy (not python) RNG: its randint upper bound is exclusive; seed differs per sample
        rng = np.random.RandomState(seed=(self.seed + idx) % 2 ** 32)
        return build_training_sample(sample, seq_length, self.max_seq_length, self.vocab_id_list,
                                     self.vocab_id_to_token_dict, self.cls_id, self.sep_id, self.mask_id,
                                     self.pad_id, self.masked_lm_prob, rng, self.binary_head)


def build_training_sample(sample, target_seq_length, max_seq_length, vocab_id_list, vocab_id_to_token_dict,
                          cls_id, sep_id, mask_id, pad_id, masked_lm_prob, np_rng, binary_head):
    """sample: list of sentences (token-id arrays). Returns the dict the BERT forward step consumes."""
    if binary_head:
        assert len(sample) > 1, "the sentence-order head needs at least two sentences"
    assert target_seq_length <= max_seq_length
    if binary_head:
        tokens_a, tokens_b, is_next_random = get_a_and_b_segments(sample, np_rng)
    else:
        tokens_a = [t for s in sample for t in s]
        tokens_b, is_next_random = [], False
    truncated = truncate_segments(tokens_a, tokens_b, len(tokens_a), len(tokens_b), target_seq_length, np_rng)
    tokens, tokentypes = create_tokens_and_tokentypes(tokens_a, tokens_b, cls_id, sep_id)
    max_predictions = masked_lm_prob * target_seq_length
    tokens, positions, labels, _, _ = create_masked_lm_predictions(
        tokens, vocab_id_list, vocab_id_to_token_dict, masked_lm_prob, cls_id, sep_id, mask_id,
        max_predictions, np_rng)
    tokens_np, types_np, labels_np, padding_mask, loss_mask = pad_and_convert_to_numpy(
        tokens, tokentypes, positions, labels, pad_id, max_seq_length)
    return {"text": tokens_np, "types": types_np, "labels": labels_np, "is_random": int(is_next_random),
            "loss_mask": loss_mask, "padding_mask": padding_mask, "truncated": int(truncated)}
