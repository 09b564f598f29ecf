"""Batch plumbing shared by the biencoder (ICT / ORQA) entry points
(megatron/data/biencoder_dataset_utils.py:13-209 surface)."""
from __future__ import annotations

import numpy as np
import torch

from .. import global_vars
from ..parallel import state as mpu
from ..parallel.data import broadcast_data
from .realm_dataset_utils import get_block_samples_mapping, get_one_epoch_dataloader  # noqa: F401


def make_attention_mask(source_block, target_block):
    return ((target_block[None, :] >= 1) * (source_block[:, None] >= 1)).astype(np.int64)


def get_ict_batch(data_iterator):
    keys = ["query_tokens", "query_mask", "context_tokens", "context_mask", "block_data"]
    data = None if data_iterator is None else next(data_iterator)
    d = broadcast_data(keys, data, torch.int64)
    return (d["query_tokens"].long(), d["query_mask"] < 0.5, d["context_tokens"].long(),
            d["context_mask"] < 0.5, d["block_data"].long())


def join_str_list(str_list):
    out = ""
    for s in str_list:
        out += s[2:] if s.startswith("##") else " " + s
    return out
