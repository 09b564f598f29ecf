"""Helpers for the ICT / REALM retrieval datasets: one-epoch loader, batch unpacking, the
(start, end, doc, block) record, and the cached block-sample mapping
(megatron/data/realm_dataset_utils.py:11-200 surface)."""
from __future__ import annotations

import os
import time

import numpy as np
import torch
import torch.distributed as dist

from .. import global_vars
from ..parallel import state as mpu
from ..parallel.data import broadcast_data
from ..utils import print_rank_0
from .dataset_utils import create_masked_lm_predictions, pad_and_convert_to_numpy


def _dev():
    return "cuda" if dist.is_initialized() and dist.get_backend() == "nccl" else "cpu"


def get_one_epoch_dataloader(dataset, micro_batch_size=None):
    """Sequential loader that walks the dataset exactly once (used to embed every block)."""
    from .data_samplers import MegatronPretrainingSampler
    args = global_vars.get_args()
    if micro_batch_size is None:
        micro_batch_size = args.micro_batch_size
    sampler = MegatronPretrainingSampler(
        total_samples=len(dataset), consumed_samples=0, micro_batch_size=micro_batch_size,
        data_parallel_rank=mpu.get_data_parallel_rank(), data_parallel_size=mpu.get_data_parallel_world_size(),
        drop_last=False)
    return torch.utils.data.DataLoader(dataset, batch_sampler=sampler, num_workers=args.num_workers,
                                       pin_memory=True)


def get_ict_batch(data_iterator):
    keys = ["query_tokens", "query_pad_mask", "block_tokens", "block_pad_mask", "block_data"]
    data = None if data_iterator is None else next(data_iterator)
    d = broadcast_data(keys, data, torch.int64)
    return (d["query_tokens"].long(), d["query_pad_mask"].long(), d["block_tokens"].long(),
            d["block_pad_mask"].long(), d["block_data"].long())


def join_str_list(str_list):
    """WordPiece tokens -> text ("##" pieces glue to the previous token)."""
    out = ""
    for s in str_list:
        out += s[2:] if s.startswith("##") else " " + s
    return out


class BlockSampleData:
    """(first sentence, end sentence, document, block id) of one retrieval block."""

    def __init__(self, start_idx, end_idx, doc_idx, block_idx):
        self.start_idx, self.end_idx, self.doc_idx, self.block_idx = start_idx, end_idx, doc_idx, block_idx

    def as_array(self):
        return np.array([self.start_idx, self.end_idx, self.doc_idx, self.block_idx]).astype(np.int64)

    def as_tuple(self):
        return self.start_idx, self.end_idx, self.doc_idx, self.block_idx


class BlockSamplesMapping:
    def __init__(self, mapping_array):
        assert mapping_array.shape[1] == 4
        self.mapping_array = mapping_array

    def __len__(self):
        return self.mapping_array.shape[0]

    def __getitem__(self, idx):
        return BlockSampleData(*self.mapping_array[idx])


def get_block_samples_mapping(block_dataset, title_dataset, data_prefix, num_epochs, max_num_samples,
                              max_seq_length, seed, name, use_one_sent_docs=False):
    """Built once on rank 0 with the native helper, cached as .npy next to the data."""
    from . import helpers
    if not num_epochs:
        if not max_num_samples:
            raise ValueError("Need to specify either max_num_samples or num_epochs")
        num_epochs = np.iinfo(np.int32).max - 1
    if not max_num_samples:
        max_num_samples = np.iinfo(np.int64).max - 1
    fname = f"{data_prefix}_{name}_indexmap"
    if num_epochs != (np.iinfo(np.int32).max - 1):
        fname += f"_{num_epochs}ep"
    if max_num_samples != (np.iinfo(np.int64).max - 1):
        fname += f"_{max_num_samples}mns"
    fname += f"_{max_seq_length}msl_{seed}s"
    if use_one_sent_docs:
        fname += "_1sentok"
    fname += ".npy"
    rank0 = not dist.is_initialized() or mpu.get_data_parallel_rank() == 0
    if rank0 and not os.path.isfile(fname):
        print(f" > WARNING: could not find index map file {fname}, building the indices on rank 0 ...")
        assert block_dataset.doc_idx.dtype == np.int64 and block_dataset.sizes.dtype == np.int32
        t0 = time.time()
        mapping = helpers.build_blocks_mapping(block_dataset.doc_idx, block_dataset.sizes, title_dataset.sizes,
                                               num_epochs, max_num_samples, max_seq_length - 3, seed,
                                               not dist.is_initialized() or dist.get_rank() == 0,
                                               use_one_sent_docs)
        np.save(fname, mapping, allow_pickle=True)
        print_rank_0(f" > elapsed time to build and save samples mapping (seconds): {time.time() - t0:4f}")
    if dist.is_initialized():
        c = torch.ones(1, dtype=torch.long, device=_dev())
        dist.all_reduce(c, group=mpu.get_data_parallel_group())
        assert int(c[0].item()) == dist.get_world_size(group=mpu.get_data_parallel_group())
    print_rank_0(f" > loading indexed mapping from {fname}")
    mapping = BlockSamplesMapping(np.load(fname, allow_pickle=True, mmap_mode="r"))
    print_rank_0(f"    total number of samples: {len(mapping)}")
    return mapping
