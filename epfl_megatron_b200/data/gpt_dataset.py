"""GPT-style dataset over an indexed token corpus (megatron/data/gpt_dataset.py:20-513
surface). Samples are seq_length+1 tokens cut from the concatenation of epoch-shuffled
documents; three index arrays (doc_idx, sample_idx, shuffle_idx) are built once on global
rank 0, cached as .npy next to the data, and memory-mapped by every rank."""
from __future__ import annotations

import os
import time

import numpy as np
import torch
import torch.distributed as dist

from ..parallel import state as mpu
from ..utils import print_rank_0
from . import helpers
from .blendable_dataset import BlendableDataset
from .dataset_utils import get_datasets_weights_and_num_samples, get_train_valid_test_split_
from .indexed_dataset import make_dataset as make_indexed_dataset


def build_train_valid_test_datasets(data_prefix, data_impl, splits_string, train_valid_test_num_samples,
                                    seq_length, seed, skip_warmup, train_data_prefix=None,
                                    valid_data_prefix=None, test_data_prefix=None):
    if data_prefix:
        print_rank_0("Single data path provided for train, valid & test")
        if len(data_prefix) == 1:
            return _build_train_valid_test_datasets(data_prefix[0], data_impl, splits_string,
                                                    train_valid_test_num_samples, seq_length, seed, skip_warmup)
        prefixes, weights, per_ds = get_datasets_weights_and_num_samples(data_prefix, train_valid_test_num_samples)
        groups = ([], [], [])
        for i, prefix in enumerate(prefixes):
            tvt = _build_train_valid_test_datasets(prefix, data_impl, splits_string, per_ds[i], seq_length, seed,
                                                   skip_warmup)
            for g, ds in zip(groups, tvt):
                if ds:
                    g.append(ds)
        return tuple(BlendableDataset(g, weights) if g else None for g in groups)
    print_rank_0("Separate data paths provided for train, valid & test. Split string will be ignored.")
    out = []
    for name, prefix, n in (("train", train_data_prefix, train_valid_test_num_samples[0]),
                            ("valid", valid_data_prefix, train_valid_test_num_samples[1]),
                            ("test", test_data_prefix, train_valid_test_num_samples[2])):
        out.append(build_dataset(name, prefix, data_impl, n, seq_length, seed, skip_warmup)
                   if prefix is not None else None)
    return tuple(out)


def build_dataset(dataset_name, data_prefix, data_impl, num_samples, seq_length, seed, skip_warmup):
    if len(data_prefix) == 1:
        return _build_dataset(dataset_name, data_prefix[0], data_impl, num_samples, seq_length, seed, skip_warmup)
    prefixes, weights, per_ds = get_datasets_weights_and_num_samples(data_prefix, num_samples)
    datasets = []
    for i, prefix in enumerate(prefixes):
        ds = _build_dataset(dataset_name, prefix, data_impl, per_ds[i], seq_length, seed, skip_warmup)
        if ds:
            datasets.append(ds)
    return BlendableDataset(datasets, weights) if datasets else None


def _build_dataset(dataset_name, data_prefix, data_impl, num_samples, seq_length, seed, skip_warmup):
    indexed = get_indexed_dataset_(data_prefix, data_impl, skip_warmup)
    total = indexed.sizes.shape[0]
    print_rank_0("    {}:".format(dataset_name))
    print_rank_0("     document indices in [0, {}) total of {} documents".format(total, total))
    documents = np.arange(start=0, stop=total, step=1, dtype=np.int32)
    return GPTDataset(dataset_name, data_prefix, documents, indexed, num_samples, seq_length, seed)


def _build_train_valid_test_datasets(data_prefix, data_impl, splits_string, train_valid_test_num_samples,
                                     seq_length, seed, skip_warmup):
    indexed = get_indexed_dataset_(data_prefix, data_impl, skip_warmup)
    total = indexed.sizes.shape[0]
    splits = get_train_valid_test_split_(splits_string, total)
    print_rank_0(" > dataset split:")
    for i, name in enumerate(("train", "validation", "test")):
        print_rank_0("    {}:".format(name))
        print_rank_0("     document indices in [{}, {}) total of {} documents".format(
            splits[i], splits[i + 1], splits[i + 1] - splits[i]))

    def build(i, name):
        if splits[i + 1] <= splits[i]:
            return None
        documents = np.arange(start=splits[i], stop=splits[i + 1], step=1, dtype=np.int32)
        return GPTDataset(name, data_prefix, documents, indexed, train_valid_test_num_samples[i], seq_length, seed)

    return build(0, "train"), build(1, "valid"), build(2, "test")


def get_indexed_dataset_(data_prefix, data_impl, skip_warmup):
    print_rank_0(" > building dataset index ...")
    t0 = time.time()
    ds = make_indexed_dataset(data_prefix, data_impl, skip_warmup)
    print_rank_0(" > finished creating indexed dataset in {:4f} seconds".format(time.time() - t0))
    print_rank_0("    number of documents: {}".format(ds.sizes.shape[0]))
    return ds


class GPTDataset(torch.utils.data.Dataset):
    def __init__(self, name, data_prefix, documents, indexed_dataset, num_samples, seq_length, seed):
        self.name, self.indexed_dataset = name, indexed_dataset
        assert np.min(documents) >= 0 and np.max(documents) < indexed_dataset.sizes.shape[0]
        self.doc_idx, self.sample_idx, self.shuffle_idx = _build_index_mappings(
            name, data_prefix, documents, indexed_dataset.sizes, num_samples, seq_length, seed)

    def __len__(self):
        return self.sample_idx.shape[0] - 1

    def __getitem__(self, idx):
        idx = self.shuffle_idx[idx]
        df, of = self.sample_idx[idx]
        dl, ol = self.sample_idx[idx + 1]
        if df == dl:
            sample = self.indexed_dataset.get(self.doc_idx[df], offset=of, length=ol - of + 1)
        else:
            parts = [self.indexed_dataset.get(self.doc_idx[df], offset=of)]
            for i in range(df + 1, dl):
                parts.append(self.indexed_dataset.get(self.doc_idx[i]))
            parts.append(self.indexed_dataset.get(self.doc_idx[dl], length=ol + 1))
            sample = np.concatenate(parts)
        return {"text": np.array(sample, dtype=np.int64)}


def _num_tokens(documents, sizes):
    return np.sum(sizes[documents])


def _num_epochs(tokens_per_epoch, seq_length, num_samples):
    epochs, total = 0, 0
    while True:
        epochs += 1
        total += tokens_per_epoch
        if (total - 1) // seq_length >= num_samples:
            return epochs


def _build_doc_idx(documents, num_epochs, np_rng, separate_last_epoch):
    if not separate_last_epoch or num_epochs == 1:
        doc_idx = np.mgrid[0:num_epochs, 0:len(documents)][1]
        doc_idx[:] = documents
        doc_idx = doc_idx.reshape(-1).astype(np.int32)
        np_rng.shuffle(doc_idx)
        return doc_idx
    first = _build_doc_idx(documents, num_epochs - 1, np_rng, False)
    last = _build_doc_idx(documents, 1, np_rng, False)
    return np.concatenate((first, last))


def _build_sample_idx(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch):
    return helpers.build_sample_idx_py(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch)


def _build_shuffle_idx(num_samples, total_size, np_rng):
    print(" > building shuffle index with split [0, {}) and [{}, {}) ...".format(num_samples, num_samples,
                                                                                 total_size), flush=True)
    dtype_ = np.uint32 if total_size < (np.iinfo(np.uint32).max - 1) else np.int64
    first = np.arange(start=0, stop=num_samples, step=1, dtype=dtype_)
    np_rng.shuffle(first)
    if num_samples == total_size:
        return first
    last = np.arange(start=num_samples, stop=total_size, step=1, dtype=dtype_)
    np_rng.shuffle(last)
    return np.concatenate((first, last))


def _build_index_mappings(name, data_prefix, documents, sizes, num_samples, seq_length, seed):
    tokens_per_epoch = _num_tokens(documents, sizes)
    num_epochs = _num_epochs(tokens_per_epoch, seq_length, num_samples)
    np_rng = np.random.RandomState(seed=seed)
    base = data_prefix + "_{}_indexmap_{}ns_{}sl_{}s".format(name, num_samples, seq_length, seed)
    doc_f, sample_f, shuffle_f = base + "_doc_idx.npy", base + "_sample_idx.npy", base + "_shuffle_idx.npy"
    rank0 = not dist.is_initialized() or dist.get_rank() == 0
    if rank0 and not (os.path.isfile(doc_f) and os.path.isfile(sample_f) and os.path.isfile(shuffle_f)):
        print_rank_0(" > WARNING: could not find index map files, building the indices on rank 0 ...")
        if num_epochs == 1:
            separate_last_epoch = False
            print(" > only one epoch required, setting separate_last_epoch to False", flush=True)
        else:
            from_prev = ((num_epochs - 1) * tokens_per_epoch - 1) // seq_length
            last_epoch_samples = num_samples - from_prev
            assert last_epoch_samples >= 0, "last epoch number of samples should be non-negative."
            per_epoch = (tokens_per_epoch - 1) // seq_length
            assert last_epoch_samples < (per_epoch + 1), "last epoch number of samples exceeded max value."
            # the last epoch is shuffled separately when it would contribute < 80 % of an epoch
            separate_last_epoch = last_epoch_samples < int(0.80 * per_epoch)
            print(" > last epoch number of samples ({}) is {} than 80% of number of samples per epoch ({}), "
                  "setting separate_last_epoch to {}".format(last_epoch_samples,
                                                             "smaller" if separate_last_epoch else "larger",
                                                             per_epoch, separate_last_epoch), flush=True)
        t0 = time.time()
        doc_idx = _build_doc_idx(documents, num_epochs, np_rng, separate_last_epoch)
        np.save(doc_f, doc_idx, allow_pickle=True)
        print_rank_0(" > elasped time to build and save doc-idx mapping (seconds): {:4f}".format(time.time() - t0))
        t0 = time.time()
        assert doc_idx.dtype == np.int32 and sizes.dtype == np.int32
        sample_idx = helpers.build_sample_idx(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch)
        np.save(sample_f, sample_idx, allow_pickle=True)
        print_rank_0(" > elasped time to build and save sample-idx mapping (seconds): {:4f}".format(time.time() - t0))
        t0 = time.time()
        n_ = from_prev if separate_last_epoch else sample_idx.shape[0] - 1
        shuffle_idx = _build_shuffle_idx(n_, sample_idx.shape[0] - 1, np_rng)
        np.save(shuffle_f, shuffle_idx, allow_pickle=True)
        print_rank_0(" > elasped time to build and save shuffle-idx mapping (seconds): {:4f}".format(time.time() - t0))
    if dist.is_initialized():
        # every rank waits for rank 0 (all-reduces over DP and PP groups act as the barrier)
        dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
        counts = torch.ones(1, dtype=torch.long, device=dev)
        dist.all_reduce(counts, group=mpu.get_data_parallel_group())
        dist.all_reduce(counts, group=mpu.get_pipeline_model_parallel_group())
        assert counts[0].item() == dist.get_world_size() // dist.get_world_size(
            group=mpu.get_tensor_model_parallel_group())
    t0 = time.time()
    print_rank_0(" > loading doc-idx mapping from {}".format(doc_f))
    doc_idx = np.load(doc_f, allow_pickle=True, mmap_mode="r")
    print_rank_0(" > loading sample-idx mapping from {}".format(sample_f))
    sample_idx = np.load(sample_f, allow_pickle=True, mmap_mode="r")
    print_rank_0(" > loading shuffle-idx mapping from {}".format(shuffle_f))
    shuffle_idx = np.load(shuffle_f, allow_pickle=True, mmap_mode="r")
    print_rank_0("    loaded indexed file in {:3.3f} seconds".format(time.time() - t0))
    print_rank_0("    total number of samples: {}".format(sample_idx.shape[0]))
    print_rank_0("    total number of epochs: {}".format(num_epochs))
    return doc_idx, sample_idx, shuffle_idx
