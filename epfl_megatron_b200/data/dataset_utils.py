"""Dataset helpers shared by the GPT / BERT / T5 / ICT pipelines
(megatron/data/dataset_utils.py:44-729 surface): blend weights, split parsing, masked-LM
example construction, sentence-window sample mappings."""
from __future__ import annotations

import collections
import math
import os
import time

import numpy as np
import torch
import torch.distributed as dist

from ..utils import print_rank_0

DSET_TYPE_BERT = "standard_bert"
DSET_TYPE_ICT = "ict"
DSET_TYPE_T5 = "t5"
DSET_TYPES = [DSET_TYPE_BERT, DSET_TYPE_ICT, DSET_TYPE_T5]


def get_datasets_weights_and_num_samples(data_prefix, train_valid_test_num_samples):
    """``[w0, path0, w1, path1, ...]`` -> normalised weights and per-dataset sample counts
    (0.5 % head-room so the blend never runs dry)."""
    assert len(data_prefix) % 2 == 0
    n = len(data_prefix) // 2
    weights = [float(data_prefix[2 * i]) for i in range(n)]
    prefixes = [data_prefix[2 * i + 1].strip() for i in range(n)]
    s = sum(weights)
    assert s > 0.0
    weights = [w / s for w in weights]
    if isinstance(train_valid_test_num_samples, list):
        per = [[int(math.ceil(v * w * 1.005)) for v in train_valid_test_num_samples] for w in weights]
    else:
        per = [int(math.ceil(train_valid_test_num_samples * w * 1.005)) for w in weights]
    return prefixes, weights, per


def get_train_valid_test_split_(splits_string, size):
    """'969, 30, 1' (or '/'-separated) -> four document boundaries."""
    if splits_string.find(",") != -1:
        splits = [float(s) for s in splits_string.split(",")]
    elif splits_string.find("/") != -1:
        splits = [float(s) for s in splits_string.split("/")]
    else:
        splits = [float(splits_string)]
    while len(splits) < 3:
        splits.append(0.0)
    splits = splits[:3]
    total = sum(splits)
    assert total > 0.0
    splits = [s / total for s in splits]
    idx = [0]
    for s in splits:
        idx.append(idx[-1] + int(round(s * float(size))))
    diff = idx[-1] - size
    for i in range(1, len(idx)):
        idx[i] -= diff
    assert len(idx) == 4 and idx[-1] == size
    return idx


def get_a_and_b_segments(sample, np_rng):
    """Split a list of sentences into segments A and B; swap them half of the time."""
    n = len(sample)
    assert n > 1, "make sure each sample has at least two sentences."
    a_end = 1
    if n >= 3:
        a_end = np_rng.randint(1, n)
    tokens_a, tokens_b = [], []
    for j in range(a_end):
        tokens_a.extend(sample[j])
    for j in range(a_end, n):
        tokens_b.extend(sample[j])
    is_next_random = False
    if np_rng.random() < 0.5:
        is_next_random = True
        tokens_a, tokens_b = tokens_b, tokens_a
    return tokens_a, tokens_b, is_next_random


def truncate_segments(tokens_a, tokens_b, len_a, len_b, max_num_tokens, np_rng):
    assert len_a > 0
    if len_a + len_b <= max_num_tokens:
        return False
    while len_a + len_b > max_num_tokens:
        if len_a > len_b:
            len_a -= 1
            tokens = tokens_a
        else:
            len_b -= 1
            tokens = tokens_b
        if np_rng.random() < 0.5:
            del tokens[0]
        else:
            tokens.pop()
    return True


def create_tokens_and_tokentypes(tokens_a, tokens_b, cls_id, sep_id):
    tokens = [cls_id] + list(tokens_a) + [sep_id]
    types = [0] * (len(tokens_a) + 2)
    if tokens_b:
        tokens += list(tokens_b) + [sep_id]
        types += [1] * (len(tokens_b) + 1)
    return tokens, types


MaskedLmInstance = collections.namedtuple("MaskedLmInstance", ["index", "label"])


def is_start_piece(piece):
    return not piece.startswith("##")


def create_masked_lm_predictions(tokens, vocab_id_list, vocab_id_to_token_dict, masked_lm_prob, cls_id, sep_id,
                                 mask_id, max_predictions_per_seq, np_rng, max_ngrams=3, do_whole_word_mask=True,
                                 favor_longer_ngram=False, do_permutation=False, geometric_dist=False,
                                 masking_style="bert"):
    """Whole-word / n-gram masking. Returns (output tokens, masked positions, masked labels,
    token boundaries, masked spans)."""
    cand_indexes, token_boundary = [], [0] * len(tokens)
    for i, tok in enumerate(tokens):
        if tok == cls_id or tok == sep_id:
            token_boundary[i] = 1
            continue
        if do_whole_word_mask and cand_indexes and not is_start_piece(vocab_id_to_token_dict[tok]):
            cand_indexes[-1].append(i)
        else:
            cand_indexes.append([i])
            if is_start_piece(vocab_id_to_token_dict[tok]):
                token_boundary[i] = 1
    output = list(tokens)
    if masked_lm_prob == 0:
        return output, [], [], token_boundary, []
    num_to_predict = min(max_predictions_per_seq, max(1, int(round(len(tokens) * masked_lm_prob))))
    ngrams = np.arange(1, max_ngrams + 1, dtype=np.int64)
    if not geometric_dist:
        pvals = 1.0 / np.arange(1, max_ngrams + 1)
        pvals /= pvals.sum(keepdims=True)
        if favor_longer_ngram:
            pvals = pvals[::-1]
    ngram_indexes = []
    for idx in range(len(cand_indexes)):
        ngram_indexes.append([cand_indexes[idx:idx + n] for n in ngrams])
    np_rng.shuffle(ngram_indexes)
    masked_lms, masked_spans, covered = [], [], set()
    for cand_set in ngram_indexes:
        if len(masked_lms) >= num_to_predict or not cand_set:
            if len(masked_lms) >= num_to_predict:
                break
            continue
        if any(i in covered for s in cand_set[0] for i in s):
            continue
        if not geometric_dist:
            n = np_rng.choice(ngrams[:len(cand_set)],
                              p=pvals[:len(cand_set)] / pvals[:len(cand_set)].sum(keepdims=True))
        else:
            n = min(np_rng.geometric(0.2), max_ngrams)
        index_set = sum(cand_set[n - 1], [])
        n -= 1
        while len(masked_lms) + len(index_set) > num_to_predict and n > 0:
            index_set = sum(cand_set[n - 1], [])
            n -= 1
        if len(masked_lms) + len(index_set) > num_to_predict or any(i in covered for i in index_set):
            continue
        for i in index_set:
            covered.add(i)
            if masking_style == "bert":
                if np_rng.random() < 0.8:
                    new = mask_id
                elif np_rng.random() < 0.5:
                    new = tokens[i]
                else:
                    new = vocab_id_list[np_rng.randint(0, len(vocab_id_list))]
            elif masking_style == "t5":
                new = mask_id
            else:
                raise ValueError("invalid value of masking style")
            output[i] = new
            masked_lms.append(MaskedLmInstance(index=i, label=tokens[i]))
        masked_spans.append(MaskedLmInstance(index=index_set, label=[tokens[i] for i in index_set]))
    assert len(masked_lms) <= num_to_predict
    np_rng.shuffle(ngram_indexes)
    if do_permutation:
        select = set()
        for cand_set in ngram_indexes:
            if len(select) >= num_to_predict:
                break
            if not cand_set or any(i in covered or i in select for s in cand_set[0] for i in s):
                continue
            n = np.random.choice(ngrams[:len(cand_set)],
                                 p=pvals[:len(cand_set)] / pvals[:len(cand_set)].sum(keepdims=True))
            index_set = sum(cand_set[n - 1], [])
            n -= 1
            while len(select) + len(index_set) > num_to_predict and n > 0:
                index_set = sum(cand_set[n - 1], [])
                n -= 1
            if len(select) + len(index_set) > num_to_predict or any(i in covered or i in select for i in index_set):
                continue
            select.update(index_set)
        select = sorted(select)
        perm = list(select)
        np_rng.shuffle(perm)
        orig = list(output)
        for src, tgt in zip(select, perm):
            output[src] = orig[tgt]
            masked_lms.append(MaskedLmInstance(index=src, label=orig[src]))
    masked_lms = sorted(masked_lms, key=lambda x: x.index)
    masked_spans = sorted(masked_spans, key=lambda x: x.index[0])
    return (output, [p.index for p in masked_lms], [p.label for p in masked_lms], token_boundary, masked_spans)


def pad_and_convert_to_numpy(tokens, tokentypes, masked_positions, masked_labels, pad_id, max_seq_length):
    n = len(tokens)
    pad = max_seq_length - n
    assert pad >= 0 and len(tokentypes) == n and len(masked_positions) == len(masked_labels)
    tokens_np = np.array(list(tokens) + [pad_id] * pad, dtype=np.int64)
    types_np = np.array(list(tokentypes) + [pad_id] * pad, dtype=np.int64)
    padding_mask = np.array([1] * n + [0] * pad, dtype=np.int64)
    labels = [-1] * max_seq_length
    loss_mask = [0] * max_seq_length
    for pos, lab in zip(masked_positions, masked_labels):
        assert pos < n
        labels[pos] = lab
        loss_mask[pos] = 1
    return tokens_np, types_np, np.array(labels, dtype=np.int64), padding_mask, np.array(loss_mask, dtype=np.int64)


def get_indexed_dataset_(data_prefix, data_impl, skip_warmup):
    from .indexed_dataset import make_dataset
    print_rank_0(" > building dataset index ...")
    t0 = time.time()
    ds = make_dataset(data_prefix, data_impl, skip_warmup)
    assert ds.sizes.shape[0] == ds.doc_idx[-1]
    print_rank_0(" > finished creating indexed dataset in {:4f} seconds".format(time.time() - t0))
    print_rank_0(" > indexed dataset stats:")
    print_rank_0("    number of documents: {}".format(ds.doc_idx.shape[0] - 1))
    print_rank_0("    number of sentences: {}".format(ds.sizes.shape[0]))
    return ds


def build_train_valid_test_datasets(data_prefix, data_impl, splits_string, train_valid_test_num_samples,
                                    max_seq_length, masked_lm_prob, short_seq_prob, seed, skip_warmup,
                                    binary_head=False, max_seq_length_dec=None, dataset_type="standard_bert"):
    if len(data_prefix) == 1:
        return _build_train_valid_test_datasets(data_prefix[0], data_impl, splits_string,
                                                train_valid_test_num_samples, max_seq_length, masked_lm_prob,
                                                short_seq_prob, seed, skip_warmup, binary_head,
                                                max_seq_length_dec, dataset_type=dataset_type)
    from .blendable_dataset import BlendableDataset
    prefixes, weights, per = get_datasets_weights_and_num_samples(data_prefix, train_valid_test_num_samples)
    groups = ([], [], [])
    for i, prefix in enumerate(prefixes):
        tvt = _build_train_valid_test_datasets(prefix, data_impl, splits_string, per[i], max_seq_length,
                                               masked_lm_prob, short_seq_prob, seed, skip_warmup, binary_head,
                                               max_seq_length_dec, dataset_type=dataset_type)
        for g, ds in zip(groups, tvt):
            if ds:
                g.append(ds)
    return tuple(BlendableDataset(g, weights) if g else None for g in groups)


def _build_train_valid_test_datasets(data_prefix, data_impl, splits_string, train_valid_test_num_samples,
                                     max_seq_length, masked_lm_prob, short_seq_prob, seed, skip_warmup,
                                     binary_head, max_seq_length_dec, dataset_type="standard_bert"):
    if dataset_type not in DSET_TYPES:
        raise ValueError("Invalid dataset_type: ", dataset_type)
    indexed = get_indexed_dataset_(data_prefix, data_impl, skip_warmup)
    title_dataset = None
    if dataset_type == DSET_TYPE_ICT:
        from .. import global_vars
        title_dataset = get_indexed_dataset_(global_vars.get_args().titles_data_path, data_impl, skip_warmup)
    total_docs = indexed.doc_idx.shape[0] - 1
    splits = get_train_valid_test_split_(splits_string, total_docs)
    print_rank_0(" > dataset split:")
    for i, name in enumerate(("train", "validation", "test")):
        print_rank_0("    {}: document indices in [{}, {}) total of {} documents".format(
            name, splits[i], splits[i + 1], splits[i + 1] - splits[i]))

    def build(i, name):
        from .bert_dataset import BertDataset
        from .t5_dataset import T5Dataset
        from .ict_dataset import ICTDataset
        from .. import global_vars
        if splits[i + 1] <= splits[i]:
            return None
        saved = indexed.get_doc_idx()
        indexed.set_doc_idx(saved[splits[i]:splits[i + 1] + 1])
        kw = dict(name=name, data_prefix=data_prefix, num_epochs=None,
                  max_num_samples=train_valid_test_num_samples[i], max_seq_length=max_seq_length, seed=seed)
        if dataset_type == DSET_TYPE_ICT:
            a = global_vars.get_args()
            ds = ICTDataset(block_dataset=indexed, title_dataset=title_dataset,
                            query_in_block_prob=a.query_in_block_prob, use_one_sent_docs=a.use_one_sent_docs,
                            binary_head=binary_head, **kw)
        elif dataset_type == DSET_TYPE_T5:
            ds = T5Dataset(indexed_dataset=indexed, masked_lm_prob=masked_lm_prob,
                           max_seq_length_dec=max_seq_length_dec, short_seq_prob=short_seq_prob, **kw)
        else:
            ds = BertDataset(indexed_dataset=indexed, masked_lm_prob=masked_lm_prob,
                             short_seq_prob=short_seq_prob, binary_head=binary_head, **kw)
        indexed.set_doc_idx(saved)
        assert indexed.doc_idx[0] == 0 and indexed.doc_idx.shape[0] == total_docs + 1
        return ds

    return build(0, "train"), build(1, "valid"), build(2, "test")


def get_samples_mapping(indexed_dataset, data_prefix, num_epochs, max_num_samples, max_seq_length, short_seq_prob,
                        seed, name, binary_head):
    """Sentence-window sample mapping, built on rank 0 with the native helper and cached."""
    from ..parallel import state as mpu
    from . import helpers
    if not num_epochs:
        if not max_num_samples:
            raise ValueError("Need to specify either max_num_samples or num_epochs")
        num_epochs = np.iinfo(np.int32).max - 1
    if not max_num_samples:
        max_num_samples = np.iinfo(np.int64).max - 1
    fname = data_prefix + "_{}_indexmap".format(name)
    if num_epochs != (np.iinfo(np.int32).max - 1):
        fname += "_{}ep".format(num_epochs)
    if max_num_samples != (np.iinfo(np.int64).max - 1):
        fname += "_{}mns".format(max_num_samples)
    fname += "_{}msl_{:0.2f}ssp_{}s.npy".format(max_seq_length, short_seq_prob, seed)
    rank0 = not dist.is_initialized() or dist.get_rank() == 0
    if rank0 and not os.path.isfile(fname):
        print(" > WARNING: could not find index map file {}, building the indices on rank 0 ...".format(fname))
        assert indexed_dataset.doc_idx.dtype == np.int64 and indexed_dataset.sizes.dtype == np.int32
        t0 = time.time()
        mapping = helpers.build_mapping(indexed_dataset.doc_idx, indexed_dataset.sizes, num_epochs, max_num_samples,
                                        max_seq_length, short_seq_prob, seed, rank0, 2 if binary_head else 1)
        np.save(fname, mapping, allow_pickle=True)
        print_rank_0(" > elasped time to build and save samples mapping (seconds): {:4f}".format(time.time() - t0))
    if dist.is_initialized():
        dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
        c = torch.ones(1, dtype=torch.long, device=dev)
        dist.all_reduce(c, group=mpu.get_data_parallel_group())
        dist.all_reduce(c, group=mpu.get_pipeline_model_parallel_group())
    print_rank_0(" > loading indexed mapping from {}".format(fname))
    mapping = np.load(fname, allow_pickle=True, mmap_mode="r")
    print_rank_0("    total number of samples: {}".format(mapping.shape[0]))
    return mapping
