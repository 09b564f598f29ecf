"""Data pipeline on CPU: indexed-dataset formats (byte layout of the mmap index included, it is part
of the public surface: SURVEY.md section 2.3), the native index builders against their Python
restatements, GPT sample construction, blending, samplers, preprocessing and merging tools.
Reference behaviour: megatron/data/{indexed_dataset,gpt_dataset,blendable_dataset,data_samplers}.py,
megatron/data/helpers.cpp, tools/{preprocess_data,merge_datasets}.py."""
import json
import os
import struct
import subprocess
import sys

import numpy as np
import pytest
import torch

from epfl_megatron_b200.data import helpers
from epfl_megatron_b200.data import indexed_dataset as ids
from epfl_megatron_b200.data.blendable_dataset import BlendableDataset
from epfl_megatron_b200.data.data_samplers import MegatronPretrainingRandomSampler, MegatronPretrainingSampler
from epfl_megatron_b200.data.gpt_dataset import GPTDataset, _num_epochs, build_train_valid_test_datasets

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _write_mmap(prefix, docs, dtype=np.uint16):
    """docs: list of documents, each a list of sentences (lists of ints)."""
    b = ids.MMapIndexedDatasetBuilder(ids.data_file_path(prefix), dtype=dtype)
    for doc in docs:
        for sent in doc:
            b.add_item(torch.tensor(sent, dtype=torch.int64))
        b.end_document()
    b.finalize(ids.index_file_path(prefix))


def _corpus(rng, n_docs=40, lo=5, hi=60, vocab=1000):
    return [[list(map(int, rng.integers(1, vocab, size=int(rng.integers(lo, hi)))))] for _ in range(n_docs)]


def test_mmap_index_byte_layout_and_roundtrip(tmp_path):
    prefix = str(tmp_path / "ds")
    docs = [[[1, 2, 3], [4, 5]], [[6]], [[7, 8, 9, 10]]]
    _write_mmap(prefix, docs)
    raw = open(ids.index_file_path(prefix), "rb").read()
    assert raw[:9] == b"MMIDIDX\x00\x00"
    version, = struct.unpack("<Q", raw[9:17])
    code, = struct.unpack("<B", raw[17:18])
    n_items, n_docs = struct.unpack("<QQ", raw[18:34])
    assert (version, n_items, n_docs) == (1, 4, 4)            # doc_idx has n_documents + 1 entries
    assert ids.code(np.uint16) == code
    sizes = np.frombuffer(raw, dtype=np.int32, count=4, offset=34)
    pointers = np.frombuffer(raw, dtype=np.int64, count=4, offset=34 + 16)
    doc_idx = np.frombuffer(raw, dtype=np.int64, count=4, offset=34 + 16 + 32)
    assert sizes.tolist() == [3, 2, 1, 4] and pointers.tolist() == [0, 6, 10, 12]
    assert doc_idx.tolist() == [0, 2, 3, 4]
    ds = ids.make_dataset(prefix, "mmap", skip_warmup=True)
    assert len(ds) == 4 and ds[3].tolist() == [7, 8, 9, 10] and ds.sizes.tolist() == [3, 2, 1, 4]
    assert ds.get(3, offset=1, length=2).tolist() == [8, 9] and ds.get(0, offset=2).tolist() == [3]
    assert [x.tolist() for x in ds[1:3]] == [[4, 5], [6]]
    assert ids.infer_dataset_impl(prefix) == "mmap" and ids.MMapIndexedDataset.exists(prefix)
    assert ids.best_fitting_dtype(32000) == np.uint16 and ids.best_fitting_dtype(70000) == np.int32


@pytest.mark.parametrize("impl", ["lazy", "cached"])
def test_legacy_indexed_formats(tmp_path, impl):
    prefix = str(tmp_path / "legacy")
    b = ids.make_builder(ids.data_file_path(prefix), "lazy", vocab_size=500)
    items = [[5, 6, 7], [8], [9, 10]]
    for it in items:
        b.add_item(torch.tensor(it, dtype=torch.int64))
        b.end_document()
    b.finalize(ids.index_file_path(prefix))
    assert open(ids.index_file_path(prefix), "rb").read(8) == b"TNTIDX\x00\x00"
    ds = ids.make_dataset(prefix, impl)
    if impl == "cached":
        ds.prefetch([0, 1, 2])
    assert [ds[i].tolist() for i in range(3)] == items and ids.infer_dataset_impl(prefix) == "cached"


def test_native_sample_idx_matches_python_restatement():
    assert helpers.native_available(), "libb200data.so must be built (python -m epfl_megatron_b200.ops.build)"
    rng = np.random.default_rng(0)
    sizes = rng.integers(1, 90, size=300).astype(np.int32)
    for seq_length, epochs in ((16, 1), (64, 3), (7, 2)):
        doc_idx = np.concatenate([rng.permutation(300) for _ in range(epochs)]).astype(np.int32)
        tokens_per_epoch = int(sizes.sum())
        a = helpers.build_sample_idx(sizes, doc_idx, seq_length, epochs, tokens_per_epoch)
        b = helpers.build_sample_idx_py(sizes, doc_idx, seq_length, epochs, tokens_per_epoch)
        assert a.shape == b.shape == ((epochs * tokens_per_epoch - 1) // seq_length + 1, 2)
        assert np.array_equal(a, b)


def test_gpt_dataset_samples_tile_the_token_stream(tmp_path):
    rng = np.random.default_rng(1)
    docs = _corpus(rng)
    prefix = str(tmp_path / "gpt")
    _write_mmap(prefix, docs)
    indexed = ids.make_dataset(prefix, "mmap", skip_warmup=True)
    seq = 32
    documents = np.arange(len(docs), dtype=np.int32)
    total = sum(len(d[0]) for d in docs)
    n = (total - 1) // seq
    ds = GPTDataset("train", prefix, documents, indexed, n, seq, seed=7)
    assert len(ds) == n and _num_epochs(total, seq, n) == 1
    # un-shuffled, consecutive samples overlap by one token and reproduce the (shuffled-doc) stream
    stream = np.concatenate([np.asarray(docs[d][0]) for d in ds.doc_idx])
    inv = np.argsort(ds.shuffle_idx)
    for k in (0, 1, n // 2, n - 1):
        s = ds[int(inv[k])]["text"]
        assert s.dtype == np.int64 and s.shape == (seq + 1,)
        assert np.array_equal(s, stream[k * seq:k * seq + seq + 1])
    # the index maps are cached next to the data and reused
    cached = [f for f in os.listdir(tmp_path) if f.endswith(".npy")]
    assert len(cached) == 3
    again = GPTDataset("train", prefix, documents, indexed, n, seq, seed=7)
    assert np.array_equal(again.shuffle_idx, ds.shuffle_idx) and np.array_equal(again.sample_idx, ds.sample_idx)
    # more samples than one epoch holds -> several epochs of documents
    multi = GPTDataset("train", prefix, documents, indexed, 3 * n, seq, seed=7)
    assert len(multi) >= 3 * n and multi.doc_idx.shape[0] % len(docs) == 0


def test_split_and_blend(tmp_path):
    rng = np.random.default_rng(2)
    pa, pb = str(tmp_path / "a"), str(tmp_path / "b")
    _write_mmap(pa, _corpus(rng, 60))
    _write_mmap(pb, _corpus(rng, 30))
    train, valid, test = build_train_valid_test_datasets([pa], "mmap", "80,10,10", [20, 4, 4], 16, 3, True)
    assert len(train) >= 20 and len(valid) >= 4 and len(test) >= 4
    blend, _, _ = build_train_valid_test_datasets(["0.75", pa, "0.25", pb], "mmap", "100,0,0", [40, 0, 0], 16, 3, True)
    assert isinstance(blend, BlendableDataset) and blend[0]["text"].shape == (17,)
    # greedy max-error assignment keeps every prefix within one sample of the requested mixture
    idx = blend.dataset_index[:40].astype(np.int64)
    frac = np.cumsum(idx == 0) / np.arange(1, 41)
    assert abs(frac[-1] - 0.75) < 0.03 and np.all(np.abs(np.cumsum(idx == 0) - 0.75 * np.arange(1, 41)) <= 1.0)
    # native blending == Python restatement
    w = np.array([0.5, 0.3, 0.2])
    di, dsi = np.zeros(200, np.uint8), np.zeros(200, np.int64)
    helpers.build_blending_indices(di, dsi, w, 3, 200, False)
    cur, ref_i, ref_s = np.zeros(3, np.int64), [], []
    for k in range(200):
        err = w * max(k, 1) - cur
        j = int(np.argmax(err))
        ref_i.append(j)
        ref_s.append(cur[j])
        cur[j] += 1
    assert di.tolist() == ref_i and dsi.tolist() == ref_s


def test_pretraining_samplers():
    got = [list(MegatronPretrainingSampler(22, 4, 2, r, 3)) for r in range(3)]
    # chunks of mbs*dp = 6 starting at the consumed sample; rank r takes its slice; tail dropped
    assert got[0] == [[4, 5], [10, 11], [16, 17]] and got[2] == [[8, 9], [14, 15], [20, 21]]
    keep = list(MegatronPretrainingSampler(9, 0, 2, 1, 2, drop_last=False))
    assert keep == [[2, 3], [6, 7], []]
    a = list(MegatronPretrainingRandomSampler(None, 40, 0, 2, 0, 2, False))
    b = list(MegatronPretrainingRandomSampler(None, 40, 0, 2, 1, 2, False))
    flat_a, flat_b = sum(a, []), sum(b, [])
    assert len(flat_a) == len(flat_b) == 20 and not set(flat_a) & set(flat_b)
    # resuming mid-epoch continues the same permutation
    resumed = list(MegatronPretrainingRandomSampler(None, 40, 8, 2, 0, 2, False))
    assert sum(resumed, []) == flat_a[4:]


def test_preprocess_and_merge_tools(tmp_path):
    jsonl = tmp_path / "in.jsonl"
    texts = ["hello world . this is a test", "another document here", "third one"]
    jsonl.write_text("\n".join(json.dumps({"text": t}) for t in texts) + "\n")
    vocab = tmp_path / "vocab.json"
    merges = tmp_path / "merges.txt"
    chars = sorted(set("".join(texts)))
    # byte-level GPT-2 vocabulary over the characters that occur (space is "Ġ")
    toks = [c if c != " " else "Ġ" for c in chars] + ["<|endoftext|>"]
    vocab.write_text(json.dumps({t: i for i, t in enumerate(toks)}))
    merges.write_text("#version: 0.2\n")
    out = str(tmp_path / "out")
    cmd = [sys.executable, os.path.join(ROOT, "tools", "preprocess_data.py"), "--input", str(jsonl),
           "--output_prefix", out, "--tokenizer_type", "GPT2BPETokenizer", "--vocab_file", str(vocab),
           "--merge_file", str(merges), "--append_eod", "--workers", "1", "--chunk_size", "1"]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    prefix = out + "_text_document"
    ds = ids.make_dataset(prefix, "mmap", skip_warmup=True)
    assert len(ds) == 3 and [int(x[-1]) for x in ds] == [len(toks) - 1] * 3
    assert len(ds[0]) == len(texts[0]) + 1
    merged = str(tmp_path / "merged")
    in_dir = tmp_path / "parts"
    in_dir.mkdir()
    for k in range(2):
        for ext in (".bin", ".idx"):
            (in_dir / f"p{k}{ext}").write_bytes(open(prefix + ext, "rb").read())
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "merge_datasets.py"), "--input", str(in_dir),
                        "--output_prefix", merged], capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    both = ids.make_dataset(merged, "mmap", skip_warmup=True)
    assert len(both) == 6 and both.get_doc_idx().tolist() == [0, 1, 2, 3, 4, 5, 6]
    assert both[4].tolist() == ds[1].tolist()


def test_retrieval_datastore_shards_and_mips_search(tmp_path):
    """REALM / ORQA evidence store: per-rank shards are merged into one pickle, and the
    maximum-inner-product index (FAISS when importable, dense torch otherwise) returns the blocks with
    the largest dot products (megatron/data/realm_index.py:17-224 behaviour)."""
    from epfl_megatron_b200.data.realm_index import FaissMIPSIndex, OpenRetreivalDataStore
    rng = np.random.default_rng(3)
    emb = rng.standard_normal((40, 16)).astype(np.float32)
    path = str(tmp_path / "blocks.pkl")
    for rank in range(2):                                 # two data-parallel ranks embed half each
        store = OpenRetreivalDataStore(path, load_from_path=False, rank=rank)
        ids = list(range(rank * 20, rank * 20 + 20))
        store.add_block_data(ids, emb[ids])
        with pytest.raises(ValueError):
            store.add_block_data(ids[:1], emb[ids[:1]])
        store.save_shard()
    merged = OpenRetreivalDataStore(path, load_from_path=False, rank=0)
    merged.merge_shards_and_save()
    assert len(merged.embed_data) == 40 and not os.path.isdir(merged.temp_dir_name)
    loaded = OpenRetreivalDataStore(path, load_from_path=True, rank=0)
    assert sorted(loaded.embed_data) == list(range(40)) and loaded.embed_data[7].dtype == np.float16
    index = FaissMIPSIndex(16, loaded)
    assert loaded.embed_data == {}                         # handed over to the index
    q = emb[[5, 31]] * 3.0
    scores, ids = index.search_mips_index(torch.from_numpy(q), top_k=4, reconstruct=False)
    want = np.argsort(-(q @ np.float32(np.float16(emb)).T), axis=1)[:, :4]
    assert ids.shape == (2, 4) and np.array_equal(ids, want) and ids[0, 0] == 5 and ids[1, 0] == 31
    assert np.all(np.diff(scores, axis=1) <= 1e-4)
    blocks = index.search_mips_index(q, top_k=2, reconstruct=True)
    got = blocks[-1] if isinstance(blocks, tuple) else blocks
    assert np.asarray(got).shape == (2, 2, 16)
