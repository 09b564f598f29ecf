"""Checkpoint layout algebra shared by the converters (weights2megatron, megatron2hf) and
the resharding tool (tools/checkpoint_util.py).

Tensor layouts (reference weights2megatron.py:80-145,
checkpoint_saver_megatron.py:190-301):
* fused QKV rows are grouped per KV head: [q_0 .. q_{n/g-1}, k, v] x g, each block hn rows;
  Q/K rows inside a head use Meta's interleaved-pair RoPE order (HF uses half-rotation);
* GLU ``dense_h_to_4h`` is cat([up (w3), gate (w1)]) on dim 0; under TP every rank holds
  [up_shard; gate_shard];
* column-parallel weights split on dim 0, row-parallel on dim 1, vocab on dim 0.
"""
from __future__ import annotations

import re
from collections import OrderedDict

import torch


# ----------------------------------------------------------------------------- RoPE layouts
def rope_hf_to_interleaved(w, n_heads, head_dim):
    """HF half-rotation rows -> Meta/Megatron interleaved pairs. w: [n_heads*head_dim, ...]."""
    rest = w.shape[1:]
    return w.view(n_heads, 2, head_dim // 2, *rest).transpose(1, 2).reshape(n_heads * head_dim, *rest)


def rope_interleaved_to_hf(w, n_heads, head_dim):
    rest = w.shape[1:]
    return w.view(n_heads, head_dim // 2, 2, *rest).transpose(1, 2).reshape(n_heads * head_dim, *rest)


def permute_qkv(qkv_w, dim, n_heads, n_heads_kv, revert=False):
    """Convert the RoPE row order of the q and k blocks of a fused QKV weight [rows, dim]
    (v untouched). revert=False: HF order -> interleaved; revert=True: interleaved -> HF."""
    head_dim = dim // n_heads
    qpg = n_heads // n_heads_kv
    f = rope_interleaved_to_hf if revert else rope_hf_to_interleaved
    groups = qkv_w.view(n_heads_kv, qpg + 2, head_dim, *qkv_w.shape[1:])
    out = groups.clone()
    for s in range(qpg + 1):          # q slots and the k slot
        blk = groups[:, s].reshape(n_heads_kv * head_dim, *qkv_w.shape[1:])
        out[:, s] = f(blk, n_heads_kv, head_dim).view(n_heads_kv, head_dim, *qkv_w.shape[1:])
    return out.reshape(qkv_w.shape)


# ----------------------------------------------------------------------------- fused QKV
def fuse_qkv(wq, wk, wv, n_heads, n_heads_kv):
    """[n*hn, h], [g*hn, h], [g*hn, h] -> grouped fused [(n + 2g)*hn, h]."""
    hn = wq.shape[0] // n_heads
    qpg = n_heads // n_heads_kv
    rest = wq.shape[1:]
    q = wq.view(n_heads_kv, qpg, hn, *rest)
    k = wk.view(n_heads_kv, 1, hn, *rest)
    v = wv.view(n_heads_kv, 1, hn, *rest)
    return torch.cat([q, k, v], dim=1).reshape((n_heads + 2 * n_heads_kv) * hn, *rest)


def split_qkv(wqkv, n_heads, n_heads_kv):
    hn = wqkv.shape[0] // (n_heads + 2 * n_heads_kv)
    qpg = n_heads // n_heads_kv
    rest = wqkv.shape[1:]
    g = wqkv.view(n_heads_kv, qpg + 2, hn, *rest)
    return (g[:, :qpg].reshape(n_heads * hn, *rest), g[:, qpg].reshape(n_heads_kv * hn, *rest),
            g[:, qpg + 1].reshape(n_heads_kv * hn, *rest))


# ----------------------------------------------------------------------------- TP split / merge
_COL = re.compile(r"(query_key_value|dense_h_to_4h|\.query|key_value)\.(weight|bias)$")
_ROW_W = re.compile(r"(self_attention\.dense|attention\.dense|inter_attention\.dense|dense_4h_to_h)\.weight$")


def partition_dim(name):
    """0 column-parallel, 1 row-parallel, None replicated."""
    if _COL.search(name):
        return 0
    if _ROW_W.search(name):
        return 1
    return None


def is_glu_h_to_4h(name, glu):
    return bool(glu) and "dense_h_to_4h" in name


def merge_tp(name, shards, glu=False):
    """Full tensor from the per-rank shards of parameter ``name``."""
    d = partition_dim(name)
    if d is None or len(shards) == 1:
        return shards[0].clone()
    if is_glu_h_to_4h(name, glu):
        ups = [s.chunk(2, dim=0)[0] for s in shards]
        gates = [s.chunk(2, dim=0)[1] for s in shards]
        return torch.cat(ups + gates, dim=0)
    return torch.cat(shards, dim=d)


def split_tp(name, full, tp, glu=False):
    d = partition_dim(name)
    if d is None or tp == 1:
        return [full.clone() for _ in range(tp)]
    if is_glu_h_to_4h(name, glu):
        up, gate = full.chunk(2, dim=0)
        return [torch.cat([u, g], dim=0) for u, g in zip(up.chunk(tp, 0), gate.chunk(tp, 0))]
    return [c.clone() for c in full.chunk(tp, dim=d)]


def pad_vocab(w, padded_size):
    """Re-pad a [V, h] table to ``padded_size`` rows (extra rows copy the last real row, as
    the reference saver does, checkpoint_saver_megatron.py:190-226)."""
    v = w.shape[0]
    if v == padded_size:
        return w
    if v > padded_size:
        return w[:padded_size].clone()
    return torch.cat([w, w[-1:].expand(padded_size - v, -1)], dim=0)


def padded_vocab_size(true_size, divisible_by, tp):
    m = divisible_by * tp
    return ((true_size + m - 1) // m) * m


# ----------------------------------------------------------------------------- key normalisation
def normalize_language_model(lm):
    """Legacy aliases -> modern spelling: 'transformer' -> 'encoder', '.attention.' ->
    '.self_attention.', flat 'word_embeddings.weight' -> nested dict."""
    out = OrderedDict()
    emb = lm.get("embedding", {})
    new_emb = OrderedDict()
    for k, v in emb.items():
        if isinstance(v, dict):
            new_emb[k] = OrderedDict(v)
        else:
            top, _, leaf = k.partition(".")
            new_emb.setdefault(top, OrderedDict())[leaf] = v
    if new_emb:
        out["embedding"] = new_emb
    enc = lm.get("encoder", lm.get("transformer"))
    if enc is not None:
        out["encoder"] = OrderedDict((k.replace(".attention.", ".self_attention."), v) for k, v in enc.items())
    for k in ("lm_head", "pooler", "decoder"):
        if k in lm:
            out[k] = lm[k]
    return out


def layer_index(key):
    m = re.match(r"layers\.(\d+)\.(.*)", key)
    return (int(m.group(1)), m.group(2)) if m else (None, key)
