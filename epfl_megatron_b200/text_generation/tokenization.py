"""Prompt tokenisation / detokenisation for generation
(megatron/text_generation/tokenization.py:13-118 surface)."""
from __future__ import annotations

import torch
import torch.distributed as dist

from .. import global_vars
from .communication import _dev, broadcast_int_list, broadcast_tensor


def detokenize_generations(tokens_gpu_tensor, lengths_gpu_tensor, return_segments):
    """-> tokens (lists), texts [, per-token text segments]. Works for any tokenizer that has
    ``detokenize`` (the reference reaches into GPT-2 BPE internals, tokenization.py:31-37)."""
    tokenizer = global_vars.get_tokenizer()
    texts, segments = [], []
    tokens = tokens_gpu_tensor.cpu().numpy().tolist()
    lengths = lengths_gpu_tensor.cpu().numpy().tolist()
    for seq, n in zip(tokens, lengths):
        seq = seq[:n]
        texts.append(tokenizer.detokenize(seq))
        if return_segments:
            inner = getattr(tokenizer, "tokenizer", None)
            if hasattr(inner, "decoder") and hasattr(inner, "byte_decoder"):
                words = []
                for t in seq:
                    w = inner.decoder[t]
                    words.append(bytearray(inner.byte_decoder[c] for c in w).decode("utf-8", errors="replace"))
            else:
                words = [tokenizer.detokenize([t]) for t in seq]
            segments.append(words)
    if return_segments:
        return tokens, texts, segments
    return tokens, texts


def _tokenize_prompts_and_batch(prompts, tokens_to_generate, add_BOS):
    tokenizer = global_vars.get_tokenizer()
    if add_BOS:
        bos = getattr(tokenizer, "bos", None)
        if bos is None:
            bos = tokenizer.eod
        toks = [[bos] + tokenizer.tokenize(p) for p in prompts]
    else:
        toks = [tokenizer.tokenize(p) for p in prompts]
    lengths = [len(t) for t in toks]
    total = max(lengths) + tokens_to_generate
    toks = [t + [tokenizer.eod] * (total - n) for t, n in zip(toks, lengths)]
    dev = _dev()
    return (torch.tensor(toks, dtype=torch.int64, device=dev),
            torch.tensor(lengths, dtype=torch.int64, device=dev))


def tokenize_prompts(prompts=None, tokens_to_generate=None, add_BOS=None, rank=0):
    """Tokenise on `rank`, pad to max prompt + tokens_to_generate with EOD, broadcast."""
    sizes, tok, lens = None, None, None
    me = dist.get_rank() if dist.is_initialized() else 0
    if me == rank:
        assert prompts is not None and tokens_to_generate is not None
        tok, lens = _tokenize_prompts_and_batch(prompts, tokens_to_generate, add_BOS)
        sizes = [tok.size(0), tok.size(1)]
    sizes = broadcast_int_list(2, int_list=sizes, rank=rank)
    b, s = int(sizes[0]), int(sizes[1])
    tok = broadcast_tensor((b, s), torch.int64, tensor=tok, rank=rank)
    lens = broadcast_tensor((b,), torch.int64, tensor=lens, rank=rank)
    return tok, lens
