"""Byte-level BPE (GPT-2 vocabulary format: vocab.json + merges.txt)
(the job of megatron/tokenizer/gpt2_tokenization.py; written from the algorithm's definition).

Text is split by the GPT-2 pre-tokenisation pattern, every piece is mapped byte-by-byte onto
printable unicode characters and merged greedily by merge rank.
"""
from __future__ import annotations

import json
from functools import lru_cache

try:
    import regex as re
except ImportError:  # pragma: no cover
    import re

PRETOKENIZE = r"""'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+"""


@lru_cache()
def bytes_to_unicode():
    """Reversible byte -> printable character table: printable latin-1 bytes map to themselves,
    the remaining 68 bytes to code points from 256 upward."""
    keep = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAC + 1)) + list(range(0xAE, 0xFF + 1))
    table, extra = {}, 0
    for b in range(256):
        if b in keep:
            table[b] = chr(b)
        else:
            table[b] = chr(256 + extra)
            extra += 1
    return table


def get_pairs(word):
    return {(a, b) for a, b in zip(word[:-1], word[1:])}


class GPT2Tokenizer:
    def __init__(self, vocab_file, merges_file, errors="replace", special_tokens=None, max_len=None):
        self.max_len = max_len if max_len is not None else int(1e12)
        with open(vocab_file, encoding="utf-8") as f:
            self.encoder = json.load(f)
        self.decoder = {v: k for k, v in self.encoder.items()}
        self.errors = errors
        self.byte_encoder = bytes_to_unicode()
        self.byte_decoder = {v: k for k, v in self.byte_encoder.items()}
        with open(merges_file, encoding="utf-8") as f:
            lines = f.read().split("\n")
        merges = [tuple(ln.split()) for ln in lines if ln and not ln.startswith("#version") and len(ln.split()) == 2]
        self.bpe_ranks = {m: i for i, m in enumerate(merges)}
        self.cache = {}
        self.pat = re.compile(PRETOKENIZE)
        self.special_tokens, self.special_tokens_decoder = {}, {}
        self.set_special_tokens(special_tokens)

    def __len__(self):
        return len(self.encoder) + len(self.special_tokens)

    def set_special_tokens(self, special_tokens):
        """Special tokens get ids after the base vocabulary, in the order given."""
        if not special_tokens:
            self.special_tokens, self.special_tokens_decoder = {}, {}
            return
        self.special_tokens = {t: len(self.encoder) + i for i, t in enumerate(special_tokens)}
        self.special_tokens_decoder = {v: k for k, v in self.special_tokens.items()}

    def bpe(self, token):
        if token in self.cache:
            return self.cache[token]
        word = tuple(token)
        while len(word) > 1:
            pairs = get_pairs(word)
            best = min(pairs, key=lambda p: self.bpe_ranks.get(p, float("inf")))
            if best not in self.bpe_ranks:
                break
            a, b = best
            merged, i = [], 0
            while i < len(word):
                if i < len(word) - 1 and word[i] == a and word[i + 1] == b:
                    merged.append(a + b)
                    i += 2
                else:
                    merged.append(word[i])
                    i += 1
            word = tuple(merged)
        out = " ".join(word)
        self.cache[token] = out
        return out

    def tokenize(self, text):
        pieces = []
        for tok in re.findall(self.pat, text):
            mapped = "".join(self.byte_encoder[b] for b in tok.encode("utf-8"))
            pieces.extend(self.bpe(mapped).split(" "))
        return pieces

    def convert_tokens_to_ids(self, tokens):
        if isinstance(tokens, str):
            return self.special_tokens.get(tokens, self.encoder.get(tokens, 0))
        return [self.special_tokens[t] if t in self.special_tokens else self.encoder.get(t, 0) for t in tokens]

    def convert_ids_to_tokens(self, ids, skip_special_tokens=False):
        out = []
        for i in ids:
            if i in self.special_tokens_decoder:
                if not skip_special_tokens:
                    out.append(self.special_tokens_decoder[i])
            else:
                out.append(self.decoder[i])
        return out

    def encode(self, text):
        return self.convert_tokens_to_ids(self.tokenize(text))

    def decode(self, tokens):
        text = "".join(self.convert_ids_to_tokens(tokens))
        return bytearray(self.byte_decoder[c] for c in text).decode("utf-8", errors=self.errors)
