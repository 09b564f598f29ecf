"""WordPiece tokenisation for BERT-style vocabularies
(the job of megatron/tokenizer/bert_tokenization.py; written from the algorithm's definition).

``FullTokenizer`` = basic tokenisation (unicode cleanup, optional lower-casing + accent
stripping, punctuation and CJK splitting) followed by greedy longest-match-first WordPiece.
"""
from __future__ import annotations

import collections
import unicodedata


def load_vocab(vocab_file):
    """One token per line; the line number is the id."""
    vocab = collections.OrderedDict()
    with open(vocab_file, "r", encoding="utf-8") as f:
        for i, line in enumerate(f):
            tok = line.rstrip("\n").strip()
            if tok == "" and line == "":
                break
            vocab[tok] = i
    return vocab


def whitespace_tokenize(text):
    text = text.strip()
    return text.split() if text else []


def _is_whitespace(ch):
    return ch in " \t\n\r" or unicodedata.category(ch) == "Zs"


def _is_control(ch):
    if ch in "\t\n\r":
        return False
    return unicodedata.category(ch) in ("Cc", "Cf")


def _is_punctuation(ch):
    cp = ord(ch)
    if 33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126:
        return True
    return unicodedata.category(ch).startswith("P")


def _is_cjk(cp):
    return (0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0x20000 <= cp <= 0x2A6DF
            or 0x2A700 <= cp <= 0x2B73F or 0x2B740 <= cp <= 0x2B81F or 0x2B820 <= cp <= 0x2CEAF
            or 0xF900 <= cp <= 0xFAFF or 0x2F800 <= cp <= 0x2FA1F)


class BasicTokenizer:
    def __init__(self, do_lower_case=True):
        self.do_lower_case = do_lower_case

    def tokenize(self, text):
        cleaned = []
        for ch in text:
            cp = ord(ch)
            if cp == 0 or cp == 0xFFFD or _is_control(ch):
                continue
            if _is_cjk(cp):
                cleaned.append(f" {ch} ")
            else:
                cleaned.append(" " if _is_whitespace(ch) else ch)
        out = []
        for tok in whitespace_tokenize("".join(cleaned)):
            if self.do_lower_case:
                tok = "".join(c for c in unicodedata.normalize("NFD", tok.lower())
                              if unicodedata.category(c) != "Mn")
            out.extend(self._split_punct(tok))
        return whitespace_tokenize(" ".join(out))

    @staticmethod
    def _split_punct(tok):
        pieces, cur = [], []
        for ch in tok:
            if _is_punctuation(ch):
                if cur:
                    pieces.append("".join(cur))
                    cur = []
                pieces.append(ch)
            else:
                cur.append(ch)
        if cur:
            pieces.append("".join(cur))
        return pieces


class WordpieceTokenizer:
    def __init__(self, vocab, unk_token="[UNK]", max_input_chars_per_word=200):
        self.vocab, self.unk_token, self.max_chars = vocab, unk_token, max_input_chars_per_word

    def tokenize(self, text):
        out = []
        for word in whitespace_tokenize(text):
            if len(word) > self.max_chars:
                out.append(self.unk_token)
                continue
            pieces, start, bad = [], 0, False
            while start < len(word):
                end = len(word)
                found = None
                while start < end:
                    sub = word[start:end]
                    if start > 0:
                        sub = "##" + sub
                    if sub in self.vocab:
                        found = sub
                        break
                    end -= 1
                if found is None:
                    bad = True
                    break
                pieces.append(found)
                start = end
            out.extend([self.unk_token] if bad else pieces)
        return out


class FullTokenizer:
    def __init__(self, vocab_file, do_lower_case=True):
        self.vocab = load_vocab(vocab_file)
        self.inv_vocab = {v: k for k, v in self.vocab.items()}
        self.basic_tokenizer = BasicTokenizer(do_lower_case=do_lower_case)
        self.wordpiece_tokenizer = WordpieceTokenizer(vocab=self.vocab)

    def tokenize(self, text):
        out = []
        for tok in self.basic_tokenizer.tokenize(text):
            out.extend(self.wordpiece_tokenizer.tokenize(tok))
        return out

    def convert_tokens_to_ids(self, tokens):
        return [self.vocab[t] for t in tokens]

    def convert_ids_to_tokens(self, ids):
        return [self.inv_vocab[i] for i in ids]

    @staticmethod
    def convert_tokens_to_string(tokens, clean_up_tokenization_spaces=True):
        text = " ".join(tokens).replace(" ##", "").strip()
        if clean_up_tokenization_spaces:
            for a, b in ((" .", "."), (" ?", "?"), (" !", "!"), (" ,", ","), (" ' ", "'"), (" n't", "n't"),
                         (" 'm", "'m"), (" 's", "'s"), (" 've", "'ve"), (" 're", "'re")):
                text = text.replace(a, b)
        return text

    def vocab_size(self):
        return len(self.vocab)
