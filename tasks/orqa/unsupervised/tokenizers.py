"""Minimal regexp tokenizer used for answer matching (the DPR / DrQA "simple" tokenizer;
tasks/orqa/unsupervised/tokenizers.py:17-243 surface, without the spaCy variant)."""
import copy

import regex


class Tokens:
    """(text, text with trailing whitespace, char span) triples with n-gram helpers."""
    TEXT, TEXT_WS, SPAN = 0, 1, 2

    def __init__(self, data, annotators=None, opts=None):
        self.data, self.annotators, self.opts = data, annotators or set(), opts or {}

    def __len__(self):
        return len(self.data)

    def slice(self, i=None, j=None):
        t = copy.copy(self)
        t.data = self.data[i:j]
        return t

    def untokenize(self):
        return "".join(t[self.TEXT_WS] for t in self.data).strip()

    def words(self, uncased=False):
        return [t[self.TEXT].lower() if uncased else t[self.TEXT] for t in self.data]

    def offsets(self):
        return [t[self.SPAN] for t in self.data]

    def ngrams(self, n=1, uncased=False, filter_fn=None, as_strings=True):
        words = self.words(uncased)
        spans = [(s, e + 1) for s in range(len(words)) for e in range(s, min(s + n, len(words)))
                 if not (filter_fn and filter_fn(words[s:e + 1]))]
        return [" ".join(words[s:e]) for s, e in spans] if as_strings else spans


class Tokenizer:
    def tokenize(self, text):
        raise NotImplementedError

    def shutdown(self):
        pass

    def __del__(self):
        self.shutdown()


class SimpleTokenizer(Tokenizer):
    ALPHA_NUM = r"[\p{L}\p{N}\p{M}]+"
    NON_WS = r"[^\p{Z}\p{C}]"

    def __init__(self, **kwargs):
        self._regexp = regex.compile("(%s)|(%s)" % (self.ALPHA_NUM, self.NON_WS),
                                     flags=regex.IGNORECASE + regex.UNICODE + regex.MULTILINE)
        self.annotators = set()

    def tokenize(self, text):
        matches = list(self._regexp.finditer(text))
        data = []
        for i, m in enumerate(matches):
            start, end = m.span()
            ws_end = matches[i + 1].span()[0] if i + 1 < len(matches) else end
            data.append((m.group(), text[start:ws_end], (start, end)))
        return Tokens(data, self.annotators)
