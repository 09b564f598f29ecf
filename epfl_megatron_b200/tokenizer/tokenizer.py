"""Tokenizer factory and wrappers (megatron/tokenizer/tokenizer.py:12-498 surface):
BERT WordPiece, GPT-2 byte-level BPE, SentencePiece (Llama; optional extra special
tokens), Falcon (HF tokenizer loaded from a local path), plus a ``NullTokenizer`` for
synthetic-data runs. ``build_tokenizer`` also derives ``args.padded_vocab_size``."""
from __future__ import annotations

from abc import ABC, abstractmethod


def build_tokenizer(args):
    if args.rank == 0:
        print(f"> building {args.tokenizer_type} tokenizer ...", flush=True)
    t = args.tokenizer_type
    if t in ("BertWordPieceLowerCase", "BertWordPieceCase"):
        assert args.vocab_file is not None
        tok = _BertWordPieceTokenizer(vocab_file=args.vocab_file, lower_case=t.endswith("LowerCase"),
                                      vocab_extra_ids=args.vocab_extra_ids)
    elif t == "GPT2BPETokenizer":
        assert args.vocab_file is not None and args.merge_file is not None
        tok = _GPT2BPETokenizer(args.vocab_file, args.merge_file)
    elif t == "SentencePieceTokenizer":
        assert args.vocab_file is not None
        tok = _SentencePieceTokenizer(args.vocab_file, vocab_extra_ids=args.vocab_extra_ids,
                                      vocab_extra_ids_list=args.vocab_extra_ids_list,
                                      new_tokens=args.new_tokens)
    elif t == "FalconTokenizer":
        tok = _FalconTokenizer(vocab_extra_ids_list=args.vocab_extra_ids_list, new_tokens=args.new_tokens,
                               path=getattr(args, "vocab_file", None) or getattr(args, "tokenizer_model", None))
    elif t == "NullTokenizer" or t is None:
        size = getattr(args, "padded_vocab_size", None) or getattr(args, "vocab_size", None) or 32000
        tok = _NullTokenizer(size)
    else:
        raise NotImplementedError(f"{t} tokenizer is not implemented.")
    args.padded_vocab_size = _vocab_size_with_padding(tok.vocab_size, args)
    return tok


def _vocab_size_with_padding(orig_vocab_size, args):
    """Pad so the vocabulary splits evenly over TP ranks in multiples of
    make_vocab_size_divisible_by (tokenizer.py:49-62 rule)."""
    multiple = args.make_vocab_size_divisible_by * args.tensor_model_parallel_size
    after = orig_vocab_size
    while after % multiple != 0:
        after += 1
    if args.rank == 0:
        print(f" > padded vocab (size: {orig_vocab_size}) with {after - orig_vocab_size} dummy tokens "
              f"(new size: {after})", flush=True)
    return after


class AbstractTokenizer(ABC):
    def __init__(self, name):
        self.name = name
        super().__init__()

    @property
    @abstractmethod
    def vocab_size(self):
        pass

    @property
    @abstractmethod
    def vocab(self):
        pass

    @property
    @abstractmethod
    def inv_vocab(self):
        pass

    @abstractmethod
    def tokenize(self, text):
        pass

    def detokenize(self, token_ids):
        raise NotImplementedError(f"detokenizer is not implemented for {self.name} tokenizer")

    def _na(self, what):
        raise NotImplementedError(f"{what} is not provided for {self.name} tokenizer")

    cls = property(lambda self: self._na("CLS"))
    sep = property(lambda self: self._na("SEP"))
    pad = property(lambda self: self._na("PAD"))
    eod = property(lambda self: self._na("EOD"))
    mask = property(lambda self: self._na("MASK"))


class _NullTokenizer(AbstractTokenizer):
    """Whitespace-separated integer ids; the last id is EOD. For synthetic-data runs."""

    def __init__(self, vocab_size):
        super().__init__("Null")
        self._vocab_size = int(vocab_size)

    vocab_size = property(lambda self: self._vocab_size)
    vocab = property(lambda self: {str(i): i for i in range(self._vocab_size)})
    inv_vocab = property(lambda self: {i: str(i) for i in range(self._vocab_size)})
    eod = property(lambda self: self._vocab_size - 1)
    pad = property(lambda self: 0)
    bos = property(lambda self: 1)
    eos = property(lambda self: self._vocab_size - 1)

    def tokenize(self, text):
        return [int(t) for t in text.split()]

    def detokenize(self, ids):
        return " ".join(str(int(i)) for i in ids)


class _BertWordPieceTokenizer(AbstractTokenizer):
    def __init__(self, vocab_file, lower_case=True, vocab_extra_ids=0):
        super().__init__("BERT Lower Case" if lower_case else "BERT Upper Case")
        from .bert_tokenization import FullTokenizer
        self.tokenizer = FullTokenizer(vocab_file, do_lower_case=lower_case)
        self.cls_id = self.tokenizer.vocab["[CLS]"]
        self.sep_id = self.tokenizer.vocab["[SEP]"]
        self.pad_id = self.tokenizer.vocab["[PAD]"]
        self.mask_id = self.tokenizer.vocab["[MASK]"]
        self._additional_special_tokens = []
        self._bos_token, self._eos_token = "[BOS]", "[EOS]"
        self.add_token(self._bos_token)
        self._bos_token_id = self.vocab.get(self._bos_token)
        self.add_token(self._eos_token)
        self._eos_token_id = self.vocab.get(self._eos_token)
        self.add_additional_special_tokens([f"<extra_id_{i}>" for i in range(vocab_extra_ids)])

    def add_token(self, token):
        if token not in self.vocab:
            self.inv_vocab[self.vocab_size] = token
            self.vocab[token] = self.vocab_size

    def add_additional_special_tokens(self, tokens):
        self._additional_special_tokens.extend(tokens)
        for t in tokens:
            self.add_token(t)

    vocab_size = property(lambda self: self.tokenizer.vocab_size())
    vocab = property(lambda self: self.tokenizer.vocab)
    inv_vocab = property(lambda self: self.tokenizer.inv_vocab)

    def tokenize(self, text):
        return self.tokenizer.convert_tokens_to_ids(self.tokenizer.tokenize(text))

    def decode(self, ids):
        return self.tokenizer.convert_tokens_to_string(self.tokenizer.convert_ids_to_tokens(ids))

    def decode_token_ids(self, token_ids):
        tokens = self.tokenizer.convert_ids_to_tokens(token_ids)
        out = ""
        for t in tokens:
            if t in ("[PAD]", "[CLS]"):
                continue
            out += t[2:] if t.startswith("##") else " " + t
        return out.strip()

    cls = property(lambda self: self.cls_id)
    sep = property(lambda self: self.sep_id)
    pad = property(lambda self: self.pad_id)
    mask = property(lambda self: self.mask_id)
    bos_token = property(lambda self: self._bos_token)
    eos_token = property(lambda self: self._eos_token)
    bos_token_id = property(lambda self: self._bos_token_id)
    eos_token_id = property(lambda self: self._eos_token_id)
    additional_special_tokens = property(lambda self: self._additional_special_tokens)
    additional_special_tokens_ids = property(
        lambda self: [self.vocab.get(t) for t in self._additional_special_tokens])


class _GPT2BPETokenizer(AbstractTokenizer):
    def __init__(self, vocab_file, merge_file):
        super().__init__("GPT2 BPE")
        from .gpt2_tokenization import GPT2Tokenizer
        self.tokenizer = GPT2Tokenizer(vocab_file, merge_file, errors="replace", special_tokens=[], max_len=None)
        self.eod_id = self.tokenizer.encoder["<|endoftext|>"]

    vocab_size = property(lambda self: len(self.tokenizer.encoder))
    vocab = property(lambda self: self.tokenizer.encoder)
    inv_vocab = property(lambda self: self.tokenizer.decoder)
    eod = property(lambda self: self.eod_id)

    def tokenize(self, text):
        return self.tokenizer.encode(text)

    def detokenize(self, token_ids):
        return self.tokenizer.decode(token_ids)


class _FalconTokenizer(AbstractTokenizer):
    """HF tokenizer for Falcon. The reference downloads 'tiiuae/falcon-40b'
    (tokenizer.py:294-295); offline, pass a local directory via --vocab_file."""

    def __init__(self, vocab_extra_ids_list=None, new_tokens=True, path=None):
        super().__init__("FalconTokenizer")
        from transformers import AutoTokenizer
        self.tokenizer = AutoTokenizer.from_pretrained(path or "tiiuae/falcon-40b")
        self._eod = self.tokenizer.vocab["<|endoftext|>"]
        if vocab_extra_ids_list and new_tokens:
            self.tokenizer.add_special_tokens({
                "additional_special_tokens": self.tokenizer.additional_special_tokens
                + vocab_extra_ids_list.split(",")})
        self._inv_vocab = {i: t for t, i in self.tokenizer.vocab.items()}

    vocab_size = property(lambda self: len(self.tokenizer.vocab))
    vocab = property(lambda self: self.tokenizer.vocab)
    inv_vocab = property(lambda self: self._inv_vocab)
    eod = property(lambda self: self._eod)

    def tokenize(self, text):
        return self.tokenizer.encode(text)

    def detokenize(self, token_ids):
        return self.tokenizer.decode(token_ids)


class _SentencePieceTokenizer(AbstractTokenizer):
    """SentencePiece model + Megatron special tokens (<CLS> <SEP> <EOD> <MASK>, pad/bos/eos
    and ``vocab_extra_ids``) appended after the model's own pieces unless new_tokens is off."""

    def __init__(self, model_file, vocab_extra_ids=0, vocab_extra_ids_list=None, new_tokens=True):
        super().__init__("SentencePieceTokenizer")
        import sentencepiece
        self._tokenizer = sentencepiece.SentencePieceProcessor(model_file=model_file)
        self._vocab, self._inv_vocab = {}, {}
        self._special_tokens, self._inv_special_tokens = {}, {}
        self._t5_tokens = []
        for i in range(len(self._tokenizer)):
            t = self._tokenizer.id_to_piece(i)
            self._inv_vocab[i] = t
            self._vocab[t] = i

        def add(t, force=False):
            if t not in self._vocab:
                if not new_tokens and not force:
                    return
                nid = len(self._vocab)
                self._vocab[t] = nid
                self._inv_vocab[nid] = t
            self._special_tokens[t] = self._vocab[t]
            self._inv_special_tokens[self._vocab[t]] = t

        for name, attr in (("<CLS>", "_cls_id"), ("<SEP>", "_sep_id"), ("<EOD>", "_eod_id"),
                           ("<MASK>", "_mask_id")):
            add(name)
            setattr(self, attr, self._vocab.get(name))
        pad_id = self._tokenizer.pad_id()
        try:
            pad_token = self._tokenizer.id_to_piece(pad_id)
        except IndexError:
            pad_token = "<PAD>"
        add(pad_token)
        self._pad_id = self._vocab.get(pad_token)
        for getter, default, attr in ((self._tokenizer.bos_id, "<BOS>", "_bos_id"),
                                      (self._tokenizer.eos_id, "<EOS>", "_eos_id")):
            try:
                tok = self._tokenizer.id_to_piece(getter())
            except IndexError:
                tok = default
            add(tok)
            setattr(self, attr, self._vocab.get(tok))
        for i in range(vocab_extra_ids):
            t = f"<extra_id_{i}>"
            add(t)
            self._t5_tokens.append(t)
        if vocab_extra_ids_list:
            for t in vocab_extra_ids_list.split(","):
                add(t)

    vocab_size = property(lambda self: len(self._vocab))
    vocab = property(lambda self: self._vocab)
    inv_vocab = property(lambda self: self._inv_vocab)

    def tokenize(self, text):
        """Split on the added special tokens, SentencePiece-encode the pieces in between."""
        ids, idx = [], 0
        while True:
            found = {}
            for tok in self._special_tokens:
                try:
                    found[tok] = text[idx:].index(tok)
                except ValueError:
                    continue
            if not found:
                break
            nxt = min(found, key=found.get)
            nidx = idx + found[nxt]
            ids.extend(self._tokenizer.encode_as_ids(text[idx:nidx]))
            ids.append(self._special_tokens[nxt])
            idx = nidx + len(nxt)
        ids.extend(self._tokenizer.encode_as_ids(text[idx:]))
        return ids

    def detokenize(self, ids):
        text, last = "", 0
        for i, tid in enumerate(ids):
            if tid in self._inv_special_tokens:
                text += self._tokenizer.decode_ids(ids[last:i]) + " "
                text += self._inv_special_tokens[tid] + " "
                last = i + 1
        text += self._tokenizer.decode_ids(ids[last:])
        return text.strip()

    cls = property(lambda self: self._cls_id)
    sep = property(lambda self: self._sep_id)
    pad = property(lambda self: self._pad_id)
    bos_token_id = property(lambda self: self._bos_id)
    bos = property(lambda self: self._bos_id)
    eod = property(lambda self: self._eod_id if self._eod_id is not None else self._eos_id)
    eos_token_id = property(lambda self: self._eos_id)
    eos = property(lambda self: self._eos_id)
    mask = property(lambda self: self._mask_id)
    additional_special_tokens_ids = property(lambda self: [self.vocab[k] for k in self._t5_tokens])
