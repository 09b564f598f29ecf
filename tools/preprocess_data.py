"""jsonl -> .bin/.idx indexed dataset (tools/preprocess_data.py:42-197 surface).

    python tools/preprocess_data.py --input corpus.jsonl --output_prefix out/corpus \
        --tokenizer_type SentencePieceTokenizer --vocab_file tokenizer.model --workers 8 --append_eod

Each worker process tokenises documents (optionally sentence-split with NLTK) and the parent
appends them to one builder per json key. Output names: <prefix>_<key>_<document|sentence>.bin/.idx
"""
import argparse
import json
import multiprocessing
import os
import sys
import time

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), os.path.pardir)))

import numpy as np

from epfl_megatron_b200.data import indexed_dataset
from epfl_megatron_b200.tokenizer import build_tokenizer

try:
    import nltk
    nltk_available = True
except ImportError:
    nltk_available = False


class IdentitySplitter:
    def tokenize(self, *text):
        return text


class Encoder:
    """Holds the per-process tokenizer / sentence splitter (created once in each pool worker)."""
    tokenizer = None
    splitter = None

    def __init__(self, args):
        self.args = args

    def initializer(self):
        Encoder.tokenizer = build_tokenizer(self.args)
        if self.args.split_sentences:
            if not nltk_available:
                raise SystemExit("NLTK is not available to split sentences.")
            lib = f"tokenizers/punkt/{self.args.lang}.pickle"
            Encoder.splitter = nltk.load(lib)
        else:
            Encoder.splitter = IdentitySplitter()

    def encode(self, line):
        data = json.loads(line)
        out = {}
        for key in self.args.json_keys:
            text = data[key]
            doc = []
            for sentence in Encoder.splitter.tokenize(text):
                ids = Encoder.tokenizer.tokenize(sentence)
                if ids:
                    doc.append(ids)
            if doc and self.args.append_eod:
                doc[-1].append(Encoder.tokenizer.eod)
            out[key] = doc
        return out, len(line)


def get_args(argv=None):
    p = argparse.ArgumentParser()
    g = p.add_argument_group(title="input data")
    g.add_argument("--input", type=str, nargs="+", required=True, help="jsonl file(s)")
    g.add_argument("--json_keys", nargs="+", default=["text"], help="keys to extract from each json line")
    g.add_argument("--split_sentences", action="store_true", help="one item per sentence (NLTK punkt)")
    g.add_argument("--lang", type=str, default="english", help="NLTK punkt language")
    g = p.add_argument_group(title="tokenizer")
    g.add_argument("--tokenizer_type", type=str, required=True,
                   choices=["BertWordPieceLowerCase", "BertWordPieceCase", "GPT2BPETokenizer",
                            "SentencePieceTokenizer", "FalconTokenizer", "NullTokenizer"])
    g.add_argument("--vocab_file", type=str, default=None)
    g.add_argument("--merge_file", type=str, default=None)
    g.add_argument("--append_eod", action="store_true", help="append <eod> to every document")
    g.add_argument("--vocab_extra_ids", type=int, default=0)
    g.add_argument("--vocab_extra_ids_list", type=str, default=None,
                   help="comma separated extra tokens for the SentencePiece tokenizer")
    g.add_argument("--no_new_tokens", action="store_false", dest="new_tokens",
                   help="do not add the Megatron special tokens to a SentencePiece vocabulary")
    g.add_argument("--vocab_size", type=int, default=None, help="NullTokenizer only")
    g = p.add_argument_group(title="output data")
    g.add_argument("--output_prefix", type=str, required=True)
    g.add_argument("--dataset_impl", type=str, default="mmap", choices=["lazy", "cached", "mmap"])
    g = p.add_argument_group(title="runtime")
    g.add_argument("--workers", type=int, required=True)
    g.add_argument("--chunk_size", type=int, required=True)
    g.add_argument("--log_interval", type=int, default=100)
    args = p.parse_args(argv)
    args.keep_empty = False
    if args.tokenizer_type.lower().startswith("bert") and not args.split_sentences:
        print("Bert tokenizer detected, are you sure you don't want to split sentences?")
    # what build_tokenizer expects beyond the flags above
    args.rank = 0
    args.make_vocab_size_divisible_by = 128
    args.tensor_model_parallel_size = 1
    return args


def main(argv=None):
    args = get_args(argv)
    t0 = time.time()
    print("Opening", args.input)
    tokenizer = build_tokenizer(args)
    encoder = Encoder(args)
    level = "sentence" if args.split_sentences else "document"
    print(f"Vocab size: {tokenizer.vocab_size}\nOutput prefix: {args.output_prefix}")
    builders, bins, idxs = {}, {}, {}
    for key in args.json_keys:
        bins[key] = f"{args.output_prefix}_{key}_{level}.bin"
        idxs[key] = f"{args.output_prefix}_{key}_{level}.idx"
        builders[key] = indexed_dataset.make_builder(bins[key], impl=args.dataset_impl,
                                                     vocab_size=tokenizer.vocab_size)

    def lines():
        for path in args.input:
            with open(path, "r", encoding="utf-8") as f:
                for line in f:
                    if line.strip():
                        yield line

    if args.workers > 1:
        pool = multiprocessing.Pool(args.workers, initializer=encoder.initializer)
        docs = pool.imap(encoder.encode, lines(), args.chunk_size)
    else:
        pool = None
        encoder.initializer()
        docs = map(encoder.encode, lines())
    proc_start = time.time()
    total_bytes = 0
    print("Time to startup:", proc_start - t0)
    i = 0
    for i, (doc, nbytes) in enumerate(docs, start=1):
        total_bytes += nbytes
        for key, sentences in doc.items():
            if not sentences:
                continue
            for s in sentences:
                builders[key].add_item(np.asarray(s, dtype=np.int64))
            builders[key].end_document()
        if i % args.log_interval == 0:
            el = time.time() - proc_start
            print(f"Processed {i} documents ({i / el:.1f} docs/s, {total_bytes / el / 1e6:.2f} MB/s).",
                  file=sys.stderr)
    if pool is not None:
        pool.close()
        pool.join()
    print("Done! Now finalizing.")
    for key in args.json_keys:
        builders[key].finalize(idxs[key])
    print(f"wrote {i} documents in {time.time() - t0:.1f} s")


if __name__ == "__main__":
    main()
