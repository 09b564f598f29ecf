"""Near-duplicate candidates with MinHash over character 5-gram shingles, banded LSH
(own implementation; the `lsh` package is not required).
usage: python find_duplicates.py --inputs corpus.json url [more.json key ...] --output dups.json"""
import argparse
import hashlib
import itertools
import json
import struct
import time

import numpy as np


def shingles(text, char_ngram=5):
    return {text[i:i + char_ngram] for i in range(0, max(len(text) - char_ngram, 1))}


def jaccard(a, b, mode="union"):
    inter = len(a & b)
    if mode == "min":
        return inter / max(min(len(a), len(b)), 1)
    if mode == "max":
        return inter / max(max(len(a), len(b)), 1)
    return inter / max(len(a | b), 1)


class MinHasher:
    def __init__(self, seeds=100, char_ngram=5, seed=1234):
        rng = np.random.RandomState(seed)
        self.a = rng.randint(1, 2 ** 31 - 1, size=seeds, dtype=np.int64)
        self.b = rng.randint(0, 2 ** 31 - 1, size=seeds, dtype=np.int64)
        self.char_ngram = char_ngram

    def fingerprint(self, text):
        hs = np.array([struct.unpack("<I", hashlib.md5(s.encode("utf-8", "ignore")).digest()[:4])[0]
                       for s in shingles(text, self.char_ngram)] or [0], dtype=np.int64)
        return ((self.a[:, None] * hs[None, :] + self.b[:, None]) % ((1 << 31) - 1)).min(axis=1)


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--seed", type=int, default=1234)
    p.add_argument("--inputs", nargs="*", default=None, help="pairs: <file> <key holding the id/url>")
    p.add_argument("--output", type=str, required=True)
    p.add_argument("--jaccard", type=str, default="union", choices=["union", "min", "max"])
    p.add_argument("--heuristic_iter", type=int, default=1)
    p.add_argument("--num_bands", type=int, default=10)
    p.add_argument("--num_seeds", type=int, default=100)
    p.add_argument("--jaccard_threshold", type=float, default=0.5)
    a = p.parse_args()
    t0 = time.time()
    hasher = MinHasher(a.num_seeds, seed=a.seed)
    rows = a.num_seeds // a.num_bands
    buckets = [dict() for _ in range(a.num_bands)]
    texts = {}
    for fname, key in zip(a.inputs[::2], a.inputs[1::2]):
        with open(fname, encoding="utf-8") as f:
            for line in f:
                doc = json.loads(line)
                uid, text = doc[key], doc["text"]
                texts[uid] = text
                fp = hasher.fingerprint(text)
                for band in range(a.num_bands):
                    sig = fp[band * rows:(band + 1) * rows].tobytes()
                    buckets[band].setdefault(sig, []).append(uid)
    seen, n_out = set(), 0
    with open(a.output, "w", encoding="utf-8") as out:
        for band in buckets:
            for ids in band.values():
                if len(ids) < 2:
                    continue
                main_id, rest = ids[0], ids[1:]
                if main_id in seen:
                    continue
                sa = shingles(texts[main_id])
                sims = []
                for other in rest:
                    if other in seen or other == main_id:
                        continue
                    j = jaccard(sa, shingles(texts[other]), a.jaccard)
                    if j > a.jaccard_threshold:
                        sims.append({other: j})
                        seen.add(other)
                if sims:
                    out.write(json.dumps({main_id: sims}, ensure_ascii=False) + "\n")
                    n_out += 1
    print(f"wrote {n_out} duplicate groups in {time.time() - t0:.1f}s")


if __name__ == "__main__":
    main()
