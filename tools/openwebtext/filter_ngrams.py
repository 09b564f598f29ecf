"""Decontaminate a training corpus against evaluation sets: every 13-gram (shorter for short
task texts) of a task document that occurs in a training document splits that document around
the match; fragments shorter than a threshold are dropped (and documents split too often).
usage: python filter_ngrams.py --tasks lambada squad ... --lambada_path ... --dedup_dataset in.json text --output out.json"""
import argparse
import json
import re
import time
from functools import partial
from multiprocessing import Pool


def get_words(text):
    words, positions = [], []
    for m in re.finditer(r"\w+", text.lower()):
        words.append(m.group(0))
        positions.append(m.start())
    return words, positions


def split_text(text, start, remove_char_each_side, seq):
    """Cut `text` around the match at `start`, trimming to sentence ends on both sides."""
    punct = ".!?"
    a = max(start - remove_char_each_side, 0)
    while a > 0 and text[a - 1] not in punct:
        a -= 1
    b = min(start + len(seq) + remove_char_each_side, len(text))
    while b < len(text) and text[b] not in punct:
        b += 1
    return text[:a], text[min(b + 1, len(text)):]


def task_ngrams(args):
    ngrams = {}

    def add(text):
        words, _ = get_words(text)
        n = args.max_ngram_size
        if len(words) < args.min_ngram_size:
            return
        if len(words) < n:
            ngrams[" ".join(words)] = 0
            return
        for i in range(len(words) - n + 1):
            ngrams[" ".join(words[i:i + n])] = 0
    for task in args.tasks:
        path = getattr(args, f"{task}_path", None)
        if task == "lambada" and path:
            with open(path, encoding="utf-8") as f:
                for line in f:
                    add(json.loads(line)["text"])
        elif task == "squad" and path:
            with open(path, encoding="utf-8") as f:
                for art in json.load(f)["data"]:
                    for para in art["paragraphs"]:
                        for qa in para["qas"]:
                            add(qa["question"])
        elif path:
            with open(path, encoding="utf-8") as f:
                for line in f:
                    try:
                        d = json.loads(line)
                        add(d.get("text") or d.get("question") or "")
                    except json.JSONDecodeError:
                        add(line)
    return ngrams


def clean_document(line, key, ngrams, args):
    doc = json.loads(line)
    pending, out, splits = [doc[key]], [], 0
    while pending:
        text = pending.pop()
        words, positions = get_words(text)
        hit = None
        n = args.max_ngram_size
        for i in range(max(len(words) - n + 1, 0)):
            seq = " ".join(words[i:i + n])
            if seq in ngrams:
                hit = (positions[i], seq)
                break
        if hit is None:
            if len(text) >= args.filter_text_char_len:
                out.append(text)
            continue
        first, second = split_text(text, hit[0], args.remove_char_each_side, hit[1])
        splits += 1
        for frag in (first, second):
            if len(frag) >= args.filter_text_char_len:
                pending.append(frag)
    if splits > args.splits_count:
        return []
    res = []
    for i, text in enumerate(out):
        d = dict(doc)
        d[key] = text
        if splits:
            d["split_id"] = i
        res.append(json.dumps(d, ensure_ascii=False))
    return res


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--tasks", nargs="*", required=True)
    for t in ("lambada", "squad", "natural_questions", "triviaqa", "webqa", "race", "piqa", "coqa"):
        p.add_argument(f"--{t}_path", type=str, default=None)
    p.add_argument("--dedup_dataset", nargs="*", default=None, help="<file> <text key>")
    p.add_argument("--output", type=str, required=True)
    p.add_argument("--num_threads", type=int, default=8)
    p.add_argument("--max_ngram_size", type=int, default=13)
    p.add_argument("--min_ngram_size", type=int, default=8)
    p.add_argument("--filter_text_char_len", type=int, default=200)
    p.add_argument("--key_threshold", type=int, default=10)
    p.add_argument("--remove_char_each_side", type=int, default=200)
    p.add_argument("--splits_count", type=int, default=10)
    a = p.parse_args()
    t0 = time.time()
    ngrams = task_ngrams(a)
    print(f"{len(ngrams)} task n-grams")
    fname, key = a.dedup_dataset[0], a.dedup_dataset[1]
    kept = 0
    with open(fname, encoding="utf-8") as fin, open(a.output, "w", encoding="utf-8") as fout, \
            Pool(a.num_threads) as pool:
        for docs in pool.imap(partial(clean_document, key=key, ngrams=ngrams, args=a), fin, 500):
            for d in docs:
                fout.write(d + "\n")
                kept += 1
    print(f"wrote {kept} documents in {time.time() - t0:.1f}s")


if __name__ == "__main__":
    main()
