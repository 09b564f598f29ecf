"""Selective clean-ups of a loose-json corpus (tasks: remove_512, remove_256_javascript,
remove_512_non_english, ftfy_fix_text, general_cleaning)."""
import argparse
import json
import re
import time
from functools import partial
from multiprocessing import Pool


def process_doc(json_line, args):
    doc = json.loads(json_line)
    text = doc["text"]
    flags = {}
    n_chars = len(text)
    if "remove_512" in args.tasks and n_chars < 512:
        return {}, text, {"remove_512": True}
    if "remove_256_javascript" in args.tasks and n_chars < 256 and "javascript" in text.lower():
        return {}, text, {"remove_256_javascript": True}
    if "remove_512_non_english" in args.tasks and n_chars < 512:
        try:
            from langdetect import detect
            if detect(text) != "en":
                return {}, text, {"remove_512_non_english": True}
        except Exception:  # noqa: BLE001
            pass
    if "ftfy_fix_text" in args.tasks:
        try:
            import ftfy
            fixed = ftfy.fix_text(text)
            flags["ftfy_fix_text"] = fixed != text
            text = fixed
        except ImportError:
            pass
    if "general_cleaning" in args.tasks:
        cleaned = re.sub(r"  +|\b\n+ |\b\n+", " ", text)
        cleaned = re.sub(r"\n\n+", "\n\n", cleaned)
        flags["general_cleaning"] = cleaned != text
        text = cleaned
    doc["text"] = text
    return doc, text, flags


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--input_files", nargs="*", required=True)
    p.add_argument("--tasks", nargs="*", required=True)
    p.add_argument("--output_path", type=str, default=None)
    p.add_argument("--log_interval", type=int, default=100)
    p.add_argument("--workers", type=int, default=4)
    a = p.parse_args()
    t0 = time.time()
    stats = {}
    for fname in a.input_files:
        out_name = (a.output_path or ".") + "/" + fname.split("/")[-1] + ".clean"
        with open(fname, encoding="utf-8") as fin, open(out_name, "w", encoding="utf-8") as fout, \
                Pool(a.workers) as pool:
            for n, (doc, _, flags) in enumerate(pool.imap(partial(process_doc, args=a), fin, 500), start=1):
                for k, v in flags.items():
                    stats[k] = stats.get(k, 0) + int(bool(v))
                if doc:
                    fout.write(json.dumps(doc, ensure_ascii=False) + "\n")
                if n % a.log_interval == 0:
                    print(f"processed {n} docs in {time.time() - t0:.1f}s {stats}", flush=True)
    print("done", stats)


if __name__ == "__main__":
    main()
