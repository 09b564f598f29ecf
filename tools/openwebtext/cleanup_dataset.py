"""Clean a loose-json corpus: fix unicode (ftfy), keep English (langdetect), drop documents with
fewer than MIN_DOCUMENT_LENGTH tokens. usage: python cleanup_dataset.py <input> <output>"""
import json
import sys
import time

MIN_DOCUMENT_LENGTH = 128


def _tools():
    try:
        import ftfy
        fix = ftfy.fix_text
    except ImportError:
        def fix(t):
            return t
    try:
        from langdetect import detect
    except ImportError:
        def detect(_):
            return "en"
    return fix, detect


def filter_corpus(filename, out_filename, print_interval=10000, tokenize=None):
    fix, detect = _tools()
    tokenize = tokenize or (lambda t: t.split())
    n = fixed = non_en = small = written = 0
    t0 = time.time()
    with open(out_filename, "wb") as out, open(filename, encoding="utf-8") as fin:
        for line in fin:
            try:
                n += 1
                doc = json.loads(line)
                text = fix(doc["text"])
                if text != doc["text"]:
                    fixed += 1
                doc["text"] = text
                if detect(text) != "en":
                    non_en += 1
                    continue
                if len(text.split()) < MIN_DOCUMENT_LENGTH or len(tokenize(text)) < MIN_DOCUMENT_LENGTH:
                    small += 1
                    continue
                out.write((json.dumps(doc, ensure_ascii=False) + "\n").encode("utf-8"))
                written += 1
                if n % print_interval == 0:
                    print(f"[PROGRESS] {time.time() - t0:.1f}s | docs {n} | fixed {fixed} | non-en {non_en} | "
                          f"small {small} | written {written}", flush=True)
            except Exception as e:  # noqa: BLE001
                print("    skipping", line[:80], e)
    print(f"[FINAL] docs {n} | fixed {fixed} | non-english {non_en} | small {small} | written {written}")


if __name__ == "__main__":
    if len(sys.argv) < 3 or sys.argv[1] in ("-h", "--help"):
        sys.exit("usage: python tools/openwebtext/cleanup_dataset.py <input jsonl> <output jsonl>")
    filter_corpus(sys.argv[1], sys.argv[2])
