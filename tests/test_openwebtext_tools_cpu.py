"""The corpus-cleaning tool chain of tools/openwebtext (reference: tools/openwebtext/*.py, README):
URL filtering -> merge -> clean -> near-duplicate detection (MinHash / LSH) -> grouping -> removal ->
ids, run through the command lines on a small synthetic crawl."""
import json
import os
import random
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OWT = os.path.join(ROOT, "tools", "openwebtext")


def _run(script, *args):
    r = subprocess.run([sys.executable, os.path.join(OWT, script), *map(str, args)], capture_output=True, text=True,
                       cwd=ROOT, timeout=300)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    return r.stdout


def _text(rng, n_words):
    words = ["alpha", "beta", "gamma", "delta", "epsilon", "zeta", "theta", "lambda", "sigma", "omega", "kernel",
             "tensor", "memory", "stream", "pipeline", "gradient", "weight", "token", "layer", "attention"]
    return " ".join(rng.choice(words) for _ in range(n_words))


def test_url_blacklist(tmp_path):
    urls = tmp_path / "urls"
    urls.mkdir()
    (urls / "a.txt").write_text("\n".join([
        "https://example.org/article/1", "https://example.org/article/1",          # duplicate
        "https://www.youtube.com/watch?v=1", "https://imgur.com/x",                # blacklisted domains
        "https://example.org/file.pdf", "https://example.org/movie.mp4?dl=1",      # extensions
        "http://a.b", "not a url at all", "https://news.example.com/story?id=7"]) + "\n")
    out = tmp_path / "clean.txt"
    log = _run("blacklist_urls.py", urls, out)
    kept = set(out.read_text().split())
    assert kept == {"https://example.org/article/1", "https://news.example.com/story?id=7"}, (kept, log)


def test_merge_clean_dedup_chain(tmp_path):
    rng = random.Random(0)
    parts = tmp_path / "parts"
    parts.mkdir()
    base = _text(rng, 300)
    docs = [{"url": "u0", "text": base},
            {"url": "u1", "text": base + " extra words appended at the very end"},       # near duplicate of u0
            {"url": "u2", "text": _text(rng, 300)},
            {"url": "u3", "text": _text(rng, 20)},                                         # too short
            {"url": "u4", "text": _text(rng, 260)}]
    (parts / "p0.json").write_text("\n".join(json.dumps(d) for d in docs[:3]) + "\n")
    (parts / "p1.json").write_text("\n".join(json.dumps(d) for d in docs[3:]) + "\n")
    merged = tmp_path / "merged.json"
    assert "merged 5 documents" in _run("merge_jsons.py", "--json_path", parts, "--output_file", merged)
    cleaned = tmp_path / "cleaned.json"
    log = _run("cleanup_dataset.py", merged, cleaned)
    kept = [json.loads(l)["url"] for l in open(cleaned)]
    assert kept == ["u0", "u1", "u2", "u4"] and "small 1" in log
    pairs = tmp_path / "pairs.json"
    _run("find_duplicates.py", "--inputs", cleaned, "url", "--output", pairs, "--jaccard_threshold", "0.7")
    found = [json.loads(l) for l in open(pairs)]
    assert len(found) == 1
    (main, sims), = found[0].items()
    assert {main, *[k for s in sims for k in s]} == {"u0", "u1"} and all(v > 0.7 for s in sims for v in s.values())
    groups = tmp_path / "groups.json"
    _run("group_duplicate_url.py", pairs, groups, "0.7")
    members = [m for l in open(groups) for m in json.loads(l).values()]
    assert len(members) == 1 and sorted(members[0]) == ["u0", "u1"]
    deduped = tmp_path / "dedup.json"
    assert "removed 1" in _run("remove_group_duplicates.py", groups, cleaned, deduped)
    left = [json.loads(l)["url"] for l in open(deduped)]
    assert len(left) == 3 and "u2" in left and "u4" in left and len({"u0", "u1"} & set(left)) == 1
    with_ids = tmp_path / "ids.json"
    _run("add_id.py", "--input_file", deduped, "--output_file", with_ids, "--id_prefix", "owt")
    ids = [json.loads(l)["adlr_id"] for l in open(with_ids)]
    assert ids == ["owt-0000000001", "owt-0000000002", "owt-0000000003"]


def test_minhash_estimates_jaccard():
    sys.path.insert(0, OWT)
    try:
        import find_duplicates as fd
    finally:
        sys.path.pop(0)
    rng = random.Random(1)
    a = _text(rng, 400)
    b = a[:len(a) * 3 // 4] + _text(rng, 100)
    true_j = fd.jaccard(fd.shingles(a), fd.shingles(b))
    h = fd.MinHasher(seeds=400, seed=5)
    est = float((h.fingerprint(a) == h.fingerprint(b)).mean())
    assert abs(est - true_j) < 0.1, (est, true_j)
    assert float((h.fingerprint(a) == h.fingerprint(a)).mean()) == 1.0
    assert fd.jaccard({1, 2, 3}, {2, 3, 4}) == 0.5 and fd.jaccard({1, 2}, {2, 3, 4}, "min") == 0.5


def test_ngram_decontamination(tmp_path):
    rng = random.Random(2)
    leak = "the quick brown fox jumps over the lazy dog while thirteen distinct words appear here in order"
    task = tmp_path / "lambada.json"
    task.write_text(json.dumps({"text": leak}) + "\n")

    def sentences(n):
        return " ".join(_text(rng, 12).capitalize() + "." for _ in range(n))
    before, after = sentences(30), sentences(30)
    corpus = tmp_path / "train.json"
    corpus.write_text("\n".join(json.dumps(d) for d in [
        {"id": "clean", "text": sentences(25)},
        {"id": "dirty", "text": before + " " + leak.capitalize() + ". " + after},
        {"id": "tiny", "text": "Short text."}]) + "\n")
    out = tmp_path / "decontaminated.json"
    log = _run("filter_ngrams.py", "--tasks", "lambada", "--lambada_path", task, "--dedup_dataset", corpus, "text",
               "--output", out, "--num_threads", "2", "--remove_char_each_side", "50")
    docs = [json.loads(l) for l in open(out)]
    by_id = {}
    for d in docs:
        by_id.setdefault(d["id"], []).append(d)
    assert "task n-grams" in log and "tiny" not in by_id                      # below the length floor
    assert len(by_id["clean"]) == 1 and "split_id" not in by_id["clean"][0]
    frags = by_id["dirty"]
    assert len(frags) == 2 and sorted(f["split_id"] for f in frags) == [0, 1]
    for f in frags:                                                           # the leak and its margins are gone
        assert "quick brown fox" not in f["text"].lower() and len(f["text"]) >= 200
        assert f["text"].strip()[0].isupper()                                 # cut at sentence boundaries
