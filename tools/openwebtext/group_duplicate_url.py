"""Connected components of the similarity pairs written by find_duplicates.py.
usage: python group_duplicate_url.py <dups.json> <groups.json> [threshold]"""
import json
import sys
import time

if __name__ == "__main__":
    t0 = time.time()
    if len(sys.argv) < 3 or sys.argv[1] in ("-h", "--help"):
        sys.exit("usage: python tools/openwebtext/group_duplicate_url.py <similar-url pairs file> <output groups jsonl> [jaccard threshold = 0.7]")
    inp, out = sys.argv[1], sys.argv[2]
    thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.7
    parent = {}

    def find(x):
        parent.setdefault(x, x)
        while parent[x] != x:
            parent[x] = parent[parent[x]]
            x = parent[x]
        return x

    with open(inp, encoding="utf-8") as f:
        for line in f:
            for main, others in json.loads(line).items():
                for o in others:
                    for other, sim in o.items():
                        if sim >= thr:
                            ra, rb = find(main), find(other)
                            if ra != rb:
                                parent[rb] = ra
    groups = {}
    for x in parent:
        groups.setdefault(find(x), []).append(x)
    with open(out, "w", encoding="utf-8") as f:
        for i, members in enumerate(g for g in groups.values() if len(g) > 1):
            f.write(json.dumps({str(i): members}, ensure_ascii=False) + "\n")
    print(f"{sum(len(g) > 1 for g in groups.values())} groups in {time.time() - t0:.1f}s")
