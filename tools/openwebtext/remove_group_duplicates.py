"""Keep the first document of every duplicate group, drop the rest.
usage: python remove_group_duplicates.py <groups.json> <data.json> <output.json>"""
import json
import sys
import time

if __name__ == "__main__":
    if len(sys.argv) < 4 or sys.argv[1] in ("-h", "--help"):
        sys.exit("usage: python tools/openwebtext/remove_group_duplicates.py <url groups jsonl> <data jsonl> <output jsonl>")
    groups, data, out = sys.argv[1], sys.argv[2], sys.argv[3]
    t0 = time.time()
    drop = set()
    with open(groups, encoding="utf-8") as f:
        for line in f:
            for members in json.loads(line).values():
                drop.update(members[1:])
    kept = removed = 0
    with open(data, encoding="utf-8") as fin, open(out, "w", encoding="utf-8") as fout:
        for line in fin:
            try:
                doc = json.loads(line)
                if doc.get("url") in drop:
                    removed += 1
                    continue
                fout.write(json.dumps(doc, ensure_ascii=False) + "\n")
                kept += 1
            except Exception as e:  # noqa: BLE001
                print("[SKIPPING]", line[:80], e)
    print(f"kept {kept}, removed {removed} in {time.time() - t0:.1f}s")
