"""Give every document of a loose-json file an id: <prefix>-<running number>."""
import argparse
import json

if __name__ == "__main__":
    p = argparse.ArgumentParser()
    p.add_argument("--input_file", required=True)
    p.add_argument("--output_file", required=True)
    p.add_argument("--id_prefix", required=True)
    p.add_argument("--log_interval", type=int, default=100)
    a = p.parse_args()
    with open(a.input_file, encoding="utf-8") as fin, open(a.output_file, "w", encoding="utf-8") as fout:
        for n, line in enumerate(fin, start=1):
            doc = json.loads(line)
            doc["adlr_id"] = f"{a.id_prefix}-{n:010d}"
            fout.write(json.dumps(doc, ensure_ascii=False) + "\n")
            if n % a.log_interval == 0:
                print(f"    processed {n:9d} documents", flush=True)
    print("done :-)")
