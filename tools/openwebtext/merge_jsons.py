"""Concatenate every *.json file of a directory into one loose-json file."""
import argparse
import glob
import json

if __name__ == "__main__":
    p = argparse.ArgumentParser()
    p.add_argument("--json_path", default=".", help="directory holding the json files")
    p.add_argument("--output_file", default="merged_output.json")
    a = p.parse_args()
    n = 0
    with open(a.output_file, "w", encoding="utf-8") as out:
        for fname in sorted(glob.glob(a.json_path + "/*.json")):
            with open(fname, encoding="utf-8") as f:
                for line in f:
                    out.write(json.dumps(json.loads(line), ensure_ascii=False) + "\n")
                    n += 1
    print(f"merged {n} documents into {a.output_file}")
