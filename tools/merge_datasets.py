"""Concatenate every .bin/.idx pair of a directory into one indexed dataset
(tools/merge_datasets.py:11-66 surface)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), os.path.pardir)))

from epfl_megatron_b200.data import indexed_dataset


def main():
    p = argparse.ArgumentParser()
    g = p.add_argument_group(title="input data")
    g.add_argument("--input", type=str, required=True, help="directory holding the datasets to merge")
    g = p.add_argument_group(title="output data")
    g.add_argument("--output_prefix", type=str, required=True, help="path of the merged files, no suffix")
    args = p.parse_args()
    assert os.path.isdir(args.input), f"{args.input} is not a directory"

    prefixes = set()
    for name in os.listdir(args.input):
        stem, ext = os.path.splitext(name)
        if ext not in (".bin", ".idx"):
            continue
        other = ".idx" if ext == ".bin" else ".bin"
        assert os.path.isfile(os.path.join(args.input, stem + other)), \
            f"{name} has no matching {other} file in {args.input}"
        prefixes.add(stem)
    assert prefixes, "no .bin/.idx pairs found"

    builder = None
    for prefix in sorted(prefixes):
        path = os.path.join(args.input, prefix)
        if builder is None:
            ds = indexed_dataset.make_dataset(path, "infer")
            dtype = ds._index.dtype if isinstance(ds, indexed_dataset.MMapIndexedDataset) else ds.dtype
            builder = indexed_dataset.MMapIndexedDatasetBuilder(
                indexed_dataset.data_file_path(args.output_prefix), dtype=dtype)
            del ds
        builder.merge_file_(path)
    builder.finalize(indexed_dataset.index_file_path(args.output_prefix))
    print(f"merged {len(prefixes)} datasets into {args.output_prefix}.bin/.idx")


if __name__ == "__main__":
    main()
