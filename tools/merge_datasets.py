"""Concatenate several indexed datasets (`<prefix>.bin` / `<prefix>.idx`) into one - the role of Megatron-LM's tools/merge_datasets.py.

    python tools/merge_datasets.py --input shard0_text_document shard1_text_document --output-prefix merged_text_document

Inputs may be in either on-disk format (mmap or legacy lazy / cached, detected from the index file); the output format is
`--dataset-impl` (default mmap) with the token dtype of the first input.  Document boundaries are preserved.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def merge(inputs, output_prefix, impl="mmap"):
    from b200dist.data import indexed_dataset as idx
    sets = [idx.make_dataset(p, "infer") for p in inputs]
    missing = [p for p, d in zip(inputs, sets) if d is None]
    if missing:
        raise FileNotFoundError(f"no indexed dataset at {missing}")
    dtype = sets[0][0].dtype.type if len(sets[0]) else np.int32
    b = idx.make_builder(idx.data_file_path(output_prefix), impl, dtype=dtype)
    n_items = n_docs = 0
    for ds in sets:
        bounds = [int(x) for x in ds.doc_idx]
        for d in range(len(bounds) - 1):
            for i in range(bounds[d], bounds[d + 1]):
                b.add_item(ds[i])
                n_items += 1
            b.end_document()
            n_docs += 1
    b.finalize(idx.index_file_path(output_prefix))
    return n_items, n_docs


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--input", nargs="+", required=True, help="dataset prefixes (without .bin / .idx)")
    ap.add_argument("--output-prefix", required=True)
    ap.add_argument("--dataset-impl", default="mmap", choices=["mmap", "lazy", "cached"])
    a = ap.parse_args(argv)
    n_items, n_docs = merge(a.input, a.output_prefix, a.dataset_impl)
    print(f"merged {len(a.input)} datasets: {n_items} sequences in {n_docs} documents -> {a.output_prefix}.bin/.idx")
    return 0


if __name__ == "__main__":
    sys.exit(main())
