"""jsonl -> Megatron `.bin/.idx` (the data-prep step of the reference's GPT notebook: CodeParrot json lines are run
through Megatron's tools/preprocess_data.py with `--tokenizer-type GPT2BPETokenizer --append-eod`, producing the
`<prefix>_text_document.{bin,idx}` pair that `--data-path` points at).

    python tools/preprocess_data.py --input data.jsonl --output-prefix codeparrot --json-keys content \
        --tokenizer-type GPT2BPETokenizer --vocab-file vocab.json --merge-file merges.txt --append-eod --workers 8
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

_TOK = None


def _build_tok(a):
    from b200dist.data import tokenizers
    if a.tokenizer_type == "NullTokenizer":
        from b200dist.engine.megatron import _NullTokenizer
        return _NullTokenizer(a.vocab_size)
    return tokenizers.build(a)


def _init(a):
    global _TOK
    _TOK = (_build_tok(a), a)


def _encode(line):
    tok, a = _TOK
    line = line.strip()
    if not line:
        return {}, 0
    rec = json.loads(line)
    out = {}
    for key in a.json_keys:
        ids = tok.tokenize(rec[key])
        if a.append_eod and ids:
            ids.append(tok.eod)
        out[key] = ids
    return out, len(line)


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--input", required=True)
    p.add_argument("--json-keys", nargs="+", default=["text"])
    p.add_argument("--output-prefix", required=True)
    p.add_argument("--tokenizer-type", default="GPT2BPETokenizer")
    p.add_argument("--vocab-file"); p.add_argument("--merge-file"); p.add_argument("--tokenizer-model")
    p.add_argument("--vocab-size", type=int, default=32767)
    p.add_argument("--append-eod", action="store_true")
    p.add_argument("--workers", type=int, default=1)
    p.add_argument("--log-interval", type=int, default=10000)
    p.add_argument("--dataset-impl", default="mmap", choices=["mmap", "lazy", "cached"])
    a = p.parse_args(argv)
    from b200dist.data.indexed_dataset import data_file_path, index_file_path, make_builder
    tok = _build_tok(a)
    dtype = np.uint16 if tok.vocab_size < 65500 else np.int32
    builders = {k: make_builder(data_file_path(f"{a.output_prefix}_{k}_document"), a.dataset_impl, dtype=dtype) for k in a.json_keys}
    t0, nbytes, ndocs = time.time(), 0, 0
    with open(a.input, encoding="utf-8") as fin:
        if a.workers > 1:
            pool = mp.Pool(a.workers, initializer=_init, initargs=(a,))
            it = pool.imap(_encode, fin, 32)
        else:
            _init(a)
            it = map(_encode, fin)
        for doc, n in it:
            nbytes += n
            for k, ids in doc.items():
                if ids:
                    builders[k].add_item(np.asarray(ids, dtype=dtype))
                    builders[k].end_document()
            ndocs += 1
            if ndocs % a.log_interval == 0:
                dt = time.time() - t0
                print(f"Processed {ndocs} documents ({ndocs / dt:.1f} docs/s, {nbytes / dt / 2**20:.2f} MB/s).", file=sys.stderr)
    for k, b in builders.items():
        b.finalize(index_file_path(f"{a.output_prefix}_{k}_document"))
    print(f"wrote {ndocs} documents to {a.output_prefix}_<key>_document.{{bin,idx}}")


if __name__ == "__main__":
    main()
