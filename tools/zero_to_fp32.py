#!/usr/bin/env python
"""Consolidate a ZeRO-3 checkpoint (the `global_stepN/zero_pp_rank_*` shards of every rank) into one fp32 state_dict - the
role of DeepSpeed's `zero_to_fp32.py`, a copy of which DeepSpeed drops into every checkpoint directory.  This
script is copied next to the shards by `ckpt.deepspeed.save_zero3_checkpoint` and needs only torch:

    python zero_to_fp32.py <checkpoint_dir> <output_file> [--tag global_step100] [--strip-unit-prefix]

`--strip-unit-prefix` rewrites the `<unit>.<param>` keys (`embed.embed_tokens.weight`, `head.lm_head.weight`) to the model's own
attribute paths (`embed_tokens.weight`, `lm_head.weight`), i.e. a `model.load_state_dict()`-able file; Hugging Face tensor names are
produced by `b200dist.ckpt.hf.save_pretrained(model, ...)` after loading it.
"""
import argparse
import os
import sys

import torch


def consolidate(ckpt_dir, tag=None):
    if tag is None:
        with open(os.path.join(ckpt_dir, "latest")) as f:
            tag = f.read().strip()
    d = os.path.join(ckpt_dir, tag)
    files = sorted(f for f in os.listdir(d) if f.endswith("_optim_states.pt"))
    world = len(files)
    shards = [None] * world
    for f in files:
        o = torch.load(os.path.join(d, f), map_location="cpu", weights_only=False)["optimizer_state_dict"]
        shards[o["rank"]] = o
    assert all(s is not None for s in shards) and shards[0]["partition_count"] == world, "missing rank shards"
    mfile = sorted(f for f in os.listdir(d) if f.endswith("_model_states.pt"))[0]
    ms = torch.load(os.path.join(d, mfile), map_location="cpu", weights_only=False)
    out = {}
    for u, uname in enumerate(ms["unit_names"]):
        flat = torch.cat([shards[r]["fp32_flat_groups"][u] for r in range(world)])        # rank order == slice order
        for pname, shape in ms["param_shapes"][u].items():
            off, n = ms["param_offsets"][u][pname], 1
            for s_ in shape:
                n *= s_
            out[f"{uname}.{pname}"] = flat[off:off + n].view(shape).clone()
    for n, t in ms["module"].items():
        out[n] = t.float()
    return out


def strip_unit_prefix(sd):
    """unit-prefixed keys (`embed.`, `layers.N.`, `head.`) -> the model's own parameter names"""
    out = {}
    for k, v in sd.items():
        if k.startswith("embed."):
            k = k[len("embed."):]
        elif k.startswith("head."):
            k = k[len("head."):]
        out[k] = v
    return out


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("checkpoint_dir")
    ap.add_argument("output_file")
    ap.add_argument("--tag", default=None)
    ap.add_argument("--strip-unit-prefix", action="store_true")
    a = ap.parse_args(argv)
    sd = consolidate(a.checkpoint_dir, a.tag)
    if a.strip_unit_prefix:
        sd = strip_unit_prefix(sd)
    torch.save(sd, a.output_file)
    print(f"saved {len(sd)} tensors, {sum(v.numel() for v in sd.values())} parameters to {a.output_file}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
