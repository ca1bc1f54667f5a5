"""Re-shard a Megatron-layout GPT checkpoint to another tensor- / pipeline-parallel size (the role of Megatron-LM's
tools/checkpoint_util.py; layout: SURVEY 5.4 / E24).

    python tools/checkpoint_util.py --load-dir ckpt_tp2 --save-dir ckpt_tp4pp2 --target-tensor-parallel-size 4 \
        --target-pipeline-parallel-size 2 [--iteration N]

Reads `<load-dir>/iter_XXXXXXX/mp_rank_TT[_PPP]/model_optim_rng.pt` of every model-parallel rank, merges the shards into
the full model (TP = PP = 1), splits it for the target layout and writes the same directory structure plus
`latest_checkpointed_iteration.txt`.  Model weights only: the optimizer state is not carried over, so continue with
`--load <save-dir> --finetune` or `--no-load-optim --no-load-rng` (the upstream tool has the same contract).  Checkpoints
written with virtual pipeline stages (`model0..k`) are merged too; the output always has one chunk per stage.

How each tensor is sharded (parallel/layers.py, models/gpt.py, models/gpt_flex.py):
  dim 0: word_embeddings, output_layer, query_key_value (a rank owns whole heads / query groups), dense_h_to_4h
         (SwiGLU: every rank stores [its gate rows; its up rows]), their biases
  dim 1: self_attention.dense.weight, dense_4h_to_h.weight
  replicated: LayerNorm parameters, row-parallel biases, position embeddings, the Switch-MLP router
"""
from __future__ import annotations

import argparse
import copy
import os
import re
import sys
from collections import OrderedDict
from typing import Dict, List, Optional, Tuple

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

_LAYER = re.compile(r"^layers\.(\d+)\.(.*)$")


def partition_dim(key: str) -> Optional[int]:
    """key: parameter name inside a section (e.g. `layers.3.mlp.dense_h_to_4h.weight`, `word_embeddings.weight`)"""
    if key.endswith(("query_key_value.weight", "query_key_value.bias", "dense_h_to_4h.weight", "dense_h_to_4h.bias")):
        return 0
    if key.endswith(("self_attention.dense.weight", "dense_4h_to_h.weight")):
        return 1
    if key.startswith(("word_embeddings.", "output_layer.")):
        return 0
    return None


def _is_swiglu_packed(key: str, swiglu: bool) -> bool:
    return swiglu and key.endswith(("dense_h_to_4h.weight", "dense_h_to_4h.bias"))


def merge_tp(shards: List[torch.Tensor], key: str, swiglu: bool) -> torch.Tensor:
    dim = partition_dim(key)
    if dim is None or len(shards) == 1:
        return shards[0].clone()
    if _is_swiglu_packed(key, swiglu):
        gates, ups = zip(*(s.chunk(2, dim=0) for s in shards))
        return torch.cat(list(gates) + list(ups), dim=0)           # full layout: [all gate rows; all up rows]
    return torch.cat(shards, dim=dim)


def split_tp(full: torch.Tensor, key: str, t: int, swiglu: bool) -> List[torch.Tensor]:
    dim = partition_dim(key)
    if dim is None or t == 1:
        return [full.clone() for _ in range(t)]
    assert full.shape[dim] % t == 0, f"{key}: size {full.shape[dim]} along dim {dim} is not divisible by the target tensor-parallel size {t}"
    if _is_swiglu_packed(key, swiglu):
        gate, up = full.chunk(2, dim=0)
        return [torch.cat([g, u], dim=0).contiguous() for g, u in zip(gate.chunk(t, 0), up.chunk(t, 0))]
    return [c.contiguous() for c in full.chunk(t, dim=dim)]


def _find_iteration(load_dir: str, iteration: Optional[int]) -> Tuple[int, str]:
    if iteration is None:
        with open(os.path.join(load_dir, "latest_checkpointed_iteration.txt")) as f:
            s = f.read().strip()
        if s == "release":
            return 0, os.path.join(load_dir, "release")
        iteration = int(s)
    return iteration, os.path.join(load_dir, "iter_{:07d}".format(iteration))


def _rank_dirs(it_dir: str) -> Tuple[int, int, Dict[Tuple[int, int], str]]:
    found = {}
    for n in sorted(os.listdir(it_dir)):
        m = re.match(r"^mp_rank_(\d\d)(?:_(\d\d\d))?$", n)
        if m and os.path.isfile(os.path.join(it_dir, n, "model_optim_rng.pt")):
            found[(int(m.group(1)), int(m.group(2) or 0))] = os.path.join(it_dir, n, "model_optim_rng.pt")
    if not found:
        raise FileNotFoundError(f"no mp_rank_* directories with model_optim_rng.pt under {it_dir}")
    tp = max(k[0] for k in found) + 1
    pp = max(k[1] for k in found) + 1
    assert len(found) == tp * pp, f"incomplete checkpoint: {len(found)} of {tp * pp} model-parallel ranks present"
    return tp, pp, found


def load_full_model(load_dir: str, iteration: Optional[int] = None):
    """-> (iteration, args of the checkpoint, full nested `language_model` dict with global layer numbers)"""
    iteration, it_dir = _find_iteration(load_dir, iteration)
    tp, pp, files = _rank_dirs(it_dir)
    sds = {k: torch.load(f, map_location="cpu", weights_only=False) for k, f in files.items()}
    ck_args = sds[(0, 0)].get("args")
    swiglu = bool(getattr(ck_args, "swiglu", False))
    num_layers = getattr(ck_args, "num_layers", None)
    # gather (global layer number -> {param: [tp shards]}) and the non-layer sections
    chunks_per_stage = max(1, sum(1 for k in sds[(0, 0)] if re.match(r"^model\d+$", k)))
    per_tp: List[Dict[str, torch.Tensor]] = [OrderedDict() for _ in range(tp)]
    for t in range(tp):
        flat = per_tp[t]
        for p in range(pp):
            sd = sds[(t, p)]
            models = [sd["model"]] if "model" in sd else [sd["model%d" % i] for i in range(chunks_per_stage)]
            layers_here = [sum(1 for k in m["language_model"]["encoder"] if k.endswith("input_layernorm.weight")) for m in models]
            for v, m in enumerate(models):
                lm = m["language_model"]
                if num_layers is not None and chunks_per_stage > 1:
                    offset = v * (num_layers // chunks_per_stage) + p * layers_here[v]      # interleaved placement (models/gpt.py::get_layer_range)
                else:
                    offset = sum(_stage_layers(sds[(t, q)]) for q in range(p))
                for sec, d in lm.get("embedding", {}).items():
                    for k, val in d.items():
                        flat.setdefault(f"embedding.{sec}.{k}", val)       # first stage's copy wins (the last stage holds a duplicate when tied)
                for k, val in lm.get("output_layer", {}).items():
                    flat[f"output_layer.{k}"] = val
                for k, val in lm["encoder"].items():
                    m_ = _LAYER.match(k)
                    if m_:
                        flat[f"encoder.layers.{offset + int(m_.group(1))}.{m_.group(2)}"] = val
                    else:
                        flat[f"encoder.{k}"] = val
    full = OrderedDict()
    for k in per_tp[0]:
        if k.startswith("encoder."):
            name = k[len("encoder."):]
        elif k.startswith("embedding."):
            name = k[len("embedding."):]
        else:
            name = k
        full[k] = merge_tp([per_tp[t][k] for t in range(tp)], name, swiglu)
    return iteration, ck_args, full


def _stage_layers(sd) -> int:
    models = [sd["model"]] if "model" in sd else [sd[k] for k in sorted(sd) if re.match(r"^model\d+$", k)]
    return sum(sum(1 for k in m["language_model"]["encoder"] if k.endswith("input_layernorm.weight")) for m in models)


def save_resharded(full: Dict[str, torch.Tensor], ck_args, iteration: int, save_dir: str, tp: int, pp: int):
    swiglu = bool(getattr(ck_args, "swiglu", False))
    untied = bool(getattr(ck_args, "untie_embeddings_and_output_weights", False))
    layer_ids = sorted({int(k.split(".")[2]) for k in full if k.startswith("encoder.layers.")})
    L = len(layer_ids)
    assert L % pp == 0, f"{L} layers are not divisible by the target pipeline-parallel size {pp}"
    per_stage = L // pp
    it_dir = os.path.join(save_dir, "iter_{:07d}".format(iteration))
    new_args = copy.copy(ck_args)
    if new_args is not None:
        new_args.tensor_model_parallel_size, new_args.pipeline_model_parallel_size = tp, pp
        new_args.virtual_pipeline_model_parallel_size = None
        if hasattr(new_args, "world_size"):
            new_args.data_parallel_size = max(1, getattr(new_args, "world_size", tp * pp) // (tp * pp))
    split_cache = {k: split_tp(v, k.split(".", 1)[1] if k.startswith(("embedding.", "encoder.")) else k, tp, swiglu) for k, v in full.items()}
    for p in range(pp):
        first, last = p == 0, p == pp - 1
        for t in range(tp):
            lm = {"encoder": OrderedDict()}
            for k, parts in split_cache.items():
                val = parts[t]
                if k.startswith("embedding."):
                    _, sec, name = k.split(".", 2)
                    keep = first or (last and not untied and sec == "word_embeddings")      # tied head: the last stage carries a copy
                    if keep:
                        lm.setdefault("embedding", OrderedDict()).setdefault(sec, OrderedDict())[name] = val
                elif k.startswith("output_layer."):
                    if last:
                        lm.setdefault("output_layer", OrderedDict())[k[len("output_layer."):]] = val
                elif k.startswith("encoder.layers."):
                    _, _, idx, name = k.split(".", 3)
                    g = layer_ids.index(int(idx))
                    if p * per_stage <= g < (p + 1) * per_stage:
                        lm["encoder"][f"layers.{g - p * per_stage}.{name}"] = val
                elif k.startswith("encoder."):
                    if last:                                                                 # final_layernorm
                        lm["encoder"][k[len("encoder."):]] = val
            sub = f"mp_rank_{t:02d}" if pp == 1 else f"mp_rank_{t:02d}_{p:03d}"
            os.makedirs(os.path.join(it_dir, sub), exist_ok=True)
            torch.save({"args": new_args, "checkpoint_version": 3.0, "iteration": iteration, "model": {"language_model": lm}},
                       os.path.join(it_dir, sub, "model_optim_rng.pt"))
    with open(os.path.join(save_dir, "latest_checkpointed_iteration.txt"), "w") as f:
        f.write(str(iteration))
    return it_dir


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--load-dir", required=True)
    ap.add_argument("--save-dir", required=True)
    ap.add_argument("--target-tensor-parallel-size", type=int, default=1)
    ap.add_argument("--target-pipeline-parallel-size", type=int, default=1)
    ap.add_argument("--iteration", type=int, default=None, help="default: latest_checkpointed_iteration.txt")
    a = ap.parse_args(argv)
    iteration, ck_args, full = load_full_model(a.load_dir, a.iteration)
    n = sum(v.numel() for v in full.values())
    print(f"merged iteration {iteration}: {len(full)} tensors, {n} parameters")
    out = save_resharded(full, ck_args, iteration, a.save_dir, a.target_tensor_parallel_size, a.target_pipeline_parallel_size)
    print(f"wrote TP={a.target_tensor_parallel_size} PP={a.target_pipeline_parallel_size} checkpoint to {out}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
