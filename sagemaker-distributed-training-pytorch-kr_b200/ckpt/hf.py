"""HuggingFace checkpoint layout: config.json + (sharded) weights `pytorch_model-0000i-of-0000N.bin` /
`model-0000i-of-0000N.safetensors` with `*.index.json` = {"metadata": {"total_size"}, "weight_map": {name: file}}
(reference reads them through AutoModelForCausalLM.from_pretrained, train.py:214-217, and writes them in
Trainer.save_model when weights are gathered).

Our models fuse q/k/v and gate/up; this module converts between the fused in-memory parameters and the HF names
(`model.layers.{i}.self_attn.{q,k,v,o}_proj.weight`, `mlp.{gate,up,down}_proj.weight`, ... / OPT's
`model.decoder.layers.{i}.self_attn.{q,k,v,out}_proj.{weight,bias}`, `fc1`, `fc2`, ...).
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Optional

import torch


# ------------------------------------------------------------------------------------------ name mapping
def to_hf_state_dict(model) -> Dict[str, torch.Tensor]:
    cfg = model.config
    sd: Dict[str, torch.Tensor] = {}
    H = cfg.hidden_size
    if cfg.model_type == "llama":
        sd["model.embed_tokens.weight"] = model.embed_tokens.data
        for i, l in enumerate(model.layers):
            p = f"model.layers.{i}."
            q, k, v = l.qkv_proj.data.split(H, 0)
            g, u = l.gate_up_proj.data.split(cfg.intermediate_size, 0)
            sd.update({p + "self_attn.q_proj.weight": q, p + "self_attn.k_proj.weight": k, p + "self_attn.v_proj.weight": v,
                       p + "self_attn.o_proj.weight": l.o_proj.data, p + "mlp.gate_proj.weight": g, p + "mlp.up_proj.weight": u,
                       p + "mlp.down_proj.weight": l.down_proj.data, p + "input_layernorm.weight": l.input_layernorm.data,
                       p + "post_attention_layernorm.weight": l.post_attention_layernorm.data})
        sd["model.norm.weight"] = model.norm.data
        sd["lm_head.weight"] = (model.lm_head if model.lm_head is not None else model.embed_tokens).data
    else:
        d = "model.decoder."
        sd[d + "embed_tokens.weight"] = model.embed_tokens.data
        sd[d + "embed_positions.weight"] = model.embed_positions.data
        for i, l in enumerate(model.layers):
            p = f"{d}layers.{i}."
            (qw, kw, vw), (qb, kb, vb) = l.qkv_proj.data.split(H, 0), l.qkv_bias.data.split(H, 0)
            sd.update({p + "self_attn.q_proj.weight": qw, p + "self_attn.k_proj.weight": kw, p + "self_attn.v_proj.weight": vw,
                       p + "self_attn.q_proj.bias": qb, p + "self_attn.k_proj.bias": kb, p + "self_attn.v_proj.bias": vb,
                       p + "self_attn.out_proj.weight": l.out_proj.data, p + "self_attn.out_proj.bias": l.out_bias.data,
                       p + "self_attn_layer_norm.weight": l.self_attn_layer_norm_w.data, p + "self_attn_layer_norm.bias": l.self_attn_layer_norm_b.data,
                       p + "fc1.weight": l.fc1.data, p + "fc1.bias": l.fc1_bias.data, p + "fc2.weight": l.fc2.data, p + "fc2.bias": l.fc2_bias.data,
                       p + "final_layer_norm.weight": l.final_layer_norm_w.data, p + "final_layer_norm.bias": l.final_layer_norm_b.data})
        sd[d + "final_layer_norm.weight"] = model.final_layer_norm_w.data
        sd[d + "final_layer_norm.bias"] = model.final_layer_norm_b.data
        sd["lm_head.weight"] = model.embed_tokens.data
    return {k: v.detach().cpu().contiguous() for k, v in sd.items()}


def load_hf_state_dict(model, sd: Dict[str, torch.Tensor], strict: bool = True):
    cfg = model.config
    get = lambda k: sd[k]
    with torch.no_grad():
        if cfg.model_type == "llama":
            model.embed_tokens.copy_(get("model.embed_tokens.weight"))
            for i, l in enumerate(model.layers):
                p = f"model.layers.{i}."
                l.qkv_proj.copy_(torch.cat([get(p + f"self_attn.{n}_proj.weight") for n in "qkv"], 0))
                l.o_proj.copy_(get(p + "self_attn.o_proj.weight"))
                l.gate_up_proj.copy_(torch.cat([get(p + "mlp.gate_proj.weight"), get(p + "mlp.up_proj.weight")], 0))
                l.down_proj.copy_(get(p + "mlp.down_proj.weight"))
                l.input_layernorm.copy_(get(p + "input_layernorm.weight"))
                l.post_attention_layernorm.copy_(get(p + "post_attention_layernorm.weight"))
            model.norm.copy_(get("model.norm.weight"))
            if model.lm_head is not None:
                model.lm_head.copy_(sd.get("lm_head.weight", sd["model.embed_tokens.weight"]))
        else:
            d = "model.decoder." if "model.decoder.embed_tokens.weight" in sd else "decoder."
            model.embed_tokens.copy_(get(d + "embed_tokens.weight"))
            model.embed_positions.copy_(get(d + "embed_positions.weight"))
            for i, l in enumerate(model.layers):
                p = f"{d}layers.{i}."
                l.qkv_proj.copy_(torch.cat([get(p + f"self_attn.{n}_proj.weight") for n in "qkv"], 0))
                l.qkv_bias.copy_(torch.cat([get(p + f"self_attn.{n}_proj.bias") for n in "qkv"], 0))
                l.out_proj.copy_(get(p + "self_attn.out_proj.weight")); l.out_bias.copy_(get(p + "self_attn.out_proj.bias"))
                l.self_attn_layer_norm_w.copy_(get(p + "self_attn_layer_norm.weight")); l.self_attn_layer_norm_b.copy_(get(p + "self_attn_layer_norm.bias"))
                l.fc1.copy_(get(p + "fc1.weight")); l.fc1_bias.copy_(get(p + "fc1.bias"))
                l.fc2.copy_(get(p + "fc2.weight")); l.fc2_bias.copy_(get(p + "fc2.bias"))
                l.final_layer_norm_w.copy_(get(p + "final_layer_norm.weight")); l.final_layer_norm_b.copy_(get(p + "final_layer_norm.bias"))
            model.final_layer_norm_w.copy_(get(d + "final_layer_norm.weight"))
            model.final_layer_norm_b.copy_(get(d + "final_layer_norm.bias"))


def rename_unit_keys(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """keys produced by Zero3Engine.gather_full_state_dict ('embed.embed_tokens', 'layers.3.qkv_proj', 'head.lm_head')
    -> attribute paths of the model ('embed_tokens', 'layers.3.qkv_proj', 'lm_head')."""
    out = {}
    for k, v in sd.items():
        for pre in ("embed.", "head."):
            if k.startswith(pre):
                k = k[len(pre):]
        out[k] = v
    return out


# ------------------------------------------------------------------------------------------ files
def _config_dict(cfg) -> Dict:
    d = {k: v for k, v in vars(cfg).items()}
    d["architectures"] = ["LlamaForCausalLM" if cfg.model_type == "llama" else "OPTForCausalLM"]
    return d


def shard_state_dict(sd: Dict[str, torch.Tensor], max_shard_bytes: int) -> List[Dict[str, torch.Tensor]]:
    shards, cur, size = [], {}, 0
    for k, v in sd.items():
        n = v.numel() * v.element_size()
        if cur and size + n > max_shard_bytes:
            shards.append(cur); cur, size = {}, 0
        cur[k] = v; size += n
    if cur:
        shards.append(cur)
    return shards


def save_pretrained(model, out_dir: str, max_shard_size: int = 10 * 2 ** 30, safe_serialization: bool = False,
                    state_dict_override: Optional[Dict[str, torch.Tensor]] = None):
    os.makedirs(out_dir, exist_ok=True)
    if state_dict_override is not None:        # gathered ZeRO-3 weights: load into a CPU skeleton view for name mapping
        import copy
        for k, v in state_dict_override.items():
            obj, parts = model, k.split(".")
            for a in parts[:-1]:
                obj = obj[int(a)] if a.isdigit() else getattr(obj, a)
            tgt = getattr(obj, parts[-1])
            if tgt.data.shape == v.shape:
                tgt.data = v.to(tgt.dtype)
            else:
                setattr(obj, parts[-1], torch.nn.Parameter(v.to(tgt.dtype)))
    sd = to_hf_state_dict(model)
    with open(os.path.join(out_dir, "config.json"), "w") as f:
        json.dump(_config_dict(model.config), f, indent=2)
    shards = shard_state_dict(sd, max_shard_size)
    ext, stem = (".safetensors", "model") if safe_serialization else (".bin", "pytorch_model")

    def write(path, d):
        if safe_serialization:
            from safetensors.torch import save_file
            save_file({k: v.clone() for k, v in d.items()}, path, metadata={"format": "pt"})
        else:
            torch.save(d, path)
    if len(shards) == 1:
        write(os.path.join(out_dir, stem + ext), shards[0])
        return
    weight_map, total = {}, 0
    for i, d in enumerate(shards):
        fn = f"{stem}-{i + 1:05d}-of-{len(shards):05d}{ext}"
        write(os.path.join(out_dir, fn), d)
        for k, v in d.items():
            weight_map[k] = fn
            total += v.numel() * v.element_size()
    with open(os.path.join(out_dir, stem + ext + ".index.json"), "w") as f:
        json.dump({"metadata": {"total_size": total}, "weight_map": weight_map}, f, indent=2)


def read_state_dict(model_dir: str) -> Dict[str, torch.Tensor]:
    def rd(path):
        if path.endswith(".safetensors"):
            from safetensors.torch import load_file
            return load_file(path)
        return torch.load(path, map_location="cpu", weights_only=True)
    for stem, ext in (("model", ".safetensors"), ("pytorch_model", ".bin")):
        idx = os.path.join(model_dir, stem + ext + ".index.json")
        one = os.path.join(model_dir, stem + ext)
        if os.path.exists(idx):
            wm = json.load(open(idx))["weight_map"]
            sd = {}
            for fn in sorted(set(wm.values())):
                sd.update(rd(os.path.join(model_dir, fn)))
            return sd
        if os.path.exists(one):
            return rd(one)
    raise FileNotFoundError(f"no HF weights (pytorch_model*.bin / model*.safetensors) under {model_dir}")


def load_pretrained(model_dir: str, device=None, dtype=torch.float32):
    from ..models.llama import LlamaConfig, LlamaForCausalLM, OPTConfig, OPTForCausalLM
    raw = json.load(open(os.path.join(model_dir, "config.json")))
    mt = raw.get("model_type", "llama")
    cls_cfg, cls = (LlamaConfig, LlamaForCausalLM) if mt == "llama" else (OPTConfig, OPTForCausalLM)
    names = {f.name for f in __import__("dataclasses").fields(cls_cfg)}
    cfg = cls_cfg(**{k: v for k, v in raw.items() if k in names})
    model = cls(cfg, device=device, dtype=dtype)
    try:
        load_hf_state_dict(model, read_state_dict(model_dir))
    except FileNotFoundError:
        print(f"[hf] no weight files under {model_dir}: keeping random init")
    return model
