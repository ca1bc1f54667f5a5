"""Hugging Face GPT-2 <-> Megatron-layout GPT checkpoints.

    python tools/convert_gpt2.py --hf-dir /models/gpt2 --save-dir ckpt_gpt2            # HF -> Megatron (TP = PP = 1)
    python tools/checkpoint_util.py --load-dir ckpt_gpt2 --save-dir ckpt_tp2 --target-tensor-parallel-size 2
    torchrun ... entrypoints/pretrain_gpt.py --load ckpt_tp2 --finetune ...              # continue on the b200dist runtime
    python tools/convert_gpt2.py --megatron-dir ckpt --hf-dir out_hf --to-hf            # Megatron -> HF (config.json + weights)

Layout differences handled here:
  * HF GPT-2 uses Conv1D: weights are stored [in, out]; Megatron linears are [out, in];
  * HF `c_attn` emits [q (all heads) | k | v]; the Megatron QKV weight is ordered per head [head][q | k | v]
    (models/gpt.py), so a tensor-parallel shard is a contiguous block of heads;
  * the vocabulary is padded to a multiple of --make-vocab-size-divisible-by x TP (pad rows are zero);
  * HF ties `lm_head` to `wte`; so does the Megatron GPT unless --untie-embeddings-and-output-weights.
Activation: GPT-2's `gelu_new` is the tanh formula of the fused bias-GeLU epilogue.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from collections import OrderedDict
from types import SimpleNamespace
from typing import Dict

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _qkv_hf_to_meg(w: torch.Tensor, nh: int) -> torch.Tensor:
    """[3h, ...] ordered [q | k | v][head][hd] -> [head][q | k | v][hd]"""
    three_h = w.shape[0]
    hd = three_h // (3 * nh)
    return w.reshape(3, nh, hd, *w.shape[1:]).transpose(0, 1).reshape(three_h, *w.shape[1:]).contiguous()


def _qkv_meg_to_hf(w: torch.Tensor, nh: int) -> torch.Tensor:
    three_h = w.shape[0]
    hd = three_h // (3 * nh)
    return w.reshape(nh, 3, hd, *w.shape[1:]).transpose(0, 1).reshape(three_h, *w.shape[1:]).contiguous()


def hf_gpt2_to_megatron(sd: Dict[str, torch.Tensor], num_layers: int, num_heads: int, padded_vocab: int) -> Dict[str, torch.Tensor]:
    """-> flat state_dict of b200dist.models.gpt.GPTModel (TP = 1)"""
    g = lambda k: sd[k] if k in sd else sd["transformer." + k]        # noqa: E731  (GPT2Model vs GPT2LMHeadModel prefixes)
    out = OrderedDict()
    wte = g("wte.weight")
    if padded_vocab > wte.shape[0]:
        wte = torch.cat([wte, torch.zeros(padded_vocab - wte.shape[0], wte.shape[1], dtype=wte.dtype)], 0)
    out["language_model.embedding.word_embeddings.weight"] = wte.clone()
    out["language_model.embedding.position_embeddings.weight"] = g("wpe.weight").clone()
    for i in range(num_layers):
        src, dst = f"h.{i}.", f"language_model.encoder.layers.{i}."
        out[dst + "input_layernorm.weight"], out[dst + "input_layernorm.bias"] = g(src + "ln_1.weight").clone(), g(src + "ln_1.bias").clone()
        out[dst + "self_attention.query_key_value.weight"] = _qkv_hf_to_meg(g(src + "attn.c_attn.weight").t().contiguous(), num_heads)
        out[dst + "self_attention.query_key_value.bias"] = _qkv_hf_to_meg(g(src + "attn.c_attn.bias"), num_heads)
        out[dst + "self_attention.dense.weight"] = g(src + "attn.c_proj.weight").t().contiguous()
        out[dst + "self_attention.dense.bias"] = g(src + "attn.c_proj.bias").clone()
        out[dst + "post_attention_layernorm.weight"], out[dst + "post_attention_layernorm.bias"] = g(src + "ln_2.weight").clone(), g(src + "ln_2.bias").clone()
        out[dst + "mlp.dense_h_to_4h.weight"] = g(src + "mlp.c_fc.weight").t().contiguous()
        out[dst + "mlp.dense_h_to_4h.bias"] = g(src + "mlp.c_fc.bias").clone()
        out[dst + "mlp.dense_4h_to_h.weight"] = g(src + "mlp.c_proj.weight").t().contiguous()
        out[dst + "mlp.dense_4h_to_h.bias"] = g(src + "mlp.c_proj.bias").clone()
    out["language_model.encoder.final_layernorm.weight"], out["language_model.encoder.final_layernorm.bias"] = g("ln_f.weight").clone(), g("ln_f.bias").clone()
    return out


def megatron_to_hf_gpt2(sd: Dict[str, torch.Tensor], num_layers: int, num_heads: int, vocab_size: int) -> Dict[str, torch.Tensor]:
    """flat GPTModel state_dict (TP = 1) -> GPT2LMHeadModel state_dict"""
    out = OrderedDict()
    out["transformer.wte.weight"] = sd["language_model.embedding.word_embeddings.weight"][:vocab_size].clone()
    out["transformer.wpe.weight"] = sd["language_model.embedding.position_embeddings.weight"].clone()
    for i in range(num_layers):
        dst, src = f"transformer.h.{i}.", f"language_model.encoder.layers.{i}."
        out[dst + "ln_1.weight"], out[dst + "ln_1.bias"] = sd[src + "input_layernorm.weight"].clone(), sd[src + "input_layernorm.bias"].clone()
        out[dst + "attn.c_attn.weight"] = _qkv_meg_to_hf(sd[src + "self_attention.query_key_value.weight"], num_heads).t().contiguous()
        out[dst + "attn.c_attn.bias"] = _qkv_meg_to_hf(sd[src + "self_attention.query_key_value.bias"], num_heads)
        out[dst + "attn.c_proj.weight"] = sd[src + "self_attention.dense.weight"].t().contiguous()
        out[dst + "attn.c_proj.bias"] = sd[src + "self_attention.dense.bias"].clone()
        out[dst + "ln_2.weight"], out[dst + "ln_2.bias"] = sd[src + "post_attention_layernorm.weight"].clone(), sd[src + "post_attention_layernorm.bias"].clone()
        out[dst + "mlp.c_fc.weight"] = sd[src + "mlp.dense_h_to_4h.weight"].t().contiguous()
        out[dst + "mlp.c_fc.bias"] = sd[src + "mlp.dense_h_to_4h.bias"].clone()
        out[dst + "mlp.c_proj.weight"] = sd[src + "mlp.dense_4h_to_h.weight"].t().contiguous()
        out[dst + "mlp.c_proj.bias"] = sd[src + "mlp.dense_4h_to_h.bias"].clone()
    out["transformer.ln_f.weight"], out["transformer.ln_f.bias"] = sd["language_model.encoder.final_layernorm.weight"].clone(), sd["language_model.encoder.final_layernorm.bias"].clone()
    out["lm_head.weight"] = out["transformer.wte.weight"]
    return out


def _nest(flat: Dict[str, torch.Tensor]) -> dict:
    lm = {"embedding": {}, "encoder": OrderedDict()}
    for k, v in flat.items():
        rest = k[len("language_model."):]
        if rest.startswith("embedding."):
            _, sec, name = rest.split(".", 2)
            lm["embedding"].setdefault(sec, OrderedDict())[name] = v
        elif rest.startswith("encoder."):
            lm["encoder"][rest[len("encoder."):]] = v
    return {"language_model": lm}


def _read_hf_weights(hf_dir: str) -> Dict[str, torch.Tensor]:
    from b200dist.ckpt import hf as hf_ckpt
    return hf_ckpt.read_state_dict(hf_dir)


def convert_hf_dir_to_megatron(hf_dir: str, save_dir: str, divisible_by: int = 128, iteration: int = 0) -> str:
    with open(os.path.join(hf_dir, "config.json")) as f:
        cfg = json.load(f)
    L, nh, H, V = cfg["n_layer"], cfg["n_head"], cfg["n_embd"], cfg["vocab_size"]
    padded = (V + divisible_by - 1) // divisible_by * divisible_by
    flat = hf_gpt2_to_megatron(_read_hf_weights(hf_dir), L, nh, padded)
    args = SimpleNamespace(num_layers=L, hidden_size=H, num_attention_heads=nh, ffn_hidden_size=cfg.get("n_inner") or 4 * H, kv_channels=H // nh,
                           max_position_embeddings=cfg["n_positions"], seq_length=cfg["n_positions"], padded_vocab_size=padded,
                           layernorm_epsilon=cfg.get("layer_norm_epsilon", 1e-5), tensor_model_parallel_size=1, pipeline_model_parallel_size=1,
                           swiglu=False, untie_embeddings_and_output_weights=False, make_vocab_size_divisible_by=divisible_by,
                           tokenizer_type="GPT2BPETokenizer", consumed_train_samples=0, consumed_valid_samples=0, iteration=iteration)
    d = os.path.join(save_dir, "release" if iteration == 0 else "iter_{:07d}".format(iteration), "mp_rank_00")
    os.makedirs(d, exist_ok=True)
    torch.save({"args": args, "checkpoint_version": 3.0, "iteration": iteration, "model": _nest(flat)}, os.path.join(d, "model_optim_rng.pt"))
    with open(os.path.join(save_dir, "latest_checkpointed_iteration.txt"), "w") as f:
        f.write("release" if iteration == 0 else str(iteration))
    return d


def convert_megatron_dir_to_hf(megatron_dir: str, hf_dir: str, vocab_size: int = None) -> str:
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import checkpoint_util as cu
    iteration, a, full = cu.load_full_model(megatron_dir)
    flat = {}
    for k, v in full.items():
        flat["language_model." + k] = v
    L, nh = a.num_layers, a.num_attention_heads
    V = vocab_size or getattr(a, "orig_vocab_size", None) or full["embedding.word_embeddings.weight"].shape[0]
    sd = megatron_to_hf_gpt2(flat, L, nh, V)
    os.makedirs(hf_dir, exist_ok=True)
    cfg = {"architectures": ["GPT2LMHeadModel"], "model_type": "gpt2", "n_layer": L, "n_head": nh, "n_embd": a.hidden_size, "n_inner": a.ffn_hidden_size,
           "n_positions": a.max_position_embeddings, "n_ctx": a.max_position_embeddings, "vocab_size": V, "activation_function": "gelu_new",
           "layer_norm_epsilon": getattr(a, "layernorm_epsilon", 1e-5), "tie_word_embeddings": True, "torch_dtype": "float32"}
    with open(os.path.join(hf_dir, "config.json"), "w") as f:
        json.dump(cfg, f, indent=2)
    torch.save({k: v.float() for k, v in sd.items() if k != "lm_head.weight"}, os.path.join(hf_dir, "pytorch_model.bin"))
    return hf_dir


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--hf-dir", required=True)
    ap.add_argument("--save-dir", help="Megatron checkpoint root to write (HF -> Megatron)")
    ap.add_argument("--megatron-dir", help="Megatron checkpoint root to read (with --to-hf)")
    ap.add_argument("--to-hf", action="store_true")
    ap.add_argument("--make-vocab-size-divisible-by", type=int, default=128)
    ap.add_argument("--vocab-size", type=int, default=None, help="unpadded vocabulary size for --to-hf")
    a = ap.parse_args(argv)
    if a.to_hf:
        assert a.megatron_dir, "--to-hf needs --megatron-dir"
        print("wrote", convert_megatron_dir_to_hf(a.megatron_dir, a.hf_dir, a.vocab_size))
    else:
        assert a.save_dir, "HF -> Megatron needs --save-dir"
        print("wrote", convert_hf_dir_to_megatron(a.hf_dir, a.save_dir, a.make_vocab_size_divisible_by))
    return 0


if __name__ == "__main__":
    sys.exit(main())
