"""Decoder-only causal LMs in the HuggingFace families the Alpaca script fine-tunes (reference:
4_training_alpaca_deepspeed/train.py:214-217 `AutoModelForCausalLM.from_pretrained`; logged run = facebook/opt-125m,
BASELINE config = LLaMA-7B):

  * ``LlamaForCausalLM`` - RMSNorm, rotary embeddings (rotate-half), SwiGLU MLP, no biases, untied lm_head.
  * ``OPTForCausalLM``   - LayerNorm, learned positions (offset 2), ReLU MLP, biases, tied lm_head.

Both are written against ops/functional.py (tcgen05 GEMMs with fused epilogues, native RMSNorm/LayerNorm, RoPE,
SwiGLU, flash attention, fused softmax-cross-entropy) with QKV and gate|up projections FUSED into one GEMM each;
``ckpt/hf.py`` converts to and from the HF parameter names / shapes (q_proj, k_proj, v_proj, gate_proj, up_proj).
Layers are exposed as ``model.layers`` so the ZeRO-3 engine can treat one decoder layer as a gather unit.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn

from ..ops import functional as OF
from ..ops import native as N


@dataclass
class LlamaConfig:
    vocab_size: int = 32000
    hidden_size: int = 4096
    intermediate_size: int = 11008
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    max_position_embeddings: int = 2048
    rms_norm_eps: float = 1e-6
    rope_theta: float = 10000.0
    initializer_range: float = 0.02
    tie_word_embeddings: bool = False
    pad_token_id: Optional[int] = None
    model_type: str = "llama"

    @staticmethod
    def llama_7b(**kw):
        return LlamaConfig(**kw)

    @staticmethod
    def tiny(**kw):
        d = dict(vocab_size=512, hidden_size=128, intermediate_size=352, num_hidden_layers=2, num_attention_heads=2, max_position_embeddings=512)
        d.update(kw)
        return LlamaConfig(**d)


@dataclass
class OPTConfig:
    vocab_size: int = 50272
    hidden_size: int = 768
    ffn_dim: int = 3072
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    max_position_embeddings: int = 2048
    layer_norm_eps: float = 1e-5
    initializer_range: float = 0.02
    pad_token_id: int = 1
    model_type: str = "opt"

    @staticmethod
    def opt_125m(**kw):
        return OPTConfig(**kw)

    @staticmethod
    def tiny(**kw):
        d = dict(vocab_size=512, hidden_size=128, ffn_dim=512, num_hidden_layers=2, num_attention_heads=2, max_position_embeddings=512)
        d.update(kw)
        return OPTConfig(**d)


def _w(o, i, std, device, dtype):
    return nn.Parameter(torch.empty(o, i, device=device, dtype=torch.float32).normal_(0.0, std).to(dtype))


class LlamaDecoderLayer(nn.Module):
    def __init__(self, cfg: LlamaConfig, device=None, dtype=torch.float32):
        super().__init__()
        H, F = cfg.hidden_size, cfg.intermediate_size
        self.cfg, self.nh, self.hd = cfg, cfg.num_attention_heads, cfg.hidden_size // cfg.num_attention_heads
        s = cfg.initializer_range
        self.input_layernorm = nn.Parameter(torch.ones(H, device=device, dtype=dtype))
        self.qkv_proj = _w(3 * H, H, s, device, dtype)          # rows: [q | k | v]
        self.o_proj = _w(H, H, s, device, dtype)
        self.post_attention_layernorm = nn.Parameter(torch.ones(H, device=device, dtype=dtype))
        self.gate_up_proj = _w(2 * F, H, s, device, dtype)      # rows: [gate | up]
        self.down_proj = _w(H, F, s, device, dtype)

    def forward(self, x, cos_t, sin_t, pos_i32):
        """x: [S, B, H] -> [S, B, H]"""
        S, B, H = x.shape
        h = OF.rms_norm(x, self.input_layernorm, self.cfg.rms_norm_eps)
        qkv = OF.linear(h, self.qkv_proj)                       # [S, B, 3H]
        qkv2 = qkv.view(S * B, 3 * H)
        if qkv.is_cuda and qkv.dtype in (torch.bfloat16, torch.float16):
            q2 = OF.rope(qkv2[:, :H], cos_t, sin_t, pos_i32, self.nh, self.hd)
            k2 = OF.rope(qkv2[:, H:2 * H], cos_t, sin_t, pos_i32, self.nh, self.hd)
            v2 = qkv2[:, 2 * H:]
        else:
            q2, k2, v2 = (_rope_ref(qkv2[:, i * H:(i + 1) * H], cos_t, sin_t, pos_i32, self.nh, self.hd) if i < 2 else qkv2[:, 2 * H:] for i in range(3))
        a = OF.attention(_v4(q2, S, B, self.nh, self.hd), _v4(k2, S, B, self.nh, self.hd), _v4(v2, S, B, self.nh, self.hd), causal=True)
        x = OF.linear(a.reshape(S, B, H), self.o_proj, residual=x)
        h = OF.rms_norm(x, self.post_attention_layernorm, self.cfg.rms_norm_eps)
        gu = OF.linear(h, self.gate_up_proj)
        return OF.linear(OF.swiglu(gu), self.down_proj, residual=x)


def _v4(t2, S, B, nh, hd):
    """[S*B, nh*hd] column slice (row stride may exceed nh*hd) -> [S, B, nh, hd] strided view"""
    return t2.as_strided((S, B, nh, hd), (B * t2.stride(0), t2.stride(0), hd, 1), t2.storage_offset())


def _rope_ref(x2, cos_t, sin_t, pos, nh, hd):
    T = x2.shape[0]
    x = x2.reshape(T, nh, hd)
    half = hd // 2
    c, s = cos_t[pos.long()][:, None, :].to(x.dtype), sin_t[pos.long()][:, None, :].to(x.dtype)
    a, b = x[..., :half], x[..., half:]
    return torch.cat([a * c - b * s, b * c + a * s], -1).reshape(T, nh * hd)


class LlamaForCausalLM(nn.Module):
    def __init__(self, cfg: LlamaConfig, device=None, dtype=torch.float32):
        super().__init__()
        self.config = cfg
        self.embed_tokens = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size, device=device, dtype=torch.float32).normal_(0, cfg.initializer_range).to(dtype))
        self.layers = nn.ModuleList([LlamaDecoderLayer(cfg, device, dtype) for _ in range(cfg.num_hidden_layers)])
        self.norm = nn.Parameter(torch.ones(cfg.hidden_size, device=device, dtype=dtype))
        self.lm_head = None if cfg.tie_word_embeddings else _w(cfg.vocab_size, cfg.hidden_size, cfg.initializer_range, device, dtype)
        self._rope = None

    def rope_tables(self, device):
        if self._rope is None or self._rope[0].device != device:
            self._rope = OF.rope_tables(self.config.max_position_embeddings, self.config.hidden_size // self.config.num_attention_heads,
                                        self.config.rope_theta, device)
        return self._rope

    def embed(self, input_ids):
        return OF.embedding(input_ids.t().contiguous(), self.embed_tokens)          # [S, B, H]

    def head_loss(self, h, labels):
        """h [S, B, H]; labels [B, S] already aligned with positions (-100 = ignore).  HF shifts inside the model:
        logits[:, :-1] vs labels[:, 1:]; we do the same shift here."""
        S, B, H = h.shape
        hn = OF.rms_norm(h, self.norm, self.config.rms_norm_eps)
        w = self.lm_head if self.lm_head is not None else self.embed_tokens
        logits = OF.linear(hn[:-1].reshape((S - 1) * B, H), w)
        tgt = labels[:, 1:].t().contiguous().view(-1)
        return OF.cross_entropy(logits, tgt, ignore_index=-100)

    def forward(self, input_ids, attention_mask=None, labels=None):
        """attention_mask: right-padding only (DataCollatorForSupervisedDataset) -> causal attention already hides
        later pads from earlier tokens, and pad positions carry label -100, so the mask is not needed."""
        B, S = input_ids.shape
        dev = input_ids.device
        cos_t, sin_t = self.rope_tables(dev)
        pos = torch.arange(S, device=dev, dtype=torch.int32).repeat_interleave(B)      # token t = s*B + b
        h = self.embed(input_ids)
        for layer in self.layers:
            h = layer(h, cos_t, sin_t, pos)
        if labels is None:
            hn = OF.rms_norm(h, self.norm, self.config.rms_norm_eps)
            w = self.lm_head if self.lm_head is not None else self.embed_tokens
            return OF.linear(hn, w).transpose(0, 1)
        return self.head_loss(h, labels)


class OPTDecoderLayer(nn.Module):
    def __init__(self, cfg: OPTConfig, device=None, dtype=torch.float32):
        super().__init__()
        H, F = cfg.hidden_size, cfg.ffn_dim
        self.cfg, self.nh, self.hd = cfg, cfg.num_attention_heads, cfg.hidden_size // cfg.num_attention_heads
        s = cfg.initializer_range
        z = lambda n: nn.Parameter(torch.zeros(n, device=device, dtype=dtype))
        o = lambda n: nn.Parameter(torch.ones(n, device=device, dtype=dtype))
        self.self_attn_layer_norm_w, self.self_attn_layer_norm_b = o(H), z(H)
        self.qkv_proj, self.qkv_bias = _w(3 * H, H, s, device, dtype), z(3 * H)
        self.out_proj, self.out_bias = _w(H, H, s, device, dtype), z(H)
        self.final_layer_norm_w, self.final_layer_norm_b = o(H), z(H)
        self.fc1, self.fc1_bias = _w(F, H, s, device, dtype), z(F)
        self.fc2, self.fc2_bias = _w(H, F, s, device, dtype), z(H)

    def forward(self, x, *unused):
        S, B, H = x.shape
        h = OF.layer_norm(x, self.self_attn_layer_norm_w, self.self_attn_layer_norm_b, self.cfg.layer_norm_eps)
        qkv2 = OF.linear(h, self.qkv_proj, self.qkv_bias).view(S * B, 3 * H)
        a = OF.attention(_v4(qkv2[:, :H], S, B, self.nh, self.hd), _v4(qkv2[:, H:2 * H], S, B, self.nh, self.hd),
                         _v4(qkv2[:, 2 * H:], S, B, self.nh, self.hd), causal=True)
        x = OF.linear(a.reshape(S, B, H), self.out_proj, self.out_bias, residual=x)
        h = OF.layer_norm(x, self.final_layer_norm_w, self.final_layer_norm_b, self.cfg.layer_norm_eps)
        h = OF.bias_act(OF.linear(h, self.fc1), self.fc1_bias, N.ACT_RELU)
        return OF.linear(h, self.fc2, self.fc2_bias, residual=x)


class OPTForCausalLM(nn.Module):
    def __init__(self, cfg: OPTConfig, device=None, dtype=torch.float32):
        super().__init__()
        self.config = cfg
        mk = lambda n, d: nn.Parameter(torch.empty(n, d, device=device, dtype=torch.float32).normal_(0, cfg.initializer_range).to(dtype))
        self.embed_tokens = mk(cfg.vocab_size, cfg.hidden_size)
        self.embed_positions = mk(cfg.max_position_embeddings + 2, cfg.hidden_size)      # HF OPT offsets positions by 2
        self.layers = nn.ModuleList([OPTDecoderLayer(cfg, device, dtype) for _ in range(cfg.num_hidden_layers)])
        self.final_layer_norm_w = nn.Parameter(torch.ones(cfg.hidden_size, device=device, dtype=dtype))
        self.final_layer_norm_b = nn.Parameter(torch.zeros(cfg.hidden_size, device=device, dtype=dtype))
        self.lm_head = None      # tied

    def embed(self, input_ids):
        B, S = input_ids.shape
        pos = torch.arange(2, S + 2, device=input_ids.device).unsqueeze(1).expand(S, B)
        return OF.embedding(input_ids.t().contiguous(), self.embed_tokens) + OF.embedding(pos.contiguous(), self.embed_positions)

    def head_loss(self, h, labels):
        S, B, H = h.shape
        hn = OF.layer_norm(h, self.final_layer_norm_w, self.final_layer_norm_b, self.config.layer_norm_eps)
        logits = OF.linear(hn[:-1].reshape((S - 1) * B, H), self.embed_tokens)
        return OF.cross_entropy(logits, labels[:, 1:].t().contiguous().view(-1), ignore_index=-100)

    def forward(self, input_ids, attention_mask=None, labels=None):
        h = self.embed(input_ids)
        for layer in self.layers:
            h = layer(h)
        if labels is None:
            hn = OF.layer_norm(h, self.final_layer_norm_w, self.final_layer_norm_b, self.config.layer_norm_eps)
            return OF.linear(hn, self.embed_tokens).transpose(0, 1)
        return self.head_loss(h, labels)


def meta_initializer(cfg, device, dtype, seed: int = 0):
    """`materialize(name, meta_param)` for models built with device="meta" (ZeRO-3 construction that never holds the whole
    model, the role of DeepSpeed's zero.Init): norm weights -> 1, biases -> 0, matrices -> N(0, initializer_range) drawn from a
    generator seeded by the parameter's NAME, so every rank creates identical values whatever the order of creation."""
    import zlib
    std = getattr(cfg, "initializer_range", 0.02)
    dev = torch.device(device)

    def materialize(name: str, p: torch.Tensor) -> torch.Tensor:
        leaf = name.rsplit(".", 1)[-1]
        if p.dim() == 1:
            ones = leaf in ("norm", "input_layernorm", "post_attention_layernorm", "weight") or leaf.endswith("_w") or leaf.endswith("norm")
            return (torch.ones if ones else torch.zeros)(p.shape, device=dev, dtype=dtype)
        g = torch.Generator(device=dev).manual_seed((zlib.crc32(name.encode()) + 1000003 * seed) & 0x7FFFFFFF)
        return torch.empty(p.shape, device=dev, dtype=torch.float32).normal_(0.0, std, generator=g).to(dtype)

    return materialize


def build_causal_lm(name_or_config, device=None, dtype=torch.float32, **config_overrides):
    """`--model_name_or_path` resolution without network access: a known name -> config with random init; a local
    directory with config.json (+ weights) -> ckpt.hf loader."""
    import os
    if isinstance(name_or_config, (LlamaConfig, OPTConfig)):
        cfg = name_or_config
    elif isinstance(name_or_config, str) and os.path.isdir(name_or_config):
        from ..ckpt import hf
        return hf.load_pretrained(name_or_config, device=device, dtype=dtype)
    else:
        n = str(name_or_config).lower()
        table = {"facebook/opt-125m": OPTConfig.opt_125m, "opt-125m": OPTConfig.opt_125m, "opt-tiny": OPTConfig.tiny,
                 "llama-7b": LlamaConfig.llama_7b, "huggyllama/llama-7b": LlamaConfig.llama_7b, "decapoda-research/llama-7b-hf": LlamaConfig.llama_7b,
                 "llama-tiny": LlamaConfig.tiny}
        if n not in table:
            raise ValueError(f"unknown model {name_or_config!r}: pass a local HF directory or one of {sorted(table)}")
        cfg = table[n](**config_overrides)
    return (LlamaForCausalLM if cfg.model_type == "llama" else OPTForCausalLM)(cfg, device=device, dtype=dtype)
