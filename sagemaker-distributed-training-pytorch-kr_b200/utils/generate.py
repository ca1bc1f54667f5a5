"""Autoregressive sampling from the in-tree causal language models (LLaMA / OPT of models/llama.py, the Megatron GPT of models/gpt.py
at TP = 1) - a checkpoint sanity tool, not a serving stack: every step re-runs the full prefix (no KV cache), which keeps it on
exactly the kernels the training forward uses.  Greedy by default; temperature / top-k / top-p sampling like transformers.generate.
"""
from __future__ import annotations

from typing import Optional

import torch


def _logits(model, ids: torch.Tensor) -> torch.Tensor:
    """[B, S] token ids -> [B, V] logits of the last position, for either model family"""
    out = model(ids)
    if hasattr(model, "language_model"):          # Megatron GPT: [S, B, V]
        return out[-1].float()
    return out[:, -1].float()                     # HF-style models: [B, S, V]


def sample_next(logits: torch.Tensor, temperature: float = 0.0, top_k: int = 0, top_p: float = 0.0, generator=None) -> torch.Tensor:
    if temperature <= 0.0:
        return logits.argmax(-1)
    logits = logits / temperature
    if top_k > 0:
        kth = logits.topk(min(top_k, logits.shape[-1]), -1).values[..., -1, None]
        logits = logits.masked_fill(logits < kth, float("-inf"))
    if 0.0 < top_p < 1.0:
        sl, si = logits.sort(-1, descending=True)
        cum = sl.softmax(-1).cumsum(-1)
        drop = cum - sl.softmax(-1) > top_p          # keep the smallest prefix whose mass reaches top_p
        sl = sl.masked_fill(drop, float("-inf"))
        logits = torch.full_like(logits, float("-inf")).scatter(-1, si, sl)
    return torch.multinomial(logits.softmax(-1), 1, generator=generator).squeeze(-1)


@torch.no_grad()
def generate(model, input_ids: torch.Tensor, max_new_tokens: int = 32, eos_token_id: Optional[int] = None, pad_token_id: int = 0,
             temperature: float = 0.0, top_k: int = 0, top_p: float = 0.0, vocab_size: Optional[int] = None, generator=None) -> torch.Tensor:
    """Returns [B, S + n] ids.  `vocab_size` cuts padded vocabulary rows (Megatron pads to a multiple of 128 x TP)."""
    was_training = model.training
    model.eval()
    ids = input_ids.clone()
    done = torch.zeros(ids.shape[0], dtype=torch.bool, device=ids.device)
    for _ in range(max_new_tokens):
        lg = _logits(model, ids)
        if vocab_size is not None:
            lg = lg[:, :vocab_size]
        nxt = sample_next(lg, temperature, top_k, top_p, generator)
        nxt = torch.where(done, torch.full_like(nxt, pad_token_id), nxt)
        ids = torch.cat([ids, nxt[:, None]], 1)
        if eos_token_id is not None:
            done |= nxt == eos_token_id
            if bool(done.all()):
                break
    model.train(was_training)
    return ids
