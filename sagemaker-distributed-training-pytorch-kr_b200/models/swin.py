"""Swin Transformer (Swin-B) - the only model the reference's Oxford-Pet script can actually train
(2_training_oxford-pet_ddp/util.py:48-54 touches `model.head`; notebook `'model_name': 'swin_b'`).

Same architecture and parameter names as torchvision.models.swin_b (so its state_dicts load): patch-embed
Conv2d(3,128,k4,s4) + LN; 4 stages with depths [2,2,18,2], dims 128/256/512/1024, heads 4/8/16/32 (head_dim 32),
7x7 shifted windows with relative-position bias and cross-window mask, MLP x4 with erf-GELU, stochastic depth
0 -> 0.5, PatchMerging between stages, final LN, average pool, `head` Linear.  87,768,224 parameters at 1000 classes.

Kernel mapping: every Linear (qkv / proj / mlp / patch-merge reduction / head) and the patch-embed convolution
(non-overlapping 4x4 patches == a [B*H/4*W/4, 48] x [48, 128] GEMM) run on the tcgen05 GEMM with bias fused in
the epilogue; LayerNorm and erf-GELU are the native kernels.  The 49-token / head_dim-32 window attention core is
batched matmul + softmax in PyTorch for now (windows are too small for the 128-row flash kernel).
"""
from __future__ import annotations

from typing import List, Optional

import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import functional as OF
from ..ops import native as N


def _relative_position_index(ws: int) -> torch.Tensor:
    coords = torch.stack(torch.meshgrid(torch.arange(ws), torch.arange(ws), indexing="ij")).flatten(1)
    rel = (coords[:, :, None] - coords[:, None, :]).permute(1, 2, 0).contiguous()
    rel[:, :, 0] += ws - 1
    rel[:, :, 1] += ws - 1
    rel[:, :, 0] *= 2 * ws - 1
    return rel.sum(-1).flatten()


class Linear(nn.Module):
    def __init__(self, i, o, bias=True):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(o, i))
        self.bias = nn.Parameter(torch.zeros(o)) if bias else None
        nn.init.trunc_normal_(self.weight, std=0.02)

    def forward(self, x, residual=None):
        return OF.linear(x, self.weight, self.bias, residual=residual)


class LayerNorm(nn.Module):
    def __init__(self, d, eps=1e-5):
        super().__init__()
        self.weight, self.bias, self.eps = nn.Parameter(torch.ones(d)), nn.Parameter(torch.zeros(d)), eps

    def forward(self, x):
        return OF.layer_norm(x, self.weight, self.bias, self.eps)


class ShiftedWindowAttention(nn.Module):
    def __init__(self, dim, window, shift, heads):
        super().__init__()
        self.dim, self.ws, self.shift, self.heads = dim, window, shift, heads
        self.qkv, self.proj = Linear(dim, 3 * dim), Linear(dim, dim)
        self.relative_position_bias_table = nn.Parameter(torch.zeros((2 * window - 1) ** 2, heads))
        nn.init.trunc_normal_(self.relative_position_bias_table, std=0.02)
        self.register_buffer("relative_position_index", _relative_position_index(window))

    def _bias(self):
        n = self.ws * self.ws
        b = self.relative_position_bias_table[self.relative_position_index].view(n, n, -1)
        return b.permute(2, 0, 1).contiguous().unsqueeze(0)

    def forward(self, x):                      # x: [B, H, W, C]
        B, H, W, C = x.shape
        ws = self.ws
        pad_r, pad_b = (ws - W % ws) % ws, (ws - H % ws) % ws
        x = F.pad(x, (0, 0, 0, pad_r, 0, pad_b))
        _, pH, pW, _ = x.shape
        shift = [self.shift if ws < pH else 0, self.shift if ws < pW else 0]
        nW = (pH // ws) * (pW // ws)
        x = OF.window_partition(x, ws, shift)               # cyclic shift + partition in one pass: [B*nW, ws*ws, C]
        qkv = self.qkv(x)                                   # [B*nW, ws*ws, 3C], last dim = [3][heads][head_dim]
        mask = None
        if sum(shift) > 0:
            mask = x.new_zeros((pH, pW))
            hs = ((0, -ws), (-ws, -shift[0]), (-shift[0], None))
            wsl = ((0, -ws), (-ws, -shift[1]), (-shift[1], None))
            c = 0
            for h in hs:
                for w in wsl:
                    mask[h[0]:h[1], w[0]:w[1]] = c
                    c += 1
            mask = mask.view(pH // ws, ws, pW // ws, ws).permute(0, 2, 1, 3).reshape(nW, ws * ws)
            mask = mask.unsqueeze(1) - mask.unsqueeze(2)
            mask = mask.masked_fill(mask != 0, float(-100.0)).masked_fill(mask == 0, float(0.0))      # [nW, T, T]
        # scores + relative-position bias + shift mask + softmax + PV in one fused kernel (ops/functional.window_attention)
        x = OF.window_attention(qkv, self._bias()[0], mask, self.heads, (C // self.heads) ** -0.5)
        x = self.proj(x)
        x = OF.window_reverse(x, B, pH, pW, ws, shift)      # un-partition + shift back in one pass
        return x if (pH == H and pW == W) else x[:, :H, :W, :].contiguous()


_FUSED_MLP = os.environ.get("B2_SWIN_FUSED_MLP", "1") != "0"


class MLP(nn.Module):
    """torchvision names: mlp.0 (Linear), mlp.1 GELU, mlp.2 Dropout, mlp.3 (Linear)"""

    def __init__(self, dim, hidden):
        super().__init__()
        self.add_module("0", Linear(dim, hidden))
        self.add_module("3", Linear(hidden, dim))

    def forward(self, x):
        fc1, fc2 = getattr(self, "0"), getattr(self, "3")
        if _FUSED_MLP:
            return OF.mlp(x, fc1.weight, fc1.bias, fc2.weight, fc2.bias, "gelu_erf")     # bias + GELU / GELU' fused in GEMM epilogues
        h = OF.bias_act(OF.linear(x, fc1.weight, None), fc1.bias, N.ACT_GELU_ERF)      # A/B arm: standalone activation kernels
        return fc2(h)


def stochastic_depth(x, p: float, training: bool):
    if p == 0.0 or not training:
        return x
    keep = 1.0 - p
    noise = torch.empty((x.shape[0],) + (1,) * (x.dim() - 1), dtype=x.dtype, device=x.device).bernoulli_(keep)
    return x * noise.div_(keep)


class SwinBlock(nn.Module):
    def __init__(self, dim, heads, window, shift, sd_prob):
        super().__init__()
        self.norm1, self.attn = LayerNorm(dim), ShiftedWindowAttention(dim, window, shift, heads)
        self.norm2, self.mlp, self.sd = LayerNorm(dim), MLP(dim, 4 * dim), sd_prob

    def forward(self, x):
        x = x + stochastic_depth(self.attn(self.norm1(x)), self.sd, self.training)
        return x + stochastic_depth(self.mlp(self.norm2(x)), self.sd, self.training)


class PatchMerging(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.reduction, self.norm = Linear(4 * dim, 2 * dim, bias=False), LayerNorm(4 * dim)

    def forward(self, x):
        B, H, W, C = x.shape
        x = F.pad(x, (0, 0, 0, W % 2, 0, H % 2))
        x = torch.cat([x[..., 0::2, 0::2, :], x[..., 1::2, 0::2, :], x[..., 0::2, 1::2, :], x[..., 1::2, 1::2, :]], -1)
        return self.reduction(self.norm(x))


class PatchEmbed(nn.Module):
    """features.0: [0] Conv2d(3, C, 4, 4) as a GEMM over non-overlapping patches, [1] permute, [2] LayerNorm"""

    def __init__(self, in_ch, dim, patch):
        super().__init__()
        conv = nn.Conv2d(in_ch, dim, patch, patch)
        self.add_module("0", conv)
        self.add_module("2", LayerNorm(dim))
        self.patch = patch

    def forward(self, x):
        conv, norm, p = getattr(self, "0"), getattr(self, "2"), self.patch
        B, C, H, W = x.shape
        if x.dtype in (torch.bfloat16, torch.float16) and H % p == 0 and W % p == 0:
            cols = x.view(B, C, H // p, p, W // p, p).permute(0, 2, 4, 1, 3, 5).reshape(B * (H // p) * (W // p), C * p * p)
            y = OF.linear(cols.contiguous(), conv.weight.view(conv.out_channels, -1), conv.bias).view(B, H // p, W // p, -1)
        else:
            y = conv(x).permute(0, 2, 3, 1)
        return norm(y)


class SwinTransformer(nn.Module):
    def __init__(self, patch=4, embed_dim=128, depths=(2, 2, 18, 2), heads=(4, 8, 16, 32), window=7, sd_prob=0.5, num_classes=1000):
        super().__init__()
        layers: List[nn.Module] = [PatchEmbed(3, embed_dim, patch)]
        total, idx = sum(depths), 0
        for s, depth in enumerate(depths):
            dim = embed_dim * 2 ** s
            blocks = []
            for i in range(depth):
                blocks.append(SwinBlock(dim, heads[s], window, 0 if i % 2 == 0 else window // 2, sd_prob * idx / (total - 1.0)))
                idx += 1
            layers.append(nn.Sequential(*blocks))
            if s < len(depths) - 1:
                layers.append(PatchMerging(dim))
        self.features = nn.Sequential(*layers)
        nf = embed_dim * 2 ** (len(depths) - 1)
        self.norm = LayerNorm(nf)
        self.head = Linear(nf, num_classes)
        nn.init.zeros_(self.head.bias)

    def forward(self, x):
        x = self.features(x)
        x = self.norm(x)
        x = x.mean(dim=(1, 2))
        return self.head(x)


def swin_b(num_classes: int = 1000, **kw) -> SwinTransformer:
    return SwinTransformer(4, 128, (2, 2, 18, 2), (4, 8, 16, 32), 7, 0.5, num_classes, **kw)


def swin_t(num_classes: int = 1000, **kw) -> SwinTransformer:
    return SwinTransformer(4, 96, (2, 2, 6, 2), (3, 6, 12, 24), 7, 0.2, num_classes, **kw)


def replace_head(model: nn.Module, num_classes: int) -> nn.Module:
    """A real head replacement (the reference only assigns `model.head.out_features`, which leaves the 1000-wide
    weight in place - SURVEY Q5).  Works for swin (`head`) and ResNet-style (`fc`) models."""
    for name in ("head", "fc"):
        old = getattr(model, name, None)
        if old is not None and hasattr(old, "weight"):
            inf = old.weight.shape[1]
            new = Linear(inf, num_classes) if isinstance(old, Linear) else nn.Linear(inf, num_classes)
            new = new.to(device=old.weight.device, dtype=old.weight.dtype)
            setattr(model, name, new)
            return model
    raise AttributeError("model has neither `head` nor `fc`")
