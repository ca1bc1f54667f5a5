"""Swin Transformer (Swin-B) - the only model the reference's Oxford-Pet script can actually train
(2_training_oxford-pet_ddp/util.py:48-54 touches `model.head`; notebook `'model_name': 'swin_b'`).

Same architecture and parameter names as torchvision.models.swin_b (so its state_dicts load): patch-embed
Conv2d(3,128,k4,s4) + LN; 4 stages with depths [2,2,18,2], dims 128/256/512/1024, heads 4/8/16/32 (head_dim 32),
7x7 shifted windows with relative-position bias and cross-window mask, MLP x4 with erf-GELU, stochastic depth
0 -> 0.5, PatchMerging between stages, final LN, average pool, `head` Linear.  87,768,224 parameters at 1000 classes.

Kernel mapping: every Linear (qkv / proj / mlp / patch-merge reduction / head) and the patch-embed convolution
(non-overlapping 4x4 patches == a [B*H/4*W/4, 48] x [48, 128] GEMM) run on the tcgen05 GEMM with bias fused in
the epilogue; LayerNorm and erf-GELU are the native kernels.  The 49-token / head_dim-32 window attention core is the fused
tensor-core kernel of csrc/window_attn_tc.cu (scores + relative-position bias + shift mask + softmax + PV); the cyclic shift and
window partition are one gather pass, and the way back (un-partition + shift back) is folded into the residual add together with
stochastic depth (csrc/window_attn.cu: rows_scale_add_kernel).
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

    def with_skip(self, x):
        """(norm(x), x'): x' == x is what the block's skip connection should use (gradients meet inside the norm backward kernel)."""
        return OF.layer_norm_residual(x, self.weight, self.bias, self.eps)


class _BiasGather(torch.autograd.Function):
    """relative_position_bias_table [(2 ws - 1)^2, heads] -> fp32 bias [heads, T, T] (T = ws^2) through the fixed index.  The backward is
    an index_add_ (atomics) instead of autograd's sort-based index_put of the plain `table[index]` expression."""

    @staticmethod
    def forward(ctx, table, index, n):
        ctx.save_for_backward(index)
        ctx.shape, ctx.dtype = table.shape, table.dtype
        return table.float().t().index_select(1, index).view(-1, n, n)

    @staticmethod
    def backward(ctx, dbias):
        (index,) = ctx.saved_tensors
        heads = ctx.shape[1]
        dt = torch.zeros((heads, ctx.shape[0]), dtype=torch.float32, device=dbias.device)
        dt.index_add_(1, index, dbias.reshape(heads, -1).float())
        return dt.t().contiguous().to(ctx.dtype), None, None          # contiguous like the parameter: AccumulateGrad keeps it as it is


_MASK_CACHE = {}


def shift_window_mask(pH: int, pW: int, ws: int, shift, device) -> torch.Tensor:
    """torchvision's cross-window attention mask of a cyclically shifted map: fp32 [nW, ws^2, ws^2], 0 where two tokens of a window come
    from the same image region and -100 where they do not.  It depends on the geometry only, so it is built once per geometry and
    device (torchvision rebuilds it with ~15 small kernels in every forward)."""
    key = (pH, pW, ws, int(shift[0]), int(shift[1]), str(device))
    m = _MASK_CACHE.get(key)
    if m is None:
        region = torch.zeros((pH, pW), dtype=torch.float32)
        hs = ((0, -ws), (-ws, -shift[0]), (-shift[0], None))
        wsl = ((0, -ws), (-ws, -shift[1]), (-shift[1], None))
        c = 0
        for h in hs:
            for w in wsl:
                region[h[0]:h[1], w[0]:w[1]] = c
                c += 1
        nW = (pH // ws) * (pW // ws)
        region = region.view(pH // ws, ws, pW // ws, ws).permute(0, 2, 1, 3).reshape(nW, ws * ws)
        diff = region.unsqueeze(1) - region.unsqueeze(2)
        m = torch.where(diff != 0, torch.tensor(-100.0), torch.tensor(0.0)).contiguous().to(device)
        _MASK_CACHE[key] = m                 # never evicted: captured CUDA graphs read it by address
    return m


class ShiftedWindowAttention(nn.Module):
    def __init__(self, dim, window, shift, heads):
        super().__init__()
        self.dim, self.ws, self.shift, self.heads = dim, window, shift, heads
        self.qkv, self.proj = Linear(dim, 3 * dim), Linear(dim, dim)
        self.relative_position_bias_table = nn.Parameter(torch.zeros((2 * window - 1) ** 2, heads))
        nn.init.trunc_normal_(self.relative_position_bias_table, std=0.02)
        self.register_buffer("relative_position_index", _relative_position_index(window))

    def _bias(self):
        """fp32 [1, heads, T, T]"""
        return _BiasGather.apply(self.relative_position_bias_table, self.relative_position_index, self.ws * self.ws).unsqueeze(0)

    def _shift_for(self, pH: int, pW: int):
        return [self.shift if self.ws < pH else 0, self.shift if self.ws < pW else 0]

    def _windows(self, xw, pH: int, pW: int, shift):
        """xw: [B*nW, ws*ws, C] (normalised, shifted, partitioned) -> projected attention output in the same window order"""
        C = xw.shape[-1]
        qkv = self.qkv(xw)                                  # [B*nW, ws*ws, 3C], last dim = [3][heads][head_dim]
        mask = shift_window_mask(pH, pW, self.ws, shift, xw.device) if sum(shift) > 0 else None      # [nW, T, T] fp32, cached
        # scores + relative-position bias + shift mask + softmax + PV in one fused kernel (ops/functional.window_attention)
        a = OF.window_attention(qkv, self._bias()[0], mask, self.heads, (C // self.heads) ** -0.5)
        return self.proj(a)

    def forward_windows(self, xw, B: int, H: int, W: int, shift, skip, sd_prob: float = 0.0):
        """The block's whole attention branch for an input that is already in window order (the norm wrote it there,
        ops.functional.layer_norm_partition): skip + stochastic_depth(window_reverse(attention(xw))), the way back fused into the add."""
        return OF.residual_sd_add(skip, self._windows(xw, H, W, shift), sd_prob, self.training, window_geom=(B, H, W, self.ws, shift))

    def forward(self, x, skip=None, sd_prob: float = 0.0):
        """x: [B, H, W, C] (already normalised).  With ``skip`` the block's residual add (skip + stochastic_depth(attention(x))) happens
        here, fused with the un-partition / shift back of the attention output."""
        B, H, W, C = x.shape
        ws = self.ws
        pad_r, pad_b = (ws - W % ws) % ws, (ws - H % ws) % ws
        x = F.pad(x, (0, 0, 0, pad_r, 0, pad_b))
        _, pH, pW, _ = x.shape
        shift = self._shift_for(pH, pW)
        x = OF.window_partition(x, ws, shift)               # cyclic shift + partition in one pass: [B*nW, ws*ws, C]
        x = self._windows(x, pH, pW, shift)
        if skip is not None and pH == H and pW == W:
            # skip + stochastic_depth(window_reverse(x)) in one pass
            return OF.residual_sd_add(skip, x, sd_prob, self.training, window_geom=(B, pH, pW, ws, shift))
        x = OF.window_reverse(x, B, pH, pW, ws, shift)      # un-partition + shift back in one pass
        x = x if (pH == H and pW == W) else x[:, :H, :W, :].contiguous()
        return x if skip is None else OF.residual_sd_add(skip, x, sd_prob, self.training)


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
        B, H, W, _ = x.shape
        ws = self.attn.ws
        if H % ws == 0 and W % ws == 0:
            # norm1 writes its rows straight into shifted-window order; the way back is folded into the residual add
            shift = self.attn._shift_for(H, W)
            xw, skip = OF.layer_norm_partition(x, self.norm1.weight, self.norm1.bias, self.norm1.eps, ws, shift)
            x = self.attn.forward_windows(xw, B, H, W, shift, skip, self.sd)   # skip + stochastic_depth(attn(norm1(x)))
        else:
            h, skip = self.norm1.with_skip(x)
            x = self.attn(h, skip=skip, sd_prob=self.sd)
        h, skip = self.norm2.with_skip(x)
        return OF.residual_sd_add(skip, self.mlp(h), self.sd, self.training)   # skip + stochastic_depth(mlp(norm2(x)))


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
        if x.dtype in (torch.bfloat16, torch.float16) and H % p == 0 and W % p == 0 and not x.requires_grad:
            cols = N.patchify(x, p)                      # [B * H/p * W/p, C*p*p]: the input images need no gradient
            y = OF.linear(cols, conv.weight.view(conv.out_channels, -1), conv.bias).view(B, H // p, W // p, -1)
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
