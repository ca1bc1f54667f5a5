"""Autograd-level ops on top of the native kernels (ops/native.py).

These are the building blocks for the models that are not hand-scheduled per layer (Swin-B, OPT, LLaMA, the
MNIST head): each Function's forward AND backward call the sm_100a kernels on CUDA tensors and the fp32 PyTorch
restatement on CPU tensors.  Weight gradients are accumulated straight into ``param.main_grad`` (fp32) when the
parameter has one - the same contract as the GPT layer - otherwise returned to autograd.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import native as N


def _lowp(t: torch.Tensor) -> bool:
    return t.dtype in (torch.bfloat16, torch.float16)


def _wgrad(param, dy2, x2):
    gb = getattr(param, "grad_buffer", None)
    if gb is not None and gb.is_cuda and gb.dtype == dy2.dtype:
        # ZeRO-3: the weight gradient is written straight into the symmetric gradient workspace
        N.gemm(dy2, x2, a_mn=True, b_mn=True, out=gb.view(param.shape[0], -1))
        param._gb_written = True
        return None
    mg = getattr(param, "main_grad", None)
    if mg is not None:
        N.gemm(dy2, x2, a_mn=True, b_mn=True, out=mg, accumulate=True)
        return None
    out = torch.zeros(param.shape, dtype=torch.float32, device=param.device)
    N.gemm(dy2, x2, a_mn=True, b_mn=True, out=out, accumulate=True)
    return out.to(param.dtype)


def _bgrad(param, dy2):
    if param is None:
        return None
    mg = getattr(param, "main_grad", None)
    acc = mg if mg is not None else torch.zeros(param.shape, dtype=torch.float32, device=param.device)
    N.colsum_accum(dy2, acc)
    return None if mg is not None else acc.to(param.dtype)


def _vgrad(param, g32):
    if param is None:
        return None
    mg = getattr(param, "main_grad", None)
    if mg is not None:
        mg.add_(g32)
        return None
    return g32.to(param.dtype)


class _Linear(torch.autograd.Function):
    """y = act(x W^T + b) (+ residual).  act in {None, 'gelu_tanh'}; bias / activation / residual ride in the GEMM epilogue."""

    @staticmethod
    def forward(ctx, x, w, b, residual, act):
        shp = x.shape
        x2 = x.reshape(-1, shp[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        z = None
        if act == "gelu_tanh":
            z = torch.empty((x2.shape[0], w.shape[0]), dtype=x.dtype, device=x.device)
            y = N.gemm(x2, w, bias=b, epilogue=N.EPI_BIAS_GELU_DUAL, out2=z)
        elif residual is not None:
            y = N.gemm(x2, w, bias=b, aux=residual.reshape(-1, w.shape[0]), epilogue=N.EPI_BIAS_RESIDUAL)
        else:
            agw = getattr(w, "_zero3_agw", None)
            if agw is not None and b is None and agw[0]._agw_pending == agw[1]:
                y = agw[0].agw_linear(agw[1], x2)      # ZeRO-3: the weight's remote shards are gathered inside this GEMM
            else:
                y = N.gemm(x2, w, bias=b, epilogue=N.EPI_BIAS if b is not None else N.EPI_NONE)
        ctx.save_for_backward(x2, w, z)
        ctx.params, ctx.shp, ctx.has_res, ctx.act = (w, b), shp, residual is not None, act
        return y.view(*shp[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w, z = ctx.saved_tensors
        wp, bp = ctx.params
        dy2 = dy.reshape(-1, w.shape[0])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        d_res = dy if ctx.has_res else None
        if ctx.act == "gelu_tanh":
            dy2 = N.bias_act_bwd(dy2, z, None, N.ACT_GELU_TANH)      # z already holds x W^T + b
        gb = _bgrad(bp, dy2)
        gw = _wgrad(wp, dy2, x2) if wp.requires_grad else None
        dx = N.gemm(dy2, w, b_mn=True).view(ctx.shp) if ctx.needs_input_grad[0] else None
        return dx, gw, gb, d_res, None


class _MLP(torch.autograd.Function):
    """y = fc2(gelu(fc1(x))) as one autograd node: bias + GELU ride in the fc1 GEMM epilogue (both z and gelu(z) are
    written), and gelu'(z) is applied in the epilogue of the fc2 data-gradient GEMM - no standalone activation kernels."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, erf):
        shp = x.shape
        x2 = x.reshape(-1, shp[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        flag = N.EPI_ERF if erf else 0
        z = torch.empty((x2.shape[0], w1.shape[0]), dtype=x.dtype, device=x.device)
        a = N.gemm(x2, w1, bias=b1, epilogue=N.EPI_BIAS_GELU_DUAL | flag, out2=z)
        y = N.gemm(a, w2, bias=b2, epilogue=N.EPI_BIAS if b2 is not None else N.EPI_NONE)
        ctx.save_for_backward(x2, w1, w2, z, a)
        ctx.params, ctx.shp, ctx.flag = (w1, b1, w2, b2), shp, flag
        return y.view(*shp[:-1], w2.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w1, w2, z, a = ctx.saved_tensors
        w1p, b1p, w2p, b2p = ctx.params
        dy2 = dy.reshape(-1, w2.shape[0])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        gb2 = _bgrad(b2p, dy2)
        gw2 = _wgrad(w2p, dy2, a) if w2p.requires_grad else None
        dz = N.gemm(dy2, w2, b_mn=True, epilogue=N.EPI_DGELU | ctx.flag, aux=z)       # (dy W2) * gelu'(z)
        gb1 = _bgrad(b1p, dz)
        gw1 = _wgrad(w1p, dz, x2) if w1p.requires_grad else None
        dx = N.gemm(dz, w1, b_mn=True).view(ctx.shp) if ctx.needs_input_grad[0] else None
        return dx, gw1, gb1, gw2, gb2, None


def mlp(x, w1, b1, w2, b2, act: str = "gelu_erf"):
    """Two-layer GELU MLP (Swin / ViT / OPT-style).  act: 'gelu_erf' (exact) or 'gelu_tanh'."""
    if not _lowp(x) or x.shape[-1] % 8 or w1.shape[0] % 8 or w2.shape[0] % 8:
        h = torch.nn.functional.gelu(torch.nn.functional.linear(x, w1, b1), approximate="tanh" if act == "gelu_tanh" else "none")
        return torch.nn.functional.linear(h, w2, b2)
    return _MLP.apply(x, w1, b1, w2, b2, act == "gelu_erf")


def linear(x, weight, bias=None, residual=None, act: Optional[str] = None):
    """Low-precision (bf16/fp16) inputs run the tcgen05 GEMM; fp32 inputs (MNIST / fp32 parity runs) use torch."""
    if not _lowp(x) or x.shape[-1] % 8 or weight.shape[0] % 8:
        y = torch.nn.functional.linear(x, weight, bias)
        if act == "gelu_tanh":
            y = torch.nn.functional.gelu(y, approximate="tanh")
        return y + residual if residual is not None else y
    return _Linear.apply(x, weight, bias, residual, act)


class _Norm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, eps, rms):
        xc = x.contiguous()
        y, mean, rstd = N.norm_fwd(xc, w, b, eps, rms)
        ctx.save_for_backward(xc, w, mean if mean is not None else rstd, rstd)
        ctx.params, ctx.rms = (w, b), rms
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, mean, rstd = ctx.saved_tensors
        dx, dg, db = N.norm_bwd(dy.contiguous(), x, w, None if ctx.rms else mean, rstd, ctx.rms)
        return dx, _vgrad(ctx.params[0], dg), (_vgrad(ctx.params[1], db) if ctx.params[1] is not None else None), None, None


class _NormResidual(torch.autograd.Function):
    """(LN(x), x): the second output aliases the input, so a pre-norm residual block hands it to its residual add and the two
    gradients that reach x (through the norm and through the skip connection) are summed INSIDE the norm backward kernel
    (norm_bwd's `dres`) instead of by a separate autograd accumulation pass."""

    @staticmethod
    def forward(ctx, x, w, b, eps):
        xc = x.contiguous()
        y, mean, rstd = N.norm_fwd(xc, w, b, eps, False)
        ctx.save_for_backward(xc, w, mean, rstd)
        ctx.params = (w, b)
        ctx.set_materialize_grads(False)
        return y, xc.view_as(xc)

    @staticmethod
    def backward(ctx, dy, dskip):
        x, w, mean, rstd = ctx.saved_tensors
        if dy is None:
            return dskip, None, None, None
        dx, dg, db = N.norm_bwd(dy.contiguous(), x, w, mean, rstd, False, dres=dskip.contiguous() if dskip is not None else None)
        return dx, _vgrad(ctx.params[0], dg), _vgrad(ctx.params[1], db), None


def layer_norm_residual(x, weight, bias, eps: float = 1e-5):
    """Returns (layer_norm(x), x') with x' == x: use x' for the skip connection of a pre-norm block (see _NormResidual)."""
    if not _lowp(x) or x.shape[-1] % 8:
        return torch.nn.functional.layer_norm(x, (x.shape[-1],), weight, bias, eps), x
    return _NormResidual.apply(x, weight, bias, eps)


_ROW_MAPS = {}


def window_row_map(B: int, H: int, W: int, ws: int, shift, device) -> torch.Tensor:
    """int32 [B*H*W]: position, in shifted-window order, of image-order row r (the inverse of the gather window_partition does).
    Geometry only, so it is built once per (geometry, device)."""
    key = (B, H, W, ws, int(shift[0]), int(shift[1]), str(device))
    m = _ROW_MAPS.get(key)
    if m is None:
        ids = torch.arange(B * H * W, dtype=torch.int32).view(B, H, W, 1)
        src = N.window_shuffle(ids, B, H, W, ws, shift, False).reshape(-1)        # image row that lands at window position p
        m = torch.empty_like(src)
        m[src.long()] = torch.arange(src.numel(), dtype=torch.int32)
        m = m.contiguous().to(device)
        _ROW_MAPS[key] = m                   # never evicted: a captured CUDA graph (engine/graphed.py) keeps reading this buffer by address
    return m


class _NormPartitionResidual(torch.autograd.Function):
    """(window_partition(LN(x)), x): the norm kernel writes each normalised row at its place in shifted-window order and its backward
    reads dy through the same permutation (b2_layernorm_fwd_map / _bwd_map), so neither direction has a shuffle pass; the second
    output is the skip connection (see _NormResidual)."""

    @staticmethod
    def forward(ctx, x, w, b, eps, row_map, ws):
        xc = x.contiguous()
        B, H, W, C = xc.shape
        y, mean, rstd = N.norm_fwd(xc, w, b, eps, False, row_map=row_map)
        ctx.save_for_backward(xc, w, mean, rstd, row_map)
        ctx.params = (w, b)
        ctx.set_materialize_grads(False)
        return y.view(B * (H // ws) * (W // ws), ws * ws, C), xc.view_as(xc)

    @staticmethod
    def backward(ctx, dy, dskip):
        x, w, mean, rstd, row_map = ctx.saved_tensors
        if dy is None:
            return dskip, None, None, None, None, None
        dx, dg, db = N.norm_bwd(dy.contiguous(), x, w, mean, rstd, False, dres=dskip.contiguous() if dskip is not None else None,
                                row_map=row_map)
        return dx, _vgrad(ctx.params[0], dg), _vgrad(ctx.params[1], db), None, None, None


def layer_norm_partition(x, weight, bias, eps: float, ws: int, shift):
    """x [B, H, W, C] (H, W multiples of ws) -> (window_partition(layer_norm(x), ws, shift) as [B*nW, ws*ws, C], x' == x for the skip
    connection).  16-bit inputs take the fused kernel pair; everything else composes the two ops."""
    B, H, W, C = x.shape
    if not _lowp(x) or C % 8:
        y = torch.nn.functional.layer_norm(x, (C,), weight, bias, eps)
        return window_partition(y, ws, shift), x
    return _NormPartitionResidual.apply(x, weight, bias, eps, window_row_map(B, H, W, ws, shift, x.device), ws)


class _ResidualScaleAdd(torch.autograd.Function):
    """out = res + scale[sample] * map(branch); map = identity or Swin's window_reverse (geom = (B, H, W, ws, shift))."""

    @staticmethod
    def forward(ctx, res, branch, scale, geom, reverse):
        B, H, W, ws, shift = geom
        ctx.geom, ctx.reverse, ctx.scale = geom, reverse, scale
        return N.rows_scale_add(branch, res, scale, B, H, W, ws, shift, 2 if reverse else 0)

    @staticmethod
    def backward(ctx, dout):
        B, H, W, ws, shift = ctx.geom
        dbranch = None
        if ctx.needs_input_grad[1]:
            if ctx.scale is None and not ctx.reverse:
                dbranch = dout
            else:
                dbranch = N.rows_scale_add(dout, None, ctx.scale, B, H, W, ws, shift, 1 if ctx.reverse else 0)
        return (dout if ctx.needs_input_grad[0] else None), dbranch, None, None, None


def residual_sd_add(res, branch, p: float = 0.0, training: bool = False, window_geom=None):
    """res + stochastic_depth(branch, p, mode='row') in one pass (torchvision semantics: each sample's branch is dropped with
    probability p and the survivors are scaled by 1 / (1 - p)).  With ``window_geom = (B, H, W, ws, shift)`` the branch is in
    Swin window order [B*nW, ws*ws, C] and is un-partitioned / shifted back on the fly (res and the result are [B, H, W, C])."""
    B = res.shape[0]
    scale = None
    if p > 0.0 and training:
        keep = 1.0 - p
        scale = torch.empty(B, dtype=torch.float32, device=res.device).bernoulli_(keep).div_(keep)
    rows = res.numel() // (B * res.shape[-1])
    geom = window_geom if window_geom is not None else (B, rows, 1, 1, (0, 0))
    if _lowp(res) and res.shape[-1] % 8 == 0:
        return _ResidualScaleAdd.apply(res, branch, scale, geom, window_geom is not None)
    if window_geom is not None:
        branch = window_reverse(branch, *window_geom)
    if scale is not None:
        branch = branch * scale.to(branch.dtype).view((B,) + (1,) * (branch.dim() - 1))
    return res + branch


def layer_norm(x, weight, bias, eps: float = 1e-5):
    if not _lowp(x) or x.shape[-1] % 8:
        return torch.nn.functional.layer_norm(x, (x.shape[-1],), weight, bias, eps)
    return _Norm.apply(x, weight, bias, eps, False)


def rms_norm(x, weight, eps: float = 1e-6):
    if not _lowp(x) or x.shape[-1] % 8:
        xf = x.float()
        return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps) * weight.float()).to(x.dtype)
    return _Norm.apply(x, weight, None, eps, True)


class _BiasAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, b, mode):
        xc = x.contiguous()
        ctx.save_for_backward(xc, b)
        ctx.mode, ctx.bparam = mode, b
        return N.bias_act_fwd(xc, b, mode)

    @staticmethod
    def backward(ctx, dy):
        x, b = ctx.saved_tensors
        dx = N.bias_act_bwd(dy.contiguous(), x, b, ctx.mode)
        gb = _bgrad(ctx.bparam, dx.reshape(-1, dx.shape[-1])) if b is not None else None
        return dx, gb, None


def bias_act(x, bias, mode: int):
    """mode: N.ACT_GELU_TANH / ACT_GELU_ERF / ACT_RELU / ACT_SILU"""
    if not _lowp(x) or x.shape[-1] % 8:
        z = x + bias if bias is not None else x
        F = torch.nn.functional
        return [lambda t: F.gelu(t, approximate="tanh"), F.gelu, F.relu, F.silu][mode](z)
    return _BiasAct.apply(x, bias, mode)


class _SwiGLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gu):
        guc = gu.contiguous()
        ctx.save_for_backward(guc)
        return N.swiglu_fwd(guc)

    @staticmethod
    def backward(ctx, dy):
        (gu,) = ctx.saved_tensors
        return N.swiglu_bwd(dy.contiguous(), gu)


def swiglu(gate_up):
    """silu(gate) * up on a fused [..., 2F] projection (gate first)."""
    if not _lowp(gate_up) or (gate_up.shape[-1] // 2) % 8:
        F_ = gate_up.shape[-1] // 2
        return torch.nn.functional.silu(gate_up[..., :F_]) * gate_up[..., F_:]
    return _SwiGLU.apply(gate_up)


class _Attention(torch.autograd.Function):
    """q, k, v: [S, B, H, D] views (D contiguous) -> o [S, B, H, D].  tcgen05 flash attention fwd/bwd."""

    @staticmethod
    def forward(ctx, q, k, v, scale, causal, p_drop=0.0, seed=0):
        o, lse = N.attn_fwd(q, k, v, scale, causal, p_drop=p_drop, seed=seed)
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.scale, ctx.causal, ctx.drop = scale, causal, (p_drop, seed)
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse = ctx.saved_tensors
        S, B, H, D = q.shape
        dqkv = torch.empty((S, B, H, 3 * D), dtype=q.dtype, device=q.device)
        dq, dk, dv = dqkv[..., :D], dqkv[..., D:2 * D], dqkv[..., 2 * D:]
        N.attn_bwd(q, k, v, o, do.contiguous(), lse, ctx.scale, ctx.causal, dq, dk, dv, p_drop=ctx.drop[0], seed=ctx.drop[1])
        return dq, dk, dv, None, None, None, None


def attention(q, k, v, scale: Optional[float] = None, causal: bool = True, dropout_p: float = 0.0, seed: int = 0):
    """Scaled-dot-product attention on [S, B, H, D] tensors.  head_dim 64 and 128 run the tcgen05 kernels
    (forward and backward) in bf16; other shapes / dtypes use the PyTorch restatement (same math, autograd backward)."""
    D = q.shape[-1]
    scale = 1.0 / math.sqrt(D) if scale is None else scale
    native_ok = q.is_cuda and q.dtype == torch.bfloat16 and D in (64, 128)
    if native_ok or (dropout_p > 0.0 and not q.is_cuda):      # CPU + dropout: the oracle in ops/native.py draws the kernels' mask
        return _Attention.apply(q, k, v, scale, causal, dropout_p, seed)
    qf, kf, vf = (t.permute(1, 2, 0, 3) for t in (q, k, v))
    o = torch.nn.functional.scaled_dot_product_attention(qf, kf, vf, is_causal=causal, scale=scale, dropout_p=dropout_p)
    return o.permute(2, 0, 1, 3).contiguous()


class _CrossEntropy(torch.autograd.Function):
    """logits [rows, V] (low precision) + labels [rows] -> per-row fp32 loss; backward overwrites logits with dlogits."""

    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        lc = logits if logits.is_contiguous() else logits.contiguous()
        m, s, t = N.xent_stats(lc, labels)
        loss = N.xent_finish(m, s, t, labels, ignore_index)
        ctx.save_for_backward(lc, labels, m, s)
        ctx.ignore_index = ignore_index
        return loss

    @staticmethod
    def backward(ctx, dloss):
        lc, labels, m, s = ctx.saved_tensors
        d = N.xent_bwd_(lc, labels, m, s, dloss.contiguous().float(), 1.0, 0, ctx.ignore_index)
        return d, None, None


def cross_entropy(logits, labels, ignore_index: int = -100, reduction: str = "mean"):
    V = logits.shape[-1]
    l2, lab = logits.reshape(-1, V), labels.reshape(-1)
    if not _lowp(logits) or V % 8:
        return torch.nn.functional.cross_entropy(l2.float(), lab, ignore_index=ignore_index, reduction=reduction)
    loss = _CrossEntropy.apply(l2, lab, ignore_index)
    if reduction == "none":
        return loss.view(labels.shape)
    if reduction == "sum":
        return loss.sum()
    return loss.sum() / (lab != ignore_index).sum().clamp(min=1)


class _Embedding(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, weight):
        flat = ids.reshape(-1)
        out = N.embedding_fwd(flat, flat, weight, None, 0)
        ctx.save_for_backward(flat)
        ctx.wparam, ctx.shape = weight, ids.shape
        return out.view(*ids.shape, weight.shape[1])

    @staticmethod
    def backward(ctx, dy):
        (flat,) = ctx.saved_tensors
        w = ctx.wparam
        mg = getattr(w, "main_grad", None)
        acc = mg if mg is not None else torch.zeros(w.shape, dtype=torch.float32, device=w.device)
        N.embedding_bwd_(dy.reshape(-1, w.shape[1]).contiguous(), flat, flat, acc, None, 0)
        return None, (None if mg is not None else acc.to(w.dtype))


def embedding(ids, weight):
    if not _lowp(weight) or weight.shape[1] % 8:
        return torch.nn.functional.embedding(ids, weight)
    return _Embedding.apply(ids, weight)


class _Rope(torch.autograd.Function):
    """rotate-half RoPE over `nheads` heads of width `hd` of x [T, nheads*hd]; out of place (native kernel on a copy)."""

    @staticmethod
    def forward(ctx, x, cos_t, sin_t, pos, nheads, hd):
        ctx.save_for_backward(cos_t, sin_t, pos)
        ctx.meta = (nheads, hd)
        y = x.contiguous().clone()
        N.rope_(y, cos_t, sin_t, pos, nheads, hd, y.stride(0), hd, backward=False)
        return y

    @staticmethod
    def backward(ctx, dy):
        cos_t, sin_t, pos = ctx.saved_tensors
        nheads, hd = ctx.meta
        d = dy.contiguous().clone()
        N.rope_(d, cos_t, sin_t, pos, nheads, hd, d.stride(0), hd, backward=True)
        return d, None, None, None, None, None


def rope(x2, cos_t, sin_t, pos_i32, nheads: int, hd: int):
    """x2: [tokens, nheads*hd] contiguous rows (may be a column slice of a wider tensor)."""
    return _Rope.apply(x2, cos_t, sin_t, pos_i32, nheads, hd)


def rope_tables(max_pos: int, hd: int, base: float = 10000.0, device=None):
    inv = 1.0 / (base ** (torch.arange(0, hd, 2, device=device).float() / hd))
    ang = torch.arange(max_pos, device=device).float()[:, None] * inv[None]
    return ang.cos().contiguous(), ang.sin().contiguous()


class _WindowAttention(torch.autograd.Function):
    """Swin window attention on the fused projection output: qkv [Bw, T, 3C] -> [Bw, T, C]."""

    @staticmethod
    def forward(ctx, qkv, bias, mask, heads, scale):
        qkv = qkv.contiguous()
        bias32 = bias.float().contiguous()
        mask32 = mask.float().contiguous() if mask is not None else None
        out, lse = N.window_attn_fwd(qkv, bias32, mask32, heads, scale)
        ctx.save_for_backward(qkv, bias32, out, lse)
        ctx.mask, ctx.heads, ctx.scale, ctx.bias_dtype = mask32, heads, scale, bias.dtype
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, bias32, out, lse = ctx.saved_tensors
        dqkv, dbias = N.window_attn_bwd(qkv, bias32, ctx.mask, out, dout, lse, ctx.heads, ctx.scale)
        return dqkv, dbias.to(ctx.bias_dtype), None, None, None


def window_attention(qkv, bias, mask, heads: int, scale: float):
    """qkv [Bw, T, 3*C] with the last dim laid out [3][heads][head_dim]; bias [heads, T, T]; mask [nW, T, T] or None.
    7x7 windows with head_dim 32 in bf16/fp16 on the GPU run the fused kernel; everything else the same math in PyTorch
    (autograd)."""
    Bw, T, C3 = qkv.shape
    hd = C3 // 3 // heads
    if T == 49 and hd == 32 and (not qkv.is_cuda or _lowp(qkv)):
        return _WindowAttention.apply(qkv, bias, mask, heads, scale)
    q, k, v = qkv.view(Bw, T, 3, heads, hd).permute(2, 0, 3, 1, 4)
    attn = (q * scale) @ k.transpose(-2, -1) + bias.to(q.dtype)[None]
    if mask is not None:
        nW = mask.shape[0]
        attn = (attn.view(Bw // nW, nW, heads, T, T) + mask.to(attn.dtype)[None, :, None]).view(Bw, heads, T, T)
    attn = torch.softmax(attn.float(), dim=-1).to(v.dtype)
    return (attn @ v).transpose(1, 2).reshape(Bw, T, C3 // 3)


class _WindowShuffle(torch.autograd.Function):
    """cyclic shift + window partition (or its inverse) as ONE gather pass; the two maps are each other's backward."""

    @staticmethod
    def forward(ctx, x, B, H, W, ws, shift, reverse):
        ctx.cfg = (B, H, W, ws, shift, reverse)
        return N.window_shuffle(x, B, H, W, ws, shift, reverse)

    @staticmethod
    def backward(ctx, g):
        B, H, W, ws, shift, reverse = ctx.cfg
        return N.window_shuffle(g, B, H, W, ws, shift, not reverse), None, None, None, None, None, None


def window_partition(x, ws: int, shift=(0, 0)):
    """[B, H, W, C] -> [B * nW, ws*ws, C] after rolling by -shift (H, W multiples of ws)."""
    B, H, W, _ = x.shape
    return _WindowShuffle.apply(x, B, H, W, ws, (int(shift[0]), int(shift[1])), False)


def window_reverse(xw, B: int, H: int, W: int, ws: int, shift=(0, 0)):
    """inverse of window_partition: [B * nW, ws*ws, C] -> [B, H, W, C], rolled back by +shift."""
    return _WindowShuffle.apply(xw, B, H, W, ws, (int(shift[0]), int(shift[1])), True)


# =============================================================================================
# scale + mask + softmax over materialised attention scores (the unfused attention path)
# =============================================================================================
class _ScaledMaskedSoftmax(torch.autograd.Function):
    """x [b, h, sq, sk] -> softmax(scale * x, masked) ; native kernel on CUDA (csrc/softmax.cu), fp32 restatement on CPU."""

    @staticmethod
    def forward(ctx, x, mask, scale, causal):
        y = N.softmax_fwd(x, mask, scale, causal)
        ctx.save_for_backward(y)
        ctx.scale = scale
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return N.softmax_bwd(dy, y, ctx.scale), None, None, None


def scaled_masked_softmax(x, mask=None, scale: float = 1.0):
    """reference: ScaledMaskedSoftmax / ScaledSoftmax (megatron/model/fused_softmax.py); mask True = masked."""
    return _ScaledMaskedSoftmax.apply(x, mask, float(scale), False)


def scaled_upper_triang_masked_softmax(x, scale: float = 1.0):
    """reference: ScaledUpperTriangMaskedSoftmax (causal); x [b, h, sq, sk] (the reference folds b and h into one dim)."""
    return _ScaledMaskedSoftmax.apply(x, None, float(scale), True)


class FusedScaleMaskSoftmax(torch.nn.Module):
    """megatron.model.fused_softmax.FusedScaleMaskSoftmax: picks the fused kernel when it applies, else scale -> mask_func ->
    torch.softmax.  The native kernel has none of the reference's shape limits (16 < sk <= 4096, sq % 4, batches % 4) - only
    sk <= 16384 - so `is_kernel_available` is true for every 16-bit input when fusion is on."""

    def __init__(self, input_in_fp16: bool = False, input_in_bf16: bool = True, attn_mask_type: str = "causal", scaled_masked_softmax_fusion: bool = True,
                 mask_func=None, softmax_in_fp32: bool = True, scale: Optional[float] = None):
        super().__init__()
        assert not (input_in_fp16 and input_in_bf16), "both fp16 and bf16 flags cannot be active at the same time."
        self.input_in_float16 = input_in_fp16 or input_in_bf16
        self.attn_mask_type, self.fusion, self.mask_func, self.softmax_in_fp32, self.scale = attn_mask_type, scaled_masked_softmax_fusion, mask_func, softmax_in_fp32, scale
        assert self.scale is None or softmax_in_fp32, "softmax should be in fp32 when scaled"

    def is_kernel_available(self, mask, b, np_, sq, sk) -> bool:
        return bool(self.fusion and self.input_in_float16 and sk <= 16384)

    def forward(self, x, mask=None):
        assert x.dim() == 4
        scale = self.scale if self.scale is not None else 1.0
        if self.is_kernel_available(mask, *x.shape):
            if self.attn_mask_type == "causal":
                return scaled_upper_triang_masked_softmax(x, scale)
            return scaled_masked_softmax(x, mask, scale)
        xf = x.float() if (self.input_in_float16 and self.softmax_in_fp32) else x
        xf = xf * scale if self.scale is not None else xf
        if self.attn_mask_type == "causal" and mask is None:
            sq, sk = x.shape[-2:]
            mask = ~torch.ones(sq, sk, dtype=torch.bool, device=x.device).tril(diagonal=sk - sq)
        if mask is not None:
            xf = self.mask_func(xf, mask) if self.mask_func is not None else xf.masked_fill(mask, -10000.0)
        p = torch.softmax(xf, dim=-1)
        return p.to(x.dtype) if (self.input_in_float16 and self.softmax_in_fp32) else p


# =============================================================================================
# fp32 MNIST ops (conv + bias + ReLU, max-pool, log-softmax): csrc/conv.cu on CUDA, ATen on CPU
# =============================================================================================
class _ConvBiasReLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, relu):
        y = N.conv2d_fwd(x, w, b, relu)
        ctx.save_for_backward(x, w, y)
        ctx.relu, ctx.has_bias = relu, b is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        dx, dw, db = N.conv2d_bwd(dy, y, x, w, ctx.relu, ctx.needs_input_grad[0], ctx.has_bias)
        return dx, dw, db, None


def conv2d_bias_relu(x, weight, bias=None, relu: bool = True):
    """valid stride-1 convolution + bias + ReLU in one kernel (fp32)"""
    return _ConvBiasReLU.apply(x, weight, bias, relu)


class _MaxPool2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y, idx = N.maxpool2_fwd(x)
        ctx.save_for_backward(idx)
        ctx.in_shape = tuple(x.shape)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        return N.maxpool2_bwd(dy, idx, ctx.in_shape)


def max_pool2x2(x):
    return _MaxPool2.apply(x)


class _LogSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        lp, _ = N.log_softmax_nll_fwd(x, None)
        ctx.save_for_backward(lp)
        return lp

    @staticmethod
    def backward(ctx, d):
        (lp,) = ctx.saved_tensors
        return N.log_softmax_nll_bwd(lp, d.contiguous())


def log_softmax_small(x):
    """log_softmax over dim 1 for <= 32 classes (one warp per row)"""
    return _LogSoftmax.apply(x)


# =============================================================================================
# BatchNorm2d / SyncBatchNorm with optional fused ReLU (csrc/batchnorm.cu)
# =============================================================================================
class _BatchNormTrain(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, eps, momentum, relu, group):
        y, mean, rstd, count = N.bn_train_fwd(x, gamma, beta, running_mean, running_var, eps, momentum, relu, group)
        ctx.save_for_backward(x, y if relu else x, gamma, mean, rstd)
        ctx.meta = (count, relu, group, gamma is not None, beta is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, gamma, mean, rstd = ctx.saved_tensors
        count, relu, group, has_g, has_b = ctx.meta
        dx, dgamma, dbeta = N.bn_bwd(dy, x, y if relu else None, gamma, mean, rstd, count, relu, group)
        return (dx, dgamma.to(gamma.dtype) if has_g else None, dbeta.to(gamma.dtype if has_g else dbeta.dtype) if has_b else None,
                None, None, None, None, None, None)


def batch_norm(x, gamma, beta, running_mean, running_var, training: bool, momentum: float = 0.1, eps: float = 1e-5, relu: bool = False, group=None):
    if training:
        return _BatchNormTrain.apply(x, gamma, beta, running_mean, running_var, float(eps), float(momentum), bool(relu), group)
    return N.bn_eval_fwd(x, gamma, beta, running_mean, running_var, eps, relu)


class BatchNorm2d(torch.nn.BatchNorm2d):
    """Drop-in nn.BatchNorm2d (same parameters / buffers / state_dict) on the native kernels; `process_group` turns it into a
    SyncBatchNorm (statistics over the group); `fuse_relu` applies ReLU in the same pass."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True, process_group=None, fuse_relu=False, **kw):
        super().__init__(num_features, eps, momentum, affine, track_running_stats, **kw)
        self.process_group, self.fuse_relu = process_group, fuse_relu

    def forward(self, x):
        if x.dim() != 4 or (x.is_cuda and not x.is_contiguous()):
            y = super().forward(x)
            return torch.relu(y) if self.fuse_relu else y
        use_batch = self.training or not self.track_running_stats
        if use_batch and self.track_running_stats and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
        mom = self.momentum if self.momentum is not None else 1.0 / max(1, int(self.num_batches_tracked)) if use_batch else 0.0
        return batch_norm(x, self.weight, self.bias, self.running_mean if self.track_running_stats else None,
                          self.running_var if self.track_running_stats else None, use_batch, mom, self.eps, self.fuse_relu,
                          self.process_group if use_batch else None)


def convert_batchnorm(module: torch.nn.Module, process_group=None) -> torch.nn.Module:
    """replace every nn.BatchNorm2d in `module` by the native one, keeping parameters and buffers (the way
    torch.nn.SyncBatchNorm.convert_sync_batchnorm does); with `process_group` the layers synchronise their statistics"""
    out = module
    if isinstance(module, torch.nn.BatchNorm2d) and not isinstance(module, BatchNorm2d):
        out = BatchNorm2d(module.num_features, module.eps, module.momentum, module.affine, module.track_running_stats, process_group)
        if module.affine:
            out.weight, out.bias = module.weight, module.bias
        if module.track_running_stats:
            out.running_mean, out.running_var, out.num_batches_tracked = module.running_mean, module.running_var, module.num_batches_tracked
        out.train(module.training)
    for name, child in list(module.named_children()):
        new = convert_batchnorm(child, process_group)
        if new is not child:
            setattr(out, name, new)
    return out
