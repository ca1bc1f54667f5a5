"""Transformer-layer variants of the Megatron GPT: grouped-query attention, rotary position embedding,
SwiGLU / squared-ReLU / erf-GELU MLPs, LayerNorm with the `1 + gamma` parametrisation, post-LayerNorm residuals and the
Switch MLP (`--num-experts`, top-1 routing, all experts on every rank as in the reference - no expert-parallel groups).

The default GPT-2 shape runs the hand-scheduled single-node layer of models/gpt.py; any of the options above selects
`FlexTransformerLayer`, which is assembled from the tensor-parallel building blocks in parallel/layers.py and the op layer
in ops/functional.py - the same sm_100a kernels (tcgen05 GEMM with fused AG / RS, flash attention, LayerNorm, RoPE, SwiGLU),
one autograd node per step instead of one per layer.  Parameter names follow the Megatron checkpoint layout
(`self_attention.query_key_value`, `self_attention.dense`, `mlp.dense_h_to_4h`, `mlp.dense_4h_to_h`, `mlp.router`,
`mlp.experts.N.*`).  reference flags: megatron/arguments.py:544-616.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn

from ..config.megatron_args import TransformerConfig
from ..ops import functional as Fn
from ..ops import native as N
from ..parallel import tensor as tp_ops
from ..parallel.layers import column_parallel_linear, init_shard, row_parallel_linear, set_tensor_model_parallel_attributes
from ..parallel.tensor import TPComm


def needs_flex_layer(cfg: TransformerConfig) -> bool:
    ng = cfg.num_query_groups
    return bool(cfg.attention_impl == "unfused" or cfg.swiglu or cfg.squared_relu or cfg.position_embedding_type == "rope" or cfg.apply_layernorm_1p
                or cfg.num_experts or cfg.apply_residual_connection_post_layernorm
                or (ng is not None and ng != cfg.num_attention_heads)
                or _activation_name(cfg) != "gelu_tanh")


class _Holder(nn.Module):
    weight: nn.Parameter
    bias: Optional[nn.Parameter]


class _NormNative(torch.autograd.Function):
    """LayerNorm over the last dim through the native kernel (CPU tensors: its fp32 restatement), any dtype."""

    @staticmethod
    def forward(ctx, x2, w, b, eps, wp, bp):
        y, mean, rstd = N.norm_fwd(x2, w, b, eps)
        ctx.save_for_backward(x2, w, mean, rstd)
        ctx.params = (wp, bp)
        return y

    @staticmethod
    def backward(ctx, dy):
        x2, w, mean, rstd = ctx.saved_tensors
        dx, dg, db = N.norm_bwd(dy.contiguous(), x2, w, mean, rstd)
        wp, bp = ctx.params
        return dx, None, None, None, Fn._vgrad(wp, dg), Fn._vgrad(bp, db)


class _BiasDropoutAdd(torch.autograd.Function):
    """y = residual + dropout(x + bias)  (the reference's jit-fused bias_dropout_add); counter-based mask from `seed`."""

    @staticmethod
    def forward(ctx, x, bias, residual, p, seed):
        ctx.p, ctx.seed, ctx.bias = p, seed, bias
        return N.bias_dropout_add(x.contiguous(), bias, residual, p, seed)

    @staticmethod
    def backward(ctx, dy):
        dy = dy.contiguous()
        dx = N.dropout_bwd(dy, ctx.p, ctx.seed) if ctx.p > 0.0 else dy
        return dx, Fn._bgrad(ctx.bias, dx.reshape(-1, dx.shape[-1])), dy, None, None


class _BiasAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bias, mode):
        xc = x.contiguous()
        ctx.save_for_backward(xc)
        ctx.bias, ctx.mode = bias, mode
        return N.bias_act_fwd(xc, bias, mode)

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        dx = N.bias_act_bwd(dy.contiguous(), x, ctx.bias, ctx.mode)
        return dx, Fn._bgrad(ctx.bias, dx.reshape(-1, dx.shape[-1])), None


class _SwiGLUNative(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gu):
        guc = gu.contiguous()
        ctx.save_for_backward(guc)
        return N.swiglu_fwd(guc)

    @staticmethod
    def backward(ctx, dy):
        (gu,) = ctx.saved_tensors
        return N.swiglu_bwd(dy.contiguous(), gu)


class _RopeNative(torch.autograd.Function):
    """rotate-half RoPE applied to `nheads` heads of x2 [T, nheads * hd]; with `rot` < hd only the first `rot` dims of every head
    rotate (--rotary-percent, arguments.py:578) and the rest pass through."""

    @staticmethod
    def forward(ctx, x2, cos_t, sin_t, pos, nheads, hd, rot=None):
        rot = hd if rot is None else rot
        ctx.save_for_backward(cos_t, sin_t, pos)
        ctx.meta = (nheads, hd, rot)
        y = x2.contiguous().clone()
        N.rope_(y, cos_t, sin_t, pos, nheads, rot, y.stride(0), hd, backward=False)
        return y

    @staticmethod
    def backward(ctx, dy):
        cos_t, sin_t, pos = ctx.saved_tensors
        nheads, hd, rot = ctx.meta
        d = dy.contiguous().clone()
        N.rope_(d, cos_t, sin_t, pos, nheads, rot, d.stride(0), hd, backward=True)
        return d, None, None, None, None, None, None


def _activation_name(cfg: TransformerConfig) -> str:
    if cfg.swiglu:
        return "swiglu"
    if cfg.squared_relu:
        return "squared_relu"
    if cfg.openai_gelu or cfg.bias_gelu_fusion:
        return "gelu_tanh"        # openai_gelu and the fused bias-GeLU are the same tanh formula
    return "gelu_erf"            # F.gelu in the reference when the fusion is off


class ParallelMLP(nn.Module):
    """h -> ffn (column parallel) -> activation -> h (row parallel).  SwiGLU keeps gate and up projections in one weight
    `[2 ffn / t, h]` (this rank's gate rows, then its up rows)."""

    def __init__(self, cfg: TransformerConfig, comm: TPComm, seed_base: int, device=None, skip_output_reduce: bool = False):
        super().__init__()
        self.cfg, self.comm, self.act = cfg, comm, _activation_name(cfg)
        self.skip_output_reduce = skip_output_reduce
        t, r = comm.size, comm.rank
        H, F = cfg.hidden_size, cfg.ffn_hidden_size
        assert F % t == 0, "ffn hidden size must be divisible by the tensor-parallel size"
        dt, std = cfg.params_dtype, cfg.init_method_std
        out_std = std / math.sqrt(2.0 * cfg.num_layers)
        self.dense_h_to_4h = _Holder()
        if self.act == "swiglu":
            gate = init_shard((F, H), std, seed_base + 3, device, dt, 0, r, t)
            up = init_shard((F, H), std, seed_base + 6, device, dt, 0, r, t)
            self.dense_h_to_4h.weight = nn.Parameter(torch.cat([gate, up], 0).contiguous())
        else:
            self.dense_h_to_4h.weight = nn.Parameter(init_shard((F, H), std, seed_base + 3, device, dt, 0, r, t))
        set_tensor_model_parallel_attributes(self.dense_h_to_4h.weight, True, 0)
        n1 = self.dense_h_to_4h.weight.shape[0]
        self.dense_h_to_4h.bias = nn.Parameter(torch.zeros(n1, device=device, dtype=dt)) if cfg.add_bias_linear else None
        if self.dense_h_to_4h.bias is not None:
            set_tensor_model_parallel_attributes(self.dense_h_to_4h.bias, True, 0)
        self.dense_4h_to_h = _Holder()
        self.dense_4h_to_h.weight = nn.Parameter(init_shard((H, F), out_std, seed_base + 4, device, dt, 1, r, t))
        set_tensor_model_parallel_attributes(self.dense_4h_to_h.weight, True, 1)
        self.dense_4h_to_h.bias = nn.Parameter(torch.zeros(H, device=device, dtype=dt)) if cfg.add_bias_linear else None
        if self.dense_4h_to_h.bias is not None:
            set_tensor_model_parallel_attributes(self.dense_4h_to_h.bias, False)
            self.dense_4h_to_h.bias.sequence_parallel = comm.sp and not skip_output_reduce

    def hidden(self, x2, comm: TPComm):
        """x2 [T_local | T, H] -> activated [T, ffn / t]"""
        b1 = self.dense_h_to_4h.bias
        if self.act == "swiglu":
            return _SwiGLUNative.apply(column_parallel_linear(x2, self.dense_h_to_4h.weight, b1, comm))
        z = column_parallel_linear(x2, self.dense_h_to_4h.weight, None, comm)
        if self.act == "squared_relu":
            a = _BiasAct.apply(z, b1, N.ACT_RELU)
            return a * a
        return _BiasAct.apply(z, b1, N.ACT_GELU_TANH if self.act == "gelu_tanh" else N.ACT_GELU_ERF)

    def forward(self, x2, residual=None, p_drop: float = 0.0, seed: int = 0):
        """returns residual + dropout(mlp(x2) + bias)  (no residual given: mlp(x2) + bias)"""
        a = self.hidden(x2, self.comm)
        b2 = self.dense_4h_to_h.bias
        if p_drop > 0.0 and residual is not None:
            out = row_parallel_linear(a, self.dense_4h_to_h.weight, None, self.comm)
            return _BiasDropoutAdd.apply(out, b2, residual, p_drop, seed)
        return row_parallel_linear(a, self.dense_4h_to_h.weight, b2, self.comm, residual)


class SwitchMLP(nn.Module):
    """Top-1 routed mixture of `num_experts` ParallelMLPs (reference: megatron/model/transformer.py SwitchMLP).  The router is a
    replicated Linear(h, E); each token goes to its arg-max expert and the expert output is scaled by the winning softmax
    probability.  Under sequence parallelism the tokens are gathered first (the routing decision must be the same on every
    TP rank because each expert is itself tensor-parallel), every expert produces PARTIAL sums for its tokens, and ONE
    reduce-scatter (all-reduce without SP) per layer combines them - not one collective per expert."""

    def __init__(self, cfg: TransformerConfig, comm: TPComm, seed_base: int, device=None):
        super().__init__()
        self.cfg, self.comm = cfg, comm
        E = int(cfg.num_experts)
        dt = cfg.params_dtype
        self.router = _Holder()
        self.router.weight = nn.Parameter(init_shard((E, cfg.hidden_size), cfg.init_method_std, seed_base + 5, device, dt))
        self.router.bias = nn.Parameter(torch.zeros(E, device=device, dtype=dt))
        for p in (self.router.weight, self.router.bias):
            set_tensor_model_parallel_attributes(p, False)
            # the router's gradient flows through `max_prob * partial expert output`, i.e. each TP rank holds a partial sum:
            # mark it for the optimizer's TP all-reduce (the one that serves sequence-parallel LayerNorm parameters)
            p.sequence_parallel = comm.size > 1
        # experts run on gathered tokens with a communicator that does no collectives of its own (partial sums)
        self.experts = nn.ModuleList()
        for e in range(E):
            sub = _ShardOnlyComm(comm)
            self.experts.append(ParallelMLP(cfg, sub, seed_base + 1000 * (e + 1), device, skip_output_reduce=True))

    def forward(self, x2, residual=None, p_drop: float = 0.0, seed: int = 0):
        comm = self.comm
        # backward of this gather / copy sums the partial input gradients of the TP-sharded experts over the group
        xf = tp_ops.gather_from_sequence_parallel_region(x2, comm.group, True) if comm.sp else \
            (tp_ops.copy_to_tensor_model_parallel_region(x2, comm.group) if comm.size > 1 else x2)
        route = torch.nn.functional.linear(xf.float(), self.router.weight.float(), self.router.bias.float())
        probs = torch.softmax(route, dim=-1)
        max_prob, max_ind = probs.max(dim=-1)
        # the bias term is computed in full on every rank, while everything upstream of max_prob is summed over the group
        prob_for_bias = _ScaleGrad.apply(max_prob, 1.0 / comm.size) if comm.size > 1 else max_prob
        out = torch.zeros_like(xf)
        bias_out = None
        for e, expert in enumerate(self.experts):
            idx = torch.nonzero(max_ind == e, as_tuple=False).flatten()
            if idx.numel() == 0:
                continue          # routing is identical on every TP rank, so all ranks skip the same experts
            xe = xf.index_select(0, idx)
            ye = row_parallel_linear(expert.hidden(xe, expert.comm), expert.dense_4h_to_h.weight, None, expert.comm)
            ye = ye * max_prob.index_select(0, idx).to(ye.dtype)[:, None]
            out = out.index_add(0, idx, ye)
            b2 = expert.dense_4h_to_h.bias
            if b2 is not None:
                be = (prob_for_bias.index_select(0, idx).to(b2.dtype)[:, None] * b2[None, :])
                bias_out = torch.zeros_like(xf).index_add(0, idx, be) if bias_out is None else bias_out.index_add(0, idx, be)
        if comm.size > 1:
            out = tp_ops.reduce_scatter_to_sequence_parallel_region(out, comm.group) if comm.sp else \
                tp_ops.reduce_from_tensor_model_parallel_region(out, comm.group)
            if bias_out is not None and comm.sp:
                bias_out = tp_ops.scatter_to_sequence_parallel_region(bias_out, comm.group)
        if bias_out is not None:
            out = out + bias_out           # biases are replicated: added once, after the reduction
        if residual is None:
            return out
        return _BiasDropoutAdd.apply(out, None, residual, p_drop, seed)


class _ScaleGrad(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, s):
        ctx.s = s
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g * ctx.s, None


class _ShardOnlyComm:
    """TPComm stand-in for the experts of a SwitchMLP: keeps the shard geometry (size / rank) of the real communicator
    but runs plain local GEMMs - the SwitchMLP does the token gather before and the reduction after all experts."""

    def __init__(self, comm: TPComm):
        self.size, self.rank, self.sp, self.group, self.backend, self.fabric = comm.size, comm.rank, False, comm.group, "none", None

    def column_fwd(self, x2, W, bias=None, epilogue=N.EPI_NONE, out2=None):
        y = N.gemm(x2, W, bias=bias, epilogue=epilogue if (bias is not None or epilogue != N.EPI_BIAS) else N.EPI_NONE, out2=out2)
        return y, x2

    def row_fwd(self, a2, W, bias=None, residual=None):
        assert bias is None and residual is None
        return N.gemm(a2, W)

    def column_dgrad(self, dy2, W):
        return N.gemm(dy2, W, b_mn=True)

    def row_dgrad(self, dy2, W, epilogue=N.EPI_NONE, aux=None):
        return N.gemm(dy2, W, b_mn=True, epilogue=epilogue, aux=aux), dy2


class FlexTransformerLayer(nn.Module):
    """x [S_local, B, H] -> [S_local, B, H]; pre-LN by default."""

    def __init__(self, cfg: TransformerConfig, layer_number: int, comm: TPComm, device=None, seed: int = 1234):
        super().__init__()
        self.cfg, self.comm, self.layer_number = cfg, comm, layer_number
        t, r = comm.size, comm.rank
        H = cfg.hidden_size
        nh, hd = cfg.num_attention_heads, cfg.kv_channels
        ng = cfg.num_query_groups or nh
        assert nh % ng == 0, "num_attention_heads must be a multiple of num_query_groups"
        assert ng % t == 0 and nh % t == 0, "query groups / heads must be divisible by the tensor-parallel size"
        self.head_dim, self.heads_local, self.groups_local, self.q_per_group = hd, nh // t, ng // t, nh // ng
        self.softmax_scale = 1.0 / math.sqrt(hd)
        dt, std = cfg.params_dtype, cfg.init_method_std
        out_std = std / math.sqrt(2.0 * cfg.num_layers)
        base = seed * 100003 + layer_number * 97
        self.input_layernorm = _LNp(H, device, dt, cfg.apply_layernorm_1p)
        sa = nn.Module()
        self.self_attention = sa
        sa.query_key_value = _Holder()
        # full QKV rows: [group][q heads of the group | k | v][hd]; a TP shard takes a contiguous block of groups (for
        # multi-head attention this is the per-head (q, k, v) interleave of the Megatron checkpoint layout)
        rows = ng * (self.q_per_group + 2) * hd
        sa.query_key_value.weight = nn.Parameter(init_shard((rows, H), std, base + 1, device, dt, 0, r, t))
        sa.query_key_value.bias = nn.Parameter(torch.zeros(rows // t, device=device, dtype=dt)) if cfg.add_bias_linear else None
        sa.dense = _Holder()
        sa.dense.weight = nn.Parameter(init_shard((H, nh * hd), out_std, base + 2, device, dt, 1, r, t))
        sa.dense.bias = nn.Parameter(torch.zeros(H, device=device, dtype=dt)) if cfg.add_bias_linear else None
        self.post_attention_layernorm = _LNp(H, device, dt, cfg.apply_layernorm_1p)      # registration order = the fused layer's
        set_tensor_model_parallel_attributes(sa.query_key_value.weight, True, 0)
        set_tensor_model_parallel_attributes(sa.dense.weight, True, 1)
        if cfg.add_bias_linear:
            set_tensor_model_parallel_attributes(sa.query_key_value.bias, True, 0)
            set_tensor_model_parallel_attributes(sa.dense.bias, False)
            sa.dense.bias.sequence_parallel = comm.sp
        for ln in (self.input_layernorm, self.post_attention_layernorm):
            for p in ln.parameters():
                set_tensor_model_parallel_attributes(p, False)
                p.sequence_parallel = comm.sp
        self.mlp = SwitchMLP(cfg, comm, base, device) if cfg.num_experts else ParallelMLP(cfg, comm, base, device)
        self.rope = cfg.position_embedding_type == "rope"
        self._rope_tables = None
        # --rotary-percent: the first int(head_dim * percent) dims of every head rotate (an even number of them)
        self.rot_dim = self.head_dim
        if self.rope and cfg.rotary_percent < 1.0:
            self.rot_dim = int(self.head_dim * cfg.rotary_percent)
            assert self.rot_dim >= 2 and self.rot_dim % 2 == 0, f"--rotary-percent {cfg.rotary_percent} leaves {self.rot_dim} rotary dims of {self.head_dim}: need an even number >= 2"

    def _tables(self, device, S):
        if self._rope_tables is None or self._rope_tables[0].device != device or self._rope_tables[0].shape[0] < S:
            n = max(S, self.cfg.max_position_embeddings)
            cos_t, sin_t = Fn.rope_tables(n, self.rot_dim, device=device)
            f = self.cfg.rotary_seq_len_interpolation_factor
            if f:
                inv = 1.0 / (10000.0 ** (torch.arange(0, self.rot_dim, 2, device=device).float() / self.rot_dim))
                ang = (torch.arange(n, device=device).float() / float(f))[:, None] * inv[None]
                cos_t, sin_t = ang.cos().contiguous(), ang.sin().contiguous()
            self._rope_tables = (cos_t, sin_t)
        return self._rope_tables

    def forward(self, x):
        cfg, comm = self.cfg, self.comm
        S_loc, B, H = x.shape
        x2 = x.reshape(S_loc * B, H)
        training = self.training
        p_h = cfg.hidden_dropout if training else 0.0
        p_a = cfg.attention_dropout if training else 0.0
        sa = self.self_attention
        hd, nh, ng, nq = self.head_dim, self.heads_local, self.groups_local, self.q_per_group
        # ---- attention half
        ln1 = self.input_layernorm(x2, cfg.layernorm_epsilon)
        qkv = column_parallel_linear(ln1, sa.query_key_value.weight, sa.query_key_value.bias, comm)        # [T, ng (nq + 2) hd]
        T = qkv.shape[0]
        S = T // B
        g = qkv.view(T, ng, (nq + 2) * hd)
        q = g[:, :, :nq * hd].reshape(T, nh * hd)
        k = g[:, :, nq * hd:(nq + 1) * hd].reshape(T, ng * hd)
        v = g[:, :, (nq + 1) * hd:].reshape(T, ng * hd)
        if self.rope:
            cos_t, sin_t = self._tables(x.device, S)
            pos = torch.arange(S, device=x.device, dtype=torch.int32).repeat_interleave(B)               # token t = s * B + b
            q = _RopeNative.apply(q, cos_t, sin_t, pos, nh, hd, self.rot_dim)
            k = _RopeNative.apply(k, cos_t, sin_t, pos, ng, hd, self.rot_dim)
        q4 = q.view(S, B, nh, hd)
        k4, v4 = k.view(S, B, ng, hd), v.view(S, B, ng, hd)
        if nq > 1:                                      # grouped-query: every K/V head serves nq query heads
            k4 = k4.unsqueeze(3).expand(S, B, ng, nq, hd).reshape(S, B, nh, hd)
            v4 = v4.unsqueeze(3).expand(S, B, ng, nq, hd).reshape(S, B, nh, hd)
        if cfg.attention_impl == "unfused":
            ctx4 = core_attention(self, q4, k4, v4, p_a)
        else:
            seed_a = tp_ops.next_dropout_seed(True) if p_a > 0.0 else 0
            ctx4 = Fn.attention(q4.contiguous(), k4.contiguous(), v4.contiguous(), self.softmax_scale, True, p_a, seed_a)
        ctx2 = ctx4.reshape(T, nh * hd)
        res1 = ln1 if cfg.apply_residual_connection_post_layernorm else x2      # both are [T_local, H]
        if p_h > 0.0:
            att = row_parallel_linear(ctx2, sa.dense.weight, None, comm)
            x1 = _BiasDropoutAdd.apply(att, sa.dense.bias, res1, p_h, tp_ops.next_dropout_seed(comm.sp))
        else:
            x1 = row_parallel_linear(ctx2, sa.dense.weight, sa.dense.bias, comm, res1)
        # ---- MLP half
        ln2 = self.post_attention_layernorm(x1, cfg.layernorm_epsilon)
        res2 = ln2 if cfg.apply_residual_connection_post_layernorm else x1
        seed2 = tp_ops.next_dropout_seed(comm.sp) if p_h > 0.0 else 0
        y2 = self.mlp(ln2, res2, p_h, seed2)
        return y2.view(S_loc, B, H)


def core_attention(self, q4, k4, v4, p_drop):
    """megatron CoreAttention (transformer.py): [s, b, h, d] -> scores [b, h, sq, sk] by batched GEMM, scale + causal mask +
    softmax in one kernel (csrc/softmax.cu; torch.softmax when --no-masked-softmax-fusion), dropout under the
    tensor-parallel RNG stream, context by batched GEMM."""
    S, B, nh, hd = q4.shape
    q = q4.permute(1, 2, 0, 3).reshape(B * nh, S, hd)
    k = k4.permute(1, 2, 3, 0).reshape(B * nh, hd, S)
    scores = torch.bmm(q, k).view(B, nh, S, S)
    doc = tp_ops.document_ids()
    if doc is not None:
        # --reset-attention-mask: causal AND same document (reference get_ltor_masks_and_position_ids: tokens after an EOD do not see
        # anything up to it); masked = True, the convention of ScaledMaskedSoftmax
        masked = (doc[:, :, None] != doc[:, None, :]) | torch.ones(S, S, dtype=torch.bool, device=scores.device).triu(1)
        masked = masked[:, None]                                                # [b, 1, sq, sk]
        if self.cfg.masked_softmax_fusion:
            probs = Fn.scaled_masked_softmax(scores, masked, self.softmax_scale)
        else:
            probs = torch.softmax((scores.float() * self.softmax_scale).masked_fill(masked, float("-inf")), dim=-1).to(scores.dtype)
    elif self.cfg.masked_softmax_fusion:
        probs = Fn.scaled_upper_triang_masked_softmax(scores, self.softmax_scale)
    else:
        keep = torch.ones(S, S, dtype=torch.bool, device=scores.device).tril()
        probs = torch.softmax((scores.float() * self.softmax_scale).masked_fill(~keep, float("-inf")), dim=-1).to(scores.dtype)
    if p_drop > 0.0:
        with tp_ops.get_cuda_rng_tracker().fork():
            probs = torch.nn.functional.dropout(probs, p_drop, training=True)
    v = v4.permute(1, 2, 0, 3).reshape(B * nh, S, hd)
    ctx = torch.bmm(probs.view(B * nh, S, S), v).view(B, nh, S, hd)
    return ctx.permute(2, 0, 1, 3).contiguous()



class _LNp(nn.Module):
    """LayerNorm holder; `one_p`: the stored weight is gamma - 1 (initialised to zero), reference apply_layernorm_1p."""

    def __init__(self, H, device, dtype, one_p: bool = False):
        super().__init__()
        self.one_p = one_p
        self.weight = nn.Parameter(torch.zeros(H, device=device, dtype=dtype) if one_p else torch.ones(H, device=device, dtype=dtype))
        self.bias = nn.Parameter(torch.zeros(H, device=device, dtype=dtype))

    def forward(self, x2, eps):
        w = self.weight
        w_eff = (w.detach() + 1) if self.one_p else w.detach()
        return _NormNative.apply(x2.contiguous(), w_eff, self.bias.detach(), eps, self.weight, self.bias)
