"""Megatron-style GPT (decoder-only, pre-LayerNorm) with tensor / sequence / pipeline parallelism.

Role of megatron.model.{gpt_model,language_model,transformer}, built by ``model_provider``
in pretrain_gpt.py:46-58.  Parameter names and shard shapes follow the Megatron checkpoint layout
(`language_model.embedding.word_embeddings.weight` [V/t, h], `...self_attention.query_key_value.weight`
[3h/t, h] with per-head (q,k,v) interleave, `...dense.weight` [h, h/t], `mlp.dense_h_to_4h` [4h/t, h],
`mlp.dense_4h_to_h` [h, 4h/t]) so `ckpt.megatron` can read and write `model_optim_rng.pt` files.

B200-first structure: a transformer layer is ONE autograd node (`_LayerFn`) with a hand-scheduled
forward and backward over the native kernels -
  LN -> [AG+]GEMM(+bias) -> tcgen05 flash attention -> GEMM[+RS](+bias+residual)
  LN -> [AG+]GEMM(+bias, GeLU, dual output) -> GEMM[+RS](+bias+residual)
so every fusion (bias/GeLU/residual epilogues, dGeLU in the dgrad epilogue, wgrad accumulated straight
into fp32 `main_grad` by TMA reduce-add) is explicit, nothing is left to an autograd graph of small ops,
and the saved-tensor set is exactly what backward needs.
"""
from __future__ import annotations

import math
from typing import List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from ..config.megatron_args import TransformerConfig
from ..ops import native as N
from ..parallel import state as mpu
from ..parallel import tensor as tp_ops
from ..parallel.layers import init_shard, set_init_method_xavier_uniform
from ..parallel.tensor import TPComm


def _wgrad(param: torch.Tensor, dy2: torch.Tensor, x2: torch.Tensor):
    """dW += dy2^T @ x2, accumulated in fp32 `main_grad` when the parameter has one (the
    gradient_accumulation_fusion path of SURVEY K18), else returned in the parameter dtype."""
    mg = getattr(param, "main_grad", None)
    if mg is not None:
        N.gemm(dy2, x2, a_mn=True, b_mn=True, out=mg, accumulate=True)
        return None
    out = torch.zeros(param.shape, dtype=torch.float32, device=param.device)
    N.gemm(dy2, x2, a_mn=True, b_mn=True, out=out, accumulate=True)
    return out.to(param.dtype)


def _vgrad(param: Optional[torch.Tensor], g32: torch.Tensor):
    """vector (bias / norm) gradient: add into fp32 main_grad or return in param dtype."""
    if param is None:
        return None
    mg = getattr(param, "main_grad", None)
    if mg is not None:
        mg.add_(g32)
        return None
    return g32.to(param.dtype)


def _norm_bwd(d_out, x, w, mean, rstd, wp, bp, dres):
    """LayerNorm backward with the residual-branch gradient folded in; dgamma/dbeta go straight into the fp32
    main_grad buffers when the parameters have them (no zero-fill / add kernels)."""
    mg_w, mg_b = getattr(wp, "main_grad", None), getattr(bp, "main_grad", None) if bp is not None else None
    if mg_w is not None and mg_b is not None:
        dx, _, _ = N.norm_bwd(d_out, x, w, mean, rstd, dgamma=mg_w, dbeta=mg_b, accumulate=True, dres=dres)
        return dx, None, None
    dx, dg, db = N.norm_bwd(d_out, x, w, mean, rstd, dres=dres)
    return dx, _vgrad(wp, dg), _vgrad(bp, db)


def _bias_grad(param: Optional[torch.Tensor], dy2: torch.Tensor):
    if param is None:
        return None
    mg = getattr(param, "main_grad", None)
    if mg is not None:
        N.colsum_accum(dy2, mg)
        return None
    acc = torch.zeros(param.shape, dtype=torch.float32, device=param.device)
    N.colsum_accum(dy2, acc)
    return acc.to(param.dtype)


def _ag_buf(comm, x2):
    """the fused all-gather->GEMM's input buffer for a LayerNorm output of x2's shape (CUDA + native TP backend), else None"""
    if not x2.is_cuda or not hasattr(comm, "column_input_buffer"):
        return None
    return comm.column_input_buffer(x2.shape[0], x2.shape[1], x2.dtype)


class _LayerFn(torch.autograd.Function):
    """x [S_local, B, H] -> y [S_local, B, H] for one transformer layer."""

    @staticmethod
    def forward(ctx, x, ln1_w, ln1_b, qkv_w, qkv_b, proj_w, proj_b, ln2_w, ln2_b, fc1_w, fc1_b, fc2_w, fc2_b, layer):
        cfg, comm = layer.cfg, layer.comm
        S_loc, B, H = x.shape
        x2 = x.reshape(S_loc * B, H)
        eps = cfg.layernorm_epsilon
        p_h = cfg.hidden_dropout if layer.training else 0.0
        # ---- attention half
        ln1, mean1, rstd1 = N.norm_fwd(x2, ln1_w, ln1_b, eps, out=_ag_buf(comm, x2))
        qkv, ln1_full = comm.column_fwd(ln1, qkv_w, bias=qkv_b, epilogue=N.EPI_BIAS if qkv_b is not None else N.EPI_NONE)
        T = qkv.shape[0]
        S = T // B
        nh, hd = layer.heads_local, layer.head_dim
        qkv4 = qkv.view(S, B, nh, 3 * hd)
        p_a = cfg.attention_dropout if layer.training else 0.0
        seed_a = tp_ops.next_dropout_seed(True) if p_a > 0.0 else 0      # heads are split over TP ranks: per-rank stream
        ctx_t, lse = N.attn_fwd(qkv4[..., :hd], qkv4[..., hd:2 * hd], qkv4[..., 2 * hd:], layer.softmax_scale, True,
                                p_drop=p_a, seed=seed_a)
        ctx2 = ctx_t.view(T, nh * hd)
        seed1 = seed2 = 0
        if p_h > 0.0:
            # under sequence parallelism every TP rank holds a different sequence shard: per-rank stream (Megatron forks the
            # model-parallel RNG for these dropouts), otherwise the stream that is identical across the TP group
            seed1, seed2 = tp_ops.next_dropout_seed(comm.sp), tp_ops.next_dropout_seed(comm.sp)
            attn_out = comm.row_fwd(ctx2, proj_w)
            x1 = N.bias_dropout_add(attn_out, proj_b, x2, p_h, seed1)
        else:
            x1 = comm.row_fwd(ctx2, proj_w, bias=proj_b, residual=x2)
        # ---- MLP half
        ln2, mean2, rstd2 = N.norm_fwd(x1, ln2_w, ln2_b, eps, out=_ag_buf(comm, x1))
        z = torch.empty((T, fc1_w.shape[0]), dtype=x.dtype, device=x.device)
        a, ln2_full = comm.column_fwd(ln2, fc1_w, bias=fc1_b, epilogue=N.EPI_BIAS_GELU_DUAL, out2=z)
        if p_h > 0.0:
            mlp_out = comm.row_fwd(a, fc2_w)
            y2 = N.bias_dropout_add(mlp_out, fc2_b, x1, p_h, seed2)
        else:
            y2 = comm.row_fwd(a, fc2_w, bias=fc2_b, residual=x1)
        ctx.layer, ctx.p_h, ctx.seeds, ctx.dims, ctx.attn_drop = layer, p_h, (seed1, seed2), (S, B, T), (p_a, seed_a)
        ctx.save_for_backward(x2, mean1, rstd1, ln1_full, qkv, ctx_t, lse, x1, mean2, rstd2, ln2_full, z, a,
                              ln1_w, qkv_w, proj_w, ln2_w, fc1_w, fc2_w)
        ctx.params = (ln1_w, ln1_b, qkv_w, qkv_b, proj_w, proj_b, ln2_w, ln2_b, fc1_w, fc1_b, fc2_w, fc2_b)
        return y2.view(S_loc, B, H)

    @staticmethod
    def backward(ctx, dy):
        layer, comm = ctx.layer, ctx.layer.comm
        (x2, mean1, rstd1, ln1_full, qkv, ctx_t, lse, x1, mean2, rstd2, ln2_full, z, a,
         ln1_w, qkv_w, proj_w, ln2_w, fc1_w, fc2_w) = ctx.saved_tensors
        (ln1_wp, ln1_bp, qkv_wp, qkv_bp, proj_wp, proj_bp, ln2_wp, ln2_bp, fc1_wp, fc1_bp, fc2_wp, fc2_bp) = ctx.params
        S, B, T = ctx.dims
        nh, hd = layer.heads_local, layer.head_dim
        H = x2.shape[1]
        dy2 = dy.reshape(-1, H).contiguous()
        p_h = ctx.p_h
        sp_vec = comm.sp   # vector params that see only the local sequence shard need a TP all-reduce later

        # ---- MLP half backward
        d_mlp = N.dropout_bwd(dy2, p_h, ctx.seeds[1]) if p_h > 0.0 else dy2
        g_fc2_b = _bias_grad(fc2_bp, d_mlp)
        dz, d_mlp_full = comm.row_dgrad(d_mlp, fc2_w, epilogue=N.EPI_DGELU, aux=z)   # (dy W2) * gelu'(z)
        g_fc2_w = _wgrad(fc2_wp, d_mlp_full, a)
        g_fc1_b = _bias_grad(fc1_bp, dz)
        g_fc1_w = _wgrad(fc1_wp, dz, ln2_full)
        d_ln2 = comm.column_dgrad(dz, fc1_w)
        dx1, g_ln2_w, g_ln2_b = _norm_bwd(d_ln2, x1, ln2_w, mean2, rstd2, ln2_wp, ln2_bp, dy2)     # dx1 = dy + dLN2

        # ---- attention half backward
        d_att = N.dropout_bwd(dx1, p_h, ctx.seeds[0]) if p_h > 0.0 else dx1
        g_proj_b = _bias_grad(proj_bp, d_att)
        d_ctx, d_att_full = comm.row_dgrad(d_att, proj_w)
        g_proj_w = _wgrad(proj_wp, d_att_full, ctx_t.view(T, nh * hd))
        qkv4 = qkv.view(S, B, nh, 3 * hd)
        dqkv = torch.empty_like(qkv)
        dqkv4 = dqkv.view(S, B, nh, 3 * hd)
        N.attn_bwd(qkv4[..., :hd], qkv4[..., hd:2 * hd], qkv4[..., 2 * hd:], ctx_t, d_ctx.view(S, B, nh, hd), lse,
                   layer.softmax_scale, True, dqkv4[..., :hd], dqkv4[..., hd:2 * hd], dqkv4[..., 2 * hd:],
                   p_drop=ctx.attn_drop[0], seed=ctx.attn_drop[1])
        g_qkv_b = _bias_grad(qkv_bp, dqkv)
        g_qkv_w = _wgrad(qkv_wp, dqkv, ln1_full)
        d_ln1 = comm.column_dgrad(dqkv, qkv_w)
        dx, g_ln1_w, g_ln1_b = _norm_bwd(d_ln1, x2, ln1_w, mean1, rstd1, ln1_wp, ln1_bp, dx1)       # dx = dx1 + dLN1
        return (dx.view(-1, B, H), g_ln1_w, g_ln1_b, g_qkv_w, g_qkv_b, g_proj_w, g_proj_b, g_ln2_w, g_ln2_b,
                g_fc1_w, g_fc1_b, g_fc2_w, g_fc2_b, None)


def _init_normal(shape, std, seed, device, dtype, full_shape=None, shard_dim=None, rank=0, world=1):
    """TP-size independent init of one parameter (parallel.layers.init_shard: full tensor from a per-parameter seed, then this rank's
    shard; normal, or Xavier-uniform under --init-method-xavier-uniform)."""
    return init_shard(full_shape or shape, std, seed, device, dtype, shard_dim, rank, world)


class ParallelTransformerLayer(nn.Module):
    def __init__(self, cfg: TransformerConfig, layer_number: int, comm: TPComm, device=None, seed: int = 1234):
        super().__init__()
        self.cfg, self.comm, self.layer_number = cfg, comm, layer_number
        t, r = comm.size, comm.rank
        H, F = cfg.hidden_size, cfg.ffn_hidden_size
        self.head_dim = cfg.kv_channels
        self.heads_local = cfg.num_attention_heads // t
        self.softmax_scale = 1.0 / math.sqrt(self.head_dim)
        dt, dev = cfg.params_dtype, device
        std = cfg.init_method_std
        out_std = std / math.sqrt(2.0 * cfg.num_layers)
        base = seed * 100003 + layer_number * 97
        nh = cfg.num_attention_heads
        self.input_layernorm = _LN(H, dev, dt)
        self.self_attention = nn.Module()
        self.self_attention.query_key_value = _Lin()
        # full QKV weight rows are ordered [head][q|k|v][hd]; a TP shard takes a contiguous block of heads
        self.self_attention.query_key_value.weight = nn.Parameter(_init_normal(None, std, base + 1, dev, dt, (nh * 3 * self.head_dim, H), 0, r, t))
        self.self_attention.query_key_value.bias = nn.Parameter(torch.zeros(3 * self.heads_local * self.head_dim, device=dev, dtype=dt)) if cfg.add_bias_linear else None
        self.self_attention.dense = _Lin()
        self.self_attention.dense.weight = nn.Parameter(_init_normal(None, out_std, base + 2, dev, dt, (H, nh * self.head_dim), 1, r, t))
        self.self_attention.dense.bias = nn.Parameter(torch.zeros(H, device=dev, dtype=dt)) if cfg.add_bias_linear else None
        self.post_attention_layernorm = _LN(H, dev, dt)
        self.mlp = nn.Module()
        self.mlp.dense_h_to_4h = _Lin()
        self.mlp.dense_h_to_4h.weight = nn.Parameter(_init_normal(None, std, base + 3, dev, dt, (F, H), 0, r, t))
        self.mlp.dense_h_to_4h.bias = nn.Parameter(torch.zeros(F // t, device=dev, dtype=dt)) if cfg.add_bias_linear else None
        self.mlp.dense_4h_to_h = _Lin()
        self.mlp.dense_4h_to_h.weight = nn.Parameter(_init_normal(None, out_std, base + 4, dev, dt, (H, F), 1, r, t))
        self.mlp.dense_4h_to_h.bias = nn.Parameter(torch.zeros(H, device=dev, dtype=dt)) if cfg.add_bias_linear else None
        # tensor-parallel attributes (checkpoint resharding, grad-norm de-duplication, SP grad all-reduce)
        for p, dim in ((self.self_attention.query_key_value.weight, 0), (self.self_attention.query_key_value.bias, 0),
                       (self.self_attention.dense.weight, 1), (self.mlp.dense_h_to_4h.weight, 0), (self.mlp.dense_h_to_4h.bias, 0),
                       (self.mlp.dense_4h_to_h.weight, 1)):
            if p is not None:
                p.tensor_model_parallel, p.partition_dim = True, dim
        for p in (self.input_layernorm.weight, self.input_layernorm.bias, self.post_attention_layernorm.weight,
                  self.post_attention_layernorm.bias, self.self_attention.dense.bias, self.mlp.dense_4h_to_h.bias):
            if p is not None:
                p.tensor_model_parallel = False
                p.sequence_parallel = comm.sp

    def forward(self, x):
        sa, m = self.self_attention, self.mlp
        return _LayerFn.apply(x, self.input_layernorm.weight, self.input_layernorm.bias, sa.query_key_value.weight,
                              sa.query_key_value.bias, sa.dense.weight, sa.dense.bias, self.post_attention_layernorm.weight,
                              self.post_attention_layernorm.bias, m.dense_h_to_4h.weight, m.dense_h_to_4h.bias,
                              m.dense_4h_to_h.weight, m.dense_4h_to_h.bias, self)


class _RecomputeFn(torch.autograd.Function):
    """`--recompute-granularity full` (Megatron's checkpointed transformer layers, tensor_parallel.random.checkpoint): the
    forward runs without recording a graph and saves only the input of a GROUP of layers; the backward re-runs the group
    with the dropout-seed counter (and torch RNG state) of the first run, then back-propagates through the re-run.
    `--distribute-saved-activations`: the saved input is split over the tensor-parallel ranks (each keeps 1/t of it) and
    all-gathered before the re-run.  `selective` needs no code here: the only activations it drops in the reference are
    the attention scores, which the flash-attention kernel never stores."""

    @staticmethod
    def forward(ctx, x, layers, distribute_group):
        ctx.layers = layers
        ctx.doc_ids = tp_ops.document_ids()
        ctx.seed_state = tp_ops.dropout_seed_state()
        ctx.cpu_rng = torch.get_rng_state()
        ctx.cuda_rng = torch.cuda.get_rng_state() if x.is_cuda else None
        ctx.dist_group, ctx.shape = distribute_group, x.shape
        if distribute_group is not None:
            t, r = dist.get_world_size(distribute_group), dist.get_rank(distribute_group)
            flat = x.detach().reshape(-1)
            assert flat.numel() % t == 0, "--distribute-saved-activations: activation size must divide by the tensor-parallel size"
            ctx.save_for_backward(flat.chunk(t)[r].clone())
        else:
            ctx.save_for_backward(x)
        with torch.no_grad():
            h = x
            for layer in layers:
                h = layer(h)
            return h

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        if ctx.dist_group is not None:
            x = tp_ops.all_gather_first_dim(x, ctx.dist_group).view(ctx.shape)
        now_seed, now_cpu = tp_ops.dropout_seed_state(), torch.get_rng_state()
        now_doc = tp_ops.document_ids()
        tp_ops.set_document_ids(ctx.doc_ids)
        now_cuda = torch.cuda.get_rng_state() if x.is_cuda else None
        tp_ops.set_dropout_seed_state(ctx.seed_state)
        torch.set_rng_state(ctx.cpu_rng)
        if ctx.cuda_rng is not None:
            torch.cuda.set_rng_state(ctx.cuda_rng)
        try:
            xin = x.detach().requires_grad_(True)
            with torch.enable_grad():
                y = xin
                for layer in ctx.layers:
                    y = layer(y)
            torch.autograd.backward(y, dy)
        finally:
            tp_ops.set_dropout_seed_state(now_seed)
            tp_ops.set_document_ids(now_doc)
            torch.set_rng_state(now_cpu)
            if now_cuda is not None:
                torch.cuda.set_rng_state(now_cuda)
        return xin.grad, None, None


def recompute_plan(num_layers: int, method: Optional[str], group: int):
    """[(first layer, one past last, recompute?)] for the layers of one model chunk (reference: ParallelTransformer.
    _checkpointed_forward, transformer.py).  `uniform`: consecutive groups of `group` layers, each re-run as a unit from its
    saved input; `block`: the first `group` layers are recomputed one by one, the rest keep their activations; no method
    given: every layer on its own (uniform with groups of one)."""
    group = max(1, int(group or 1))
    if method == "block":
        return [(i, i + 1, i < group) for i in range(num_layers)]
    if method not in (None, "uniform"):
        raise ValueError("Invalid activation recompute method.")
    if method is None:
        group = 1
    return [(i, min(i + group, num_layers), True) for i in range(0, num_layers, group)]


class _Lin(nn.Module):
    """parameter holder (weight, bias) so state_dict keys read `<name>.weight` / `<name>.bias`."""
    weight: nn.Parameter
    bias: Optional[nn.Parameter]


class _LN(nn.Module):
    def __init__(self, H, device, dtype):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(H, device=device, dtype=dtype))
        self.bias = nn.Parameter(torch.zeros(H, device=device, dtype=dtype))


class _NormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, eps):
        y, mean, rstd = N.norm_fwd(x, w, b, eps)
        ctx.save_for_backward(x, w, mean, rstd)
        ctx.params = (w, b)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, mean, rstd = ctx.saved_tensors
        dx, dg, db = N.norm_bwd(dy.contiguous(), x, w, mean, rstd)
        return dx, _vgrad(ctx.params[0], dg), _vgrad(ctx.params[1], db), None


class _EmbeddingFn(torch.autograd.Function):
    """tokens [B, S] -> [S_local, B, H]: vocab-parallel lookup + position embedding, reduce(-scatter)."""

    @staticmethod
    def forward(ctx, tokens, position_ids, wte, wpe, model):
        comm = model.comm
        B, S = tokens.shape
        tok = tokens.t().contiguous().view(-1)            # [S*B], sequence-major like the activations
        pos = position_ids.t().contiguous().view(-1)
        vstart = model.vocab_start
        use_pos = wpe if comm.rank == 0 else None          # added once across the TP group
        part = N.embedding_fwd(tok, pos, wte, use_pos, vstart)
        out = comm.reduce_scatter_seq(part) if comm.size > 1 else part
        ctx.save_for_backward(tok, pos)
        ctx.model, ctx.params, ctx.shape = model, (wte, wpe), (S, B)
        return out.view(-1, B, wte.shape[1])

    @staticmethod
    def backward(ctx, dy):
        tok, pos = ctx.saved_tensors
        model, (wte, wpe), (S, B) = ctx.model, ctx.params, ctx.shape
        comm = model.comm
        H = wte.shape[1]
        dy2 = dy.reshape(-1, H).contiguous()
        dy_full = comm.gather_seq(dy2)
        g_wte = g_wpe = None
        wte_mg, wpe_mg = getattr(wte, "main_grad", None), getattr(wpe, "main_grad", None) if wpe is not None else None
        dwte = wte_mg if wte_mg is not None else torch.zeros(wte.shape, dtype=torch.float32, device=wte.device)
        dwpe = None
        if wpe is not None:
            dwpe = wpe_mg if wpe_mg is not None else torch.zeros(wpe.shape, dtype=torch.float32, device=wpe.device)
        N.embedding_bwd_(dy_full, tok, pos, dwte, dwpe, model.vocab_start)
        if wte_mg is None:
            g_wte = dwte.to(wte.dtype)
        if wpe is not None and wpe_mg is None:
            g_wpe = dwpe.to(wpe.dtype)
        return None, None, g_wte, g_wpe, None


class _LMHeadLossFn(torch.autograd.Function):
    """final hidden [S_local, B, H] + labels [B, S] -> per-token loss [B, S] (fp32).
    logits = gather_seq(h) @ wte_shard^T  (parallel_lm_logits, K19) followed by vocab-parallel softmax
    cross-entropy (K20/C13) whose backward overwrites the logits buffer with dlogits."""

    @staticmethod
    def forward(ctx, h, wte, labels, model):
        comm = model.comm
        S_loc, B, H = h.shape
        h2 = h.reshape(S_loc * B, H)
        logits, h_full = comm.column_fwd(h2, wte)
        T, V = logits.shape
        lab = labels.t().contiguous().view(-1)
        if hasattr(comm, "xent_forward"):
            m, s, loss = comm.xent_forward(logits, lab, model.vocab_start, model.ignore_index)
        else:
            m, s, t = N.xent_stats(logits, lab, model.vocab_start)
            m, s, t = comm.combine_xent_stats(m, s, t)
            loss = N.xent_finish(m, s, t, lab, model.ignore_index)
        ctx.save_for_backward(logits, h_full, lab, m, s, wte)
        ctx.model, ctx.params, ctx.dims = model, (wte,), (T // B, B)
        return loss.view(T // B, B).t().contiguous()

    @staticmethod
    def backward(ctx, dloss):
        logits, h_full, lab, m, s, wte = ctx.saved_tensors
        model, comm = ctx.model, ctx.model.comm
        S, B = ctx.dims
        row_scale = dloss.t().contiguous().view(-1).float()
        dlogits = N.xent_bwd_(logits, lab, m, s, row_scale, 1.0, model.vocab_start, model.ignore_index)
        g_wte = _wgrad(ctx.params[0], dlogits, h_full)
        dh = comm.column_dgrad(dlogits, wte)
        return dh.view(-1, B, wte.shape[1]), g_wte, None, None


class GPTModel(nn.Module):
    """pre_process: owns the embedding; post_process: owns final LN + (tied) output head + loss."""

    def __init__(self, cfg: TransformerConfig, num_tokentypes: int = 0, parallel_output: bool = True, pre_process: bool = True,
                 post_process: bool = True, device=None, seed: int = 1234, layer_offset: Optional[int] = None,
                 num_layers_here: Optional[int] = None, comm: Optional[TPComm] = None):
        super().__init__()
        self.cfg, self.pre_process, self.post_process, self.parallel_output = cfg, pre_process, post_process, parallel_output
        self.comm = comm if comm is not None else TPComm(mpu.get_tensor_model_parallel_group(), cfg.sequence_parallel, cfg.comm_backend)
        t, r = self.comm.size, self.comm.rank
        self.vocab_local = cfg.padded_vocab_size // t
        self.vocab_start = r * self.vocab_local
        self.ignore_index = -100
        self.input_tensor = None
        dt, H = cfg.params_dtype, cfg.hidden_size
        set_init_method_xavier_uniform(getattr(cfg, "init_method_xavier_uniform", False))
        if layer_offset is None or num_layers_here is None:
            layer_offset, num_layers_here = get_layer_range(cfg)
        self.layer_offset, self.num_layers_here = layer_offset, num_layers_here
        lm = nn.Module()
        self.language_model = lm
        if pre_process or (post_process and not cfg.untie_embeddings_and_output_weights):
            lm.embedding = nn.Module()
            lm.embedding.word_embeddings = _Lin()
            lm.embedding.word_embeddings.weight = nn.Parameter(
                _init_normal(None, cfg.init_method_std, seed * 100003 + 7, device, dt, (cfg.padded_vocab_size, H), 0, r, t))
            lm.embedding.word_embeddings.weight.tensor_model_parallel = True
            lm.embedding.word_embeddings.weight.partition_dim = 0
            lm.embedding.word_embeddings.weight.shared_embedding = not cfg.untie_embeddings_and_output_weights
        if pre_process and cfg.position_embedding_type == "learned_absolute":
            lm.embedding.position_embeddings = _Lin()
            lm.embedding.position_embeddings.weight = nn.Parameter(
                _init_normal((cfg.max_position_embeddings, H), cfg.init_method_std, seed * 100003 + 11, device, dt))
            lm.embedding.position_embeddings.weight.tensor_model_parallel = False
        lm.encoder = nn.Module()
        from .gpt_flex import FlexTransformerLayer, needs_flex_layer
        # GPT-2 shaped layers run the hand-scheduled single-node layer; GQA / RoPE / SwiGLU / Switch-MLP ... the modular one
        layer_cls = FlexTransformerLayer if (needs_flex_layer(cfg) or getattr(cfg, "force_flex_layer", False)) else ParallelTransformerLayer
        lm.encoder.layers = nn.ModuleList(
            [layer_cls(cfg, layer_offset + i + 1, self.comm, device, seed) for i in range(num_layers_here)])
        if post_process:
            lm.encoder.final_layernorm = _LN(H, device, dt)
            for p in lm.encoder.final_layernorm.parameters():
                p.tensor_model_parallel, p.sequence_parallel = False, self.comm.sp
            if cfg.untie_embeddings_and_output_weights:
                lm.output_layer = _Lin()
                lm.output_layer.weight = nn.Parameter(
                    _init_normal(None, cfg.init_method_std, seed * 100003 + 13, device, dt, (cfg.padded_vocab_size, H), 0, r, t))
                lm.output_layer.weight.tensor_model_parallel, lm.output_layer.weight.partition_dim = True, 0

    # ---- pipeline plumbing (megatron module API)
    def set_input_tensor(self, t):
        self.input_tensor = t

    def shared_embedding_or_output_weight(self):
        if self.cfg.untie_embeddings_and_output_weights and self.post_process:
            return self.language_model.output_layer.weight
        return self.language_model.embedding.word_embeddings.weight

    def word_embeddings_weight(self):
        return self.language_model.embedding.word_embeddings.weight

    def forward(self, tokens, position_ids=None, attention_mask=None, labels=None):
        """attention_mask is accepted for API parity and ignored: causality is implicit in the kernel
        (SURVEY K23).  Returns per-token loss [B, S] on the last stage when labels are given, logits
        [S, B, V/t] when they are not, and hidden states [S_local, B, H] on other stages."""
        lm = self.language_model
        # --reset-attention-mask: [b, s] document ids (entrypoints/pretrain_gpt.get_batch) for the unfused attention; anything else
        # (None, or a reference-style [b, 1, s, s] mask) means plain causal attention
        tp_ops.set_document_ids(attention_mask if (torch.is_tensor(attention_mask) and attention_mask.dim() == 2) else None)
        if self.pre_process:
            if position_ids is None:
                position_ids = torch.arange(tokens.shape[1], device=tokens.device).unsqueeze(0).expand_as(tokens)
            wpe = lm.embedding.position_embeddings.weight if hasattr(lm.embedding, "position_embeddings") else None
            h = _EmbeddingFn.apply(tokens, position_ids, lm.embedding.word_embeddings.weight, wpe, self)
            p = self.cfg.hidden_dropout if self.training else 0.0
            if p > 0.0:
                h = _DropoutFn.apply(h, p, tp_ops.next_dropout_seed(self.comm.sp))
        else:
            h = self.input_tensor
        if self.cfg.recompute_granularity == "full" and self.training:
            dgroup = None
            if self.cfg.distribute_saved_activations and self.comm.size > 1 and not self.comm.sp:
                dgroup = self.comm.group          # under sequence parallelism the activation is already 1/t per rank
            layers = list(lm.encoder.layers)
            for a, b, re in recompute_plan(len(layers), self.cfg.recompute_method, self.cfg.recompute_num_layers):
                if re:
                    h = _RecomputeFn.apply(h, layers[a:b], dgroup)      # keep only the group's input; re-run in backward
                else:
                    for layer in layers[a:b]:
                        h = layer(h)
        else:
            for layer in lm.encoder.layers:
                h = layer(h)
        if not self.post_process:
            return h
        ln = lm.encoder.final_layernorm
        S_loc, B, H = h.shape
        hn = _NormFn.apply(h.reshape(S_loc * B, H), ln.weight, ln.bias, self.cfg.layernorm_epsilon).view(S_loc, B, H)
        w = self.shared_embedding_or_output_weight()
        if labels is None:
            logits, _ = self.comm.column_fwd(hn.reshape(S_loc * B, H), w)
            return logits.view(-1, B, logits.shape[-1])
        return _LMHeadLossFn.apply(hn, w, labels, self)


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, seed):
        ctx.p, ctx.seed = p, seed
        return N.bias_dropout_add(x.contiguous(), None, None, p, seed)

    @staticmethod
    def backward(ctx, dy):
        return N.dropout_bwd(dy.contiguous(), ctx.p, ctx.seed), None, None


def get_layer_range(cfg: TransformerConfig):
    """(offset, count) of the transformer layers owned by this (pipeline, virtual-pipeline) position
    (megatron.model.transformer._get_num_layers + the interleaved offset rule)."""
    pp = mpu.get_pipeline_model_parallel_world_size()
    pr = mpu.get_pipeline_model_parallel_rank()
    vpp = mpu.get_virtual_pipeline_model_parallel_world_size()
    if getattr(cfg, "standalone_embedding_stage", False) and pp > 1:
        # --standalone-embedding-stage (arguments.py:1054): the first pipeline stage holds the input embedding only, the layers are
        # divided over the remaining pp - 1 stages
        assert vpp is None, "--standalone-embedding-stage cannot be combined with virtual pipeline stages"
        assert cfg.num_layers % (pp - 1) == 0, f"num_layers ({cfg.num_layers}) must be divisible by the {pp - 1} transformer pipeline stages"
        per_stage = cfg.num_layers // (pp - 1)
        return (0, 0) if pr == 0 else ((pr - 1) * per_stage, per_stage)
    per_stage = cfg.num_layers // pp
    if vpp is not None:
        per_chunk = per_stage // vpp
        vr = mpu.get_virtual_pipeline_model_parallel_rank()
        return vr * (cfg.num_layers // vpp) + pr * per_chunk, per_chunk
    return pr * per_stage, per_stage


def count_parameters(model: nn.Module) -> int:
    return sum(p.numel() for p in model.parameters())
