"""Stand-alone tensor-parallel layers: ColumnParallelLinear, RowParallelLinear, VocabParallelEmbedding and the
vocab-parallel cross-entropy (role of megatron.core.tensor_parallel.{layers,cross_entropy}).

The flagship GPT layer (models/gpt.py::_LayerFn) schedules these steps by hand inside one autograd node; the classes
here expose the same steps as modules so that other layer shapes (SwiGLU, GQA, RoPE, Switch-MLP - models/gpt_flex.py)
and user models can be assembled from them.  Forward and backward go through `TPComm`, i.e. with the `native` backend
a sequence-parallel ColumnParallelLinear is ONE fused all-gather -> GEMM kernel and a RowParallelLinear ONE
GEMM -> reduce-scatter kernel (csrc/comm.cu); the backward of each is the mirrored fused kernel plus a wgrad GEMM that
accumulates into fp32 `main_grad` (the reference's `LinearWithGradAccumulationAndAsyncCommunication`).

Weights are initialised from a per-parameter seed by drawing the FULL tensor and slicing this rank's shard, so a model
has the same initial state at every TP size (the effect of the reference's --use-cpu-initialization path).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from ..ops import native as N
from ..ops.functional import _bgrad, _wgrad
from . import state as mpu
from .tensor import TPComm


_XAVIER_UNIFORM = False        # --init-method-xavier-uniform (arguments.py:435-437, :870): every weight matrix, output projections included


def set_init_method_xavier_uniform(on: bool) -> None:
    global _XAVIER_UNIFORM
    _XAVIER_UNIFORM = bool(on)


def init_shard(full_shape, std: float, seed: int, device, dtype, shard_dim: Optional[int] = None, rank: int = 0, world: int = 1):
    """Deterministic init that does not depend on the TP size: draw the FULL tensor from a per-parameter seed, then slice this rank's
    shard (the effect of Megatron's --use-cpu-initialization path).  Normal(0, std), or Xavier-uniform over the full [out, in] matrix
    when --init-method-xavier-uniform is set."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    fs = tuple(full_shape)
    if _XAVIER_UNIFORM and len(fs) == 2:
        import math
        bound = math.sqrt(6.0 / (fs[0] + fs[1]))
        full = torch.empty(fs, dtype=torch.float32).uniform_(-bound, bound, generator=g)
    else:
        full = torch.empty(fs, dtype=torch.float32).normal_(0.0, std, generator=g)
    if shard_dim is not None and world > 1:
        full = full.chunk(world, dim=shard_dim)[rank]
    return full.to(device=device, dtype=dtype).contiguous()


def set_tensor_model_parallel_attributes(p: torch.Tensor, is_parallel: bool, dim: int = -1, stride: int = 1):
    """reference: megatron.core.tensor_parallel.layers.set_tensor_model_parallel_attributes"""
    p.tensor_model_parallel, p.partition_dim, p.partition_stride = is_parallel, dim, stride


def default_comm(cfg_sp: bool = False, backend: str = "nccl") -> TPComm:
    return TPComm(mpu.get_tensor_model_parallel_group(), cfg_sp, backend)


class _ColumnFn(torch.autograd.Function):
    """x2 [T_local | T, K] , W [N/t, K] -> y [T, N/t].  SP: the input is all-gathered along tokens (fused with the GEMM on
    the native backend) and kept for the wgrad; dgrad is reduce-scattered (SP) or all-reduced (no SP)."""

    @staticmethod
    def forward(ctx, x2, w, b, comm):
        y, x_full = comm.column_fwd(x2.contiguous(), w, bias=b, epilogue=N.EPI_BIAS if b is not None else N.EPI_NONE)
        ctx.save_for_backward(x_full, w)
        ctx.comm, ctx.params = comm, (w, b)
        return y

    @staticmethod
    def backward(ctx, dy):
        x_full, w = ctx.saved_tensors
        wp, bp = ctx.params
        dy2 = dy.contiguous()
        gb = _bgrad(bp, dy2)
        gw = _wgrad(wp, dy2, x_full)
        dx = ctx.comm.column_dgrad(dy2, w)
        return dx, gw, gb, None


class _RowFn(torch.autograd.Function):
    """a2 [T, K/t], W [N, K/t] -> y [T_local | T, N] = sum over ranks (+ bias + residual), reduce-scatter under SP."""

    @staticmethod
    def forward(ctx, a2, w, b, residual, comm):
        a2 = a2.contiguous()
        y = comm.row_fwd(a2, w, bias=b, residual=residual)
        ctx.save_for_backward(a2, w)
        ctx.comm, ctx.params, ctx.has_res = comm, (w, b), residual is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        a2, w = ctx.saved_tensors
        wp, bp = ctx.params
        dy2 = dy.contiguous()
        da, dy_full = ctx.comm.row_dgrad(dy2, w)
        gw = _wgrad(wp, dy_full, a2)
        gb = _bgrad(bp, dy2)          # SP: local tokens only - the parameter carries `sequence_parallel` for the TP all-reduce
        return da, gw, gb, (dy2 if ctx.has_res else None), None


def column_parallel_linear(x2, weight, bias, comm: TPComm):
    return _ColumnFn.apply(x2, weight, bias, comm)


def row_parallel_linear(a2, weight, bias, comm: TPComm, residual=None):
    return _RowFn.apply(a2, weight, bias, residual, comm)


class ColumnParallelLinear(nn.Module):
    """Y = X A^T + b with A split along its output rows: this rank holds [output_size / t, input_size].
    `gather_output` all-gathers Y along the last dim (off inside transformer layers, where a RowParallelLinear follows)."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True, gather_output: bool = False, init_std: float = 0.02,
                 seed: int = 1234, comm: Optional[TPComm] = None, device=None, dtype=torch.float32, skip_bias_add: bool = False):
        super().__init__()
        self.comm = comm if comm is not None else default_comm()
        t, r = self.comm.size, self.comm.rank
        assert output_size % t == 0, f"output size {output_size} is not divisible by the tensor-parallel size {t}"
        self.input_size, self.output_size, self.gather_output, self.skip_bias_add = input_size, output_size, gather_output, skip_bias_add
        self.weight = nn.Parameter(init_shard((output_size, input_size), init_std, seed, device, dtype, 0, r, t))
        set_tensor_model_parallel_attributes(self.weight, True, 0)
        self.bias = None
        if bias:
            self.bias = nn.Parameter(torch.zeros(output_size // t, device=device, dtype=dtype))
            set_tensor_model_parallel_attributes(self.bias, True, 0)

    def forward(self, x):
        lead = x.shape[:-1]
        y = column_parallel_linear(x.reshape(-1, x.shape[-1]), self.weight, None if self.skip_bias_add else self.bias, self.comm)
        if self.comm.sp:          # tokens were gathered: leading dim grows by t (sequence-major activations [s, b, h])
            lead = (lead[0] * self.comm.size,) + tuple(lead[1:])
        y = y.view(*lead, y.shape[-1])
        if self.gather_output:
            from .tensor import gather_from_tensor_model_parallel_region
            y = gather_from_tensor_model_parallel_region(y, self.comm.group)
        return (y, self.bias) if self.skip_bias_add else y


class RowParallelLinear(nn.Module):
    """Y = X A^T + b with A split along its input columns: this rank holds [output_size, input_size / t]; the partial
    products are summed over the TP group (reduce-scatter along tokens under sequence parallelism)."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True, input_is_parallel: bool = True, init_std: float = 0.02,
                 seed: int = 1234, comm: Optional[TPComm] = None, device=None, dtype=torch.float32, skip_bias_add: bool = False):
        super().__init__()
        self.comm = comm if comm is not None else default_comm()
        t, r = self.comm.size, self.comm.rank
        assert input_size % t == 0, f"input size {input_size} is not divisible by the tensor-parallel size {t}"
        self.input_size, self.output_size, self.input_is_parallel, self.skip_bias_add = input_size, output_size, input_is_parallel, skip_bias_add
        self.weight = nn.Parameter(init_shard((output_size, input_size), init_std, seed, device, dtype, 1, r, t))
        set_tensor_model_parallel_attributes(self.weight, True, 1)
        self.bias = None
        if bias:
            self.bias = nn.Parameter(torch.zeros(output_size, device=device, dtype=dtype))
            set_tensor_model_parallel_attributes(self.bias, False)
            self.bias.sequence_parallel = self.comm.sp

    def forward(self, x, residual=None):
        if not self.input_is_parallel:
            from .tensor import scatter_to_tensor_model_parallel_region
            x = scatter_to_tensor_model_parallel_region(x, self.comm.group)
        lead = x.shape[:-1]
        res2 = residual.reshape(-1, residual.shape[-1]) if residual is not None else None
        y = row_parallel_linear(x.reshape(-1, x.shape[-1]), self.weight, None if self.skip_bias_add else self.bias, self.comm, res2)
        if self.comm.sp:
            lead = (lead[0] // self.comm.size,) + tuple(lead[1:])
        y = y.view(*lead, y.shape[-1])
        return (y, self.bias) if self.skip_bias_add else y


class _VocabEmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, weight, vocab_start, comm, to_sp):
        flat = ids.reshape(-1).contiguous()
        part = N.embedding_fwd(flat, flat, weight, None, vocab_start)          # rows outside this shard's range come back zero
        if comm.size > 1:
            out = comm.reduce_scatter_seq(part) if to_sp else _all_reduce(part, comm)
        else:
            out = part
        ctx.save_for_backward(flat)
        ctx.meta = (weight, vocab_start, comm, to_sp)
        return out

    @staticmethod
    def backward(ctx, dy):
        (flat,) = ctx.saved_tensors
        w, vocab_start, comm, to_sp = ctx.meta
        dy2 = dy.reshape(-1, w.shape[1]).contiguous()
        if to_sp and comm.size > 1:
            dy2 = comm.gather_seq(dy2)
        mg = getattr(w, "main_grad", None)
        acc = mg if mg is not None else torch.zeros(w.shape, dtype=torch.float32, device=w.device)
        N.embedding_bwd_(dy2, flat, flat, acc, None, vocab_start)
        return None, (None if mg is not None else acc.to(w.dtype)), None, None, None


def _all_reduce(x, comm):
    from .tensor import all_reduce
    return all_reduce(x, comm.group)


class VocabParallelEmbedding(nn.Module):
    """Embedding table split along the vocabulary: this rank holds rows [r V/t, (r+1) V/t); tokens outside the range
    contribute zeros and the partial lookups are summed over the TP group (all-reduce, or reduce-scatter along the
    token dim when the consumer is sequence-parallel).  Token ids are taken in the order given (flattened)."""

    def __init__(self, num_embeddings: int, embedding_dim: int, init_std: float = 0.02, seed: int = 1234,
                 comm: Optional[TPComm] = None, device=None, dtype=torch.float32, reduce_scatter_output: Optional[bool] = None):
        super().__init__()
        self.comm = comm if comm is not None else default_comm()
        t, r = self.comm.size, self.comm.rank
        assert num_embeddings % t == 0, "pad the vocabulary to a multiple of the tensor-parallel size (pad_vocab_size)"
        self.num_embeddings, self.embedding_dim = num_embeddings, embedding_dim
        self.vocab_start, self.vocab_end = r * (num_embeddings // t), (r + 1) * (num_embeddings // t)
        self.to_sp = self.comm.sp if reduce_scatter_output is None else bool(reduce_scatter_output)
        self.weight = nn.Parameter(init_shard((num_embeddings, embedding_dim), init_std, seed, device, dtype, 0, r, t))
        set_tensor_model_parallel_attributes(self.weight, True, 0)

    def forward(self, ids):
        out = _VocabEmbedFn.apply(ids, self.weight, self.vocab_start, self.comm, self.to_sp)
        if self.to_sp and self.comm.size > 1:
            return out            # [tokens / t, H]: the caller knows the token layout
        return out.view(*ids.shape, self.embedding_dim)


class _VocabParallelCrossEntropy(torch.autograd.Function):
    """logits [rows, V/t] (this rank's vocabulary shard) + target [rows] (global ids) -> per-row loss (fp32).
    One exchange of (max, sum-exp, target logit[, logit sum]) per row replaces the reference's three all-reduces
    (cross_entropy.py: MAX of the row max, SUM of the predicted logit, SUM of sum-exp)."""

    @staticmethod
    def forward(ctx, logits, target, comm, vocab_start, label_smoothing, ignore_index, global_vocab):
        lc = logits.contiguous()
        lab = target.reshape(-1).contiguous()
        m, s, t = N.xent_stats(lc, lab, vocab_start)
        m, s, t = comm.combine_xent_stats(m, s, t)
        loss = N.xent_finish(m, s, t, lab, ignore_index)
        if label_smoothing > 0.0:
            # loss_ls = (1 - a) * nll - a * mean_v log p_v,  a = ls * V / (V - 1)   (reference cross_entropy.py label_smoothing
            # branch; the mean runs over the GLOBAL vocabulary here - the reference averages each rank's shard only)
            a = label_smoothing * global_vocab / (global_vocab - 1)
            rowsum = lc.float().sum(-1)
            if comm.size > 1:
                rowsum = _all_reduce(rowsum, comm)
            mean_logp = rowsum / global_vocab - (m + torch.log(s))
            keep = (lab != ignore_index)
            loss = torch.where(keep, (1.0 - a) * loss - a * mean_logp, torch.zeros_like(loss))
        ctx.save_for_backward(lc, lab, m, s)
        ctx.meta = (vocab_start, label_smoothing, ignore_index, global_vocab)
        return loss.view(target.shape)

    @staticmethod
    def backward(ctx, dloss):
        lc, lab, m, s = ctx.saved_tensors
        vocab_start, ls, ignore_index, V = ctx.meta
        g = dloss.reshape(-1).contiguous().float()
        d = N.xent_bwd_(lc, lab, m, s, g, 1.0, vocab_start, ignore_index)      # (p - onehot) * g, in place over the logits
        if ls > 0.0:
            a = ls * V / (V - 1)
            keep = lab != ignore_index
            gk = torch.where(keep, g, torch.zeros_like(g))
            local = lab - vocab_start
            inr = keep & (local >= 0) & (local < lc.shape[1])
            rows = torch.nonzero(inr, as_tuple=False).flatten()
            d[rows, local[rows]] += (a * gk[rows]).to(d.dtype)                 # d = p - (1 - a) onehot - a / V
            d.sub_((a / V * gk).to(d.dtype)[:, None])
        return d, None, None, None, None, None, None


def vocab_parallel_cross_entropy(vocab_parallel_logits, target, label_smoothing: float = 0.0, comm: Optional[TPComm] = None,
                                 ignore_index: int = -100):
    """Per-token cross-entropy of vocabulary-sharded logits `[..., V/t]` against global target ids `[...]`.
    NOTE: like the fused LM-head loss, the backward reuses the logits buffer for dlogits."""
    comm = comm if comm is not None else default_comm()
    Vl = vocab_parallel_logits.shape[-1]
    return _VocabParallelCrossEntropy.apply(vocab_parallel_logits.reshape(-1, Vl), target, comm, comm.rank * Vl, float(label_smoothing),
                                            ignore_index, Vl * comm.size)
