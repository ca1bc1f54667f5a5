"""Tensor / sequence-parallel primitives.

Counterparts of megatron.core.tensor_parallel:
  * ``TPComm`` - the four GEMM+collective compositions the transformer layer needs
      column_fwd      : [all-gather over sequence ->] GEMM               (ColumnParallelLinear forward)
      row_fwd         : GEMM -> reduce-scatter (SP) | all-reduce          (RowParallelLinear forward)
      column_dgrad    : GEMM -> reduce-scatter (SP) | all-reduce          (ColumnParallelLinear backward)
      row_dgrad       : [all-gather ->] GEMM                             (RowParallelLinear backward)
    With ``backend="native"`` on CUDA each is ONE fused kernel that moves tiles over NVLink from inside
    the GEMM (parallel.fabric: ag_gemm / gemm_rs); ``backend="nccl"`` is the reference's semantics
    (NCCL collective, then / before a separate GEMM) and is what the gloo CPU tier runs.
  * autograd-paired region mappings (copy/reduce/scatter/gather and the three sequence-parallel ones),
  * ``broadcast_data`` (pretrain_gpt.py:75), the model-parallel RNG tracker, vocab range helpers.
"""
from __future__ import annotations

import contextlib
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from ..ops import native as N
from . import state as mpu


# ------------------------------------------------------------------------------------------
# raw collectives on the TP group (work for gloo + nccl)
# ------------------------------------------------------------------------------------------
def _ws(group) -> int:
    return dist.get_world_size(group) if (group is not None and dist.is_initialized()) else 1


def all_gather_first_dim(x: torch.Tensor, group) -> torch.Tensor:
    t = _ws(group)
    if t == 1:
        return x
    out = torch.empty((x.shape[0] * t,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x.contiguous(), group=group)
    return out


def reduce_scatter_first_dim(x: torch.Tensor, group) -> torch.Tensor:
    t = _ws(group)
    if t == 1:
        return x
    assert x.shape[0] % t == 0
    out = torch.empty((x.shape[0] // t,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    if x.is_cuda:
        dist.reduce_scatter_tensor(out, x.contiguous(), group=group)
    else:  # gloo has no reduce_scatter: all-reduce + slice (CPU tier only)
        y = x.contiguous().clone()
        dist.all_reduce(y, group=group)
        r = dist.get_rank(group)
        out.copy_(y[r * out.shape[0]:(r + 1) * out.shape[0]])
    return out


def all_reduce(x: torch.Tensor, group, op=dist.ReduceOp.SUM) -> torch.Tensor:
    if _ws(group) > 1:
        dist.all_reduce(x, op=op, group=group)
    return x


# ------------------------------------------------------------------------------------------
# GEMM + collective compositions
# ------------------------------------------------------------------------------------------
class TPComm:
    def __init__(self, group=None, sequence_parallel: bool = False, backend: str = "nccl"):
        self.group = group
        self.size = _ws(group)
        self.rank = dist.get_rank(group) if self.size > 1 else 0
        self.sp = bool(sequence_parallel and self.size > 1)
        self.backend = backend if self.size > 1 else "none"
        self.fabric = None
        if self.backend == "native":
            from . import fabric
            self.fabric = fabric.get_fabric(group)
            if not self.fabric.usable:
                self.backend, self.fabric = "nccl", None

    def column_input_buffer(self, rows: int, K: int, dtype) -> Optional[torch.Tensor]:
        """[rows, K] buffer a producer may write the NEXT column_fwd / row_dgrad input into (the symmetric shard buffer of
        the fused all-gather->GEMM: saves its staging copy), or None when there is none.  Only valid until that call."""
        if self.sp and self.fabric is not None and rows % 128 == 0:
            return self.fabric.shard_buffer(rows, K, dtype)
        return None

    # x: [T_local, K] (SP) or [T, K];  W: [N_local, K]  ->  (y [T, N_local], x_full [T, K])
    def column_fwd(self, x2, W, bias=None, epilogue=N.EPI_NONE, out2=None):
        if self.sp:
            if self.fabric is not None:
                return self.fabric.ag_gemm(x2, W, bias=bias, epilogue=epilogue, out2=out2)
            x_full = all_gather_first_dim(x2, self.group)
        else:
            x_full = x2
        y = N.gemm(x_full, W, bias=bias, epilogue=epilogue if (bias is not None or epilogue != N.EPI_BIAS) else N.EPI_NONE, out2=out2)
        return y, x_full

    # a: [T, K_local]; W: [N, K_local]; residual: [T_local, N] -> y = residual + (sum_ranks a W^T) + bias
    def row_fwd(self, a2, W, bias=None, residual=None):
        if self.size == 1:
            if residual is not None:
                return N.gemm(a2, W, bias=bias, aux=residual, epilogue=N.EPI_BIAS_RESIDUAL)
            return N.gemm(a2, W, bias=bias, epilogue=N.EPI_BIAS if bias is not None else N.EPI_NONE)
        if self.fabric is not None and self.sp:
            return self.fabric.gemm_rs(a2, W, bias=bias, residual=residual)
        if self.fabric is not None and a2.shape[0] % (128 * self.size) == 0:       # no SP: GEMM -> all-reduce in one kernel
            return self.fabric.gemm_ar(a2, W, bias=bias, residual=residual)
        part = N.gemm(a2, W)
        red = reduce_scatter_first_dim(part, self.group) if self.sp else all_reduce(part, self.group)
        if bias is None and residual is None:
            return red
        return N.bias_dropout_add(red, bias, residual)

    # dy: [T, N_local]; W: [N_local, K] -> dx [T_local, K] (SP) or [T, K] (all-reduced)
    def column_dgrad(self, dy2, W):
        if self.size == 1:
            return N.gemm(dy2, W, b_mn=True)
        if self.fabric is not None and self.sp:
            return self.fabric.gemm_rs(dy2, W, b_mn=True)
        if self.fabric is not None and dy2.shape[0] % (128 * self.size) == 0:
            return self.fabric.gemm_ar(dy2, W, b_mn=True)
        part = N.gemm(dy2, W, b_mn=True)
        return reduce_scatter_first_dim(part, self.group) if self.sp else all_reduce(part, self.group)

    # dy: [T_local, N] (SP) or [T, N]; W: [N, K_local] -> (da [T, K_local], dy_full [T, N])
    def row_dgrad(self, dy2, W, epilogue=N.EPI_NONE, aux=None):
        if self.sp:
            if self.fabric is not None:
                return self.fabric.ag_gemm(dy2, W, b_mn=True, epilogue=epilogue, aux=aux)
            dy_full = all_gather_first_dim(dy2, self.group)
        else:
            dy_full = dy2
        da = N.gemm(dy_full, W, b_mn=True, epilogue=epilogue, aux=aux)
        return da, dy_full

    def gather_seq(self, x):
        if not self.sp:
            return x
        if self.fabric is not None and (x.numel() * x.element_size()) % 16 == 0:
            return self.fabric.all_gather_rows(x)
        return all_gather_first_dim(x, self.group)

    def reduce_scatter_seq(self, x):
        if self.sp and self.fabric is not None and (x.numel() * x.element_size()) % (16 * self.size) == 0:
            return self.fabric.reduce_scatter_rows(x)
        return reduce_scatter_first_dim(x, self.group) if self.sp else all_reduce(x, self.group)

    def xent_forward(self, logits2, labels, vocab_start: int, ignore_index: int = -100):
        """vocab-parallel cross-entropy forward -> (global row max, global sum-exp, loss).  Native backend on GPUs: ONE kernel
        (shard statistics, exchange over NVLink, loss); otherwise statistics kernel + collective + finish."""
        if self.size > 1 and self.fabric is not None and logits2.is_cuda:
            return self.fabric.xent_vocab_parallel(logits2, labels, vocab_start, ignore_index)
        m, s, t = N.xent_stats(logits2, labels, vocab_start)
        m, s, t = self.combine_xent_stats(m, s, t)
        return m, s, N.xent_finish(m, s, t, labels, ignore_index)

    def combine_xent_stats(self, m, s, t):
        """per-rank (row max, sum-exp relative to it, target logit) of a vocab shard -> global (max, sum-exp, target).
        One exchange of 3 scalars per row instead of Megatron's three all-reduces over [s, b]."""
        if self.size == 1:
            return m, s, t
        if self.fabric is not None and (m.numel() * 4) % 16 == 0:
            allst = self.fabric.all_gather_rows(torch.stack([m, s, t]).unsqueeze(0)).view(self.size, 3, -1)
            gm = allst[:, 0].max(0).values
            gs = (allst[:, 1] * torch.exp(allst[:, 0] - gm)).sum(0)
            return gm, gs, allst[:, 2].sum(0)
        gm = m.clone()
        all_reduce(gm, self.group, op=dist.ReduceOp.MAX)
        st = torch.stack([s * torch.exp(m - gm), t])
        all_reduce(st, self.group)
        return gm, st[0].contiguous(), st[1].contiguous()


# ------------------------------------------------------------------------------------------
# autograd-paired region mappings (megatron.core.tensor_parallel.mappings)
# ------------------------------------------------------------------------------------------
def _split_last(x, group):
    t, r = _ws(group), (dist.get_rank(group) if _ws(group) > 1 else 0)
    return x.chunk(t, dim=-1)[r].contiguous() if t > 1 else x


def _gather_last(x, group):
    t = _ws(group)
    if t == 1:
        return x
    parts = [torch.empty_like(x) for _ in range(t)]
    dist.all_gather(parts, x.contiguous(), group=group)
    return torch.cat(parts, dim=-1)


def _split_first(x, group):
    t, r = _ws(group), (dist.get_rank(group) if _ws(group) > 1 else 0)
    return x.chunk(t, dim=0)[r].contiguous() if t > 1 else x


def _mk(fwd, bwd, name):
    class _Fn(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x, group):
            ctx.group = group
            return fwd(x, group)

        @staticmethod
        def backward(ctx, g):
            return bwd(g, ctx.group), None
    _Fn.__name__ = name
    return _Fn


_ident = lambda x, g: x
_allreduce = lambda x, g: all_reduce(x.contiguous().clone(), g)
_CopyTo = _mk(_ident, _allreduce, "CopyToModelParallelRegion")
_ReduceFrom = _mk(_allreduce, _ident, "ReduceFromModelParallelRegion")
_ScatterTo = _mk(_split_last, _gather_last, "ScatterToModelParallelRegion")
_GatherFrom = _mk(_gather_last, _split_last, "GatherFromModelParallelRegion")
_ScatterToSP = _mk(_split_first, all_gather_first_dim, "ScatterToSequenceParallelRegion")
_GatherFromSP = _mk(all_gather_first_dim, reduce_scatter_first_dim, "GatherFromSequenceParallelRegion")
_GatherFromSPNoRS = _mk(all_gather_first_dim, _split_first, "GatherFromSequenceParallelRegionToAll")
_ReduceScatterToSP = _mk(reduce_scatter_first_dim, all_gather_first_dim, "ReduceScatterToSequenceParallelRegion")


def _grp(group):
    return group if group is not None else mpu.get_tensor_model_parallel_group()


def copy_to_tensor_model_parallel_region(x, group=None): return _CopyTo.apply(x, _grp(group))
def reduce_from_tensor_model_parallel_region(x, group=None): return _ReduceFrom.apply(x, _grp(group))
def scatter_to_tensor_model_parallel_region(x, group=None): return _ScatterTo.apply(x, _grp(group))
def gather_from_tensor_model_parallel_region(x, group=None): return _GatherFrom.apply(x, _grp(group))
def scatter_to_sequence_parallel_region(x, group=None): return _ScatterToSP.apply(x, _grp(group))
def reduce_scatter_to_sequence_parallel_region(x, group=None): return _ReduceScatterToSP.apply(x, _grp(group))


def gather_from_sequence_parallel_region(x, group=None, tensor_parallel_output_grad: bool = True):
    return (_GatherFromSP if tensor_parallel_output_grad else _GatherFromSPNoRS).apply(x, _grp(group))


# ------------------------------------------------------------------------------------------
# helpers
# ------------------------------------------------------------------------------------------
def vocab_range(global_vocab: int, rank: int, world: int) -> Tuple[int, int]:
    per = global_vocab // world
    return rank * per, (rank + 1) * per


def divide(a: int, b: int) -> int:
    assert a % b == 0, f"{a} is not divisible by {b}"
    return a // b


def broadcast_data(keys: List[str], data: Optional[Dict[str, torch.Tensor]], datatype: torch.dtype, group=None, device=None):
    """TP-rank 0 broadcasts a dict of same-dtype tensors to its TP group: key sizes first, then one
    flattened tensor (megatron.core.tensor_parallel.data.broadcast_data)."""
    group = _grp(group)
    t = _ws(group)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    if t == 1:
        return {k: data[k].to(device) for k in keys}
    src = dist.get_global_rank(group, 0)
    is_src = dist.get_rank() == src
    max_dim = 5
    sizes = torch.zeros(len(keys) * max_dim, dtype=torch.long, device=device)
    if is_src:
        for i, k in enumerate(keys):
            assert data[k].dim() < max_dim
            for d, s in enumerate(data[k].shape):
                sizes[i * max_dim + d] = s
    dist.broadcast(sizes, src, group=group)
    sizes_cpu = sizes.cpu().tolist()
    shapes, total = {}, 0
    for i, k in enumerate(keys):
        shp = [s for s in sizes_cpu[i * max_dim:(i + 1) * max_dim] if s > 0]
        shapes[k] = shp
        n = 1
        for s in shp:
            n *= s
        total += n
    if is_src:
        flat = torch.cat([data[k].to(device=device, dtype=datatype).reshape(-1) for k in keys])
    else:
        flat = torch.empty(total, dtype=datatype, device=device)
    dist.broadcast(flat, src, group=group)
    out, off = {}, 0
    for k in keys:
        n = 1
        for s in shapes[k]:
            n *= s
        out[k] = flat[off:off + n].view(shapes[k])
        off += n
    return out


class RNGStatesTracker:
    """Named RNG streams (megatron.core.tensor_parallel.random.CudaRNGStatesTracker):
    the default stream is identical across a TP group (dropout in sequence-replicated regions), the
    ``model-parallel-rng`` stream is seeded with seed + 2718 + tp_rank (dropout inside TP regions).
    Works on CPU (torch generator state) and CUDA (device generator state)."""

    def __init__(self):
        self.states: Dict[str, torch.Tensor] = {}
        self.seeds = set()

    def reset(self):
        self.states, self.seeds = {}, set()

    def _get(self):
        return torch.cuda.get_rng_state() if torch.cuda.is_available() else torch.get_rng_state()

    def _set(self, s):
        if torch.cuda.is_available():
            torch.cuda.set_rng_state(s)
        else:
            torch.set_rng_state(s)

    def add(self, name: str, seed: int):
        if seed in self.seeds:
            raise RuntimeError(f"seed {seed} already exists")
        if name in self.states:
            raise RuntimeError(f"rng state {name} already exists")
        self.seeds.add(seed)
        orig = self._get()
        if torch.cuda.is_available():
            torch.cuda.manual_seed(seed)
        else:
            torch.manual_seed(seed)
        self.states[name] = self._get()
        self._set(orig)

    def get_states(self):
        """tracker streams + the counter behind next_dropout_seed (so a resumed run draws the same masks)."""
        st = {k: v.clone() for k, v in self.states.items()}
        st["__dropout_seed_counter__"] = torch.tensor([_SEED_COUNTER["base"], _SEED_COUNTER["calls"]], dtype=torch.int64)
        return st

    def set_states(self, states):
        states = dict(states)
        ctr = states.pop("__dropout_seed_counter__", None)
        if ctr is not None:
            _SEED_COUNTER["base"], _SEED_COUNTER["calls"] = int(ctr[0]), int(ctr[1])
        self.states = states

    @contextlib.contextmanager
    def fork(self, name: str = "model-parallel-rng"):
        if name not in self.states:
            raise RuntimeError(f"rng state {name} is not added")
        orig = self._get()
        self._set(self.states[name])
        try:
            yield
        finally:
            self.states[name] = self._get()
            self._set(orig)


_RNG_TRACKER = RNGStatesTracker()
_SEED_COUNTER = {"base": 1234, "calls": 0}


def get_cuda_rng_tracker() -> RNGStatesTracker:
    return _RNG_TRACKER


def model_parallel_cuda_manual_seed(seed: int):
    """default RNG = seed (same across TP), tracker stream = seed + 2718 + tp_rank."""
    offset = seed + 2718
    tp_seed = offset + mpu.get_tensor_model_parallel_rank()
    _RNG_TRACKER.reset()
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed(seed)
    _RNG_TRACKER.add("model-parallel-rng", tp_seed)
    _SEED_COUNTER["base"], _SEED_COUNTER["calls"] = seed, 0


# --reset-attention-mask: document id of every token of the micro-batch in flight ([b, s] int64, or None).  The unfused attention
# (models/gpt_flex.core_attention) masks pairs from different documents; GPTModel.forward installs it, _RecomputeFn re-installs the
# one of its own forward when it re-runs layers in the backward.
_DOC_IDS = None


def set_document_ids(doc_ids):
    global _DOC_IDS
    _DOC_IDS = doc_ids


def document_ids():
    return _DOC_IDS


def dropout_seed_state():
    """(base, calls) of the counter behind next_dropout_seed - saved by activation recomputation so the re-run of a
    layer draws the masks of its first run."""
    return _SEED_COUNTER["base"], _SEED_COUNTER["calls"]


def set_dropout_seed_state(state) -> None:
    _SEED_COUNTER["base"], _SEED_COUNTER["calls"] = int(state[0]), int(state[1])


def next_dropout_seed(tensor_parallel_region: bool) -> int:
    """Counter-based seeds for the Philox dropout kernels: deterministic per call site order; inside TP
    regions the seed differs per TP rank (the tracker's contract), outside it is shared."""
    _SEED_COUNTER["calls"] += 1
    s = (_SEED_COUNTER["base"] * 1000003 + _SEED_COUNTER["calls"]) & 0x7FFFFFFFFFFF
    if tensor_parallel_region:
        s = s * 31 + 2718 + mpu.get_tensor_model_parallel_rank()
    return s


# ------------------------------------------------------------------------------------------
# activation checkpointing with RNG restore (megatron.core.tensor_parallel.random.checkpoint)
# ------------------------------------------------------------------------------------------
class _CheckpointFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, run_function, distribute_saved_activations, *args):
        ctx.run_function, ctx.distribute = run_function, bool(distribute_saved_activations)
        ctx.cpu_rng = torch.get_rng_state()
        ctx.cuda_rng = torch.cuda.get_rng_state() if torch.cuda.is_available() else None
        ctx.tracker = get_cuda_rng_tracker().get_states()
        with torch.no_grad():
            outputs = run_function(*args)
        ctx.first_shape = None
        saved = list(args)
        if ctx.distribute and torch.is_tensor(args[0]) and _ws(mpu.get_tensor_model_parallel_group()) > 1:
            # keep only this TP rank's 1/t slice of the first argument (the hidden states); re-gathered in backward
            g = mpu.get_tensor_model_parallel_group()
            t, r = _ws(g), dist.get_rank(g)
            ctx.first_shape = args[0].shape
            saved[0] = args[0].detach().reshape(-1).chunk(t)[r].clone()
        ctx.tensor_idx = [i for i, a in enumerate(saved) if torch.is_tensor(a)]
        ctx.others = [None if torch.is_tensor(a) else a for a in saved]
        ctx.save_for_backward(*[saved[i] for i in ctx.tensor_idx])
        return outputs

    @staticmethod
    def backward(ctx, *grads):
        if not torch.autograd._is_checkpoint_valid():
            raise RuntimeError("Checkpointing is not compatible with .grad(), please use .backward() if possible")
        args = list(ctx.others)
        for i, t in zip(ctx.tensor_idx, ctx.saved_tensors):
            args[i] = t
        if ctx.first_shape is not None:
            args[0] = all_gather_first_dim(args[0], mpu.get_tensor_model_parallel_group()).view(ctx.first_shape)
        now_cpu, now_tracker = torch.get_rng_state(), get_cuda_rng_tracker().get_states()
        now_cuda = torch.cuda.get_rng_state() if torch.cuda.is_available() else None
        torch.set_rng_state(ctx.cpu_rng)
        if ctx.cuda_rng is not None:
            torch.cuda.set_rng_state(ctx.cuda_rng)
        get_cuda_rng_tracker().set_states(ctx.tracker)
        try:
            detached = [a.detach().requires_grad_(a.requires_grad) if torch.is_tensor(a) else a for a in args]
            with torch.enable_grad():
                outputs = ctx.run_function(*detached)
        finally:
            torch.set_rng_state(now_cpu)
            if now_cuda is not None:
                torch.cuda.set_rng_state(now_cuda)
            get_cuda_rng_tracker().set_states(now_tracker)
        outputs = (outputs,) if torch.is_tensor(outputs) else tuple(outputs)
        pairs = [(o, g) for o, g in zip(outputs, grads) if torch.is_tensor(o) and o.requires_grad and g is not None]
        torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
        return (None, None) + tuple(a.grad if torch.is_tensor(a) else None for a in detached)


def checkpoint(function, distribute_saved_activations, *args):
    """Run `function(*args)` without keeping its intermediate activations; the backward pass re-runs it under the RNG states
    (default generator, CUDA generator, tracker streams and the dropout-seed counter) of the first run, so dropout masks
    repeat.  Same signature as the reference's tensor_parallel.checkpoint."""
    return _CheckpointFunction.apply(function, distribute_saved_activations, *args)
