"""Mixed-precision optimizer over flat arenas + the data-parallel gradient buffer.

One object covers what Megatron splits between its local DistributedDataParallel (contiguous fp32 grad
buffer with `main_grad` views, one DP all-reduce after the last micro-batch), Float16OptimizerWith-
Float16Params (fp32 master weights, unscale + inf check, clip, step, copy-back), the grad scalers and
apex FusedAdam (C14, C15, C16, C17; K21, K22):

  model params (bf16/fp16) : ONE arena, parameters re-pointed at views of it
  main params  (fp32)      : arena (aliases the model arena when training in fp32)
  main_grad    (fp32)      : arena, `param.main_grad` views; wgrad GEMMs accumulate straight into it
  exp_avg / exp_avg_sq     : arenas

so a step is: [DP all-reduce of the grad arena] -> [SP / embedding grad all-reduces] -> sum-of-squares
kernels (+ model-parallel all-reduce of one scalar) -> ONE fused AdamW launch per weight-decay group that
reads the clip coefficient and the found-inf flag on the device and writes the low-precision weights.
No host synchronisation happens inside a step unless fp16 dynamic loss scaling needs the overflow flag.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import native as N
from ..parallel import state as mpu


def _align(n: int, a: int = 64) -> int:
    return (n + a - 1) // a * a


class ConstantGradScaler:
    def __init__(self, scale: float):
        self.scale = float(scale)

    def update(self, found_inf: bool):
        pass

    def state_dict(self):
        return {"scale": self.scale}

    def load_state_dict(self, sd):
        self.scale = float(sd["scale"])


class DynamicGradScaler(ConstantGradScaler):
    """Megatron DynamicGradScaler: back off by `backoff` after `hysteresis` overflowing steps, grow by
    `growth` after `growth_interval` clean steps (flags arguments.py:972-979; trace 3_...ipynb:4347-4348)."""

    def __init__(self, initial_scale=2.0 ** 32, min_scale=1.0, growth_factor=2.0, backoff_factor=0.5, growth_interval=1000, hysteresis=2):
        super().__init__(initial_scale)
        self.min_scale, self.growth_factor, self.backoff_factor = float(min_scale), growth_factor, backoff_factor
        self.growth_interval, self.hysteresis = int(growth_interval), int(hysteresis)
        self._growth_tracker, self._hysteresis_tracker = 0, self.hysteresis

    def update(self, found_inf: bool):
        if found_inf:
            self._growth_tracker = 0
            self._hysteresis_tracker -= 1
            if self._hysteresis_tracker <= 0:
                self.scale = max(self.scale * self.backoff_factor, self.min_scale)
        else:
            self._growth_tracker += 1
            if self._growth_tracker == self.growth_interval:
                self._growth_tracker = 0
                self._hysteresis_tracker = self.hysteresis
                self.scale *= self.growth_factor

    def state_dict(self):
        return {"scale": self.scale, "growth_tracker": self._growth_tracker, "hysteresis_tracker": self._hysteresis_tracker}

    def load_state_dict(self, sd):
        self.scale = float(sd["scale"])
        self._growth_tracker, self._hysteresis_tracker = sd["growth_tracker"], sd["hysteresis_tracker"]


def default_no_weight_decay(name: str, p: torch.Tensor) -> bool:
    """Megatron's rule: biases and 1-D (norm) parameters get no weight decay."""
    return name.endswith(".bias") or p.dim() == 1


class MixedPrecisionOptimizer:
    def __init__(self, model_chunks: List[nn.Module], *, lr: float, weight_decay: float = 0.01, betas=(0.9, 0.999), eps: float = 1e-8,
                 clip_grad: float = 1.0, optimizer: str = "adam", sgd_momentum: float = 0.9, grad_scaler: Optional[ConstantGradScaler] = None,
                 dp_group=None, tp_group=None, mp_group=None, embedding_group=None, comm_backend: str = "nccl",
                 no_wd_fn: Callable[[str, torch.Tensor], bool] = default_no_weight_decay, log_num_zeros: bool = False,
                 overlap_grad_reduce: bool = False, arena_multiple: int = 1, symmetric_model_arena: bool = False):
        self.chunks = list(model_chunks)
        self.lr, self.wd, self.betas, self.eps, self.clip_grad = lr, weight_decay, betas, eps, clip_grad
        self.kind, self.sgd_momentum = optimizer, sgd_momentum
        self.scaler = grad_scaler
        self.dp_group, self.tp_group, self.mp_group, self.embedding_group = dp_group, tp_group, mp_group, embedding_group
        self.dp_size = dist.get_world_size(dp_group) if (dp_group is not None and dist.is_initialized()) else 1
        self.tp_size = dist.get_world_size(tp_group) if (tp_group is not None and dist.is_initialized()) else 1
        self.tp_rank = dist.get_rank(tp_group) if self.tp_size > 1 else 0
        self.step_count = 0
        self.comm_backend = comm_backend
        self.log_num_zeros = log_num_zeros
        # ---- collect parameters, sort into segments: (weight-decay?, de-dup class)
        named = []
        seen = set()
        for ci, ch in enumerate(self.chunks):
            for n, p in ch.named_parameters():
                if id(p) in seen or not p.requires_grad:
                    continue
                seen.add(id(p))
                named.append((f"{ci}.{n}" if len(self.chunks) > 1 else n, p))
        def klass(p):
            if getattr(p, "shared_duplicate", False):
                return 2          # copy of a parameter that lives on another pipeline stage: no norm contribution
            return 0 if getattr(p, "tensor_model_parallel", False) else 1   # 0 sharded, 1 replicated across TP
        keyed = [((0 if not no_wd_fn(n, p) else 1), klass(p), i, n, p) for i, (n, p) in enumerate(named)]
        # order: wd-sharded, wd-replicated, wd-dup, nowd-dup, nowd-replicated, nowd-sharded -> replicated ones adjacent
        keyed.sort(key=lambda k: (k[0], k[1] if k[0] == 0 else -k[1], k[2]))
        self.names = [k[3] for k in keyed]
        self._module_order = [k[2] for k in keyed]          # position of each entry in module (named_parameters) order
        self.params: List[nn.Parameter] = [k[4] for k in keyed]
        self.is_wd = [k[0] == 0 for k in keyed]
        self.klass = [k[1] for k in keyed]
        dev = self.params[0].device
        self.model_dtype = self.params[0].dtype
        for p in self.params:
            assert p.dtype == self.model_dtype, "all parameters of a model chunk set must share one dtype"
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += _align(p.numel())
        m = 64 * max(1, int(arena_multiple))        # (distributed optimizer: arena length divisible by the DP size)
        self.total = (off + m - 1) // m * m
        self.n_wd = sum(_align(p.numel()) for p, w in zip(self.params, self.is_wd) if w)
        # ---- arenas
        self.fabric = None
        if comm_backend == "native" and self.dp_size > 1 and dev.type == "cuda":
            from ..parallel import fabric
            fb = fabric.get_fabric(dp_group)
            if fb.usable:
                self.fabric = fb
        self._model_symm = None
        if symmetric_model_arena and self.fabric is not None:     # the distributed optimizer all-gathers updated slices in place
            self._model_symm = self.fabric.alloc(self.total, self.model_dtype)
            self.model_arena = self._model_symm.tensor
            self.model_arena.zero_()
        else:
            self.model_arena = torch.zeros(self.total, dtype=self.model_dtype, device=dev)
        for p, o in zip(self.params, self.offsets):
            v = self.model_arena[o:o + p.numel()].view_as(p)
            v.copy_(p.data)
            p.data = v
        self.low_precision = self.model_dtype != torch.float32
        self.main_param = self.model_arena.float() if self.low_precision else self.model_arena
        if self.fabric is not None:
            self._grad_symm = self.fabric.alloc(self.total, torch.float32)
            self.main_grad = self._grad_symm.tensor
            self.main_grad.zero_()
        else:
            self._grad_symm = None
            self.main_grad = torch.zeros(self.total, dtype=torch.float32, device=dev)
        for p, o in zip(self.params, self.offsets):
            p.main_grad = self.main_grad[o:o + p.numel()].view_as(p)
        self.exp_avg = torch.zeros(self.total, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(self.total, dtype=torch.float32, device=dev) if optimizer == "adam" else None
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.found_inf = torch.zeros(1, dtype=torch.float32, device=dev)
        self._norm_spans = self._spans(lambda i: self.klass[i] == 0 or (self.klass[i] == 1 and self.tp_rank == 0))
        self._sp_spans = self._spans(lambda i: getattr(self.params[i], "sequence_parallel", False))
        self.last_grad_norm: Optional[torch.Tensor] = None
        self.skipped_last = False
        self.param_groups = [{"lr": lr, "weight_decay": weight_decay}]   # scheduler-facing view

    # ------------------------------------------------------------------ helpers
    def _spans(self, pred):
        spans, start, end = [], None, None
        for i, (p, o) in enumerate(zip(self.params, self.offsets)):
            if pred(i):
                if start is None:
                    start = o
                end = o + _align(p.numel())
            elif start is not None:
                spans.append((start, end))
                start = None
        if start is not None:
            spans.append((start, end))
        return spans

    def zero_grad(self, set_to_none: bool = True):
        self.main_grad.zero_()
        for p in self.params:
            p.grad = None

    def _fold_autograd_grads(self):
        """Parameters whose gradient came back through autograd (.grad) instead of the fused main_grad
        path are folded into the fp32 arena."""
        for p in self.params:
            if p.grad is not None:
                N.accum_fp32_(p.main_grad.view(-1), p.grad.reshape(-1).contiguous()) if p.grad.dtype != torch.float32 else p.main_grad.add_(p.grad)
                p.grad = None

    # ------------------------------------------------------------------ gradient reduction
    def reduce_model_grads(self):
        """DP all-reduce (mean) of the flat grad arena, then TP all-reduce of sequence-parallel vector
        grads and the tied-embedding all-reduce across the first/last pipeline stages."""
        self._fold_autograd_grads()
        if self.dp_size > 1:
            if self.fabric is not None:
                self.fabric.all_reduce_(self._grad_symm, scale=1.0 / self.dp_size)
            else:
                self.main_grad.div_(self.dp_size)
                dist.all_reduce(self.main_grad, group=self.dp_group)
        if self.tp_size > 1:
            for (a, b) in self._sp_spans:
                dist.all_reduce(self.main_grad[a:b], group=self.tp_group)
        if self.embedding_group is not None and dist.get_world_size(self.embedding_group) > 1:
            for p in self.params:
                if getattr(p, "shared_embedding", False) and getattr(p, "in_embedding_group", False):
                    dist.all_reduce(p.main_grad, group=self.embedding_group)

    # ------------------------------------------------------------------ step
    def step(self, lr: Optional[float] = None, weight_decay: Optional[float] = None):
        """Returns (update_successful, grad_norm tensor or None, num_zeros or None)."""
        lr = self.param_groups[0]["lr"] if lr is None else lr
        wd = self.param_groups[0]["weight_decay"] if weight_decay is None else weight_decay
        grad_scale = 1.0
        fi = None
        if self.scaler is not None:
            self.found_inf.zero_()
            N.scale_check_(self.main_grad, 1.0 / self.scaler.scale, self.found_inf)
            if self.mp_group is not None and dist.get_world_size(self.mp_group) > 1:
                dist.all_reduce(self.found_inf, op=dist.ReduceOp.MAX, group=self.mp_group)
            fi = self.found_inf
            bad = bool(self.found_inf.item() > 0) if isinstance(self.scaler, DynamicGradScaler) else False
            self.scaler.update(bad)
            if bad:
                self.skipped_last = True
                return False, None, None
        self.skipped_last = False
        num_zeros = None
        if self.log_num_zeros:          # --log-num-zeros-in-grad: zeros of the (unscaled) gradient, TP-duplicated parameters counted once
            nz = torch.zeros(1, dtype=torch.float32, device=self.main_grad.device)
            for (a, b) in self._norm_spans:
                nz += (self.main_grad[a:b] == 0).sum()
            if self.mp_group is not None and dist.get_world_size(self.mp_group) > 1:
                dist.all_reduce(nz, group=self.mp_group)
            num_zeros = nz
        ss = None
        if self.clip_grad and self.clip_grad > 0:
            first = True
            for (a, b) in self._norm_spans:
                N.sumsq(self.main_grad[a:b], self.sumsq, accumulate=not first)
                first = False
            if first:
                self.sumsq.zero_()
            if self.mp_group is not None and dist.get_world_size(self.mp_group) > 1:
                dist.all_reduce(self.sumsq, group=self.mp_group)
            ss = self.sumsq
            self.last_grad_norm = self.sumsq
        self.step_count += 1
        segs = [(0, self.n_wd, wd), (self.n_wd, self.total, 0.0)]
        for (a, b, w) in segs:
            if b <= a:
                continue
            plp = self.model_arena[a:b] if self.low_precision else None
            if self.kind == "adam":
                N.adam_(self.main_param[a:b], self.exp_avg[a:b], self.exp_avg_sq[a:b], self.main_grad[a:b], plp, lr=lr,
                        beta1=self.betas[0], beta2=self.betas[1], eps=self.eps, wd=w, step=self.step_count, adamw=True,
                        grad_scale=grad_scale, sumsq=ss, max_norm=float(self.clip_grad or 0.0), found_inf=fi)
            else:
                coef = 1.0
                if ss is not None:
                    coef = min(1.0, self.clip_grad / (math.sqrt(float(ss)) + 1e-6))
                N.sgd_(self.main_param[a:b], self.exp_avg[a:b], self.main_grad[a:b], lr=lr, momentum=self.sgd_momentum, wd=w,
                       grad_scale=coef, first_step=self.step_count == 1)
                if plp is not None:
                    N.cast_from_fp32_(plp, self.main_param[a:b])
        return True, self.last_grad_norm, num_zeros

    def params_norm(self) -> float:
        """L2 norm of the fp32 master parameters over the model-parallel group (--log-params-norm)"""
        tot = torch.zeros(1, dtype=torch.float32, device=self.main_param.device)
        for (a, b) in self._norm_spans:
            N.sumsq(self.main_param[a:b], tot, accumulate=True)
        if self.mp_group is not None and dist.get_world_size(self.mp_group) > 1:
            dist.all_reduce(tot, group=self.mp_group)
        return math.sqrt(float(tot))

    def grad_norm_value(self) -> float:
        return math.sqrt(float(self.last_grad_norm)) if self.last_grad_norm is not None else 0.0

    def get_loss_scale(self) -> float:
        return self.scaler.scale if self.scaler is not None else 1.0

    def reload_model_params(self):
        if self.low_precision:
            self.main_param.copy_(self.model_arena.float())

    # ------------------------------------------------------------------ checkpoint
    def state_dict(self) -> Dict:
        per = {}
        for n, p, o in zip(self.names, self.params, self.offsets):
            sl = slice(o, o + p.numel())
            per[n] = {"exp_avg": self.exp_avg[sl].view_as(p).clone(),
                      "exp_avg_sq": self.exp_avg_sq[sl].view_as(p).clone() if self.exp_avg_sq is not None else None,
                      "fp32_param": self.main_param[sl].view_as(p).clone()}
        return {"step": self.step_count, "state": per, "grad_scaler": self.scaler.state_dict() if self.scaler else None,
                "param_groups": [dict(g) for g in self.param_groups]}

    def _load_reference_layout(self, sd: Dict):
        """Optimizer section written by the reference stack (Megatron Float16OptimizerWithFloat16Params.state_dict()):
        {'optimizer': torch Adam state_dict {state: {idx: {exp_avg, exp_avg_sq[, step]}}, param_groups: [{params: [idx..],
        step?}, ..]}, 'grad_scaler': .., 'fp32_from_fp16_params': [[fp32 tensors of group 0], [group 1]]}.  Megatron builds two
        groups in module order - weight-decay parameters, then no-weight-decay ones (biases and 1-D tensors) - which is the
        rule `no_wd_fn` applies here, so the i-th parameter of a group is matched by position and checked by shape."""
        inner = sd["optimizer"]
        groups = inner["param_groups"]
        if len(groups) != 2:
            raise ValueError(f"reference-layout optimizer state has {len(groups)} parameter groups (expected 2: weight decay / no weight decay); "
                             "load the model only with --no-load-optim")
        order = sorted(range(len(self.params)), key=lambda i: self._module_order[i])
        mine = [[i for i in order if self.is_wd[i]], [i for i in order if not self.is_wd[i]]]
        fp32 = sd.get("fp32_from_fp16_params")
        step = None
        for g, idxs in enumerate(mine):
            ref_ids = groups[g]["params"]
            if len(ref_ids) != len(idxs):
                raise ValueError(f"reference-layout optimizer state: group {g} has {len(ref_ids)} parameters, this model has {len(idxs)}; "
                                 "load the model only with --no-load-optim")
            step = groups[g].get("step", step)
            for k, i in enumerate(idxs):
                p, o = self.params[i], self.offsets[i]
                st = inner["state"].get(ref_ids[k])
                sl = slice(o, o + p.numel())
                if st is not None:
                    if tuple(st["exp_avg"].shape) != tuple(p.shape):
                        raise ValueError(f"reference-layout optimizer state: {self.names[i]} is {tuple(p.shape)} here but {tuple(st['exp_avg'].shape)} in the "
                                         "checkpoint (different tensor-parallel size?); load the model only with --no-load-optim")
                    self.exp_avg[sl].copy_(st["exp_avg"].reshape(-1))
                    if self.exp_avg_sq is not None and st.get("exp_avg_sq") is not None:
                        self.exp_avg_sq[sl].copy_(st["exp_avg_sq"].reshape(-1))
                    if st.get("step") is not None:
                        step = st["step"]
                if self.low_precision and fp32 is not None:
                    self.main_param[sl].copy_(fp32[g][k].reshape(-1))
        if step is not None:
            self.step_count = int(step)
        if self.scaler is not None and sd.get("grad_scaler"):
            gs = sd["grad_scaler"]
            self.scaler.load_state_dict({"scale": float(gs["scale"]), "growth_tracker": int(gs.get("growth_tracker", 0)),
                                         "hysteresis_tracker": int(gs.get("hysteresis_tracker", getattr(self.scaler, "hysteresis", 2)))})
        if not (self.low_precision and fp32 is not None):
            self.reload_model_params()

    def load_state_dict(self, sd: Dict):
        if "optimizer" in sd and "state" in sd.get("optimizer", {}):       # written by the reference stack
            return self._load_reference_layout(sd)
        self.step_count = sd["step"]
        for n, p, o in zip(self.names, self.params, self.offsets):
            st = sd["state"].get(n)
            if st is None:
                continue
            sl = slice(o, o + p.numel())
            self.exp_avg[sl].copy_(st["exp_avg"].reshape(-1))
            if self.exp_avg_sq is not None and st.get("exp_avg_sq") is not None:
                self.exp_avg_sq[sl].copy_(st["exp_avg_sq"].reshape(-1))
            if self.low_precision:
                self.main_param[sl].copy_(st["fp32_param"].reshape(-1))
        if self.scaler is not None and sd.get("grad_scaler"):
            self.scaler.load_state_dict(sd["grad_scaler"])
