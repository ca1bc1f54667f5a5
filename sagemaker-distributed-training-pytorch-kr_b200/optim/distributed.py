"""`--use-distributed-optimizer`: optimizer state sharded over the data-parallel group (the role of Megatron's
distrib_optimizer.py; flag at arguments.py via T9).

Same arenas as MixedPrecisionOptimizer (flat bf16 model arena, fp32 main_grad arena that the wgrad GEMMs accumulate
into), but every DP rank keeps the fp32 master copy and the Adam moments only for its contiguous 1/dp slice of the arena:

    backward ...                      main_grad (full, fp32) on every rank
    reduce-scatter (mean) over DP     -> this rank's gradient slice
    grad norm                         sum of squares of the slice, all-reduced over DP and the model-parallel group
    AdamW on the slice                fp32 master slice -> bf16 slice of the model arena
    all-gather over DP                every rank has the updated bf16 parameters again

Memory per rank: 2 + 4 (main_grad) + 12 / dp bytes per parameter instead of 18.  With `--comm-backend native` on GPUs both
collectives are the fabric kernels of csrc/comm.cu on symmetric arenas - `b2_reduce_scatter` (multimem.ld_reduce of this
rank's slice, scaled by 1/dp, straight into the fp32 gradient shard) and `b2_allgather` (multimem.st of the updated bf16
slice into every rank's model arena, in place); otherwise torch.distributed calls (reduce_scatter_tensor /
all_gather_into_tensor on NCCL, all-reduce + slice on gloo).
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.distributed as dist

from ..ops import native as N
from .mixed import DynamicGradScaler, MixedPrecisionOptimizer


class DistributedOptimizer(MixedPrecisionOptimizer):
    def __init__(self, model_chunks, **kw):
        dp_group = kw.get("dp_group")
        dp = dist.get_world_size(dp_group) if (dp_group is not None and dist.is_initialized()) else 1
        kw["arena_multiple"] = dp
        kw["symmetric_model_arena"] = True
        super().__init__(model_chunks, **kw)
        self.dp_rank = dist.get_rank(dp_group) if dp > 1 else 0
        self.shard = self.total // dp
        self.lo, self.hi = self.dp_rank * self.shard, (self.dp_rank + 1) * self.shard
        dev = self.model_arena.device
        # shard-sized optimizer state replaces the full-size arenas of the base class
        self.main_param = self.model_arena[self.lo:self.hi].float()
        self.exp_avg = torch.zeros(self.shard, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(self.shard, dtype=torch.float32, device=dev) if self.kind == "adam" else None
        self.grad_shard = torch.zeros(self.shard, dtype=torch.float32, device=dev)
        self._norm_spans_local = [(max(a, self.lo) - self.lo, min(b, self.hi) - self.lo) for a, b in self._norm_spans if min(b, self.hi) > max(a, self.lo)]

    # gradients: model-parallel reductions on the full arena (base class), data-parallel mean as a reduce-scatter
    def reduce_model_grads(self):
        dp, grp = self.dp_size, self.dp_group
        self.dp_size = 1                        # the base class must not all-reduce over DP
        try:
            super().reduce_model_grads()
        finally:
            self.dp_size = dp
        if dp > 1:
            if self.fabric is not None and self._grad_symm is not None:
                g = self._grad_symm
                N._call("b2_reduce_scatter", g.ptrs, g.ctrls, g.mc_ptr or None, self.total, N.DT_FP32, 1.0 / dp, self.dp_rank, dp, 0,
                        N._p(self.grad_shard), N._st())
            elif self.main_grad.is_cuda:
                dist.reduce_scatter_tensor(self.grad_shard, self.main_grad, group=grp)
                self.grad_shard.div_(dp)
            else:                               # gloo has no reduce-scatter: all-reduce, keep the slice
                dist.all_reduce(self.main_grad, group=grp)
                self.grad_shard.copy_(self.main_grad[self.lo:self.hi]).div_(dp)
        else:
            self.grad_shard.copy_(self.main_grad[self.lo:self.hi])

    def _all_reduce_scalar(self, t, op=dist.ReduceOp.SUM):
        if self.dp_size > 1:
            dist.all_reduce(t, op=op, group=self.dp_group)
        if self.mp_group is not None and dist.get_world_size(self.mp_group) > 1:
            dist.all_reduce(t, op=op, group=self.mp_group)

    def step(self, lr: Optional[float] = None, weight_decay: Optional[float] = None):
        lr = self.param_groups[0]["lr"] if lr is None else lr
        wd = self.param_groups[0]["weight_decay"] if weight_decay is None else weight_decay
        fi = None
        if self.scaler is not None:
            self.found_inf.zero_()
            N.scale_check_(self.grad_shard, 1.0 / self.scaler.scale, self.found_inf)
            self._all_reduce_scalar(self.found_inf, dist.ReduceOp.MAX)
            fi = self.found_inf
            bad = bool(self.found_inf.item() > 0) if isinstance(self.scaler, DynamicGradScaler) else False
            self.scaler.update(bad)
            if bad:
                self.skipped_last = True
                return False, None, None
        self.skipped_last = False
        ss = None
        if self.clip_grad and self.clip_grad > 0:
            self.sumsq.zero_()
            for (a, b) in self._norm_spans_local:
                N.sumsq(self.grad_shard[a:b], self.sumsq, accumulate=True)
            self._all_reduce_scalar(self.sumsq)
            ss = self.sumsq
            self.last_grad_norm = self.sumsq
        self.step_count += 1
        for (a, b, w) in ((0, self.n_wd, wd), (self.n_wd, self.total, 0.0)):
            a, b = max(a, self.lo), min(b, self.hi)
            if b <= a:
                continue
            la, lb = a - self.lo, b - self.lo
            plp = self.model_arena[a:b] if self.low_precision else None
            if self.kind == "adam":
                N.adam_(self.main_param[la:lb], self.exp_avg[la:lb], self.exp_avg_sq[la:lb], self.grad_shard[la:lb], plp, lr=lr,
                        beta1=self.betas[0], beta2=self.betas[1], eps=self.eps, wd=w, step=self.step_count, adamw=True,
                        grad_scale=1.0, sumsq=ss, max_norm=float(self.clip_grad or 0.0), found_inf=fi)
            else:
                coef = 1.0 if ss is None else min(1.0, self.clip_grad / (math.sqrt(float(ss)) + 1e-6))
                N.sgd_(self.main_param[la:lb], self.exp_avg[la:lb], self.grad_shard[la:lb], lr=lr, momentum=self.sgd_momentum, wd=w,
                       grad_scale=coef, first_step=self.step_count == 1)
                if plp is not None:
                    N.cast_from_fp32_(plp, self.main_param[la:lb])
            if not self.low_precision:
                self.model_arena[a:b].copy_(self.main_param[la:lb])
        if self.dp_size > 1:                    # every rank gets the other ranks' updated slices
            mine = self.model_arena[self.lo:self.hi]
            if self._model_symm is not None:
                m = self._model_symm
                N._call("b2_allgather", m.ptrs, m.ctrls, m.mc_ptr or None, self.total, N._dt(self.model_arena), self.dp_rank, self.dp_size, 0, N._st())
            elif self.model_arena.is_cuda:
                dist.all_gather_into_tensor(self.model_arena, mine, group=self.dp_group)
            else:
                parts = [torch.empty_like(mine) for _ in range(self.dp_size)]
                dist.all_gather(parts, mine.clone(), group=self.dp_group)
                for r, t in enumerate(parts):
                    self.model_arena[r * self.shard:(r + 1) * self.shard].copy_(t)
        return True, self.last_grad_norm, None

    def reload_model_params(self):
        self.main_param.copy_(self.model_arena[self.lo:self.hi].float())

    # per-DP-rank optimizer shard (Megatron writes one distrib_optim file per DP rank too)
    def state_dict(self) -> Dict:
        return {"step": self.step_count, "distributed": True, "dp_size": self.dp_size, "shard": (self.lo, self.hi),
                "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone() if self.exp_avg_sq is not None else None,
                "fp32_param": self.main_param.clone(), "grad_scaler": self.scaler.state_dict() if self.scaler else None,
                "param_groups": [dict(g) for g in self.param_groups]}

    def load_state_dict(self, sd: Dict):
        assert sd.get("distributed") and tuple(sd["shard"]) == (self.lo, self.hi) and sd["dp_size"] == self.dp_size, \
            "distributed-optimizer checkpoints are tied to the data-parallel size and rank they were written with"
        self.step_count = sd["step"]
        self.exp_avg.copy_(sd["exp_avg"])
        if self.exp_avg_sq is not None and sd.get("exp_avg_sq") is not None:
            self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.main_param.copy_(sd["fp32_param"])
        if self.scaler is not None and sd.get("grad_scaler"):
            self.scaler.load_state_dict(sd["grad_scaler"])
