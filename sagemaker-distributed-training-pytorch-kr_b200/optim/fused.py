"""torch.optim-compatible optimizers whose update math runs in the native kernels (csrc/optim.cu) on
CUDA and in the fp32 PyTorch restatement on CPU.

Parameters whose storage (and gradient storage) is contiguous in memory - e.g. the flat arenas built by
``utils.flat.flatten_parameters`` or DDP's gradient buckets - are merged into runs, so a whole model is
typically updated by a handful of launches instead of one per tensor (the job apex/DeepSpeed
multi_tensor_apply does for the reference).
"""
from __future__ import annotations

from typing import List

import torch

from ..ops import native as N


def _runs(params: List[torch.Tensor]):
    """Group consecutive fp32 params whose data AND grads follow each other in memory with the same
    (small, alignment-padding) gap; such a run is updated by one launch over the whole span."""
    runs, cur = [], []
    for p in params:
        if p.grad is None:
            continue
        if cur:
            q = cur[-1]
            ok = (p.dtype == q.dtype == torch.float32 and p.grad.dtype == torch.float32
                  and p.is_contiguous() and p.grad.is_contiguous() and q.is_contiguous() and q.grad.is_contiguous())
            if ok:   # a run must live inside ONE storage (arena / bucket), not merely be adjacent in the address space
                ok = (p.untyped_storage().data_ptr() == q.untyped_storage().data_ptr()
                      and p.grad.untyped_storage().data_ptr() == q.grad.untyped_storage().data_ptr())
            if ok:
                gap_p = p.data_ptr() - (q.data_ptr() + q.numel() * 4)
                gap_g = p.grad.data_ptr() - (q.grad.data_ptr() + q.grad.numel() * 4)
                ok = 0 <= gap_p == gap_g < 256
            if ok:
                cur.append(p)
                continue
            runs.append(cur)
        cur = [p]
    if cur:
        runs.append(cur)
    return runs


def _span(ts: List[torch.Tensor]) -> torch.Tensor:
    """1-D view from the first element of ts[0] to the last element of ts[-1] (padding included)."""
    if len(ts) == 1:
        return ts[0].reshape(-1)
    n = (ts[-1].data_ptr() + ts[-1].numel() * ts[-1].element_size() - ts[0].data_ptr()) // ts[0].element_size()
    return torch.as_strided(ts[0], (n,), (1,), ts[0].storage_offset())


class _Base(torch.optim.Optimizer):
    _state_names: tuple = ()

    def _run_state(self, run):
        key = run[0]
        st = self.state[key]
        n = _span([q.data for q in run]).numel()
        if "n" not in st or st["n"] != n:
            st.clear()
            st["n"] = n
            st["step"] = 0
            for name in self._state_names:
                st[name] = torch.zeros(n, dtype=torch.float32, device=key.device)
        return st


class Adadelta(_Base):
    """torch.optim.Adadelta semantics (pytorch_mnist_ddp.py:267)."""
    _state_names = ("square_avg", "acc_delta")

    def __init__(self, params, lr=1.0, rho=0.9, eps=1e-6, weight_decay=0.0):
        super().__init__(params, dict(lr=lr, rho=rho, eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        for g in self.param_groups:
            for run in _runs(g["params"]):
                st = self._run_state(run)
                p, gr = _span([q.data for q in run]), _span([q.grad for q in run])
                N.adadelta_(p, st["square_avg"], st["acc_delta"], gr.float() if gr.dtype != torch.float32 else gr,
                            lr=g["lr"], rho=g["rho"], eps=g["eps"], wd=g["weight_decay"])
                st["step"] += 1


class Adam(_Base):
    """torch.optim.Adam / AdamW semantics (pytorch_oxford_ddp.py:258; DeepSpeed FusedAdam adam_w_mode)."""
    _state_names = ("exp_avg", "exp_avg_sq")

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, adamw=False):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, adamw=adamw))

    @torch.no_grad()
    def step(self, closure=None):
        for g in self.param_groups:
            for run in _runs(g["params"]):
                st = self._run_state(run)
                st["step"] += 1
                p, gr = _span([q.data for q in run]), _span([q.grad for q in run])
                if p.dtype != torch.float32:
                    raise TypeError("fused Adam expects fp32 parameters (use optim.mixed for bf16 models)")
                N.adam_(p, st["exp_avg"], st["exp_avg_sq"], gr, None, lr=g["lr"], beta1=g["betas"][0], beta2=g["betas"][1],
                        eps=g["eps"], wd=g["weight_decay"], step=st["step"], adamw=g["adamw"])


class AdamW(Adam):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2):
        super().__init__(params, lr, betas, eps, weight_decay, adamw=True)


class SGD(_Base):
    _state_names = ("momentum_buffer",)

    def __init__(self, params, lr=1e-3, momentum=0.0, weight_decay=0.0):
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        for g in self.param_groups:
            for run in _runs(g["params"]):
                st = self._run_state(run)
                p, gr = _span([q.data for q in run]), _span([q.grad for q in run])
                N.sgd_(p, st["momentum_buffer"], gr, lr=g["lr"], momentum=g["momentum"], wd=g["weight_decay"],
                       first_step=st["step"] == 0)
                st["step"] += 1


class StepLR:
    """torch.optim.lr_scheduler.StepLR (pytorch_mnist_ddp.py:268)."""

    def __init__(self, optimizer, step_size: int, gamma: float = 0.1):
        self.opt, self.step_size, self.gamma, self.epoch = optimizer, step_size, gamma, 0
        self.base = [g["lr"] for g in optimizer.param_groups]

    def step(self):
        self.epoch += 1
        for g, b in zip(self.opt.param_groups, self.base):
            g["lr"] = b * self.gamma ** (self.epoch // self.step_size)

    def get_last_lr(self):
        return [g["lr"] for g in self.opt.param_groups]

    def state_dict(self):
        return {"epoch": self.epoch, "base": self.base}

    def load_state_dict(self, sd):
        self.epoch, self.base = sd["epoch"], sd["base"]
