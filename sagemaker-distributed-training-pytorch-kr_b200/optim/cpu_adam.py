"""Host-side Adam / AdamW for offloaded optimizer state (role of deepspeed.ops.adam.DeepSpeedCPUAdam; selected
by `offload_optimizer.device = cpu` in configs/default_offload_opt_param-original.json:24-27).

The arithmetic runs in csrc/host/cpu_adam.cpp (AVX-512 when the CPU has it, threads over 16 Ki-element tiles); one pass
updates the fp32 master shard and both moments and writes the bf16 / fp16 copy that goes back to the GPU.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from ..utils import hostlib

_LP = {None: 0, torch.bfloat16: 1, torch.float16: 2}


def _lib():
    lib = hostlib.lib()
    if not getattr(lib, "_cpu_adam_ready", False):
        lib.b2h_cpu_adam.restype = ctypes.c_int
        lib.b2h_cpu_adam.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int, ctypes.c_int64] + [ctypes.c_float] * 5 + \
            [ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        lib.b2h_cpu_sumsq.restype = ctypes.c_double
        lib.b2h_cpu_sumsq.argtypes = [ctypes.c_void_p, ctypes.c_int64]
        lib.b2h_cpu_adam_simd_width.restype = ctypes.c_int
        lib._cpu_adam_ready = True
    return lib


def simd_width() -> int:
    return int(_lib().b2h_cpu_adam_simd_width())


def _check(t: torch.Tensor, name: str):
    assert t.device.type == "cpu" and t.dtype == torch.float32 and t.is_contiguous(), f"{name}: contiguous fp32 host tensor expected"


def cpu_adam_(p32: torch.Tensor, m: torch.Tensor, v: torch.Tensor, g: torch.Tensor, p_lp: Optional[torch.Tensor] = None, *, lr: float,
              beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, wd: float = 0.0, step: int, adamw: bool = True,
              grad_scale: float = 1.0, bias_correction: bool = True, threads: int = 0, force_generic: bool = False) -> None:
    """In-place step on flat host tensors; `p_lp` (bf16 / fp16, same numel) receives the rounded parameters."""
    for t, n in ((p32, "p32"), (m, "m"), (v, "v"), (g, "g")):
        _check(t, n)
    n = p32.numel()
    assert m.numel() == n and v.numel() == n and g.numel() == n
    lp_dt = 0
    if p_lp is not None:
        assert p_lp.device.type == "cpu" and p_lp.is_contiguous() and p_lp.numel() == n and p_lp.dtype in _LP
        lp_dt = _LP[p_lp.dtype]
    rc = _lib().b2h_cpu_adam(p32.data_ptr(), m.data_ptr(), v.data_ptr(), g.data_ptr(), p_lp.data_ptr() if p_lp is not None else None,
                             lp_dt, n, lr, beta1, beta2, eps, wd, int(step), int(adamw), grad_scale, int(bias_correction), int(threads),
                             int(force_generic))
    if rc != 0:
        raise ValueError("b2h_cpu_adam: bad arguments")


def cpu_sumsq(g: torch.Tensor) -> float:
    _check(g, "g")
    return float(_lib().b2h_cpu_sumsq(g.data_ptr(), g.numel()))


class CPUAdam(torch.optim.Optimizer):
    """torch.optim-style wrapper (constructor surface of DeepSpeedCPUAdam): parameters are fp32 host tensors."""

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, adamw_mode=True):
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay))
        self.adamw_mode = adamw_mode

    @torch.no_grad()
    def step(self, closure=None, fp16_param_groups=None):
        loss = closure() if closure is not None else None
        for gi, group in enumerate(self.param_groups):
            for pi, p in enumerate(group["params"]):
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p.data, memory_format=torch.contiguous_format)
                    st["exp_avg_sq"] = torch.zeros_like(p.data, memory_format=torch.contiguous_format)
                st["step"] += 1
                lp = fp16_param_groups[gi][pi].data.view(-1) if fp16_param_groups is not None else None
                lp_host = lp if (lp is not None and lp.device.type == "cpu") else (torch.empty(p.numel(), dtype=lp.dtype) if lp is not None else None)
                cpu_adam_(p.data.view(-1), st["exp_avg"].view(-1), st["exp_avg_sq"].view(-1), p.grad.data.contiguous().view(-1), lp_host,
                          lr=group["lr"], beta1=group["betas"][0], beta2=group["betas"][1], eps=group["eps"], wd=group["weight_decay"],
                          step=st["step"], adamw=self.adamw_mode, bias_correction=group["bias_correction"])
                if lp is not None and lp_host is not lp:
                    lp.copy_(lp_host, non_blocking=True)
        return loss
