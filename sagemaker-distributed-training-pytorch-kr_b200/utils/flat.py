"""Flat parameter arenas: re-point every parameter of a module at a slice of one contiguous buffer per
dtype so optimizers, gradient reducers and checkpoints can treat a model as a few flat tensors."""
from __future__ import annotations

from typing import Dict, List, Tuple

import torch
import torch.nn as nn


def flatten_parameters(module: nn.Module, align: int = 32) -> Dict[torch.dtype, torch.Tensor]:
    groups: Dict[torch.dtype, List[nn.Parameter]] = {}
    seen = set()
    for p in module.parameters():
        if id(p) in seen:
            continue
        seen.add(id(p))
        groups.setdefault(p.dtype, []).append(p)
    arenas = {}
    for dt, ps in groups.items():
        total = sum((p.numel() + align - 1) // align * align for p in ps)
        flat = torch.zeros(total, dtype=dt, device=ps[0].device)
        off = 0
        for p in ps:
            v = flat[off: off + p.numel()].view_as(p)
            v.copy_(p.data)
            p.data = v
            off += (p.numel() + align - 1) // align * align
        arenas[dt] = flat
    return arenas


def param_layout(params: List[nn.Parameter], align: int = 32) -> Tuple[List[int], int]:
    offs, off = [], 0
    for p in params:
        offs.append(off)
        off += (p.numel() + align - 1) // align * align
    return offs, off
