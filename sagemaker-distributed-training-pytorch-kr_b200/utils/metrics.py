"""Structured observability beside the reference's stdout-only logging:

* `JsonlSink`      - one JSON object per log point (iteration, loss, lr, step ms, tokens/s or images/s, MFU ...), rank 0 only.
                     Enabled with `--metrics-jsonl PATH` (Megatron engine) or the env var B200DIST_METRICS_JSONL (all engines).
* `ScalarWriter`   - `--tensorboard-dir`: torch.utils.tensorboard when the `tensorboard` package is importable, otherwise the
                     same scalars as `<dir>/scalars.jsonl` (the image this framework is developed in has no tensorboard).
* `nvtx_range`     - NVTX push/pop around a phase (no-op without CUDA); the ranges show up in nsys / ncu timelines, and
                     `--profile --profile-step-start/-end --profile-ranks` bracket iterations with cudaProfilerStart/Stop as in
                     the reference (arguments.py:775-787).
* `model_flops_per_token` / `mfu` - the 6 N + attention estimate against the MEASURED cuBLAS bf16 peak (MEASURED_PEAKS.json)
                     so the number is comparable across rounds.
"""
from __future__ import annotations

import contextlib
import json
import os
import time
from typing import Dict, Optional

import torch

_REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
NOMINAL_BF16_FLOPS = 2.25e15          # B200 dense bf16


def measured_peak_flops() -> float:
    """sustained cuBLAS bf16 FLOP/s from MEASURED_PEAKS.json (driver-written), else the nominal dense peak"""
    try:
        with open(os.path.join(_REPO, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)

        v = d.get("bf16_tflops_sustained") or d.get("bf16_tflops")
        return float(v) * 1e12 if v else NOMINAL_BF16_FLOPS
    except Exception:
        return NOMINAL_BF16_FLOPS


def model_flops_per_token(num_layers: int, hidden: int, ffn: int, vocab: int, seq: int, swiglu: bool = False, kv_ratio: float = 1.0) -> float:
    """forward + backward FLOPs per token of a decoder-only transformer (causal attention counted at half the square)"""
    qkv = 2 * hidden * hidden * (1 + 2 * kv_ratio)
    proj = 2 * hidden * hidden
    mlp = 2 * hidden * ffn * (3 if swiglu else 2)
    attn = 2 * 2 * seq * hidden * 0.5
    head = 2 * hidden * vocab
    return 3.0 * (num_layers * (qkv + proj + mlp + attn) + head)


def mfu(tokens_per_s_per_gpu: float, flops_per_token: float, peak: Optional[float] = None) -> float:
    return tokens_per_s_per_gpu * flops_per_token / (peak or measured_peak_flops())


class JsonlSink:
    def __init__(self, path: Optional[str] = None, rank: int = 0):
        path = path or os.environ.get("B200DIST_METRICS_JSONL")
        self._f = None
        if path and rank == 0:
            d = os.path.dirname(os.path.abspath(path))
            os.makedirs(d, exist_ok=True)
            self._f = open(path, "a", buffering=1)

    @property
    def enabled(self) -> bool:
        return self._f is not None

    def write(self, **record):
        if self._f is None:
            return
        record.setdefault("time", round(time.time(), 3))
        self._f.write(json.dumps({k: (float(v) if torch.is_tensor(v) else v) for k, v in record.items()}) + "\n")

    def close(self):
        if self._f is not None:
            self._f.close()
            self._f = None


class ScalarWriter:
    """add_scalar(tag, value, step) like SummaryWriter"""

    def __init__(self, log_dir: Optional[str], rank: int = 0, queue_size: int = 1000):
        self._tb, self._f = None, None
        if not log_dir or rank != 0:
            return
        os.makedirs(log_dir, exist_ok=True)
        try:
            from torch.utils.tensorboard import SummaryWriter          # needs the `tensorboard` package
            self._tb = SummaryWriter(log_dir=log_dir, max_queue=queue_size)
        except Exception:
            print(f"> tensorboard is not installed: writing scalars to {os.path.join(log_dir, 'scalars.jsonl')}", flush=True)
            self._f = open(os.path.join(log_dir, "scalars.jsonl"), "a", buffering=1)

    @property
    def enabled(self) -> bool:
        return self._tb is not None or self._f is not None

    def add_scalar(self, tag: str, value, step: int):
        v = float(value)
        if self._tb is not None:
            self._tb.add_scalar(tag, v, step)
        elif self._f is not None:
            self._f.write(json.dumps({"tag": tag, "value": v, "step": int(step)}) + "\n")

    def flush(self):
        if self._tb is not None:
            self._tb.flush()


_NVTX_ON = [os.environ.get("B200DIST_NVTX", "0") == "1"]


def enable_nvtx(on: bool = True):
    _NVTX_ON[0] = bool(on)


@contextlib.contextmanager
def nvtx_range(name: str):
    on = _NVTX_ON[0] and torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


def nvtx_push(name: str):
    if _NVTX_ON[0] and torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)


def nvtx_pop():
    if _NVTX_ON[0] and torch.cuda.is_available():
        torch.cuda.nvtx.range_pop()
