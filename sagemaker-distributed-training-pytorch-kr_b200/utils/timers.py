"""Named timers with log levels and min/max-across-ranks reporting (megatron.timers; flags
--timing-log-level / --timing-log-option / --no-barrier-with-level-1-timing, arguments.py:626-651; in-tree
use: timers('batch-generator', log_level=2) at pretrain_gpt.py:109-112)."""
from __future__ import annotations

import time
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from .metrics import nvtx_pop, nvtx_push


class _Dummy:
    def start(self, barrier=False): pass
    def stop(self, barrier=False): pass
    def reset(self): pass
    def elapsed(self, reset=True, barrier=False): return 0.0


class _Timer:
    def __init__(self, name):
        self.name, self._elapsed, self._started, self._t0 = name, 0.0, False, 0.0
        self._group = None

    def _sync(self, barrier):
        if barrier and dist.is_initialized():
            dist.barrier(group=self._group)
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    def start(self, barrier=False):
        assert not self._started, f"timer {self.name} has already been started"
        self._sync(barrier)
        self._t0, self._started = time.time(), True
        nvtx_push(self.name)          # visible in nsys / ncu timelines when --profile (or B200DIST_NVTX=1) is on

    def stop(self, barrier=False):
        assert self._started, f"timer {self.name} is not started"
        nvtx_pop()
        self._sync(barrier)
        self._elapsed += time.time() - self._t0
        self._started = False

    def reset(self):
        self._elapsed, self._started = 0.0, False

    def elapsed(self, reset=True, barrier=False):
        was = self._started
        if was:
            self.stop(barrier)
        e = self._elapsed
        if reset:
            self.reset()
        if was:
            self.start(barrier)
        return e


class Timers:
    def __init__(self, log_level: int = 0, log_option: str = "minmax"):
        self._log_level, self._log_option = log_level, log_option
        self._timers: Dict[str, _Timer] = {}
        self._levels: Dict[str, int] = {}
        self._dummy = _Dummy()

    def __call__(self, name: str, log_level: Optional[int] = None):
        if name in self._timers:
            return self._timers[name]
        lvl = 2 if log_level is None else log_level   # undeclared level = most verbose, like the reference
        assert lvl <= 2
        if lvl > self._log_level:
            return self._dummy
        self._timers[name] = _Timer(name)
        self._levels[name] = lvl
        return self._timers[name]

    def _gather(self, names, reset, barrier):
        world = dist.get_world_size() if dist.is_initialized() else 1
        rank = dist.get_rank() if dist.is_initialized() else 0
        if barrier and dist.is_initialized():
            dist.barrier()
        local = torch.zeros(len(names))
        for i, n in enumerate(names):
            if n in self._timers:
                local[i] = self._timers[n].elapsed(reset=reset)
        if world == 1:
            return local[None, :]
        dev = torch.device("cuda", torch.cuda.current_device()) if (torch.cuda.is_available() and dist.get_backend() == "nccl") else torch.device("cpu")
        allv = torch.zeros(world, len(names), device=dev)
        dist.all_gather_into_tensor(allv.view(-1), local.to(dev)) if dev.type == "cuda" else dist.all_gather(list(allv.unbind(0)), local)
        return allv.cpu()

    def write(self, names: List[str], writer, iteration: int, normalizer: float = 1.0, reset: bool = False, barrier: bool = False) -> None:
        """--log-timers-to-tensorboard: the maximum over ranks of each named timer as `<name>-time` scalars (megatron.timers.Timers.write)."""
        assert normalizer > 0.0
        allv = self._gather(names, reset, barrier) / normalizer
        for i, n in enumerate(names):
            col = allv[:, i]
            if writer is not None and float(col.max()) > 0.0:
                writer.add_scalar(n + "-time", float(col.max()), iteration)
        if writer is not None:
            writer.flush()

    def log(self, names: List[str], rank=None, normalizer: float = 1.0, reset=True, barrier=False) -> Optional[str]:
        assert normalizer > 0.0
        allv = self._gather(names, reset, barrier) * (1000.0 / normalizer)
        out = None
        if self._log_option in ("max", "minmax"):
            title = "(min, max) time across ranks (ms):" if self._log_option == "minmax" else "max time across ranks (ms):"
            lines = [title]
            for i, n in enumerate(names):
                col = allv[:, i]
                col = col[col > 0]
                if col.numel() == 0:
                    continue
                if self._log_option == "minmax":
                    lines.append("    {}: ({:.2f}, {:.2f})".format((n + " ").ljust(48, "."), col.min().item(), col.max().item()))
                else:
                    lines.append("    {}: {:.2f}".format((n + " ").ljust(48, "."), col.max().item()))
            out = "\n".join(lines) if len(lines) > 1 else None
        else:
            lines = ["times across ranks (ms):"]
            for i, n in enumerate(names):
                lines.append(f"  {n}:")
                for r in range(allv.shape[0]):
                    if allv[r, i] > 0:
                        lines.append("     rank {:2d}: {:.2f}".format(r, allv[r, i].item()))
            out = "\n".join(lines)
        if out is None:
            return None
        world = dist.get_world_size() if dist.is_initialized() else 1
        my = dist.get_rank() if dist.is_initialized() else 0
        if rank is None:
            rank = world - 1
        if my == rank:
            print(out, flush=True)
        return out
