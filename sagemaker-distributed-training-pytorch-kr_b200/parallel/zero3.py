"""ZeRO stage 3: parameters, gradients and optimizer state partitioned over the data-parallel group.

Counterpart of DeepSpeed's zero/{partition_parameters,parameter_offload,partitioned_param_coordinator,stage3}.py
 configured by configs/default_offload_opt_param.json.  B200-first layout:

  * a *unit* (one decoder layer, the embeddings, the head) owns ONE flat low-precision buffer; each rank keeps only
    its 1/world slice of it (+ fp32 master / exp_avg / exp_avg_sq / fp32 grad-accumulation slices);
  * two full-size gather workspaces and two gradient workspaces live in SYMMETRIC memory and rotate between units:
    the parameters of unit L are views into workspace L % n_ws, so their addresses never change;
  * fetch   = copy my slice into the workspace + the fabric all-gather kernel (NVSwitch multicast stores) on a
    side stream, issued one unit ahead (static prefetch order, forward and backward);
  * release = nothing to free: the workspace is simply overwritten by the next fetch;
  * reduce  = ONE fused kernel per unit and micro-step (csrc/comm.cu rs_update): multimem.ld_reduce of my slice of
    the gradient workspace, scaled, accumulated into the fp32 gradient slice; on the optimizer boundary, when no
    global-norm clipping is requested, the same kernel applies AdamW to the fp32 slice and writes the bf16 slice
    (gradient reduce-scatter fused with the partitioned Adam update).  With clipping the update cannot start before
    the last unit's gradient is reduced; the reduce kernels leave sum(g^2) in a device scalar, and the Adam pass is ONE
    launch over the flat arena of all slices that derives the clip coefficient from that scalar on the device (no host
    round trip between backward and the update; the norm is only read back when somebody logs it);
  * ``stage3_max_live_parameters`` sizes the ring of gather workspaces (more than two: the last units of the forward are
    still resident when the backward starts, DeepSpeed's ``stage3_max_reuse_distance`` effect) and
    ``stage3_prefetch_bucket_size`` the look-ahead depth of the prefetcher;
  * parameters below ``stage3_param_persistence_threshold`` stay replicated (their grads are all-reduced).

On CPU / gloo (tests) the same bookkeeping runs with torch.distributed collectives.
``offload_param.device = cpu`` keeps the low-precision slices in pinned host memory and stages them H2D before each
fetch (slow path).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import native as N


def _align(n, a):
    return (n + a - 1) // a * a


def _assign(p: nn.Parameter, t: torch.Tensor):
    """p.data = t, also when p still lives on the meta device (where `.data =` refuses the device change)"""
    if p.device.type == "meta":
        keep = dict(p.__dict__)
        torch.utils.swap_tensors(p, nn.Parameter(t, requires_grad=p.requires_grad))
        p.__dict__.update(keep)
    else:
        p.data = t


class _Unit:
    def __init__(self, index: int, name: str, module: nn.Module, params: List[nn.Parameter], names: List[str]):
        self.index, self.name, self.module, self.params, self.names = index, name, module, params, names
        self.offsets: List[int] = []
        self.numel = 0          # padded flat size (multiple of world * 128)
        self.slice = 0


class _FetchFn(torch.autograd.Function):
    """identity on the activation; forward = make sure unit `idx` is gathered (and prefetch the next one);
    backward = the unit's backward has finished -> reduce-scatter its gradients."""

    @staticmethod
    def forward(ctx, x, engine, idx):
        ctx.engine, ctx.idx = engine, idx
        engine._ensure(idx, forward=True)
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        ctx.engine._reduce_unit(ctx.idx)
        return g, None, None


class _PreBackwardFn(torch.autograd.Function):
    """identity on the unit output; backward = re-gather the unit before its backward runs (prefetch the previous)."""

    @staticmethod
    def forward(ctx, y, engine, idx):
        ctx.engine, ctx.idx = engine, idx
        return y.view_as(y)

    @staticmethod
    def backward(ctx, g):
        ctx.engine._ensure(ctx.idx, forward=False)
        return g, None, None


class _CheckpointFn(torch.autograd.Function):
    """activation checkpointing of one gather unit (HF `--gradient_checkpointing`): the forward runs without a graph and
    keeps only the unit's input; the backward - which autograd runs between `_PreBackwardFn.backward` (unit re-gathered)
    and `_FetchFn.backward` (unit's gradients reduced) - re-runs the unit with a graph and back-propagates through it, so
    parameter gradients land exactly where the un-checkpointed path puts them."""

    @staticmethod
    def forward(ctx, x, fn, args):
        ctx.fn, ctx.args = fn, args
        ctx.save_for_backward(x)
        ctx.cpu_rng = torch.get_rng_state()
        ctx.cuda_rng = torch.cuda.get_rng_state() if x.is_cuda else None
        with torch.no_grad():
            return fn(x, *args)

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        now_cpu = torch.get_rng_state()
        now_cuda = torch.cuda.get_rng_state() if x.is_cuda else None
        torch.set_rng_state(ctx.cpu_rng)
        if ctx.cuda_rng is not None:
            torch.cuda.set_rng_state(ctx.cuda_rng)
        try:
            xin = x.detach().requires_grad_(True)
            with torch.enable_grad():
                y = ctx.fn(xin, *ctx.args)
            torch.autograd.backward(y, g)
        finally:
            torch.set_rng_state(now_cpu)
            if now_cuda is not None:
                torch.cuda.set_rng_state(now_cuda)
        return xin.grad, None, None


class Zero3Engine:
    def __init__(self, units: Sequence[nn.Module], unit_names: Optional[Sequence[str]] = None, group=None, *, lr: float = 1e-5,
                 betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0, max_grad_norm: float = 1.0,
                 gradient_accumulation_steps: int = 1, persistence_threshold: int = 0, comm_backend: str = "native",
                 offload_param: str = "none", pin_memory: bool = True, grad_dtype: Optional[torch.dtype] = None,
                 offload_optimizer: str = "none", checkpoint_units: Sequence[int] = (), max_live_parameters: float = 0,
                 prefetch_bucket_size: int = 0, max_reuse_distance: float = 0, fuse_gather_gemm_units: Sequence[int] = (),
                 materialize=None, device=None):
        """materialize / device: construction without ever holding the whole model (DeepSpeed `zero.Init`): build the modules on
        the `meta` device and pass `materialize(qualified_name, meta_param) -> tensor on device` (deterministic and identical on
        every rank, e.g. models.llama.meta_initializer); the engine then creates one unit's parameters at a time, keeps this
        rank's slice and drops the rest, so the peak is one unit plus the partitioned state."""
        self.checkpoint_units = set(checkpoint_units)      # unit indices whose activations are recomputed in backward
        self.group = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        self.world = dist.get_world_size(self.group) if self.group is not None else 1
        self.rank = dist.get_rank(self.group) if self.group is not None else 0
        self.lr, self.betas, self.eps, self.wd, self.max_grad_norm = lr, betas, eps, weight_decay, max_grad_norm
        self.ga, self.micro = max(1, gradient_accumulation_steps), 0
        self.step_count = 0
        self.offload_param = offload_param
        # offload_optimizer = cpu (configs/default_offload_opt_param-original.json:24-27): fp32 master shard and both Adam
        # moments live in pinned host memory and the step runs in csrc/host/cpu_adam.cpp (ZeRO-Offload); per step one D2H
        # of the reduced gradient slice and one H2D of the updated low-precision slice per unit
        self.offload_optimizer = offload_optimizer == "cpu"
        first = next(p for u in units for p in u.parameters())
        self.dev, self.dtype = first.device, first.dtype
        self._materialize = materialize
        if self.dev.type == "meta":
            if materialize is None or device is None:
                raise ValueError("ZeRO-3: parameters on the meta device need `materialize` and `device`")
            self.dev = torch.device(device)
        self.grad_dtype = grad_dtype or self.dtype
        self.fabric = None
        if comm_backend == "native" and self.world > 1 and self.dev.type == "cuda":
            from . import fabric
            fb = fabric.get_fabric(self.group)
            if fb.usable:
                self.fabric = fb
        # ---- units, persistent (replicated) parameters
        self.units: List[_Unit] = []
        self.persistent: List[nn.Parameter] = []
        names = list(unit_names) if unit_names is not None else [f"unit{i}" for i in range(len(units))]
        seen = set()
        pad = self.world * 128
        for i, (m, nm) in enumerate(zip(units, names)):
            ps, pn = [], []
            for n, p in m.named_parameters():
                if id(p) in seen or not p.requires_grad:
                    continue
                seen.add(id(p))
                if p.numel() <= persistence_threshold or getattr(p, "_zero3_persistent", False):
                    if p.device.type == "meta":            # replicated parameters are created right away (they are small)
                        _assign(p, materialize(f"{nm}.{n}" if nm else n, p).to(device=self.dev, dtype=self.dtype))
                    self.persistent.append(p)
                    continue
                ps.append(p); pn.append(n)
            u = _Unit(i, nm, m, ps, pn)
            off = 0
            for p in ps:
                u.offsets.append(off)
                off += _align(p.numel(), 64)
            u.numel = _align(max(off, pad), pad)
            u.slice = u.numel // self.world
            self.units.append(u)
        self.num_persistent = sum(p.numel() for p in self.persistent)
        max_numel = max(u.numel for u in self.units)
        # ---- rotating workspaces (symmetric when the fabric is up).  Gather workspaces form a ring of n_ws >= 2:
        # stage3_max_live_parameters bounds how many gathered elements may be alive at once (DeepSpeed's meaning), so
        # n_ws = max_live // largest unit; stage3_max_reuse_distance caps it the same way (a unit that comes back within
        # that many elements - the last units at the forward/backward turn - is still resident and is not re-gathered).
        # The prefetcher looks `depth` units ahead: stage3_prefetch_bucket_size elements, at least one unit, at most n_ws - 1.
        n_units = len(self.units)
        n_ws = 2
        if max_live_parameters and max_live_parameters > 0:
            n_ws = int(max_live_parameters // max_numel)
            if max_reuse_distance and max_reuse_distance > 0:
                n_ws = min(n_ws, max(2, int(max_reuse_distance // max_numel)))
        self.n_ws = max(2, min(n_ws, n_units, 8))
        depth = 1
        if prefetch_bucket_size and prefetch_bucket_size > 0:
            depth = -(-int(prefetch_bucket_size) // max_numel)
        self.prefetch_depth = max(1, min(depth, self.n_ws - 1))
        self._ws_param, self._ws_grad = [], []
        for k in range(self.n_ws):
            if self.fabric is not None:
                self._ws_param.append(self.fabric.alloc(max_numel, self.dtype))
            else:
                self._ws_param.append(torch.zeros(max_numel, dtype=self.dtype, device=self.dev))
        for k in range(2):
            if self.fabric is not None:
                self._ws_grad.append(self.fabric.alloc(max_numel, self.grad_dtype))
            else:
                self._ws_grad.append(torch.zeros(max_numel, dtype=self.grad_dtype, device=self.dev))
        # ---- partition: per-unit slices
        self.lp_slices, self.fp32, self.exp_avg, self.exp_avg_sq, self.gacc = [], [], [], [], []
        self._gacc_host, self._lp_host = [], []
        # every rank-local slice (fp32 master, both moments, fp32 gradient accumulator, low-precision copy) is a view of ONE
        # flat arena per kind, so the clipped optimizer step is a single launch over the arena
        tot_slice = sum(u.slice for u in self.units)
        self._arena = None
        if not self.offload_optimizer:
            self._arena = {k: torch.zeros(tot_slice, dtype=torch.float32, device=self.dev) for k in ("fp32", "m", "v", "gacc")}
            self._arena["lp"] = torch.zeros(tot_slice, dtype=self.dtype, device=self.dev) if offload_param != "cpu" else None
        a_off = 0
        for u in self.units:
            flat = torch.zeros(u.numel, dtype=self.dtype, device=self.dev)
            for p, o, n in zip(u.params, u.offsets, u.names):
                src = p.data if p.device.type != "meta" else materialize(f"{u.name}.{n}" if u.name else n, p)
                flat[o:o + p.numel()].copy_(src.reshape(-1))
                del src
            sl = flat[self.rank * u.slice:(self.rank + 1) * u.slice].clone()
            if offload_param == "cpu":
                sl = sl.cpu().pin_memory() if (pin_memory and torch.cuda.is_available()) else sl.cpu()
            elif self._arena is not None:
                v_ = self._arena["lp"][a_off:a_off + u.slice]
                v_.copy_(sl)
                sl = v_
            self.lp_slices.append(sl)
            if self.offload_optimizer:
                pin = (lambda t: t.pin_memory()) if (pin_memory and torch.cuda.is_available()) else (lambda t: t)
                self.fp32.append(pin(flat[self.rank * u.slice:(self.rank + 1) * u.slice].float().cpu()))
                self.exp_avg.append(pin(torch.zeros(u.slice, dtype=torch.float32)))
                self.exp_avg_sq.append(pin(torch.zeros(u.slice, dtype=torch.float32)))
                self._gacc_host.append(pin(torch.zeros(u.slice, dtype=torch.float32)))
                self._lp_host.append(sl if sl.device.type == "cpu" else pin(torch.zeros(u.slice, dtype=self.dtype)))
                self.gacc.append(torch.zeros(u.slice, dtype=torch.float32, device=self.dev))
            else:
                self._arena["fp32"][a_off:a_off + u.slice].copy_(flat[self.rank * u.slice:(self.rank + 1) * u.slice])
                self.fp32.append(self._arena["fp32"][a_off:a_off + u.slice])
                self.exp_avg.append(self._arena["m"][a_off:a_off + u.slice])
                self.exp_avg_sq.append(self._arena["v"][a_off:a_off + u.slice])
                self.gacc.append(self._arena["gacc"][a_off:a_off + u.slice])
            a_off += u.slice
            ws_p, ws_g = self._param_ws(u.index), self._grad_ws(u.index)
            for p, o in zip(u.params, u.offsets):      # parameters become views of the rotating workspace
                _assign(p, ws_p[o:o + p.numel()].view(p.shape))
                p.grad_buffer = ws_g[o:o + p.numel()].view(p.shape)
                p.zero3_unit = u.index
            del flat
        self.persist_m = [torch.zeros_like(p, dtype=torch.float32) for p in self.persistent]
        self.persist_v = [torch.zeros_like(p, dtype=torch.float32) for p in self.persistent]
        self.persist_fp32 = [p.data.float().clone() for p in self.persistent]
        # units whose ONE partitioned parameter is a 2-D weight cut into whole rows (the LM head): their all-gather is fused
        # into the GEMM that consumes the weight (csrc/gemm.cu b2_agw_gemm) instead of a stand-alone gather
        self._agw_units = set()
        self._agw_pending: Optional[int] = None
        # (the caller names the candidates: the weight must reach the GEMM through ops.functional.linear - an embedding
        # lookup over the same kind of parameter cannot pull its rows on demand)
        if fuse_gather_gemm_units and self.fabric is not None and offload_param != "cpu" and self.dtype in (torch.bfloat16, torch.float16):
            for u in self.units:
                if u.index in set(fuse_gather_gemm_units) and len(u.params) == 1 and u.params[0].dim() == 2:
                    p = u.params[0]
                    rows, K = p.shape
                    if rows % self.world == 0 and (rows // self.world) % 8 == 0 and K % 8 == 0 and u.numel == p.numel():
                        self._agw_units.add(u.index)
                        p._zero3_agw = (self, u.index)
        self._resident = [None] * self.n_ws    # which unit currently occupies each workspace
        self._events: Dict[int, torch.cuda.Event] = {}
        self._stream = torch.cuda.Stream() if self.dev.type == "cuda" else None
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=self.dev)
        self._norm_sq: Optional[torch.Tensor] = None       # global sum(g^2) of the last step, on the device
        self.param_groups = [{"lr": lr, "weight_decay": weight_decay}]
        self._reduce_at_end: List[int] = []

    # ------------------------------------------------------------------ workspace helpers
    @property
    def last_grad_norm(self) -> float:
        """global gradient norm of the last optimizer step (reads the device scalar: synchronises, so log sparingly)"""
        return float(self._norm_sq.sqrt()) if self._norm_sq is not None else 0.0

    def _param_ws(self, idx) -> torch.Tensor:
        w = self._ws_param[idx % self.n_ws]
        return w.tensor if hasattr(w, "tensor") else w

    def _grad_ws(self, idx) -> torch.Tensor:
        w = self._ws_grad[idx % 2]
        return w.tensor if hasattr(w, "tensor") else w

    def total_params(self) -> int:
        return sum(sum(p.numel() for p in u.params) for u in self.units) + self.num_persistent

    # ------------------------------------------------------------------ fetch / prefetch
    def _gather(self, idx: int):
        """all-gather unit idx into workspace idx % n_ws (on the current stream)."""
        u = self.units[idx]
        ws = self._param_ws(idx)
        sl = self.lp_slices[idx]
        mine = ws[self.rank * u.slice:(self.rank + 1) * u.slice]
        if sl.device != mine.device:
            sl = sl.to(mine.device, non_blocking=True)          # offload_param=cpu: pinned H2D staging
        if mine.is_cuda:
            N.axpby(sl, None, 1.0, 0.0, out=mine)                # raw-pointer copy: no autograd version bump on the workspace
        else:
            mine.data.copy_(sl)
        if self.world > 1:
            if self.fabric is not None:
                buf = self._ws_param[idx % self.n_ws]
                N._call("b2_allgather", buf.ptrs, buf.ctrls, buf.mc_ptr or None, u.numel, N._dt(ws), self.rank, self.world, 0, N._st())
            else:
                out = ws[:u.numel]
                dist.all_gather_into_tensor(out.data, mine.data.clone(), group=self.group)
        self._resident[idx % self.n_ws] = idx

    def _stage_own_slice(self, idx: int):
        """fused gather + GEMM: only this rank's slice is put in place; the consuming GEMM pulls the rest (agw_linear)"""
        u = self.units[idx]
        ws = self._param_ws(idx)
        mine = ws[self.rank * u.slice:(self.rank + 1) * u.slice]
        N.axpby(self.lp_slices[idx], None, 1.0, 0.0, out=mine)
        self._resident[idx % self.n_ws] = idx
        self._agw_pending = idx

    def agw_linear(self, idx: int, x2: torch.Tensor) -> torch.Tensor:
        """x2 @ W^T for the single weight W of unit idx whose remote shards have not been gathered yet: ONE kernel gathers
        them over NVLink and multiplies (own rows first); afterwards the workspace holds the whole weight."""
        assert self._agw_pending == idx
        u = self.units[idx]
        p = u.params[0]
        y = self.fabric.agw_gemm(x2, self._ws_param[idx % self.n_ws], u.offsets[0], p.shape[0] // self.world, p.shape[1])
        self._agw_pending = None
        return y

    def _ensure(self, idx: int, forward: bool):
        step = 1 if forward else -1
        if self._agw_pending is not None and self._agw_pending != idx:
            raise RuntimeError(f"ZeRO-3: unit {self._agw_pending} was staged for a fused gather+GEMM that never ran")
        if forward and idx in self._agw_units and self._resident[idx % self.n_ws] != idx:
            if self._stream is not None:
                torch.cuda.current_stream().wait_stream(self._stream)      # earlier side-stream gathers into this ring are done
            self._stage_own_slice(idx)
            return
        if self._stream is None:
            if self._resident[idx % self.n_ws] != idx:
                self._gather(idx)
            return
        cur = torch.cuda.current_stream()
        if self._resident[idx % self.n_ws] != idx:
            self._stream.wait_stream(cur)
            with torch.cuda.stream(self._stream):
                self._gather(idx)
                ev = torch.cuda.Event(); ev.record()
                self._events[idx] = ev
        ev = self._events.pop(idx, None)
        if ev is not None:
            cur.wait_event(ev)
        for k in range(1, self.prefetch_depth + 1):
            nxt = idx + step * k
            if forward and nxt in self._agw_units:
                continue          # gathered by its own first GEMM
            if 0 <= nxt < len(self.units) and self._resident[nxt % self.n_ws] != nxt:
                # prefetch into a workspace whose previous occupant is at least n_ws - depth units behind: its compute was
                # issued earlier on the current stream, which the side stream waits for
                self._stream.wait_stream(cur)
                with torch.cuda.stream(self._stream):
                    self._gather(nxt)
                    ev = torch.cuda.Event(); ev.record()
                    self._events[nxt] = ev

    def wrap(self, idx: int, fn, x, *args):
        """run unit idx: y = fn(x, *args) with fetch before forward/backward and gradient reduction after backward."""
        if torch.is_tensor(x) and x.is_floating_point() and x.requires_grad:
            x = _FetchFn.apply(x, self, idx)          # its backward runs right after the unit's backward
        else:
            self._ensure(idx, forward=True)           # first unit (token ids in): reduced when loss.backward() returns
            if torch.is_grad_enabled():
                self._reduce_at_end.append(idx)
        if idx in self.checkpoint_units and torch.is_grad_enabled() and torch.is_tensor(x) and x.requires_grad:
            y = _CheckpointFn.apply(x, fn, args)
        else:
            y = fn(x, *args)
        if self._agw_pending == idx:
            raise RuntimeError(f"ZeRO-3: unit {idx} ({self.units[idx].name}) did not run its weight through ops.functional.linear; "
                               "do not list it in fuse_gather_gemm_units")
        if torch.is_grad_enabled() and y.requires_grad:
            y = _PreBackwardFn.apply(y, self, idx)
        return y

    # ------------------------------------------------------------------ gradient reduction (+ fused Adam)
    def _reduce_unit(self, idx: int):
        u = self.units[idx]
        # gradients that came back through autograd are moved into the (symmetric) gradient workspace; weight
        # gradients written there directly by the GEMM (ops/functional._wgrad) are already in place
        for p in u.params:
            if p.grad is not None:
                p.grad_buffer.data.copy_(p.grad.to(p.grad_buffer.dtype))
                p.grad = None
            elif not getattr(p, "_gb_written", False):
                p.grad_buffer.data.zero_()
            p._gb_written = False
        boundary = (self.micro + 1) % self.ga == 0
        fuse_adam = boundary and not (self.max_grad_norm and self.max_grad_norm > 0) and not self.offload_optimizer
        gscale = 1.0 / (self.world * self.ga)
        gws = self._grad_ws(idx)
        acc = self.micro % self.ga != 0
        lr, wd = self.param_groups[0]["lr"], self.param_groups[0]["weight_decay"]
        if self.fabric is not None:
            buf = self._ws_grad[idx % 2]
            lp = self.lp_slices[idx]
            N._call("b2_rs_update", buf.ptrs, buf.ctrls, buf.mc_ptr or None, u.numel, N._dt(gws), self.rank, self.world, float(gscale),
                    N._p(self.gacc[idx]), int(acc), N._p(self.sumsq) if boundary else None, int(fuse_adam), N._p(self.fp32[idx]), N._p(self.exp_avg[idx]),
                    N._p(self.exp_avg_sq[idx]), N._p(lp) if (fuse_adam and lp.is_cuda) else None, float(lr), float(self.betas[0]),
                    float(self.betas[1]), float(self.eps), float(wd), self.step_count + 1, 0, N._st())
            if fuse_adam and not lp.is_cuda:
                lp.copy_(self.fp32[idx].to(lp.dtype), non_blocking=True)
        else:
            full = gws[:u.numel].float()
            if self.world > 1:
                dist.all_reduce(full, group=self.group)
            mine = full[self.rank * u.slice:(self.rank + 1) * u.slice] * gscale
            if acc:
                self.gacc[idx].add_(mine)
            else:
                self.gacc[idx].copy_(mine)
            self.sumsq.add_(self.gacc[idx].pow(2).sum() if boundary else 0.0)
            if fuse_adam:
                self._adam_slice(idx, 1.0)
        if boundary and self.offload_optimizer:
            self._gacc_host[idx].copy_(self.gacc[idx], non_blocking=True)      # D2H of the reduced slice, overlaps the rest of backward

    def _adam_slice(self, idx, grad_scale):
        lp = self.lp_slices[idx]
        lr, wd = self.param_groups[0]["lr"], self.param_groups[0]["weight_decay"]
        if self.offload_optimizer:
            from ..optim.cpu_adam import cpu_adam_
            lph = self._lp_host[idx]
            lowp = lph.dtype in (torch.bfloat16, torch.float16)
            cpu_adam_(self.fp32[idx], self.exp_avg[idx], self.exp_avg_sq[idx], self._gacc_host[idx], lph if lowp else None, lr=lr,
                      beta1=self.betas[0], beta2=self.betas[1], eps=self.eps, wd=wd, step=self.step_count + 1, adamw=True, grad_scale=grad_scale)
            if not lowp:
                lp.copy_(self.fp32[idx], non_blocking=True)       # fp32 model (parity runs): the master shard is the parameter shard
            elif lph is not lp:
                lp.copy_(lph, non_blocking=True)
            return
        N.adam_(self.fp32[idx], self.exp_avg[idx], self.exp_avg_sq[idx], self.gacc[idx], lp if lp.device == self.fp32[idx].device else None,
                lr=lr, beta1=self.betas[0], beta2=self.betas[1], eps=self.eps, wd=wd, step=self.step_count + 1, adamw=True, grad_scale=grad_scale)
        if lp.device != self.fp32[idx].device:
            lp.copy_(self.fp32[idx].to(lp.dtype), non_blocking=True)

    # ------------------------------------------------------------------ engine API (DeepSpeedEngine.backward / step)
    def backward(self, loss: torch.Tensor):
        loss.backward()
        for idx in self._reduce_at_end:
            self._reduce_unit(idx)
        self._reduce_at_end = []
        self.micro += 1

    def is_gradient_accumulation_boundary(self) -> bool:
        return self.micro % self.ga == 0 and self.micro > 0

    def step(self):
        """acts on the gradient-accumulation boundary only (DeepSpeedEngine.step semantics)."""
        if not self.is_gradient_accumulation_boundary():
            return False
        lr, wd = self.param_groups[0]["lr"], self.param_groups[0]["weight_decay"]
        # persistent (replicated) parameters: all-reduce their grads, include them in the norm
        for p in self.persistent:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
            if self.world > 1:
                dist.all_reduce(p.grad, group=self.group)
            p.grad.div_(self.world * self.ga)
        clip = bool(self.max_grad_norm and self.max_grad_norm > 0)
        # global sum(g^2): the reduce kernels accumulated this rank's slices on the boundary micro-step; everything below stays
        # on the device (the Adam kernels derive the clip coefficient from the scalar themselves)
        tot = self.sumsq.clone()
        if self.world > 1:
            dist.all_reduce(tot, group=self.group)
        for p in self.persistent:
            tot += p.grad.float().pow(2).sum()
        self._norm_sq = tot
        mx = float(self.max_grad_norm) if clip else 0.0
        if self.offload_optimizer:
            coef = 1.0
            if clip:
                coef = min(1.0, self.max_grad_norm / (float(tot.sqrt()) + 1e-6))
            if self.dev.type == "cuda":
                torch.cuda.current_stream().synchronize()        # the gradient slices' D2H copies have landed
            for i in range(len(self.units)):
                self._adam_slice(i, coef)
        elif clip:
            a = self._arena
            if a["lp"] is not None:
                N.adam_(a["fp32"], a["m"], a["v"], a["gacc"], a["lp"], lr=lr, beta1=self.betas[0], beta2=self.betas[1], eps=self.eps, wd=wd,
                        step=self.step_count + 1, adamw=True, grad_scale=1.0, sumsq=tot, max_norm=mx)
            else:       # offload_param=cpu: low-precision slices live in pinned host memory
                N.adam_(a["fp32"], a["m"], a["v"], a["gacc"], None, lr=lr, beta1=self.betas[0], beta2=self.betas[1], eps=self.eps, wd=wd,
                        step=self.step_count + 1, adamw=True, grad_scale=1.0, sumsq=tot, max_norm=mx)
                for i, lp in enumerate(self.lp_slices):
                    lp.copy_(self.fp32[i].to(lp.dtype), non_blocking=True)
        for p, m, v, p32 in zip(self.persistent, self.persist_m, self.persist_v, self.persist_fp32):
            N.adam_(p32.view(-1), m.view(-1), v.view(-1), p.grad.float().view(-1).contiguous(), None, lr=lr, beta1=self.betas[0], beta2=self.betas[1],
                    eps=self.eps, wd=0.0, step=self.step_count + 1, adamw=True, grad_scale=1.0, sumsq=tot if clip else None, max_norm=mx)
            p.data.copy_(p32.to(p.dtype))
            p.grad = None
        self.step_count += 1
        self.micro = 0
        self.sumsq.zero_()
        self._resident = [None] * self.n_ws   # parameter slices changed: workspaces are stale
        self._events.clear()
        return True

    def zero_grad(self):
        for p in self.persistent:
            p.grad = None

    # ------------------------------------------------------------------ checkpoint helpers
    def gather_full_state_dict(self) -> Dict[str, torch.Tensor]:
        """stage3_gather_16bit_weights_on_model_save: every unit is gathered in turn; returns CPU tensors."""
        out = {}
        for u in self.units:
            self._resident[u.index % self.n_ws] = None
            self._ensure(u.index, forward=True) if self._stream is None else self._gather(u.index)
            if self.dev.type == "cuda":
                torch.cuda.synchronize()
            for n, p in zip(u.names, u.params):
                out[f"{u.name}.{n}" if u.name else n] = p.data.detach().cpu().clone()
        self._resident = [None] * self.n_ws
        return out

    def shard_state(self) -> Dict:
        return {"zero_stage": 3, "partition_count": self.world, "rank": self.rank, "step": self.step_count,
                "fp32_flat_groups": [t.detach().cpu() for t in self.fp32],
                "exp_avg": [t.detach().cpu() for t in self.exp_avg], "exp_avg_sq": [t.detach().cpu() for t in self.exp_avg_sq],
                "unit_names": [u.name for u in self.units], "unit_numel": [u.numel for u in self.units],
                "param_shapes": [{n: tuple(p.shape) for n, p in zip(u.names, u.params)} for u in self.units],
                "param_offsets": [dict(zip(u.names, u.offsets)) for u in self.units],
                # replicated (persistent) parameters: master copy and moments, identical on every rank
                "persist_fp32": [t.detach().cpu() for t in self.persist_fp32], "persist_m": [t.detach().cpu() for t in self.persist_m],
                "persist_v": [t.detach().cpu() for t in self.persist_v]}

    def load_shard_state(self, sd: Dict):
        assert sd["partition_count"] == self.world, "ZeRO-3 checkpoint was written with a different world size"
        self.step_count = sd["step"]
        for i in range(len(self.units)):
            self.fp32[i].copy_(sd["fp32_flat_groups"][i])
            self.exp_avg[i].copy_(sd["exp_avg"][i])
            self.exp_avg_sq[i].copy_(sd["exp_avg_sq"][i])
            self.lp_slices[i].copy_(self.fp32[i].to(self.lp_slices[i].dtype))
        if sd.get("persist_fp32") is not None and len(sd["persist_fp32"]) == len(self.persistent):
            for p, p32, m, v, s32, sm, sv in zip(self.persistent, self.persist_fp32, self.persist_m, self.persist_v,
                                                 sd["persist_fp32"], sd["persist_m"], sd["persist_v"]):
                p32.copy_(s32); m.copy_(sm); v.copy_(sv)
                p.data.copy_(p32.to(p.dtype))
        self._resident = [None] * self.n_ws
