"""Bucketed data-parallel wrapper (the DDP of pytorch_mnist_ddp.py:262-264 / pytorch_oxford_ddp.py:251-253).

What torch's DistributedDataParallel + C++ ``Reducer`` do for the reference is done
here by:

* ``csrc/host/ddp_reducer.cpp`` - bucket assignment (reverse registration order, [first, rest] byte caps,
  one bucket family per dtype), per-parameter ready tracking, in-order bucket launch, ready-order
  rebuild after the first iteration;
* this module - flat bucket storage with ``param.grad`` as views into it (no copy-in / copy-out),
  post-accumulate-grad hooks, launch of the per-bucket all-reduce while backward is still running,
  ``no_sync()``, buffer broadcast policy, parameter verification + rank-0 broadcast at construction.

Communication per bucket:
  * CUDA + ``comm_backend="native"``: ONE fused kernel per bucket on a side stream that all-reduces
    through NVSwitch multicast (``multimem.ld_reduce`` / ``multimem.st``), scales by 1/world and casts
    (``parallel.fabric``); the bucket lives in symmetric memory.
  * CUDA + ``"nccl"``: NCCL all-reduce of the pre-divided bucket (the reference's behaviour, kept as the
    measured baseline).
  * CPU: gloo all-reduce (``--backend`` fallback of pytorch_mnist_ddp.py:92).
"""
from __future__ import annotations

import contextlib
import ctypes
from typing import List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from ..utils import hostlib

_MiB = 1024 * 1024


def assign_buckets(nbytes: List[int], keys: List[int], first_cap: int, cap: int) -> List[int]:
    """bucket index for each tensor, tensors given in visit order (C++ planner)."""
    n = len(nbytes)
    if n == 0:
        return []
    L = hostlib.lib()
    a = (ctypes.c_int64 * n)(*nbytes)
    k = (ctypes.c_int32 * n)(*keys)
    out = (ctypes.c_int32 * n)()
    L.b2h_ddp_assign_buckets(a, k, ctypes.c_int32(n), ctypes.c_int64(first_cap), ctypes.c_int64(cap), out)
    return list(out)


class _Bucket:
    __slots__ = ("index", "params", "param_ids", "flat", "views", "work", "offsets", "symm")

    def __init__(self, index):
        self.index = index
        self.params: List[nn.Parameter] = []
        self.param_ids: List[int] = []
        self.flat: Optional[torch.Tensor] = None
        self.views: List[torch.Tensor] = []
        self.offsets: List[int] = []
        self.work = None
        self.symm = None


class GradBucket:
    """what a communication hook sees of one bucket (torch.distributed.GradBucket)"""

    def __init__(self, b: _Bucket, last: bool):
        self._b, self._last = b, last

    def index(self) -> int:
        return self._b.index

    def buffer(self) -> torch.Tensor:
        return self._b.flat

    def gradients(self) -> List[torch.Tensor]:
        return list(self._b.views)

    def parameters(self) -> List[nn.Parameter]:
        return list(self._b.params)

    def is_last(self) -> bool:
        return self._last

    def set_buffer(self, t: torch.Tensor):
        self._b.flat.copy_(t.view_as(self._b.flat))


class _HookWork:
    def __init__(self, out, flat):
        self.out, self.flat = out, flat

    def wait(self):
        r = self.out
        if hasattr(r, "wait") and not torch.is_tensor(r):       # torch.futures.Future / torch._C.Future / Work
            r = r.wait()
            if isinstance(r, (list, tuple)):
                r = r[0]
        if r is not None and r.data_ptr() != self.flat.data_ptr():
            self.flat.copy_(r.view_as(self.flat))


def allreduce_hook(process_group, bucket: GradBucket):
    """default hook: average over the group (torch ddp_comm_hooks.default_hooks.allreduce_hook)"""
    g = process_group if process_group is not None else dist.group.WORLD
    t = bucket.buffer()
    t.div_(dist.get_world_size(g))
    return dist.all_reduce(t, group=g, async_op=True).get_future().then(lambda f: f.value()[0])


def _compress_hook(dtype):
    def hook(process_group, bucket: GradBucket):
        g = process_group if process_group is not None else dist.group.WORLD
        buf = bucket.buffer()
        c = buf.to(dtype).div_(dist.get_world_size(g))
        fut = dist.all_reduce(c, group=g, async_op=True).get_future()
        return fut.then(lambda f: buf.copy_(f.value()[0]))
    return hook


fp16_compress_hook = _compress_hook(torch.float16)     # torch ddp_comm_hooks.default_hooks.fp16_compress_hook
bf16_compress_hook = _compress_hook(torch.bfloat16)


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, process_group=None,
                 bucket_cap_mb: float = 25, first_bucket_cap_mb: float = 1, broadcast_buffers: bool = True,
                 find_unused_parameters: bool = False, gradient_as_bucket_view: bool = True,
                 comm_backend: str = "auto", skip_constant_buffers: bool = True, rebuild_buckets: bool = True,
                 static_graph: bool = False):
        super().__init__()
        self.module = module
        self.static_graph = static_graph        # accepted for API parity: the bucket plan is static after the first rebuild anyway
        self._comm_hook = None
        self.process_group = process_group if process_group is not None else dist.group.WORLD
        self.world_size = dist.get_world_size(self.process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.process_group) if dist.is_initialized() else 0
        self.broadcast_buffers = broadcast_buffers
        self.skip_constant_buffers = skip_constant_buffers
        self.find_unused_parameters = find_unused_parameters
        self.first_cap = int(first_bucket_cap_mb * _MiB)
        self.cap = int(bucket_cap_mb * _MiB)
        self.require_backward_grad_sync = True
        self._rebuild_pending = rebuild_buckets
        self._callback_queued = False
        self._launched = 0
        self._params = [p for p in module.parameters() if p.requires_grad]
        self._param_index = {id(p): i for i, p in enumerate(self._params)}
        first = self._params[0] if self._params else None
        self.on_cuda = bool(first is not None and first.is_cuda)
        if comm_backend == "auto":
            comm_backend = "native" if self.on_cuda else "gloo"
        self.comm_backend = comm_backend
        self._fabric = None
        if self.comm_backend == "native" and self.world_size > 1:
            from . import fabric
            self._fabric = fabric.get_fabric(self.process_group)
            if not self._fabric.usable:
                self.comm_backend = "nccl"
                self._fabric = None
        self._comm_stream = torch.cuda.Stream() if self.on_cuda else None
        if self.world_size > 1:
            self._verify_and_broadcast()
        self._hooks = []
        order = list(range(len(self._params)))[::-1]   # reverse registration ~ gradient ready order
        self._build_buckets(order)
        for p in self._params:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(p)))

    # ------------------------------------------------------------------ construction helpers
    def _verify_and_broadcast(self):
        """c10d `_verify_param_shape_across_processes` + `_sync_module_states`."""
        sig = [tuple(p.shape) for p in self.module.parameters()]
        gathered = [None] * self.world_size
        dist.all_gather_object(gathered, sig, group=self.process_group)
        for r, other in enumerate(gathered):
            if other != sig:
                raise RuntimeError(f"DDP: parameter shapes differ between rank {self.rank} and rank {r}")
        tensors = [p.data for p in self.module.parameters()] + [b for b in self.module.buffers()]
        self._broadcast_coalesced(tensors, self._src_rank())

    def _src_rank(self) -> int:
        """global rank of the group's rank 0 (torch.distributed.broadcast takes a GLOBAL source rank)"""
        return dist.get_global_rank(self.process_group, 0) if self.process_group is not dist.group.WORLD else 0

    def _broadcast_coalesced(self, tensors, src, chunk_bytes=250 * _MiB):
        by_dtype = {}
        for t in tensors:
            by_dtype.setdefault(t.dtype, []).append(t)
        for dt, ts in by_dtype.items():
            batch, size = [], 0
            for t in ts + [None]:
                if t is not None:
                    batch.append(t)
                    size += t.numel() * t.element_size()
                if batch and (t is None or size >= chunk_bytes):
                    flat = torch.cat([x.reshape(-1) for x in batch])
                    dist.broadcast(flat, src, group=self.process_group)
                    off = 0
                    for x in batch:
                        x.copy_(flat[off:off + x.numel()].view_as(x))
                        off += x.numel()
                    batch, size = [], 0

    def _build_buckets(self, visit_order: List[int]):
        params = [self._params[i] for i in visit_order]
        dtype_keys = {}
        keys = [dtype_keys.setdefault((p.dtype, str(p.device)), len(dtype_keys)) for p in params]
        nbytes = [p.numel() * p.element_size() for p in params]
        assignment = assign_buckets(nbytes, keys, self.first_cap, self.cap)
        nb = (max(assignment) + 1) if assignment else 0
        old = {id(p): p.grad for p in self._params if p.grad is not None}
        self.buckets = [_Bucket(i) for i in range(nb)]
        bucket_of = [0] * len(self._params)
        for pos, pi in enumerate(visit_order):
            b = self.buckets[assignment[pos]]
            b.params.append(self._params[pi])
            b.param_ids.append(pi)
            bucket_of[pi] = assignment[pos]
        for b in self.buckets:
            total, offs = 0, []
            for p in b.params:
                offs.append(total)
                total += (p.numel() + 31) // 32 * 32     # keep every view 64/128-byte aligned
            b.offsets = offs
            p0 = b.params[0]
            if self._fabric is not None:
                b.symm = self._fabric.alloc(total, p0.dtype)
                b.flat = b.symm.tensor
                b.flat.zero_()
            else:
                b.flat = torch.zeros(total, dtype=p0.dtype, device=p0.device)
            b.views = [b.flat[o:o + p.numel()].view_as(p) for o, p in zip(offs, b.params)]
            for p, v in zip(b.params, b.views):
                g = old.get(id(p))
                if g is not None:
                    v.copy_(g)
                p.grad = v
        L = hostlib.lib()
        arr = (ctypes.c_int32 * len(bucket_of))(*bucket_of)
        if getattr(self, "_core", None):
            L.b2h_reducer_destroy(ctypes.c_void_p(self._core))
        self._core = L.b2h_reducer_create(arr, ctypes.c_int32(len(bucket_of)), ctypes.c_int32(nb))
        self._bucket_of = bucket_of
        self._launch_buf = (ctypes.c_int32 * max(nb, 1))()

    # ------------------------------------------------------------------ backward-time machinery
    def _make_hook(self, p):
        pi = self._param_index[id(p)]

        def hook(param):
            if not self.require_backward_grad_sync or self.world_size == 1:
                return
            b = self.buckets[self._bucket_of[pi]]
            view = b.views[b.param_ids.index(pi)]
            if param.grad is not None and param.grad.data_ptr() != view.data_ptr():
                view.copy_(param.grad)      # e.g. after zero_grad(set_to_none=True)
                param.grad = view
            if not self._callback_queued:
                torch.autograd.Variable._execution_engine.queue_callback(self._finalize)
                self._callback_queued = True
            n = hostlib.lib().b2h_reducer_mark_ready(ctypes.c_void_p(self._core), ctypes.c_int32(pi), self._launch_buf)
            if n < 0:
                raise RuntimeError("DDP: a parameter was marked ready twice in one backward pass "
                                   "(shared parameter used in several autograd graphs without no_sync())")
            for i in range(n):
                self._launch(self.buckets[self._launch_buf[i]])
        return hook

    def _launch(self, b: _Bucket):
        self._launched += 1
        if self._comm_hook is not None:
            state, hook = self._comm_hook
            out = hook(state, GradBucket(b, b.index == len(self.buckets) - 1))
            b.work = _HookWork(out, b.flat)
            return
        if self.comm_backend == "native":
            ev = torch.cuda.Event()
            ev.record()
            self._comm_stream.wait_event(ev)
            with torch.cuda.stream(self._comm_stream):
                self._fabric.all_reduce_(b.symm, scale=1.0 / self.world_size)
            b.work = self._comm_stream
        else:
            b.flat.div_(self.world_size)
            b.work = dist.all_reduce(b.flat, group=self.process_group, async_op=True)

    def _finalize(self):
        self._callback_queued = False
        L = hostlib.lib()
        nun = ctypes.c_int32(0)
        unused = (ctypes.c_int32 * max(len(self._params), 1))()
        n = L.b2h_reducer_finalize(ctypes.c_void_p(self._core), self._launch_buf, unused, ctypes.byref(nun))
        if nun.value and not self.find_unused_parameters:
            names = [k for k, v in self.module.named_parameters() if id(v) in {id(self._params[unused[i]]) for i in range(nun.value)}]
            raise RuntimeError(f"DDP: {nun.value} parameters received no gradient ({names[:5]}...); "
                               "pass find_unused_parameters=True")
        for i in range(n):
            self._launch(self.buckets[self._launch_buf[i]])
        waited_stream = False
        for b in self.buckets:
            if b.work is None:
                continue
            if isinstance(b.work, _HookWork):
                b.work.wait()
            elif isinstance(b.work, torch.cuda.Stream):
                if not waited_stream:
                    torch.cuda.current_stream().wait_stream(b.work)
                    waited_stream = True
            else:
                b.work.wait()
            b.work = None
        if self._rebuild_pending:
            # rebuild buckets in the order gradients actually arrived (torch Reducer::rebuild_buckets)
            order = (ctypes.c_int32 * max(len(self._params), 1))()
            m = L.b2h_reducer_ready_order(ctypes.c_void_p(self._core), order)
            self._rebuild_pending = False
            ready = [order[i] for i in range(m)]
            if self.find_unused_parameters:
                ready = None       # torch skips the rebuild here: a parameter unused in iteration 1 would have no bucket later
            else:
                # every rank must rebuild with the SAME layout (the rebuild allocates symmetric buffers collectively and
                # same-index buckets are reduced together): rank 0's arrival order wins, as in torch's Reducer
                seen = set(ready)
                ready += [i for i in range(len(self._params)) if i not in seen]     # never-ready parameters keep a bucket
                if self.world_size > 1:
                    t = torch.tensor(ready, dtype=torch.int32, device=self._params[0].device)
                    dist.broadcast(t, src=self._src_rank(), group=self.process_group)
                    ready = [int(v) for v in t.tolist()]
            if ready is not None and ready != [pid for b in self.buckets for pid in b.param_ids]:
                self._build_buckets(ready)
                L.b2h_reducer_reset(ctypes.c_void_p(self._core))
                return
        L.b2h_reducer_reset(ctypes.c_void_p(self._core))

    # ------------------------------------------------------------------ public API
    def register_comm_hook(self, state, hook):
        """torch DDP's communication-hook contract (default_comm_hooks.cpp / ddp_comm_hooks):
        `hook(state, bucket: GradBucket) -> Future[Tensor] | Tensor`; the returned tensor replaces the bucket's gradients
        (it must already be averaged).  Registering a hook bypasses the fused NVLS kernel for the buckets."""
        if self._comm_hook is not None:
            raise RuntimeError("register_comm_hook: a communication hook is already registered")
        if not callable(hook):
            raise TypeError("communication hook must be callable")
        self._comm_hook = (state, hook)

    @contextlib.contextmanager
    def no_sync(self):
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def _sync_buffers(self):
        """torch DDP broadcasts every module buffer from rank 0 before each forward.  Integer /
        bool buffers (swin's 24 relative_position_index tables) never change, so they are skipped."""
        bufs = [b for b in self.module.buffers()
                if not (self.skip_constant_buffers and not (b.is_floating_point() or b.is_complex()))]
        if bufs:
            self._broadcast_coalesced(bufs, self._src_rank())

    def forward(self, *args, **kwargs):
        if self.world_size > 1 and self.broadcast_buffers and self.require_backward_grad_sync:
            self._sync_buffers()
        return self.module(*args, **kwargs)

    def zero_grad(self, set_to_none: bool = False):
        for b in self.buckets:
            b.flat.zero_()
        for b in self.buckets:
            for p, v in zip(b.params, b.views):
                p.grad = v

    def state_dict(self, *args, **kwargs):
        # keys carry the ``module.`` prefix exactly like the reference's saved DDP wrappers
        return super().state_dict(*args, **kwargs)

    def bucket_summary(self):
        return [(b.index, len(b.params), b.flat.numel() * b.flat.element_size()) for b in self.buckets]

    def __del__(self):
        try:
            if getattr(self, "_core", None):
                hostlib.lib().b2h_reducer_destroy(ctypes.c_void_p(self._core))
                self._core = None
        except Exception:
            pass
