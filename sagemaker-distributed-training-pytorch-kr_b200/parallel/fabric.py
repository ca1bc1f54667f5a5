"""NVLink / NVSwitch fabric: symmetric buffers + the fused in-kernel collectives (csrc/comm.cu, gemm.cu).

Bootstrap uses torch.distributed only for plumbing: ``torch.distributed._symmetric_memory`` allocates a
buffer on every rank of a group, exchanges the handles, maps every peer's copy into this process and - on
NVSwitch systems - binds them to one multicast address.  From then on all traffic is issued by our own
kernels through those addresses (peer ld/st, TMA on peer addresses, multimem.ld_reduce / multimem.st);
NCCL is not involved on these paths.

A symmetric buffer = payload + a 256-byte control block (arrival slots, call epochs) that the kernels use
to stay in lock-step across GPUs; because the epochs live in device memory, calls can be captured into a
CUDA graph and replayed.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..ops import native as N

_CTRL_BYTES = 256
_FABRICS: Dict[int, "Fabric"] = {}

c_p, c_i, c_ll, c_f = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float


class SymmBuffer:
    """payload tensor (this rank's copy) + peer / multicast addresses."""

    def __init__(self, fabric: "Fabric", numel: int, dtype: torch.dtype):
        import torch.distributed._symmetric_memory as symm_mem
        esz = torch.empty((), dtype=dtype).element_size()
        self.payload_bytes = (numel * esz + 255) // 256 * 256
        total = self.payload_bytes + _CTRL_BYTES
        dev = torch.device("cuda", torch.cuda.current_device())
        self.raw = symm_mem.empty(total, dtype=torch.uint8, device=dev)
        self.raw.zero_()
        self.handle = symm_mem.rendezvous(self.raw, group=fabric.group)
        self.tensor = self.raw[: numel * esz].view(dtype)
        self.numel, self.dtype = numel, dtype
        w = fabric.world
        base = [int(p) for p in self.handle.buffer_ptrs]
        self.ptrs = (c_p * w)(*base)
        self.ctrls = (c_p * w)(*[b + self.payload_bytes for b in base])
        mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        self.mc_ptr = mc if (mc and fabric.use_multicast) else 0
        torch.cuda.synchronize()
        dist.barrier(group=fabric.group)     # every rank has zeroed its control block before the first kernel


class Fabric:
    def __init__(self, group=None):
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.usable = False
        self.use_multicast = True
        self.error: Optional[str] = None
        if not torch.cuda.is_available() or self.world < 2 or self.world > 8 or dist.get_backend(self.group) != "nccl":
            self.error = "needs CUDA, an NCCL-bootstrapped group and 2..8 ranks"
            return
        try:
            L = N.lib()
            L.b2_allreduce.argtypes = [c_p, c_p, c_p, c_ll, c_i, c_f, c_i, c_i, c_i, c_p]
            L.b2_allreduce.restype = c_i
            L.b2_ag_gemm.argtypes = [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_ll, c_ll, c_i, c_i, c_i, c_p, c_p, c_ll, c_p, c_ll, c_i, c_i, c_i, c_p]
            L.b2_ag_gemm.restype = c_i
            L.b2_gemm_rs.argtypes = [c_p, c_p, c_i, c_i, c_i, c_ll, c_ll, c_i, c_i, c_p, c_p, c_p, c_p, c_ll, c_p, c_p, c_ll, c_i, c_i, c_p]
            L.b2_gemm_rs.restype = c_i
            L.b2_make_ar_maps.argtypes = [c_p, c_p, c_i, c_i, c_i, c_ll, c_i]
            L.b2_make_ar_maps.restype = c_i
            L.b2_gemm_ar.argtypes = [c_p, c_p, c_i, c_i, c_i, c_ll, c_ll, c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_ll, c_p, c_p, c_ll, c_i, c_i, c_p]
            L.b2_gemm_ar.restype = c_i
            L.b2_xent_vocab_parallel.argtypes = [c_p, c_p, c_p, c_p, c_i, c_i, c_p, c_p, c_p, c_ll, c_i, c_ll, c_ll, c_ll, c_i, c_p]
            L.b2_xent_vocab_parallel.restype = c_i
            L.b2_agw_gemm.argtypes = [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_ll, c_ll, c_i, c_i, c_i, c_i, c_p]
            L.b2_agw_gemm.restype = c_i
            L.b2_fused_config.argtypes = [c_i, c_i]
            L.b2_fused_config.restype = None
            L.b2_reduce_scatter.argtypes = [c_p, c_p, c_p, c_ll, c_i, c_f, c_i, c_i, c_i, c_p, c_p]
            L.b2_reduce_scatter.restype = c_i
            L.b2_allgather.argtypes = [c_p, c_p, c_p, c_ll, c_i, c_i, c_i, c_i, c_p]
            L.b2_allgather.restype = c_i
            L.b2_rs_update.argtypes = [c_p, c_p, c_p, c_ll, c_i, c_i, c_i, c_f, c_p, c_i, c_p, c_i, c_p, c_p, c_p, c_p,
                                       c_f, c_f, c_f, c_f, c_f, c_i, c_i, c_p]
            L.b2_rs_update.restype = c_i
            import os
            L.b2_fused_config(int(os.environ.get("B2_FUSED_PAIR", "-1")), int(os.environ.get("B2_FUSED_NCOMM", "0")))     # tuning / A-B runs
            if os.environ.get("B2_SPIN_TIMEOUT_MS"):
                L.b2_set_spin_timeout_ms.argtypes = [c_ll]
                L.b2_set_spin_timeout_ms.restype = c_i
                L.b2_set_spin_timeout_ms(int(os.environ["B2_SPIN_TIMEOUT_MS"]))
            self._ag, self._rs, self._coll = {}, {}, {}
            probe = SymmBuffer(self, 1024, torch.float32)
            probe.tensor.fill_(float(self.rank + 1))
            self.all_reduce_(probe, scale=1.0)
            torch.cuda.synchronize()
            want = self.world * (self.world + 1) / 2
            ok = bool((probe.tensor == want).all().item())
            flag = torch.tensor([int(ok)], device=probe.tensor.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
            self.usable = bool(flag.item())
            if not self.usable:
                self.error = "probe all-reduce returned wrong values"
        except Exception as e:  # symmetric memory not available on this system -> callers fall back to NCCL
            self.error = f"{type(e).__name__}: {e}"
            self.usable = False

    def alloc(self, numel: int, dtype: torch.dtype) -> SymmBuffer:
        return SymmBuffer(self, numel, dtype)

    def all_reduce_(self, buf: SymmBuffer, scale: float = 1.0, numel: Optional[int] = None, max_ctas: int = 0):
        """in place: buf <- scale * sum over ranks (fused NVLS reduce + scale + broadcast)."""
        n = buf.numel if numel is None else numel
        esz = buf.tensor.element_size()
        q = 16 * self.world                             # every rank reduces a 16-byte aligned slice
        n_pad = (n * esz + q - 1) // q * q // esz       # payload is 256-byte padded, so rounding up is safe
        N._call("b2_allreduce", buf.ptrs, buf.ctrls, c_p(buf.mc_ptr) if buf.mc_ptr else None, n_pad, N._dt(buf.tensor),
                float(scale), self.rank, self.world, max_ctas, N._st())


    # ------------------------------------------------------------------ standalone row collectives
    def _coll_buf(self, numel: int, dtype, tag: str) -> SymmBuffer:
        key = (numel, dtype, tag)
        if key not in self._coll:
            self._coll[key] = self.alloc(numel, dtype)
        return self._coll[key]

    def reduce_scatter_rows(self, x2: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
        """[world*rows, C] partial sums on every rank -> [rows, C] reduced shard * scale (multimem.ld_reduce of my slice)."""
        n = x2.numel()
        buf = self._coll_buf(n, x2.dtype, "rs")
        buf.tensor.view_as(x2).copy_(x2)
        out = torch.empty((x2.shape[0] // self.world,) + tuple(x2.shape[1:]), dtype=x2.dtype, device=x2.device)
        N._call("b2_reduce_scatter", buf.ptrs, buf.ctrls, c_p(buf.mc_ptr) if buf.mc_ptr else None, n, N._dt(x2), float(scale), self.rank,
                self.world, 0, N._p(out), N._st())
        return out

    def all_gather_rows(self, x2: torch.Tensor) -> torch.Tensor:
        """[rows, C] shard -> [world*rows, C] (multimem.st broadcast).  Returns a private copy."""
        n = x2.numel() * self.world
        buf = self._coll_buf(n, x2.dtype, "ag")
        full = buf.tensor.view((x2.shape[0] * self.world,) + tuple(x2.shape[1:]))
        full[self.rank * x2.shape[0]:(self.rank + 1) * x2.shape[0]].copy_(x2)
        N._call("b2_allgather", buf.ptrs, buf.ctrls, c_p(buf.mc_ptr) if buf.mc_ptr else None, n, N._dt(x2), self.rank, self.world, 0, N._st())
        return full.clone()

    # ------------------------------------------------------------------ all-gather -> GEMM
    def _ag_state(self, rows: int, K: int, dtype):
        key = (rows, K, dtype)
        st = self._ag.get(key)
        if st is None:
            symm = self.alloc(rows * K, dtype)
            flags = torch.zeros(self.world * (rows // 128), dtype=torch.int32, device=symm.tensor.device)
            st = self._ag[key] = {"symm": symm, "flags": flags}
        return st

    def shard_buffer(self, rows: int, K: int, dtype) -> torch.Tensor:
        """[rows, K] view of the symmetric shard buffer: a producer (e.g. LayerNorm) that writes here saves the
        staging copy in ag_gemm."""
        return self._ag_state(rows, K, dtype)["symm"].tensor.view(rows, K)

    def ag_gemm(self, x2: torch.Tensor, W: torch.Tensor, b_mn: bool = False, bias=None, epilogue: int = N.EPI_NONE, aux=None,
                out2=None, n_comm: int = 0):
        """y = epilogue(all_gather_rows(x2) @ op(W)); returns (y [world*rows, N], x_full [world*rows, K])."""
        rows, K = x2.shape
        if rows % 128:
            raise ValueError(f"ag_gemm: rows per rank ({rows}) must be a multiple of 128")
        st = self._ag_state(rows, K, x2.dtype)
        shard = st["symm"].tensor.view(rows, K)
        if x2.data_ptr() != shard.data_ptr():
            shard.copy_(x2)
        Nn = W.shape[1] if b_mn else W.shape[0]
        M = rows * self.world
        xg = torch.empty((M, K), dtype=x2.dtype, device=x2.device)
        out = torch.empty((M, Nn), dtype=x2.dtype, device=x2.device)
        N._call("b2_ag_gemm", st["symm"].ptrs, st["symm"].ctrls, N._p(xg), N._p(st["flags"]), N._p(W), N._p(out), rows, Nn, K,
                W.stride(0), out.stride(0), int(b_mn), N._dt(x2), epilogue, N._p(bias), N._p(aux),
                aux.stride(0) if aux is not None else 0, N._p(out2), out2.stride(0) if out2 is not None else 0,
                self.rank, self.world, n_comm, N._st())
        return out, xg

    # ------------------------------------------------------------------ GEMM -> reduce-scatter
    def _rs_state(self, rows: int, Nn: int, dtype):
        key = (rows, Nn, dtype)
        st = self._rs.get(key)
        if st is None:
            stage = self.alloc(2 * self.world * rows * ((Nn + 63) // 64) * 64, dtype)     # [parity][source][row block][64-col chunk][128 x 128 B]
            cnt = self.alloc(max(rows // 128, 1), torch.int32)
            st = self._rs[key] = {"stage": stage, "cnt": cnt}
        return st

    def gemm_rs(self, a2: torch.Tensor, W: torch.Tensor, b_mn: bool = False, bias=None, residual=None):
        """out[rows, N] = residual + bias + reduce_scatter_rows(a2 @ op(W)) in one kernel."""
        M, K = a2.shape
        if M % (128 * self.world):
            raise ValueError(f"gemm_rs: M ({M}) must be a multiple of 128 * world")
        rows = M // self.world
        Nn = W.shape[1] if b_mn else W.shape[0]
        st = self._rs_state(rows, Nn, a2.dtype)
        out = torch.empty((rows, Nn), dtype=a2.dtype, device=a2.device)
        N._call("b2_gemm_rs", N._p(a2), N._p(W), M, Nn, K, a2.stride(0), W.stride(0), int(b_mn), N._dt(a2), st["cnt"].ptrs,
                st["stage"].ctrls, st["stage"].ptrs, N._p(out), out.stride(0), N._p(bias),
                N._p(residual), residual.stride(0) if residual is not None else 0, self.rank, self.world, N._st())
        return out


def _gemm_ar(self, a2: torch.Tensor, W: torch.Tensor, b_mn: bool = False, bias=None, residual=None):
    """out[M, N] = residual + bias + all_reduce(a2 @ op(W)) in one kernel; every rank gets the full result (a private copy)."""
    M, K = a2.shape
    if M % (128 * self.world):
        raise ValueError(f"gemm_ar: M ({M}) must be a multiple of 128 * world")
    rows = M // self.world
    Nn = W.shape[1] if b_mn else W.shape[0]
    st = self._rs_state(rows, Nn, a2.dtype)
    if "ar_out" not in st:
        out = self.alloc(M * Nn, a2.dtype)
        maps = torch.empty(self.world * 128, dtype=torch.uint8, device=a2.device)
        rc = N.lib().b2_make_ar_maps(N._p(maps), out.ptrs, self.world, M, Nn, Nn, N._dt(a2))
        if rc != 0:
            raise RuntimeError("b2_make_ar_maps: " + N.lib().b2_last_error().decode())
        st["ar_out"], st["ar_maps"] = out, maps
    out = st["ar_out"]
    N._call("b2_gemm_ar", N._p(a2), N._p(W), M, Nn, K, a2.stride(0), W.stride(0), int(b_mn), N._dt(a2), st["cnt"].ptrs,
            st["stage"].ctrls, st["stage"].ptrs, N._p(st["ar_maps"]), N._p(out.tensor), Nn, N._p(bias),
            N._p(residual), residual.stride(0) if residual is not None else 0, self.rank, self.world, N._st())
    return out.tensor.view(M, Nn).clone()


Fabric.gemm_ar = _gemm_ar


def _xent_vocab_parallel(self, logits2: torch.Tensor, labels: torch.Tensor, vocab_start: int, ignore_index: int = -100):
    """(global max, global sum-exp, loss) per row of vocab-sharded logits in one kernel (stats -> exchange -> finish)."""
    rows, V = logits2.shape
    buf = self._coll_buf(2 * 3 * rows, torch.float32, "xent")
    gmax = torch.empty(rows, dtype=torch.float32, device=logits2.device)
    gsum, loss = torch.empty_like(gmax), torch.empty_like(gmax)
    N._call("b2_xent_vocab_parallel", N._p(logits2), N._p(labels), buf.ptrs, buf.ctrls, self.rank, self.world, N._p(gmax), N._p(gsum),
            N._p(loss), rows, V, logits2.stride(0), int(vocab_start), int(ignore_index), N._dt(logits2), N._st())
    return gmax, gsum, loss


Fabric.xent_vocab_parallel = _xent_vocab_parallel


def _agw_gemm(self, x2: torch.Tensor, buf: SymmBuffer, elem_offset: int, shard_rows: int, K: int, n_comm: int = 0) -> torch.Tensor:
    """y [M, world * shard_rows] = x2 @ Wfull^T where Wfull lives in the symmetric buffer `buf` at `elem_offset`, row-sharded
    over the ranks, and only this rank's shard is in place: the kernel gathers the other shards while it multiplies.  On
    return (stream order) the buffer holds the complete weight."""
    M = x2.shape[0]
    assert x2.shape[1] == K and x2.is_contiguous()
    esz = buf.tensor.element_size()
    w = self.world
    key = ("agw", id(buf), elem_offset, shard_rows, K)
    st = self._coll.get(key)
    if st is None:
        bpr = (shard_rows + 127) // 128
        st = self._coll[key] = {
            "flags": torch.zeros(w * bpr, dtype=torch.int32, device=x2.device),
            "shards": (c_p * w)(*[int(buf.ptrs[r]) + (elem_offset + r * shard_rows * K) * esz for r in range(w)])}
    out = torch.empty((M, w * shard_rows), dtype=x2.dtype, device=x2.device)
    N._call("b2_agw_gemm", st["shards"], buf.ctrls, int(buf.ptrs[self.rank]) + elem_offset * esz, N._p(st["flags"]), N._p(x2), N._p(out),
            M, shard_rows, K, x2.stride(0), out.stride(0), N._dt(x2), self.rank, w, n_comm, N._st())
    return out


Fabric.agw_gemm = _agw_gemm


def get_fabric(group=None) -> Fabric:
    g = group if group is not None else dist.group.WORLD
    key = id(g)
    if key not in _FABRICS:
        _FABRICS[key] = Fabric(g)
        if dist.get_rank() == 0:
            f = _FABRICS[key]
            print(f"> NVLink fabric for group of {f.world}: {'ready' if f.usable else 'unavailable (' + str(f.error) + ')'}", flush=True)
    return _FABRICS[key]
