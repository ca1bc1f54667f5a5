"""Raw (non-autograd) bindings to ``libb200dist.so`` plus plain-PyTorch fp32 reference versions.

Every op here has two bodies:

* CUDA tensors -> the hand-written sm_100a kernel through the C ABI declared in ``include/b200dist.h``
  (ctypes; pointers + the current CUDA stream).  If the shared library is missing on a GPU machine
  the op raises - there is no silent eager fallback on the GPU.
* CPU tensors -> a straightforward fp32 PyTorch restatement of the same math.  This is the oracle
  the GPU tests compare against, and what the gloo/CPU tiers run.
"""
from __future__ import annotations

import ctypes
import math
import os
from typing import Optional

import torch
import torch.nn.functional as F

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB_PATH = os.path.join(_PKG, "lib", "libb200dist.so")
_lib = None
_launches = 0  # kernels launched through this module (bench.py reports it)

DT_FP16, DT_BF16, DT_FP32 = 0, 1, 2
EPI_NONE, EPI_BIAS, EPI_BIAS_GELU, EPI_BIAS_GELU_DUAL, EPI_BIAS_RESIDUAL, EPI_DGELU = 0, 1, 2, 3, 4, 5
EPI_ERF = 0x100      # OR into a GELU epilogue: exact erf GELU (Swin / torch default) instead of the tanh approximation
ACT_GELU_TANH, ACT_GELU_ERF, ACT_RELU, ACT_SILU = 0, 1, 2, 3

c_p, c_i, c_ll, c_f, c_ull = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_ulonglong

_SIGS = {
    "b2_gemm": [c_p, c_p, c_p, c_i, c_i, c_i, c_ll, c_ll, c_ll, c_i, c_i, c_i, c_i, c_p, c_p, c_ll, c_p, c_ll, c_i, c_i, c_i, c_f, c_i, c_p],
    "b2_norm_bwd_num_parts": [c_i, c_i],
    "b2_layernorm_fwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_f, c_i, c_i, c_p],
    "b2_layernorm_bwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_p],
    "b2_cast_rows_from_fp32": [c_p, c_ll, c_p, c_ll, c_i, c_i, c_p],
    "b2_layernorm_bwd_res": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_p],
    "b2_layernorm_fwd_map": [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_f, c_p, c_i, c_i, c_p],
    "b2_layernorm_bwd_map": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_p, c_i, c_i, c_i, c_p],
    "b2_bias_dropout_add": [c_p, c_p, c_p, c_p, c_ll, c_i, c_f, c_ull, c_i, c_p],
    "b2_dropout_bwd": [c_p, c_p, c_ll, c_f, c_ull, c_i, c_p],
    "b2_bias_act_fwd": [c_p, c_p, c_p, c_ll, c_i, c_i, c_i, c_p],
    "b2_bias_act_bwd": [c_p, c_p, c_p, c_p, c_ll, c_i, c_i, c_i, c_p],
    "b2_swiglu_fwd": [c_p, c_p, c_ll, c_i, c_i, c_p],
    "b2_swiglu_bwd": [c_p, c_p, c_p, c_ll, c_i, c_i, c_p],
    "b2_colsum": [c_p, c_p, c_i, c_i, c_ll, c_i, c_p],
    "b2_rope": [c_p, c_p, c_p, c_p, c_ll, c_i, c_i, c_ll, c_i, c_i, c_i, c_p],
    "b2_axpby": [c_p, c_p, c_p, c_ll, c_f, c_f, c_i, c_p],
    "b2_accum_fp32": [c_p, c_p, c_ll, c_i, c_p],
    "b2_cast_from_fp32": [c_p, c_p, c_ll, c_i, c_p],
    "b2_xent_stats": [c_p, c_p, c_p, c_p, c_p, c_ll, c_i, c_ll, c_ll, c_i, c_p],
    "b2_xent_finish": [c_p, c_p, c_p, c_p, c_p, c_ll, c_ll, c_p],
    "b2_xent_bwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_f, c_ll, c_i, c_ll, c_ll, c_ll, c_i, c_p],
    "b2_adam": [c_p, c_p, c_p, c_p, c_p, c_ll, c_f, c_f, c_f, c_f, c_f, c_i, c_i, c_f, c_p, c_f, c_p, c_i, c_i, c_p],
    "b2_adadelta": [c_p, c_p, c_p, c_p, c_ll, c_f, c_f, c_f, c_f, c_p],
    "b2_sgd": [c_p, c_p, c_p, c_ll, c_f, c_f, c_f, c_f, c_i, c_p],
    "b2_sumsq_scratch_floats": [],
    "b2_sumsq": [c_p, c_ll, c_p, c_p, c_p, c_i, c_i, c_p],
    "b2_scale_check": [c_p, c_ll, c_f, c_p, c_p],
    "b2_embedding_fwd": [c_p, c_p, c_p, c_p, c_p, c_ll, c_i, c_ll, c_ll, c_i, c_p],
    "b2_embedding_bwd": [c_p, c_p, c_p, c_p, c_p, c_ll, c_i, c_ll, c_ll, c_i, c_p],
    "b2_window_shuffle": [c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_p],
    "b2_patchify4": [c_p, c_p, c_i, c_i, c_i, c_i, c_p],
    "b2_rows_scale_add": [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_p],
    "b2_window_attn_fwd": [c_p, c_p, c_p, c_p, c_p, c_ll, c_i, c_i, c_i, c_i, c_f, c_i, c_p],
    "b2_window_attn_bwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_ll, c_i, c_i, c_i, c_i, c_f, c_i, c_p],
    "b2_attn_fwd": [c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i] + [c_ll] * 12 + [c_f, c_i, c_f, c_ull, c_p],
    "b2_attn_bwd": [c_p] * 10 + [c_i, c_i, c_i, c_i] + [c_ll] * 21 + [c_f, c_i, c_f, c_ull, c_p],
    "b2_bn_reduce": [c_p, c_p, c_i, c_i, c_i, c_i, c_p],
    "b2_bn_finalize": [c_p, ctypes.c_double, c_p, c_p, c_f, c_f, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_p],
    "b2_bn_eval_coeffs": [c_p, c_p, c_p, c_p, c_f, c_p, c_p, c_i, c_p],
    "b2_bn_apply": [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_p],
    "b2_bn_bwd_reduce": [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_p],
    "b2_bn_bwd_apply": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, ctypes.c_double, c_p, c_i, c_i, c_i, c_i, c_i, c_p],
    "b2_conv2d_fwd": [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_p],
    "b2_conv2d_bwd": [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_p],
    "b2_maxpool2_fwd": [c_p, c_p, c_p, c_ll, c_i, c_i, c_p],
    "b2_maxpool2_bwd": [c_p, c_p, c_p, c_ll, c_i, c_i, c_p],
    "b2_log_softmax_nll_fwd": [c_p, c_p, c_p, c_p, c_ll, c_i, c_p],
    "b2_log_softmax_nll_bwd": [c_p, c_p, c_p, c_p, c_f, c_p, c_ll, c_i, c_p],
    "b2_softmax_fwd": [c_p, c_p, c_p, c_ll, c_i, c_i, c_i, c_i, c_f, c_i, c_i, c_p],
    "b2_softmax_bwd": [c_p, c_p, c_p, c_ll, c_i, c_f, c_i, c_p],
    "b2_num_sms": [],
    "b2_counter_add": [c_p, c_ull, c_p],
}


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def lib():
    """Load (once) and return the ctypes handle; raises if the library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(
                f"{_LIB_PATH} not found: run `python -m b200dist.build_native` (or __graft_entry__.build()); "
                "the CUDA path has no eager fallback on purpose")
        L = ctypes.CDLL(_LIB_PATH)
        L.b2_last_error.restype = ctypes.c_char_p
        for name, sig in _SIGS.items():
            fn = getattr(L, name, None)
            if fn is None:
                continue
            fn.argtypes = sig
            fn.restype = c_i
        _lib = L
        if os.environ.get("B2_GEMM_PAIR_MODE"):       # A/B switch: 0 = single-CTA tiles only, 2 = CTA pairs wherever possible
            L.b2_gemm_set_pair_mode.argtypes = [ctypes.c_int]
            L.b2_gemm_set_pair_mode.restype = None
            L.b2_gemm_set_pair_mode(int(os.environ["B2_GEMM_PAIR_MODE"]))
        if os.environ.get("B2_WATTN_IMPL"):           # A/B switch: 0 = CUDA-core window attention, 1 = tensor-core kernels (default)
            L.b2_window_attn_set_impl.argtypes = [ctypes.c_int]
            L.b2_window_attn_set_impl(int(os.environ["B2_WATTN_IMPL"]))
    return _lib


def set_window_attn_impl(impl: int) -> int:
    """Swin window attention kernels: 1 = tensor-core (mma.sync, csrc/window_attn_tc.cu; default), 0 = CUDA-core.  Returns the previous value."""
    L = lib()
    L.b2_window_attn_set_impl.argtypes = [ctypes.c_int]
    L.b2_window_attn_set_impl.restype = ctypes.c_int
    return L.b2_window_attn_set_impl(int(impl))


def set_window_attn_grid(fwd_ctas_per_sm: int = 0, bwd_ctas_per_sm: int = 0) -> None:
    """Tuning knob of the tensor-core window attention: CTAs per SM the grids are sized for (0 = occupancy query, the default)."""
    L = lib()
    L.b2_window_attn_set_grid.argtypes = [ctypes.c_int, ctypes.c_int]
    L.b2_window_attn_set_grid.restype = None
    L.b2_window_attn_set_grid(int(fwd_ctas_per_sm), int(bwd_ctas_per_sm))


def gemm_plan(M: int, N: int, K: int, out_fp32: bool = False, accumulate: bool = False, split_k: int = 0, sms: int = 148, pair_mode: int = 1) -> dict:
    """The tile width, CTA pairing and split-K `gemm` would pick for this problem on `sms` SMs (host arithmetic only: works without a GPU)."""
    L = lib()
    out = (ctypes.c_int * 7)()
    L.b2_gemm_plan.argtypes = [ctypes.c_int] * 8 + [ctypes.POINTER(ctypes.c_int)]
    L.b2_gemm_plan.restype = ctypes.c_int
    if L.b2_gemm_plan(M, N, K, int(out_fp32), int(accumulate), split_k, sms, pair_mode, out) != 0:
        raise ValueError("gemm_plan: empty problem")
    return dict(zip(("bn", "pair", "tiles_m", "tiles_n", "num_kb", "split_k", "kb_per_split"), list(out)))


def set_gemm_pair_mode(mode: int) -> None:
    """CTA-pair (cta_group::2) tile selection of `gemm`: 0 never, 1 heuristic (default), 2 whenever the shape allows."""
    L = lib()
    L.b2_gemm_set_pair_mode.argtypes = [ctypes.c_int]
    L.b2_gemm_set_pair_mode.restype = None
    L.b2_gemm_set_pair_mode(int(mode))


def set_reserved_sms(n: int) -> None:
    """leave `n` SMs free of the persistent kernels (needed when NCCL p2p runs next to the spin-waiting fused kernels)"""
    L = lib()
    L.b2_set_reserved_sms.argtypes = [ctypes.c_int]
    L.b2_set_reserved_sms.restype = None
    L.b2_set_reserved_sms(int(n))


_replay_word: Optional[torch.Tensor] = None


def dropout_replay_counter(device=None) -> torch.Tensor:
    """Installs (once) the device word that every dropout kernel mixes into its seed and returns it.  A CUDA graph that
    contains dropout captures ``bump_dropout_replay_counter()`` as its first node, so each replay draws new masks."""
    global _replay_word
    if _replay_word is None:
        dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        _replay_word = torch.zeros(1, dtype=torch.int64, device=dev)
        L = lib()
        L.b2_set_dropout_replay_counter.argtypes = [ctypes.c_void_p]
        L.b2_set_dropout_replay_counter.restype = None
        L.b2_set_dropout_replay_counter(_p(_replay_word))
    return _replay_word


def bump_dropout_replay_counter() -> None:
    _call("b2_counter_add", _p(dropout_replay_counter()), 1, _st())


def launches() -> int:
    return _launches


def count_launches(n: int) -> None:
    """credit `n` kernel launches that did not go through _call (CUDA-graph replays of captured kernels)"""
    global _launches
    _launches += int(n)


def _call(name: str, *args) -> None:
    global _launches
    L = lib()
    rc = getattr(L, name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed: {L.b2_last_error().decode()}")
    _launches += 1


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _st():
    return torch.cuda.current_stream().cuda_stream


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return DT_BF16
    if t.dtype == torch.float16:
        return DT_FP16
    if t.dtype == torch.float32:
        return DT_FP32
    raise TypeError(f"unsupported dtype {t.dtype}")


def num_sms() -> int:
    return lib().b2_num_sms() if torch.cuda.is_available() else 148


# =============================================================================================
# GEMM
# =============================================================================================
def gelu_tanh_ref(x: torch.Tensor) -> torch.Tensor:
    return F.gelu(x, approximate="tanh")


def gelu_tanh_grad_ref(x: torch.Tensor) -> torch.Tensor:
    k0, k1 = 0.7978845608028654, 0.044715
    t = torch.tanh(k0 * x * (1 + k1 * x * x))
    return 0.5 * (1 + t) + 0.5 * x * (1 - t * t) * k0 * (1 + 3 * k1 * x * x)


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         out: Optional[torch.Tensor] = None, epilogue: int = EPI_NONE, bias: Optional[torch.Tensor] = None,
         aux: Optional[torch.Tensor] = None, out2: Optional[torch.Tensor] = None, out_fp32: bool = False,
         accumulate: bool = False, alpha: float = 1.0, split_k: int = 0, max_ctas: int = 0) -> torch.Tensor:
    """C[M,N] (+)= alpha * op(A) @ op(B).

    a: [M,K] (a_mn=False) or [K,M] (a_mn=True); b: [N,K] (b_mn=False, an nn.Linear weight) or [K,N].
    Rows may be strided (last dim contiguous).  ``out_fp32``/``accumulate`` give an fp32 result that is
    added into ``out`` (main_grad semantics).
    """
    assert a.dim() == 2 and b.dim() == 2 and a.stride(1) == 1 and b.stride(1) == 1
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    assert K == Kb, f"gemm K mismatch {K} vs {Kb}"
    if out is None:
        assert not accumulate
        out = torch.empty((M, N), dtype=torch.float32 if out_fp32 else a.dtype, device=a.device)
    if not a.is_cuda:
        A = a.float().t() if a_mn else a.float()
        Bm = b.float() if b_mn else b.float().t()
        c = alpha * (A @ Bm)
        if bias is not None:
            c = c + bias.float()
        erf, epi = bool(epilogue & EPI_ERF), epilogue & ~EPI_ERF
        act = (lambda t: F.gelu(t)) if erf else gelu_tanh_ref
        if epi == EPI_BIAS_GELU:
            c = act(c)
        elif epi == EPI_BIAS_GELU_DUAL:
            out2.copy_(c.to(out2.dtype))
            c = act(c)
        elif epi == EPI_BIAS_RESIDUAL:
            c = c + aux.float()
        elif epi == EPI_DGELU:
            z = aux.float()
            c = c * ((0.5 * (1 + torch.erf(z * 0.7071067811865476)) + z * torch.exp(-0.5 * z * z) * 0.3989422804014327) if erf
                     else gelu_tanh_grad_ref(z))
        if accumulate:
            out.add_(c.to(out.dtype))
        else:
            out.copy_(c.to(out.dtype))
        return out
    assert out.stride(1) == 1
    _call("b2_gemm", _p(a), _p(b), _p(out), M, N, K, a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn),
          _dt(a), epilogue, _p(bias), _p(aux), aux.stride(0) if aux is not None else 0, _p(out2),
          out2.stride(0) if out2 is not None else 0, int(out.dtype == torch.float32), int(accumulate), split_k,
          float(alpha), max_ctas, _st())
    return out


# =============================================================================================
# LayerNorm / RMSNorm
# =============================================================================================
def norm_fwd(x: torch.Tensor, gamma: torch.Tensor, beta: Optional[torch.Tensor], eps: float, rms: bool = False,
             out: Optional[torch.Tensor] = None, row_map: Optional[torch.Tensor] = None):
    """Returns (y, mean, rstd); mean is None for RMSNorm.  ``out`` (CUDA only): a contiguous [rows, H] buffer that receives y
    (e.g. the symmetric shard buffer the fused all-gather->GEMM reads, which saves its staging copy).  ``row_map`` (int32 [rows], a
    permutation): the normalised row r is written to y[row_map[r]] (y is then returned as [rows, H])."""
    H = x.shape[-1]
    x2 = x.reshape(-1, H)
    rows = x2.shape[0]
    if row_map is not None:
        assert row_map.dtype == torch.int32 and row_map.numel() == rows and row_map.is_contiguous() and out is None
        if not x.is_cuda:
            y, mean, rstd = norm_fwd(x2, gamma, beta, eps, rms)
            yp = torch.empty_like(y)
            yp[row_map.long()] = y
            return yp, mean, rstd
        assert x2.is_contiguous()
        y = torch.empty_like(x2)
        mean = None if rms else torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        _call("b2_layernorm_fwd_map", _p(x2), _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd), rows, H, float(eps), _p(row_map), _dt(x),
              int(rms), _st())
        return y, mean, rstd
    if not x.is_cuda:
        xf = x2.float()
        if rms:
            mean = None
            rstd = torch.rsqrt(xf.pow(2).mean(-1) + eps)
            y = xf * rstd[:, None] * gamma.float()
        else:
            mean = xf.mean(-1)
            rstd = torch.rsqrt(xf.var(-1, unbiased=False) + eps)
            y = (xf - mean[:, None]) * rstd[:, None] * gamma.float() + (beta.float() if beta is not None else 0)
        return y.to(x.dtype).view_as(x), mean, rstd
    assert x2.is_contiguous()
    if out is not None:
        assert out.is_contiguous() and out.shape == x2.shape and out.dtype == x2.dtype
        y = out
    else:
        y = torch.empty_like(x2)
    mean = None if rms else torch.empty(rows, dtype=torch.float32, device=x.device)
    rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
    _call("b2_layernorm_fwd", _p(x2), _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd), rows, H, float(eps), _dt(x),
          int(rms), _st())
    return y.view_as(x), mean, rstd


def norm_bwd(dy, x, gamma, mean, rstd, rms: bool = False, dgamma: Optional[torch.Tensor] = None,
             dbeta: Optional[torch.Tensor] = None, accumulate: bool = False, dres: Optional[torch.Tensor] = None,
             row_map: Optional[torch.Tensor] = None):
    """Returns (dx, dgamma_fp32, dbeta_fp32).  With ``accumulate`` the fp32 grads are added in place (main_grad);
    ``dres`` is the gradient arriving through the residual branch: dx = dres + d(norm)/dx in the same pass;
    ``row_map``: the forward wrote row r to y[row_map[r]], so the gradient of row r is dy[row_map[r]]."""
    H = x.shape[-1]
    x2, dy2 = x.reshape(-1, H), dy.reshape(-1, H)
    rows = x2.shape[0]
    if row_map is not None and not x.is_cuda:
        dy2, row_map = dy2[row_map.long()], None
    if dgamma is None:
        dgamma = torch.empty(H, dtype=torch.float32, device=x.device)
        accumulate = False
    if dbeta is None and not rms:
        assert not accumulate, "accumulate needs both dgamma and dbeta"
        dbeta = torch.empty(H, dtype=torch.float32, device=x.device)
    if not x.is_cuda:
        xf, dyf, g = x2.float(), dy2.float(), gamma.float()
        m = 0 if rms else mean[:, None]
        xh = (xf - m) * rstd[:, None]
        gy = dyf * g
        m1 = 0 if rms else gy.mean(-1, keepdim=True)
        m2 = (gy * xh).mean(-1, keepdim=True)
        dx = rstd[:, None] * (gy - m1 - xh * m2)
        if dres is not None:
            dx = dx + dres.reshape(-1, H).float()
        dg = (dyf * xh).sum(0)
        if accumulate:
            dgamma.add_(dg)
        else:
            dgamma.copy_(dg)
        if not rms:
            db = dyf.sum(0)
            if accumulate:
                dbeta.add_(db)
            else:
                dbeta.copy_(db)
        return dx.to(x.dtype).view_as(x), dgamma, dbeta
    assert x2.is_contiguous() and dy2.is_contiguous()
    dx = torch.empty_like(x2)
    parts = lib().b2_norm_bwd_num_parts(rows, H)
    ws = torch.empty(parts * 2 * H, dtype=torch.float32, device=x.device)
    if dres is not None:
        dres = dres.reshape(-1, H)
        assert dres.is_contiguous() and dres.dtype == x.dtype
    if row_map is not None:
        assert row_map.dtype == torch.int32 and row_map.numel() == rows and row_map.is_contiguous()
        _call("b2_layernorm_bwd_map", _p(dy2), _p(x2), _p(gamma), _p(mean), _p(rstd), _p(dres), _p(dx), _p(dgamma), _p(dbeta),
              _p(ws), rows, H, _p(row_map), _dt(x), int(rms), int(accumulate), _st())
    else:
        _call("b2_layernorm_bwd_res", _p(dy2), _p(x2), _p(gamma), _p(mean), _p(rstd), _p(dres), _p(dx), _p(dgamma), _p(dbeta),
              _p(ws), rows, H, _dt(x), int(rms), int(accumulate), _st())
    return dx.view_as(x), dgamma, dbeta


# =============================================================================================
# elementwise
# =============================================================================================
def _philox_keep_ref(n: int, p: float, seed: int) -> torch.Tensor:
    # CPU oracle uses torch's generator; masks only need to be self-consistent fwd<->bwd per backend
    g = torch.Generator().manual_seed(seed & 0x7FFFFFFF)
    return torch.rand(n, generator=g) >= p


def bias_dropout_add(x, bias, residual, p: float = 0.0, seed: int = 0):
    H = x.shape[-1]
    if not x.is_cuda:
        t = x.float() + (bias.float() if bias is not None else 0)
        if p > 0:
            keep = _philox_keep_ref(t.numel(), p, seed).view_as(t)
            t = torch.where(keep, t / (1 - p), torch.zeros_like(t))
        if residual is not None:
            t = t + residual.float()
        return t.to(x.dtype)
    out = torch.empty_like(x)
    _call("b2_bias_dropout_add", _p(x), _p(bias), _p(residual), _p(out), x.numel(), H, float(p), seed, _dt(x), _st())
    return out


def dropout_bwd(dy, p: float, seed: int):
    if not dy.is_cuda:
        keep = _philox_keep_ref(dy.numel(), p, seed).view_as(dy)
        return torch.where(keep, dy.float() / (1 - p), torch.zeros_like(dy, dtype=torch.float32)).to(dy.dtype)
    dx = torch.empty_like(dy)
    _call("b2_dropout_bwd", _p(dy), _p(dx), dy.numel(), float(p), seed, _dt(dy), _st())
    return dx


def _act_ref(z, mode):
    if mode == ACT_GELU_TANH:
        return gelu_tanh_ref(z)
    if mode == ACT_GELU_ERF:
        return F.gelu(z)
    if mode == ACT_RELU:
        return F.relu(z)
    return F.silu(z)


def bias_act_fwd(x, bias, mode: int):
    if not x.is_cuda:
        return _act_ref(x.float() + (bias.float() if bias is not None else 0), mode).to(x.dtype)
    out = torch.empty_like(x)
    _call("b2_bias_act_fwd", _p(x), _p(bias), _p(out), x.numel(), x.shape[-1], mode, _dt(x), _st())
    return out


def bias_act_bwd(dy, x, bias, mode: int):
    if not x.is_cuda:
        z = (x.float() + (bias.float() if bias is not None else 0)).requires_grad_(True)
        with torch.enable_grad():
            y = _act_ref(z, mode)
        (g,) = torch.autograd.grad(y, z, dy.float())
        return g.to(x.dtype)
    dx = torch.empty_like(x)
    _call("b2_bias_act_bwd", _p(dy), _p(x), _p(bias), _p(dx), x.numel(), x.shape[-1], mode, _dt(x), _st())
    return dx


def swiglu_fwd(gu: torch.Tensor) -> torch.Tensor:
    Fdim = gu.shape[-1] // 2
    if not gu.is_cuda:
        g, u = gu.float()[..., :Fdim], gu.float()[..., Fdim:]
        return (F.silu(g) * u).to(gu.dtype)
    g2 = gu.reshape(-1, 2 * Fdim)
    out = torch.empty(g2.shape[0], Fdim, dtype=gu.dtype, device=gu.device)
    _call("b2_swiglu_fwd", _p(g2), _p(out), g2.shape[0], Fdim, _dt(gu), _st())
    return out.view(*gu.shape[:-1], Fdim)


def swiglu_bwd(dy: torch.Tensor, gu: torch.Tensor) -> torch.Tensor:
    Fdim = gu.shape[-1] // 2
    if not gu.is_cuda:
        g, u, d = gu.float()[..., :Fdim], gu.float()[..., Fdim:], dy.float()
        s = torch.sigmoid(g)
        return torch.cat([d * u * s * (1 + g * (1 - s)), d * g * s], -1).to(gu.dtype)
    g2, d2 = gu.reshape(-1, 2 * Fdim), dy.reshape(-1, Fdim)
    out = torch.empty_like(g2)
    _call("b2_swiglu_bwd", _p(d2), _p(g2), _p(out), g2.shape[0], Fdim, _dt(gu), _st())
    return out.view_as(gu)


def colsum_accum(dy2: torch.Tensor, out_fp32: torch.Tensor) -> torch.Tensor:
    """out_fp32[n] += sum_m dy2[m, n]"""
    if not dy2.is_cuda:
        out_fp32.add_(dy2.float().sum(0))
        return out_fp32
    _call("b2_colsum", _p(dy2), _p(out_fp32), dy2.shape[0], dy2.shape[1], dy2.stride(0), _dt(dy2), _st())
    return out_fp32


def rope_(x: torch.Tensor, cos_t: torch.Tensor, sin_t: torch.Tensor, pos: torch.Tensor, nheads: int, hd: int,
          tok_stride: int, head_stride: int, backward: bool = False) -> torch.Tensor:
    """In-place rotate-half RoPE over ``nheads`` heads of width ``hd`` starting at x's data pointer.
    x: any tensor whose first element is head 0 / dim 0 of token 0; pos: int32 [tokens]."""
    tokens = pos.numel()
    if not x.is_cuda:
        flat = x.view(-1)
        half = hd // 2
        idx_t = torch.arange(tokens)[:, None, None] * tok_stride
        idx_h = torch.arange(nheads)[None, :, None] * head_stride
        idx_d = torch.arange(half)[None, None, :]
        ia = (idx_t + idx_h + idx_d).reshape(-1)
        ib = ia + half
        a, b = flat[ia].float(), flat[ib].float()
        c = cos_t[pos.long()][:, None, :].expand(tokens, nheads, half).reshape(-1)
        s = sin_t[pos.long()][:, None, :].expand(tokens, nheads, half).reshape(-1) * (-1.0 if backward else 1.0)
        flat[ia] = (a * c - b * s).to(x.dtype)
        flat[ib] = (b * c + a * s).to(x.dtype)
        return x
    _call("b2_rope", _p(x), _p(cos_t), _p(sin_t), _p(pos), tokens, nheads, hd, tok_stride, head_stride, int(backward),
          _dt(x), _st())
    return x


def axpby(a, b, alpha: float = 1.0, beta: float = 1.0, out=None):
    if not a.is_cuda:
        r = a.float() * alpha + (b.float() * beta if b is not None else 0)
        if out is None:
            return r.to(a.dtype)
        out.copy_(r.to(a.dtype))
        return out
    if out is None:
        out = torch.empty_like(a)
    _call("b2_axpby", _p(a), _p(b), _p(out), a.numel(), float(alpha), float(beta), _dt(a), _st())
    return out


def accum_fp32_(dst: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    if not dst.is_cuda:
        dst.add_(src.float())
        return dst
    _call("b2_accum_fp32", _p(dst), _p(src), dst.numel(), _dt(src), _st())
    return dst


def cast_from_fp32_(dst: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    if not dst.is_cuda:
        dst.copy_(src.to(dst.dtype))
        return dst
    _call("b2_cast_from_fp32", _p(dst), _p(src), dst.numel(), _dt(dst), _st())
    return dst


# =============================================================================================
# cross entropy
# =============================================================================================
def xent_stats(logits2: torch.Tensor, labels: torch.Tensor, vocab_start: int = 0):
    """Per-row (max, sumexp relative to max, target logit or 0) over this vocab shard."""
    rows, V = logits2.shape
    if not logits2.is_cuda:
        lf = logits2.float()
        m = lf.max(-1).values
        s = torch.exp(lf - m[:, None]).sum(-1)
        lab = labels - vocab_start
        ok = (lab >= 0) & (lab < V)
        t = torch.where(ok, lf.gather(1, lab.clamp(0, V - 1)[:, None])[:, 0], torch.zeros_like(m))
        return m, s, t
    m = torch.empty(rows, dtype=torch.float32, device=logits2.device)
    s = torch.empty_like(m)
    t = torch.empty_like(m)
    _call("b2_xent_stats", _p(logits2), _p(labels), _p(m), _p(s), _p(t), rows, V, logits2.stride(0), vocab_start,
          _dt(logits2), _st())
    return m, s, t


def xent_finish(m, s, t, labels, ignore_index: int = -100):
    if not m.is_cuda:
        loss = torch.log(s) + m - t
        return torch.where(labels == ignore_index, torch.zeros_like(loss), loss)
    loss = torch.empty_like(m)
    _call("b2_xent_finish", _p(m), _p(s), _p(t), _p(labels), _p(loss), m.numel(), ignore_index, _st())
    return loss


def xent_bwd_(logits2, labels, gmax, gsum, row_scale: Optional[torch.Tensor], scalar_scale: float,
              vocab_start: int = 0, ignore_index: int = -100):
    """Overwrites logits2 with dlogits = (softmax - onehot) * row_scale * scalar_scale."""
    rows, V = logits2.shape
    if not logits2.is_cuda:
        p = torch.exp(logits2.float() - gmax[:, None]) / gsum[:, None]
        lab = labels - vocab_start
        ok = (lab >= 0) & (lab < V)
        onehot = torch.zeros_like(p)
        onehot[ok, lab[ok]] = 1.0
        sc = torch.where(labels == ignore_index, torch.zeros_like(gmax), torch.ones_like(gmax)) * scalar_scale
        if row_scale is not None:
            sc = sc * row_scale
        logits2.copy_(((p - onehot) * sc[:, None]).to(logits2.dtype))
        return logits2
    _call("b2_xent_bwd", _p(logits2), _p(logits2), _p(labels), _p(gmax), _p(gsum), _p(row_scale), float(scalar_scale),
          rows, V, logits2.stride(0), vocab_start, ignore_index, _dt(logits2), _st())
    return logits2


# =============================================================================================
# optimizers (flat fp32 arenas)
# =============================================================================================
def adam_(p32, m, v, g, p_lp: Optional[torch.Tensor], *, lr, beta1, beta2, eps, wd, step: int, adamw: bool = True,
          grad_scale: float = 1.0, sumsq: Optional[torch.Tensor] = None, max_norm: float = 0.0,
          found_inf: Optional[torch.Tensor] = None):
    if not p32.is_cuda:
        if found_inf is not None and float(found_inf) != 0:
            return
        coef = 1.0
        if sumsq is not None and max_norm > 0:
            coef = min(1.0, max_norm / (math.sqrt(float(sumsq)) + 1e-6))
        gr = g.float() * (grad_scale * coef)
        if not adamw:
            gr = gr + wd * p32
        m.mul_(beta1).add_(gr, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(gr, gr, value=1 - beta2)
        bc1, bc2 = 1 - beta1 ** step, 1 - beta2 ** step
        upd = (m / bc1) / ((v / bc2).sqrt() + eps)
        if adamw:
            upd = upd + wd * p32
        p32.add_(upd, alpha=-lr)
        if p_lp is not None:
            p_lp.copy_(p32.to(p_lp.dtype))
        return
    _call("b2_adam", _p(p32), _p(m), _p(v), _p(g), _p(p_lp), p32.numel(), float(lr), float(beta1), float(beta2),
          float(eps), float(wd), int(step), int(adamw), float(grad_scale), _p(sumsq), float(max_norm), _p(found_inf),
          _dt(g), _dt(p_lp) if p_lp is not None else DT_FP32, _st())


def adadelta_(p32, sq, acc, g, *, lr, rho, eps, wd=0.0):
    if not p32.is_cuda:
        gr = g + wd * p32
        sq.mul_(rho).addcmul_(gr, gr, value=1 - rho)
        d = (acc + eps).sqrt() / (sq + eps).sqrt() * gr
        acc.mul_(rho).addcmul_(d, d, value=1 - rho)
        p32.add_(d, alpha=-lr)
        return
    _call("b2_adadelta", _p(p32), _p(sq), _p(acc), _p(g), p32.numel(), float(lr), float(rho), float(eps), float(wd), _st())


def sgd_(p32, mom, g, *, lr, momentum=0.0, wd=0.0, grad_scale=1.0, first_step=False):
    if not p32.is_cuda:
        gr = g * grad_scale + wd * p32
        if first_step:
            mom.copy_(gr)
        else:
            mom.mul_(momentum).add_(gr)
        p32.add_(mom if momentum != 0 else gr, alpha=-lr)
        return
    _call("b2_sgd", _p(p32), _p(mom), _p(g), p32.numel(), float(lr), float(momentum), float(wd), float(grad_scale),
          int(first_step), _st())


_sumsq_ws = {}


def sumsq(x: torch.Tensor, out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    """out[0] (+)= sum(x^2) as an fp32 device scalar (no host sync)."""
    if out is None:
        out = torch.zeros(1, dtype=torch.float32, device=x.device)
        accumulate = False
    if not x.is_cuda:
        val = x.float().pow(2).sum()
        if accumulate:
            out.add_(val)
        else:
            out.fill_(val)
        return out
    key = (x.device.index, torch.cuda.current_stream().cuda_stream)
    if key not in _sumsq_ws:
        n = lib().b2_sumsq_scratch_floats()
        _sumsq_ws[key] = (torch.empty(n, dtype=torch.float32, device=x.device),
                          torch.zeros(1, dtype=torch.int32, device=x.device))
    scratch, counter = _sumsq_ws[key]
    _call("b2_sumsq", _p(x), x.numel(), _p(scratch), _p(counter), _p(out), int(accumulate), _dt(x), _st())
    return out


def scale_check_(x32: torch.Tensor, scale: float, found_inf: torch.Tensor):
    if not x32.is_cuda:
        x32.mul_(scale)
        if not torch.isfinite(x32).all():
            found_inf.fill_(1.0)
        return
    _call("b2_scale_check", _p(x32), x32.numel(), float(scale), _p(found_inf), _st())


# =============================================================================================
# embedding
# =============================================================================================
def embedding_fwd(tok, pos, wte, wpe, vocab_start: int = 0):
    """tok/pos: int64 [tokens]; returns [tokens, H]"""
    tokens, H = tok.numel(), wte.shape[1]
    if not wte.is_cuda:
        idx = tok - vocab_start
        ok = (idx >= 0) & (idx < wte.shape[0])
        out = torch.where(ok[:, None], wte.float()[idx.clamp(0, wte.shape[0] - 1)], torch.zeros(1))
        if wpe is not None:
            out = out + wpe.float()[pos]
        return out.to(wte.dtype)
    out = torch.empty(tokens, H, dtype=wte.dtype, device=wte.device)
    _call("b2_embedding_fwd", _p(tok), _p(pos), _p(wte), _p(wpe), _p(out), tokens, H, vocab_start, wte.shape[0],
          _dt(wte), _st())
    return out


def embedding_bwd_(dy2, tok, pos, dwte32, dwpe32, vocab_start: int = 0):
    tokens, H = dy2.shape
    if not dy2.is_cuda:
        idx = tok - vocab_start
        ok = (idx >= 0) & (idx < dwte32.shape[0])
        dwte32.index_add_(0, idx[ok], dy2.float()[ok])
        if dwpe32 is not None:
            dwpe32.index_add_(0, pos, dy2.float())
        return
    _call("b2_embedding_bwd", _p(dy2), _p(tok), _p(pos), _p(dwte32), _p(dwpe32), tokens, H, vocab_start,
          dwte32.shape[0], _dt(dy2), _st())


# =============================================================================================
# attention.  q/k/v/o are 4-D views [seq, batch, heads, D] with D contiguous (any strides otherwise)
# =============================================================================================
def attn_dropout_mask(B: int, Hh: int, S: int, p_drop: float, seed: int, device="cpu") -> torch.Tensor:
    """[B, H, S(query), S(key)] multiplier (0 or 1/keep) - the decision rule of csrc/attention.cu restated: one hash per PAIR of
    adjacent keys, h = lowbias32(seed32 + q*CQ + (k>>1)*CK + (b*H+h)*CB) (mod 2^32); key k is kept iff its 16-bit half of h (low
    half for even k, high half for odd k) >= round(p * 65536); kept values are scaled by 65536 / (65536 - threshold)."""
    M = 0xFFFFFFFF
    seed32 = (seed ^ (seed >> 32)) & M
    q = torch.arange(S, dtype=torch.int64, device=device)[:, None]
    k = torch.arange(S, dtype=torch.int64, device=device)[None, :]
    bh = torch.arange(B * Hh, dtype=torch.int64, device=device).view(B, Hh, 1, 1)
    x = (seed32 + q * 0x9E3779B1 + (k >> 1) * 0x85EBCA77 + bh * 0xC2B2AE3D) & M
    x = x ^ (x >> 16); x = (x * 0x7FEB352D) & M
    x = x ^ (x >> 15); x = (x * 0x846CA68B) & M
    x = x ^ (x >> 16)
    half = torch.where((k & 1) == 1, x >> 16, x & 0xFFFF)
    t = max(0, min(int(float(p_drop) * 65536.0 + 0.5), 65535))
    return (half >= t).float() * (65536.0 / (65536 - t))


def _attn_ref(q, k, v, scale, causal, p_drop: float = 0.0, seed: int = 0):
    qf, kf, vf = (t.float().permute(1, 2, 0, 3) for t in (q, k, v))  # [b, h, s, d]
    s = torch.matmul(qf, kf.transpose(-1, -2)) * scale
    if causal:
        S = s.shape[-1]
        mask = torch.ones(S, S, dtype=torch.bool, device=s.device).tril()
        s = s.masked_fill(~mask, float("-inf"))
    lse = torch.logsumexp(s, -1)
    p = torch.exp(s - lse[..., None])
    mk = attn_dropout_mask(s.shape[0], s.shape[1], s.shape[-1], p_drop, seed, s.device) if p_drop > 0.0 else None
    o = torch.matmul(p * mk if mk is not None else p, vf)
    return o, lse, p, mk


def attn_fwd(q, k, v, scale: float, causal: bool, out: Optional[torch.Tensor] = None, p_drop: float = 0.0, seed: int = 0):
    """Returns (o [seq, batch, heads, D] contiguous, lse [batch, heads, seq] fp32).  p_drop > 0: attention dropout on
    the softmax output (counter-based mask from `seed`; give the same pair to attn_bwd)."""
    S, B, Hh, D = q.shape
    if out is None:
        out = torch.empty((S, B, Hh, D), dtype=q.dtype, device=q.device)
    if not q.is_cuda:
        o, lse, _, _ = _attn_ref(q, k, v, scale, causal, p_drop, seed)
        out.copy_(o.permute(2, 0, 1, 3).to(q.dtype))
        return out, lse.contiguous()
    assert q.dtype == torch.bfloat16, "flash attention kernel is bf16"
    for t in (q, k, v, out):
        assert t.stride(3) == 1
    lse = torch.empty((B, Hh, S), dtype=torch.float32, device=q.device)
    _call("b2_attn_fwd", _p(q), _p(k), _p(v), _p(out), _p(lse), S, B, Hh, D,
          q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
          v.stride(0), v.stride(1), v.stride(2), out.stride(0), out.stride(1), out.stride(2),
          float(scale), int(causal), float(p_drop), int(seed) & 0xFFFFFFFFFFFFFFFF, _st())
    return out, lse


def attn_bwd(q, k, v, o, dout, lse, scale: float, causal: bool, dq: torch.Tensor, dk: torch.Tensor, dv: torch.Tensor,
             p_drop: float = 0.0, seed: int = 0):
    """Writes dq/dk/dv (4-D views like q/k/v, typically slices of one packed dQKV buffer)."""
    S, B, Hh, D = q.shape
    if not q.is_cuda:
        _, _, p, mk = _attn_ref(q, k, v, scale, causal, p_drop, seed)
        qf, kf, vf, dof = (t.float().permute(1, 2, 0, 3) for t in (q, k, v, dout))
        of = o.float().permute(1, 2, 0, 3)
        dvf = torch.matmul((p * mk if mk is not None else p).transpose(-1, -2), dof)
        dp = torch.matmul(dof, vf.transpose(-1, -2))
        if mk is not None:
            dp = dp * mk
        delta = (dof * of).sum(-1, keepdim=True)
        ds = p * (dp - delta) * scale
        dqf = torch.matmul(ds, kf)
        dkf = torch.matmul(ds.transpose(-1, -2), qf)
        dq.copy_(dqf.permute(2, 0, 1, 3).to(dq.dtype))
        dk.copy_(dkf.permute(2, 0, 1, 3).to(dk.dtype))
        dv.copy_(dvf.permute(2, 0, 1, 3).to(dv.dtype))
        return
    delta = torch.empty((B, Hh, S), dtype=torch.float32, device=q.device)
    dq32 = torch.zeros((S, B, Hh, D), dtype=torch.float32, device=q.device)
    _call("b2_attn_bwd", _p(q), _p(k), _p(v), _p(o), _p(dout), _p(lse), _p(delta), _p(dq32), _p(dk), _p(dv), S, B, Hh, D,
          q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
          v.stride(0), v.stride(1), v.stride(2), o.stride(0), o.stride(1), o.stride(2),
          dout.stride(0), dout.stride(1), dout.stride(2), dk.stride(0), dk.stride(1), dk.stride(2),
          dv.stride(0), dv.stride(1), dv.stride(2), float(scale), int(causal), float(p_drop), int(seed) & 0xFFFFFFFFFFFFFFFF, _st())
    # fp32 dQ accumulator -> the (usually strided) 16-bit dq slice
    if dq.stride(3) == 1 and dq.stride(1) == Hh * dq.stride(2) and dq.stride(0) == B * dq.stride(1) and dq.stride(2) % 8 == 0 and D % 8 == 0:
        _call("b2_cast_rows_from_fp32", _p(dq), dq.stride(2), _p(dq32), S * B * Hh, D, _dt(dq), _st())
    else:
        dq.copy_(dq32)


# =============================================================================================
# BatchNorm2d (NCHW), optional fused ReLU, optional cross-rank statistics (SyncBatchNorm)     (csrc/batchnorm.cu)
# =============================================================================================
def _f32(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    return t if (t is None or t.dtype == torch.float32) else t.float()


def _bn_sync(t: torch.Tensor, group):
    if group is not None and torch.distributed.is_initialized() and torch.distributed.get_world_size(group) > 1:
        torch.distributed.all_reduce(t, group=group)
    return t


def bn_train_fwd(x, gamma, beta, running_mean, running_var, eps: float, momentum: float, relu: bool = False, group=None):
    """x [N, C, H, W] -> (y, mean[C], rstd[C], count): batch statistics over N*H*W (x the group's ranks when `group` is given);
    running statistics are updated in place like torch (unbiased variance)."""
    Nn, C, H, W = x.shape
    HW = H * W
    if not x.is_cuda:
        xf = x.float()
        sums = torch.stack([xf.sum((0, 2, 3)), (xf * xf).sum((0, 2, 3))], 1).double()
        cnt = torch.tensor([float(Nn * HW)], dtype=torch.float64)
        _bn_sync(sums, group); _bn_sync(cnt, group)
        count = float(cnt)
        mean = (sums[:, 0] / count)
        var = (sums[:, 1] / count - mean * mean).clamp_(min=0)
        rstd = torch.rsqrt(var.float() + eps)
        mean = mean.float()
        g = gamma.float() if gamma is not None else torch.ones(C)
        b = beta.float() if beta is not None else torch.zeros(C)
        y = (xf - mean[None, :, None, None]) * (rstd * g)[None, :, None, None] + b[None, :, None, None]
        if relu:
            y = torch.relu(y)
        if running_mean is not None:
            unb = var.float() * (count / (count - 1)) if count > 1 else var.float()
            running_mean.mul_(1 - momentum).add_(momentum * mean)
            running_var.mul_(1 - momentum).add_(momentum * unb)
        return y.to(x.dtype), mean, rstd, count
    x = x.contiguous()
    dev = x.device
    # the kernels take fp32 parameters and statistics; a model cast to bf16 carries 16-bit ones: widen, and narrow back after
    rm_in, rv_in = running_mean, running_var
    gamma, beta = _f32(gamma), _f32(beta)
    running_mean, running_var = _f32(running_mean), _f32(running_var)
    sums = torch.zeros(C, 2, dtype=torch.float64, device=dev)
    _call("b2_bn_reduce", _p(x), _p(sums), Nn, C, HW, _dt(x), _st())
    count = float(Nn * HW)
    if group is not None and torch.distributed.is_initialized() and torch.distributed.get_world_size(group) > 1:
        _bn_sync(sums, group)
        count *= torch.distributed.get_world_size(group)          # equal per-rank batch shapes (DDP)
    mean, rstd = torch.empty(C, dtype=torch.float32, device=dev), torch.empty(C, dtype=torch.float32, device=dev)
    scale, shift = torch.empty_like(mean), torch.empty_like(mean)
    _call("b2_bn_finalize", _p(sums), count, _p(gamma), _p(beta), float(eps), float(momentum), _p(running_mean), _p(running_var), _p(mean), _p(rstd),
          _p(scale), _p(shift), C, _st())
    if rm_in is not None and rm_in.dtype != torch.float32:
        rm_in.copy_(running_mean); rv_in.copy_(running_var)
    y = torch.empty_like(x)
    _call("b2_bn_apply", _p(x), _p(scale), _p(shift), _p(y), Nn, C, HW, int(relu), _dt(x), _st())
    return y, mean, rstd, count


def bn_eval_fwd(x, gamma, beta, running_mean, running_var, eps: float, relu: bool = False):
    Nn, C, H, W = x.shape
    if not x.is_cuda:
        y = F.batch_norm(x.float(), running_mean, running_var, gamma.float() if gamma is not None else None, beta.float() if beta is not None else None,
                         False, 0.0, eps)
        return (torch.relu(y) if relu else y).to(x.dtype)
    x = x.contiguous()
    gamma, beta, running_mean, running_var = _f32(gamma), _f32(beta), _f32(running_mean), _f32(running_var)
    scale, shift = torch.empty(C, dtype=torch.float32, device=x.device), torch.empty(C, dtype=torch.float32, device=x.device)
    _call("b2_bn_eval_coeffs", _p(running_mean), _p(running_var), _p(gamma), _p(beta), float(eps), _p(scale), _p(shift), C, _st())
    y = torch.empty_like(x)
    _call("b2_bn_apply", _p(x), _p(scale), _p(shift), _p(y), Nn, C, H * W, int(relu), _dt(x), _st())
    return y


def bn_bwd(dy, x, y, gamma, mean, rstd, count: float, relu: bool = False, group=None):
    """-> (dx, dgamma[C] fp32, dbeta[C] fp32); `count` = elements per channel over the whole group (from bn_train_fwd)"""
    Nn, C, H, W = x.shape
    HW = H * W
    if not x.is_cuda:
        g = dy.float()
        if relu:
            g = g * (y > 0)
        xhat = (x.float() - mean[None, :, None, None]) * rstd[None, :, None, None]
        sums = torch.stack([g.sum((0, 2, 3)), (g * xhat).sum((0, 2, 3))], 1).double()
        dbeta, dgamma = sums[:, 0].float().clone(), sums[:, 1].float().clone()      # local sums: DDP averages parameter gradients
        _bn_sync(sums, group)
        mg, mgx = (sums[:, 0] / count).float(), (sums[:, 1] / count).float()
        gm = gamma.float() if gamma is not None else torch.ones(C)
        dx = (gm * rstd)[None, :, None, None] * (g - mg[None, :, None, None] - xhat * mgx[None, :, None, None])
        return dx.to(x.dtype), dgamma, dbeta
    dy, x = dy.contiguous(), x.contiguous()
    gamma = _f32(gamma)
    if dy.dtype != x.dtype:
        dy = dy.to(x.dtype)
    sums = torch.zeros(C, 2, dtype=torch.float64, device=x.device)
    _call("b2_bn_bwd_reduce", _p(dy), _p(x), _p(y), _p(mean), _p(rstd), _p(sums), Nn, C, HW, int(relu), _dt(x), _st())
    dbeta, dgamma = sums[:, 0].float(), sums[:, 1].float()
    _bn_sync(sums, group)
    dx = torch.empty_like(x)
    _call("b2_bn_bwd_apply", _p(dy), _p(x), _p(y), _p(mean), _p(rstd), _p(gamma), _p(sums), float(count), _p(dx), Nn, C, HW, int(relu), _dt(x), _st())
    return dx, dgamma, dbeta


# =============================================================================================
# fp32 MNIST kernels: valid 3x3 convolution (+bias+ReLU), 2x2 max-pool, log-softmax + NLL   (csrc/conv.cu)
# =============================================================================================
def conv2d_fwd(x, w, bias, relu: bool):
    """x [B, Cin, H, W] fp32, w [Cout, Cin, K, K] -> relu?(conv(x, w) + bias), valid, stride 1"""
    if not x.is_cuda:
        y = F.conv2d(x, w, bias)
        return F.relu(y) if relu else y
    assert x.dtype == torch.float32 and w.dtype == torch.float32 and (bias is None or bias.dtype == torch.float32), "conv kernels are fp32"
    x, w = x.contiguous(), w.contiguous()
    B, Cin, H, W = x.shape
    Cout, _, K, _ = w.shape
    y = torch.empty((B, Cout, H - K + 1, W - K + 1), dtype=torch.float32, device=x.device)
    _call("b2_conv2d_fwd", _p(x), _p(w), _p(bias), _p(y), B, Cin, H, W, Cout, K, int(relu), _st())
    return y


def conv2d_bwd(dy, y, x, w, relu: bool, need_dx: bool, has_bias: bool = True):
    """-> (dx | None, dw, db | None); y is the forward output (ReLU mask)"""
    if not x.is_cuda:
        g = dy * (y > 0) if relu else dy
        dx = torch.nn.grad.conv2d_input(x.shape, w, g) if need_dx else None
        dw = torch.nn.grad.conv2d_weight(x, w.shape, g)
        return dx, dw, (g.sum((0, 2, 3)) if has_bias else None)
    assert dy.dtype == torch.float32 and x.dtype == torch.float32 and w.dtype == torch.float32, "conv kernels are fp32"
    dy, y, x, w = dy.contiguous(), y.contiguous(), x.contiguous(), w.contiguous()
    B, Cin, H, W = x.shape
    Cout, _, K, _ = w.shape
    dx = torch.empty_like(x) if need_dx else None
    dw = torch.empty_like(w)
    db = torch.empty(Cout, dtype=torch.float32, device=x.device) if has_bias else None
    _call("b2_conv2d_bwd", _p(dy), _p(y), _p(x), _p(w), _p(dx), _p(dw), _p(db), B, Cin, H, W, Cout, K, int(relu), 0, _st())
    return dx, dw, db


def maxpool2_fwd(x):
    """x [..., H, W] fp32 -> (y [..., H//2, W//2], idx uint8 same shape as y)"""
    H, W = x.shape[-2:]
    if not x.is_cuda:
        y, ind = F.max_pool2d(x.reshape(-1, 1, H, W), 2, return_indices=True)
        iy, ix = (ind // W) % 2, (ind % W) % 2
        return y.view(*x.shape[:-2], H // 2, W // 2), (iy * 2 + ix).to(torch.uint8).view(*x.shape[:-2], H // 2, W // 2)
    assert x.dtype == torch.float32, "max-pool kernel is fp32"
    x = x.contiguous()
    y = torch.empty((*x.shape[:-2], H // 2, W // 2), dtype=torch.float32, device=x.device)
    idx = torch.empty(y.shape, dtype=torch.uint8, device=x.device)
    _call("b2_maxpool2_fwd", _p(x), _p(y), _p(idx), x.numel() // (H * W), H, W, _st())
    return y, idx


def maxpool2_bwd(dy, idx, in_shape):
    H, W = in_shape[-2:]
    if not dy.is_cuda:
        dx = torch.zeros(in_shape, dtype=dy.dtype)
        Ho, Wo = H // 2, W // 2
        v = dx[..., :2 * Ho, :2 * Wo].reshape(*in_shape[:-2], Ho, 2, Wo, 2).permute(*range(len(in_shape) - 2), -4, -2, -3, -1)
        oh = torch.nn.functional.one_hot(idx.long(), 4).view(*idx.shape, 2, 2).to(dy.dtype)
        v.copy_(oh * dy[..., None, None])
        return dx
    dy = dy.contiguous()
    dx = torch.empty(in_shape, dtype=torch.float32, device=dy.device)
    _call("b2_maxpool2_bwd", _p(dy), _p(idx), _p(dx), dx.numel() // (H * W), H, W, _st())
    return dx


def log_softmax_nll_fwd(x, target=None):
    """x [rows, C <= 32] fp32 -> (logp, loss[rows] | None)"""
    if not x.is_cuda:
        lp = F.log_softmax(x, dim=1)
        return lp, (F.nll_loss(lp, target, reduction="none") if target is not None else None)
    assert x.dtype == torch.float32 and x.dim() == 2 and x.shape[1] <= 32, "log-softmax kernel: fp32, at most 32 classes"
    x = x.contiguous()
    lp = torch.empty_like(x)
    loss = torch.empty(x.shape[0], dtype=torch.float32, device=x.device) if target is not None else None
    _call("b2_log_softmax_nll_fwd", _p(x), _p(target), _p(lp), _p(loss), x.shape[0], x.shape[1], _st())
    return lp, loss


def log_softmax_nll_bwd(logp, dlogp=None, target=None, gloss=None, gscalar: float = 1.0):
    """gradient wrt the logits of  sum(dlogp * logp) + sum(gloss * gscalar * nll(logp, target))"""
    if not logp.is_cuda:
        d = dlogp.clone() if dlogp is not None else torch.zeros_like(logp)
        if target is not None:
            g = (gloss if gloss is not None else torch.ones(logp.shape[0])) * gscalar
            d[torch.arange(logp.shape[0]), target] -= g
        return d - logp.exp() * d.sum(1, keepdim=True)
    dx = torch.empty_like(logp)
    _call("b2_log_softmax_nll_bwd", _p(logp), _p(dlogp.contiguous() if dlogp is not None else None), _p(target),
          _p(gloss.contiguous() if gloss is not None else None), float(gscalar), _p(dx), logp.shape[0], logp.shape[1], _st())
    return dx


# =============================================================================================
# Stand-alone scale + mask + softmax (unfused attention path; Megatron fused_kernels scaled_*_softmax)
# =============================================================================================
def softmax_ref(x, mask, scale: float, causal: bool):
    """fp32 restatement: x [b, h, sq, sk]; mask bool/uint8 [b | 1, 1, sq, sk] (True = masked -> -10000 after scaling)."""
    s = x.float() * scale
    if mask is not None:
        s = s.masked_fill(mask.bool(), -10000.0)
    if causal:
        sq, sk = x.shape[-2], x.shape[-1]
        keep = torch.ones(sq, sk, dtype=torch.bool, device=x.device).tril(diagonal=sk - sq)
        s = s.masked_fill(~keep, float("-inf"))
    return torch.softmax(s, dim=-1)


def softmax_fwd(x: torch.Tensor, mask: Optional[torch.Tensor], scale: float, causal: bool) -> torch.Tensor:
    """x: [b, h, sq, sk] contiguous scores -> probabilities (same dtype)."""
    assert x.dim() == 4
    if not x.is_cuda:
        return softmax_ref(x, mask, scale, causal).to(x.dtype)
    x = x.contiguous()
    b, h, sq, sk = x.shape
    m8 = None
    if mask is not None:
        assert mask.shape[-2:] == (sq, sk) and mask.shape[1] == 1 and mask.shape[0] in (1, b)
        m8 = mask.to(torch.uint8).contiguous()
    y = torch.empty_like(x)
    _call("b2_softmax_fwd", _p(x), _p(m8), _p(y), b * h * sq, sq, sk, h, m8.shape[0] if m8 is not None else 1, float(scale), int(causal),
          _dt(x), _st())
    return y


def softmax_bwd(dy: torch.Tensor, y: torch.Tensor, scale: float) -> torch.Tensor:
    if not y.is_cuda:
        yf, df = y.float(), dy.float()
        return (scale * yf * (df - (df * yf).sum(-1, keepdim=True))).to(y.dtype)
    dy, y = dy.contiguous(), y.contiguous()
    dx = torch.empty_like(y)
    _call("b2_softmax_bwd", _p(dy), _p(y), _p(dx), y.numel() // y.shape[-1], y.shape[-1], float(scale), _dt(y), _st())
    return dx


# =============================================================================================
# Swin window attention.  qkv [Bw, T, 3*C] (last dim = [3][heads][hd]), bias fp32 [heads, T, T], mask fp32 [nW, T, T] | None
# =============================================================================================
def _wattn_ref(qkv, bias, mask, heads, scale):
    Bw, T, C3 = qkv.shape
    hd = C3 // 3 // heads
    q, k, v = qkv.float().view(Bw, T, 3, heads, hd).permute(2, 0, 3, 1, 4)          # [Bw, heads, T, hd]
    s = torch.matmul(q * scale, k.transpose(-1, -2)) + bias.float()[None]
    if mask is not None:
        nW = mask.shape[0]
        s = (s.view(Bw // nW, nW, heads, T, T) + mask.float()[None, :, None]).view(Bw, heads, T, T)
    lse = torch.logsumexp(s, -1)
    p = torch.exp(s - lse[..., None])
    return q, k, v, p, lse


def window_attn_fwd(qkv, bias, mask, heads: int, scale: float):
    """Returns (out [Bw, T, C], lse [Bw, heads, T])."""
    Bw, T, C3 = qkv.shape
    C = C3 // 3
    if not qkv.is_cuda:
        q, k, v, p, lse = _wattn_ref(qkv, bias, mask, heads, scale)
        o = torch.matmul(p, v).transpose(1, 2).reshape(Bw, T, C)
        return o.to(qkv.dtype), lse
    assert qkv.is_contiguous() and bias.dtype == torch.float32 and bias.is_contiguous()
    out = torch.empty((Bw, T, C), dtype=qkv.dtype, device=qkv.device)
    lse = torch.empty((Bw, heads, T), dtype=torch.float32, device=qkv.device)
    nW = mask.shape[0] if mask is not None else 1
    _call("b2_window_attn_fwd", _p(qkv), _p(bias), _p(mask), _p(out), _p(lse), Bw, nW, heads, T, C // heads, float(scale), _dt(qkv), _st())
    return out, lse


def window_attn_bwd(qkv, bias, mask, out, dout, lse, heads: int, scale: float):
    """Returns (dqkv like qkv, dbias fp32 [heads, T, T])."""
    Bw, T, C3 = qkv.shape
    C = C3 // 3
    if not qkv.is_cuda:
        q, k, v, p, _ = _wattn_ref(qkv, bias, mask, heads, scale)
        do = dout.float().view(Bw, T, heads, C // heads).transpose(1, 2)
        of = out.float().view(Bw, T, heads, C // heads).transpose(1, 2)
        dv = torch.matmul(p.transpose(-1, -2), do)
        dp = torch.matmul(do, v.transpose(-1, -2))
        ds = p * (dp - (do * of).sum(-1, keepdim=True))
        dq = torch.matmul(ds, k) * scale
        dk = torch.matmul(ds.transpose(-1, -2), q) * scale
        dqkv = torch.stack([dq, dk, dv], 0).permute(1, 3, 0, 2, 4).reshape(Bw, T, C3)
        return dqkv.to(qkv.dtype), ds.sum(0)
    dqkv = torch.empty_like(qkv)
    dbias = torch.zeros((heads, T, T), dtype=torch.float32, device=qkv.device)
    nW = mask.shape[0] if mask is not None else 1
    _call("b2_window_attn_bwd", _p(qkv), _p(bias), _p(mask), _p(out), _p(dout.contiguous()), _p(lse), _p(dqkv), _p(dbias), Bw, nW, heads, T,
          C // heads, float(scale), _dt(qkv), _st())
    return dqkv, dbias


def patchify(x, p: int):
    """x [B, C, H, W] -> [B * (H/p) * (W/p), C * p * p]: non-overlapping p x p patches as GEMM rows, columns ordered (c, i, j) like a
    Conv2d weight viewed [out, C*p*p].  p == 4 with 16-bit elements on the GPU is one coalesced pass (csrc/window_attn.cu: patchify4)."""
    B, C, H, W = x.shape
    if x.is_cuda and p == 4 and x.dtype in (torch.bfloat16, torch.float16) and H % 4 == 0 and W % 4 == 0:
        x = x.contiguous()
        out = torch.empty((B * (H // 4) * (W // 4), C * 16), dtype=x.dtype, device=x.device)
        _call("b2_patchify4", _p(x), _p(out), B, C, H, W, _st())
        return out
    return x.view(B, C, H // p, p, W // p, p).permute(0, 2, 4, 1, 3, 5).reshape(B * (H // p) * (W // p), C * p * p).contiguous()


def rows_scale_add(x, res, scale, B: int, H: int, W: int, ws: int, shift, mode: int):
    """out = res + scale[sample] * map(x) over B * H * W rows of C elements (res / scale may be None: 0 / 1).
    mode 0: map = identity (out has x's shape); 1: window partition of x (x [B, H, W, C] -> out [B*nW, ws*ws, C]);
    2: window reverse (x [B*nW, ws*ws, C] -> out [B, H, W, C]).  scale: fp32 [B].  The stochastic-depth residual add with
    Swin's window_reverse folded in (csrc/window_attn.cu: rows_scale_add_kernel); its backward is mode 1 <-> 2 without res."""
    C = x.shape[-1]
    sh, sw = int(shift[0]), int(shift[1])
    oshape = tuple(x.shape) if mode == 0 else ((B * (H // ws) * (W // ws), ws * ws, C) if mode == 1 else (B, H, W, C))
    if x.is_cuda and x.dtype in (torch.bfloat16, torch.float16) and C % 8 == 0:
        x = x.contiguous()
        res = res.contiguous() if res is not None else None
        assert x.numel() == B * H * W * C and (res is None or (res.dtype == x.dtype and res.numel() == x.numel()))
        assert scale is None or (scale.dtype == torch.float32 and scale.numel() == B and scale.is_contiguous())
        out = torch.empty(oshape, dtype=x.dtype, device=x.device)
        _call("b2_rows_scale_add", _p(x), _p(res), _p(scale), _p(out), B, H, W, C, ws if mode else 1, sh, sw, int(mode), _dt(x), _st())
        return out
    y = x if mode == 0 else window_shuffle(x, B, H, W, ws, (sh, sw), mode == 2)
    y = y.reshape(B, -1).float()
    if scale is not None:
        y = y * scale.float().reshape(B, 1)
    if res is not None:
        y = y + res.reshape(B, -1).float()
    return y.to(x.dtype).reshape(oshape)


def window_shuffle(x, B: int, H: int, W: int, ws: int, shift, reverse: bool):
    """reverse=False: x [B, H, W, C] -> [B*nW, ws*ws, C] (roll by -shift, then window partition);
    reverse=True : x [B*nW, ws*ws, C] -> [B, H, W, C] (un-partition, then roll by +shift)."""
    C = x.shape[-1]
    sh, sw = int(shift[0]), int(shift[1])
    if x.is_cuda and x.dtype in (torch.bfloat16, torch.float16) and C % 8 == 0:
        x = x.contiguous()
        out = torch.empty((B * (H // ws) * (W // ws), ws * ws, C) if not reverse else (B, H, W, C), dtype=x.dtype, device=x.device)
        _call("b2_window_shuffle", _p(x), _p(out), B, H, W, C, ws, sh, sw, int(reverse), _st())
        return out
    if not reverse:
        if sh or sw:
            x = torch.roll(x, shifts=(-sh, -sw), dims=(1, 2))
        return x.reshape(B, H // ws, ws, W // ws, ws, C).permute(0, 1, 3, 2, 4, 5).reshape(-1, ws * ws, C)
    x = x.reshape(B, H // ws, W // ws, ws, ws, C).permute(0, 1, 3, 2, 4, 5).reshape(B, H, W, C)
    if sh or sw:
        x = torch.roll(x, shifts=(sh, sw), dims=(1, 2))
    return x
