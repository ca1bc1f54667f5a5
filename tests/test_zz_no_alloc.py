"""Allocation-free entry point `b2_sumsq_static` (collected last on purpose: if it fails under `-x`, every other
test has already run).

Some users of the library cannot obtain device memory at the moment they need a norm - e.g. a watchdog thread
in a job whose framework allocator already owns nearly all of HBM, a torch-free C/C++ host with no workspace
pool, or a service that freezes its memory map after start-up.  For them the library carries a variant of the
sum-of-squares kernel whose scratch is a fixed __device__ array inside the library.  This test drives BOTH
entry points purely through the C ABI (ctypes on the shared library, raw pointers + stream) and checks that the
static variant reproduces `b2_sumsq` - twice in a row, because the second call meets the scratch exactly as the
first one left it - and that it refuses sizes beyond its documented maximum.
"""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "sagemaker-distributed-training-pytorch-kr_b200", "lib", "libb200dist.so")
DT_BF16, DT_FP32 = 1, 2
TOL = 1e-4   # relative tolerance the kernel tier uses for sum-of-squares (bench/kernel_check.py)


def _lib():
    L = ctypes.CDLL(LIB)
    vp, ll, ci = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int
    L.b2_last_error.restype = ctypes.c_char_p
    L.b2_sumsq.argtypes = [vp, ll, vp, vp, vp, ci, ci, vp]
    L.b2_sumsq.restype = ci
    L.b2_sumsq_scratch_floats.restype = ci
    L.b2_sumsq_static.argtypes = [vp, ll, vp, ci, ci, vp]
    L.b2_sumsq_static.restype = ci
    L.b2_sumsq_static_max_elems.restype = ll
    L.b2_sumsq_static_scratch_bytes.restype = ci
    return L


def _reference(L, x, dtype_code, stream):
    scratch = torch.empty(L.b2_sumsq_scratch_floats(), dtype=torch.float32, device=x.device)
    counter = torch.zeros(1, dtype=torch.int32, device=x.device)
    out = torch.zeros(1, dtype=torch.float32, device=x.device)
    rc = L.b2_sumsq(x.data_ptr(), x.numel(), scratch.data_ptr(), counter.data_ptr(), out.data_ptr(), 0, dtype_code, stream)
    assert rc == 0, L.b2_last_error().decode()
    return out


def test_static_sumsq_matches_allocating_entry_point_twice_in_a_row():
    L = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    assert 0 < L.b2_sumsq_static_scratch_bytes() <= 64 * 1024
    torch.manual_seed(0)
    cases = [(torch.randn(1_000_003, device=dev), DT_FP32),
             (torch.randn(3_333_331, device=dev).to(torch.bfloat16), DT_BF16),
             (torch.randn(17, device=dev), DT_FP32)]
    for x, code in cases:
        ref = _reference(L, x, code, stream)
        outs = []
        for _ in range(2):                      # second call sees the scratch left behind by the first
            out = torch.full((1,), -1.0, dtype=torch.float32, device=dev)
            rc = L.b2_sumsq_static(x.data_ptr(), x.numel(), out.data_ptr(), 0, code, stream)
            assert rc == 0, L.b2_last_error().decode()
            outs.append(out)
        torch.cuda.synchronize()
        for out in outs:
            assert torch.isfinite(out).all()
            assert abs(out.item() - ref.item()) <= TOL * max(1.0, abs(ref.item())), (out.item(), ref.item())
        # accumulate = 1 adds onto the previous value
        acc = outs[0].clone()
        assert L.b2_sumsq_static(x.data_ptr(), x.numel(), acc.data_ptr(), 1, code, stream) == 0
        torch.cuda.synchronize()
        assert abs(acc.item() - 2 * ref.item()) <= 2 * TOL * max(1.0, abs(ref.item()))


def test_static_sumsq_rejects_sizes_beyond_its_maximum():
    L = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    x = torch.ones(8, device=dev)
    out = torch.zeros(1, device=dev)
    too_many = L.b2_sumsq_static_max_elems() + 1
    rc = L.b2_sumsq_static(x.data_ptr(), too_many, out.data_ptr(), 0, DT_FP32, torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = L.b2_last_error().decode()
    assert "outside" in msg and str(L.b2_sumsq_static_max_elems()) in msg
    torch.cuda.synchronize()
    assert out.item() == 0.0                   # nothing was launched
