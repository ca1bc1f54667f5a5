"""Timing experiment: attention backward with the dQ reduce-add skipped (results wrong on purpose)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes
import torch
from b200dist.ops import native as N
dev = "cuda"
S, B, H, D = 1024, 12, 12, 64
qkv = torch.randn(S, B, H, 3 * D, device=dev, dtype=torch.bfloat16)
q, k, v = qkv[..., :D], qkv[..., D:2 * D], qkv[..., 2 * D:]
o, lse = N.attn_fwd(q, k, v, 0.125, True)
do = torch.randn_like(o); dqkv = torch.empty_like(qkv)
L = N.lib(); L.b2_attn_set_debug.argtypes = [ctypes.c_int]; L.b2_attn_set_debug.restype = None
def t(bits):
    L.b2_attn_set_debug(bits)
    f = lambda: N.attn_bwd(q, k, v, o, do, lse, 0.125, True, dqkv[..., :D], dqkv[..., D:2 * D], dqkv[..., 2 * D:])
    for _ in range(3): f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s): f()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(10): f()
    torch.cuda.synchronize(); g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); g.replay(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / 10
for bits in (0, 1, 2, 3, 0):
    print(f"attn bwd (delta + fill + kernel + cast), dbg bits {bits}: {t(bits):.4f} ms", flush=True)
L.b2_attn_set_debug(0)
