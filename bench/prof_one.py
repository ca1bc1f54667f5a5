"""Tiny driver for ncu captures of one kernel at the flagship shape:  python bench/prof_one.py attn_fwd|attn_bwd|gemm|wgrad|wattn"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dist.ops import native as N
what = sys.argv[1] if len(sys.argv) > 1 else "attn_fwd"
dev = "cuda"
torch.manual_seed(0)
S, B, H, D = 1024, 12, 12, 64
qkv = torch.randn(S, B, H, 3 * D, device=dev, dtype=torch.bfloat16)
q, k, v = qkv[..., :D], qkv[..., D:2 * D], qkv[..., 2 * D:]
if what.startswith("attn"):
    o, lse = N.attn_fwd(q, k, v, 0.125, True)
    do = torch.randn_like(o); dqkv = torch.empty_like(qkv)
    for _ in range(2 if what == "attn_both" else 4):
        if what == "attn_both":
            N.attn_fwd(q, k, v, 0.125, True, o)
            N.attn_bwd(q, k, v, o, do, lse, 0.125, True, dqkv[..., :D], dqkv[..., D:2 * D], dqkv[..., 2 * D:])
        elif what == "attn_fwd":
            N.attn_fwd(q, k, v, 0.125, True, o)
        else:
            N.attn_bwd(q, k, v, o, do, lse, 0.125, True, dqkv[..., :D], dqkv[..., D:2 * D], dqkv[..., 2 * D:])
elif what in ("gemm", "gemm_pair", "gemm_1cta"):
    N.set_gemm_pair_mode(2 if what == "gemm_pair" else 0)
    a = torch.randn(12288, 768, device=dev, dtype=torch.bfloat16); b = torch.randn(3072, 768, device=dev, dtype=torch.bfloat16)
    out = torch.empty(12288, 3072, device=dev, dtype=torch.bfloat16)
    for _ in range(4):
        N.gemm(a, b, out=out)
elif what == "wgrad":
    dy = torch.randn(12288, 3072, device=dev, dtype=torch.bfloat16); x = torch.randn(12288, 768, device=dev, dtype=torch.bfloat16)
    acc = torch.zeros(3072, 768, device=dev)
    for _ in range(4):
        N.gemm(dy, x, a_mn=True, b_mn=True, out=acc, accumulate=True)
elif what == "wattn":      # Swin-B stage 3 at batch 64: 256 windows x 16 heads, shifted (masked) block
    Bw, heads, nW = 256, 16, 4
    wqkv = torch.randn(Bw, 49, 3 * heads * 32, device=dev, dtype=torch.bfloat16)
    bias = torch.randn(heads, 49, 49, device=dev)
    mask = torch.where(torch.rand(nW, 49, 49, device=dev) < 0.3, -100.0, 0.0)
    for _ in range(2):
        wo, wlse = N.window_attn_fwd(wqkv, bias, mask, heads, 32 ** -0.5)
        N.window_attn_bwd(wqkv, bias, mask, wo, torch.ones_like(wo), wlse, heads, 32 ** -0.5)
torch.cuda.synchronize()
print("done", what)
