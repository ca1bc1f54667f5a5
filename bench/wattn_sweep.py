"""Timing sweep of the tensor-core Swin window attention (csrc/window_attn_tc.cu) over the four Swin-B stages at batch 64 per GPU and
over the number of CTAs per SM the grids are sized for (0 = the occupancy query's answer, the default):

    python bench/wattn_sweep.py [--batch 64]

CUDA events on the launching stream, median of 20 after 5 warm-up calls, a 256 MiB memset between calls (L2 flush).  GB/s counts the
compulsory traffic (forward: qkv in, out out; backward: qkv, dO, O in, dqkv out)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from b200dist.ops import native as N


def timeit(fn, flush, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--per-sm", default="0,2,3,4,6,8")
    a = ap.parse_args()
    dev = "cuda"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    scale = 32 ** -0.5
    # Swin-B at 224 px: (windows per image, heads) per stage
    for stage, (nW, heads) in enumerate([(64, 4), (16, 8), (4, 16), (1, 32)], 1):
        Bw = a.batch * nW
        qkv = torch.randn(Bw, 49, 3 * heads * 32, device=dev, dtype=torch.bfloat16)
        bias = torch.randn(heads, 49, 49, device=dev)
        byt = Bw * 49 * heads * 32 * 2
        for masked in (False, True):
            mask = torch.where(torch.rand(nW, 49, 49, device=dev) < 0.3, -100.0, 0.0) if masked else None
            N.set_window_attn_grid(0, 0)
            out, lse = N.window_attn_fwd(qkv, bias, mask, heads, scale)
            do = torch.randn_like(out)
            cells = []
            for per_sm in [int(x) for x in a.per_sm.split(",")]:
                N.set_window_attn_grid(per_sm, per_sm)
                f = timeit(lambda: N.window_attn_fwd(qkv, bias, mask, heads, scale), flush)
                b = timeit(lambda: N.window_attn_bwd(qkv, bias, mask, out, do, lse, heads, scale), flush)
                cells.append(f"{per_sm}: {f * 1e3:6.1f}/{b * 1e3:6.1f}")
                if per_sm == 0:
                    head = f"fwd {f * 1e3:6.1f} us {4 * byt / f / 1e6:5.0f} GB/s  bwd {b * 1e3:6.1f} us {8 * byt / b / 1e6:5.0f} GB/s"
            N.set_window_attn_grid(0, 0)
            print(f"stage {stage} Bw{Bw:5d} heads{heads:3d} mask={int(masked)} | {head} | CTAs/SM fwd/bwd us: " + "  ".join(cells), flush=True)


if __name__ == "__main__":
    main()
