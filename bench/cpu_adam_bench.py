"""Host-side micro-benchmark of csrc/host/cpu_adam.cpp (offload_optimizer.device = cpu) against torch.optim.AdamW on the same
CPU: time per step over an fp32 shard, effective memory bandwidth (7 fp32 streams + 1 bf16 stream per element), both code
paths (AVX-512 / generic), thread scaling.   python bench/cpu_adam_bench.py [--elems 33554432] [--steps 5]"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--elems", type=int, default=32 * 1024 * 1024)
    ap.add_argument("--steps", type=int, default=9)
    a = ap.parse_args()
    from b200dist.optim.cpu_adam import cpu_adam_, simd_width
    n = a.elems
    p, m, v, g = torch.randn(n), torch.zeros(n), torch.zeros(n), torch.randn(n)
    lp = torch.empty(n, dtype=torch.bfloat16)
    bytes_per_elem = 4 * 4 + 3 * 4 + 2                  # read p, m, v, g; write p, m, v; write bf16 copy
    rows = []

    def run(label, fn):
        fn(1)                                           # warm-up (page faults, thread start)
        times = []
        for s in range(a.steps):
            t0 = time.perf_counter()
            fn(s + 2)
            times.append(time.perf_counter() - t0)
        dt, med = min(times), sorted(times)[len(times) // 2]      # shared host: report the best and the median step
        rows.append({"impl": label, "best_ms": round(dt * 1e3, 2), "median_ms": round(med * 1e3, 2), "best_GB_per_s": round(n * bytes_per_elem / dt / 1e9, 1),
                     "best_Melem_per_s": round(n / dt / 1e6, 1)})
        print(rows[-1], flush=True)

    ncpu = os.cpu_count() or 1
    # long warm-up first: on a fresh allocation the first dozen passes run at a fraction of the steady-state rate on this
    # host (page placement / huge-page promotion), whichever implementation runs them
    for s in range(16):
        cpu_adam_(p, m, v, g, lp, lr=1e-3, wd=0.01, step=s + 1, threads=ncpu)
    for threads in sorted({1, max(1, ncpu // 2), ncpu}):
        run(f"b2h_cpu_adam simd{simd_width()} threads={threads}",
            lambda s, t=threads: cpu_adam_(p, m, v, g, lp, lr=1e-3, wd=0.01, step=s, threads=t))
    run(f"b2h_cpu_adam generic threads={ncpu}", lambda s: cpu_adam_(p, m, v, g, lp, lr=1e-3, wd=0.01, step=s, threads=ncpu, force_generic=True))
    torch.set_num_threads(ncpu)
    ref = torch.nn.Parameter(p.clone())
    ref.grad = g.clone()
    for foreach in (True, False):
        opt = torch.optim.AdamW([ref], lr=1e-3, weight_decay=0.01, foreach=foreach)

        def step(s, opt=opt):
            opt.step()
            lp.copy_(ref.data)                          # the bf16 copy the offload path needs as well
        for s in range(16):                             # same long warm-up for its (new) state tensors
            step(s)
        run(f"torch.optim.AdamW foreach={foreach} + bf16 copy, {ncpu} threads", step)
    print(json.dumps({"elems": n, "cpu_count": ncpu, "rows": rows}))


if __name__ == "__main__":
    main()
