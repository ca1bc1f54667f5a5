"""Multi-GPU check of the fabric kernels against NCCL (run under torchrun, ALWAYS inside `timeout`):
    timeout 120 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 bench/comm_check.py
"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from b200dist.launch import env as launch_env

info = launch_env.init_distributed("nccl")
rank, world = info.rank, info.world_size
dev = torch.device("cuda", torch.cuda.current_device())
from b200dist.parallel import fabric as fb
from b200dist.ops import native as N
FAILS = []


def p0(*a):
    if rank == 0:
        print(*a, flush=True)


def report(name, got, ref, tol):
    err = (got.float() - ref.float()).abs().max().item()
    sc = ref.float().abs().max().item() + 1e-9
    bad = torch.tensor([int(not err <= tol * max(1.0, sc))], device=dev)
    dist.all_reduce(bad, op=dist.ReduceOp.MAX)
    p0(f"{'FAIL' if bad.item() else 'ok  '} {name:50s} max_err={err:.3e} ref_max={sc:.3e}")
    if bad.item():
        FAILS.append(name)


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record(); torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def graph_time(fn, reps=10, iters=5):
    """GPU time per call with launch overhead removed: `reps` calls captured in one CUDA graph, replayed `iters` times."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn(); fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize(); dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    torch.cuda.synchronize(); dist.barrier()
    g.replay(); torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        g.replay()
    b.record(); torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / (iters * reps)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


F = fb.get_fabric(dist.group.WORLD)
p0("fabric usable:", F.usable, "error:", F.error)
if not F.usable:
    sys.exit(3)
which = sys.argv[1] if len(sys.argv) > 1 else "all"

if which in ("all", "allreduce"):
    for dtype, tol in ((torch.float32, 1e-5), (torch.bfloat16, 2e-2)):
        for n in (1000, 1 << 20, 25 * (1 << 20) // 4 + 13):
            buf = F.alloc(n, dtype)
            p0("  multicast" if buf.mc_ptr else "  peer ld/st", dtype, n)
            torch.manual_seed(rank + 1)
            x = torch.randn(n, device=dev).to(dtype)
            buf.tensor.copy_(x)
            ref = x.clone().float()
            dist.all_reduce(ref)
            ref = ref / world
            F.all_reduce_(buf, scale=1.0 / world)
            torch.cuda.synchronize()
            report(f"allreduce {dtype} n={n}", buf.tensor[:n], ref, tol)
            buf.tensor.copy_(x)
            F.all_reduce_(buf, scale=1.0 / world)       # second call on the same buffer (epochs advance)
            torch.cuda.synchronize()
            report(f"allreduce {dtype} n={n} (2nd call)", buf.tensor[:n], ref, tol)
    # reduce-scatter with a scale factor, fp32 and bf16 (the bf16 multicast branch once dropped it)
    for dtype, tol in ((torch.float32, 1e-5), (torch.bfloat16, 2e-2)):
        torch.manual_seed(31 + rank)
        x = torch.randn(world * 512, 384, device=dev).to(dtype)
        ref = x.clone().float()
        dist.all_reduce(ref)
        ref = ref[rank * 512:(rank + 1) * 512] * 0.37
        got = F.reduce_scatter_rows(x, scale=0.37)
        torch.cuda.synchronize()
        report(f"reduce_scatter_rows scale=0.37 {dtype}", got, ref, tol)
    for mb in (1, 4, 25, 64, 224, 497):
        n = mb * (1 << 20) // 4
        buf = F.alloc(n, torch.float32)
        t = torch.randn(n, device=dev)
        ms = timeit(lambda: F.all_reduce_(buf, 1.0 / world))
        ms_n = timeit(lambda: (t.div_(world), dist.all_reduce(t)))
        busbw = 2 * (world - 1) / world * n * 4 / ms / 1e6
        p0(f"time allreduce+scale {mb:4d} MiB fp32: fused {ms:.3f} ms ({busbw:.0f} GB/s busbw) | nccl div+allreduce {ms_n:.3f} ms")

if which in ("all", "ag"):
    for (rows, K, Nn, bmn) in [(256, 128, 256, False), (1024, 768, 1152, False), (6144, 768, 1536, False), (1024, 768, 384, True)]:
        torch.manual_seed(100 + rank)
        x = torch.randn(rows, K, device=dev, dtype=torch.bfloat16)
        torch.manual_seed(7)
        W = (torch.randn((K, Nn) if bmn else (Nn, K), device=dev, dtype=torch.bfloat16) * 0.05)
        bias = torch.randn(Nn, device=dev, dtype=torch.bfloat16) if not bmn else None
        y, xg = F.ag_gemm(x, W, b_mn=bmn, bias=bias, epilogue=N.EPI_BIAS if bias is not None else N.EPI_NONE)
        torch.cuda.synchronize()
        xs = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(xs, x)
        xf = torch.cat(xs)
        ref = xf.float() @ (W.float() if bmn else W.float().t()) + (bias.float() if bias is not None else 0)
        report(f"ag_gemm gathered rows{rows} K{K}", xg, xf, 0.0)
        report(f"ag_gemm out rows{rows} K{K} N{Nn} b_mn={int(bmn)}", y, ref, 2e-2)
        y2, _ = F.ag_gemm(x, W, b_mn=bmn, bias=bias, epilogue=N.EPI_BIAS if bias is not None else N.EPI_NONE)
        torch.cuda.synchronize()
        report(f"ag_gemm (2nd call) rows{rows}", y2, ref, 2e-2)
    rows, K, Nn = 6144, 768, 1536
    x = torch.randn(rows, K, device=dev, dtype=torch.bfloat16); W = torch.randn(Nn, K, device=dev, dtype=torch.bfloat16)
    xf = torch.empty(rows * world, K, device=dev, dtype=torch.bfloat16)
    ms = timeit(lambda: F.ag_gemm(x, W))
    ms_n = timeit(lambda: (dist.all_gather_into_tensor(xf, x), N.gemm(xf, W)))
    ms_g = timeit(lambda: N.gemm(xf, W))
    p0(f"time ag_gemm rows/rank {rows} K{K} N{Nn}: fused {ms:.3f} ms | nccl all_gather + gemm {ms_n:.3f} ms | gemm alone {ms_g:.3f} ms (host-issue bound)")
    for (rows, K, Nn, tag) in [(6144, 768, 1152, "qkv"), (6144, 768, 1536, "fc1"), (6144, 768, 16384, "logits"), (12288, 768, 1152, "qkv"), (12288, 768, 1536, "fc1")]:
        x = F.shard_buffer(rows, K, torch.bfloat16); x.normal_()
        W = torch.randn(Nn, K, device=dev, dtype=torch.bfloat16)
        xf = torch.randn(rows * world, K, device=dev, dtype=torch.bfloat16)
        gf = graph_time(lambda: F.ag_gemm(x, W))
        gg = graph_time(lambda: N.gemm(xf, W))
        roof = max(2.0 * rows * world * Nn * K / 1.39e15, rows * K * 2 * (world - 1) / 770e9) * 1e3
        p0(f"graph-timed ag_gemm {tag:7s} rows/rank {rows} K{K} N{Nn}: fused {gf*1e3:7.1f} us | plain gemm on gathered input {gg*1e3:7.1f} us | roofline {roof*1e3:6.1f} us -> {roof/gf:.2f} of roofline")

if which in ("all", "rs"):
    for (M, K, Nn, bmn) in [(512, 128, 256, False), (2048, 384, 768, False), (12288, 1536, 768, False), (2048, 576, 768, True)]:
        torch.manual_seed(200 + rank)
        a = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        W = torch.randn((K, Nn) if bmn else (Nn, K), device=dev, dtype=torch.bfloat16) * 0.05
        torch.manual_seed(9)
        bias = torch.randn(Nn, device=dev, dtype=torch.bfloat16)
        rows = M // world
        res = torch.randn(rows, Nn, device=dev, dtype=torch.bfloat16)
        out = F.gemm_rs(a, W, b_mn=bmn, bias=bias, residual=res)
        torch.cuda.synchronize()
        part = a.float() @ (W.float() if bmn else W.float().t())
        dist.all_reduce(part)
        ref = part[rank * rows:(rank + 1) * rows] + bias.float() + res.float()
        report(f"gemm_rs M{M} K{K} N{Nn} b_mn={int(bmn)}", out, ref, 3e-2)
        out2 = F.gemm_rs(a, W, b_mn=bmn, bias=bias, residual=res)
        out3 = F.gemm_rs(a, W, b_mn=bmn, bias=bias, residual=res)
        torch.cuda.synchronize()
        report(f"gemm_rs (3rd call) M{M}", out3, ref, 3e-2)
    # back-to-back calls with fresh inputs every time (a signal that overtook its data would show up as a stale partial)
    M, K, Nn = 12288, 384, 768
    W = torch.randn(Nn, K, device=dev, dtype=torch.bfloat16) * 0.05
    outs, refs = [], []
    for it in range(24):
        torch.manual_seed(1000 + 17 * it + rank)
        a = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        outs.append(F.gemm_rs(a, W))
        part = a.float() @ W.float().t()
        dist.all_reduce(part)
        refs.append(part[rank * (M // world):(rank + 1) * (M // world)])
    torch.cuda.synchronize()
    worst = max((o.float() - r).abs().max().item() for o, r in zip(outs, refs))
    report("gemm_rs 24 back-to-back calls, fresh data (worst)", torch.full((1,), worst, device=dev), torch.zeros(1, device=dev), 8e-2)
    del outs, refs
    M, K, Nn = 12288, 1536, 768
    a = torch.randn(M, K, device=dev, dtype=torch.bfloat16); W = torch.randn(Nn, K, device=dev, dtype=torch.bfloat16)
    o = torch.empty(M // world, Nn, device=dev, dtype=torch.bfloat16)
    ms = timeit(lambda: F.gemm_rs(a, W))
    ms_n = timeit(lambda: dist.reduce_scatter_tensor(o, N.gemm(a, W)))
    ms_g = timeit(lambda: N.gemm(a, W))
    p0(f"time gemm_rs M{M} K{K} N{Nn}: fused {ms:.3f} ms | gemm + nccl reduce_scatter {ms_n:.3f} ms | gemm alone {ms_g:.3f} ms (host-issue bound)")
    for (M, K, Nn, tag) in [(12288, 384, 768, "proj"), (12288, 1536, 768, "fc2"), (24576, 384, 768, "proj"), (24576, 1536, 768, "fc2")]:
        a = torch.randn(M, K, device=dev, dtype=torch.bfloat16); W = torch.randn(Nn, K, device=dev, dtype=torch.bfloat16)
        gf = graph_time(lambda: F.gemm_rs(a, W))
        gg = graph_time(lambda: N.gemm(a, W))
        roof = max(2.0 * M * Nn * K / 1.39e15, M * Nn * 2 * (world - 1) / world / 770e9) * 1e3
        p0(f"graph-timed gemm_rs {tag:5s} M{M} K{K} N{Nn}: fused {gf*1e3:7.1f} us | plain gemm {gg*1e3:7.1f} us | roofline {roof*1e3:6.1f} us -> {roof/gf:.2f} of roofline")

if which in ("all", "ar_gemm"):
    for (M, K, Nn, bmn) in [(512, 128, 256, False), (2048, 384, 768, False), (12288, 1536, 768, False), (2048, 576, 768, True)]:
        if M % (128 * world):
            continue
        torch.manual_seed(300 + rank)
        a = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        W = torch.randn((K, Nn) if bmn else (Nn, K), device=dev, dtype=torch.bfloat16) * 0.05
        torch.manual_seed(9)
        bias = torch.randn(Nn, device=dev, dtype=torch.bfloat16)
        res = torch.randn(M, Nn, device=dev, dtype=torch.bfloat16)
        out = F.gemm_ar(a, W, b_mn=bmn, bias=bias, residual=res)
        torch.cuda.synchronize()
        part = a.float() @ (W.float() if bmn else W.float().t())
        dist.all_reduce(part)
        ref = part + bias.float() + res.float()
        report(f"gemm_ar M{M} K{K} N{Nn} b_mn={int(bmn)}", out, ref, 3e-2)
        out2 = F.gemm_ar(a, W, b_mn=bmn, bias=bias, residual=res)
        out3 = F.gemm_ar(a, W, b_mn=bmn)
        torch.cuda.synchronize()
        report(f"gemm_ar (2nd call) M{M}", out2, ref, 3e-2)
        report(f"gemm_ar (3rd call, no bias / residual) M{M}", out3, part, 3e-2)
    for (M, K, Nn, tag) in [(12288, 192, 768, "proj tp4"), (12288, 768, 768, "fc2 tp4"), (12288, 384, 768, "proj tp2")]:
        a = torch.randn(M, K, device=dev, dtype=torch.bfloat16); W = torch.randn(Nn, K, device=dev, dtype=torch.bfloat16)
        t = torch.empty(M, Nn, device=dev, dtype=torch.bfloat16)
        gf = graph_time(lambda: F.gemm_ar(a, W))
        gg = graph_time(lambda: N.gemm(a, W))
        ms_n = timeit(lambda: dist.all_reduce(N.gemm(a, W)))
        p0(f"graph-timed gemm_ar {tag:9s} M{M} K{K} N{Nn}: fused {gf*1e3:7.1f} us (incl. the private copy of the result) | plain gemm {gg*1e3:7.1f} us | gemm + nccl all_reduce {ms_n*1e3:7.1f} us (stream-timed)")

if which in ("all", "agw"):
    # all-gather of a row-sharded weight fused into its GEMM (ZeRO-3 head): shard rows need not be a multiple of 128
    for (M, K, srows) in [(126, 64, 256), (300, 256, 104), (2044, 1024, 4000), (2044, 4096, 4000)]:
        Nn = srows * world
        wbuf = F.alloc(Nn * K, torch.bfloat16)
        torch.manual_seed(500 + rank)
        mine = (torch.randn(srows, K, device=dev) * 0.05).to(torch.bfloat16)
        shards = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(shards, mine)
        wfull = torch.cat(shards)
        torch.manual_seed(77)
        x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        for call in range(2):
            wbuf.tensor.zero_()
            wbuf.tensor.view(Nn, K)[rank * srows:(rank + 1) * srows].copy_(mine)
            torch.cuda.synchronize(); dist.barrier()
            y = F.agw_gemm(x, wbuf, 0, srows, K)
            torch.cuda.synchronize()
            report(f"agw_gemm out M{M} K{K} shard_rows{srows} (call {call + 1})", y, x.float() @ wfull.float().t(), 2e-2)
            report(f"agw_gemm gathered weight shard_rows{srows} (call {call + 1})", wbuf.tensor.view(Nn, K), wfull, 0.0)
    gf = graph_time(lambda: F.agw_gemm(x, wbuf, 0, srows, K))
    gg = graph_time(lambda: N.gemm(x, wfull))
    p0(f"graph-timed agw_gemm M{M} K{K} N{Nn}: fused {gf*1e3:.1f} us | plain gemm on the gathered weight {gg*1e3:.1f} us | remote shards {srows * (world - 1) * K * 2 / 1e6:.1f} MB")

if which in ("all", "xent"):
    for (rows, V) in [(512, 1024), (3000, 4096), (24576, 16384)]:
        torch.manual_seed(400 + rank)
        logits = (torch.randn(rows, V, device=dev) * 3).to(torch.bfloat16)
        torch.manual_seed(5)
        labels = torch.randint(0, V * world, (rows,), device=dev)
        labels[::7] = -100
        gm, gs, loss = F.xent_vocab_parallel(logits, labels, rank * V, -100)
        torch.cuda.synchronize()
        full = [torch.empty_like(logits) for _ in range(world)]
        dist.all_gather(full, logits)
        lf = torch.cat(full, dim=1).float()
        ref = torch.nn.functional.cross_entropy(lf, labels, ignore_index=-100, reduction="none")
        report(f"xent_vocab_parallel loss rows{rows} V{V}x{world}", loss, ref, 2e-3)
        report(f"xent_vocab_parallel max rows{rows}", gm, lf.max(-1).values, 1e-6)
        gm2, gs2, loss2 = F.xent_vocab_parallel(logits, labels, rank * V, -100)      # second call: other parity of the stats buffer
        torch.cuda.synchronize()
        report(f"xent_vocab_parallel (2nd call) rows{rows}", loss2, ref, 2e-3)
    gt = graph_time(lambda: F.xent_vocab_parallel(logits, labels, rank * V, -100))
    p0(f"graph-timed xent_vocab_parallel rows 24576 V {V} per rank: {gt*1e3:.1f} us ({rows * V * 2 / (gt * 1e-3) / 1e9:.0f} GB/s of logits)")

p0("FAILED:" if FAILS else "ALL OK", FAILS)
dist.barrier()
dist.destroy_process_group()
sys.exit(1 if FAILS else 0)
