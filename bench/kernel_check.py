"""Numerics + timing sweep of every native kernel against the fp32 PyTorch oracle (runs on one B200).

    python bench/kernel_check.py [--quick] [--only gemm,attn,...]

Prints one line per case: max abs error, reference scale, verdict, and (for the hot kernels) CUDA-event
time with the achieved fraction of the measured peak from MEASURED_PEAKS.json.
"""
import argparse
import json
import os
import sys
import traceback

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from b200dist.ops import native as N

dev = torch.device("cuda")
torch.manual_seed(0)
PEAKS = {"bf16_tflops": 1662.1, "hbm_gbs": 6545.0}
try:
    PEAKS.update(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))))
except Exception:
    pass
FAILS = []


def cpu(t):
    return None if t is None else t.detach().cpu()


def report(name, got, ref, tol):
    got, ref = got.detach().float().cpu(), ref.detach().float().cpu()
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-12
    # two bounds: the worst element against the tensor's scale, and the relative RMS error (a wrong tile or a systematic bias
    # moves it far beyond rounding noise, which for 16-bit outputs sits near 2^-9 / sqrt(3) ~ 1e-3 of the RMS value)
    rms_ref = ref.pow(2).mean().sqrt().item()
    rel_rms = (got - ref).pow(2).mean().sqrt().item() / (rms_ref + 1e-12)
    rms_tol = max(0.35 * tol, 1e-6) if rms_ref > 1e-9 else float("inf")
    bad = not (err <= tol * max(1.0, scale)) or not torch.isfinite(got).all() or not (rel_rms <= rms_tol)
    print(f"{'FAIL' if bad else 'ok  '} {name:58s} max_err={err:.3e} ref_max={scale:.3e} tol={tol:g} rel_rms={rel_rms:.2e} (<= {rms_tol:.1e})", flush=True)
    if bad:
        FAILS.append(name)


def graph_timeit(fn, reps=20, iters=5):
    """GPU time per call without host launch cost: `reps` back-to-back calls captured in one CUDA graph."""
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        fn(); fn()
    torch.cuda.current_stream().wait_stream(st)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    torch.cuda.synchronize()
    g.replay(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); g.replay(); e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) / reps)
    ts.sort()
    return ts[len(ts) // 2]


def timeit(fn, iters=20, warm=5, flush=None):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def _gemm_numerics(tag):
    cases = [
        # M, N, K, a_mn, b_mn
        (256, 256, 128, False, False), (128, 128, 64, False, False), (384, 512, 320, False, False),
        (200, 264, 136, False, False), (256, 256, 128, False, True), (200, 264, 136, False, True),
        (256, 256, 128, True, True), (200, 264, 136, True, True), (256, 256, 128, True, False),
        (2048, 2304, 768, False, False), (2048, 768, 2304, False, True), (2304, 768, 2048, True, True),
    ]
    for (M, Nn, K, amn, bmn) in cases:
        a = torch.randn((K, M) if amn else (M, K), device=dev, dtype=torch.bfloat16)
        b = torch.randn((K, Nn) if bmn else (Nn, K), device=dev, dtype=torch.bfloat16)
        got = N.gemm(a, b, a_mn=amn, b_mn=bmn)
        ref = N.gemm(cpu(a), cpu(b), a_mn=amn, b_mn=bmn, out=torch.empty(M, Nn))
        report(f"gemm[{tag}] M{M} N{Nn} K{K} a_mn={int(amn)} b_mn={int(bmn)}", got, ref, 2e-2)
    # epilogues
    M, Nn, K = 384, 512, 256
    a = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
    b = torch.randn(Nn, K, device=dev, dtype=torch.bfloat16) * 0.05
    bias = torch.randn(Nn, device=dev, dtype=torch.bfloat16)
    aux = torch.randn(M, Nn, device=dev, dtype=torch.bfloat16)
    for epi, nm in [(N.EPI_BIAS, "bias"), (N.EPI_BIAS_GELU, "bias_gelu"), (N.EPI_BIAS_RESIDUAL, "bias_residual"), (N.EPI_DGELU, "dgelu")]:
        bb = None if epi == N.EPI_DGELU else bias
        ax = aux if epi in (N.EPI_BIAS_RESIDUAL, N.EPI_DGELU) else None
        got = N.gemm(a, b, epilogue=epi, bias=bb, aux=ax)
        ref = N.gemm(cpu(a), cpu(b), epilogue=epi, bias=cpu(bb), aux=cpu(ax), out=torch.empty(M, Nn))
        report(f"gemm[{tag}] epilogue {nm}", got, ref, 2e-2)
    for epi, nm in [(N.EPI_BIAS_GELU | N.EPI_ERF, "bias_gelu(erf)"), (N.EPI_DGELU | N.EPI_ERF, "dgelu(erf)")]:
        bb = None if (epi & 0xFF) == N.EPI_DGELU else bias
        ax = aux if (epi & 0xFF) == N.EPI_DGELU else None
        got = N.gemm(a, b, epilogue=epi, bias=bb, aux=ax)
        ref = N.gemm(cpu(a), cpu(b), epilogue=epi, bias=cpu(bb), aux=cpu(ax), out=torch.empty(M, Nn))
        report(f"gemm[{tag}] epilogue {nm}", got, ref, 2e-2)
    z = torch.empty(M, Nn, device=dev, dtype=torch.bfloat16)
    got = N.gemm(a, b, epilogue=N.EPI_BIAS_GELU_DUAL, bias=bias, out2=z)
    zr = torch.empty(M, Nn)
    ref = N.gemm(cpu(a), cpu(b), epilogue=N.EPI_BIAS_GELU_DUAL, bias=cpu(bias), out2=zr, out=torch.empty(M, Nn))
    report(f"gemm[{tag}] epilogue gelu_dual (act)", got, ref, 2e-2)
    report(f"gemm[{tag}] epilogue gelu_dual (pre)", z, zr, 2e-2)
    # fp32 accumulate + split-K (wgrad)
    Mo, No, Kr = 768, 768, 4096
    dy = torch.randn(Kr, Mo, device=dev, dtype=torch.bfloat16)
    x = torch.randn(Kr, No, device=dev, dtype=torch.bfloat16)
    acc = torch.ones(Mo, No, device=dev, dtype=torch.float32)
    N.gemm(dy, x, a_mn=True, b_mn=True, out=acc, accumulate=True)
    ref = 1.0 + dy.float().t() @ x.float()
    report(f"gemm[{tag}] wgrad fp32 accumulate auto split-K", acc, ref, 5e-3)
    acc2 = torch.zeros(Mo, No, device=dev, dtype=torch.float32)
    N.gemm(dy, x, a_mn=True, b_mn=True, out=acc2, accumulate=True, split_k=1)
    report(f"gemm[{tag}] wgrad fp32 accumulate split_k=1", acc2, ref - 1.0, 5e-3)
    out32 = N.gemm(a, b, out_fp32=True)
    report(f"gemm[{tag}] fp32 out (no accumulate)", out32, a.float() @ b.float().t(), 5e-3)


def check_gemm(quick):
    for mode, tag in ((0, "1cta"), (2, "pair")):       # single-CTA tiles, then cta_group::2 pairs wherever the shape allows
        N.set_gemm_pair_mode(mode)
        _gemm_numerics(tag)
    N.set_gemm_pair_mode(1)
    print("co-resident CTA pairs of the GEMM kernel:", N.lib().b2_gemm_resident_pairs(), "on", torch.cuda.get_device_properties(0).multi_processor_count, "SMs", flush=True)
    # timing on the GPT-2-small / 1.5B / LLaMA shapes
    if not quick:
        flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
        for (M, Nn, K, amn, bmn, tag) in [(12288, 2304, 768, False, False, "qkv fwd"), (12288, 3072, 768, False, False, "fc1 fwd"),
                                          (12288, 768, 3072, False, False, "fc2 fwd"), (12288, 768, 2304, False, True, "qkv dgrad"),
                                          (12288, 32768, 768, False, False, "logits"), (8192, 8192, 8192, False, False, "square 8k"),
                                          (4096, 11008, 4096, False, False, "llama gate")]:
            a = torch.randn((K, M) if amn else (M, K), device=dev, dtype=torch.bfloat16)
            b = torch.randn((K, Nn) if bmn else (Nn, K), device=dev, dtype=torch.bfloat16)
            o = torch.empty(M, Nn, device=dev, dtype=torch.bfloat16)
            N.set_gemm_pair_mode(0)
            ms1 = graph_timeit(lambda: N.gemm(a, b, a_mn=amn, b_mn=bmn, out=o))
            N.set_gemm_pair_mode(2)
            ms = graph_timeit(lambda: N.gemm(a, b, a_mn=amn, b_mn=bmn, out=o))
            N.set_gemm_pair_mode(1)
            tf = 2.0 * M * Nn * K / ms / 1e9
            bt = b if bmn else b.t()
            ms_ref = graph_timeit(lambda: torch.matmul(a.t() if amn else a, bt, out=o))
            print(f"time gemm {tag:12s} M{M} N{Nn} K{K}: pair {ms:.4f} ms {tf:7.1f} TFLOP/s = {tf / PEAKS['bf16_tflops']:.3f} of measured cuBLAS peak | 1-CTA {ms1:.4f} ms | torch.matmul {ms_ref:.4f} ms  (graph-timed, back to back)", flush=True)
        # fused epilogues on the GPT-2-small shapes (the model's actual fwd / dgrad calls)
        for (M, Nn, K, bmn, epi, tag) in [(12288, 2304, 768, False, N.EPI_BIAS, "qkv  bias"), (12288, 768, 768, False, N.EPI_BIAS_RESIDUAL, "proj bias+res"),
                                          (12288, 3072, 768, False, N.EPI_BIAS_GELU_DUAL, "fc1  gelu dual"), (12288, 768, 3072, False, N.EPI_BIAS_RESIDUAL, "fc2  bias+res"),
                                          (12288, 3072, 768, True, N.EPI_DGELU, "fc2 dgrad dgelu"), (12288, 768, 768, True, N.EPI_NONE, "proj dgrad")]:
            a = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
            b = torch.randn((K, Nn) if bmn else (Nn, K), device=dev, dtype=torch.bfloat16)
            o = torch.empty(M, Nn, device=dev, dtype=torch.bfloat16)
            o2 = torch.empty(M, Nn, device=dev, dtype=torch.bfloat16) if epi == N.EPI_BIAS_GELU_DUAL else None
            bias = torch.randn(Nn, device=dev, dtype=torch.bfloat16) if epi in (N.EPI_BIAS, N.EPI_BIAS_RESIDUAL, N.EPI_BIAS_GELU_DUAL) else None
            aux = torch.randn(M, Nn, device=dev, dtype=torch.bfloat16) if epi in (N.EPI_BIAS_RESIDUAL, N.EPI_DGELU) else None
            N.set_gemm_pair_mode(0)
            ms1 = graph_timeit(lambda: N.gemm(a, b, b_mn=bmn, out=o, epilogue=epi, bias=bias, aux=aux, out2=o2))
            N.set_gemm_pair_mode(2)
            ms = graph_timeit(lambda: N.gemm(a, b, b_mn=bmn, out=o, epilogue=epi, bias=bias, aux=aux, out2=o2))
            ms0 = graph_timeit(lambda: N.gemm(a, b, b_mn=bmn, out=o))
            N.set_gemm_pair_mode(1)
            tf = 2.0 * M * Nn * K / ms / 1e9
            print(f"time gemm+epi {tag:16s} M{M} N{Nn} K{K}: pair {ms:.4f} ms {tf:7.1f} TFLOP/s | 1-CTA {ms1:.4f} ms | pair, plain epilogue {ms0:.4f} ms  (graph-timed)", flush=True)
        for (Mo, No, Kr, tag) in [(2304, 768, 12288, "qkv wgrad"), (768, 768, 12288, "proj wgrad"), (3072, 768, 12288, "fc1 wgrad")]:
            dy = torch.randn(Kr, Mo, device=dev, dtype=torch.bfloat16)
            x = torch.randn(Kr, No, device=dev, dtype=torch.bfloat16)
            acc = torch.zeros(Mo, No, device=dev, dtype=torch.float32)
            N.set_gemm_pair_mode(0)
            ms1 = graph_timeit(lambda: N.gemm(dy, x, a_mn=True, b_mn=True, out=acc, accumulate=True))
            N.set_gemm_pair_mode(2)
            ms = graph_timeit(lambda: N.gemm(dy, x, a_mn=True, b_mn=True, out=acc, accumulate=True))
            N.set_gemm_pair_mode(1)
            tf = 2.0 * Mo * No * Kr / ms / 1e9
            print(f"time gemm {tag:12s} M{Mo} N{No} K{Kr}: pair {ms:.4f} ms {tf:7.1f} TFLOP/s = {tf / PEAKS['bf16_tflops']:.3f} of measured cuBLAS peak | 1-CTA {ms1:.4f} ms  (graph-timed)", flush=True)


def check_norm(quick):
    for (rows, H, rms) in [(1000, 768, False), (513, 1024, False), (300, 4096, True), (257, 1536, False), (64, 128, False),
                           (1001, 256, False), (777, 128, False), (4099, 128, True), (301, 512, False), (131, 64, False), (515, 200, False)]:
        x = torch.randn(rows, H, device=dev, dtype=torch.bfloat16)
        g = torch.randn(H, device=dev, dtype=torch.bfloat16)
        b = None if rms else torch.randn(H, device=dev, dtype=torch.bfloat16)
        dy = torch.randn(rows, H, device=dev, dtype=torch.bfloat16)
        y, mean, rstd = N.norm_fwd(x, g, b, 1e-5, rms)
        yr, meanr, rstdr = N.norm_fwd(cpu(x), cpu(g), cpu(b), 1e-5, rms)
        report(f"norm fwd rows{rows} H{H} rms={int(rms)}", y, yr, 2e-2)
        dx, dg, db = N.norm_bwd(dy, x, g, mean, rstd, rms)
        dxr, dgr, dbr = N.norm_bwd(cpu(dy), cpu(x), cpu(g), meanr, rstdr, rms)
        report(f"norm bwd dx rows{rows} H{H} rms={int(rms)}", dx, dxr, 3e-2)
        report(f"norm bwd dgamma rows{rows} H{H}", dg, dgr, 2e-3)
        if not rms:
            report(f"norm bwd dbeta rows{rows} H{H}", db, dbr, 2e-3)
        # residual-gradient fusion + accumulate into existing fp32 buffers (main_grad path)
        dres = torch.randn(rows, H, device=dev, dtype=torch.bfloat16)
        mg_w, mg_b = torch.ones(H, device=dev), (None if rms else torch.ones(H, device=dev))
        dx2, _, _ = N.norm_bwd(dy, x, g, mean, rstd, rms, dgamma=mg_w, dbeta=mg_b, accumulate=not rms, dres=dres)
        dx2r, dg2r, db2r = N.norm_bwd(cpu(dy), cpu(x), cpu(g), meanr, rstdr, rms, dres=cpu(dres))
        report(f"norm bwd +dres dx rows{rows} H{H} rms={int(rms)}", dx2, dx2r, 3e-2)
        report(f"norm bwd +dres dgamma(acc) rows{rows} H{H}", mg_w, dg2r + (0.0 if rms else 1.0), 2e-3)
    if not quick:
        # device time without host launch cost: 20 calls captured in a CUDA graph, cycling over 4 input sets so that consecutive calls
        # never find their rows in L2 (the sets of one shape add up to more than the 126 MB L2, except the 768-wide GPT shape: 75 MB)
        for (rows, H, tag) in [(12288, 768, "GPT-2 small micro-batch"), (200704, 128, "swin_b stage 1"), (50176, 256, "swin_b stage 2"),
                               (12544, 512, "swin_b stage 3")]:
            xs = [torch.randn(rows, H, device=dev, dtype=torch.bfloat16) for _ in range(4)]
            dys = [torch.randn(rows, H, device=dev, dtype=torch.bfloat16) for _ in range(4)]
            g = torch.ones(H, device=dev, dtype=torch.bfloat16)
            b = torch.zeros(H, device=dev, dtype=torch.bfloat16)
            _, mean, rstd = N.norm_fwd(xs[0], g, b, 1e-5)
            mg_w, mg_b = torch.zeros(H, device=dev), torch.zeros(H, device=dev)
            k = [0]

            def fwd():
                k[0] += 1
                return N.norm_fwd(xs[k[0] % 4], g, b, 1e-5)

            def bwd():
                k[0] += 1
                return N.norm_bwd(dys[k[0] % 4], xs[k[0] % 4], g, mean, rstd, dgamma=mg_w, dbeta=mg_b, accumulate=True, dres=dys[(k[0] + 1) % 4])
            ms_f, ms_b = graph_timeit(fwd), graph_timeit(bwd)
            gf, gbw = 2 * rows * H * 2 / ms_f / 1e6, 4 * rows * H * 2 / ms_b / 1e6
            print(f"time layernorm {rows}x{H} ({tag}): fwd {ms_f * 1e3:.1f} us {gf:.0f} GB/s = {gf / PEAKS['hbm_gbs']:.3f} of measured copy bw | "
                  f"bwd(+dres, +fold) {ms_b * 1e3:.1f} us {gbw:.0f} GB/s = {gbw / PEAKS['hbm_gbs']:.3f}  (graph-timed)", flush=True)


def check_elementwise(quick):
    x = torch.randn(512, 768, device=dev, dtype=torch.bfloat16)
    bias = torch.randn(768, device=dev, dtype=torch.bfloat16)
    res = torch.randn(512, 768, device=dev, dtype=torch.bfloat16)
    report("bias_dropout_add p=0", N.bias_dropout_add(x, bias, res), N.bias_dropout_add(cpu(x), cpu(bias), cpu(res)), 2e-2)
    y = N.bias_dropout_add(x, bias, None, p=0.25, seed=1234)
    kept = (y != 0).float().mean().item()
    print(f"{'ok  ' if abs(kept - 0.75) < 0.02 else 'FAIL'} dropout keep fraction {kept:.4f} (want 0.75)")
    dy = torch.ones_like(x)
    dxm = N.dropout_bwd(dy, 0.25, 1234)
    same = ((dxm != 0) == (y != 0)).float().mean().item()
    print(f"{'ok  ' if same > 0.999 else 'FAIL'} dropout fwd/bwd mask agreement {same:.5f}")
    if same <= 0.999 or abs(kept - 0.75) >= 0.02:
        FAILS.append("dropout")
    for mode in range(4):
        report(f"bias_act fwd mode{mode}", N.bias_act_fwd(x, bias, mode), N.bias_act_fwd(cpu(x), cpu(bias), mode), 2e-2)
        report(f"bias_act bwd mode{mode}", N.bias_act_bwd(res, x, bias, mode), N.bias_act_bwd(cpu(res), cpu(x), cpu(bias), mode), 3e-2)
    gu = torch.randn(300, 2 * 1408, device=dev, dtype=torch.bfloat16)
    dyy = torch.randn(300, 1408, device=dev, dtype=torch.bfloat16)
    report("swiglu fwd", N.swiglu_fwd(gu), N.swiglu_fwd(cpu(gu)), 2e-2)
    report("swiglu bwd", N.swiglu_bwd(dyy, gu), N.swiglu_bwd(cpu(dyy), cpu(gu)), 3e-2)
    acc = torch.zeros(768, device=dev)
    xx = torch.randn(3000, 768, device=dev, dtype=torch.bfloat16)
    N.colsum_accum(xx, acc)
    report("colsum", acc, xx.float().sum(0), 2e-3)
    report("axpby", N.axpby(x, res, 0.5, 2.0), N.axpby(cpu(x), cpu(res), 0.5, 2.0), 2e-2)
    # rope round trip + reference
    T, nh, hd = 64, 4, 64
    q = torch.randn(T, nh * hd, device=dev, dtype=torch.bfloat16)
    inv = 1.0 / (10000 ** (torch.arange(0, hd, 2, device=dev).float() / hd))
    ang = torch.arange(128, device=dev).float()[:, None] * inv[None]
    cos_t, sin_t = ang.cos().contiguous(), ang.sin().contiguous()
    pos = torch.arange(T, device=dev, dtype=torch.int32)
    q_ref = N.rope_(cpu(q).clone(), cpu(cos_t), cpu(sin_t), cpu(pos), nh, hd, nh * hd, hd)
    q_gpu = N.rope_(q.clone(), cos_t, sin_t, pos, nh, hd, nh * hd, hd)
    report("rope fwd", q_gpu, q_ref, 2e-2)
    q_back = N.rope_(q_gpu.clone(), cos_t, sin_t, pos, nh, hd, nh * hd, hd, backward=True)
    report("rope fwd->bwd round trip", q_back, q, 3e-2)


def check_xent(quick):
    for (rows, V, vs) in [(300, 32768, 0), (128, 16384, 16384), (77, 50688, 0)]:
        logits = torch.randn(rows, V, device=dev, dtype=torch.bfloat16) * 3
        labels = torch.randint(0, V + vs, (rows,), device=dev)
        labels[::7] = -100
        m, s, t = N.xent_stats(logits, labels, vs)
        mr, sr, tr = N.xent_stats(cpu(logits), cpu(labels), vs)
        report(f"xent stats max rows{rows} V{V}", m, mr, 1e-5)
        report(f"xent stats sumexp rows{rows} V{V}", s, sr, 1e-3)
        report(f"xent stats tgt rows{rows} V{V}", t, tr, 1e-5)
        loss = N.xent_finish(m, s, t, labels)
        report(f"xent loss rows{rows} V{V}", loss, N.xent_finish(mr, sr, tr, cpu(labels)), 1e-3)
        lr_ = cpu(logits).clone()
        N.xent_bwd_(lr_, cpu(labels), mr, sr, None, 0.5, vs)
        lg = logits.clone()
        N.xent_bwd_(lg, labels, m, s, None, 0.5, vs)
        report(f"xent bwd rows{rows} V{V}", lg, lr_, 2e-2)


def check_optim(quick):
    n = 1_000_003
    p = torch.randn(n, device=dev)
    m = torch.zeros(n, device=dev)
    v = torch.zeros(n, device=dev)
    g = torch.randn(n, device=dev)
    plp = torch.empty(n, device=dev, dtype=torch.bfloat16)
    pc, mc, vc, gc = cpu(p).clone(), cpu(m).clone(), cpu(v).clone(), cpu(g)
    ss = N.sumsq(g)
    report("sumsq", ss, gc.pow(2).sum()[None], 1e-4)
    ss2 = N.sumsq(g)
    report("sumsq (second call, counter reset)", ss2, gc.pow(2).sum()[None], 1e-4)
    for step in (1, 2):
        N.adam_(p, m, v, g, plp, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, wd=0.1, step=step, sumsq=ss, max_norm=1.0)
        N.adam_(pc, mc, vc, gc, None, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, wd=0.1, step=step, sumsq=cpu(ss), max_norm=1.0)
    report("adamw p (2 steps, clip)", p, pc, 1e-5)
    report("adamw m", m, mc, 1e-5)
    report("adamw low-precision copy", plp, pc, 1e-2)
    sq, acc = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    sqc, accc = torch.zeros(n), torch.zeros(n)
    N.adadelta_(p, sq, acc, g, lr=1.0, rho=0.9, eps=1e-6)
    N.adadelta_(pc, sqc, accc, gc, lr=1.0, rho=0.9, eps=1e-6)
    report("adadelta", p, pc, 1e-4)
    fi = torch.zeros(1, device=dev)
    gg = g.clone(); gg[12345] = float("inf")
    N.scale_check_(gg, 0.5, fi)
    print(f"{'ok  ' if fi.item() == 1.0 else 'FAIL'} scale_check found_inf={fi.item()}")
    if not quick:
        n = 124_000_000
        p = torch.randn(n, device=dev); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev); g = torch.randn(n, device=dev)
        plp = torch.empty(n, device=dev, dtype=torch.bfloat16)
        ms = timeit(lambda: N.adam_(p, m, v, g, plp, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, wd=0.1, step=3), iters=10)
        gb = n * (4 * 4 + 3 * 4 + 2) / ms / 1e6
        print(f"time adam 124M params: {ms:.3f} ms {gb:.0f} GB/s = {gb / PEAKS['hbm_gbs']:.3f} of measured copy bw")


def check_embedding(quick):
    V, H, T = 1000, 768, 512
    wte = torch.randn(V, H, device=dev, dtype=torch.bfloat16)
    wpe = torch.randn(1024, H, device=dev, dtype=torch.bfloat16)
    tok = torch.randint(0, 2 * V, (T,), device=dev)
    pos = torch.arange(T, device=dev) % 1024
    report("embedding fwd (vocab shard)", N.embedding_fwd(tok, pos, wte, wpe, 500), N.embedding_fwd(cpu(tok), cpu(pos), cpu(wte), cpu(wpe), 500), 2e-2)
    dy = torch.randn(T, H, device=dev, dtype=torch.bfloat16)
    dw, dp = torch.zeros(V, H, device=dev), torch.zeros(1024, H, device=dev)
    dwr, dpr = torch.zeros(V, H), torch.zeros(1024, H)
    N.embedding_bwd_(dy, tok, pos, dw, dp, 500)
    N.embedding_bwd_(cpu(dy), cpu(tok), cpu(pos), dwr, dpr, 500)
    report("embedding bwd dwte", dw, dwr, 1e-3)
    report("embedding bwd dwpe", dp, dpr, 1e-3)


def check_wattn(quick):
    """Swin window attention (7x7 windows, head_dim 32): fused kernels vs the fp32 PyTorch restatement (tensor-core kernels, then the
    CUDA-core ones they replaced)."""
    for impl, tag in ((1, "tc"), (0, "simt")):
        N.set_window_attn_impl(impl)
        _check_wattn_numerics(tag)
    N.set_window_attn_impl(1)
    _check_wattn_rest(quick)


def _check_wattn_numerics(tag):
    for (Bw, heads, nW, dt) in [(8, 4, 4, torch.bfloat16), (30, 3, 1, torch.bfloat16), (128, 8, 16, torch.bfloat16), (12, 2, 4, torch.float16)]:
        qkv = torch.randn(Bw, 49, 3 * heads * 32, device=dev, dtype=dt)
        bias = torch.randn(heads, 49, 49, device=dev)
        mask = None if nW == 1 else torch.where(torch.rand(nW, 49, 49, device=dev) < 0.3, -100.0, 0.0)
        scale = 32 ** -0.5
        out, lse = N.window_attn_fwd(qkv, bias, mask, heads, scale)
        outr, lser = N.window_attn_fwd(cpu(qkv), cpu(bias), cpu(mask), heads, scale)
        report(f"wattn[{tag}] fwd Bw{Bw} heads{heads} nW{nW}", out, outr, 2e-2)
        report(f"wattn[{tag}] lse Bw{Bw} heads{heads} nW{nW}", lse, lser, 2e-3)
        do = torch.randn_like(out)
        dqkv, dbias = N.window_attn_bwd(qkv, bias, mask, out, do, lse, heads, scale)
        dqkvr, dbiasr = N.window_attn_bwd(cpu(qkv), cpu(bias), cpu(mask), outr.to(dt), cpu(do), lser, heads, scale)
        for part, name in enumerate(("dq", "dk", "dv")):
            report(f"wattn[{tag}] bwd {name} Bw{Bw} heads{heads}", dqkv.view(Bw, 49, 3, -1)[:, :, part], dqkvr.view(Bw, 49, 3, -1)[:, :, part], 4e-2)
        report(f"wattn[{tag}] bwd dbias Bw{Bw} heads{heads}", dbias, dbiasr, 2e-2)


def _check_wattn_rest(quick):
    # cyclic shift + window partition and its inverse (one gather pass each) against roll + permute
    for (B, H, W, C, sh) in [(2, 14, 14, 64, (3, 3)), (3, 28, 56, 128, (3, 0)), (1, 7, 7, 32, (0, 0))]:
        x = torch.randn(B, H, W, C, device=dev, dtype=torch.bfloat16)
        xw = N.window_shuffle(x, B, H, W, 7, sh, False)
        report(f"window partition B{B} {H}x{W} C{C} shift{sh}", xw, N.window_shuffle(cpu(x), B, H, W, 7, sh, False), 0.0)
        back = N.window_shuffle(xw, B, H, W, 7, sh, True)
        report(f"window reverse  B{B} {H}x{W} C{C} shift{sh} (round trip)", back, x, 0.0)
    # 4 x 4 patches as GEMM rows (patch-embedding convolution as one GEMM) against view / permute / reshape
    for (B, C, H, W, dt) in [(2, 3, 32, 48, torch.bfloat16), (3, 5, 8, 4, torch.float16), (8, 3, 224, 224, torch.bfloat16)]:
        x = torch.randn(B, C, H, W, device=dev, dtype=dt)
        report(f"patchify4 B{B} C{C} {H}x{W}", N.patchify(x, 4), N.patchify(cpu(x), 4), 0.0)
    # LayerNorm writing its rows through a permutation (Swin: norm + cyclic shift + window partition in one pass) and its backward
    for (B, H, W, C, sh) in [(2, 14, 14, 128, (3, 3)), (3, 28, 56, 64, (3, 0)), (2, 7, 7, 1024, (0, 0))]:
        from b200dist.ops import functional as OF
        rm = OF.window_row_map(B, H, W, 7, sh, torch.device(dev))
        x = torch.randn(B, H, W, C, device=dev, dtype=torch.bfloat16)
        g, b = torch.randn(C, device=dev, dtype=torch.bfloat16), torch.randn(C, device=dev, dtype=torch.bfloat16)
        y, mean, rstd = N.norm_fwd(x, g, b, 1e-5, False, row_map=rm)
        yr, meanr, rstdr = N.norm_fwd(cpu(x), cpu(g), cpu(b), 1e-5, False, row_map=cpu(rm))
        report(f"norm fwd + window row map B{B} {H}x{W} C{C} shift{sh}", y, yr, 2e-2)
        dy, dres = torch.randn_like(y), torch.randn_like(x)
        dx, dg, db = N.norm_bwd(dy, x, g, mean, rstd, False, dres=dres, row_map=rm)
        dxr, dgr, dbr = N.norm_bwd(cpu(dy), cpu(x), cpu(g), meanr, rstdr, False, dres=cpu(dres), row_map=cpu(rm))
        report(f"norm bwd + window row map dx B{B} {H}x{W} C{C}", dx, dxr, 3e-2)
        report(f"norm bwd + window row map dgamma B{B} {H}x{W} C{C}", dg, dgr, 2e-3)
        report(f"norm bwd + window row map dbeta B{B} {H}x{W} C{C}", db, dbr, 2e-3)
    # residual add + per-sample scale (stochastic depth), identity map and with the window maps folded in
    for (B, H, W, C, sh, dt) in [(3, 14, 14, 64, (3, 3), torch.bfloat16), (2, 28, 56, 128, (3, 0), torch.float16), (5, 7, 7, 32, (0, 0), torch.bfloat16)]:
        res = torch.randn(B, H, W, C, device=dev, dtype=dt)
        xw = torch.randn(B * (H // 7) * (W // 7), 49, C, device=dev, dtype=dt)
        sc = torch.empty(B, device=dev).bernoulli_(0.6).div_(0.6)
        for mode, xin, r in ((2, xw, res), (1, res, None), (0, res, res.flip(0)), (0, xw, None)):
            for scale in (sc, None):
                got = N.rows_scale_add(xin, r, scale, B, H, W, 7, sh, mode)
                want = N.rows_scale_add(cpu(xin).float(), cpu(r).float() if r is not None else None, cpu(scale), B, H, W, 7, sh, mode)
                report(f"rows_scale_add mode{mode} B{B} {H}x{W} C{C} shift{sh} res={int(r is not None)} scale={int(scale is not None)}", got, want, 8e-3)
    if not quick:
        for (B, H, W, C) in [(64, 56, 56, 128), (64, 14, 14, 512)]:
            res = torch.randn(B, H, W, C, device=dev, dtype=torch.bfloat16)
            xw = torch.randn(B * (H // 7) * (W // 7), 49, C, device=dev, dtype=torch.bfloat16)
            sc = torch.empty(B, device=dev).bernoulli_(0.6).div_(0.6)
            ms2 = timeit(lambda: N.rows_scale_add(xw, res, sc, B, H, W, 7, (3, 3), 2))
            ms1 = timeit(lambda: N.rows_scale_add(res, None, sc, B, H, W, 7, (3, 3), 1))
            ms_e = timeit(lambda: res + N.window_shuffle(xw, B, H, W, 7, (3, 3), True) * sc.to(res.dtype).view(B, 1, 1, 1))
            byt = res.numel() * 2
            print(f"time rows_scale_add B{B} {H}x{W} C{C}: fwd (reverse + scale + add) {ms2:.4f} ms {3 * byt / ms2 / 1e6:.0f} GB/s | bwd (partition + scale) "
                  f"{ms1:.4f} ms {2 * byt / ms1 / 1e6:.0f} GB/s | window_shuffle kernel + ATen multiply + add {ms_e:.4f} ms", flush=True)
        for (Bw, heads, nW, tag) in [(4096, 4, 64, "swin_b stage 1, batch 64"), (256, 16, 4, "swin_b stage 3, batch 64")]:
            qkv = torch.randn(Bw, 49, 3 * heads * 32, device=dev, dtype=torch.bfloat16)
            bias = torch.randn(heads, 49, 49, device=dev)
            mask = torch.zeros(nW, 49, 49, device=dev)
            out, lse = N.window_attn_fwd(qkv, bias, mask, heads, 32 ** -0.5)
            do = torch.randn_like(out)
            ms_f = timeit(lambda: N.window_attn_fwd(qkv, bias, mask, heads, 32 ** -0.5))
            ms_b = timeit(lambda: N.window_attn_bwd(qkv, bias, mask, out, do, lse, heads, 32 ** -0.5))
            N.set_window_attn_impl(0)
            ms_f0 = timeit(lambda: N.window_attn_fwd(qkv, bias, mask, heads, 32 ** -0.5))
            ms_b0 = timeit(lambda: N.window_attn_bwd(qkv, bias, mask, out, do, lse, heads, 32 ** -0.5))
            N.set_window_attn_impl(1)

            def eager():
                q, k, v = qkv.view(Bw, 49, 3, heads, 32).permute(2, 0, 3, 1, 4)
                a = (q * 32 ** -0.5) @ k.transpose(-2, -1) + bias.to(q.dtype)[None]
                a = (a.view(Bw // nW, nW, heads, 49, 49) + mask.to(a.dtype)[None, :, None]).view(Bw, heads, 49, 49)
                return (torch.softmax(a.float(), -1).to(v.dtype) @ v).transpose(1, 2).reshape(Bw, 49, heads * 32)
            ms_e = timeit(eager)
            byt = Bw * 49 * heads * 32 * 2
            print(f"time wattn {tag}: fwd {ms_f:.4f} ms ({4 * byt / ms_f / 1e6:.0f} GB/s) bwd {ms_b:.4f} ms ({8 * byt / ms_b / 1e6:.0f} GB/s) | "
                  f"CUDA-core kernels fwd {ms_f0:.4f} ms bwd {ms_b0:.4f} ms | eager PyTorch chain fwd {ms_e:.4f} ms", flush=True)


def check_attn(quick):
    for (S, B, Hh, D, causal) in [(128, 1, 1, 64, True), (256, 2, 3, 64, True), (384, 1, 2, 64, False), (200, 1, 2, 64, True),
                                  (1024, 2, 4, 64, True), (256, 1, 2, 128, True), (512, 2, 3, 128, True), (384, 1, 2, 128, False),
                                  (200, 1, 2, 128, True)]:
        qkv = torch.randn(S, B, Hh, 3 * D, device=dev, dtype=torch.bfloat16)
        q, k, v = qkv[..., :D], qkv[..., D:2 * D], qkv[..., 2 * D:]
        scale = D ** -0.5
        o, lse = N.attn_fwd(q, k, v, scale, causal)
        qc, kc, vc = cpu(q), cpu(k), cpu(v)
        orf, lser = N.attn_fwd(qc, kc, vc, scale, causal)
        report(f"attn fwd S{S} B{B} H{Hh} D{D} causal={int(causal)}", o, orf, 2e-2)
        report(f"attn lse S{S} B{B} H{Hh} D{D}", lse, lser, 2e-3)
        if D in (64, 128):
            do = torch.randn(S, B, Hh, D, device=dev, dtype=torch.bfloat16)
            dqkv = torch.zeros_like(qkv)
            N.attn_bwd(q, k, v, o, do, lse, scale, causal, dqkv[..., :D], dqkv[..., D:2 * D], dqkv[..., 2 * D:])
            dref = torch.zeros(S, B, Hh, 3 * D)
            N.attn_bwd(qc, kc, vc, orf, cpu(do), lser, scale, causal, dref[..., :D], dref[..., D:2 * D], dref[..., 2 * D:])
            report(f"attn bwd dq S{S} B{B} H{Hh} D{D}", dqkv[..., :D], dref[..., :D], 4e-2)
            report(f"attn bwd dk S{S} B{B} H{Hh} D{D}", dqkv[..., D:2 * D], dref[..., D:2 * D], 4e-2)
            report(f"attn bwd dv S{S} B{B} H{Hh} D{D}", dqkv[..., 2 * D:], dref[..., 2 * D:], 4e-2)
    # attention dropout: the kernels' counter-based mask is restated in ops/native.py, so fwd and bwd are checked exactly
    for (S, B, Hh, D, pd) in [(256, 2, 3, 64, 0.1), (384, 1, 2, 128, 0.25), (200, 1, 2, 64, 0.5)]:
        qkv = torch.randn(S, B, Hh, 3 * D, device=dev, dtype=torch.bfloat16)
        q, k, v = qkv[..., :D], qkv[..., D:2 * D], qkv[..., 2 * D:]
        scale, seed = D ** -0.5, 0x1234567 + S
        o, lse = N.attn_fwd(q, k, v, scale, True, p_drop=pd, seed=seed)
        qc, kc, vc = cpu(q), cpu(k), cpu(v)
        orf, lser = N.attn_fwd(qc, kc, vc, scale, True, p_drop=pd, seed=seed)
        report(f"attn fwd dropout {pd} S{S} D{D}", o, orf, 2e-2)
        report(f"attn lse dropout {pd} S{S} D{D}", lse, lser, 2e-3)
        keep = N.attn_dropout_mask(B, Hh, S, pd, seed).ne(0).float().mean().item()
        print(f"{'ok  ' if abs(keep - (1 - pd)) < 0.01 else 'FAIL'} attn dropout keep fraction {keep:.4f} (want {1 - pd})")
        do = torch.randn(S, B, Hh, D, device=dev, dtype=torch.bfloat16)
        dqkv = torch.zeros_like(qkv)
        N.attn_bwd(q, k, v, o, do, lse, scale, True, dqkv[..., :D], dqkv[..., D:2 * D], dqkv[..., 2 * D:], p_drop=pd, seed=seed)
        dref = torch.zeros(S, B, Hh, 3 * D)
        N.attn_bwd(qc, kc, vc, orf, cpu(do), lser, scale, True, dref[..., :D], dref[..., D:2 * D], dref[..., 2 * D:], p_drop=pd, seed=seed)
        for nm, sl in (("dq", slice(0, D)), ("dk", slice(D, 2 * D)), ("dv", slice(2 * D, 3 * D))):
            report(f"attn bwd {nm} dropout {pd} S{S} D{D}", dqkv[..., sl], dref[..., sl], 4e-2)
    if not quick:
        S, B, Hh, D = 1024, 12, 12, 64
        qkv = torch.randn(S, B, Hh, 3 * D, device=dev, dtype=torch.bfloat16)
        q, k, v = qkv[..., :D], qkv[..., D:2 * D], qkv[..., 2 * D:]
        out = torch.empty(S, B, Hh, D, device=dev, dtype=torch.bfloat16)
        ms = timeit(lambda: N.attn_fwd(q, k, v, 0.125, True, out))
        fl = 4.0 * S * S * D * B * Hh / 2
        print(f"time attn fwd S1024 B12 H12 D64 causal: {ms:.4f} ms {fl / ms / 1e9:.1f} TFLOP/s = {fl / ms / 1e9 / PEAKS['bf16_tflops']:.3f} of measured cuBLAS peak")
        o, lse = N.attn_fwd(q, k, v, 0.125, True)
        do = torch.randn_like(o)
        dqkv = torch.empty_like(qkv)
        ms = timeit(lambda: N.attn_bwd(q, k, v, o, do, lse, 0.125, True, dqkv[..., :D], dqkv[..., D:2 * D], dqkv[..., 2 * D:]))
        print(f"time attn bwd S1024 B12 H12 D64 causal: {ms:.4f} ms {2.5 * fl / ms / 1e9:.1f} TFLOP/s = {2.5 * fl / ms / 1e9 / PEAKS['bf16_tflops']:.3f} of measured cuBLAS peak")
        ms = timeit(lambda: N.attn_fwd(q, k, v, 0.125, True, out, p_drop=0.1, seed=7))
        print(f"time attn fwd S1024 B12 H12 D64 causal dropout 0.1: {ms:.4f} ms")
        ms = timeit(lambda: N.attn_bwd(q, k, v, o, do, lse, 0.125, True, dqkv[..., :D], dqkv[..., D:2 * D], dqkv[..., 2 * D:], p_drop=0.1, seed=7))
        print(f"time attn bwd S1024 B12 H12 D64 causal dropout 0.1: {ms:.4f} ms")
        S, B, Hh, D = 512, 16, 32, 128                  # LLaMA-7B attention at the ZeRO-3 benchmark shape
        qkv = torch.randn(S, B, Hh, 3 * D, device=dev, dtype=torch.bfloat16)
        q, k, v = qkv[..., :D], qkv[..., D:2 * D], qkv[..., 2 * D:]
        fl = 4.0 * S * S * D * B * Hh / 2
        o, lse = N.attn_fwd(q, k, v, D ** -0.5, True)
        ms = timeit(lambda: N.attn_fwd(q, k, v, D ** -0.5, True, o))
        print(f"time attn fwd S512 B16 H32 D128 causal: {ms:.4f} ms {fl / ms / 1e9:.1f} TFLOP/s")
        do = torch.randn_like(o)
        dqkv = torch.empty_like(qkv)
        ms = timeit(lambda: N.attn_bwd(q, k, v, o, do, lse, D ** -0.5, True, dqkv[..., :D], dqkv[..., D:2 * D], dqkv[..., 2 * D:]))
        print(f"time attn bwd S512 B16 H32 D128 causal: {ms:.4f} ms {2.5 * fl / ms / 1e9:.1f} TFLOP/s")
        try:
            from flash_attn import flash_attn_func
            qb, kb, vb = (t.permute(1, 0, 2, 3).contiguous() for t in (q, k, v))
            ms = timeit(lambda: flash_attn_func(qb, kb, vb, causal=True))
            print(f"  (context) flash_attn 2.8 library fwd same shape: {ms:.4f} ms")
        except Exception as e:  # library only used as a yard-stick
            print("  flash_attn yard-stick unavailable:", type(e).__name__)


ALL = {"gemm": check_gemm, "norm": check_norm, "elementwise": check_elementwise, "xent": check_xent,
       "optim": check_optim, "embedding": check_embedding, "attn": check_attn, "wattn": check_wattn}

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    names = [n for n in args.only.split(",") if n] or list(ALL)
    print("device:", torch.cuda.get_device_name(0), "| SMs:", N.num_sms(), "| lib:", N.lib_path(), flush=True)
    for n in names:
        print(f"=== {n} ===", flush=True)
        try:
            ALL[n](args.quick)
            torch.cuda.synchronize()
        except Exception:
            traceback.print_exc()
            FAILS.append(n + " (exception)")
            try:
                torch.cuda.synchronize()
            except Exception:
                print("CUDA context is broken; stopping")
                break
    print("FAILED:" if FAILS else "ALL OK", FAILS)
    sys.exit(1 if FAILS else 0)
