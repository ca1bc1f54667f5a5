"""Property tests (hypothesis) of the native host components against their oracles: random shapes instead of hand-picked ones
(SURVEY 4.3 items 1 and 3)."""
import numpy as np
import pytest
import torch

hyp = pytest.importorskip("hypothesis")
from hypothesis import given, settings, strategies as st  # noqa: E402

FAST = settings(max_examples=40, deadline=None)


@FAST
@given(st.lists(st.integers(1, 60), min_size=1, max_size=40), st.integers(1, 33), st.integers(1, 3), st.integers(0, 2 ** 31 - 1))
def test_sample_idx_cpp_equals_python_twin(doc_sizes, seq_length, num_epochs, seed):
    from b200dist.data import gpt_dataset as gd
    sizes = np.asarray(doc_sizes, dtype=np.int32)
    tokens_per_epoch = int(sizes.sum())
    if (num_epochs * tokens_per_epoch - 1) // seq_length < 1:
        return
    rs = np.random.RandomState(seed)
    doc_idx = np.concatenate([rs.permutation(len(sizes)) for _ in range(num_epochs)]).astype(np.int32)
    a = gd.build_sample_idx(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch)
    b = gd.build_sample_idx_py(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch)
    assert a.shape == b.shape and (a == b).all()
    # every sample spans exactly seq_length + 1 tokens (the last one is shared with the next sample)
    flat = np.concatenate([[0], np.cumsum(sizes[doc_idx])])
    pos = flat[a[:, 0]] + a[:, 1]
    assert (np.diff(pos) == seq_length).all()


@FAST
@given(st.lists(st.floats(0.05, 1.0), min_size=1, max_size=6), st.integers(1, 400))
def test_blending_indices_track_the_weights(weights, size):
    from b200dist.data import gpt_dataset as gd
    w = np.asarray(weights, dtype=np.float64)
    w = w / w.sum()
    di, dsi = gd.build_blending_indices(w, size)
    counts = np.bincount(di, minlength=len(w))
    assert counts.sum() == size and (np.abs(counts - w * size) <= 1.0 + 1e-9).all()        # greedy error bound: within one sample
    for d in range(len(w)):
        assert (dsi[di == d] == np.arange(counts[d])).all()                                  # per-dataset running index


@FAST
@given(st.lists(st.integers(1, 300_000), min_size=1, max_size=60), st.integers(1, 1 << 21), st.integers(1, 1 << 23))
def test_bucket_assignment_equals_torch_reducer(sizes, first, cap):
    from torch._C._distributed_c10d import _compute_bucket_assignment_by_size
    from b200dist.parallel.ddp import assign_buckets
    tensors = [torch.empty(s) for s in sizes]
    ref, _ = _compute_bucket_assignment_by_size(tensors, [first, cap])
    got = assign_buckets([s * 4 for s in sizes], [0] * len(sizes), first, cap)
    mine = [[i for i, b in enumerate(got) if b == k] for k in range(max(got) + 1)]
    assert mine == [list(b) for b in ref]


@FAST
@given(st.integers(2, 8), st.integers(1, 4), st.integers(1, 6), st.booleans())
def test_pipeline_schedules_and_batched_plan(P, V, k, forward_only):
    from b200dist.parallel import pipeline as pp
    M = k * P if V > 1 else k
    for r in range(P):
        assert pp.build_schedule(P, r, V, M, forward_only) == pp.build_schedule_py(P, r, V, M, forward_only)
    assert pp.simulate_batched(P, V, M, forward_only)
    if not forward_only:
        sim = pp.simulate(P, V, M)
        assert sim["ok"]


@FAST
@given(st.integers(1, 70000), st.sampled_from([True, False]), st.floats(0.0, 0.2), st.integers(1, 50))
def test_cpu_adam_matches_reference_formula(n, adamw, wd, step):
    from b200dist.optim.cpu_adam import cpu_adam_
    g = torch.Generator().manual_seed(n)
    p, m, v, gr = torch.randn(n, generator=g), torch.rand(n, generator=g) * 0.1, torch.rand(n, generator=g) * 0.01, torch.randn(n, generator=g)
    lr, b1, b2, eps = 1e-2, 0.9, 0.98, 1e-8
    gg = gr.double() if adamw else gr.double() + wd * p.double()
    m2 = b1 * m.double() + (1 - b1) * gg
    v2 = b2 * v.double() + (1 - b2) * gg * gg
    pp_ = p.double() * (1 - lr * wd) if adamw else p.double()
    want = pp_ - lr / (1 - b1 ** step) * m2 / (v2.sqrt() / (1 - b2 ** step) ** 0.5 + eps)
    lp = torch.empty(n, dtype=torch.bfloat16)
    cpu_adam_(p, m, v, gr, lp, lr=lr, beta1=b1, beta2=b2, eps=eps, wd=wd, step=step, adamw=adamw, threads=2)
    assert torch.allclose(p.double(), want, rtol=2e-5, atol=2e-6)
    assert torch.equal(lp, p.to(torch.bfloat16))


@settings(max_examples=12, deadline=None)
@given(st.sampled_from(["adam", "adamw", "sgd", "sgd_momentum", "adadelta"]), st.floats(0.0, 0.1), st.integers(0, 10 ** 6))
def test_fused_optimizers_track_torch_optim(kind, wd, seed):
    """the flat-arena optimizers of optim/fused.py (CPU: the fp32 restatement of the multi-tensor kernels) against torch.optim over
    several parameter tensors of odd sizes and 4 steps, plus StepLR"""
    from b200dist.optim import fused
    g = torch.Generator().manual_seed(seed)
    shapes = [(7, 5), (13,), (3, 4, 2), (1,)]
    ps = [torch.randn(*s, generator=g) for s in shapes]
    a = [torch.nn.Parameter(p.clone()) for p in ps]
    b = [torch.nn.Parameter(p.clone()) for p in ps]
    mk = {"adam": (lambda q: fused.Adam(q, lr=1e-2, weight_decay=wd), lambda q: torch.optim.Adam(q, lr=1e-2, weight_decay=wd)),
          "adamw": (lambda q: fused.AdamW(q, lr=1e-2, weight_decay=wd), lambda q: torch.optim.AdamW(q, lr=1e-2, weight_decay=wd)),
          "sgd": (lambda q: fused.SGD(q, lr=0.1, weight_decay=wd), lambda q: torch.optim.SGD(q, lr=0.1, weight_decay=wd)),
          "sgd_momentum": (lambda q: fused.SGD(q, lr=0.1, momentum=0.9, weight_decay=wd), lambda q: torch.optim.SGD(q, lr=0.1, momentum=0.9, weight_decay=wd)),
          "adadelta": (lambda q: fused.Adadelta(q, lr=1.0, weight_decay=wd), lambda q: torch.optim.Adadelta(q, lr=1.0, weight_decay=wd))}[kind]
    oa, ob = mk[0](a), mk[1](b)
    sa, sb = fused.StepLR(oa, 1, 0.7), torch.optim.lr_scheduler.StepLR(ob, 1, 0.7)
    for _ in range(4):
        for x, y in zip(a, b):
            gr = torch.randn(x.shape, generator=g)
            x.grad, y.grad = gr.clone(), gr.clone()
        oa.step(); ob.step()
        sa.step(); sb.step()
        assert abs(sa.get_last_lr()[0] - sb.get_last_lr()[0]) < 1e-12
    for x, y in zip(a, b):
        assert torch.allclose(x, y, rtol=2e-5, atol=2e-6), (kind, (x - y).abs().max())
