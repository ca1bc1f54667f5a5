"""DDP tier: bucket planner vs torch's, reducer ready-tracking, 2-process gloo
equivalence with single-process training, no_sync accumulation, MNIST parameter count."""
import copy

import pytest
import torch
import torch.nn as nn

from tests._dist_util import run_distributed


def test_mnist_param_count():
    from b200dist.models.mnist import Net, count_parameters
    assert count_parameters(Net()) == 1_199_882


def test_bucket_assignment_matches_torch():
    from torch._C._distributed_c10d import _compute_bucket_assignment_by_size
    from b200dist.parallel.ddp import assign_buckets
    torch.manual_seed(0)
    sizes = [int(x) for x in torch.randint(1, 400_000, (57,))]
    tensors = [torch.empty(s) for s in sizes]
    for first, cap in ((1 << 20, 25 << 20), (4096, 1 << 20), (1, 1)):
        ref, _ = _compute_bucket_assignment_by_size(tensors, [first, cap])
        got = assign_buckets([s * 4 for s in sizes], [0] * len(sizes), first, cap)
        mine = [[i for i, b in enumerate(got) if b == k] for k in range(max(got) + 1)]
        assert mine == [list(b) for b in ref]


def test_bucket_assignment_separates_dtypes():
    from b200dist.parallel.ddp import assign_buckets
    got = assign_buckets([100, 100, 100, 100], [0, 1, 0, 1], 1 << 20, 1 << 20)
    assert got[0] == got[2] and got[1] == got[3] and got[0] != got[1]


def test_reducer_core_in_order_launch():
    import ctypes
    from b200dist.utils import hostlib
    L = hostlib.lib()
    bucket_of = (ctypes.c_int32 * 5)(0, 0, 1, 1, 2)
    h = L.b2h_reducer_create(bucket_of, 5, 3)
    launch = (ctypes.c_int32 * 3)()
    assert L.b2h_reducer_mark_ready(ctypes.c_void_p(h), 4, launch) == 0      # bucket 2 complete but 0,1 pending
    assert L.b2h_reducer_mark_ready(ctypes.c_void_p(h), 2, launch) == 0
    assert L.b2h_reducer_mark_ready(ctypes.c_void_p(h), 3, launch) == 0      # bucket 1 complete, 0 pending
    assert L.b2h_reducer_mark_ready(ctypes.c_void_p(h), 0, launch) == 0
    n = L.b2h_reducer_mark_ready(ctypes.c_void_p(h), 1, launch)
    assert n == 3 and list(launch[:3]) == [0, 1, 2]
    assert L.b2h_reducer_mark_ready(ctypes.c_void_p(h), 1, launch) == -1     # double ready
    L.b2h_reducer_destroy(ctypes.c_void_p(h))


class _Tiny(nn.Module):
    def __init__(self):
        super().__init__()
        self.a, self.b, self.c = nn.Linear(16, 64), nn.Linear(64, 64), nn.Linear(64, 4)

    def forward(self, x):
        return self.c(torch.relu(self.b(torch.relu(self.a(x)))))


def _ddp_worker(rank, world, steps):
    import torch.distributed as dist
    from b200dist.optim.fused import Adadelta
    from b200dist.parallel.ddp import DistributedDataParallel as DDP
    torch.manual_seed(1234 + rank)          # different init per rank: ctor must broadcast rank 0's
    model = _Tiny()
    ddp = DDP(model, bucket_cap_mb=0.004, first_bucket_cap_mb=0.001)
    assert len(ddp.buckets) > 1
    opt = Adadelta(ddp.parameters(), lr=1.0)
    torch.manual_seed(7)
    X, Y = torch.randn(steps, 8, 16), torch.randint(0, 4, (steps, 8))
    for s in range(steps):
        xs, ys = X[s].chunk(world)[rank], Y[s].chunk(world)[rank]
        opt.zero_grad()
        nn.functional.cross_entropy(ddp(xs), ys).backward()
        opt.step()
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    return flat, all(torch.equal(gathered[0], g) for g in gathered)


def test_ddp_two_process_equals_single_process():
    steps = 4
    outs = run_distributed(_ddp_worker, 2, steps)
    assert all(o[1] for o in outs), "parameters diverged across ranks"
    # single-process oracle on the full batch, started from rank 0's init
    from b200dist.optim.fused import Adadelta
    torch.manual_seed(1234)
    model = _Tiny()
    opt = Adadelta(model.parameters(), lr=1.0)
    torch.manual_seed(7)
    X, Y = torch.randn(steps, 8, 16), torch.randint(0, 4, (steps, 8))
    for s in range(steps):
        opt.zero_grad()
        nn.functional.cross_entropy(model(X[s]), Y[s]).backward()
        opt.step()
    ref = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    assert torch.allclose(torch.from_numpy(outs[0][0]), ref, atol=2e-5, rtol=1e-4)


def _nosync_worker(rank, world):
    from b200dist.parallel.ddp import DistributedDataParallel as DDP
    torch.manual_seed(0)
    model = _Tiny()
    ddp = DDP(model)
    torch.manual_seed(100 + rank)
    x1, x2 = torch.randn(4, 16), torch.randn(4, 16)
    with ddp.no_sync():
        ddp(x1).sum().backward()
    ddp(x2).sum().backward()
    g = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    # oracle: mean over ranks of (grad(x1) + grad(x2))
    torch.manual_seed(0)
    ref = _Tiny()
    tot = None
    for r in range(world):
        torch.manual_seed(100 + r)
        a, b = torch.randn(4, 16), torch.randn(4, 16)
        ref.zero_grad()
        (ref(a).sum() + ref(b).sum()).backward()
        gr = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
        tot = gr if tot is None else tot + gr
    return torch.allclose(g, tot / world, atol=1e-5)


def test_ddp_no_sync_accumulates_then_reduces():
    assert all(run_distributed(_nosync_worker, 2))


def test_adadelta_matches_torch():
    from b200dist.optim.fused import Adadelta
    torch.manual_seed(0)
    m1 = _Tiny(); m2 = copy.deepcopy(m1)
    o1, o2 = Adadelta(m1.parameters(), lr=1.0), torch.optim.Adadelta(m2.parameters(), lr=1.0)
    for _ in range(3):
        x = torch.randn(8, 16)
        for m, o in ((m1, o1), (m2, o2)):
            o.zero_grad(); m(x).pow(2).sum().backward(); o.step()
    for p, q in zip(m1.parameters(), m2.parameters()):
        assert torch.allclose(p, q, atol=1e-6)


def test_distributed_sampler_matches_torch():
    from torch.utils.data import DistributedSampler as TS
    from b200dist.data.mnist import DistributedSampler
    ds = list(range(103))
    for rank in range(4):
        a = DistributedSampler(103, 4, rank, shuffle=True, seed=0); a.set_epoch(3)
        b = TS(ds, 4, rank, shuffle=True, seed=0); b.set_epoch(3)
        assert list(a) == list(b)


def _hook_worker(rank, world, which):
    """communication hooks (torch DDP register_comm_hook contract) and find_unused_parameters"""
    import torch.distributed as dist
    from b200dist.parallel import ddp as D
    torch.manual_seed(0)
    model = _Tiny()
    extra = nn.Linear(16, 4)                      # never used in forward
    model.add_module("unused_branch", extra)
    ddp = D.DistributedDataParallel(model, bucket_cap_mb=0.004, first_bucket_cap_mb=0.001, find_unused_parameters=True)
    seen = []
    if which == "allreduce":
        ddp.register_comm_hook(None, lambda st, b: (seen.append((b.index(), b.is_last(), len(b.parameters()))), D.allreduce_hook(st, b))[1])
    elif which == "bf16":
        ddp.register_comm_hook(None, D.bf16_compress_hook)
    elif which == "custom":
        def hook(state, bucket):               # synchronous hook returning a tensor
            t = bucket.buffer().clone()
            dist.all_reduce(t)
            return t / world
        ddp.register_comm_hook(None, hook)
    torch.manual_seed(100 + rank)
    x = torch.randn(4, 16)
    ddp(x).sum().backward()
    g = torch.cat([p.grad.reshape(-1) for n, p in model.named_parameters() if "unused_branch" not in n])
    unused_zero = all(float(p.grad.abs().max()) == 0 for p in extra.parameters())
    torch.manual_seed(0)
    ref = _Tiny()
    tot = None
    for r in range(world):
        torch.manual_seed(100 + r)
        a = torch.randn(4, 16)
        ref.zero_grad()
        ref(a).sum().backward()
        gr = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
        tot = gr if tot is None else tot + gr
    tol = 2e-2 if which == "bf16" else 1e-5
    ok = torch.allclose(g, tot / world, atol=tol, rtol=tol)
    try:
        ddp.register_comm_hook(None, D.allreduce_hook)
        twice = which == "none"
    except RuntimeError:
        twice = True
    return ok, unused_zero, twice, len(seen)


def test_ddp_comm_hooks_and_unused_parameters():
    for which in ("none", "allreduce", "bf16", "custom"):
        outs = run_distributed(_hook_worker, 2, which)
        for ok, unused_zero, twice, nseen in outs:
            assert ok and unused_zero and twice, (which, ok, unused_zero, twice)
            if which == "allreduce":
                assert nseen >= 2


def _vs_torch_ddp_worker(rank, world, steps):
    """our wrapper and torch.nn.parallel.DistributedDataParallel side by side: same init, same per-rank data, same optimizer"""
    import copy
    from torch.nn.parallel import DistributedDataParallel as TorchDDP
    from b200dist.parallel.ddp import DistributedDataParallel as DDP
    torch.manual_seed(1234 + rank)
    base = _Tiny()
    ours = DDP(copy.deepcopy(base), bucket_cap_mb=0.004, first_bucket_cap_mb=0.001)
    theirs = TorchDDP(copy.deepcopy(base), bucket_cap_mb=0.004)
    o1 = torch.optim.SGD(ours.parameters(), lr=0.1, momentum=0.9)
    o2 = torch.optim.SGD(theirs.parameters(), lr=0.1, momentum=0.9)
    torch.manual_seed(7)
    X, Y = torch.randn(steps, 8, 16), torch.randint(0, 4, (steps, 8))
    for s in range(steps):
        xs, ys = X[s].chunk(world)[rank], Y[s].chunk(world)[rank]
        for m, o in ((ours, o1), (theirs, o2)):
            o.zero_grad()
            nn.functional.cross_entropy(m(xs), ys).backward()
            o.step()
    a = torch.cat([p.detach().reshape(-1) for p in ours.module.parameters()])
    b = torch.cat([p.detach().reshape(-1) for p in theirs.module.parameters()])
    keys_equal = list(ours.state_dict()) == list(theirs.state_dict())
    return float((a - b).abs().max()), keys_equal


def test_ddp_wrapper_tracks_torch_ddp():
    for err, keys_equal in run_distributed(_vs_torch_ddp_worker, 2, 5):
        assert err < 1e-6 and keys_equal


class _TwoBranch(nn.Module):
    def __init__(self):
        super().__init__()
        self.a, self.b, self.head = nn.Linear(16, 32), nn.Linear(16, 32), nn.Linear(32, 4)

    def forward(self, x, swap):
        # the two branches reach autograd in a rank-dependent order, so gradient hooks fire in different orders per rank
        u, v = (self.b(x), self.a(x)) if swap else (self.a(x), self.b(x))
        return self.head(torch.relu(u) + torch.relu(v) * 1.0)


def _rebuild_order_worker(rank, world):
    import torch.distributed as dist
    from b200dist.parallel.ddp import DistributedDataParallel as DDP
    torch.manual_seed(0)
    model = _TwoBranch()
    ddp = DDP(model, bucket_cap_mb=0.002, first_bucket_cap_mb=0.001)
    torch.manual_seed(50 + rank)
    xs = [torch.randn(4, 16) for _ in range(3)]
    for x in xs:
        ddp.zero_grad()
        ddp(x, swap=(rank == 1)).sum().backward()
    layout = [list(b.param_ids) for b in ddp.buckets]
    objs = [None] * world
    dist.all_gather_object(objs, layout)
    g = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    # oracle for the last step: mean over ranks of the local gradients
    torch.manual_seed(0)
    ref = _TwoBranch()
    tot = None
    for r in range(world):
        torch.manual_seed(50 + r)
        x = [torch.randn(4, 16) for _ in range(3)][-1]
        ref.zero_grad()
        ref(x, swap=(r == 1)).sum().backward()
        gr = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
        tot = gr if tot is None else tot + gr
    return all(o == objs[0] for o in objs), bool(torch.allclose(g, tot / world, atol=1e-5)), sorted(i for b in layout for i in b) == list(range(6))


def test_ddp_bucket_rebuild_uses_rank0_order_on_every_rank():
    """ranks whose gradients arrive in different orders still rebuild identical buckets (rank 0's order is broadcast), every
    parameter keeps a bucket, and the reduced gradients are the cross-rank mean"""
    for same_layout, grads_ok, complete in run_distributed(_rebuild_order_worker, 2):
        assert same_layout and grads_ok and complete
