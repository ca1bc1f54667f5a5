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
    for mb in (1, 25, 224, 497):
        n = mb * (1 << 20) // 4
        buf = F.alloc(n, torch.float32)
        t = torch.randn(n, device=dev)
        ms = timeit(lambda: F.all_reduce_(buf, 1.0 / world))
        ms_n = timeit(lambda: (t.div_(world), dist.all_reduce(t)))
        busbw = 2 * (world - 1) / world * n * 4 / ms / 1e6
        p0(f"time allreduce+scale {mb:4d} MiB fp32: fused {ms:.3f} ms ({busbw:.0f} GB/s busbw) | nccl div+allreduce {ms_n:.3f} ms")

p0("FAILED:" if FAILS else "ALL OK", FAILS)
dist.barrier()
dist.destroy_process_group()
sys.exit(1 if FAILS else 0)
