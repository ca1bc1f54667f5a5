"""Oxford-Pet style DDP throughput (BASELINE config 2: Swin-B, bf16, 224 px, synthetic images) - img/s.

    timeout 600 python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/bench_ddp.py \
        --model swin_b --batch-size 64 --steps 10 --warmup 3 [--fp32]

The training step is the one of engine/oxford.py (forward, cross-entropy, backward, gradient all-reduce, Adam).
bf16: gradients accumulate in the optimizer's fp32 arena and are reduced by one fused all-reduce+scale launch per
step; fp32: bucketed DDP (C++ reducer core, one fused kernel per bucket, overlapped with backward).
Every step copies its image batch from pinned host memory.  One JSON line on rank 0, CUDA-event timed, max over ranks.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from b200dist.launch import env as launch_env


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="swin_b")
    ap.add_argument("--num-classes", type=int, default=37)
    ap.add_argument("--batch-size", type=int, default=64, help="per GPU")
    ap.add_argument("--size", type=int, default=224)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--fp32", action="store_true")
    ap.add_argument("--comm-backend", default="native")
    ap.add_argument("--no-cuda-graph", action="store_true")
    ap.add_argument("--kernel-profile", default="", help="after the timed steps, trace 2 more with torch.profiler (CUPTI) on rank 0 and write a "
                                                          "per-kernel table here (diagnostic: never a bench value)")
    a = ap.parse_args()
    real_fd = os.dup(1)
    os.dup2(2, 1)
    info = launch_env.init_distributed("nccl")
    rank, world = info.rank, info.world_size
    dev = torch.device("cuda", torch.cuda.current_device())
    from b200dist.ops import functional as OF
    from b200dist.ops import native as N
    from b200dist.parallel.ddp import DistributedDataParallel as DDP
    from b200dist.utils import train_util as util
    torch.manual_seed(0)
    dtype = torch.float32 if a.fp32 else torch.bfloat16
    model = util.torch_model(a.model, num_classes=a.num_classes, pretrained=False, local_rank=info.local_rank).to(dev).to(dtype)
    nparams = sum(p.numel() for p in model.parameters())
    if a.fp32:
        from b200dist.optim.fused import Adam
        if world > 1:
            model = DDP(model, bucket_cap_mb=25.0, comm_backend=a.comm_backend)
        opt, mixed = Adam(model.parameters(), lr=1e-4), False
    else:
        from b200dist.optim.mixed import MixedPrecisionOptimizer
        opt = MixedPrecisionOptimizer([model], lr=1e-4, weight_decay=0.0, betas=(0.9, 0.999), eps=1e-8, clip_grad=0.0,
                                      dp_group=dist.group.WORLD if world > 1 else None, comm_backend=a.comm_backend, no_wd_fn=lambda n, p: True)
        mixed = True
    model.train()
    g = torch.Generator().manual_seed(rank)
    host_x = [torch.randn(a.batch_size, 3, a.size, a.size, generator=g).to(dtype).pin_memory() for _ in range(3)]
    host_y = [torch.randint(0, a.num_classes, (a.batch_size,), generator=g).pin_memory() for _ in range(3)]

    def step(i):
        x = host_x[i % 3].to(dev, non_blocking=True)
        y = host_y[i % 3].to(dev, non_blocking=True)
        opt.zero_grad()
        loss = OF.cross_entropy(model(x), y)
        loss.backward()
        if mixed:
            opt.reduce_model_grads()
        opt.step()
        return loss

    graphed = None
    if not a.no_cuda_graph and mixed and (world == 1 or getattr(opt, "fabric", None) is not None):
        # ~2000 small launches per step are host-bound from Python: capture zero-grad + forward + backward + the fused
        # gradient all-reduce once and replay it; the optimizer step stays outside (its step count is a kernel argument)
        from b200dist.engine.graphed import GraphedStep

        def fwd_bwd(x, y):
            opt.zero_grad()
            loss = OF.cross_entropy(model(x), y)
            loss.backward()
            opt.reduce_model_grads()
            return loss.detach()
        graphed = GraphedStep(fwd_bwd, (host_x[0].to(dev), host_y[0].to(dev)))

        def step(i):                                   # noqa: F811
            loss = graphed(host_x[i % 3], host_y[i % 3])      # pinned host -> static device buffers, then replay
            opt.step()
            return loss

    for i in range(a.warmup):
        step(i)
    torch.cuda.synchronize(); dist.barrier()
    sampler = None
    if rank == 0:
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        from bench import ClockSampler
        sampler = ClockSampler(torch.cuda.current_device())
    l0 = N.launches()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.time()
    s.record()
    for i in range(a.steps):
        loss = step(i)
    e.record()
    last = float(loss.detach())
    torch.cuda.synchronize(); dist.barrier()
    wall = (time.time() - w0) * 1e3
    t = torch.tensor([max(s.elapsed_time(e), wall)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item() / a.steps
    clocks = sampler.stop() if sampler is not None else None
    if a.kernel_profile:
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for i in range(2):
                step(i)
            torch.cuda.synchronize()
        if rank == 0:
            rows = {}
            for ev in prof.events():
                if ev.device_type is not None and "cuda" in str(ev.device_type).lower():
                    r = rows.setdefault(ev.name, [0, 0.0])
                    r[0] += 1; r[1] += ev.device_time if hasattr(ev, "device_time") else ev.cuda_time
            tot = sum(v[1] for v in rows.values())
            with open(a.kernel_profile, "w") as f:
                f.write(f"# per-kernel device time over 2 steps (torch.profiler / CUPTI, rank 0): total {tot / 2e3:.3f} ms per step of kernel time\n")
                for name, (cnt, us) in sorted(rows.items(), key=lambda kv: -kv[1][1]):
                    f.write(f"{us / 2e3:9.3f} ms/step {100 * us / tot:5.1f}%  x{cnt / 2:7.1f}  avg {us / cnt:8.1f} us  {name[:150]}\n")
        dist.barrier()
    if rank == 0:
        out = {"metric": "DDP img/s", "value": world * a.batch_size / (ms / 1e3), "unit": "images/s", "n_gpus": world, "steps": a.steps,
               "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "dtype": "fp32" if a.fp32 else "bf16",
               "data": "synthetic images (pinned host -> device every step), random-init weights",
               "config": {"model": a.model, "params": nparams, "image": a.size, "batch_per_gpu": a.batch_size, "global_batch": world * a.batch_size,
                          "parallelism": f"dp{world}", "comm_backend": a.comm_backend, "optimizer": "Adam", "cuda_graph": graphed is not None},
               "last_loss": last, "clocks": clocks, "gpu_launches": N.launches() - l0,
               "e2e": {"included": True, "h2d_bytes_per_step": a.batch_size * 3 * a.size * a.size * (4 if a.fp32 else 2) + a.batch_size * 8, "d2h_bytes_per_step": 4}, "peak_mem_gib": torch.cuda.max_memory_allocated() / 2**30}
        os.write(real_fd, (json.dumps(out) + "\n").encode())
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
