"""Spawn helper for gloo multi-process tests (world_size > 1 on CPU)."""
import os
import socket
import sys
import traceback

import torch.multiprocessing as mp


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _plain(x):
    """tensors -> numpy so results cross the process boundary by value (no shared-memory handles)"""
    import torch
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy()
    if isinstance(x, (list, tuple)):
        return type(x)(_plain(v) for v in x)
    if isinstance(x, dict):
        return {k: _plain(v) for k, v in x.items()}
    return x


def _worker(rank, world, port, fn, args, q):
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank), "WORLD_SIZE": str(world),
                       "LOCAL_RANK": str(rank), "OMP_NUM_THREADS": "1"})
    try:
        import torch
        torch.set_num_threads(1)
        import torch.distributed as dist
        dist.init_process_group("gloo", rank=rank, world_size=world)
        out = _plain(fn(rank, world, *args))
        q.put((rank, "ok", out))
        dist.barrier()
        dist.destroy_process_group()
    except Exception:
        q.put((rank, "error", traceback.format_exc()))


def run_distributed(fn, world, *args, timeout=300):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, fn, args, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = {}
    try:
        for _ in range(world):
            r, status, out = q.get(timeout=timeout)
            if status != "ok":
                raise AssertionError(f"rank {r} failed:\n{out}")
            results[r] = out
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()
    return [results[r] for r in range(world)]
