"""Headline benchmark: GPT tok/s on the BASELINE.json config "Megatron GPT-2 small (CodeParrot-style, seq 1024,
vocab 32768) pretrain_gpt.py, TP=2 with sequence parallel, DP=4" - the full training step (forward, backward,
gradient reduction, clipping, AdamW) in bf16 on synthetic tokens and random-init weights.

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 10 --warmup 3

Weak scaling: 24 sequences x 1024 tokens per GPU per step, 12 x 1024 tokens per GPU per micro-batch (the reference
run's micro-batch 12 at TP=1; a TP=2 group shares micro-batches of 24 sequences); at 8 GPUs this is the reference's
global batch 192 with TP=2 (+SP) x DP=4.  N=1 runs TP=1.
Prints ONE JSON line on rank 0 (see README / DESIGN.md for the field contract).

--impl reference : the reference's GPT path needs Megatron-LM + apex, which are neither vendored in
                   /root/reference nor installable offline -> prints {"impl": "reference", "unavailable": ...}.
--impl eager     : (not the reference) the same model written with stock PyTorch ops (cuBLAS, SDPA flash,
                   torch LayerNorm, fused AdamW) + NCCL collectives - the on-box library baseline.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "entrypoints"))

PUBLISHED_TOK_S = None   # no number is published for this config (772k tok/s is 16xA100, TP=4, fp16, vocab 50688)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "eager"])
    ap.add_argument("--micro-batch-size", type=int, default=0, help="0 = 12 x tp: 12 sequences (12288 tokens) per GPU per micro-batch, as in the reference run")
    ap.add_argument("--seqs-per-gpu", type=int, default=24)
    ap.add_argument("--num-layers", type=int, default=12)
    ap.add_argument("--hidden-size", type=int, default=768)
    ap.add_argument("--heads", type=int, default=12)
    ap.add_argument("--seq-length", type=int, default=1024)
    ap.add_argument("--vocab", type=int, default=32768)
    ap.add_argument("--tp", type=int, default=0, help="0 = min(2, gpus) (the BASELINE config)")
    ap.add_argument("--comm-backend", default="native", choices=["native", "nccl"])
    ap.add_argument("--dropout", type=float, default=0.1, help="hidden AND attention dropout (reference run: 0.1 each, 3_training_megatron-lm.ipynb:985,1056)")
    ap.add_argument("--no-clocks", action="store_true")
    ap.add_argument("--kernel-profile", default="", help="after the timed runs, trace 2 more steps with torch.profiler (CUPTI) on rank 0 and write a per-kernel table here (diagnostic: never a bench value)")
    ap.add_argument("--no-cuda-graph", action="store_true", help="issue every kernel from Python instead of replaying the captured micro-batch graph")
    return ap.parse_args()


class ClockSampler:
    FIELDS = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.proc, self.path = None, f"/tmp/b200dist_clocks_{os.getpid()}.csv"
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx = float(parts[1])
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.remove(self.path)
        except OSError:
            pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def library_baseline(n_gpus: int, value: float):
    """archived same-box comparison arms (profiles/library_baselines.json: this stack over NCCL collectives, and stock PyTorch);
    never the reference - see BASELINE.md section 3"""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "library_baselines.json")))
        row = d["tokens_per_s"].get(str(n_gpus))
        if not row:
            return None
        out = {"source": "archived runs on this pool (profiles/library_baselines.json), same config and timing rules; not the reference"}
        for k in ("nccl_backend", "stock_pytorch"):
            if row.get(k):
                out[k] = {"tokens_per_s": row[k], "ours_over_it": value / row[k]}
        return out
    except Exception:
        return None


_REAL_STDOUT_FD = None


def emit(line: str):
    """write the result line to the process's original stdout"""
    sys.stdout.flush()
    if _REAL_STDOUT_FD is not None:
        os.write(_REAL_STDOUT_FD, (line + "\n").encode())
    else:
        print(line, flush=True)


def main():
    a = parse()
    if a.impl == "reference":
        print(json.dumps({"impl": "reference", "unavailable": "reference GPT path = Megatron-LM + apex (cloned by its notebook, not vendored in /root/reference; no wheel in /opt/wheelhouse, no network); `pip install /root/reference` has nothing to build (no setup.py/pyproject)"}))
        return 0
    # stdout carries exactly ONE JSON line: everything else that writes to fd 1 (NCCL's version banner is printed from
    # C, below Python's sys.stdout) is sent to stderr until the result is ready
    sys.stdout.flush()
    global _REAL_STDOUT_FD
    _REAL_STDOUT_FD = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    assert world == a.gpus, f"--gpus {a.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {a.gpus}"
    tp = a.tp or min(2, a.gpus)
    dp = a.gpus // tp
    gbs = a.seqs_per_gpu * a.gpus
    if a.micro_batch_size <= 0:
        a.micro_batch_size = 12 * tp        # a TP group shares its micro-batch: per-GPU tokens per micro-batch stay at 12 x 1024
    assert gbs % (a.micro_batch_size * dp) == 0
    if a.impl == "eager":
        from baseline.eager_gpt import run_eager
        return run_eager(a, tp, dp, gbs, ClockSampler, emit)

    import pretrain_gpt as entry
    from b200dist.engine import megatron as eng
    from b200dist.ops import native as N
    from b200dist.parallel import pipeline as pp

    argv = ["--num-layers", str(a.num_layers), "--hidden-size", str(a.hidden_size), "--num-attention-heads", str(a.heads),
            "--seq-length", str(a.seq_length), "--max-position-embeddings", str(a.seq_length), "--micro-batch-size", str(a.micro_batch_size),
            "--global-batch-size", str(gbs), "--tensor-model-parallel-size", str(tp), "--bf16", "--tokenizer-type", "NullTokenizer",
            "--vocab-size", str(a.vocab - 1), "--mock-data", "--train-iters", str(a.steps + a.warmup + 2), "--lr", "3e-4", "--min-lr", "3e-5",
            "--lr-decay-style", "cosine", "--lr-warmup-iters", "2", "--weight-decay", "0.1", "--adam-beta2", "0.95", "--clip-grad", "1.0",
            "--hidden-dropout", str(a.dropout), "--attention-dropout", str(a.dropout), "--log-interval", "1000000", "--eval-iters", "0",
            "--comm-backend", a.comm_backend, "--num-workers", "0"]
    if tp > 1:
        argv.append("--sequence-parallel")
    if not a.no_cuda_graph:
        argv.append("--cuda-graph")
    _stdout = sys.stdout
    sys.stdout = sys.stderr      # keep stdout for the single JSON line
    args = eng.initialize_megatron(argv=argv)
    model, optimizer, sched = eng.setup_model_and_optimizer(entry.model_provider)
    for m in model:
        m.train()
    rank = dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    nmb = gbs // (a.micro_batch_size * dp)

    # synthetic token batches in PINNED host memory; every step copies its micro-batches host->device
    g = torch.Generator().manual_seed(1234 + rank)
    host_batches = [torch.randint(0, a.vocab, (a.micro_batch_size, a.seq_length + 1), generator=g).pin_memory() for _ in range(nmb * 2)]

    class HostIter:
        """yields {'text': pinned int64 [mbs, seq+1]} - get_batch() performs the H2D copy (broadcast_data)"""
        def __init__(self): self.i = 0
        def __iter__(self): return self
        def __next__(self):
            b = host_batches[self.i % len(host_batches)]; self.i += 1
            return {"text": b}

    class DeviceIter(HostIter):
        def __init__(self):
            super().__init__()
            self.dev = [b.to(dev) for b in host_batches]
        def __next__(self):
            b = self.dev[self.i % len(self.dev)]; self.i += 1
            return {"text": b}

    from b200dist.parallel import state as mpu
    is_src = mpu.get_tensor_model_parallel_rank() == 0

    def run(n_steps, it, read_loss):
        last = None
        for _ in range(n_steps):
            loss_dict, skipped, gn, _ = eng.train_step(entry.forward_step, it if is_src else None, model, optimizer, sched)
            if read_loss and loss_dict:
                last = float(loss_dict["lm loss"])       # device->host read of the step's loss
            elif read_loss:
                torch.cuda.current_stream().synchronize()
        return last

    def timed(n_steps, it, read_loss):
        dist.barrier(); torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        s.record()
        last = run(n_steps, it, read_loss)
        e.record()
        timed.host_issue_ms = (time.perf_counter() - t0) * 1e3 / n_steps    # host time to ISSUE a step (no sync)
        torch.cuda.synchronize(); dist.barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ms = torch.tensor([s.elapsed_time(e), wall], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms[0].item(), ms[1].item(), last

    dit, hit = DeviceIter(), HostIter()
    run(a.warmup, dit, False)
    sampler = None if (a.no_clocks or rank != 0) else ClockSampler(torch.cuda.current_device())
    l0 = N.launches()
    ms_dev, _, _ = timed(a.steps, dit, False)
    host_issue = timed.host_issue_ms
    launches = N.launches() - l0
    run(1, hit, True)
    ms_e2e, wall_e2e, last_loss = timed(a.steps, hit, True)
    clocks = sampler.stop() if sampler is not None else {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    if a.kernel_profile:
        from torch.profiler import profile, ProfilerActivity
        dist.barrier(); torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            run(2, dit, False)
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
                    f.write(f"{us / 2e3:9.3f} ms/step {100 * us / tot:5.1f}%  x{cnt // 2:5d}  avg {us / cnt:8.1f} us  {name[:150]}\n")
        dist.barrier()
    sys.stdout = _stdout
    tokens = gbs * a.seq_length
    value = tokens * a.steps / (ms_dev / 1e3)
    e2e = tokens * a.steps / (max(ms_e2e, wall_e2e) / 1e3)
    if rank == 0:
        out = {"metric": "GPT tok/s (Megatron TP/PP)", "value": value, "unit": "tokens/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
               "ms_per_step": ms_dev / a.steps, "higher_is_better": True, "scaling": "weak",
               "vs_baseline": (value / PUBLISHED_TOK_S) if PUBLISHED_TOK_S else None, "dtype": "bf16", "data": "synthetic tokens, random-init weights",
               "impl": "ours",
               "config": {"model": f"GPT-2 small {a.num_layers}L h{a.hidden_size} {a.heads} heads (Megatron GPT, CodeParrot-style)", "global_batch": gbs,
                          "seq_len": a.seq_length, "vocab": a.vocab, "micro_batch": a.micro_batch_size, "num_microbatches": nmb,
                          "parallelism": f"tp{tp}{'+sp' if tp > 1 else ''} dp{dp}", "comm_backend": a.comm_backend, "dropout": a.dropout, "cuda_graph": bool(not a.no_cuda_graph),
                          "optimizer": "AdamW fp32 master + grad clip 1.0", "l2": "per-step working set (weights+activations, GBs) exceeds the 126 MB L2"},
               "clocks": clocks,
               "e2e": {"value": e2e, "unit": "tokens/s", "h2d_bytes_per_step": nmb * a.micro_batch_size * (a.seq_length + 1) * 8 if is_src else 0,
                       "d2h_bytes_per_step": 4, "ms_per_step": max(ms_e2e, wall_e2e) / a.steps, "last_loss": last_loss},
               "gpu_launches": launches, "host_issue_ms_per_step": host_issue, "library_baseline": library_baseline(a.gpus, value)}
        emit(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
