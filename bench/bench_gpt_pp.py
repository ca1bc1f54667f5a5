"""BASELINE config 4: Megatron GPT 1.5B, TP=4 x PP=2, interleaved 1F1B, on 8 GPUs - tokens/s with the JSON contract of bench.py.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        bench/bench_gpt_pp.py --gpus 8 --steps 6 --warmup 3
    ... --gpus 4 --tp 2 --pp 2      (smaller boxes)        ... --cpu --tiny --gpus 2 --tp 1 --pp 2   (logic check on gloo)

Model shape (SURVEY App. C.4 option b; GPT-2-XL's 25 heads do not divide by TP=4): 48 layers, hidden 1536, 12 heads of
head_dim 128 (the native flash-attention kernels), ffn 6144, seq 1024, vocab 32768 -> 1.41 B parameters.  24 layers per
stage, 6 per virtual stage -> 4 model chunks per rank.  Data: synthetic tokens in pinned host memory, copied H2D every
micro-batch on the stages that need them (first: tokens, last: labels); bf16, AdamW + clipping inside the timed region.
Device-timed with CUDA events, max over ranks (wall clock in --cpu mode).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "entrypoints"))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=8)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--tp", type=int, default=4)
    ap.add_argument("--pp", type=int, default=2)
    ap.add_argument("--layers-per-virtual-stage", type=int, default=6, help="0 = plain 1F1B")
    ap.add_argument("--num-layers", type=int, default=48)
    ap.add_argument("--hidden-size", type=int, default=1536)
    ap.add_argument("--heads", type=int, default=12)
    ap.add_argument("--seq-length", type=int, default=1024)
    ap.add_argument("--vocab", type=int, default=32768)
    ap.add_argument("--micro-batch-size", type=int, default=4)
    ap.add_argument("--num-microbatches", type=int, default=8, help="per step and data-parallel replica (multiple of --pp for interleaving)")
    ap.add_argument("--sequence-parallel", action="store_true", default=True)
    ap.add_argument("--no-sequence-parallel", dest="sequence_parallel", action="store_false")
    ap.add_argument("--comm-backend", default="native", choices=["native", "nccl"])
    ap.add_argument("--cpu", action="store_true", help="gloo + wall clock (logic check of this script without a GPU)")
    ap.add_argument("--tiny", action="store_true", help="4 layers, hidden 64, 4 heads, seq 32, vocab 256 (with --cpu)")
    return ap.parse_args()


def main():
    a = parse()
    if a.tiny:
        a.num_layers, a.hidden_size, a.heads, a.seq_length, a.vocab, a.micro_batch_size = 4, 64, 4, 32, 256, 2
        a.layers_per_virtual_stage = min(a.layers_per_virtual_stage, 1) if a.layers_per_virtual_stage else 0
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    assert world == a.gpus, f"--gpus {a.gpus} but WORLD_SIZE={world}"
    assert a.gpus % (a.tp * a.pp) == 0
    dp = a.gpus // (a.tp * a.pp)
    gbs = a.micro_batch_size * a.num_microbatches * dp
    import pretrain_gpt as entry
    from b200dist.engine import megatron as eng
    from b200dist.parallel import state as mpu
    argv = ["--num-layers", str(a.num_layers), "--hidden-size", str(a.hidden_size), "--num-attention-heads", str(a.heads),
            "--seq-length", str(a.seq_length), "--max-position-embeddings", str(a.seq_length), "--micro-batch-size", str(a.micro_batch_size),
            "--global-batch-size", str(gbs), "--tensor-model-parallel-size", str(a.tp), "--pipeline-model-parallel-size", str(a.pp),
            "--tokenizer-type", "NullTokenizer", "--vocab-size", str(a.vocab - 1), "--mock-data", "--train-iters", str(a.steps + a.warmup + 2),
            "--lr", "1.5e-4", "--min-lr", "1.5e-5", "--lr-decay-style", "cosine", "--lr-warmup-iters", "2", "--weight-decay", "0.1", "--adam-beta2", "0.95",
            "--clip-grad", "1.0", "--hidden-dropout", "0.0", "--attention-dropout", "0.0", "--log-interval", "1000000", "--eval-iters", "0",
            "--comm-backend", a.comm_backend, "--num-workers", "0"]
    if not a.cpu:
        argv.append("--bf16")
    else:
        argv += ["--distributed-backend", "gloo"]
    if a.layers_per_virtual_stage:
        argv += ["--num-layers-per-virtual-pipeline-stage", str(a.layers_per_virtual_stage)]
    if a.tp > 1 and a.sequence_parallel:
        argv.append("--sequence-parallel")
    real_stdout = sys.stdout
    sys.stdout = sys.stderr
    args = eng.initialize_megatron(argv=argv)
    model, optimizer, sched = eng.setup_model_and_optimizer(entry.model_provider)
    for m in model:
        m.train()
    rank = dist.get_rank()
    cuda = torch.cuda.is_available() and not a.cpu
    dev = torch.device("cuda", torch.cuda.current_device()) if cuda else torch.device("cpu")
    nmb = a.num_microbatches
    g = torch.Generator().manual_seed(1234 + mpu.get_data_parallel_rank())
    host = [torch.randint(0, a.vocab, (a.micro_batch_size, a.seq_length + 1), generator=g) for _ in range(nmb * 2)]
    if cuda:
        host = [t.pin_memory() for t in host]

    class HostIter:
        def __init__(self): self.i = 0
        def __iter__(self): return self
        def __next__(self):
            b = host[self.i % len(host)]; self.i += 1
            return {"text": b}

    needs_data = mpu.get_tensor_model_parallel_rank() == 0 and (mpu.is_pipeline_first_stage(ignore_virtual=True) or mpu.is_pipeline_last_stage(ignore_virtual=True))

    def iters():
        if not needs_data:
            return [None] * len(model) if len(model) > 1 else None
        return [HostIter() for _ in model] if len(model) > 1 else HostIter()

    def run(n, read_loss):
        last = None
        for _ in range(n):
            loss_dict, skipped, gn, _ = eng.train_step(entry.forward_step, iters(), model, optimizer, sched)
            if read_loss and loss_dict:
                last = float(loss_dict["lm loss"])
        return last

    def timed(n):
        dist.barrier()
        if cuda:
            torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
        t0 = time.perf_counter()
        last = run(n, True)
        if cuda:
            e.record()
            torch.cuda.synchronize()
            ms = s.elapsed_time(e)
        else:
            ms = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item(), last

    run(a.warmup, False)
    sampler = None
    if cuda and rank == 0:
        from bench import ClockSampler
        sampler = ClockSampler(torch.cuda.current_device())
    ms, last = timed(a.steps)
    clocks = sampler.stop() if sampler is not None else {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    # the loss lives on the last pipeline stage; share it for the report
    lt = torch.tensor([last if last is not None else -1.0], device=dev)
    dist.all_reduce(lt, op=dist.ReduceOp.MAX)
    sys.stdout = real_stdout
    tokens = gbs * a.seq_length
    if rank == 0:
        n_params = a.num_layers * (12 * a.hidden_size ** 2 + 13 * a.hidden_size) + a.vocab * a.hidden_size + a.seq_length * a.hidden_size
        print(json.dumps({
            "metric": "GPT tok/s (Megatron TP/PP)", "value": tokens * a.steps / (ms / 1e3), "unit": "tokens/s", "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if cuda else "fp32 (cpu logic check)", "data": "synthetic tokens (pinned host -> device every micro-batch), random-init weights",
            "impl": "ours", "e2e_included": True, "last_loss": lt.item(), "clocks": clocks,
            "tp_kernels": "fused NVLink kernels" if (args.comm_backend == "native" and a.tp > 1) else "NCCL collectives + tcgen05 GEMMs",
            "config": {"model": f"GPT {a.num_layers}L h{a.hidden_size} {a.heads} heads ({n_params / 1e9:.2f} B parameters)", "global_batch": gbs,
                       "seq_len": a.seq_length, "vocab": a.vocab, "micro_batch": a.micro_batch_size, "num_microbatches": nmb,
                       "parallelism": f"tp{a.tp}{'+sp' if (a.tp > 1 and a.sequence_parallel) else ''} pp{a.pp}"
                                      f"{' interleaved x%d' % (a.num_layers // a.pp // a.layers_per_virtual_stage) if a.layers_per_virtual_stage else ' 1f1b'} dp{dp}",
                       "p2p": os.environ.get("B2_PP_P2P", "batched"), "comm_backend": args.comm_backend,
                       "l2": "per-step working set exceeds the 126 MB L2"}}), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
