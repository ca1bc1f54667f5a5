"""LLaMA / OPT ZeRO-3 fine-tuning throughput (BASELINE config 5: LLaMA-7B, bf16, seq 512, synthetic instruction batches).

    timeout 900 python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/bench_zero3.py \
        --model llama-7b --seq 512 --micro-batch 4 --steps 5 --warmup 2 [--comm-backend native|nccl]

The reference's schedule (4_training_alpaca_deepspeed.ipynb:470-472): 4 sequences per GPU per micro-batch x 8 gradient-
accumulation steps per optimizer step (`--micro-batch 4 --grad-accum 8`, the defaults), AdamW + clipping 1.0, fp32 master
weights / moments partitioned over the ranks, parameters gathered layer by layer; the DeepSpeed JSON's stage3 knobs
(max_live_parameters 1e9, prefetch bucket 0.9 hidden^2, max_reuse_distance 1e9) are passed to the engine.  A "step" is one
OPTIMIZER step (all its micro-batches).  Prints one JSON line on rank 0; CUDA-event timed, max over ranks; every micro-batch
copies its token batch from pinned host memory and the step's loss is read back, so the number is end to end.  L2: a 7B step touches > 100 GB, far beyond the 126 MB L2.
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
    ap.add_argument("--model", default="llama-7b")
    ap.add_argument("--seq", type=int, default=512)
    ap.add_argument("--micro-batch", type=int, default=4)
    ap.add_argument("--grad-accum", type=int, default=8)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3, help="warm-up OPTIMIZER steps")
    ap.add_argument("--max-live-parameters", type=float, default=1e9)
    ap.add_argument("--max-reuse-distance", type=float, default=1e9)
    ap.add_argument("--no-clocks", action="store_true")
    ap.add_argument("--full-init", action="store_true", help="materialise the whole model on every rank before partitioning (default: meta-device "
                    "construction, one unit at a time - the zero.Init way)")
    ap.add_argument("--no-fuse-gather-gemm", action="store_true", help="gather the LM head with the stand-alone kernel instead of inside its GEMM")
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (0 = the model's own)")
    ap.add_argument("--comm-backend", default="native")
    ap.add_argument("--clip", type=float, default=1.0)
    ap.add_argument("--offload-optimizer", default="none", choices=["none", "cpu"], help="cpu = the reference's -original config (C++ CPU Adam)")
    ap.add_argument("--offload-param", default="none", choices=["none", "cpu"], help="cpu = the reference's shipped config (pinned host shards)")
    ap.add_argument("--gradient-checkpointing", action="store_true")
    a = ap.parse_args()
    real_fd = os.dup(1)
    os.dup2(2, 1)
    info = launch_env.init_distributed("nccl")
    rank, world = info.rank, info.world_size
    dev = torch.device("cuda", torch.cuda.current_device())
    from b200dist.engine.hf_trainer import Zero3CausalLM
    from b200dist.models.llama import build_causal_lm
    from b200dist.ops import native as N
    t0 = time.time()
    torch.manual_seed(0)
    kw = {"num_hidden_layers": a.layers} if a.layers else {}
    from b200dist.models.llama import meta_initializer
    model = build_causal_lm(a.model, device=dev if a.full_init else "meta", dtype=torch.bfloat16, **kw)
    nparams = sum(p.numel() for p in model.parameters())
    init_kw = {} if a.full_init else {"materialize": meta_initializer(model.config, dev, torch.bfloat16, seed=0), "device": dev}
    zm = Zero3CausalLM(model, **init_kw, lr=2e-5, weight_decay=0.0, max_grad_norm=a.clip, gradient_accumulation_steps=a.grad_accum,
                       max_live_parameters=a.max_live_parameters, max_reuse_distance=a.max_reuse_distance,
                       prefetch_bucket_size=int(0.9 * model.config.hidden_size ** 2),
                       persistence_threshold=10 * model.config.hidden_size, comm_backend=a.comm_backend if world > 1 else "nccl",
                       offload_optimizer=a.offload_optimizer, offload_param=a.offload_param, gradient_checkpointing=a.gradient_checkpointing)
    if a.no_fuse_gather_gemm:
        zm.engine._agw_units = set()
    torch.cuda.synchronize()
    setup_s = time.time() - t0
    V = model.config.vocab_size
    g = torch.Generator().manual_seed(1 + rank)
    host = [torch.randint(3, V, (a.micro_batch, a.seq), generator=g).pin_memory() for _ in range(4)]
    ids_dev = torch.empty((a.micro_batch, a.seq), dtype=torch.long, device=dev)

    def step(i):
        loss = None
        for k in range(a.grad_accum):
            ids_dev.copy_(host[(i * a.grad_accum + k) % len(host)], non_blocking=True)
            labels = ids_dev.clone()
            labels[:, : a.seq // 3] = -100            # prompt part masked like the SFT collator does
            loss = zm(ids_dev, labels=labels)
            zm.engine.backward(loss)
            zm.engine.step()                          # acts on the accumulation boundary only
        return loss

    for i in range(a.warmup):
        step(i)
    torch.cuda.synchronize(); dist.barrier()
    sampler = None
    if rank == 0 and not a.no_clocks:
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        from bench import ClockSampler
        sampler = ClockSampler(torch.cuda.current_device())
    l0 = N.launches()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.time()
    s.record()
    last = None
    for i in range(a.steps):
        last = step(i)
    e.record()
    last_loss = float(last.detach())                      # D2H of the step result
    torch.cuda.synchronize(); dist.barrier()
    wall = (time.time() - w0) * 1e3
    t = torch.tensor([max(s.elapsed_time(e), wall)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item() / a.steps
    clocks = sampler.stop() if sampler is not None else {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    tokens = world * a.micro_batch * a.seq * a.grad_accum
    mem = torch.cuda.max_memory_allocated() / 2**30
    if rank == 0:
        flops = 6.0 * nparams * tokens / world + 12.0 * model.config.num_hidden_layers * model.config.hidden_size * a.seq * a.micro_batch * a.grad_accum * a.seq / 2 * 3
        out = {"metric": "LLaMA-7B ZeRO-3 tok/s" if a.model == "llama-7b" and not a.layers else f"{a.model} ZeRO-3 tok/s", "value": tokens / (ms / 1e3),
               "unit": "tokens/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
               "dtype": "bf16", "data": "synthetic instruction batches (pinned host -> device every step), random-init weights",
               "config": {"model": a.model, "layers": model.config.num_hidden_layers, "params": nparams, "seq_len": a.seq, "micro_batch_per_gpu": a.micro_batch, "grad_accum": a.grad_accum,
                          "global_batch": world * a.micro_batch * a.grad_accum, "construction": "whole model per rank" if a.full_init else "meta device, one unit at a time", "gather_workspaces": zm.engine.n_ws, "fused_gather_gemm_units": sorted(zm.engine._agw_units), "prefetch_depth": zm.engine.prefetch_depth, "parallelism": f"zero3 dp{world}", "comm_backend": a.comm_backend,
                          "fabric": zm.engine.fabric is not None, "optimizer": f"AdamW fp32 partitioned, clip {a.clip}",
                          "offload_optimizer": a.offload_optimizer, "offload_param": a.offload_param, "gradient_checkpointing": a.gradient_checkpointing},
               "model_tflops_per_gpu": flops / (ms / 1e3) / 1e12, "peak_mem_gib": mem, "setup_s": setup_s, "last_loss": last_loss,
               "gpu_launches": N.launches() - l0, "clocks": clocks, "grad_norm": zm.engine.last_grad_norm,
               "e2e": {"included": True, "h2d_bytes_per_step": a.grad_accum * a.micro_batch * a.seq * 8, "d2h_bytes_per_step": 4}}
        os.write(real_fd, (json.dumps(out) + "\n").encode())
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
