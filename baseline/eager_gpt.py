"""On-box LIBRARY baseline (not the reference, not the product): the same GPT-2-small training step written
with stock PyTorch - nn.Linear (cuBLAS), F.scaled_dot_product_attention (library flash attention),
nn.LayerNorm, F.cross_entropy, torch.optim.AdamW(fused=True), bf16 autocast with fp32 master weights - and
NCCL for data parallelism (torch DDP), the reference step's hidden / attention dropout (--dropout), logits and cross-
entropy kept in bf16 (fp32 accumulation inside the kernels, no fp32 copy of the logits).  It is what "a path that only calls cuBLAS/NCCL" achieves on this
machine, measured with the same timing rules as bench.py.  Tensor parallelism is not implemented in this
baseline: it always runs pure data parallel (dp = gpus), which needs no TP communication at all and is
therefore an upper bound for a library TP=2 run of this small model.
"""
from __future__ import annotations

import json
import os
import sys
import time

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


class Block(nn.Module):
    def __init__(self, h, nh, p=0.0):
        super().__init__()
        self.p = p
        self.ln1, self.ln2 = nn.LayerNorm(h), nn.LayerNorm(h)
        self.qkv, self.proj = nn.Linear(h, 3 * h), nn.Linear(h, h)
        self.fc1, self.fc2 = nn.Linear(h, 4 * h), nn.Linear(4 * h, h)
        self.nh = nh

    def forward(self, x):
        B, S, H = x.shape
        q, k, v = self.qkv(self.ln1(x)).view(B, S, 3, self.nh, H // self.nh).permute(2, 0, 3, 1, 4)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True, dropout_p=self.p if self.training else 0.0).transpose(1, 2).reshape(B, S, H)
        x = x + F.dropout(self.proj(a), self.p, self.training)
        return x + F.dropout(self.fc2(F.gelu(self.fc1(self.ln2(x)), approximate="tanh")), self.p, self.training)


class EagerGPT(nn.Module):
    def __init__(self, L, h, nh, vocab, seq, p=0.0):
        super().__init__()
        self.p = p
        self.wte, self.wpe = nn.Embedding(vocab, h), nn.Embedding(seq, h)
        self.blocks = nn.ModuleList([Block(h, nh, p) for _ in range(L)])
        self.lnf = nn.LayerNorm(h)

    def forward(self, tokens, labels):
        pos = torch.arange(tokens.shape[1], device=tokens.device)
        x = F.dropout(self.wte(tokens) + self.wpe(pos), self.p, self.training)
        for b in self.blocks:
            x = b(x)
        logits = F.linear(self.lnf(x), self.wte.weight)
        return F.cross_entropy(logits.view(-1, logits.shape[-1]), labels.reshape(-1))


def run_eager(a, tp, dp, gbs, ClockSampler, emit=print):
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    lr_ = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr_)
    dev = torch.device("cuda", lr_)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29500")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.manual_seed(1234)
    model = EagerGPT(a.num_layers, a.hidden_size, a.heads, a.vocab, a.seq_length, getattr(a, "dropout", 0.0)).to(dev)
    model.train()
    ddp = nn.parallel.DistributedDataParallel(model, device_ids=[lr_]) if world > 1 else model
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1, fused=True)
    nmb = gbs // (a.micro_batch_size * world)
    g = torch.Generator().manual_seed(1234 + rank)
    host = [torch.randint(0, a.vocab, (a.micro_batch_size, a.seq_length + 1), generator=g).pin_memory() for _ in range(nmb * 2)]
    devb = [b.to(dev) for b in host]

    def step(i, from_host, read):
        opt.zero_grad(set_to_none=True)
        tot = None
        for m in range(nmb):
            b = host[(i * nmb + m) % len(host)].to(dev, non_blocking=True) if from_host else devb[(i * nmb + m) % len(devb)]
            ctx = ddp.no_sync() if (world > 1 and m < nmb - 1) else torch.autocast("cuda", enabled=False)
            with ctx:
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    loss = ddp(b[:, :-1], b[:, 1:]) / nmb
                loss.backward()
            tot = loss.detach() if tot is None else tot + loss.detach()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        return float(tot) if read else None

    def timed(n, from_host, read):
        dist.barrier(); torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); s.record()
        last = None
        for i in range(n):
            last = step(i, from_host, read)
        e.record(); torch.cuda.synchronize(); dist.barrier()
        ms = torch.tensor([s.elapsed_time(e), (time.perf_counter() - t0) * 1e3], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms[0].item(), ms[1].item(), last

    for i in range(a.warmup):
        step(i, False, False)
    sampler = None if (a.no_clocks or rank != 0) else ClockSampler(lr_)
    ms_dev, _, _ = timed(a.steps, False, False)
    ms_e2e, wall, last = timed(a.steps, True, True)
    clocks = sampler.stop() if sampler else {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    tokens = gbs * a.seq_length
    if rank == 0:
        emit(json.dumps({"metric": "GPT tok/s (Megatron TP/PP)", "value": tokens * a.steps / (ms_dev / 1e3), "unit": "tokens/s", "n_gpus": world,
                          "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_dev / a.steps, "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "bf16 autocast", "data": "synthetic tokens, random-init weights", "impl": "eager-library-baseline",
                          "config": {"model": "GPT-2 small (stock PyTorch ops)", "global_batch": gbs, "seq_len": a.seq_length, "vocab": a.vocab,
                                     "parallelism": f"dp{world} (torch DDP/NCCL)", "dropout": getattr(a, "dropout", 0.0)},
                          "clocks": clocks, "e2e": {"value": tokens * a.steps / (max(ms_e2e, wall) / 1e3), "unit": "tokens/s",
                                                    "h2d_bytes_per_step": nmb * a.micro_batch_size * (a.seq_length + 1) * 8, "d2h_bytes_per_step": 4, "last_loss": last},
                          "gpu_launches": 0}))
    dist.barrier(); dist.destroy_process_group()
    return 0
