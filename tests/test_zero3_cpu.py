"""ZeRO-3 partition / fetch / reduce / step bookkeeping on CPU (gloo, world 2) against plain single-process AdamW."""
import numpy as np
import torch

from tests._dist_util import run_distributed


def _build(seed=0):
    from b200dist.models.llama import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(seed)
    return LlamaForCausalLM(LlamaConfig.tiny())


def _run_engine(rank, world, steps, ga, clip):
    from b200dist.engine.hf_trainer import Zero3CausalLM
    model = _build()
    zm = Zero3CausalLM(model, lr=1e-2, weight_decay=0.0, max_grad_norm=clip, gradient_accumulation_steps=ga, persistence_threshold=200, comm_backend="nccl")
    torch.manual_seed(3)
    data = torch.randint(0, 512, (steps * ga, world, 2, 16))
    losses = []
    for i in range(steps * ga):
        ids = data[i, rank]
        loss = zm(ids, labels=ids)
        zm.engine.backward(loss)
        zm.engine.step()
        losses.append(float(loss))
    sd = zm.engine.gather_full_state_dict()
    key = sorted(sd)[3]
    return losses, sd[key].float().sum().item(), zm.engine.total_params()


def _run_reference(world, steps, ga, clip):
    model = _build()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.0)
    torch.manual_seed(3)
    data = torch.randint(0, 512, (steps * ga, world, 2, 16))
    losses = []
    for s in range(steps):
        opt.zero_grad()
        for g in range(ga):
            per_rank = []
            for r in range(world):
                ids = data[s * ga + g, r]
                loss = model(ids, labels=ids)
                (loss / (world * ga)).backward()
                per_rank.append(float(loss))
            losses.append(per_rank)
        if clip:
            torch.nn.utils.clip_grad_norm_(model.parameters(), clip)
        opt.step()
    return losses, model


def test_zero3_matches_adamw_with_clipping_and_accumulation():
    world, steps, ga, clip = 2, 3, 2, 1.0
    outs = run_distributed(_run_engine, world, steps, ga, clip)
    ref_losses, ref_model = _run_reference(world, steps, ga, clip)
    for r in range(world):
        mine = outs[r][0]
        want = [ref_losses[i][r] for i in range(steps * ga)]
        assert np.allclose(mine, want, rtol=2e-4, atol=2e-4), (r, mine, want)
    assert outs[0][2] == sum(p.numel() for p in ref_model.parameters())


def test_zero3_fused_adam_path_without_clipping():
    world, steps, ga, clip = 2, 2, 1, 0.0
    outs = run_distributed(_run_engine, world, steps, ga, clip)
    ref_losses, _ = _run_reference(world, steps, ga, clip)
    for r in range(world):
        want = [ref_losses[i][r] for i in range(steps * ga)]
        assert np.allclose(outs[r][0], want, rtol=2e-4, atol=2e-4)
