"""ZeRO-3 partition / fetch / reduce / step bookkeeping on CPU (gloo, world 2) against plain single-process AdamW."""
import numpy as np
import torch

from tests._dist_util import run_distributed


def _build(seed=0):
    from b200dist.models.llama import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(seed)
    return LlamaForCausalLM(LlamaConfig.tiny())


def _run_engine(rank, world, steps, ga, clip, offload_optimizer="none", gradient_checkpointing=False):
    from b200dist.engine.hf_trainer import Zero3CausalLM
    model = _build()
    zm = Zero3CausalLM(model, lr=1e-2, weight_decay=0.0, max_grad_norm=clip, gradient_accumulation_steps=ga, persistence_threshold=200, comm_backend="nccl",
                       offload_optimizer=offload_optimizer, gradient_checkpointing=gradient_checkpointing)
    torch.manual_seed(3)
    data = torch.randint(0, 512, (steps * ga, world, 2, 16))
    losses = []
    for i in range(steps * ga):
        ids = data[i, rank]
        loss = zm(ids, labels=ids)
        zm.engine.backward(loss)
        zm.engine.step()
        losses.append(float(loss.detach()))
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


def test_zero3_cpu_offloaded_optimizer_matches_adamw():
    """offload_optimizer.device = cpu (the reference's `-original` config): host fp32 shards + the C++ CPU Adam"""
    world, steps, ga, clip = 2, 3, 2, 1.0
    outs = run_distributed(_run_engine, world, steps, ga, clip, "cpu")
    ref_losses, _ = _run_reference(world, steps, ga, clip)
    for r in range(world):
        want = [ref_losses[i][r] for i in range(steps * ga)]
        assert np.allclose(outs[r][0], want, rtol=2e-4, atol=2e-4), (r, outs[r][0], want)


def test_cpu_adam_kernel_matches_torch_both_code_paths():
    from b200dist.optim.cpu_adam import CPUAdam, cpu_adam_, cpu_sumsq, simd_width
    assert simd_width() in (8, 16)
    torch.manual_seed(0)
    n = 3 * 16384 + 37                      # several tiles + a ragged tail
    for adamw, wd in ((True, 0.1), (False, 0.05), (True, 0.0)):
        for generic in (False, True):
            p0 = torch.randn(n)
            ref = p0.clone().requires_grad_()
            opt = (torch.optim.AdamW if adamw else torch.optim.Adam)([ref], lr=3e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=wd)
            p, m, v = p0.clone(), torch.zeros(n), torch.zeros(n)
            lp = torch.empty(n, dtype=torch.bfloat16)
            lp16 = torch.empty(n, dtype=torch.float16)
            for step in range(1, 4):
                g = torch.randn(n)
                ref.grad = g.clone() * 0.5
                opt.step()
                cpu_adam_(p, m, v, g, lp if step < 3 else lp16, lr=3e-3, beta1=0.9, beta2=0.95, eps=1e-8, wd=wd, step=step, adamw=adamw,
                          grad_scale=0.5, threads=3, force_generic=generic)
            assert torch.allclose(p, ref.detach(), rtol=1e-5, atol=1e-6), (adamw, wd, generic, (p - ref.detach()).abs().max())
            assert torch.equal(lp16, p.to(torch.float16))
            assert abs(cpu_sumsq(g) - float(g.double().pow(2).sum())) < 1e-6 * n
    # bf16 copy is round-to-nearest-even of the fp32 result
    p, m, v, g = torch.randn(n), torch.zeros(n), torch.zeros(n), torch.randn(n)
    lp = torch.empty(n, dtype=torch.bfloat16)
    cpu_adam_(p, m, v, g, lp, lr=1e-2, step=1)
    assert torch.equal(lp, p.to(torch.bfloat16))
    # optimizer wrapper
    w = torch.nn.Parameter(torch.randn(1000))
    w2 = torch.nn.Parameter(w.detach().clone())
    a, b = CPUAdam([w], lr=1e-2, weight_decay=0.01), torch.optim.AdamW([w2], lr=1e-2, weight_decay=0.01)
    for _ in range(3):
        gr = torch.randn(1000)
        w.grad, w2.grad = gr.clone(), gr.clone()
        a.step(); b.step()
    assert torch.allclose(w, w2, rtol=1e-5, atol=1e-6)


def test_zero3_gradient_checkpointing_is_loss_identical():
    """--gradient_checkpointing True: decoder layers keep only their input and are re-run in backward"""
    world, steps, ga, clip = 2, 3, 2, 1.0
    plain = run_distributed(_run_engine, world, steps, ga, clip)
    ckpt = run_distributed(_run_engine, world, steps, ga, clip, "none", True)
    for r in range(world):
        assert plain[r][0] == ckpt[r][0], (plain[r][0], ckpt[r][0])
        assert plain[r][1] == ckpt[r][1]


def test_zero3_meta_construction_never_holds_the_model_and_matches_the_initializer():
    """zero.Init-style construction: modules on the meta device + a name-seeded initializer; the engine creates one unit at a
    time.  Every gathered parameter equals what the initializer returns for its name, and training runs."""
    import torch
    from b200dist.engine.hf_trainer import Zero3CausalLM
    from b200dist.models.llama import build_causal_lm, meta_initializer
    for name in ("llama-tiny", "opt-tiny"):
        model = build_causal_lm(name, device="meta", dtype=torch.float32)
        assert all(p.device.type == "meta" for p in model.parameters())
        init = meta_initializer(model.config, "cpu", torch.float32, seed=3)
        zm = Zero3CausalLM(model, lr=1e-2, max_grad_norm=1.0, persistence_threshold=200, comm_backend="nccl", materialize=init, device="cpu")
        sd = zm.engine.gather_full_state_dict()
        assert len(sd) > 4
        for k, v in sd.items():
            want = init(k, torch.empty(v.shape, device="meta"))
            assert torch.equal(v, want), k
        for p in zm.engine.persistent:
            assert p.device.type == "cpu"
        ids = torch.randint(0, 512, (2, 32), generator=torch.Generator().manual_seed(0))
        losses = []
        for _ in range(4):
            loss = zm(ids, labels=ids)
            zm.engine.backward(loss)
            zm.engine.step()
            losses.append(float(loss.detach()))
        assert losses[-1] < losses[0], losses
