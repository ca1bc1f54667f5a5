"""GPT layer variants (GQA, RoPE, SwiGLU, squared-ReLU, erf-GELU, LayerNorm-1p, Switch MLP) and the stand-alone
tensor-parallel layers (E10 / E12) on CPU: forward + backward against plain-autograd restatements, and TP=2 (with and
without sequence parallelism) against the single-process model over gloo."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from tests._dist_util import run_distributed


def _cfg(**kw):
    from b200dist.config.megatron_args import TransformerConfig
    base = dict(num_layers=2, hidden_size=64, num_attention_heads=4, ffn_hidden_size=128, kv_channels=16, seq_length=16,
                max_position_embeddings=16, padded_vocab_size=128, hidden_dropout=0.0, attention_dropout=0.0,
                params_dtype=torch.float32, comm_backend="nccl")
    base.update(kw)
    return TransformerConfig(**base)


def _rope_ref(x, S):
    """x [B, S, heads, hd], rotate-half with base 10000"""
    hd = x.shape[-1]
    inv = 1.0 / (10000.0 ** (torch.arange(0, hd, 2).float() / hd))
    ang = torch.arange(S).float()[:, None] * inv[None]
    cos, sin = ang.cos()[None, :, None, :], ang.sin()[None, :, None, :]
    a, b = x[..., :hd // 2], x[..., hd // 2:]
    return torch.cat([a * cos - b * sin, b * cos + a * sin], -1)


def _ref_forward(model, tokens, labels):
    """plain autograd restatement for the flex layer at TP=1 using the model's own parameters"""
    lm, cfg = model.language_model, model.cfg
    B, S = tokens.shape
    H = cfg.hidden_size
    x = lm.embedding.word_embeddings.weight[tokens]
    if hasattr(lm.embedding, "position_embeddings"):
        x = x + lm.embedding.position_embeddings.weight[torch.arange(S)]
    nh, hd = cfg.num_attention_heads, cfg.kv_channels
    ng = cfg.num_query_groups or nh
    nq = nh // ng
    one = 1.0 if cfg.apply_layernorm_1p else 0.0

    def act(z):
        if cfg.swiglu:
            g, u = z.chunk(2, -1)
            return F.silu(g) * u
        if cfg.squared_relu:
            return F.relu(z) ** 2
        if cfg.openai_gelu or cfg.bias_gelu_fusion:
            return F.gelu(z, approximate="tanh")
        return F.gelu(z)

    def mlp(m, h):
        return F.linear(act(F.linear(h, m.dense_h_to_4h.weight, m.dense_h_to_4h.bias)), m.dense_4h_to_h.weight, m.dense_4h_to_h.bias)

    for layer in lm.encoder.layers:
        sa = layer.self_attention
        h = F.layer_norm(x, (H,), layer.input_layernorm.weight + one, layer.input_layernorm.bias, cfg.layernorm_epsilon)
        qkv = F.linear(h, sa.query_key_value.weight, sa.query_key_value.bias).view(B, S, ng, (nq + 2) * hd)
        q = qkv[..., :nq * hd].reshape(B, S, nh, hd)
        k = qkv[..., nq * hd:(nq + 1) * hd]
        v = qkv[..., (nq + 1) * hd:]
        if cfg.position_embedding_type == "rope":
            q, k = _rope_ref(q, S), _rope_ref(k, S)
        k = k.repeat_interleave(nq, dim=2)
        v = v.repeat_interleave(nq, dim=2)
        q, k, v = (t.permute(0, 2, 1, 3) for t in (q, k, v))
        att = (q @ k.transpose(-1, -2)) / math.sqrt(hd)
        att = att.masked_fill(~torch.ones(S, S, dtype=torch.bool).tril(), float("-inf")).softmax(-1)
        ctx = (att @ v).permute(0, 2, 1, 3).reshape(B, S, nh * hd)
        res = h if cfg.apply_residual_connection_post_layernorm else x
        x = res + F.linear(ctx, sa.dense.weight, sa.dense.bias)
        h = F.layer_norm(x, (H,), layer.post_attention_layernorm.weight + one, layer.post_attention_layernorm.bias, cfg.layernorm_epsilon)
        res = h if cfg.apply_residual_connection_post_layernorm else x
        if cfg.num_experts:
            m = layer.mlp
            probs = F.linear(h, m.router.weight, m.router.bias).softmax(-1)
            mp, mi = probs.max(-1)
            out = torch.zeros_like(h)
            for e, ex in enumerate(m.experts):
                sel = (mi == e)
                out = out + torch.where(sel[..., None], mlp(ex, h) * mp[..., None], torch.zeros_like(h))
            x = res + out
        else:
            x = res + mlp(layer.mlp, h)
    fl = lm.encoder.final_layernorm
    x = F.layer_norm(x, (H,), fl.weight, fl.bias, cfg.layernorm_epsilon)
    logits = x @ model.shared_embedding_or_output_weight().t()
    return F.cross_entropy(logits.view(-1, logits.shape[-1]), labels.reshape(-1), reduction="none").view(B, S)


VARIANTS = {
    "swiglu": dict(swiglu=True),
    "rope": dict(position_embedding_type="rope"),
    "gqa": dict(num_query_groups=2),
    "mqa_rope_swiglu_nobias": dict(num_query_groups=1, position_embedding_type="rope", swiglu=True, add_bias_linear=False, bias_gelu_fusion=False),
    "gqa2_rope_swiglu_nobias": dict(num_query_groups=2, position_embedding_type="rope", swiglu=True, add_bias_linear=False, bias_gelu_fusion=False),
    "squared_relu": dict(squared_relu=True),
    "gelu_erf": dict(bias_gelu_fusion=False),
    "layernorm_1p": dict(apply_layernorm_1p=True),
    "post_ln_residual": dict(apply_residual_connection_post_layernorm=True),
    "switch_mlp": dict(num_experts=4),
    "untied_head": dict(untie_embeddings_and_output_weights=True, swiglu=True),
}


@pytest.mark.parametrize("name", sorted(VARIANTS))
def test_variant_matches_autograd_reference(name):
    from b200dist.models.gpt import GPTModel
    from b200dist.models.gpt_flex import FlexTransformerLayer
    torch.manual_seed(0)
    model = GPTModel(_cfg(**VARIANTS[name]), device="cpu", layer_offset=0, num_layers_here=2)
    assert all(isinstance(l, FlexTransformerLayer) for l in model.language_model.encoder.layers)
    with torch.no_grad():
        for n, p in model.named_parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
            if "router.weight" in n:
                p.mul_(20.0)          # spread the routing so several experts receive tokens
    tokens = torch.randint(0, 128, (3, 16))
    labels = torch.randint(0, 128, (3, 16))
    w = torch.rand(3, 16)
    (model(tokens, labels=labels) * w).sum().backward()
    got = {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None}
    out = model(tokens, labels=labels).detach()
    model.zero_grad()
    ref = _ref_forward(model, tokens, labels)
    assert torch.allclose(out, ref, rtol=1e-4, atol=1e-4), (out - ref).abs().max()
    (ref * w).sum().backward()
    checked = 0
    for n, p in model.named_parameters():
        if p.grad is None:
            assert n not in got or got[n].abs().max() == 0, n
            continue
        if n not in got:               # an expert that received no token
            assert "experts" in n and p.grad.abs().max() == 0, f"no gradient for {n}"
            continue
        assert torch.allclose(got[n], p.grad, rtol=2e-3, atol=2e-4), (n, (got[n] - p.grad).abs().max())
        checked += 1
    assert checked >= 10


def test_flex_layer_equals_fused_layer_on_the_default_shape():
    """same seeds, same parameter layout: the modular layer and the hand-scheduled layer are interchangeable"""
    from b200dist.models.gpt import GPTModel
    tokens = torch.randint(0, 128, (2, 16))
    labels = torch.randint(0, 128, (2, 16))
    a = GPTModel(_cfg(), device="cpu", layer_offset=0, num_layers_here=2)
    cfg = _cfg()
    cfg.force_flex_layer = True
    b = GPTModel(cfg, device="cpu", layer_offset=0, num_layers_here=2)
    assert [n for n, _ in a.named_parameters()] == [n for n, _ in b.named_parameters()]
    b.load_state_dict(a.state_dict())
    la, lb = model_loss(a, tokens, labels), model_loss(b, tokens, labels)
    assert torch.allclose(la, lb, rtol=1e-5, atol=1e-5)
    la.sum().backward(); lb.sum().backward()
    for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.allclose(p.grad, q.grad, rtol=1e-4, atol=1e-5), n


def model_loss(m, tokens, labels):
    return m(tokens, labels=labels)


# ---------------------------------------------------------------------------------------------------------------
# TP=2 (+SP) against one process
# ---------------------------------------------------------------------------------------------------------------
def _train_worker(rank, world, tp, sp, variant, steps):
    from b200dist.models.gpt import GPTModel
    from b200dist.optim.mixed import MixedPrecisionOptimizer
    from b200dist.parallel import state as mpu
    mpu.initialize_model_parallel(tp, 1, None)
    cfg = _cfg(sequence_parallel=sp, tensor_model_parallel_size=tp, **VARIANTS[variant])
    model = GPTModel(cfg, device="cpu", seed=11)
    with torch.no_grad():
        for n, p in model.named_parameters():
            if "router.weight" in n:
                p.mul_(20.0)
    opt = MixedPrecisionOptimizer([model], lr=1e-2, weight_decay=0.01, clip_grad=1.0, dp_group=mpu.get_data_parallel_group(),
                                  tp_group=mpu.get_tensor_model_parallel_group(), mp_group=mpu.get_model_parallel_group())
    torch.manual_seed(5)
    data = torch.randint(0, 128, (steps, 2, 17))
    losses = []
    for s in range(steps):
        opt.zero_grad()
        t = data[s]
        loss = model(t[:, :-1].contiguous(), labels=t[:, 1:].contiguous()).mean()
        loss.backward()
        opt.reduce_model_grads()
        opt.step()
        losses.append(float(loss.detach()))
    return losses


@pytest.mark.parametrize("variant,sp", [("gqa2_rope_swiglu_nobias", False), ("gqa", True), ("swiglu", True), ("switch_mlp", False),
                                        ("switch_mlp", True)])
def test_variant_tp2_matches_single(variant, sp):
    ref = run_distributed(_train_worker, 1, 1, False, variant, 3)[0]
    outs = run_distributed(_train_worker, 2, 2, sp, variant, 3)
    assert np.allclose(outs[0], ref, rtol=3e-4, atol=3e-4), (outs[0], ref)
    assert np.allclose(outs[1], ref, rtol=3e-4, atol=3e-4), (outs[1], ref)


# ---------------------------------------------------------------------------------------------------------------
# stand-alone layers
# ---------------------------------------------------------------------------------------------------------------
def _layers_worker(rank, world, sp):
    import torch.distributed as dist
    from b200dist.parallel import state as mpu
    from b200dist.parallel.layers import ColumnParallelLinear, RowParallelLinear, VocabParallelEmbedding, vocab_parallel_cross_entropy
    from b200dist.parallel.tensor import TPComm
    mpu.initialize_model_parallel(world, 1, None)
    comm = TPComm(mpu.get_tensor_model_parallel_group(), sp, "nccl")
    V, H, Fh, S, B = 32, 16, 24, 8, 2
    emb = VocabParallelEmbedding(V, H, seed=1, comm=comm)
    col = ColumnParallelLinear(H, Fh, seed=2, comm=comm)
    row = RowParallelLinear(Fh, V, seed=3, comm=comm)      # output: logits over the full vocabulary, then re-sharded below
    head = ColumnParallelLinear(V, V, bias=False, seed=4, comm=comm)
    with torch.no_grad():
        col.bias.add_(0.1 * torch.arange(Fh).float().cos().chunk(world)[rank])      # column-parallel bias is sharded
        row.bias.add_(0.1 * torch.arange(V).float().cos())
    torch.manual_seed(3)
    ids = torch.randint(0, V, (S, B))
    tgt = torch.randint(0, V, (S, B))
    tgt[0, 0] = -100
    x = emb(ids.reshape(-1))                                   # [S*B (/t), H]
    h = col(x.view(-1, B, H))                                  # [S, B, Fh/t]
    y = row(torch.tanh(h))                                     # [S (/t), B, V]
    logits = head(y)                                           # [S, B, V/t]
    loss = vocab_parallel_cross_entropy(logits, tgt, label_smoothing=0.1, comm=comm)
    loss.sum().backward()
    grads = {}
    for name, m, dim in (("emb", emb, 0), ("col", col, 0), ("row", row, 1), ("head", head, 0)):
        g = m.weight.grad
        if world > 1:
            parts = [torch.empty_like(g) for _ in range(world)]
            dist.all_gather(parts, g.contiguous())
            g = torch.cat(parts, dim)
        grads[name] = g
    for name, m in (("colb", col),):
        g = m.bias.grad
        if world > 1:
            parts = [torch.empty_like(g) for _ in range(world)]
            dist.all_gather(parts, g.contiguous())
            g = torch.cat(parts, 0)
        grads[name] = g
    gb = row.bias.grad.clone()
    if sp and world > 1:
        dist.all_reduce(gb)            # sequence-parallel bias: the optimizer's TP all-reduce
    grads["rowb"] = gb
    return loss.detach(), grads


def _layers_ref():
    from b200dist.parallel.layers import init_shard
    V, H, Fh, S, B = 32, 16, 24, 8, 2
    We = init_shard((V, H), 0.02, 1, "cpu", torch.float32).requires_grad_()
    Wc = init_shard((Fh, H), 0.02, 2, "cpu", torch.float32).requires_grad_()
    Wr = init_shard((V, Fh), 0.02, 3, "cpu", torch.float32).requires_grad_()
    Wh = init_shard((V, V), 0.02, 4, "cpu", torch.float32).requires_grad_()
    bc = (0.1 * torch.arange(Fh).float().cos()).requires_grad_()
    br = (0.1 * torch.arange(V).float().cos()).requires_grad_()
    torch.manual_seed(3)
    ids = torch.randint(0, V, (S, B))
    tgt = torch.randint(0, V, (S, B))
    tgt[0, 0] = -100
    logits = F.linear(F.linear(torch.tanh(F.linear(We[ids], Wc, bc)), Wr, br), Wh)
    # the reference's definition (megatron/core/tensor_parallel/cross_entropy.py): a = ls V / (V - 1), (1 - a) nll - a mean(log p)
    logp = F.log_softmax(logits.view(-1, V), -1)
    nll = F.nll_loss(logp, tgt.view(-1), reduction="none", ignore_index=-100)
    a = 0.1 * V / (V - 1)
    loss = torch.where(tgt.view(-1) != -100, (1 - a) * nll - a * logp.mean(-1), torch.zeros_like(nll)).view(S, B)
    loss.sum().backward()
    return loss.detach(), {"emb": We.grad, "col": Wc.grad, "row": Wr.grad, "head": Wh.grad, "colb": bc.grad, "rowb": br.grad}


@pytest.mark.parametrize("world,sp", [(1, False), (2, False), (2, True)])
def test_parallel_layers_and_label_smoothed_loss(world, sp):
    ref_loss, ref_g = _layers_ref()
    outs = run_distributed(_layers_worker, world, sp)
    for loss, g in outs:
        assert np.allclose(loss, ref_loss.numpy(), rtol=1e-4, atol=1e-5)
        for k in ref_g:
            assert np.allclose(g[k], ref_g[k].numpy(), rtol=1e-3, atol=1e-5), (k, np.abs(g[k] - ref_g[k].numpy()).max())


def test_unfused_core_attention_matches_flash_path():
    """--attention-impl unfused (the reference's CoreAttention) against the default attention on the same weights"""
    from b200dist.models.gpt import GPTModel
    tokens, labels = torch.randint(0, 128, (2, 16)), torch.randint(0, 128, (2, 16))
    a = GPTModel(_cfg(position_embedding_type="rope"), device="cpu", layer_offset=0, num_layers_here=2)
    for fusion in (True, False):
        b = GPTModel(_cfg(position_embedding_type="rope", attention_impl="unfused", masked_softmax_fusion=fusion), device="cpu", layer_offset=0, num_layers_here=2)
        b.load_state_dict(a.state_dict())
        a.zero_grad()
        la, lb = a(tokens, labels=labels), b(tokens, labels=labels)
        assert torch.allclose(la, lb, rtol=1e-5, atol=1e-5)
        la.sum().backward(); lb.sum().backward()
        for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
            assert torch.allclose(p.grad, q.grad, rtol=1e-3, atol=1e-5), n


def test_scale_mask_softmax_op_and_module():
    from b200dist.ops import functional as Fn
    torch.manual_seed(0)
    x = torch.randn(2, 3, 8, 12, requires_grad=True)
    mask = torch.rand(2, 1, 8, 12) > 0.7
    y = Fn.scaled_masked_softmax(x, mask, 0.5)
    ref = torch.softmax((x * 0.5).masked_fill(mask, -10000.0), -1)
    assert torch.allclose(y, ref, atol=1e-6)
    g = torch.randn_like(y)
    (gx,) = torch.autograd.grad(y, x, g)
    (gr,) = torch.autograd.grad(ref, x, g)
    assert torch.allclose(gx, gr, atol=1e-6)
    yc = Fn.scaled_upper_triang_masked_softmax(x, 2.0)           # sq < sk: keys j <= i + (sk - sq) visible
    keep = torch.ones(8, 12, dtype=torch.bool).tril(diagonal=4)
    refc = torch.softmax((x * 2.0).masked_fill(~keep, float("-inf")), -1)
    assert torch.allclose(yc, refc, atol=1e-6) and float(yc.detach()[..., 0, 5:].abs().max()) == 0.0
    m = Fn.FusedScaleMaskSoftmax(input_in_bf16=True, attn_mask_type="padding", scale=0.5)
    xb = x.detach().to(torch.bfloat16)
    assert m.is_kernel_available(mask, *xb.shape)
    assert torch.allclose(m(xb, mask).float(), ref.detach(), atol=2e-2)
    m2 = Fn.FusedScaleMaskSoftmax(input_in_bf16=False, attn_mask_type="causal", scaled_masked_softmax_fusion=False, scale=2.0)
    assert torch.allclose(m2(x.detach()), refc.detach(), atol=1e-6)


def test_partial_rotary_rotates_only_the_leading_dims_of_every_head():
    """--rotary-percent < 1 (arguments.py:578): the first int(head_dim * percent) dims of each head get rotate-half RoPE with frequencies
    over that sub-dimension, the rest pass through; forward and backward against the formula"""
    import torch
    from b200dist.models.gpt_flex import _RopeNative
    from b200dist.ops import functional as Fn
    torch.manual_seed(0)
    T, nh, hd, rot = 6, 3, 16, 8
    x = torch.randn(T, nh * hd, requires_grad=True)
    pos = torch.tensor([0, 0, 1, 1, 2, 5], dtype=torch.int32)
    cos_t, sin_t = Fn.rope_tables(8, rot)
    y = _RopeNative.apply(x, cos_t, sin_t, pos, nh, hd, rot)
    xr = x.detach().clone().requires_grad_()
    xh = xr.view(T, nh, hd)
    a, b, rest = xh[..., :rot // 2], xh[..., rot // 2:rot], xh[..., rot:]
    c, s = cos_t[pos.long()][:, None, :], sin_t[pos.long()][:, None, :]
    ref = torch.cat([a * c - b * s, b * c + a * s, rest], -1).reshape(T, nh * hd)
    assert torch.allclose(y, ref, atol=1e-6)
    g = torch.randn_like(ref)
    y.backward(g); ref.backward(g)
    assert torch.allclose(x.grad, xr.grad, atol=1e-6)
    assert torch.equal(y.view(T, nh, hd)[..., rot:], x.detach().view(T, nh, hd)[..., rot:])


def test_init_method_xavier_uniform_covers_every_weight_matrix():
    """--init-method-xavier-uniform (arguments.py:435-437): weights are uniform within sqrt(6 / (fan_in + fan_out)) of the FULL matrix in both
    layer implementations, and the switch does not leak into the next model"""
    import math
    import torch
    from b200dist.config.megatron_args import TransformerConfig
    from b200dist.models.gpt import GPTModel
    kw = dict(num_layers=1, hidden_size=64, num_attention_heads=4, ffn_hidden_size=256, kv_channels=16, seq_length=16, max_position_embeddings=16,
              padded_vocab_size=128, params_dtype=torch.float32)
    for extra in ({}, {"swiglu": True}):
        m = GPTModel(TransformerConfig(init_method_xavier_uniform=True, **kw, **extra), device="cpu", layer_offset=0, num_layers_here=1)
        for n, p in m.named_parameters():
            if p.dim() == 2 and "position_embeddings" not in n:
                bound = math.sqrt(6.0 / (p.shape[0] + p.shape[1]))
                if "dense_h_to_4h" in n and extra:          # gate and up are drawn as two [F, H] matrices
                    bound = math.sqrt(6.0 / (256 + 64))
                assert p.abs().max().item() <= bound + 1e-6 and p.std().item() > 0.5 * bound, n
    plain = GPTModel(TransformerConfig(**kw), device="cpu", layer_offset=0, num_layers_here=1)
    assert abs(plain.language_model.encoder.layers[0].mlp.dense_h_to_4h.weight.std().item() - 0.02) < 2e-3


@pytest.mark.parametrize("fused_softmax", [True, False])
def test_reset_attention_mask_isolates_documents(fused_softmax):
    """--reset-attention-mask + --reset-position-ids (pretrain_gpt.py:83-88): with the document ids installed, every document of a packed
    sequence gets the logits it gets when it is run alone; also through the full-recompute backward"""
    import importlib.util
    import os
    import torch
    from b200dist.config.megatron_args import TransformerConfig
    from b200dist.models.gpt import GPTModel
    spec = importlib.util.spec_from_file_location("pretrain_gpt_entry2", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "entrypoints", "pretrain_gpt.py"))
    entry = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(entry)
    torch.manual_seed(0)
    eod, S = 99, 24
    kw = dict(num_layers=2, hidden_size=64, num_attention_heads=4, ffn_hidden_size=128, kv_channels=16, seq_length=S, max_position_embeddings=S,
              padded_vocab_size=128, params_dtype=torch.float32, hidden_dropout=0.0, attention_dropout=0.0, attention_impl="unfused",
              masked_softmax_fusion=fused_softmax)
    m = GPTModel(TransformerConfig(**kw), device="cpu", layer_offset=0, num_layers_here=2).eval()
    tokens = torch.randint(0, 90, (2, S))
    tokens[0, 7], tokens[0, 15], tokens[1, 11] = eod, eod, eod
    pos = entry.reset_position_ids_after_eod(tokens, eod)
    is_eod = (tokens == eod).long()
    doc = torch.cumsum(is_eod, 1) - is_eod
    with torch.no_grad():
        full = m(tokens, pos, doc)                                   # [S, B, V]
        for b, (a, e) in ((0, (0, 8)), (0, (8, 16)), (0, (16, S)), (1, (0, 12)), (1, (12, S))):
            alone = m(tokens[b:b + 1, a:e].contiguous(), None, None)
            assert torch.allclose(full[a:e, b], alone[:, 0], atol=2e-5), (b, a, e, (full[a:e, b] - alone[:, 0]).abs().max())
        causal = m(tokens, pos, None)
    assert not torch.allclose(full[8:16, 0], causal[8:16, 0], atol=1e-3)            # without the mask the documents leak into each other
    # backward with full recompute re-installs the document ids of its own forward
    mt = GPTModel(TransformerConfig(recompute_granularity="full", **kw), device="cpu", layer_offset=0, num_layers_here=2).train()
    mt.load_state_dict(m.state_dict())
    labels = torch.randint(0, 90, (2, S))
    loss = mt(tokens, pos, doc, labels=labels).mean()
    from b200dist.parallel import tensor as tp_ops
    tp_ops.set_document_ids(None)                                    # a later micro-batch without document ids is now "in flight"
    loss.backward()
    ref = GPTModel(TransformerConfig(**kw), device="cpu", layer_offset=0, num_layers_here=2).train()
    ref.load_state_dict(m.state_dict())
    ref(tokens, pos, doc, labels=labels).mean().backward()
    for (n, p), (_, q) in zip(mt.named_parameters(), ref.named_parameters()):
        if q.grad is not None:
            assert torch.allclose(p.grad, q.grad, atol=1e-6), n
