"""Model-parity tier on CPU: the hand-scheduled GPT layer (forward AND backward) against a
plain-autograd PyTorch restatement; TP=2 with and without sequence parallelism, and PP=2 (1F1B and interleaved)
against the single-process model, over gloo."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from tests._dist_util import run_distributed


def _cfg(**kw):
    from b200dist.config.megatron_args import TransformerConfig
    base = dict(num_layers=2, hidden_size=64, num_attention_heads=4, ffn_hidden_size=256, kv_channels=16, seq_length=16,
                max_position_embeddings=16, padded_vocab_size=128, hidden_dropout=0.0, attention_dropout=0.0,
                params_dtype=torch.float32, comm_backend="nccl")
    base.update(kw)
    return TransformerConfig(**base)


def _ref_forward(model, tokens, labels):
    """Plain autograd GPT using the SAME parameters (TP=1 layout)."""
    lm = model.language_model
    cfg = model.cfg
    B, S = tokens.shape
    pos = torch.arange(S)
    x = lm.embedding.word_embeddings.weight[tokens] + lm.embedding.position_embeddings.weight[pos]   # [B,S,H]
    nh, hd = cfg.num_attention_heads, cfg.kv_channels
    for layer in lm.encoder.layers:
        sa, mlp = layer.self_attention, layer.mlp
        h = F.layer_norm(x, (cfg.hidden_size,), layer.input_layernorm.weight, layer.input_layernorm.bias, cfg.layernorm_epsilon)
        qkv = F.linear(h, sa.query_key_value.weight, sa.query_key_value.bias).view(B, S, nh, 3 * hd)
        q, k, v = (t.permute(0, 2, 1, 3) for t in qkv.split(hd, dim=-1))
        att = (q @ k.transpose(-1, -2)) / math.sqrt(hd)
        att = att.masked_fill(~torch.ones(S, S, dtype=torch.bool).tril(), float("-inf")).softmax(-1)
        ctx = (att @ v).permute(0, 2, 1, 3).reshape(B, S, nh * hd)
        x = x + F.linear(ctx, sa.dense.weight, sa.dense.bias)
        h = F.layer_norm(x, (cfg.hidden_size,), layer.post_attention_layernorm.weight, layer.post_attention_layernorm.bias, cfg.layernorm_epsilon)
        x = x + F.linear(F.gelu(F.linear(h, mlp.dense_h_to_4h.weight, mlp.dense_h_to_4h.bias), approximate="tanh"), mlp.dense_4h_to_h.weight, mlp.dense_4h_to_h.bias)
    fl = lm.encoder.final_layernorm
    x = F.layer_norm(x, (cfg.hidden_size,), fl.weight, fl.bias, cfg.layernorm_epsilon)
    logits = x @ lm.embedding.word_embeddings.weight.t()
    return F.cross_entropy(logits.view(-1, logits.shape[-1]), labels.reshape(-1), reduction="none").view(B, S)


def test_gpt_forward_backward_matches_autograd_reference():
    from b200dist.models.gpt import GPTModel
    torch.manual_seed(0)
    model = GPTModel(_cfg(), device="cpu", layer_offset=0, num_layers_here=2)
    with torch.no_grad():    # non-trivial biases / norm params so their gradients are exercised
        for n, p in model.named_parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    tokens = torch.randint(0, 128, (3, 16))
    labels = torch.randint(0, 128, (3, 16))
    w = torch.rand(3, 16)
    loss = (model(tokens, labels=labels) * w).sum()
    loss.backward()
    got = {n: p.grad.clone() for n, p in model.named_parameters()}
    model.zero_grad()
    ref = (_ref_forward(model, tokens, labels) * w).sum()
    ref.backward()
    assert abs(loss.item() - ref.item()) < 1e-3 * abs(ref.item())
    for n, p in model.named_parameters():
        assert torch.allclose(got[n], p.grad, atol=2e-4, rtol=2e-3), (n, (got[n] - p.grad).abs().max())


def test_gpt2_small_param_count_tp4():
    """the logged reference run: 31,825,152 parameters per rank at TP=4, vocab 50688 (3_...ipynb:1844)."""
    H, L, V, S, t = 768, 12, 50688, 1024, 4
    per_layer = (2 * H + 2 * H) + (3 * H * H // t + 3 * H // t) + (H * H // t + H) + (4 * H * H // t + 4 * H // t) + (4 * H * H // t + H)
    total = V // t * H + S * H + L * per_layer + 2 * H
    assert total == 31_825_152
    from b200dist.models.gpt import GPTModel, count_parameters
    m = GPTModel(_cfg(), device="cpu", layer_offset=0, num_layers_here=2)
    H, V, S = 64, 128, 16
    per = 4 * H + 3 * H * H + 3 * H + H * H + H + 4 * H * H + 4 * H + 4 * H * H + H
    assert count_parameters(m) == V * H + S * H + 2 * per + 2 * H


def _train_worker(rank, world, tp, pp_size, vpp_layers, sp, steps):
    """runs `steps` optimizer steps; returns the per-step losses (last pipeline stage) and a parameter checksum"""
    import torch.distributed as dist
    from b200dist.models.gpt import GPTModel
    from b200dist.optim.mixed import MixedPrecisionOptimizer
    from b200dist.parallel import pipeline as pp
    from b200dist.parallel import state as mpu
    from b200dist.parallel import tensor as tpo
    L = 4
    vpp = None
    if vpp_layers:
        vpp = (L // pp_size) // vpp_layers
    mpu.initialize_model_parallel(tp, pp_size, vpp)
    pp.reset_links()
    cfg = _cfg(num_layers=L, sequence_parallel=sp, tensor_model_parallel_size=tp, pipeline_model_parallel_size=pp_size,
               virtual_pipeline_model_parallel_size=vpp)
    chunks = []
    for v in range(vpp or 1):
        if vpp:
            mpu.set_virtual_pipeline_model_parallel_rank(v)
        chunks.append(GPTModel(cfg, pre_process=mpu.is_pipeline_first_stage(), post_process=mpu.is_pipeline_last_stage(), device="cpu", seed=11))
    if vpp:
        mpu.set_virtual_pipeline_model_parallel_rank(0)
    if pp_size > 1:
        for m in chunks:
            if m.post_process and not m.pre_process:
                w = m.word_embeddings_weight(); w.shared_duplicate = True; w.in_embedding_group = True
                with torch.no_grad():
                    w.zero_()
            elif m.pre_process:
                m.word_embeddings_weight().in_embedding_group = True
        if mpu.is_rank_in_embedding_group(ignore_virtual=True):
            for m in chunks:
                if hasattr(m.language_model, "embedding"):
                    dist.all_reduce(m.word_embeddings_weight().data, group=mpu.get_embedding_group())
    emb_group = mpu.get_embedding_group() if (pp_size > 1 and mpu.is_rank_in_embedding_group(ignore_virtual=True)) else None
    opt = MixedPrecisionOptimizer(chunks, lr=1e-2, weight_decay=0.01, clip_grad=1.0, dp_group=mpu.get_data_parallel_group(),
                                  tp_group=mpu.get_tensor_model_parallel_group(), mp_group=mpu.get_model_parallel_group(), embedding_group=emb_group)
    torch.manual_seed(5)
    M, B, S = 4, 2, 16
    data = torch.randint(0, 128, (steps, M, B, S + 1))
    losses = []

    for s in range(steps):
        it = iter([data[s, m] for m in range(M)])
        its = [iter([data[s, m] for m in range(M)]) for _ in chunks]

        def fwd_step(data_iter, model):
            t = next(data_iter)
            out = model(t[:, :-1].contiguous(), labels=t[:, 1:].contiguous())
            return out, lambda o: (o.mean(), {"lm loss": o.mean().detach()})
        opt.zero_grad()
        Sl = S // (tp if sp else 1)
        red = pp.forward_backward(fwd_step, its if len(chunks) > 1 else its[0], chunks if len(chunks) > 1 else chunks[0], M, (Sl, B, 64), torch.float32)
        opt.reduce_model_grads()
        opt.step()
        if red:
            losses.append(float(sum(r["lm loss"] for r in red) / len(red)))
    # checksum of the first layer's LN weight + (tp-gathered) qkv weight norm for cross-config comparison
    return losses


def _single_process_losses(steps):
    out = run_distributed(_train_worker, 1, 1, 1, 0, False, steps)
    return out[0]


@pytest.fixture(scope="module")
def ref_losses():
    return _single_process_losses(3)


def test_tp2_matches_single(ref_losses):
    outs = run_distributed(_train_worker, 2, 2, 1, 0, False, 3)
    assert np.allclose(outs[0], ref_losses, rtol=2e-4, atol=2e-4), (outs[0], ref_losses)


def test_tp2_sequence_parallel_matches_single(ref_losses):
    outs = run_distributed(_train_worker, 2, 2, 1, 0, True, 3)
    assert np.allclose(outs[0], ref_losses, rtol=2e-4, atol=2e-4), (outs[0], ref_losses)


def test_pp2_1f1b_matches_single(ref_losses):
    outs = run_distributed(_train_worker, 2, 1, 2, 0, False, 3)
    assert np.allclose(outs[1], ref_losses, rtol=2e-4, atol=2e-4), (outs[1], ref_losses)


def test_pp2_interleaved_matches_single(ref_losses):
    # 4 layers, PP=2, 1 layer per virtual stage -> 2 chunks per rank: the PP=2 interleaved case the reference rejects
    outs = run_distributed(_train_worker, 2, 1, 2, 1, False, 3)
    assert np.allclose(outs[1], ref_losses, rtol=2e-4, atol=2e-4), (outs[1], ref_losses)


def test_dp2_matches_single_on_same_global_batch():
    # dp=2: each rank sees the same data here, so the averaged gradient equals the single-process one
    ref = _single_process_losses(2)
    outs = run_distributed(_train_worker, 2, 1, 1, 0, False, 2)
    assert np.allclose(outs[0], ref, rtol=2e-4, atol=2e-4)


def test_optimizer_loads_the_reference_stack_optimizer_layout():
    """an `optimizer` checkpoint section in Megatron's layout (torch-Adam state_dict + fp32_from_fp16_params, two parameter
    groups in module order) lands in the flat arenas exactly like this framework's own layout"""
    import torch
    from b200dist.config.megatron_args import TransformerConfig
    from b200dist.models.gpt import GPTModel
    from b200dist.optim.mixed import MixedPrecisionOptimizer, default_no_weight_decay
    cfg = TransformerConfig(num_layers=2, hidden_size=32, num_attention_heads=4, ffn_hidden_size=64, kv_channels=8, seq_length=16,
                            max_position_embeddings=16, padded_vocab_size=64, hidden_dropout=0.0, attention_dropout=0.0,
                            params_dtype=torch.bfloat16, bf16=True)
    torch.manual_seed(0)
    model = GPTModel(cfg, device=torch.device("cpu"), layer_offset=0, num_layers_here=2)
    a = MixedPrecisionOptimizer([model], lr=1e-3, weight_decay=0.1, clip_grad=1.0)
    tokens = torch.randint(0, 64, (2, 17))
    for _ in range(2):
        a.zero_grad()
        model(tokens[:, :-1].contiguous(), labels=tokens[:, 1:].contiguous()).mean().backward()
        a.reduce_model_grads(); a.step()
    own = a.state_dict()
    # restate it the way Megatron's Float16OptimizerWithFloat16Params writes it
    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    groups = [[(n, p) for n, p in named if not default_no_weight_decay(n, p)], [(n, p) for n, p in named if default_no_weight_decay(n, p)]]
    state, pg, fp32, idx = {}, [], [], 0
    for g in groups:
        ids = []
        for n, p in g:
            state[idx] = {"exp_avg": own["state"][n]["exp_avg"], "exp_avg_sq": own["state"][n]["exp_avg_sq"]}
            ids.append(idx); idx += 1
        pg.append({"params": ids, "step": own["step"], "lr": 1e-3})
        fp32.append([own["state"][n]["fp32_param"] for n, p in g])
    ref_layout = {"optimizer": {"state": state, "param_groups": pg}, "grad_scaler": None, "fp32_from_fp16_params": fp32}
    torch.manual_seed(1)
    model_b = GPTModel(cfg, device=torch.device("cpu"), layer_offset=0, num_layers_here=2)
    b = MixedPrecisionOptimizer([model_b], lr=1e-3, weight_decay=0.1, clip_grad=1.0)
    b.load_state_dict(ref_layout)
    assert b.step_count == a.step_count
    assert torch.equal(b.exp_avg, a.exp_avg) and torch.equal(b.exp_avg_sq, a.exp_avg_sq) and torch.equal(b.main_param, a.main_param)
    bad = {"optimizer": {"state": state, "param_groups": pg[:1]}, "fp32_from_fp16_params": fp32}
    import pytest
    with pytest.raises(ValueError, match="--no-load-optim"):
        b.load_state_dict(bad)
