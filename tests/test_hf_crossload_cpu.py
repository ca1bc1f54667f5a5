"""Cross-load with the REAL Hugging Face implementation: a checkpoint written by `transformers` loads
into our LLaMA / OPT models and produces the same logits, and a checkpoint written by ckpt/hf.py loads into `transformers`
(sharded safetensors + index, and the legacy .bin format)."""
import os

import pytest
import torch

transformers = pytest.importorskip("transformers")


def _hf_tiny(kind):
    if kind == "llama":
        cfg = transformers.LlamaConfig(vocab_size=512, hidden_size=128, intermediate_size=352, num_hidden_layers=2, num_attention_heads=2,
                                       num_key_value_heads=2, max_position_embeddings=512, rms_norm_eps=1e-6, tie_word_embeddings=False,
                                       attention_bias=False, mlp_bias=False)
        return transformers.LlamaForCausalLM(cfg)
    cfg = transformers.OPTConfig(vocab_size=512, hidden_size=128, ffn_dim=512, num_hidden_layers=2, num_attention_heads=2, max_position_embeddings=512,
                                 word_embed_proj_dim=128, do_layer_norm_before=True, dropout=0.0, attention_dropout=0.0, activation_function="relu")
    return transformers.OPTForCausalLM(cfg)


@pytest.mark.parametrize("kind", ["llama", "opt"])
def test_checkpoint_written_by_transformers_loads_here(tmp_path, kind):
    from b200dist.ckpt import hf as hf_ckpt
    torch.manual_seed(0)
    ref = _hf_tiny(kind).eval()
    ref.save_pretrained(str(tmp_path), safe_serialization=True, max_shard_size="300KB")        # forces several shards + index.json
    assert any(f.endswith(".index.json") for f in os.listdir(tmp_path))
    ours = hf_ckpt.load_pretrained(str(tmp_path)).eval()
    ids = torch.randint(2, 512, (2, 24))
    with torch.no_grad():
        want = ref(input_ids=ids).logits
        got = ours(ids)
    assert got.shape == want.shape and torch.allclose(got, want, atol=2e-4, rtol=1e-3), (got - want).abs().max()
    # loss with label shifting and ignore_index, as the Trainer uses it
    labels = ids.clone()
    labels[:, :5] = -100
    with torch.no_grad():
        assert abs(float(ours(ids, labels=labels)) - float(ref(input_ids=ids, labels=labels).loss)) < 2e-4


@pytest.mark.parametrize("kind,safe", [("llama", True), ("llama", False), ("opt", True)])
def test_checkpoint_written_here_loads_in_transformers(tmp_path, kind, safe):
    from b200dist.ckpt import hf as hf_ckpt
    from b200dist.models.llama import build_causal_lm
    torch.manual_seed(1)
    ours = build_causal_lm(f"{kind}-tiny").eval()
    hf_ckpt.save_pretrained(ours, str(tmp_path), max_shard_size=300 * 1024, safe_serialization=safe)
    cls = transformers.LlamaForCausalLM if kind == "llama" else transformers.OPTForCausalLM
    ref = cls.from_pretrained(str(tmp_path), torch_dtype=torch.float32).eval()
    ids = torch.randint(2, 512, (2, 24))
    with torch.no_grad():
        assert torch.allclose(ours(ids), ref(input_ids=ids).logits, atol=2e-4, rtol=1e-3)


@pytest.mark.parametrize("kind", ["opt", "llama"])
def test_zero3_training_loss_curve_matches_eager_transformers(tmp_path, kind):
    """SURVEY 7.2 P10 / P11 exit test: the same weights trained for 5 steps by (a) transformers eager + torch.optim.AdamW and
    (b) our model under the ZeRO-3 engine give the same loss curve"""
    from b200dist.ckpt import hf as hf_ckpt
    from b200dist.engine.hf_trainer import Zero3CausalLM
    torch.manual_seed(0)
    ref = _hf_tiny(kind).train()
    ref.save_pretrained(str(tmp_path), safe_serialization=True)
    ours = hf_ckpt.load_pretrained(str(tmp_path)).train()
    zm = Zero3CausalLM(ours, lr=2e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01, max_grad_norm=1.0, gradient_accumulation_steps=1,
                       persistence_threshold=0, comm_backend="nccl")
    decay = [p for n, p in ref.named_parameters()]
    opt = torch.optim.AdamW(decay, lr=2e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01)
    g = torch.Generator().manual_seed(3)
    a, b = [], []
    for step in range(5):
        ids = torch.randint(2, 512, (2, 24), generator=g)
        labels = ids.clone()
        labels[:, :6] = -100
        opt.zero_grad()
        loss = ref(input_ids=ids, labels=labels).loss
        loss.backward()
        torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
        opt.step()
        a.append(float(loss.detach()))
        l2 = zm(ids, labels=labels)
        zm.engine.backward(l2)
        zm.engine.step()
        b.append(float(l2.detach()))
    assert all(abs(x - y) < 3e-4 * max(1.0, abs(x)) for x, y in zip(a, b)), (a, b)
    assert len(set(round(x, 4) for x in b)) == 5          # the weights moved: five different batches, five different losses


def test_greedy_generation_matches_transformers_generate(tmp_path):
    """utils/generate.py (prefix re-run, no KV cache) produces the tokens of transformers' cached greedy decoding on the same weights"""
    from b200dist.ckpt import hf as hf_ckpt
    from b200dist.utils.generate import generate, sample_next
    torch.manual_seed(0)
    ref = _hf_tiny("llama").eval()
    ref.save_pretrained(str(tmp_path), safe_serialization=True)
    ours = hf_ckpt.load_pretrained(str(tmp_path)).eval()
    prompt = torch.randint(3, 512, (2, 5))
    want = ref.generate(prompt, max_new_tokens=8, do_sample=False, pad_token_id=0)
    got = generate(ours, prompt, max_new_tokens=8)
    assert torch.equal(got, want[:, :got.shape[1]])
    # sampling filters: top-k keeps k candidates, top-p keeps the smallest prefix reaching p
    lg = torch.tensor([[2.0, 1.0, 0.5, -1.0, -3.0]])
    g = torch.Generator().manual_seed(1)
    draws = {int(sample_next(lg, 1.0, top_k=2, generator=g)) for _ in range(50)}
    assert draws <= {0, 1} and len(draws) == 2
    draws = {int(sample_next(lg, 1.0, top_p=0.5, generator=g)) for _ in range(50)}
    assert draws == {0}
