"""Single-GPU kernel tier: every native kernel vs the plain PyTorch fp32 oracle of the same op,
including non-tile-multiple shapes, all operand majors, every fused epilogue, split-K accumulation, head_dim 64/128,
causal / full attention and sequence lengths that are not multiples of the 128-row tile."""
import importlib.util
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def kc():
    spec = importlib.util.spec_from_file_location("kernel_check", os.path.join(ROOT, "bench", "kernel_check.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("section", ["gemm", "norm", "elementwise", "xent", "optim", "embedding", "attn", "wattn"])
def test_kernel_section(kc, section):
    from b200dist.ops import native as N
    assert N.available(), "libb200dist.so missing on a GPU box"
    kc.FAILS.clear()
    before = N.launches()
    kc.ALL[section](True)
    torch.cuda.synchronize()
    assert N.launches() > before, "no native kernel was launched"
    assert not kc.FAILS, kc.FAILS


def test_gpt_layer_gpu_matches_cpu_oracle():
    """bf16 GPT (2 layers) on the native kernels vs the same model evaluated by the fp32 CPU reference path."""
    from b200dist.config.megatron_args import TransformerConfig
    from b200dist.models.gpt import GPTModel
    kw = dict(num_layers=2, hidden_size=256, num_attention_heads=4, ffn_hidden_size=1024, kv_channels=64, seq_length=256,
              max_position_embeddings=256, padded_vocab_size=512, hidden_dropout=0.0, attention_dropout=0.0)
    gpu = GPTModel(TransformerConfig(params_dtype=torch.bfloat16, bf16=True, **kw), device="cuda", layer_offset=0, num_layers_here=2, seed=3)
    cpu = GPTModel(TransformerConfig(params_dtype=torch.float32, **kw), device="cpu", layer_offset=0, num_layers_here=2, seed=3)
    with torch.no_grad():
        for (n, p), (_, q) in zip(gpu.named_parameters(), cpu.named_parameters()):
            q.copy_(p.float().cpu())
    tok = torch.randint(0, 512, (2, 257))
    lg = gpu(tok[:, :-1].cuda().contiguous(), labels=tok[:, 1:].cuda().contiguous()).mean()
    lc = cpu(tok[:, :-1].contiguous(), labels=tok[:, 1:].contiguous()).mean()
    lg.backward(); lc.backward()
    assert abs(lg.item() - lc.item()) < 3e-2 * abs(lc.item())
    worst = 0.0
    for (n, p), (_, q) in zip(gpu.named_parameters(), cpu.named_parameters()):
        g, r = p.grad.float().cpu(), q.grad
        err = (g - r).abs().max().item() / (r.abs().max().item() + 1e-6)
        worst = max(worst, err)
        assert err < 0.08, (n, err)


def test_smoke_entry():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    ge.smoke()


def test_dropout_masks_change_between_graph_replays_and_agree_within_one():
    """dropout under a CUDA graph: the captured seed is scrambled with a device-side replay counter that the graph's first
    node bumps, so (1) two replays draw different masks, (2) forward and backward of one replay use the same mask, for the
    Philox element-wise dropout and for attention dropout"""
    import torch
    from b200dist.ops import native as N
    dev = torch.device("cuda", 0)
    x = torch.ones(4096, 256, device=dev, dtype=torch.bfloat16)
    S, B, H, D = 256, 2, 4, 64
    torch.manual_seed(0)
    qkv = torch.randn(S, B, H, 3 * D, device=dev, dtype=torch.bfloat16)
    q, k, v = qkv[..., :D], qkv[..., D:2 * D], qkv[..., 2 * D:]
    N.dropout_replay_counter(dev)
    out = {}

    def step():
        N.bump_dropout_replay_counter()
        out["y"] = N.bias_dropout_add(x, None, None, 0.25, 1234)
        out["dx"] = N.dropout_bwd(x, 0.25, 1234)
        out["o"], out["lse"] = N.attn_fwd(q, k, v, 0.125, True, p_drop=0.2, seed=77)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        step()
    seen = []
    for _ in range(3):
        g.replay()
        torch.cuda.synchronize()
        y, dx, o = out["y"].float(), out["dx"].float(), out["o"].float().clone()
        assert torch.equal(y == 0, dx == 0), "forward and backward masks of one replay differ"
        keep = (y != 0).float().mean().item()
        assert abs(keep - 0.75) < 0.01, keep
        seen.append(((y != 0).clone(), o))
    assert not torch.equal(seen[0][0], seen[1][0]) and not torch.equal(seen[1][0], seen[2][0]), "replays repeated the element-wise mask"
    assert not torch.equal(seen[0][1], seen[1][1]), "replays repeated the attention-dropout mask"
