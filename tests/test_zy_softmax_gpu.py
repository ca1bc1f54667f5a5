"""Stand-alone scale + mask + softmax kernel (csrc/softmax.cu, the reference's scaled_*_softmax_cuda extensions) against the
fp32 PyTorch restatement, and the unfused CoreAttention path of the modular GPT layer on the GPU.  Collected after the
other kernel tiers (file name) because these kernels were added after the last GPU session of round 1."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape,causal,masked", [((2, 4, 128, 128), True, False), ((3, 2, 96, 200), False, True), ((1, 5, 77, 77), True, False),
                                                  ((2, 2, 64, 4096), False, True), ((1, 1, 16, 16384), False, False), ((2, 3, 40, 72), True, False)])
def test_softmax_kernel_matches_fp32_reference(dtype, shape, causal, masked):
    from b200dist.ops import native as N
    assert N.available(), "libb200dist.so missing on a GPU box"
    torch.manual_seed(0)
    b, h, sq, sk = shape
    x = (torch.randn(shape, device="cuda") * 3).to(dtype)
    mask = (torch.rand(b, 1, sq, sk, device="cuda") > 0.8) if masked else None
    before = N.launches()
    y = N.softmax_fwd(x, mask, 0.37, causal)
    ref = N.softmax_ref(x, mask, 0.37, causal)
    tol = 1e-5 if dtype == torch.float32 else 8e-3
    assert torch.allclose(y.float(), ref, atol=tol, rtol=tol), (y.float() - ref).abs().max()
    assert torch.allclose(y.float().sum(-1), torch.ones(b, h, sq, device="cuda"), atol=2e-2)
    dy = torch.randn(shape, device="cuda").to(dtype)
    dx = N.softmax_bwd(dy, y, 0.37)
    yf, df = y.float(), dy.float()
    dref = 0.37 * yf * (df - (df * yf).sum(-1, keepdim=True))
    assert torch.allclose(dx.float(), dref, atol=tol * 4, rtol=tol * 4), (dx.float() - dref).abs().max()
    assert N.launches() >= before + 2, "the native softmax kernels did not run"
    if masked:        # broadcast mask over the batch
        y1 = N.softmax_fwd(x, mask[:1], 0.37, causal)
        assert torch.allclose(y1.float(), N.softmax_ref(x, mask[:1], 0.37, causal), atol=tol, rtol=tol)


def test_unfused_core_attention_layer_matches_flash_on_gpu():
    from b200dist.config.megatron_args import TransformerConfig
    from b200dist.models.gpt import GPTModel
    kw = dict(num_layers=2, hidden_size=256, num_attention_heads=4, ffn_hidden_size=1024, kv_channels=64, seq_length=256,
              max_position_embeddings=256, padded_vocab_size=512, hidden_dropout=0.0, attention_dropout=0.0, params_dtype=torch.bfloat16, bf16=True,
              position_embedding_type="rope", swiglu=True, num_query_groups=2)
    dev = torch.device("cuda", 0)
    a = GPTModel(TransformerConfig(**kw), device=dev, layer_offset=0, num_layers_here=2)
    b = GPTModel(TransformerConfig(attention_impl="unfused", **kw), device=dev, layer_offset=0, num_layers_here=2)
    b.load_state_dict(a.state_dict())
    tokens = torch.randint(0, 512, (2, 256), device=dev)
    labels = torch.randint(0, 512, (2, 256), device=dev)
    la, lb = a(tokens, labels=labels), b(tokens, labels=labels)
    assert torch.isfinite(la).all() and torch.allclose(la, lb, atol=5e-2, rtol=5e-2), (la - lb).abs().max()
    la.mean().backward(); lb.mean().backward()
    for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
        ga, gb = p.grad.float(), q.grad.float()
        assert torch.isfinite(ga).all() and (ga - gb).norm() <= 0.08 * ga.norm() + 1e-4, n
