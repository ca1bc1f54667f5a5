"""Swin-specific ops on CPU (the oracles the GPU kernels are checked against): fused window attention autograd vs the eager
chain, shift+partition / reverse as inverse permutations, the fused GELU MLP node, and torchvision forward parity."""
import pytest
import torch

from b200dist.ops import functional as OF


def test_window_attention_matches_eager_chain():
    torch.manual_seed(0)
    Bw, T, heads = 8, 49, 3
    qkv = torch.randn(Bw, T, 3 * heads * 32, requires_grad=True)
    bias = torch.randn(heads, T, T, requires_grad=True)
    mask = torch.where(torch.rand(4, T, T) < 0.3, -100.0, 0.0)
    out = OF.window_attention(qkv, bias, mask, heads, 0.2)
    g = torch.randn_like(out)
    out.backward(g)
    g_qkv, g_bias = qkv.grad.clone(), bias.grad.clone()
    qkv.grad = bias.grad = None
    q, k, v = qkv.view(Bw, T, 3, heads, 32).permute(2, 0, 3, 1, 4)
    a = (q * 0.2) @ k.transpose(-2, -1) + bias[None]
    a = (a.view(2, 4, heads, T, T) + mask[None, :, None]).view(Bw, heads, T, T)
    ref = (torch.softmax(a, -1) @ v).transpose(1, 2).reshape(Bw, T, heads * 32)
    ref.backward(g)
    assert torch.allclose(out, ref, atol=1e-5) and torch.allclose(g_qkv, qkv.grad, atol=1e-5) and torch.allclose(g_bias, bias.grad, atol=1e-4)


@pytest.mark.parametrize("shift", [(0, 0), (3, 3), (3, 0)])
def test_window_partition_reverse_are_inverse(shift):
    x = torch.randn(2, 14, 21, 8, requires_grad=True)
    xw = OF.window_partition(x, 7, shift)
    assert xw.shape == (2 * 2 * 3, 49, 8)
    ref = torch.roll(x, shifts=(-shift[0], -shift[1]), dims=(1, 2)).view(2, 2, 7, 3, 7, 8).permute(0, 1, 3, 2, 4, 5).reshape(12, 49, 8)
    assert torch.equal(xw, ref)
    back = OF.window_reverse(xw, 2, 14, 21, 7, shift)
    assert torch.equal(back, x)
    g = torch.randn_like(xw)
    xw.backward(g)                    # gradient of a permutation = the inverse permutation of the upstream gradient
    assert torch.equal(x.grad, OF.window_reverse(g, 2, 14, 21, 7, shift))


@pytest.mark.parametrize("act", ["gelu_erf", "gelu_tanh"])
def test_fused_mlp_node_matches_autograd(act):
    torch.manual_seed(1)
    t = [torch.randn(s, dtype=torch.bfloat16) for s in ((6, 16), (32, 16), (32,), (16, 32), (16,))]
    a = [x.clone().requires_grad_() for x in t]
    y = OF.mlp(*a, act=act)
    g = torch.randn_like(y)
    y.backward(g)
    r = [x.float().requires_grad_() for x in t]
    h = torch.nn.functional.gelu(torch.nn.functional.linear(r[0], r[1], r[2]), approximate="tanh" if act == "gelu_tanh" else "none")
    yr = torch.nn.functional.linear(h, r[3], r[4])
    yr.backward(g.float())
    for mine, ref in zip([y] + [x.grad for x in a], [yr] + [x.grad for x in r]):
        scale = ref.abs().max().item() + 1e-6
        assert (mine.float() - ref).abs().max().item() <= 0.03 * scale


def test_swin_t_forward_matches_torchvision():
    tv = pytest.importorskip("torchvision")
    from b200dist.models import swin
    torch.manual_seed(0)
    m = swin.swin_t(num_classes=10).eval()
    ref = tv.models.swin_t(weights=None, num_classes=10).eval()
    ref.load_state_dict(m.state_dict())
    x = torch.randn(2, 3, 64, 64)
    with torch.no_grad():
        assert torch.allclose(m(x), ref(x), atol=1e-5)


def test_mnist_native_op_layer_matches_aten_on_cpu():
    """the autograd wrappers around csrc/conv.cu (CPU: their restatements) against plain ATen ops, and the MNIST model with
    native_ops on / off (same parameters, eval mode so dropout is inert)"""
    import torch
    import torch.nn.functional as F
    from b200dist.models.mnist import Net
    from b200dist.ops import functional as Fn
    from b200dist.ops import native as N
    torch.manual_seed(0)
    x = torch.randn(3, 2, 9, 9, requires_grad=True)
    w = torch.randn(5, 2, 3, 3, requires_grad=True)
    b = torch.randn(5, requires_grad=True)
    y = Fn.max_pool2x2(Fn.conv2d_bias_relu(x, w, b))
    ref = F.max_pool2d(F.relu(F.conv2d(x, w, b)), 2)
    assert torch.allclose(y, ref, atol=1e-6)
    g = torch.randn_like(y)
    got = torch.autograd.grad(y, (x, w, b), g)
    want = torch.autograd.grad(ref, (x, w, b), g)
    for a_, b_ in zip(got, want):
        assert torch.allclose(a_, b_, atol=1e-5)
    z = torch.randn(7, 10, requires_grad=True)
    lp = Fn.log_softmax_small(z)
    t = torch.randint(0, 10, (7,))
    (gz,) = torch.autograd.grad(F.nll_loss(lp, t), z)
    (gr,) = torch.autograd.grad(F.nll_loss(F.log_softmax(z, 1), t), z)
    assert torch.allclose(gz, gr, atol=1e-6)
    # fused loss form of the kernel pair
    lp2, loss = N.log_softmax_nll_fwd(z.detach(), t)
    assert torch.allclose(loss, F.nll_loss(F.log_softmax(z.detach(), 1), t, reduction="none"), atol=1e-6)
    assert torch.allclose(N.log_softmax_nll_bwd(lp2, None, t, None, 1.0 / 7), gr, atol=1e-6)
    # odd spatial size: the last row / column is dropped by the pool and receives zero gradient
    xo = torch.randn(2, 3, 7, 5, requires_grad=True)
    yo = Fn.max_pool2x2(xo)
    ro = F.max_pool2d(xo, 2)
    assert torch.equal(yo, ro)
    assert torch.equal(torch.autograd.grad(yo.sum(), xo)[0], torch.autograd.grad(ro.sum(), xo)[0])
    m1, m2 = Net().eval(), Net(native_ops=True).eval()
    m2.load_state_dict(m1.state_dict())
    img = torch.randn(4, 1, 28, 28)
    assert torch.allclose(m1(img), m2(img), atol=1e-5)


def test_native_batchnorm_matches_torch_and_converts_resnet():
    import torch
    from b200dist.ops import functional as Fn
    torch.manual_seed(0)
    x = (torch.randn(6, 5, 7, 9) * 2 + 3).requires_grad_()
    ref = torch.nn.BatchNorm2d(5)
    ours = Fn.BatchNorm2d(5)
    with torch.no_grad():
        ref.weight.uniform_(0.5, 1.5); ref.bias.normal_()
    ours.load_state_dict(ref.state_dict())
    for step in range(2):                      # two training steps: running statistics evolve identically
        y, r = ours(x), ref(x)
        assert torch.allclose(y, r, atol=1e-5)
        g = torch.randn_like(r)
        got = torch.autograd.grad(y, (x, ours.weight, ours.bias), g)
        want = torch.autograd.grad(r, (x, ref.weight, ref.bias), g)
        for a, b in zip(got, want):
            assert torch.allclose(a, b, atol=2e-5), (a - b).abs().max()
    assert torch.allclose(ours.running_mean, ref.running_mean, atol=1e-6) and torch.allclose(ours.running_var, ref.running_var, atol=1e-5)
    assert int(ours.num_batches_tracked) == 2
    ours.eval(); ref.eval()
    assert torch.allclose(ours(x), ref(x), atol=1e-5)
    # fused ReLU, forward and backward
    fr = Fn.BatchNorm2d(5, fuse_relu=True)
    fr.load_state_dict(ref.state_dict()); fr.train(); ref.train()
    y, r = fr(x), torch.relu(ref(x))
    assert torch.allclose(y, r, atol=1e-5)
    g = torch.randn_like(r)
    for a, b in zip(torch.autograd.grad(y, (x, fr.weight, fr.bias), g), torch.autograd.grad(r, (x, ref.weight, ref.bias), g)):
        assert torch.allclose(a, b, atol=2e-5)
    # torchvision ResNet with every BatchNorm2d swapped for the native module
    import torchvision
    net = torchvision.models.resnet18(weights=None, num_classes=7)
    twin = Fn.convert_batchnorm(__import__("copy").deepcopy(net))
    assert sum(isinstance(m, Fn.BatchNorm2d) for m in twin.modules()) == 20 and list(twin.state_dict()) == list(net.state_dict())
    img = torch.randn(4, 3, 64, 64)
    a, b = net(img), twin(img)
    assert torch.allclose(a, b, atol=1e-4)
    a.square().mean().backward(); b.square().mean().backward()
    ga, gb = net.layer2[0].conv1.weight.grad, twin.layer2[0].conv1.weight.grad
    assert torch.allclose(ga, gb, atol=1e-5, rtol=1e-3)


def _syncbn_worker(rank, world):
    import torch
    from b200dist.ops import functional as Fn
    torch.manual_seed(0)
    x = torch.randn(8, 3, 5, 5) * 1.5 + 0.5
    g = torch.randn(8, 3, 5, 5)
    bn = Fn.BatchNorm2d(3, process_group=torch.distributed.group.WORLD)
    with torch.no_grad():
        bn.weight.copy_(torch.tensor([0.5, 1.0, 1.5])); bn.bias.copy_(torch.tensor([0.1, -0.2, 0.3]))
    xs = x.chunk(world)[rank].clone().requires_grad_()
    y = bn(xs)
    dx, dw, db = torch.autograd.grad(y, (xs, bn.weight, bn.bias), g.chunk(world)[rank])
    return y.detach(), dx, dw, db, bn.running_mean.clone(), bn.running_var.clone()


def test_sync_batchnorm_over_two_ranks_equals_full_batch():
    import numpy as np
    import torch
    from tests._dist_util import run_distributed
    outs = run_distributed(_syncbn_worker, 2)
    torch.manual_seed(0)
    x = (torch.randn(8, 3, 5, 5) * 1.5 + 0.5).requires_grad_()
    g = torch.randn(8, 3, 5, 5)
    ref = torch.nn.BatchNorm2d(3)
    with torch.no_grad():
        ref.weight.copy_(torch.tensor([0.5, 1.0, 1.5])); ref.bias.copy_(torch.tensor([0.1, -0.2, 0.3]))
    y = ref(x)
    dx, dw, db = torch.autograd.grad(y, (x, ref.weight, ref.bias), g)
    assert np.allclose(np.concatenate([outs[0][0], outs[1][0]]), y.detach().numpy(), atol=1e-5)
    assert np.allclose(np.concatenate([outs[0][1], outs[1][1]]), dx.numpy(), atol=2e-5)
    assert np.allclose(outs[0][2] + outs[1][2], dw.numpy(), atol=2e-5) and np.allclose(outs[0][3] + outs[1][3], db.numpy(), atol=2e-5)
    assert np.allclose(outs[0][4], ref.running_mean.numpy(), atol=1e-6) and np.allclose(outs[1][5], ref.running_var.numpy(), atol=1e-5)
