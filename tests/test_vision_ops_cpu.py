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


def test_rows_scale_add_and_residual_sd_add_match_composition():
    """ops.native.rows_scale_add (CPU restatement of csrc/window_attn.cu: rows_scale_add_kernel) and the autograd function around it
    against roll / partition / broadcast-multiply / add composed from plain torch ops, forward and backward."""
    from b200dist.ops import functional as OF
    from b200dist.ops import native as N
    torch.manual_seed(0)
    B, H, W, C, ws, shift = 3, 14, 14, 16, 7, (3, 3)
    res = torch.randn(B, H, W, C)
    xw = torch.randn(B * 4, ws * ws, C)
    scale = torch.tensor([0.0, 2.0, 1.0])
    # mode 2: window reverse folded in
    want = res + N.window_shuffle(xw, B, H, W, ws, shift, True) * scale.view(B, 1, 1, 1)
    assert torch.allclose(N.rows_scale_add(xw, res, scale, B, H, W, ws, shift, 2), want, atol=1e-6)
    # mode 1 is its adjoint: <rows_scale_add(x, mode 2), g> == <x, rows_scale_add(g, mode 1)>
    g = torch.randn(B, H, W, C)
    lhs = (N.rows_scale_add(xw, None, scale, B, H, W, ws, shift, 2) * g).sum()
    rhs = (xw * N.rows_scale_add(g, None, scale, B, H, W, ws, shift, 1)).sum()
    assert torch.allclose(lhs, rhs, rtol=1e-5)
    # mode 0 on [B, rows, C]
    x3, r3 = torch.randn(B, 10, C), torch.randn(B, 10, C)
    assert torch.allclose(N.rows_scale_add(x3, r3, scale, B, 10, 1, 1, (0, 0), 0), r3 + x3 * scale.view(B, 1, 1), atol=1e-6)
    assert torch.allclose(N.rows_scale_add(x3, r3, None, B, 10, 1, 1, (0, 0), 0), r3 + x3, atol=1e-6)
    # autograd function in bf16 (the dtype that takes the fused path) against the same math composed in fp32
    for geom in (None, (B, H, W, ws, shift)):
        r16 = res.bfloat16().requires_grad_()
        b16 = (xw if geom else torch.randn(B, H, W, C)).bfloat16().requires_grad_()
        torch.manual_seed(5)
        out = OF.residual_sd_add(r16, b16, 0.4, True, window_geom=geom)
        torch.manual_seed(5)
        sc = torch.empty(B).bernoulli_(0.6).div_(0.6)
        rf, bf = r16.detach().float().requires_grad_(), b16.detach().float().requires_grad_()
        ref = rf + (N.window_shuffle(bf, B, H, W, ws, shift, True) if geom else bf) * sc.view(B, 1, 1, 1)
        assert torch.allclose(out.float(), ref, atol=3e-2, rtol=2e-2)
        go = torch.randn_like(ref)
        out.backward(go.bfloat16())
        ref.backward(go.bfloat16().float())
        assert torch.allclose(r16.grad.float(), rf.grad, atol=1e-6)
        assert torch.allclose(b16.grad.float(), bf.grad, atol=3e-2, rtol=2e-2)
    # eval / p = 0: a plain add
    assert torch.equal(OF.residual_sd_add(res, g, 0.3, False), res + g)


def test_layer_norm_residual_sums_both_gradients():
    from b200dist.ops import functional as OF
    torch.manual_seed(1)
    x = torch.randn(4, 9, 32).bfloat16().requires_grad_()
    w, b = torch.randn(32).bfloat16().requires_grad_(), torch.randn(32).bfloat16().requires_grad_()
    y, skip = OF.layer_norm_residual(x, w, b)
    g1, g2 = torch.randn_like(y), torch.randn_like(y)
    (y * g1 + skip * g2).sum().backward()
    xf, wf, bf = (t.detach().float().requires_grad_() for t in (x, w, b))
    yr = torch.nn.functional.layer_norm(xf, (32,), wf, bf)
    (yr * g1.float() + xf * g2.float()).sum().backward()
    assert torch.allclose(y.float(), yr, atol=3e-2, rtol=2e-2)
    assert torch.allclose(x.grad.float(), xf.grad, atol=6e-2, rtol=3e-2)
    assert torch.allclose(w.grad.float(), wf.grad, atol=0.25, rtol=3e-2) and torch.allclose(b.grad.float(), bf.grad, atol=0.1, rtol=3e-2)
    # only the skip output used: the gradient passes straight through
    x2 = torch.randn(2, 3, 16).bfloat16().requires_grad_()
    _, s2 = OF.layer_norm_residual(x2, torch.ones(16).bfloat16(), torch.zeros(16).bfloat16())
    s2.sum().backward()
    assert torch.equal(x2.grad, torch.ones_like(x2))


def test_swin_bias_gather_and_cached_mask_match_torchvision_expressions():
    from b200dist.models import swin
    torch.manual_seed(2)
    ws, heads = 7, 3
    table = torch.randn((2 * ws - 1) ** 2, heads, requires_grad=True)
    index = swin._relative_position_index(ws)
    got = swin._BiasGather.apply(table, index, ws * ws)
    t2 = table.detach().clone().requires_grad_()
    ref = t2[index].view(ws * ws, ws * ws, -1).permute(2, 0, 1)
    assert torch.equal(got, ref)
    g = torch.randn_like(ref)
    got.backward(g)
    ref.backward(g)
    assert torch.allclose(table.grad, t2.grad, atol=1e-5)
    # the cached mask against torchvision's construction
    pH, pW, shift = 14, 21, (3, 3)
    m = swin.shift_window_mask(pH, pW, ws, shift, torch.device("cpu"))
    assert swin.shift_window_mask(pH, pW, ws, shift, torch.device("cpu")) is m
    am = torch.zeros((pH, pW))
    c = 0
    for h in ((0, -ws), (-ws, -shift[0]), (-shift[0], None)):
        for w in ((0, -ws), (-ws, -shift[1]), (-shift[1], None)):
            am[h[0]:h[1], w[0]:w[1]] = c
            c += 1
    am = am.view(pH // ws, ws, pW // ws, ws).permute(0, 2, 1, 3).reshape(-1, ws * ws)
    am = am.unsqueeze(1) - am.unsqueeze(2)
    am = am.masked_fill(am != 0, -100.0).masked_fill(am == 0, 0.0)
    assert torch.equal(m, am)


def test_swin_block_training_step_matches_torchvision_gradients():
    """train mode, stochastic depth off (its draws are not comparable): the restructured block (skip connection through the norm
    function, residual add fused with window_reverse) gives torchvision's loss and gradients in fp32"""
    tv = pytest.importorskip("torchvision")
    from b200dist.models import swin
    torch.manual_seed(3)
    m = swin.SwinTransformer(4, 32, (2, 2), (1, 2), 7, 0.0, 5).train()
    ref = tv.models.swin_transformer.SwinTransformer([4, 4], 32, [2, 2], [1, 2], [7, 7], stochastic_depth_prob=0.0, num_classes=5).train()
    ref.load_state_dict(m.state_dict())
    x = torch.randn(2, 3, 56, 56)
    y = torch.tensor([1, 3])
    l1 = torch.nn.functional.cross_entropy(m(x), y)
    l2 = torch.nn.functional.cross_entropy(ref(x), y)
    assert torch.allclose(l1, l2, atol=1e-5)
    l1.backward(); l2.backward()
    gr = dict(ref.named_parameters())
    for n, p in m.named_parameters():
        assert torch.allclose(p.grad, gr[n].grad, atol=2e-4, rtol=1e-3), n


def test_layer_norm_partition_matches_norm_then_partition():
    """the norm that writes its rows in shifted-window order (row map folded into the kernel pair; CPU: the restatement) against
    layer_norm followed by window_partition, forward and backward including the skip output"""
    from b200dist.ops import functional as OF
    torch.manual_seed(6)
    B, H, W, C, ws = 2, 14, 21, 16, 7
    for shift in ((0, 0), (3, 3), (3, 0)):
        rm = OF.window_row_map(B, H, W, ws, shift, torch.device("cpu"))
        assert rm.dtype == torch.int32 and sorted(rm.tolist()) == list(range(B * H * W))
        assert OF.window_row_map(B, H, W, ws, shift, torch.device("cpu")) is rm
        x = torch.randn(B, H, W, C).bfloat16().requires_grad_()
        w, b = (torch.rand(C) + 0.5).bfloat16().requires_grad_(), torch.randn(C).bfloat16().requires_grad_()
        yw, skip = OF.layer_norm_partition(x, w, b, 1e-5, ws, shift)
        xf, wf, bf = (t.detach().float().requires_grad_() for t in (x, w, b))
        ref = OF.window_partition(torch.nn.functional.layer_norm(xf, (C,), wf, bf), ws, shift)
        assert yw.shape == ref.shape and torch.allclose(yw.float(), ref, atol=3e-2, rtol=2e-2)
        g1, g2 = torch.randn_like(ref), torch.randn(B, H, W, C)
        (yw.float() * g1).sum().backward(retain_graph=True)
        (skip.float() * g2).sum().backward()
        ((ref * g1).sum() + (xf * g2).sum()).backward()
        assert torch.allclose(x.grad.float(), xf.grad, atol=8e-2, rtol=3e-2)
        assert torch.allclose(w.grad.float(), wf.grad, atol=0.3, rtol=3e-2) and torch.allclose(b.grad.float(), bf.grad, atol=0.15, rtol=3e-2)
    # fp32 inputs compose the two ops
    x32 = torch.randn(B, H, W, C)
    y32, s32 = OF.layer_norm_partition(x32, torch.ones(C), torch.zeros(C), 1e-5, ws, (3, 3))
    assert s32 is x32 and torch.allclose(y32, OF.window_partition(torch.nn.functional.layer_norm(x32, (C,)), ws, (3, 3)))


def test_patchify_matches_conv_as_gemm():
    from b200dist.ops import native as N
    torch.manual_seed(7)
    x = torch.randn(2, 3, 8, 12)
    conv = torch.nn.Conv2d(3, 5, 4, 4)
    cols = N.patchify(x, 4)
    y = (cols @ conv.weight.view(5, -1).t() + conv.bias).view(2, 2, 3, 5)
    assert torch.allclose(y, conv(x).permute(0, 2, 3, 1), atol=1e-5)


def test_gpu_augment_pipeline_semantics():
    """data/gpu_augment.py on the CPU: the degenerate configuration is resize + normalise, the flip is a flip, HSV agrees with OpenCV,
    a seed reproduces a batch, and the full pipeline returns normalised images of the requested size and dtype"""
    cv2 = pytest.importorskip("cv2")
    import numpy as np
    from b200dist.data.gpu_augment import GpuAugment, hsv_to_rgb, rgb_to_hsv
    torch.manual_seed(0)
    u8 = (torch.rand(5, 40, 50, 3) * 255).byte()
    mean, std = torch.tensor([0.485, 0.456, 0.406])[None, :, None, None], torch.tensor([0.229, 0.224, 0.225])[None, :, None, None]
    ref = (u8.permute(0, 3, 1, 2).float() / 255 - mean) / std
    off = dict(p_noise=0, p_blur=0, p_tone=0, p_hsv=0, scale=(1.0, 1.0), ratio=(1.0, 1.0))
    assert torch.allclose(GpuAugment(40, 50, seed=1, p_flip=0, **off)(u8), ref, atol=1e-4)
    assert torch.allclose(GpuAugment(40, 50, seed=1, p_flip=1.0, **off)(u8), ref.flip(2), atol=1e-4)
    x = torch.rand(3, 3, 9, 9)
    assert torch.allclose(hsv_to_rgb(rgb_to_hsv(x)), x, atol=1e-5)
    img = (torch.rand(8, 8, 3) * 255).byte().numpy()
    cv = cv2.cvtColor(img, cv2.COLOR_RGB2HSV_FULL).astype(np.float32) / 255.0
    mine = rgb_to_hsv(torch.from_numpy(img).permute(2, 0, 1)[None].float() / 255)[0].permute(1, 2, 0).numpy()
    assert np.abs(((mine - cv + 0.5) % 1.0 - 0.5)[..., 0]).max() < 0.01 and np.abs(mine - cv)[..., 1:].max() < 0.01      # 8-bit quantisation of cv2's result
    a, b = GpuAugment(32, 24, seed=7)(u8), GpuAugment(32, 24, seed=7)(u8)
    assert torch.equal(a, b) and not torch.equal(a, GpuAugment(32, 24, seed=8)(u8))
    y = GpuAugment(32, 24, seed=3, dtype=torch.bfloat16)(u8)
    assert y.shape == (5, 3, 32, 24) and y.dtype == torch.bfloat16 and torch.isfinite(y.float()).all()
    lo, hi = (0 - 0.485) / 0.229 - 1e-2, (1 - 0.406) / 0.225 + 1e-2
    assert y.float().min() >= lo and y.float().max() <= hi
    # crops stay inside the image: a constant image stays constant through crop + flip
    const = torch.full((4, 30, 30, 3), 128, dtype=torch.uint8)
    z = GpuAugment(16, 16, seed=2, p_noise=0, p_blur=0, p_tone=0, p_hsv=0)(const)
    assert torch.allclose(z, ((torch.full((4, 3, 16, 16), 128 / 255.0) - mean) / std), atol=1e-4)
