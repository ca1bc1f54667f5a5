"""Bodies of the GPU checks for csrc/conv.cu, csrc/batchnorm.cu, csrc/softmax.cu and the modular GPT layer on the GPU.
tests/test_zy_*_gpu.py runs each of them in a CHILD process: a fault in one kernel tier must not poison the CUDA context of the
pytest process (and with it every test collected later).

    python -m tests._zy_checks <name>
"""
import sys

import torch
import torch.nn.functional as F



def conv_kernels():
    for shape in [(64, 1, 28, 28, 32, 3), (64, 32, 26, 26, 64, 3), (3, 5, 11, 9, 13, 5), (2, 3, 8, 8, 4, 1), (5, 7, 10, 12, 9, 2)]:
        for relu in (True, False):
            _conv_one(shape, relu)


def _conv_one(shape, relu):
    from b200dist.ops import native as N
    assert N.available()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    B, Cin, H, W, Cout, K = shape
    torch.manual_seed(0)
    x = torch.randn(B, Cin, H, W, device="cuda", requires_grad=True)
    w = (torch.randn(Cout, Cin, K, K, device="cuda") / (Cin * K * K) ** 0.5).requires_grad_()
    b = torch.randn(Cout, device="cuda", requires_grad=True)
    before = N.launches()
    y = N.conv2d_fwd(x.detach(), w.detach(), b.detach(), relu)
    ref = F.conv2d(x, w, b)
    ref = F.relu(ref) if relu else ref
    assert torch.allclose(y, ref, atol=2e-5, rtol=1e-4), (y - ref).abs().max()
    dy = torch.randn_like(ref)
    dx, dw, db = N.conv2d_bwd(dy, y, x.detach(), w.detach(), relu, True, True)
    rx, rw, rb = torch.autograd.grad(ref, (x, w, b), dy)
    assert torch.allclose(dx, rx, atol=1e-4, rtol=1e-3), (dx - rx).abs().max()
    assert torch.allclose(dw, rw, atol=2e-3, rtol=2e-3), (dw - rw).abs().max()
    assert torch.allclose(db, rb, atol=2e-3, rtol=2e-3), (db - rb).abs().max()
    assert N.launches() >= before + 2, "the native conv kernels did not run"     # one C call forward, one backward (dgrad + wgrad)


def pool_and_log_softmax():
    from b200dist.ops import native as N
    torch.manual_seed(1)
    for shape in ((64, 64, 24, 24), (3, 5, 7, 9)):
        x = torch.randn(*shape, device="cuda", requires_grad=True)
        y, idx = N.maxpool2_fwd(x.detach())
        ref = F.max_pool2d(x, 2)
        assert torch.equal(y, ref)
        g = torch.randn_like(ref)
        assert torch.equal(N.maxpool2_bwd(g, idx, tuple(x.shape)), torch.autograd.grad(ref, x, g)[0])
    for rows, C in ((1000, 10), (7, 32), (33, 1)):
        z = torch.randn(rows, C, device="cuda", requires_grad=True)
        t = torch.randint(0, C, (rows,), device="cuda")
        lp, loss = N.log_softmax_nll_fwd(z.detach(), t)
        rlp = F.log_softmax(z, 1)
        assert torch.allclose(lp, rlp, atol=1e-5) and torch.allclose(loss, F.nll_loss(rlp, t, reduction="none"), atol=1e-5)
        d = torch.randn_like(lp)
        gl = torch.rand(rows, device="cuda")
        got = N.log_softmax_nll_bwd(lp, d, t, gl, 0.5)
        (want,) = torch.autograd.grad((rlp * d).sum() + (F.nll_loss(rlp, t, reduction="none") * gl * 0.5).sum(), z)
        assert torch.allclose(got, want, atol=1e-5), (got - want).abs().max()


def mnist_step():
    from b200dist.models.mnist import Net
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)
    a, b = Net().cuda().eval(), Net(native_ops=True).cuda().eval()       # eval: dropout inert, gradients still flow
    b.load_state_dict(a.state_dict())
    x = torch.randn(64, 1, 28, 28, device="cuda")
    t = torch.randint(0, 10, (64,), device="cuda")
    la, lb = F.nll_loss(a(x), t), F.nll_loss(b(x), t)
    assert torch.allclose(la, lb, atol=1e-5)
    la.backward(); lb.backward()
    for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.allclose(p.grad, q.grad, atol=2e-4, rtol=2e-3), (n, (p.grad - q.grad).abs().max())


def softmax_kernels():
    for dtype in (torch.bfloat16, torch.float16, torch.float32):
        for shape, causal, masked in [((2, 4, 128, 128), True, False), ((3, 2, 96, 200), False, True), ((1, 5, 77, 77), True, False),
                                      ((2, 2, 64, 4096), False, True), ((1, 1, 16, 16384), False, False), ((2, 3, 40, 72), True, False)]:
            _softmax_one(dtype, shape, causal, masked)


def _softmax_one(dtype, shape, causal, masked):
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


def unfused_attention_layer():
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


def batchnorm_kernels():
    from b200dist.ops import functional as Fn
    from b200dist.ops import native as N
    torch.manual_seed(0)
    for dtype, tol in ((torch.float32, 2e-5), (torch.bfloat16, 3e-2)):
        for shape in ((64, 64, 56, 56), (8, 2048, 7, 7), (3, 5, 7, 9)):
            x = (torch.randn(*shape, device="cuda") * 2 + 3).to(dtype).requires_grad_()
            ref = torch.nn.BatchNorm2d(shape[1]).cuda()
            with torch.no_grad():
                ref.weight.uniform_(0.5, 1.5); ref.bias.normal_()
            for relu in (False, True):
                ours = Fn.BatchNorm2d(shape[1], fuse_relu=relu).cuda()
                ours.load_state_dict(ref.state_dict())
                ref_t = torch.nn.BatchNorm2d(shape[1]).cuda()
                ref_t.load_state_dict(ref.state_dict())
                before = N.launches()
                y = ours(x)
                pre = ref_t(x.float())
                # fused ReLU: an activation that rounds to +-1e-7 may land on either side of zero in two fp32 implementations; the
                # oracle takes OUR mask where the pre-activation is within rounding of zero, so one flipped element (whose
                # gradient is O(1)) cannot fail the comparison, and the mask itself is checked everywhere else
                if relu:
                    mask = (y.float() > 0)
                    sure = pre.detach().abs() > 1e-4 * max(1.0, float(pre.detach().abs().max()))
                    assert torch.equal(mask[sure], (pre.detach() > 0)[sure]), (dtype, shape, "relu mask")
                    r = pre * mask
                else:
                    r = pre
                assert N.launches() >= before + 3, "native batch-norm kernels did not run"
                assert torch.allclose(y.float(), r, atol=tol, rtol=tol), (dtype, shape, relu, (y.float() - r).abs().max())
                g = torch.randn_like(r)
                got = torch.autograd.grad(y, (x, ours.weight, ours.bias), g.to(dtype))
                want = torch.autograd.grad(r, (x, ref_t.weight, ref_t.bias), g.to(dtype).float())
                scale = max(1.0, float(want[1].abs().max()), float(want[2].abs().max()))
                assert torch.allclose(got[0].float(), want[0].float(), atol=tol * 2, rtol=tol * 2), (dtype, shape, relu, "dx", (got[0].float() - want[0].float()).abs().max())
                assert torch.allclose(got[1].float(), want[1], atol=tol * 4 * scale, rtol=tol * 4), (dtype, shape, relu, "dgamma")
                assert torch.allclose(got[2].float(), want[2], atol=tol * 4 * scale, rtol=tol * 4), (dtype, shape, relu, "dbeta")
                assert torch.allclose(ours.running_mean, ref_t.running_mean, atol=tol) and torch.allclose(ours.running_var, ref_t.running_var, atol=tol * 4, rtol=tol * 4)
                ours.eval(); ref_t.eval()
                re = ref_t(x.detach().float())
                assert torch.allclose(ours(x.detach()).float(), torch.relu(re) if relu else re, atol=tol, rtol=tol)
    import torchvision
    net = torchvision.models.resnet50(weights=None, num_classes=37).cuda()
    twin = Fn.convert_batchnorm(__import__("copy").deepcopy(net))
    img = torch.randn(8, 3, 128, 128, device="cuda")
    # 53 batch-norm layers deep, training mode: the convolutions in between must not amplify last-bit differences of the
    # normalised activations (TF32 rounds them at 2^-11), so both networks run true fp32 convolutions here
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    a, b = net(img), twin(img)
    assert (a - b).abs().max() <= 2e-3 * max(1.0, float(a.abs().max())), (a - b).abs().max()
    net.eval(); twin.eval()
    a, b = net(img), twin(img)
    assert (a - b).abs().max() <= 2e-3 * max(1.0, float(a.abs().max())), (a - b).abs().max()


CHECKS = {"conv_kernels": conv_kernels, "pool_and_log_softmax": pool_and_log_softmax, "mnist_step": mnist_step,
          "batchnorm_kernels": batchnorm_kernels, "softmax_kernels": softmax_kernels, "unfused_attention_layer": unfused_attention_layer}

if __name__ == "__main__":
    CHECKS[sys.argv[1]]()
    torch.cuda.synchronize()
    print("ok", sys.argv[1])
