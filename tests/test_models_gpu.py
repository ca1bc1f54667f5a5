"""Model-level GPU checks (1 GPU): the native bf16 path of each model family against its own fp32 CPU oracle path, and the
CUDA-graph replay of a DDP-style training step against the eagerly issued one."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda", 0)


def test_swin_t_bf16_gpu_close_to_fp32_cpu():
    dev = _dev()
    from b200dist.models import swin
    torch.manual_seed(0)
    m = swin.swin_t(num_classes=10).eval()
    x = torch.randn(4, 3, 224, 224)
    with torch.no_grad():
        ref = m(x)
        got = m.to(dev).to(torch.bfloat16)(x.to(dev).to(torch.bfloat16)).float().cpu()
    assert torch.isfinite(got).all()
    assert (got - ref).abs().max().item() < 0.06 * (ref.abs().max().item() + 1.0)


@pytest.mark.parametrize("name", ["llama-tiny", "opt-tiny"])
def test_causal_lm_bf16_gpu_loss_close_to_fp32_cpu(name):
    dev = _dev()
    from b200dist.models.llama import build_causal_lm
    torch.manual_seed(0)
    m = build_causal_lm(name)
    ids = torch.randint(0, 512, (2, 128))
    ref = float(m(ids, labels=ids))
    mg = build_causal_lm(name, device=dev, dtype=torch.bfloat16)
    mg.load_state_dict({k: v.to(dev).to(torch.bfloat16) for k, v in m.state_dict().items()})
    loss = mg(ids.to(dev), labels=ids.to(dev))
    loss.backward()
    assert abs(float(loss) - ref) < 0.05
    assert all(p.grad is None or torch.isfinite(p.grad).all() for p in mg.parameters())


def test_graphed_step_matches_eager_step():
    dev = _dev()
    from b200dist.engine.graphed import GraphedStep
    from b200dist.models import swin
    from b200dist.ops import functional as OF
    from b200dist.optim.mixed import MixedPrecisionOptimizer

    def run(graph):
        torch.manual_seed(0)
        model = swin.SwinTransformer(4, 96, (2, 2, 6, 2), (3, 6, 12, 24), 7, 0.0, 10).to(dev).to(torch.bfloat16).train()
        opt = MixedPrecisionOptimizer([model], lr=1e-3, weight_decay=0.0, betas=(0.9, 0.999), eps=1e-8, clip_grad=0.0,
                                      dp_group=None, comm_backend="native", no_wd_fn=lambda n, p: True)
        g = torch.Generator().manual_seed(1)
        x = torch.randn(8, 3, 64, 64, generator=g).to(dev).to(torch.bfloat16)
        y = torch.randint(0, 10, (8,), generator=g).to(dev)

        def fwd_bwd(xx, yy):
            opt.zero_grad()
            loss = OF.cross_entropy(model(xx), yy)
            loss.backward()
            opt.reduce_model_grads()
            return loss.detach()
        step = GraphedStep(fwd_bwd, (x, y)) if graph else fwd_bwd
        out = []
        for _ in range(5):
            out.append(float(step(x, y)))
            opt.step()
        return out

    eager, graphed = run(False), run(True)
    assert eager[-1] < eager[0]
    assert max(abs(a - b) for a, b in zip(eager, graphed)) < 0.05, (eager, graphed)
