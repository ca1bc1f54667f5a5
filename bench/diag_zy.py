"""Per-case error report for the conv / batch-norm kernel tiers against float64 PyTorch references (diagnostic; prints, never asserts)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from b200dist.ops import native as N
from b200dist.ops import functional as Fn

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def err(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max()), float((a - b).norm() / (b.norm() + 1e-30))


def conv():
    for shape in [(64, 1, 28, 28, 32, 3), (64, 32, 26, 26, 64, 3), (3, 5, 11, 9, 13, 5), (2, 3, 8, 8, 4, 1), (5, 7, 10, 12, 9, 2)]:
        for relu in (True, False):
            B, Cin, H, W, Cout, K = shape
            torch.manual_seed(0)
            x = torch.randn(B, Cin, H, W, device="cuda")
            w = torch.randn(Cout, Cin, K, K, device="cuda") / (Cin * K * K) ** 0.5
            b = torch.randn(Cout, device="cuda")
            try:
                y = N.conv2d_fwd(x, w, b, relu)
                xd, wd, bd = x.double().requires_grad_(), w.double().requires_grad_(), b.double().requires_grad_()
                ref = F.conv2d(xd, wd, bd)
                ref = F.relu(ref) if relu else ref
                x32, w32, b32 = x.clone().requires_grad_(), w.clone().requires_grad_(), b.clone().requires_grad_()
                r32 = F.conv2d(x32, w32, b32)
                r32 = F.relu(r32) if relu else r32
                dy = torch.randn_like(y)
                dx, dw, db = N.conv2d_bwd(dy, y, x, w, relu, True, True)
                rx, rw, rb = torch.autograd.grad(ref, (xd, wd, bd), dy.double())
                cx, cw, cb = torch.autograd.grad(r32, (x32, w32, b32), dy)
                print(f"conv {shape} relu={int(relu)}: y {err(y, ref)} (cudnn32 {err(r32, ref)}) dx {err(dx, rx)} (cudnn32 {err(cx, rx)}) "
                      f"dw {err(dw, rw)} (cudnn32 {err(cw, rw)}) db {err(db, rb)} (cudnn32 {err(cb, rb)})", flush=True)
            except Exception as e:
                print(f"conv {shape} relu={int(relu)}: EXC {type(e).__name__}: {e}", flush=True)


def bn():
    for dtype in (torch.float32, torch.bfloat16):
        for shape in ((64, 64, 56, 56), (8, 2048, 7, 7), (3, 5, 7, 9)):
            torch.manual_seed(0)
            x = (torch.randn(*shape, device="cuda") * 2 + 3).to(dtype).requires_grad_()
            ref = torch.nn.BatchNorm2d(shape[1]).cuda()
            with torch.no_grad():
                ref.weight.uniform_(0.5, 1.5); ref.bias.normal_()
            for relu in (False, True):
                try:
                    ours = Fn.BatchNorm2d(shape[1], fuse_relu=relu).cuda()
                    ours.load_state_dict(ref.state_dict())
                    ref_t = torch.nn.BatchNorm2d(shape[1]).cuda().double()
                    ref_t.load_state_dict(ref.state_dict())
                    y = ours(x)
                    xd = x.detach().double().requires_grad_()
                    r = ref_t(xd)
                    r = torch.relu(r) if relu else r
                    g = torch.randn_like(r).to(dtype)
                    got = torch.autograd.grad(y, (x, ours.weight, ours.bias), g)
                    want = torch.autograd.grad(r, (xd, ref_t.weight, ref_t.bias), g.double())
                    print(f"bn {dtype} {shape} relu={int(relu)}: y {err(y, r)} dx {err(got[0], want[0])} dgamma {err(got[1], want[1])} dbeta {err(got[2], want[2])} "
                          f"rmean {err(ours.running_mean, ref_t.running_mean)} rvar {err(ours.running_var, ref_t.running_var)}", flush=True)
                    ours.eval(); ref_t.eval()
                    re = ref_t(x.detach().double())
                    print(f"    eval: {err(ours(x.detach()), torch.relu(re) if relu else re)}", flush=True)
                except Exception as e:
                    import traceback; traceback.print_exc()
                    print(f"bn {dtype} {shape} relu={int(relu)}: EXC {type(e).__name__}: {e}", flush=True)
    import torchvision, copy
    net = torchvision.models.resnet50(weights=None, num_classes=37).cuda()
    twin = Fn.convert_batchnorm(copy.deepcopy(net))
    img = torch.randn(8, 3, 128, 128, device="cuda")
    a, b = net(img), twin(img)
    print("resnet50 train-mode twin:", err(b, a), "out absmax", float(a.abs().max()), flush=True)
    net.eval(); twin.eval()
    a, b = net(img), twin(img)
    print("resnet50 eval-mode twin:", err(b, a), "out absmax", float(a.abs().max()), flush=True)


if __name__ == "__main__":
    conv()
    bn()
    torch.cuda.synchronize()
    print("diag done")
