"""MNIST training loops (single process and DDP).  Same flags, log lines and control flow as
1_training_mnist_ddp/pytorch_mnist_ddp.py (train :43-63, test :66-84, dist_setting :87-104,
check_sagemaker :107-112, main :115-277) and its single-process control pytorch_mnist.py, with the
reference's latent faults fixed: per-rank batch uses max(world//8, 1), DDP construction
is CPU-safe, the model file is written by rank 0 into the model dir."""
from __future__ import annotations

import argparse
import os
import time

import torch
import torch.distributed as dist
import torch.nn.functional as F

from ..data.mnist import DistributedSampler, MNISTTensors, TensorBatchLoader
from ..launch import env as launch_env
from ..models.mnist import Net
from ..optim.fused import Adadelta, StepLR
from ..parallel.ddp import DistributedDataParallel as DDP


def build_parser(ddp: bool = True) -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="PyTorch MNIST Example")
    p.add_argument("--batch-size", type=int, default=64, metavar="N", help="input batch size for training (default: 64)")
    p.add_argument("--test-batch-size", type=int, default=1000, metavar="N", help="input batch size for testing (default: 1000)")
    p.add_argument("--epochs", type=int, default=2, metavar="N", help="number of epochs to train (default: 2)")
    p.add_argument("--lr", type=float, default=1.0, metavar="LR", help="learning rate (default: 1.0)")
    p.add_argument("--gamma", type=float, default=0.7, metavar="M", help="Learning rate step gamma (default: 0.7)")
    p.add_argument("--seed", type=int, default=1, metavar="S", help="random seed (default: 1)")
    p.add_argument("--log-interval", type=int, default=10, metavar="N", help="how many batches to wait before logging training status")
    p.add_argument("--save-model", action="store_true", default=False, help="For Saving the current Model")
    p.add_argument("--verbose", action="store_true", default=False, help="For displaying smdistributed.dataparallel-specific logs")
    p.add_argument("--data-path", type=str, default="../data", help="Path for downloading the MNIST dataset")
    if ddp:
        p.add_argument("--backend", type=str, default="nccl", help="backend for distributed operations (nccl | gloo | smddp | native)")
    # additions (not in the reference)
    p.add_argument("--model-dir", type=str, default=None, help="where mnist_cnn.pt is written (default: cwd / SM_MODEL_DIR)")
    p.add_argument("--native-ops", dest="native_ops", action="store_true", default=None,
                   help="run conv+bias+ReLU / max-pool / log-softmax on the fp32 kernels of csrc/conv.cu instead of cuDNN / ATen "
                        "(default on a CUDA device: the kernels match cuDNN to rounding on a B200, tests/test_zy_conv_gpu.py)")
    p.add_argument("--no-native-ops", dest="native_ops", action="store_false", help="keep the convolutions / pooling / log-softmax on cuDNN / ATen")
    p.add_argument("--synthetic", action="store_true", default=False, help="use synthetic MNIST-shaped data")
    p.add_argument("--synthetic-size", type=int, default=None)
    p.add_argument("--max-steps", type=int, default=None, help="stop each epoch after this many steps (smoke tests)")
    p.add_argument("--set-epoch", action="store_true", default=False, help="reshuffle per epoch (reference never calls set_epoch)")
    return p


def check_sagemaker(args):
    dirs = launch_env.sagemaker_dirs()
    if "model_dir" in dirs:
        args.data_path = dirs.get("channel_training", args.data_path)
        args.model_dir = dirs["model_dir"]
        args.save_model = True
    return args


_SINK = []


def _sink():
    """JSONL metrics sink (env B200DIST_METRICS_JSONL), created on first use by the logging rank"""
    if not _SINK:
        from ..utils.metrics import JsonlSink
        _SINK.append(JsonlSink())
    return _SINK[0]


def train(args, model, device, train_loader, optimizer, epoch):
    model.train()
    for batch_idx, (data, target) in enumerate(train_loader):
        data, target = data.to(device), target.to(device)
        optimizer.zero_grad()
        output = model(data)
        loss = F.nll_loss(output, target)
        loss.backward()
        optimizer.step()
        if batch_idx % args.log_interval == 0 and args.rank == 0:
            print("Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}".format(
                epoch, batch_idx * len(data) * args.world_size, len(train_loader.dataset),
                100.0 * batch_idx / len(train_loader), loss.item()), flush=True)
            _sink().write(workload="mnist", epoch=epoch, batch_idx=batch_idx, loss=loss.item())
        if args.verbose:
            print("Batch", batch_idx, "from rank", args.rank)
        if args.max_steps is not None and batch_idx + 1 >= args.max_steps:
            break


def test(model, device, test_loader):
    model.eval()
    test_loss, correct = 0.0, 0
    with torch.no_grad():
        for data, target in test_loader:
            data, target = data.to(device), target.to(device)
            output = model(data)
            test_loss += F.nll_loss(output, target, reduction="sum").item()
            pred = output.argmax(dim=1, keepdim=True)
            correct += pred.eq(target.view_as(pred)).sum().item()
    n = len(test_loader.dataset)
    test_loss /= n
    print("\nTest set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n".format(test_loss, correct, n, 100.0 * correct / n), flush=True)
    return test_loss, correct / n


def main(argv=None, ddp: bool = True):
    args = build_parser(ddp).parse_args(argv)
    args = check_sagemaker(args)
    if ddp:
        info = launch_env.init_distributed(args.backend)
        args.world_size, args.rank, args.local_rank = info.world_size, info.rank, info.local_rank
        # reference rule: per-rank batch = flag / number of 8-GPU nodes (pytorch_mnist_ddp.py:102-103, fixed)
        args.batch_size //= max(args.world_size // 8, 1)
        args.batch_size = max(args.batch_size, 1)
    else:
        args.world_size, args.rank, args.local_rank = 1, 0, 0
    torch.manual_seed(args.seed)
    device = launch_env.device_for(launch_env.RankInfo(args.rank, args.local_rank, args.world_size))
    if args.verbose:
        print("Hello from rank", args.rank, "of world_size", args.world_size)
    train_ds = MNISTTensors(args.data_path, train=True, synthetic=True if args.synthetic else None, synthetic_size=args.synthetic_size)
    sampler = DistributedSampler(len(train_ds), args.world_size, args.rank) if ddp else None
    train_loader = TensorBatchLoader(train_ds, args.batch_size, sampler=sampler, shuffle=not ddp, pin=device.type == "cuda")
    test_loader = None
    if args.rank == 0:
        test_ds = MNISTTensors(args.data_path, train=False, synthetic=True if args.synthetic else None,
                               synthetic_size=None if args.synthetic_size is None else max(args.synthetic_size // 6, 1))
        test_loader = TensorBatchLoader(test_ds, args.test_batch_size)
    native_ops = (device.type == "cuda") if args.native_ops is None else bool(args.native_ops)
    model = Net(native_ops=native_ops).to(device)
    if ddp and args.world_size > 1:
        model = DDP(model, comm_backend="auto" if args.backend in ("smddp", "native") else ("nccl" if device.type == "cuda" else "gloo"))
    optimizer = Adadelta(model.parameters(), lr=args.lr)
    scheduler = StepLR(optimizer, step_size=1, gamma=args.gamma)
    t0 = time.time()
    for epoch in range(1, args.epochs + 1):
        if sampler is not None and args.set_epoch:
            sampler.set_epoch(epoch)
        train(args, model, device, train_loader, optimizer, epoch)
        if args.rank == 0:
            test(model, device, test_loader)
        scheduler.step()
    if args.rank == 0:
        print(f"Total training time: {time.time() - t0:.1f}s")
    if args.save_model and args.rank == 0:
        out_dir = args.model_dir or "."
        os.makedirs(out_dir, exist_ok=True)
        torch.save(model.state_dict(), os.path.join(out_dir, "mnist_cnn.pt"))
    if ddp and dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return model
