"""Oxford-IIIT-Pet image classification under bucketed DDP: same flags, log lines, history file and checkpoint
files as the reference's 2_training_oxford-pet_ddp/pytorch_oxford_ddp.py (args_fn :92-137, data loaders :139-203,
dist_setting / check_sagemaker :206-231, train :234-372, validate :375-439), with its faults fixed:
  * true step time (the reference divides every iteration's time by log_interval -> 5x inflated img/s, Q6);
  * metrics are read on the host only every --log-interval steps (Q7); validation is sharded and all-reduced (Q8);
  * per-rank batch = flag / max(world // 8, 1) (Q1); CPU-safe DDP construction (Q2); real head resize (Q5);
  * bf16 training (BASELINE config 2): bf16 model + fp32 master weights in the fused optimizer, DDP buckets
    all-reduced by the fused NVLS kernel.
"""
from __future__ import annotations

import argparse
import logging
import os
import sys
import time

import torch
import torch.distributed as dist
import torch.nn as nn

from ..data.imagefolder import EvalTransform, ImageFolderDataset, SyntheticImageDataset, TrainAugment
from ..launch import env as launch_env
from ..ops import functional as OF
from ..parallel.ddp import DistributedDataParallel as DDP
from ..utils import train_util as util

logger = logging.getLogger(__name__)
logger.setLevel(logging.DEBUG)
if not logger.handlers:
    logger.addHandler(logging.StreamHandler(sys.stdout))


def _strbool(v):
    return str(v).lower() in ("1", "true", "yes", "y", "t")


def args_fn(argv=None):
    p = argparse.ArgumentParser(description="PyTorch Resnet50 Example")
    p.add_argument("--log-interval", type=int, default=5, metavar="N", help="how many batches to wait before logging training status")
    p.add_argument("--backend", type=str, default="nccl", help="backend for distributed training (nccl | gloo | smddp | native)")
    p.add_argument("--channels-last", type=_strbool, default=True)
    p.add_argument("--seed", type=int, default=1, metavar="S", help="random seed (default: 1)")
    p.add_argument("-p", "--print-freq", default=10, type=int, metavar="N", help="print frequency (default: 10)")
    p.add_argument("--native-bn", action="store_true", default=False,
                   help="models with BatchNorm2d (resnet*): run it on the kernels of csrc/batchnorm.cu instead of cuDNN / ATen")
    p.add_argument("--sync-bn", action="store_true", default=False, help="with --native-bn: batch statistics over all ranks (SyncBatchNorm)")
    p.add_argument("--model_name", type=str, default="resnet50")
    p.add_argument("--height", type=int, default=224)
    p.add_argument("--width", type=int, default=224)
    p.add_argument("--lr", type=float, default=0.0001)
    p.add_argument("--num-classes", type=int, default=10)
    p.add_argument("--num-epochs", type=int, default=3)
    p.add_argument("--batch-size", type=int, default=64)
    p.add_argument("--test-batch-size", type=int, default=200)
    # additions so the script runs off SageMaker / with synthetic data / in bf16
    p.add_argument("--data-dir", type=str, default=None, help="<dir>/{train,val}/<class>/*.jpg (SM_CHANNEL_TRAINING on SageMaker)")
    p.add_argument("--model-dir", type=str, default="./model_out")
    p.add_argument("--synthetic", action="store_true")
    p.add_argument("--synthetic-train-size", type=int, default=5899)
    p.add_argument("--synthetic-val-size", type=int, default=734)
    p.add_argument("--bf16", type=_strbool, nargs="?", const=True, default=False)
    p.add_argument("--no-pretrained", action="store_true")
    p.add_argument("--pretrained-path", type=str, default=None, help="torchvision-format state_dict file to start from (no download here)")
    p.add_argument("--max-steps", type=int, default=None)
    p.add_argument("--num-workers", type=int, default=2)
    p.add_argument("--bucket-cap-mb", type=float, default=25.0)
    p.add_argument("--gpu-augment", action="store_true", help="real images: the loader workers only decode + resize (uint8 over PCIe); the random "
                                                               "augmentations run as batched tensor ops on the training device (data/gpu_augment.py)")
    p.add_argument("--cuda-graph", action="store_true", help="bf16 path: capture zero-grad + forward + backward + gradient all-reduce once, replay per step")
    return p.parse_args(argv)


def check_sagemaker(args):
    d = launch_env.sagemaker_dirs()
    if "model_dir" in d:
        args.data_dir = d.get("channel_training", args.data_dir)
        args.model_dir = d["model_dir"]
    return args


def dist_setting(args):
    info = launch_env.init_distributed(args.backend)
    args.world_size, args.rank, args.local_rank = info.world_size, info.rank, info.local_rank
    args.batch_size //= max(args.world_size // 8, 1)
    args.batch_size = max(args.batch_size, 1)
    return args


def _loaders(args):
    if args.synthetic or not args.data_dir:
        tr = SyntheticImageDataset(args.synthetic_train_size, args.num_classes, args.height, args.width, seed=args.seed)
        va = SyntheticImageDataset(args.synthetic_val_size, args.num_classes, args.height, args.width, seed=args.seed + 1)
    else:
        if getattr(args, "gpu_augment", False):
            from ..data.gpu_augment import DecodeResize
            tr = ImageFolderDataset(os.path.join(args.data_dir, "train"), DecodeResize(args.height, args.width), raw_uint8=True)
        else:
            tr = ImageFolderDataset(os.path.join(args.data_dir, "train"), TrainAugment(args.height, args.width, args.seed + args.rank))
        va = ImageFolderDataset(os.path.join(args.data_dir, "val"), EvalTransform(args.height, args.width))
    pin = torch.cuda.is_available()
    tr_s = torch.utils.data.distributed.DistributedSampler(tr, num_replicas=args.world_size, rank=args.rank) if args.world_size > 1 else None
    va_s = torch.utils.data.distributed.DistributedSampler(va, num_replicas=args.world_size, rank=args.rank, shuffle=False) if args.world_size > 1 else None
    nw = 0 if isinstance(tr, SyntheticImageDataset) else args.num_workers
    tl = torch.utils.data.DataLoader(tr, batch_size=args.batch_size, shuffle=tr_s is None, sampler=tr_s, num_workers=nw, pin_memory=pin, drop_last=True)
    vl = torch.utils.data.DataLoader(va, batch_size=args.test_batch_size, shuffle=False, sampler=va_s, num_workers=nw, pin_memory=pin)
    return tl, tr_s, vl


class _Criterion(nn.Module):
    def forward(self, output, target):
        return OF.cross_entropy(output, target)


_SINK = []


def _sink():
    """JSONL metrics sink (env B200DIST_METRICS_JSONL), created on first use by the logging rank"""
    if not _SINK:
        from ..utils.metrics import JsonlSink
        _SINK.append(JsonlSink())
    return _SINK[0]


def train(args):
    best_acc1 = -1.0
    model_history = util.init_modelhistory({})
    dev = args.device
    dtype = torch.bfloat16 if (args.bf16 and dev.type == "cuda") else torch.float32
    model = util.torch_model(args.model_name, num_classes=args.num_classes, pretrained=not args.no_pretrained and not args.pretrained_path, local_rank=args.local_rank,
                             pretrained_path=args.pretrained_path).to(dev)
    if args.native_bn:
        from ..ops.functional import convert_batchnorm
        model = convert_batchnorm(model, dist.group.WORLD if (args.sync_bn and dist.is_initialized() and args.world_size > 1) else None)
    criterion = _Criterion()
    if dtype != torch.float32:
        from ..optim.mixed import MixedPrecisionOptimizer
        model = model.to(dtype)
    if args.world_size > 1:
        backend = "native" if (args.backend in ("smddp", "native", "nccl") and dev.type == "cuda") else ("nccl" if dev.type == "cuda" else "gloo")
        if dtype != torch.float32:
            ddp = None          # bf16: gradients live in the optimizer's fp32 arena, reduced there in one fused launch
        else:
            model = DDP(model, bucket_cap_mb=args.bucket_cap_mb, comm_backend=backend)
    torch.backends.cudnn.benchmark = False if args.seed else True
    if dtype != torch.float32:
        optimizer = MixedPrecisionOptimizer([model], lr=args.lr, weight_decay=0.0, betas=(0.9, 0.999), eps=1e-8, clip_grad=0.0,
                                            dp_group=dist.group.WORLD if args.world_size > 1 else None,
                                            comm_backend="native" if dev.type == "cuda" else "nccl", no_wd_fn=lambda n, p: True)
        mixed = True
    else:
        from ..optim.fused import Adam
        optimizer, mixed = Adam(model.parameters(), lr=args.lr), False
    train_loader, train_sampler, test_loader = _loaders(args)
    n_local = len(train_sampler) if train_sampler is not None else len(train_loader.dataset)
    logger.info("Processes {}/{} ({:.0f}%) of train data".format(n_local, len(train_loader.dataset), 100.0 * n_local / len(train_loader.dataset)))
    nv = len(test_loader.sampler) if hasattr(test_loader, "sampler") and test_loader.sampler is not None else len(test_loader.dataset)
    logger.info("Processes {}/{} ({:.0f}%) of test data".format(nv, len(test_loader.dataset), 100.0 * nv / len(test_loader.dataset)))
    print(" local_rank : {}, local_batch_size : {}".format(args.local_rank, args.batch_size))

    gpu_aug = None
    if getattr(args, "gpu_augment", False) and not args.synthetic:
        from ..data.gpu_augment import GpuAugment
        gpu_aug = GpuAugment(args.height, args.width, seed=args.seed + args.rank, dtype=dtype)
    graphed = None      # built lazily from the first full batch (static shapes: the loader drops nothing, so the last partial batch runs eagerly)

    def eager_step(data, target):
        optimizer.zero_grad()
        output = model(data)
        loss = criterion(output, target)
        loss.backward()
        if mixed:
            optimizer.reduce_model_grads()
        return loss, output

    for epoch in range(1, args.num_epochs + 1):
        batch_time, losses = util.AverageMeter("Time", ":6.3f"), util.AverageMeter("Loss", ":.4e")
        top1, top5 = util.AverageMeter("Acc@1", ":6.2f"), util.AverageMeter("Acc@5", ":6.2f")
        model.train()
        if train_sampler is not None:
            train_sampler.set_epoch(epoch)
        end = time.time()
        since = 0
        for batch_idx, (data, target) in enumerate(train_loader):
            data = data.to(dev, non_blocking=True)
            data = gpu_aug(data) if gpu_aug is not None else data.to(dtype)
            target = target.to(dev, non_blocking=True)
            use_graph = (getattr(args, "cuda_graph", False) and mixed and dev.type == "cuda" and data.size(0) == args.batch_size
                         and (args.world_size == 1 or getattr(optimizer, "fabric", None) is not None))
            if use_graph:
                if graphed is None:
                    from .graphed import GraphedStep
                    graphed = GraphedStep(lambda d, t: tuple(x.detach() for x in eager_step(d, t)), (data, target))
                loss, output = graphed(data, target)
            else:
                loss, output = eager_step(data, target)
            optimizer.step()
            since += 1
            if args.rank == 0 and (batch_idx % args.log_interval == args.log_interval - 1 or batch_idx == len(train_loader) - 1):
                prec1, prec5 = util.accuracy(output, target, topk=(1, 5))
                losses.update(util.to_python_float(loss.float()), data.size(0))
                top1.update(util.to_python_float(prec1), data.size(0))
                top5.update(util.to_python_float(prec5), data.size(0))
                if dev.type == "cuda":
                    torch.cuda.synchronize()
                batch_time.update((time.time() - end) / since)      # true per-step time over the logging window
                end, since = time.time(), 0
                print("Epoch: [{0}][{1}/{2}] Train_Time={bt.val:.3f}: avg-{bt.avg:.3f}, Train_Speed={3:.3f} ({4:.3f}), "
                      "Train_Loss={loss.val:.10f}:({loss.avg:.4f}), Train_Prec@1={top1.val:.3f}:({top1.avg:.3f}), "
                      "Train_Prec@5={top5.val:.3f}:({top5.avg:.3f})".format(
                          epoch, batch_idx, len(train_loader), args.world_size * args.batch_size / batch_time.val,
                          args.world_size * args.batch_size / batch_time.avg, bt=batch_time, loss=losses, top1=top1, top5=top5), flush=True)
                for k, v in (("epoch", epoch), ("batch_idx", batch_idx), ("batch_time", batch_time.val), ("losses", losses.val), ("top1", top1.val), ("top5", top5.val)):
                    model_history[k].append(v)
                _sink().write(workload="oxford", epoch=epoch, batch_idx=batch_idx, step_ms=batch_time.val * 1000.0,
                              images_per_s=args.world_size * args.batch_size / batch_time.val, loss=losses.val, top1=top1.val, top5=top5.val)
            if args.max_steps is not None and batch_idx + 1 >= args.max_steps:
                break
        acc1 = validate(test_loader, model, criterion, epoch, model_history, args, dtype)
        if args.rank == 0:
            is_best = acc1 > best_acc1
            best_acc1 = max(acc1, best_acc1)
            os.makedirs(args.model_dir, exist_ok=True)
            util.save_history(os.path.join(args.model_dir, "model_history.p"), model_history)
            util.save_model({"epoch": epoch + 1, "model_name": args.model_name, "state_dict": model.state_dict(), "best_acc1": best_acc1,
                             "optimizer": optimizer.state_dict(), "class_to_idx": train_loader.dataset.class_to_idx}, is_best, args)
        if args.world_size > 1:
            dist.barrier()
    return model, model_history


def validate(val_loader, model, criterion, epoch, model_history, args, dtype=torch.float32):
    batch_time, losses = util.AverageMeter("Time", ":6.3f"), util.AverageMeter("Loss", ":.4e")
    top1, top5 = util.AverageMeter("Acc@1", ":6.2f"), util.AverageMeter("Acc@5", ":6.2f")
    model.eval()
    end = time.time()
    dev = args.device
    tot = torch.zeros(4, device=dev)           # sum loss*n, sum top1*n, sum top5*n, n  (sharded validation, all-reduced)
    for batch_idx, (data, target) in enumerate(val_loader):
        data, target = data.to(dev).to(dtype), target.to(dev)
        with torch.no_grad():
            output = model(data)
            loss = criterion(output, target)
        prec1, prec5 = util.accuracy(output, target, topk=(1, 5))
        n = data.size(0)
        tot += torch.stack([loss.float() * n, prec1[0] * n, prec5[0] * n, torch.tensor(float(n), device=dev)])
        losses.update(util.to_python_float(loss.float()), n)
        top1.update(util.to_python_float(prec1), n)
        top5.update(util.to_python_float(prec5), n)
        batch_time.update(time.time() - end)
        end = time.time()
        if args.rank == 0:
            print("Test: [{0}/{1}]  Test_Time={bt.val:.3f}:({bt.avg:.3f}), Test_Speed={2:.3f}:({3:.3f}), Test_Loss={loss.val:.4f}:({loss.avg:.4f}), "
                  "Test_Prec@1={top1.val:.3f}:({top1.avg:.3f}), Test_Prec@5={top5.val:.3f}:({top5.avg:.3f})".format(
                      batch_idx, len(val_loader), args.world_size * n / max(batch_time.val, 1e-9), args.world_size * n / max(batch_time.avg, 1e-9),
                      bt=batch_time, loss=losses, top1=top1, top5=top5), flush=True)
            for k, v in (("val_epoch", epoch), ("val_batch_idx", batch_idx), ("val_batch_time", batch_time.val), ("val_losses", losses.val),
                         ("val_top1", top1.val), ("val_top5", top5.val)):
                model_history[k].append(v)
    if args.world_size > 1:
        dist.all_reduce(tot)
    n = max(tot[3].item(), 1.0)
    avg_loss, avg1, avg5 = tot[0].item() / n, tot[1].item() / n, tot[2].item() / n
    for k, v in (("val_avg_epoch", epoch), ("val_avg_batch_time", batch_time.avg), ("val_avg_losses", avg_loss), ("val_avg_top1", avg1), ("val_avg_top5", avg5)):
        model_history[k].append(v)
    return avg1


def main(argv=None):
    print("start main function")
    args = args_fn(argv)
    args = check_sagemaker(args)
    args = dist_setting(args)
    args.device = launch_env.device_for(launch_env.RankInfo(args.rank, args.local_rank, args.world_size))
    torch.manual_seed(args.seed)
    out = train(args)
    if dist.is_initialized():
        dist.destroy_process_group()
    return out
