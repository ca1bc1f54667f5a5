"""Training utilities of the Oxford-Pet workload (reference: 2_training_oxford-pet_ddp/util.py): model factory,
top-k accuracy, checkpoint writer, meters, step-decay LR with warm-up, history (17 keys, JSON) and the S3
checkpoint sync helpers (thin wrappers over the `aws` CLI; no-ops off AWS)."""
from __future__ import annotations

import codecs
import json
import os
import shutil
import subprocess
from typing import Optional, Dict, List, Sequence

import torch


def torch_model(model_name: str, num_classes: int = 0, pretrained: bool = True, local_rank: int = 0, model_type: str = "custom", dtype=None,
                pretrained_path: Optional[str] = None):
    """Model factory (util.py:34-66).  `pretrained` weights are used when torchvision has them cached locally;
    this box has no network, so the default is random init.  Unlike the reference (Q5) the classification head is
    really resized to `num_classes`."""
    from ..models import swin
    name = model_name.lower()
    if name in ("swin_b", "swin_t"):
        model = getattr(swin, name)()
        if pretrained:
            try:
                import torchvision
                ref = getattr(torchvision.models, name)(weights="DEFAULT")
                model.load_state_dict(ref.state_dict())
            except Exception as e:      # offline: keep random init, say so once
                if local_rank == 0:
                    print(f"[torch_model] pretrained weights for {name} unavailable ({type(e).__name__}); using random init")
        if num_classes > 0:
            swin.replace_head(model, num_classes)
    else:
        import torchvision
        model = torchvision.models.__dict__[model_name](weights=None)
        if num_classes > 0:
            swin.replace_head(model, num_classes)
    if pretrained_path:
        # a torchvision-format state_dict file (`torch.save(torchvision.models.swin_b(weights=...).state_dict(), path)`); the
        # classifier rows are skipped when their shape differs from the replaced head
        sd = torch.load(pretrained_path, map_location="cpu", weights_only=True)
        sd = sd.get("state_dict", sd)
        own = model.state_dict()
        keep = {k.removeprefix("module."): v for k, v in sd.items()}
        keep = {k: v for k, v in keep.items() if k in own and own[k].shape == v.shape}
        missing = [k for k in own if k not in keep]
        model.load_state_dict(keep, strict=False)
        if local_rank == 0:
            print(f"[torch_model] loaded {len(keep)} tensors from {pretrained_path}; left at their initial values: {missing[:6]}{'...' if len(missing) > 6 else ''}")
    if dtype is not None:
        model = model.to(dtype)
    return model


def accuracy(output: torch.Tensor, target: torch.Tensor, topk: Sequence[int] = (1,)) -> List[torch.Tensor]:
    """precision@k in percent (util.py:69-83).  k is capped at the number of classes (the reference's top-5 raises on a head with
    fewer than five outputs)."""
    with torch.no_grad():
        maxk = min(max(topk), output.shape[1])
        _, pred = output.float().topk(maxk, 1, True, True)
        correct = pred.t().eq(target.view(1, -1).expand_as(pred.t()))
        return [correct[:k].reshape(-1).float().sum(0, keepdim=True).mul_(100.0 / target.size(0)) for k in topk]


def save_model(state: Dict, is_best: bool, args) -> None:
    """checkpoint.pth (+ model_best.pth), legacy pickle serialization like util.py:86-94."""
    os.makedirs(args.model_dir, exist_ok=True)
    fn = os.path.join(args.model_dir, "checkpoint.pth")
    torch.save(state, fn, _use_new_zipfile_serialization=False)
    if is_best:
        shutil.copyfile(fn, os.path.join(args.model_dir, "model_best.pth"))


class AverageMeter:
    def __init__(self, name, fmt=":f"):
        self.name, self.fmt = name, fmt
        self.reset()

    def reset(self):
        self.val = self.avg = self.sum = 0.0
        self.count = 0

    def update(self, val, n=1):
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count

    def __str__(self):
        return ("{name} {val" + self.fmt + "} ({avg" + self.fmt + "})").format(**self.__dict__)


class ProgressMeter:
    def __init__(self, num_batches, meters, prefix=""):
        digits = len(str(num_batches // 1))
        self.batch_fmtstr = "[{:" + str(digits) + "d}/" + "{:" + str(digits) + "d}".format(num_batches) + "]"
        self.meters, self.prefix = meters, prefix

    def display(self, batch):
        print("\t".join([self.prefix + self.batch_fmtstr.format(batch)] + [str(m) for m in self.meters]))


def adjust_learning_rate(optimizer, epoch, step, len_epoch, args):
    """step decay (x0.1 every 30 epochs, extra step at 80) with a 5-epoch linear warm-up (util.py:138-154)."""
    factor = epoch // 30 + (1 if epoch >= 80 else 0)
    lr = args.lr * (0.1 ** factor)
    if epoch < 5:
        lr = lr * float(1 + step + epoch * len_epoch) / (5.0 * len_epoch)
    if getattr(args, "rank", 0) == 0:
        print("epoch = {}, step = {}, lr = {}".format(epoch, step, lr))
    for g in optimizer.param_groups:
        g["lr"] = lr
    return lr


def save_history(path, history):
    out = {k: list(map(float, v)) for k, v in history.items()}
    with codecs.open(path, "w", encoding="utf-8") as f:
        json.dump(out, f, separators=(",", ":"), sort_keys=True, indent=4)


def to_python_float(t):
    if hasattr(t, "item"):
        return t.item()
    if hasattr(t, "index"):
        return t[0]
    return t


HISTORY_KEYS = ("epoch", "batch_idx", "batch_time", "losses", "top1", "top5", "val_epoch", "val_batch_idx", "val_batch_time", "val_losses",
                "val_top1", "val_top5", "val_avg_epoch", "val_avg_batch_time", "val_avg_losses", "val_avg_top1", "val_avg_top5")


def init_modelhistory(model_history: Dict) -> Dict:
    for k in HISTORY_KEYS:
        model_history[k] = []
    return model_history


def aws_s3_sync(source, destination):
    cmd = ["aws", "s3", "sync", "--quiet", source, destination]
    print(f"Syncing files from {source} to {destination}")
    if shutil.which("aws") is None:
        print("aws CLI not present: skipping sync")
        return False
    p = subprocess.run(cmd, capture_output=True)
    return p.returncode == 0


def sync_local_checkpoints_to_s3(local_path="/opt/ml/checkpoints", s3_path=None):
    s3_path = s3_path or os.path.dirname(os.path.dirname(os.getenv("SM_MODULE_DIR", ""))) + "/checkpoints"
    return aws_s3_sync(local_path, s3_path)


def sync_s3_checkpoints_to_local(local_path="/opt/ml/checkpoints", s3_path=None):
    s3_path = s3_path or os.path.dirname(os.path.dirname(os.getenv("SM_MODULE_DIR", ""))) + "/checkpoints"
    return aws_s3_sync(s3_path, local_path)
