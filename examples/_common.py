import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def parse(description):
    ap = argparse.ArgumentParser(description=description)
    ap.add_argument("--full", action="store_true", help="the notebook's hyper-parameters instead of the one-minute smoke settings")
    ap.add_argument("--data", default=None, help="dataset directory (the notebook's S3 channel)")
    ap.add_argument("--processes", type=int, default=None, help="processes (GPUs) to use; default: all GPUs, or 2 on a CPU-only box")
    ap.add_argument("--output", default=os.path.join(ROOT, "local_jobs"))
    return ap.parse_args()


def nproc(a):
    if a.processes:
        return a.processes
    import torch
    return torch.cuda.device_count() or 2


def on_gpu():
    import torch
    return torch.cuda.is_available()
