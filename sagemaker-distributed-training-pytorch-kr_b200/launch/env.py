"""Launcher contract: rank identity from torchrun, OpenMPI or SageMaker env, plus process-group init.

The reference scripts read ``OMPI_COMM_WORLD_{SIZE,RANK,LOCAL_RANK}`` published by SageMaker's mpirun
launcher (pytorch_mnist_ddp.py:94-96, pretrain_gpt.py:10-14, train.py:21-25) and SageMaker channel
variables (pytorch_mnist_ddp.py:107-112).  Here either env family works - ``torchrun`` on one 8xB200
box is the native launcher - and nothing requires ``SM_*``.
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class RankInfo:
    rank: int = 0
    local_rank: int = 0
    world_size: int = 1
    node_rank: int = 0
    launcher: str = "single"


def detect_rank_env(env=None) -> RankInfo:
    env = os.environ if env is None else env
    if "RANK" in env and "WORLD_SIZE" in env:
        r, w = int(env["RANK"]), int(env["WORLD_SIZE"])
        lr = int(env.get("LOCAL_RANK", r % max(1, int(env.get("LOCAL_WORLD_SIZE", 8)))))
        return RankInfo(r, lr, w, int(env.get("GROUP_RANK", env.get("NODE_RANK", r // 8))), "torchrun")
    if "OMPI_COMM_WORLD_RANK" in env:
        r, w = int(env["OMPI_COMM_WORLD_RANK"]), int(env["OMPI_COMM_WORLD_SIZE"])
        lr = int(env.get("OMPI_COMM_WORLD_LOCAL_RANK", r % 8))
        return RankInfo(r, lr, w, r // 8, "mpi")
    return RankInfo()


def export_torch_env(info: Optional[RankInfo] = None) -> RankInfo:
    """The OMPI->torch shim of pretrain_gpt.py:10-14 / train.py:21-25 (RANK, LOCAL_RANK, WORLD_SIZE,
    NODE_RANK, CUDA_DEVICE_MAX_CONNECTIONS=1), without requiring mpirun."""
    info = info or detect_rank_env()
    os.environ.setdefault("RANK", str(info.rank))
    os.environ.setdefault("LOCAL_RANK", str(info.local_rank))
    os.environ.setdefault("WORLD_SIZE", str(info.world_size))
    os.environ.setdefault("NODE_RANK", str(info.node_rank))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "1")
    install_hang_dump()
    return info


def install_hang_dump() -> None:
    """B2_HANG_DUMP_S=<seconds>: every rank prints the Python stack of all its threads to stderr every <seconds> (and on
    SIGUSR1), so a stalled multi-rank run shows WHERE each rank is waiting before an outer `timeout` kills it."""
    s = os.environ.get("B2_HANG_DUMP_S")
    if not s:
        return
    import faulthandler
    import signal
    import sys
    try:
        faulthandler.enable(file=sys.stderr)
        faulthandler.dump_traceback_later(float(s), repeat=True, file=sys.stderr)
        faulthandler.register(signal.SIGUSR1, file=sys.stderr, all_threads=True)
    except (ValueError, OSError, AttributeError):
        pass


def sagemaker_dirs(env=None) -> dict:
    env = os.environ if env is None else env
    out = {}
    if env.get("SM_MODEL_DIR") is not None:
        out["model_dir"] = env["SM_MODEL_DIR"]
    for k, v in env.items():
        if k.startswith("SM_CHANNEL_"):
            out["channel_" + k[len("SM_CHANNEL_"):].lower()] = v
    return out


def resolve_backend(requested: str) -> str:
    """``smddp`` (closed-source AWS backend) maps onto the native NVLink fabric backend = nccl bootstrap
    + in-kernel NVLS collectives; without CUDA everything falls back to gloo (pytorch_mnist_ddp.py:92)."""
    if not torch.cuda.is_available():
        return "gloo"
    return "nccl" if requested in ("nccl", "smddp", "native") else requested


def init_distributed(backend: str = "nccl", timeout_minutes: int = 10) -> RankInfo:
    info = export_torch_env()
    be = resolve_backend(backend)
    if torch.cuda.is_available():
        torch.cuda.set_device(info.local_rank % torch.cuda.device_count())
    if not dist.is_initialized():
        kw = {}
        if be == "nccl" and torch.cuda.is_available():
            kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
        timeout = datetime.timedelta(minutes=timeout_minutes)
        restart = os.environ.get("TORCHELASTIC_RESTART_COUNT")
        if restart is not None and os.environ.get("TORCHELASTIC_USE_AGENT_STORE") == "True":
            # elastic restarts (`torchrun --max-restarts N`): the agent's TCPStore outlives the failed attempt, and with it the
            # transport addresses the dead workers published; a rank of the new attempt that reads them before its peers have
            # overwritten them connects to a closed port.  Give every attempt its own key space.
            store = dist.TCPStore(os.environ["MASTER_ADDR"], int(os.environ["MASTER_PORT"]), info.world_size, False, timeout)
            kw["store"] = dist.PrefixStore(f"/b200dist/attempt_{restart}", store)
        dist.init_process_group(backend=be, rank=info.rank, world_size=info.world_size, timeout=timeout, **kw)
    return info


def device_for(info: RankInfo) -> torch.device:
    if torch.cuda.is_available():
        return torch.device("cuda", info.local_rank % torch.cuda.device_count())
    return torch.device("cpu")
