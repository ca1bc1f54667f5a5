"""CUDA-graph execution of the micro-batch forward+backward.

A GPT-2-small micro-batch is ~300 kernel launches of 10-100 us each; issued one by one from Python the
host cannot keep a B200 busy (measured: 33 ms of host issue time per 42 ms step on one GPU, and fully
host-bound with tensor parallelism).  Instead of a tracing compiler, the step is captured ONCE into a CUDA
graph - the native kernels take raw pointers, the TMA descriptors are baked into the kernel parameters, the
gradient arena and the token buffer are static - and replayed per micro-batch.  Everything that varies
between steps (learning rate, Adam bias correction, loss-scale bookkeeping) stays outside the graph in the
optimizer step, which is only a handful of launches.

Constraints (checked by ``supported``): pipeline size 1, fixed micro-batch shape.  Dropout: the counter-based seeds are
drawn on the host, so a captured kernel node would repeat its mask on every replay; the graph therefore starts with a
one-thread kernel that bumps a device-side replay counter which every dropout kernel mixes into its seed
(``ops.native.dropout_replay_counter``) - forward and backward of one replay agree, successive replays differ.  Collectives inside the captured region (tensor-parallel NCCL or the fused fabric
kernels, the data-parallel loss average) are captured with it.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional

import torch

from ..ops import native as N
import torch.distributed as dist

from ..parallel import state as mpu


class _StaticIter:
    def __init__(self, batch: Dict[str, torch.Tensor]):
        self.batch = batch

    def __iter__(self):
        return self

    def __next__(self):
        return self.batch


def supported(args, model=None) -> bool:
    ok = (torch.cuda.is_available() and args.pipeline_model_parallel_size == 1
          and not args.fp16 and args.recompute_granularity is None
          and getattr(args, "num_experts", None) is None)          # Switch-MLP routing (`nonzero`) is data dependent: not capturable
    if ok and args.tensor_model_parallel_size > 1:
        # tensor-parallel collectives inside the captured region must be the fabric kernels (plain launches);
        # NCCL calls are kept out of captured graphs
        comm = getattr(model, "comm", None)
        ok = comm is not None and comm.fabric is not None and comm.sp
    return ok


class GraphedMicrobatch:
    def __init__(self, forward_step_func: Callable, model: torch.nn.Module, num_microbatches: int, micro_batch_size: int,
                 seq_length: int, device: torch.device):
        self.fwd, self.model, self.M = forward_step_func, model, num_microbatches
        self.tokens = torch.zeros(micro_batch_size, seq_length + 1, dtype=torch.int64, device=device)
        self.loss_acc: Dict[str, torch.Tensor] = {}
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.tp_group = mpu.get_tensor_model_parallel_group()
        self.tp_size = mpu.get_tensor_model_parallel_world_size()
        self.tp_src = mpu.get_tensor_model_parallel_src_rank()
        self.is_src = mpu.get_tensor_model_parallel_rank() == 0
        self._static_iter = _StaticIter({"text": self.tokens})
        self.dropout = any(getattr(m, "training", False) for m in [model]) and (
            getattr(getattr(model, "cfg", None), "hidden_dropout", 0.0) > 0.0 or getattr(getattr(model, "cfg", None), "attention_dropout", 0.0) > 0.0)
        if self.dropout:
            N.dropout_replay_counter(device)

    def _one(self):
        if self.dropout:
            N.bump_dropout_replay_counter()        # first node of the captured graph: new dropout masks per replay
        out, loss_func = self.fwd(self._static_iter, self.model)
        loss, reduced = loss_func(out)
        (loss / self.M).backward()
        for k, v in reduced.items():
            if k not in self.loss_acc:
                self.loss_acc[k] = torch.zeros((), dtype=torch.float32, device=self.tokens.device)
            self.loss_acc[k].add_(v.detach().float() / self.M)

    def capture(self):
        # eager warm-up on a side stream (allocator / lazy-init / cudaFuncSetAttribute paths), then capture
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):
                self._one()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        n0 = N.launches()
        with torch.cuda.graph(self.graph):
            self._one()
        self.kernels_per_replay = N.launches() - n0     # libb200dist kernel nodes in the captured micro-batch
        torch.cuda.synchronize()

    def load_tokens(self, data_iterator):
        """TP-source rank copies its (pinned) host batch into the static device buffer; the other TP ranks
        receive it by broadcast.  No host synchronisation."""
        if self.is_src:
            b = next(data_iterator)["text"]
            self.tokens.copy_(b, non_blocking=True)
        if self.tp_size > 1:
            dist.broadcast(self.tokens, self.tp_src, group=self.tp_group)

    def run_step(self, data_iterator) -> List[Dict[str, torch.Tensor]]:
        for v in self.loss_acc.values():
            v.zero_()
        for _ in range(self.M):
            self.load_tokens(data_iterator)
            self.graph.replay()
            N.count_launches(self.kernels_per_replay)      # replayed kernels are launches too (bench.py reports the total)
        out = {k: v.clone() for k, v in self.loss_acc.items()}
        dp = mpu.get_data_parallel_world_size()
        if dp > 1:     # data-parallel average of the logged loss: once per step, outside the captured region
            for v in out.values():
                dist.all_reduce(v, group=mpu.get_data_parallel_group())
                v.div_(dp)
        return [out]


class GraphedStep:
    """Any fixed-shape training closure (zero-grad + forward + backward + fabric gradient reduction) captured once and
    replayed: used by the DDP workloads, whose ~2000 small launches per step are host-bound when issued from Python.
    `fn(*static_inputs)` returns a tensor or a tuple of tensors (loss, logits, ...) that live in the graph's memory
    pool and are overwritten by the next replay; inputs are copied into static buffers before each replay.
    NCCL collectives must stay outside `fn` (only the fabric kernels are captured)."""

    def __init__(self, fn: Callable, example_inputs):
        self.fn = fn
        self.static_in = [torch.empty_like(t) for t in example_inputs]
        for s, t in zip(self.static_in, example_inputs):
            s.copy_(t)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                fn(*self.static_in)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        n0 = N.launches()
        with torch.cuda.graph(self.graph):
            self.static_out = fn(*self.static_in)
        self.kernels_per_replay = N.launches() - n0
        torch.cuda.synchronize()

    def __call__(self, *inputs):
        for s, t in zip(self.static_in, inputs):
            s.copy_(t, non_blocking=True)
        self.graph.replay()
        N.count_launches(self.kernels_per_replay)
        return self.static_out
