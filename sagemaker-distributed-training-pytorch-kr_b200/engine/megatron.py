"""Megatron-style pre-training runtime: initialize -> build model/optimizer/scheduler -> data iterators ->
train loop with logging / evaluation / checkpointing.

Role of megatron.{initialize,training,global_vars} as driven by pretrain_gpt.py:143-149:
``pretrain(train_valid_test_datasets_provider, model_provider, model_type, forward_step_func,
args_defaults=...)`` keeps the reference's callback contract, so an entry script written against
Megatron's API (model_provider / get_batch / loss_func / forward_step) ports by changing its imports.
"""
from __future__ import annotations

import math
import os
import signal
import sys
import time
from typing import Callable, List, Optional

import torch
import torch.distributed as dist

from ..config import megatron_args
from ..launch import env as launch_env
from ..optim.mixed import ConstantGradScaler, DynamicGradScaler, MixedPrecisionOptimizer
from ..optim.schedulers import OptimizerParamScheduler
from ..parallel import pipeline as pp
from ..utils import microbatches as mb
from ..parallel import state as mpu
from ..parallel import tensor as tp_ops
from ..utils.timers import Timers

_GLOBAL = {"args": None, "timers": None, "tokenizer": None, "signal": None, "start_time": time.time(), "sink": None, "tb": None,
           "optimizer": None}


def get_args():
    assert _GLOBAL["args"] is not None, "args is not initialized."
    return _GLOBAL["args"]


def get_timers() -> Timers:
    return _GLOBAL["timers"]


def get_tokenizer():
    return _GLOBAL["tokenizer"]


def print_rank_0(msg):
    if not dist.is_initialized() or dist.get_rank() == 0:
        print(msg, flush=True)


def print_rank_last(msg):
    if not dist.is_initialized() or dist.get_rank() == dist.get_world_size() - 1:
        print(msg, flush=True)


def is_last_rank():
    return not dist.is_initialized() or dist.get_rank() == dist.get_world_size() - 1


class _NullTokenizer:
    """`--tokenizer-type NullTokenizer --vocab-size N`: ids are the tokens, eod = N."""

    def __init__(self, vocab_size):
        self._v = int(vocab_size)
        self.eod = self._v
        self.vocab_size = self._v + 1

    def tokenize(self, text):
        return [int(x) for x in text.split(" ")]

    def detokenize(self, ids):
        return " ".join(str(i) for i in ids)


def build_tokenizer(args):
    if args.tokenizer_type in (None, "NullTokenizer"):
        vs = args.vocab_size if args.vocab_size is not None else 32767
        tok = _NullTokenizer(vs)
    elif args.tokenizer_type == "GPT2BPETokenizer":
        from ..data.tokenizers import GPT2BPETokenizer
        tok = GPT2BPETokenizer(args.vocab_file, args.merge_file)
    else:
        from ..data.tokenizers import build as build_other
        tok = build_other(args)
    args.padded_vocab_size = megatron_args.pad_vocab_size(tok.vocab_size, args.make_vocab_size_divisible_by,
                                                          args.tensor_model_parallel_size, verbose=args.rank == 0)
    return tok


class _SignalHandler:
    """--exit-signal-handler: SIGTERM -> save a checkpoint and exit cleanly (arguments.py:804-811)."""

    def __init__(self, sig=signal.SIGTERM):
        self.sig, self.received = sig, False
        self._old = signal.signal(sig, self._h)

    def _h(self, *_):
        self.received = True

    def signals_received(self) -> bool:
        if not dist.is_initialized():
            return self.received
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
        t = torch.tensor([int(self.received)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return bool(t.item())


def initialize_megatron(extra_args_provider=None, args_defaults=None, ignore_unknown_args=False, argv=None, allow_no_cuda=True):
    info = launch_env.export_torch_env()
    args = megatron_args.parse_args(extra_args_provider, ignore_unknown_args, argv)
    args.rank, args.world_size = info.rank, info.world_size
    args = megatron_args.validate_args(args, args_defaults or {}, verbose=True)
    _GLOBAL["args"] = args
    mb.build_num_microbatches_calculator(args)
    _GLOBAL["timers"] = Timers(args.timing_log_level, args.timing_log_option)
    backend = args.distributed_backend if torch.cuda.is_available() else "gloo"
    if args.rank == 0:
        print("> initializing torch distributed ...", flush=True)
    launch_env.init_distributed(backend, args.distributed_timeout_minutes)
    args.local_rank = info.local_rank
    mpu.initialize_model_parallel(args.tensor_model_parallel_size, args.pipeline_model_parallel_size,
                                  args.virtual_pipeline_model_parallel_size, args.pipeline_model_parallel_split_rank,
                                  verbose=True)
    pp.reset_links()
    if args.pipeline_model_parallel_size > 1 and torch.cuda.is_available() and args.comm_backend == "native":
        # pipeline p2p is NCCL: its send/recv kernels must find a free SM while one of the fused TP kernels (persistent,
        # spinning on the TP peer) is running, or two stages can wait on each other forever.  The fused kernels therefore size
        # their grids to (SMs - B2_RESERVED_SMS, default 16) under PP.  Measured on 4 x B200 (GPT 1.41 B, TP2 x PP2 interleaved):
        # 172 k tokens/s with the fused kernels against 131 k with NCCL tensor-parallel collectives; B2_PP_NATIVE_TP=0 selects
        # the latter (all-NCCL) path.
        if os.environ.get("B2_PP_NATIVE_TP", "1") == "1":
            from ..ops import native as _N
            _N.set_reserved_sms(int(os.environ.get("B2_RESERVED_SMS", "16")))
            if args.rank == 0:
                print("> pipeline parallel + native TP kernels: reserving SMs for NCCL p2p", flush=True)
        else:
            args.comm_backend = "nccl"
            if args.rank == 0:
                print("> pipeline parallel: tensor-parallel collectives use NCCL (B2_PP_NATIVE_TP=0)", flush=True)
    if args.rank == 0:
        print("> setting random seeds to {} ...".format(args.seed), flush=True)
    seed = args.seed + 100 * mpu.get_pipeline_model_parallel_rank()
    if args.data_parallel_random_init:
        seed += 10 * mpu.get_data_parallel_rank()
    import random
    import numpy as np
    random.seed(seed)
    np.random.seed(seed)
    tp_ops.model_parallel_cuda_manual_seed(seed)
    _GLOBAL["tokenizer"] = build_tokenizer(args)
    from ..utils import metrics as _metrics
    last = args.world_size - 1                       # the rank that prints the training log also owns the structured sinks
    _GLOBAL["sink"] = _metrics.JsonlSink(args.metrics_jsonl, rank=0 if args.rank == last else 1)
    _GLOBAL["tb"] = _metrics.ScalarWriter(args.tensorboard_dir, rank=0 if args.rank == last else 1, queue_size=args.tensorboard_queue_size)
    if args.profile:
        _metrics.enable_nvtx(True)                   # timers double as NVTX ranges inside the cudaProfilerStart/Stop window
    if args.exit_signal_handler:
        _GLOBAL["signal"] = _SignalHandler()
    # native code: dataset helpers + kernels are prebuilt in-tree; build them here if a fresh checkout
    t0 = time.time()
    from ..utils import hostlib
    hostlib.lib()
    if torch.cuda.is_available():
        from ..ops import native
        native.lib()
    print_rank_0(">>> done with native library load. time: {:.3f} seconds".format(time.time() - t0))
    return args


def get_model(model_provider_func, wrap_with_ddp=True) -> List[torch.nn.Module]:
    args = get_args()
    chunks = []
    vpp = args.virtual_pipeline_model_parallel_size
    if mpu.get_pipeline_model_parallel_world_size() > 1 and vpp is not None:
        for i in range(vpp):
            mpu.set_virtual_pipeline_model_parallel_rank(i)
            pre, post = mpu.is_pipeline_first_stage(), mpu.is_pipeline_last_stage()
            chunks.append(model_provider_func(pre_process=pre, post_process=post))
        mpu.set_virtual_pipeline_model_parallel_rank(0)
    else:
        chunks.append(model_provider_func(pre_process=mpu.is_pipeline_first_stage(), post_process=mpu.is_pipeline_last_stage()))
    n = sum(p.numel() for m in chunks for p in m.parameters())
    if mpu.get_data_parallel_rank() == 0:
        print(" > number of parameters on (tensor, pipeline) model parallel rank ({}, {}): {}".format(
            mpu.get_tensor_model_parallel_rank(), mpu.get_pipeline_model_parallel_rank(), n), flush=True)
    # tied embeddings across the first / last pipeline stage: mark the last-stage copy and sync its value
    if mpu.get_pipeline_model_parallel_world_size() > 1 and not args.untie_embeddings_and_output_weights:
        for m in chunks:
            if getattr(m, "post_process", False) and not getattr(m, "pre_process", False):
                w = m.word_embeddings_weight()
                w.shared_duplicate = True
                w.in_embedding_group = True
                with torch.no_grad():
                    w.zero_()
            elif getattr(m, "pre_process", False):
                m.word_embeddings_weight().in_embedding_group = True
        if mpu.is_rank_in_embedding_group(ignore_virtual=True):
            for m in chunks:
                if getattr(m, "pre_process", False) or getattr(m, "post_process", False):
                    if hasattr(m.language_model, "embedding"):
                        dist.all_reduce(m.word_embeddings_weight().data, group=mpu.get_embedding_group())
    return chunks


def get_optimizer_param_scheduler(optimizer):
    args = get_args()
    if args.train_iters:
        if args.lr_decay_iters is None:
            args.lr_decay_iters = args.train_iters
        lr_decay_steps = args.lr_decay_iters * args.global_batch_size
        wd_incr_steps = args.train_iters * args.global_batch_size
        lr_warmup_steps = (args.lr_warmup_fraction * lr_decay_steps) if args.lr_warmup_fraction is not None else args.lr_warmup_iters * args.global_batch_size
    elif args.train_samples:
        update_train_iters(args)
        if args.lr_decay_samples is None:
            args.lr_decay_samples = args.train_samples
        lr_decay_steps, wd_incr_steps = args.lr_decay_samples, args.train_samples
        lr_warmup_steps = (args.lr_warmup_fraction * lr_decay_steps) if args.lr_warmup_fraction is not None else args.lr_warmup_samples
    else:
        raise Exception("either train-iters or train-samples should be provided.")
    return OptimizerParamScheduler(optimizer, max_lr=args.lr, min_lr=args.min_lr, lr_warmup_steps=lr_warmup_steps, lr_decay_steps=lr_decay_steps,
                                   lr_decay_style=args.lr_decay_style, start_wd=args.start_weight_decay, end_wd=args.end_weight_decay,
                                   wd_incr_steps=wd_incr_steps, wd_incr_style=args.weight_decay_incr_style,
                                   use_checkpoint_opt_param_scheduler=args.use_checkpoint_opt_param_scheduler,
                                   override_opt_param_scheduler=args.override_opt_param_scheduler)


def update_train_iters(args):
    """sample-based training: the iteration count, including the shorter iterations of a batch-size ramp-up (training.py)"""
    if args.train_iters:
        return
    if args.rampup_batch_size is None:
        args.train_iters = args.train_samples // args.global_batch_size
    else:
        iterations, consumed = 0, 0
        while consumed <= int(args.rampup_batch_size[2]):
            mb.update_num_microbatches(consumed, consistency_check=False)
            consumed += mb.get_current_global_batch_size()
            iterations += 1
        mb.update_num_microbatches(0, consistency_check=False)
        iterations += (args.train_samples - consumed) // args.global_batch_size
        args.train_iters = iterations
    print_rank_0("setting training iterations to {}".format(args.train_iters))


def get_megatron_optimizer(model: List[torch.nn.Module]) -> MixedPrecisionOptimizer:
    args = get_args()
    scaler = None
    if args.fp16:
        scaler = ConstantGradScaler(args.loss_scale) if args.loss_scale else DynamicGradScaler(
            args.initial_loss_scale, args.min_loss_scale, 2.0, 0.5, args.loss_scale_window, args.hysteresis)
    cls = MixedPrecisionOptimizer
    if args.use_distributed_optimizer:
        from ..optim.distributed import DistributedOptimizer
        cls = DistributedOptimizer
    return cls(model, lr=args.lr, weight_decay=args.weight_decay, betas=(args.adam_beta1, args.adam_beta2), eps=args.adam_eps,
                                   clip_grad=args.clip_grad, optimizer=args.optimizer, sgd_momentum=args.sgd_momentum, grad_scaler=scaler,
                                   dp_group=mpu.get_data_parallel_group(), tp_group=mpu.get_tensor_model_parallel_group(),
                                   mp_group=mpu.get_model_parallel_group(),
                                   embedding_group=mpu.get_embedding_group() if mpu.is_rank_in_embedding_group(ignore_virtual=True) and mpu.get_pipeline_model_parallel_world_size() > 1 else None,
                                   comm_backend=args.comm_backend, log_num_zeros=bool(args.log_num_zeros_in_grad))


def setup_model_and_optimizer(model_provider_func, no_optimizer=False):
    args = get_args()
    model = get_model(model_provider_func)
    optimizer = None if no_optimizer else get_megatron_optimizer(model)
    sched = None if no_optimizer else get_optimizer_param_scheduler(optimizer)
    args.iteration = 0
    if args.load is not None:
        from ..ckpt import megatron as ckpt
        timers = get_timers()
        timers("load-checkpoint", log_level=0).start(barrier=True)
        args.iteration = ckpt.load_checkpoint(model, optimizer, sched, args)
        timers("load-checkpoint").stop(barrier=True)
        timers.log(["load-checkpoint"])
    return model, optimizer, sched


_GRAPH = {"runner": None, "tried": False}


def get_graph_runner(forward_step_func, model):
    """Lazily captures the micro-batch forward+backward into a CUDA graph (--cuda-graph; engine/graphed.py)."""
    args = get_args()
    if not getattr(args, "cuda_graph", False) or _GRAPH["tried"]:
        return _GRAPH["runner"]
    _GRAPH["tried"] = True
    from . import graphed
    if len(model) != 1 or not graphed.supported(args, model[0]):
        print_rank_0("> --cuda-graph requested but not applicable (needs PP=1, bf16/fp32, no recompute, fabric collectives for TP): running eagerly")
        return None
    if args.rampup_batch_size is not None:
        print_rank_0("> --cuda-graph replays a fixed number of micro-batches: not used with --rampup-batch-size")
        return None
    nmb = mb.get_num_microbatches()
    dev = torch.device("cuda", torch.cuda.current_device())
    args._static_batch = True
    r = graphed.GraphedMicrobatch(forward_step_func, model[0], nmb, args.micro_batch_size, args.seq_length, dev)
    r.capture()
    _GRAPH["runner"] = r
    print_rank_0("> captured micro-batch forward+backward into a CUDA graph")
    return r


def train_step(forward_step_func, data_iterator, model, optimizer, opt_param_scheduler):
    args, timers = get_args(), get_timers()
    runner = get_graph_runner(forward_step_func, model)
    optimizer.zero_grad()
    S = args.seq_length // (args.tensor_model_parallel_size if args.sequence_parallel else 1)
    shape = (S, args.micro_batch_size, args.hidden_size)
    nmb = mb.get_num_microbatches()
    timers("forward-backward", log_level=1).start(barrier=args.barrier_with_L1_time)
    if runner is not None:
        losses = runner.run_step(data_iterator)
    else:
        losses = pp.forward_backward(forward_step_func, data_iterator, model, nmb, shape, args.params_dtype, forward_only=False,
                                     grad_scale=optimizer.get_loss_scale(), timers=timers if args.timing_log_level >= 2 else None)
    timers("forward-backward").stop()
    if getattr(args, "empty_unused_memory_level", 0) >= 1 and torch.cuda.is_available():      # --empty-unused-memory-level (training.py train_step)
        torch.cuda.empty_cache()
    timers("all-grads-sync", log_level=1).start(barrier=args.barrier_with_L1_time)
    optimizer.reduce_model_grads()
    timers("all-grads-sync").stop()
    timers("optimizer", log_level=1).start(barrier=args.barrier_with_L1_time)
    ok, grad_norm, num_zeros = optimizer.step()
    timers("optimizer").stop()
    if getattr(args, "empty_unused_memory_level", 0) >= 2 and torch.cuda.is_available():
        torch.cuda.empty_cache()
    if ok:
        increment = nmb * args.micro_batch_size * args.data_parallel_size
        opt_param_scheduler.step(increment=increment)
        skipped = 0
    else:
        skipped = 1
    loss_reduced = {}
    if mpu.is_pipeline_last_stage(ignore_virtual=True) and losses:
        for k in losses[0]:
            loss_reduced[k] = sum(x[k] for x in losses) / len(losses)
    return loss_reduced, skipped, grad_norm, num_zeros


def training_log(loss_dict, total_loss_dict, learning_rate, iteration, loss_scale, report_memory_flag, skipped_iter, grad_norm,
                 params_norm=None, num_zeros_in_grad=None):
    """'iteration N/ M | consumed samples: ... | elapsed time per iteration (ms): ... | learning rate: ... |
    global batch size: ... | lm loss: ... | loss scale: ... | grad norm: ... | number of skipped iterations: ... |
    number of nan iterations: ... |'  (format of 3_...ipynb:4347)."""
    args, timers = get_args(), get_timers()
    adv, skp, nan = "advanced iterations", "skipped iterations", "nan iterations"
    total_loss_dict[adv] = total_loss_dict.get(adv, 0) + (0 if skipped_iter else 1)
    total_loss_dict[skp] = total_loss_dict.get(skp, 0) + skipped_iter
    got_nan = False
    for k, v in loss_dict.items():
        if not skipped_iter:
            total_loss_dict[k] = total_loss_dict.get(k, 0.0) + v
    total_loss_dict[nan] = total_loss_dict.get(nan, 0) + int(got_nan)
    if iteration % args.log_interval == 0:
        elapsed = timers("interval-time").elapsed(barrier=True)
        per_iter = elapsed / (total_loss_dict[adv] + total_loss_dict[skp])
        s = " iteration {:8d}/{:8d} |".format(iteration, args.train_iters)
        logged = {}
        s += " consumed samples: {:12d} |".format(args.consumed_train_samples)
        s += " elapsed time per iteration (ms): {:.1f} |".format(per_iter * 1000.0)
        s += " learning rate: {:.3E} |".format(learning_rate)
        batch_size = mb.get_current_global_batch_size()
        s += " global batch size: {:5d} |".format(batch_size)
        for k in list(total_loss_dict.keys()):
            if k in (adv, skp, nan):
                continue
            v = total_loss_dict[k]
            avg = float(v) / float(max(1, total_loss_dict[adv]))
            if avg > 0.0:
                s += " {}: {:.6E} |".format(k, avg)
                logged[k] = avg
            total_loss_dict[k] = 0.0
        s += " loss scale: {:.1f} |".format(loss_scale)
        rec = {"iteration": iteration, "consumed_samples": args.consumed_train_samples, "ms_per_iteration": per_iter * 1000.0,
               "learning_rate": learning_rate, "global_batch_size": batch_size, "loss_scale": loss_scale,
               "skipped_iterations": total_loss_dict[skp], **{k.replace(" ", "_"): v for k, v in logged.items()}}
        if grad_norm is not None:
            gn = math.sqrt(float(grad_norm)) if torch.is_tensor(grad_norm) else float(grad_norm)
            s += " grad norm: {:.3f} |".format(gn)
            rec["grad_norm"] = gn
        if num_zeros_in_grad is not None:
            s += " num zeros: {:.1f} |".format(float(num_zeros_in_grad))
            rec["num_zeros_in_grad"] = float(num_zeros_in_grad)
        if params_norm is not None:
            s += " params norm: {:.3f} |".format(params_norm)
            rec["params_norm"] = params_norm
        s += " number of skipped iterations: {:3d} |".format(total_loss_dict[skp])
        s += " number of nan iterations: {:3d} |".format(total_loss_dict[nan])
        tokens_per_s = batch_size * args.seq_length / max(per_iter, 1e-9)
        s += " tokens/s: {:.0f} |".format(tokens_per_s)
        sink, tb = _GLOBAL["sink"], _GLOBAL["tb"]
        if (sink is not None and sink.enabled) or (tb is not None and tb.enabled):
            from ..utils import metrics as _metrics
            fpt = _metrics.model_flops_per_token(args.num_layers, args.hidden_size, args.ffn_hidden_size, args.padded_vocab_size, args.seq_length,
                                                 swiglu=bool(args.swiglu), kv_ratio=(args.num_query_groups / args.num_attention_heads)
                                                 if args.group_query_attention else 1.0)
            rec.update(tokens_per_s=tokens_per_s, mfu=_metrics.mfu(tokens_per_s / max(1, args.world_size), fpt), world_size=args.world_size)
            if sink is not None:
                sink.write(**rec)
            if tb is not None and tb.enabled and iteration % args.tensorboard_log_interval == 0:
                # the reference's switches (arguments.py:652-679): learning rate and loss scale on by default, batch size / world size /
                # timers / memory on request
                gate = {"learning_rate": getattr(args, "log_learning_rate_to_tensorboard", True),
                        "loss_scale": getattr(args, "log_loss_scale_to_tensorboard", True),
                        "global_batch_size": getattr(args, "log_batch_size_to_tensorboard", False),
                        "world_size": getattr(args, "log_world_size_to_tensorboard", False)}
                for k, v in rec.items():
                    if k not in ("iteration", "time") and isinstance(v, (int, float)) and gate.get(k, True):
                        tb.add_scalar(k.replace("_", "-"), v, iteration)
                        if k.endswith("loss") or k in ("learning_rate", "loss_scale", "global_batch_size"):
                            tb.add_scalar(k.replace("_", "-") + " vs samples", v, args.consumed_train_samples)
                if getattr(args, "log_memory_to_tensorboard", False) and torch.cuda.is_available():
                    st = torch.cuda.memory_stats()
                    for tag, key in (("mem-reserved-bytes", "reserved_bytes.all.current"), ("mem-allocated-bytes", "allocated_bytes.all.current"),
                                     ("mem-allocated-count", "allocation.all.current")):
                        tb.add_scalar(tag, st.get(key, 0), iteration)
                tb.flush()
        total_loss_dict[adv] = total_loss_dict[skp] = total_loss_dict[nan] = 0
        print_rank_last(s)
        if report_memory_flag and learning_rate > 0.0 and torch.cuda.is_available():
            mib = 1024.0 * 1024.0
            print("[Rank {}] (after {} iterations) memory (MB) | allocated: {} | max allocated: {} | reserved: {} | max reserved: {}".format(
                dist.get_rank() if dist.is_initialized() else 0, iteration, torch.cuda.memory_allocated() / mib, torch.cuda.max_memory_allocated() / mib,
                torch.cuda.memory_reserved() / mib, torch.cuda.max_memory_reserved() / mib), flush=True)
            report_memory_flag = False
        names = ["forward-backward", "forward-compute", "backward-compute", "batch-generator", "all-grads-sync", "optimizer"]
        if getattr(args, "log_timers_to_tensorboard", False) and args.tensorboard_dir and iteration % args.tensorboard_log_interval == 0:
            # collective (max over ranks): every rank calls it, the rank that owns the writer records
            timers.write(names, tb if (tb is not None and tb.enabled) else None, iteration, normalizer=args.log_interval)
        timers.log(names, normalizer=args.log_interval)
    return report_memory_flag


def evaluate(forward_step_func, data_iterator, model, verbose=False):
    args = get_args()
    for m in model:
        m.eval()
    total = {}
    S = args.seq_length // (args.tensor_model_parallel_size if args.sequence_parallel else 1)
    shape = (S, args.micro_batch_size, args.hidden_size)
    nmb = mb.get_num_microbatches()
    with torch.no_grad():
        for it in range(args.eval_iters):
            if verbose and (it + 1) % args.log_interval == 0:
                print_rank_0("Evaluating iter {}/{}".format(it + 1, args.eval_iters))
            losses = pp.forward_backward(forward_step_func, data_iterator, model, nmb, shape, args.params_dtype, forward_only=True)
            if mpu.is_pipeline_last_stage(ignore_virtual=True):
                for d in losses:
                    for k, v in d.items():
                        total[k] = total.get(k, 0.0) + v
            args.consumed_valid_samples += mb.get_current_global_batch_size()
    for m in model:
        m.train()
    for k in total:
        total[k] = total[k] / (args.eval_iters * nmb)
    return total


def evaluate_and_print_results(prefix, forward_step_func, data_iterator, model, iteration, verbose=False):
    total = evaluate(forward_step_func, data_iterator, model, verbose)
    s = " validation loss at {} | ".format(prefix)
    for k, v in total.items():
        val = float(v)
        s += "{} value: {:.6E} | ".format(k, val)
        s += "{} PPL: {:.6E} | ".format(k, math.exp(min(20, val)))
    tb = _GLOBAL.get("tb")
    if tb is not None and tb.enabled and isinstance(iteration, int):
        args = get_args()
        for k, v in total.items():
            tb.add_scalar(f"{k} validation", float(v), iteration)
            tb.add_scalar(f"{k} validation vs samples", float(v), args.consumed_train_samples)
            if getattr(args, "log_validation_ppl_to_tensorboard", False):
                tb.add_scalar(f"{k} validation ppl", math.exp(min(20, float(v))), iteration)
                tb.add_scalar(f"{k} validation ppl vs samples", math.exp(min(20, float(v))), args.consumed_train_samples)
        tb.flush()
    print_rank_last("-" * (len(s) + 1))
    print_rank_last(s)
    print_rank_last("-" * (len(s) + 1))
    return total


def save_checkpoint_and_time(iteration, model, optimizer, opt_param_scheduler):
    from ..ckpt import megatron as ckpt
    timers = get_timers()
    timers("save-checkpoint", log_level=0).start(barrier=True)
    ckpt.save_checkpoint(iteration, model, optimizer, opt_param_scheduler, get_args())
    timers("save-checkpoint").stop(barrier=True)
    timers.log(["save-checkpoint"])


def train(forward_step_func, model, optimizer, opt_param_scheduler, train_data_iterator, valid_data_iterator, process_non_loss_data_func=None):
    args, timers = get_args(), get_timers()
    for m in model:
        m.train()
    total_loss_dict = {}
    iteration = args.iteration
    timers("interval-time", log_level=0).start(barrier=True)
    report_memory_flag = True
    prof_on = False
    while iteration < args.train_iters:
        if args.profile and torch.cuda.is_available() and (dist.get_rank() if dist.is_initialized() else 0) in args.profile_ranks:
            if iteration == args.profile_step_start:
                torch.cuda.cudart().cudaProfilerStart(); prof_on = True
            if iteration == args.profile_step_end and prof_on:
                torch.cuda.cudart().cudaProfilerStop(); prof_on = False
        args.curr_iteration = iteration
        mb.update_num_microbatches(args.consumed_train_samples)
        loss_dict, skipped, grad_norm, num_zeros = train_step(forward_step_func, train_data_iterator, model, optimizer, opt_param_scheduler)
        iteration += 1
        args.consumed_train_samples += mb.get_current_global_batch_size()
        loss_scale = optimizer.get_loss_scale()
        lr = optimizer.param_groups[0]["lr"]
        # loss values stay device tensors between log points; they are only read on the host at a log boundary
        params_norm = optimizer.params_norm() if (args.log_params_norm and iteration % args.log_interval == 0 and hasattr(optimizer, "params_norm")) else None
        report_memory_flag = training_log(loss_dict, total_loss_dict, lr, iteration, loss_scale, report_memory_flag, skipped, grad_norm,
                                          params_norm, num_zeros)
        if args.eval_interval and iteration % args.eval_interval == 0 and args.do_valid:
            evaluate_and_print_results("iteration {}".format(iteration), forward_step_func, valid_data_iterator, model, iteration, False)
        saved = False
        if args.save and args.save_interval and iteration % args.save_interval == 0:
            save_checkpoint_and_time(iteration, model, optimizer, opt_param_scheduler)
            saved = True
        if _GLOBAL["signal"] is not None and _GLOBAL["signal"].signals_received():
            if args.save and not saved:
                save_checkpoint_and_time(iteration, model, optimizer, opt_param_scheduler)
            print_rank_0("exiting program after receiving SIGTERM.")
            sys.exit()
        if args.exit_duration_in_mins:
            mins = (time.time() - _GLOBAL["start_time"]) / 60.0
            if mins > args.exit_duration_in_mins:
                if args.save and not saved:
                    save_checkpoint_and_time(iteration, model, optimizer, opt_param_scheduler)
                print_rank_0("exiting program after {} minutes".format(mins))
                sys.exit()
        if args.exit_interval and iteration % args.exit_interval == 0:
            if args.save and not saved:
                save_checkpoint_and_time(iteration, model, optimizer, opt_param_scheduler)
            print_rank_0("exiting program at iteration {}".format(iteration))
            sys.exit()
        inj = args.inject_fault or os.environ.get("B200DIST_INJECT_FAULT")     # "rank:iteration" test hook
        if inj and os.environ.get("TORCHELASTIC_RESTART_COUNT", "0") == "0":      # only in the first attempt of an elastic job
            r, it = (int(x) for x in inj.split(":"))
            if (dist.get_rank() if dist.is_initialized() else 0) == r and iteration == it:
                raise RuntimeError(f"injected fault on rank {r} at iteration {it}")
    return iteration


def cyclic_iter(it):
    while True:
        for x in it:
            yield x


def build_train_valid_test_data_iterators(provider):
    args = get_args()
    print_rank_0("> building train, validation, and test datasets ...")
    if args.iteration > 0 and args.consumed_train_samples == 0:
        args.consumed_train_samples = args.iteration * args.global_batch_size
    if args.train_samples:
        train_samples = args.train_samples
    else:
        train_samples = args.train_iters * args.global_batch_size
    eval_iters = (args.train_iters // max(args.eval_interval, 1) + 1) * args.eval_iters
    nums = [train_samples, eval_iters * args.global_batch_size, args.eval_iters * args.global_batch_size]
    print_rank_0(" > datasets target sizes (minimum size):")
    print_rank_0("    train:      {}".format(nums[0]))
    print_rank_0("    validation: {}".format(nums[1]))
    print_rank_0("    test:       {}".format(nums[2]))
    from ..data.samplers import build_pretraining_data_loader
    iters = [None, None, None]
    if mpu.get_tensor_model_parallel_rank() == 0:
        tr, va, te = provider(nums)
        loaders = [build_pretraining_data_loader(tr, args.consumed_train_samples, args),
                   build_pretraining_data_loader(va, args.consumed_valid_samples, args),
                   build_pretraining_data_loader(te, 0, args)]
        for i, l in enumerate(loaders):
            if l is not None:
                iters[i] = iter(l) if args.dataloader_type == "single" else iter(cyclic_iter(l))
    # only TP rank 0 owns loaders; every rank must agree on which phases run (flags travel over the TP group)
    flags = torch.tensor([int(iters[0] is not None and args.train_iters > 0), int(iters[1] is not None and args.eval_iters > 0),
                          int(iters[2] is not None and args.eval_iters > 0)], dtype=torch.int64,
                         device="cuda" if (torch.cuda.is_available() and args.distributed_backend == "nccl") else "cpu")
    if mpu.get_tensor_model_parallel_world_size() > 1:
        dist.broadcast(flags, mpu.get_tensor_model_parallel_src_rank(), group=mpu.get_tensor_model_parallel_group())
    args.do_train, args.do_valid, args.do_test = (bool(x) for x in flags.tolist())
    return iters


def pretrain(train_valid_test_dataset_provider, model_provider, model_type, forward_step_func, process_non_loss_data_func=None,
             extra_args_provider=None, args_defaults=None, argv=None):
    args = initialize_megatron(extra_args_provider, args_defaults or {}, argv=argv)
    timers = get_timers()
    print_rank_0("time to initialize megatron (seconds): {:.3f}".format(time.time() - _GLOBAL["start_time"]))
    timers("model-and-optimizer-setup", log_level=0).start(barrier=True)
    model, optimizer, sched = setup_model_and_optimizer(model_provider)
    timers("model-and-optimizer-setup").stop()
    timers("train/valid/test-data-iterators-setup", log_level=0).start(barrier=True)
    if args.virtual_pipeline_model_parallel_size is not None and len(model) > 1:
        # interleaved schedule: every model chunk draws the SAME micro-batch sequence from its own iterator (first chunk
        # needs the tokens, last chunk the labels of one and the same batch)
        train_it, valid_it, test_it = [], [], []
        for i in range(len(model)):
            mpu.set_virtual_pipeline_model_parallel_rank(i)
            t_i, v_i, te_i = build_train_valid_test_data_iterators(train_valid_test_dataset_provider)
            train_it.append(t_i); valid_it.append(v_i); test_it.append(te_i)
        mpu.set_virtual_pipeline_model_parallel_rank(0)
    else:
        train_it, valid_it, test_it = build_train_valid_test_data_iterators(train_valid_test_dataset_provider)
    timers("train/valid/test-data-iterators-setup").stop()
    print_rank_0("done with setup ...")
    timers.log(["model-and-optimizer-setup", "train/valid/test-data-iterators-setup"], barrier=True)
    iteration = args.iteration
    if not args.skip_train:
        print_rank_0("training ...")
        if args.train_iters > 0:
            iteration = train(forward_step_func, model, optimizer, sched, train_it, valid_it, process_non_loss_data_func)
        print_rank_0("[after training is done]")
    if args.save and iteration != 0 and (args.save_interval is None or iteration % args.save_interval != 0):
        save_checkpoint_and_time(iteration, model, optimizer, sched)
    if args.do_valid:
        evaluate_and_print_results("the end of training for val data", forward_step_func, valid_it, model, iteration, False)
    if args.do_test:
        evaluate_and_print_results("the end of training for test data", forward_step_func, test_it, model, iteration, False)
    return model, optimizer
