"""Megatron checkpoint layout:

    <save>/latest_checkpointed_iteration.txt
    <save>/iter_{it:07d}/mp_rank_{tp:02d}[_{pp:03d}]/model_optim_rng.pt

written by data-parallel rank 0 of every (tp, pp) position.  The file is a dict with keys ``args``,
``checkpoint_version`` (3.0), ``iteration``, ``model`` (or ``model0..k`` with virtual pipeline stages),
``optimizer``, ``opt_param_scheduler``, ``rng_state``.  The ``model`` entry nests parameters as
language_model -> {embedding -> {word_embeddings, position_embeddings}, encoder -> {layers.N....,
final_layernorm.*}} exactly like Megatron's ``state_dict_for_save_checkpoint`` so files written by the
reference stack load here and vice versa (flags: arguments.py:922-956).

``optimizer``: this framework writes {step, state[name] -> {exp_avg, exp_avg_sq, fp32_param}, grad_scaler, param_groups}; a file
written by the reference stack carries Megatron's {optimizer: torch-Adam state_dict, grad_scaler, fp32_from_fp16_params}
instead and is recognised and mapped onto the flat arenas by `MixedPrecisionOptimizer.load_state_dict` (same tensor-parallel
size required; otherwise load the weights only with --no-load-optim).  Megatron itself cannot read this framework's optimizer
section or the per-DP-rank `distrib_optim_dpNNN.pt` files: pass --no-load-optim there."""
from __future__ import annotations

import os
import random
import sys
from typing import List

import numpy as np
import torch
import torch.distributed as dist

from ..parallel import state as mpu
from ..parallel import tensor as tp_ops


def _p0(m):
    if not dist.is_initialized() or dist.get_rank() == 0:
        print(m, flush=True)


def get_checkpoint_name(root, iteration, release=False, tp_rank=None, pp_rank=None, pp_world=None):
    d = "release" if release else "iter_{:07d}".format(iteration)
    tp_rank = mpu.get_tensor_model_parallel_rank() if tp_rank is None else tp_rank
    pp_rank = mpu.get_pipeline_model_parallel_rank() if pp_rank is None else pp_rank
    pp_world = mpu.get_pipeline_model_parallel_world_size() if pp_world is None else pp_world
    sub = f"mp_rank_{tp_rank:02d}" if pp_world == 1 else f"mp_rank_{tp_rank:02d}_{pp_rank:03d}"
    return os.path.join(root, d, sub, "model_optim_rng.pt")


def get_checkpoint_tracker_filename(root):
    return os.path.join(root, "latest_checkpointed_iteration.txt")


def model_state_dict_nested(model: torch.nn.Module) -> dict:
    """flat `language_model.encoder.layers.0.x.weight` keys -> Megatron's nested sections."""
    lm = {"embedding": {}, "encoder": {}}
    for k, v in model.state_dict().items():
        assert k.startswith("language_model."), k
        rest = k[len("language_model."):]
        if rest.startswith("embedding.word_embeddings."):
            lm["embedding"].setdefault("word_embeddings", {})[rest[len("embedding.word_embeddings."):]] = v
        elif rest.startswith("embedding.position_embeddings."):
            lm["embedding"].setdefault("position_embeddings", {})[rest[len("embedding.position_embeddings."):]] = v
        elif rest.startswith("encoder."):
            lm["encoder"][rest[len("encoder."):]] = v
        elif rest.startswith("output_layer."):
            lm.setdefault("output_layer", {})[rest[len("output_layer."):]] = v
        else:
            lm[rest] = v
    if not lm["embedding"]:
        del lm["embedding"]
    return {"language_model": lm}


def load_nested_state_dict(model: torch.nn.Module, nested: dict, strict: bool = True):
    lm = nested["language_model"]
    flat = {}
    for sec, d in lm.get("embedding", {}).items():
        for k, v in d.items():
            flat[f"language_model.embedding.{sec}.{k}"] = v
    enc = lm.get("encoder", lm.get("transformer", {}))
    for k, v in enc.items():
        flat[f"language_model.encoder.{k}"] = v
    for k, v in lm.get("output_layer", {}).items():
        flat[f"language_model.output_layer.{k}"] = v
    own = model.state_dict()
    missing = [k for k in own if k not in flat]
    unexpected = [k for k in flat if k not in own]
    if strict and (missing or unexpected):
        raise RuntimeError(f"checkpoint mismatch: missing {missing[:4]} unexpected {unexpected[:4]}")
    with torch.no_grad():
        for k, v in flat.items():
            if k in own:
                own[k].copy_(v.to(own[k].dtype))


def get_rng_state():
    st = {"random_rng_state": random.getstate(), "np_rng_state": np.random.get_state(), "torch_rng_state": torch.get_rng_state(),
          "cuda_rng_state": torch.cuda.get_rng_state() if torch.cuda.is_available() else None,
          "rng_tracker_states": tp_ops.get_cuda_rng_tracker().get_states()}
    return [st]


def save_checkpoint(iteration, model: List[torch.nn.Module], optimizer, opt_param_scheduler, args):
    _p0("saving checkpoint at iteration {:7d} to {}".format(iteration, args.save))
    name = get_checkpoint_name(args.save, iteration)
    if getattr(args, "use_distributed_optimizer", False) and optimizer is not None and not args.no_save_optim:
        # optimizer state is sharded over DP: every DP rank writes its own slice next to the model file
        os.makedirs(os.path.dirname(name), exist_ok=True)
        torch.save(optimizer.state_dict(), os.path.join(os.path.dirname(name), "distrib_optim_dp{:03d}.pt".format(mpu.get_data_parallel_rank())))
    if mpu.get_data_parallel_rank() == 0:
        sd = {"args": args, "checkpoint_version": 3.0, "iteration": iteration}
        if len(model) == 1:
            sd["model"] = model_state_dict_nested(model[0])
        else:
            for i, m in enumerate(model):
                sd["model%d" % i] = model_state_dict_nested(m)
        if not args.no_save_optim:
            if optimizer is not None and not getattr(args, "use_distributed_optimizer", False):
                sd["optimizer"] = optimizer.state_dict()
            if opt_param_scheduler is not None:
                sd["opt_param_scheduler"] = opt_param_scheduler.state_dict()
        if not args.no_save_rng:
            sd["rng_state"] = get_rng_state()
        os.makedirs(os.path.dirname(name), exist_ok=True)
        torch.save(sd, name)
    if dist.is_initialized():
        dist.barrier()
    _p0("  successfully saved checkpoint at iteration {:7d} to {}".format(iteration, args.save))
    if not dist.is_initialized() or dist.get_rank() == 0:
        with open(get_checkpoint_tracker_filename(args.save), "w") as f:
            f.write(str(iteration))
    if dist.is_initialized():
        dist.barrier()


def read_metadata(tracker):
    with open(tracker) as f:
        s = f.read().strip()
    if s == "release":
        return 0, True
    return int(s), False


def load_checkpoint(model: List[torch.nn.Module], optimizer, opt_param_scheduler, args, load_arg="load", strict=True) -> int:
    load_dir = getattr(args, load_arg)
    tracker = get_checkpoint_tracker_filename(load_dir)
    if not os.path.isfile(tracker):
        _p0("WARNING: could not find the metadata file {} ".format(tracker))
        _p0("    will not load any checkpoints and will start from random")
        if args.exit_on_missing_checkpoint:
            _p0(">> '--exit-on-missing-checkpoint' set ... exiting. <<")
            if dist.is_initialized():
                dist.barrier()
            sys.exit()
        return 0
    iteration, release = read_metadata(tracker)
    name = get_checkpoint_name(load_dir, iteration, release)
    _p0(f" loading checkpoint from {load_dir} at iteration {iteration}")
    sd = torch.load(name, map_location="cpu", weights_only=False)
    if args.finetune or release:
        iteration = 0
    else:
        iteration = sd.get("iteration", sd.get("total_iters", 0))
    if "args" in sd and not args.finetune:
        ca = sd["args"]
        args.consumed_train_samples = getattr(ca, "consumed_train_samples", 0)
        args.consumed_valid_samples = getattr(ca, "consumed_valid_samples", 0)
        if args.use_checkpoint_args:
            for k in ("num_layers", "hidden_size", "ffn_hidden_size", "seq_length", "num_attention_heads", "kv_channels",
                      "max_position_embeddings", "padded_vocab_size", "position_embedding_type"):
                if hasattr(ca, k):
                    setattr(args, k, getattr(ca, k))
    if len(model) == 1:
        load_nested_state_dict(model[0], sd["model"], strict)
    else:
        for i, m in enumerate(model):
            load_nested_state_dict(m, sd["model%d" % i], strict)
    if not release and not args.finetune and not args.no_load_optim:
        shard_file = os.path.join(os.path.dirname(name), "distrib_optim_dp{:03d}.pt".format(mpu.get_data_parallel_rank()))
        if optimizer is not None and getattr(args, "use_distributed_optimizer", False) and os.path.isfile(shard_file):
            optimizer.load_state_dict(torch.load(shard_file, map_location="cpu", weights_only=False))
        elif optimizer is not None and "optimizer" in sd:
            optimizer.load_state_dict(sd["optimizer"])
        if opt_param_scheduler is not None and "opt_param_scheduler" in sd:
            opt_param_scheduler.load_state_dict(sd["opt_param_scheduler"])
    if not release and not args.finetune and not args.no_load_rng and "rng_state" in sd:
        rs = sd["rng_state"][0]
        random.setstate(rs["random_rng_state"])
        np.random.set_state(rs["np_rng_state"])
        torch.set_rng_state(rs["torch_rng_state"])
        if torch.cuda.is_available() and rs.get("cuda_rng_state") is not None:
            torch.cuda.set_rng_state(rs["cuda_rng_state"])
        if rs.get("rng_tracker_states"):
            tp_ops.get_cuda_rng_tracker().set_states(rs["rng_tracker_states"])
    if optimizer is not None and (release or args.finetune or args.no_load_optim):
        optimizer.reload_model_params()        # fp32 master copy <- the freshly loaded model weights
    if dist.is_initialized():
        dist.barrier()
    _p0(f"  successfully loaded checkpoint from {load_dir} at iteration {iteration}")
    return iteration
