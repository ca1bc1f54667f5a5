"""DeepSpeed JSON config reader with HF-style "auto" resolution (reference: configs/default_offload_opt_param.json
T14 / -original.json T15; HfTrainerDeepSpeedConfig in transformers/integrations/deepspeed.py).

"auto" fields are filled from the TrainingArguments and the model's hidden size exactly as the HF integration does:
  optimizer.params {lr, betas, eps, weight_decay}  <- learning_rate, (adam_beta1, adam_beta2), adam_epsilon, weight_decay
  scheduler.params {warmup_min_lr=0, warmup_max_lr=lr, warmup_num_steps, total_num_steps}
  bf16.enabled / fp16.enabled                      <- --bf16 / --fp16
  train_micro_batch_size_per_gpu, gradient_accumulation_steps, train_batch_size = micro * GA * world
  gradient_clipping                                <- max_grad_norm
  zero_optimization: reduce_bucket_size = hidden^2, stage3_prefetch_bucket_size = 0.9 hidden^2,
                     stage3_param_persistence_threshold = 10 hidden
"""
from __future__ import annotations

import copy
import json
from dataclasses import dataclass, field
from typing import Any, Dict, Optional


def _walk_auto(d, path=""):
    for k, v in d.items():
        p = f"{path}.{k}" if path else k
        if isinstance(v, dict):
            yield from _walk_auto(v, p)
        elif v == "auto":
            yield p


@dataclass
class ZeroConfig:
    stage: int = 0
    offload_optimizer_device: str = "none"
    offload_param_device: str = "none"
    pin_memory: bool = True
    overlap_comm: bool = True
    contiguous_gradients: bool = True
    sub_group_size: float = 1e9
    reduce_bucket_size: int = 0
    stage3_prefetch_bucket_size: int = 0
    stage3_param_persistence_threshold: int = 0
    stage3_max_live_parameters: float = 1e9
    stage3_max_reuse_distance: float = 1e9
    stage3_gather_16bit_weights_on_model_save: bool = False


@dataclass
class DSConfig:
    raw: Dict[str, Any] = field(default_factory=dict)
    bf16: bool = False
    fp16: bool = False
    optimizer_type: Optional[str] = None
    optimizer_params: Dict[str, Any] = field(default_factory=dict)
    scheduler_type: Optional[str] = None
    scheduler_params: Dict[str, Any] = field(default_factory=dict)
    zero: ZeroConfig = field(default_factory=ZeroConfig)
    gradient_accumulation_steps: int = 1
    gradient_clipping: float = 0.0
    steps_per_print: int = 10
    train_batch_size: int = 0
    train_micro_batch_size_per_gpu: int = 0
    wall_clock_breakdown: bool = False


def load_ds_config(path_or_dict) -> Dict[str, Any]:
    if isinstance(path_or_dict, dict):
        return copy.deepcopy(path_or_dict)
    with open(path_or_dict) as f:
        return json.load(f)


def resolve_auto(cfg: Dict[str, Any], args, hidden_size: int, world_size: int, num_training_steps: int) -> DSConfig:
    """fills every "auto" and returns the typed view; `cfg` is updated in place (it is what gets saved as ds_config)."""
    def setp(path, value):
        node = cfg
        keys = path.split(".")
        for k in keys[:-1]:
            node = node.get(k)
            if node is None:
                return
        if node.get(keys[-1]) == "auto":
            node[keys[-1]] = value

    setp("bf16.enabled", bool(args.bf16))
    setp("fp16.enabled", bool(args.fp16))
    setp("optimizer.params.lr", args.learning_rate)
    setp("optimizer.params.betas", [args.adam_beta1, args.adam_beta2])
    setp("optimizer.params.eps", args.adam_epsilon)
    setp("optimizer.params.weight_decay", args.weight_decay)
    setp("scheduler.params.warmup_min_lr", 0)
    setp("scheduler.params.warmup_max_lr", args.learning_rate)
    setp("scheduler.params.warmup_num_steps", args.get_warmup_steps(num_training_steps))
    setp("scheduler.params.total_num_steps", num_training_steps)
    setp("train_micro_batch_size_per_gpu", args.per_device_train_batch_size)
    setp("gradient_accumulation_steps", args.gradient_accumulation_steps)
    setp("train_batch_size", args.per_device_train_batch_size * args.gradient_accumulation_steps * world_size)
    setp("gradient_clipping", args.max_grad_norm)
    setp("zero_optimization.reduce_bucket_size", hidden_size * hidden_size)
    setp("zero_optimization.stage3_prefetch_bucket_size", int(0.9 * hidden_size * hidden_size))
    setp("zero_optimization.stage3_param_persistence_threshold", 10 * hidden_size)
    left = list(_walk_auto(cfg))
    if left:
        raise ValueError(f"unresolved \"auto\" entries in the DeepSpeed config: {left}")

    z = cfg.get("zero_optimization", {})
    zc = ZeroConfig(
        stage=int(z.get("stage", 0)),
        offload_optimizer_device=(z.get("offload_optimizer") or {}).get("device", "none"),
        offload_param_device=(z.get("offload_param") or {}).get("device", "none"),
        pin_memory=bool((z.get("offload_param") or {}).get("pin_memory", (z.get("offload_optimizer") or {}).get("pin_memory", True))),
        overlap_comm=bool(z.get("overlap_comm", True)), contiguous_gradients=bool(z.get("contiguous_gradients", True)),
        sub_group_size=float(z.get("sub_group_size", 1e9)), reduce_bucket_size=int(z.get("reduce_bucket_size", 5e8)),
        stage3_prefetch_bucket_size=int(z.get("stage3_prefetch_bucket_size", 5e7)),
        stage3_param_persistence_threshold=int(z.get("stage3_param_persistence_threshold", 1e5)),
        stage3_max_live_parameters=float(z.get("stage3_max_live_parameters", 1e9)),
        stage3_max_reuse_distance=float(z.get("stage3_max_reuse_distance", 1e9)),
        stage3_gather_16bit_weights_on_model_save=bool(z.get("stage3_gather_16bit_weights_on_model_save", False)))
    opt, sch = cfg.get("optimizer") or {}, cfg.get("scheduler") or {}
    return DSConfig(raw=cfg, bf16=bool((cfg.get("bf16") or {}).get("enabled", False)), fp16=bool((cfg.get("fp16") or {}).get("enabled", False)),
                    optimizer_type=opt.get("type"), optimizer_params=dict(opt.get("params") or {}), scheduler_type=sch.get("type"),
                    scheduler_params=dict(sch.get("params") or {}), zero=zc,
                    gradient_accumulation_steps=int(cfg.get("gradient_accumulation_steps", 1)),
                    gradient_clipping=float(cfg.get("gradient_clipping", 0.0)), steps_per_print=int(cfg.get("steps_per_print", 10)),
                    train_batch_size=int(cfg.get("train_batch_size", 0)),
                    train_micro_batch_size_per_gpu=int(cfg.get("train_micro_batch_size_per_gpu", 0)),
                    wall_clock_breakdown=bool(cfg.get("wall_clock_breakdown", False)))
