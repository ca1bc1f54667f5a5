"""HF-style dataclass argument system for the Alpaca script (reference: 4_training_alpaca_deepspeed/train.py:72-89,
210-211 - `HfArgumentParser((ModelArguments, DataArguments, TrainingArguments))`).

Values arrive as `--flag True|False|value` strings because the SageMaker toolkit turns the hyper-parameter dict
into argv (4_...ipynb:877), so booleans accept both a bare flag and a string.
"""
from __future__ import annotations

import argparse
import dataclasses
import math
from dataclasses import dataclass, field, fields
from typing import Any, Optional, Tuple, get_args, get_origin, Union


def _strbool(v):
    if isinstance(v, bool):
        return v
    s = str(v).lower()
    if s in ("yes", "true", "t", "y", "1"):
        return True
    if s in ("no", "false", "f", "n", "0"):
        return False
    raise argparse.ArgumentTypeError(f"Truthy value expected: got {v}")


@dataclass
class ModelArguments:
    model_name_or_path: Optional[str] = field(default="facebook/opt-125m")


@dataclass
class DataArguments:
    data_path: Optional[str] = field(default=None, metadata={"help": "Path to the training data."})


@dataclass
class TrainingArguments:
    """The subset of transformers.TrainingArguments the notebook passes or that feeds the DeepSpeed "auto" fields
    (4_...ipynb:464-485), plus the three fields train.py adds (cache_dir, optim, model_max_length)."""
    output_dir: str = "./output"
    cache_dir: Optional[str] = None
    optim: str = "adamw_torch"
    model_max_length: int = field(default=512, metadata={"help": "Maximum sequence length. Sequences will be right padded (and possibly truncated)."})
    bf16: bool = False
    fp16: bool = False
    tf32: Optional[bool] = None
    num_train_epochs: float = 3.0
    max_steps: int = -1
    per_device_train_batch_size: int = 8
    per_device_eval_batch_size: int = 8
    gradient_accumulation_steps: int = 1
    evaluation_strategy: str = "no"
    eval_steps: Optional[int] = None
    resume_from_checkpoint: Optional[str] = None
    save_strategy: str = "steps"
    save_steps: int = 500
    save_total_limit: Optional[int] = None
    learning_rate: float = 5e-5
    weight_decay: float = 0.0
    adam_beta1: float = 0.9
    adam_beta2: float = 0.999
    adam_epsilon: float = 1e-8
    max_grad_norm: float = 1.0
    warmup_ratio: float = 0.0
    warmup_steps: int = 0
    lr_scheduler_type: str = "linear"
    logging_steps: int = 500
    seed: int = 42
    deepspeed: Optional[str] = None
    report_to: str = "none"
    dataloader_num_workers: int = 0
    local_rank: int = -1
    gradient_checkpointing: bool = False
    # additions
    comm_backend: str = field(default="native", metadata={"help": "native (fabric kernels) | nccl"})
    synthetic_data: bool = field(default=False, metadata={"help": "fixed-length synthetic instruction batches (BASELINE config 5)"})
    synthetic_samples: int = 4096

    world_size: int = 1

    def get_warmup_steps(self, num_training_steps: int) -> int:
        return self.warmup_steps if self.warmup_steps > 0 else math.ceil(num_training_steps * self.warmup_ratio)

    @property
    def train_batch_size(self) -> int:
        return self.per_device_train_batch_size


def _base_type(tp):
    if get_origin(tp) is Union:
        args = [a for a in get_args(tp) if a is not type(None)]
        return args[0] if args else str
    return tp


# transformers.TrainingArguments fields (4.28 .. 5.x) that this trainer does not act on: accepted on the command line with a
# one-line notice, so launch lines written for the reference keep working.  Fields the trainer honours are
# dataclass fields above; anything in neither set is an error, as in transformers.
HF_ACCEPTED_IGNORED = frozenset("""
overwrite_output_dir do_train do_eval do_predict prediction_loss_only per_gpu_train_batch_size per_gpu_eval_batch_size eval_accumulation_steps
eval_delay torch_empty_cache_steps lr_scheduler_kwargs log_level log_level_replica log_on_each_node logging_dir logging_strategy
logging_first_step logging_nan_inf_filter save_safetensors save_on_each_node save_only_model restore_callback_states_from_checkpoint no_cuda
use_cpu use_mps_device data_seed jit_mode_eval use_ipex bf16_full_eval fp16_full_eval fp16_opt_level half_precision_backend ddp_backend
tpu_num_cores tpu_metrics_debug debug dataloader_drop_last dataloader_prefetch_factor dataloader_pin_memory dataloader_persistent_workers past_index
run_name disable_tqdm remove_unused_columns label_names load_best_model_at_end metric_for_best_model greater_is_better ignore_data_skip sharded_ddp
fsdp fsdp_min_num_params fsdp_config fsdp_transformer_layer_cls_to_wrap accelerator_config parallelism_config label_smoothing_factor optim_args
adafactor group_by_length length_column_name ddp_find_unused_parameters ddp_bucket_cap_mb ddp_broadcast_buffers ddp_timeout skip_memory_metrics
use_legacy_prediction_loop push_to_hub hub_model_id hub_strategy hub_token hub_private_repo hub_always_push hub_revision gradient_checkpointing_kwargs
include_inputs_for_metrics include_for_metrics eval_do_concat_batches eval_use_gather_object batch_eval_metrics eval_on_start auto_find_batch_size
full_determinism torchdynamo ray_scope torch_compile torch_compile_backend torch_compile_mode dispatch_batches split_batches include_tokens_per_second
include_num_input_tokens_seen neftune_noise_alpha optim_target_modules use_liger_kernel liger_kernel_config use_cache average_tokens_across_devices
project trackio_space_id enable_jit_checkpoint train_sampling_strategy mp_parameters xpu_backend push_to_hub_model_id push_to_hub_organization
push_to_hub_token
""".split())
HF_ALIASES = {"eval_strategy": "evaluation_strategy"}          # renamed upstream in 4.41


def _strip_accepted(remaining):
    """drop `--known_hf_flag [value]` pairs from the unparsed tokens; returns (still unknown, names that were ignored)"""
    out, ignored, i = [], [], 0
    while i < len(remaining):
        tok = remaining[i]
        name = tok[2:].split("=", 1)[0] if tok.startswith("--") else None
        if name in HF_ACCEPTED_IGNORED:
            ignored.append(name)
            i += 1
            if "=" not in tok and i < len(remaining) and not remaining[i].startswith("--"):
                i += 1
            continue
        out.append(tok)
        i += 1
    return out, ignored


class HfArgumentParser(argparse.ArgumentParser):
    """dataclass types -> argparse; parse_args_into_dataclasses() returns one instance per type."""

    def __init__(self, dataclass_types, **kw):
        super().__init__(formatter_class=argparse.ArgumentDefaultsHelpFormatter, **kw)
        self.dataclass_types = list(dataclass_types) if isinstance(dataclass_types, (list, tuple)) else [dataclass_types]
        for dt in self.dataclass_types:
            for f in fields(dt):
                if not f.init:
                    continue
                tp = _base_type(f.type if not isinstance(f.type, str) else eval(f.type, vars(__import__("typing")) | {"Optional": Optional}))
                default = f.default if f.default is not dataclasses.MISSING else None
                kwargs: dict = {"help": f.metadata.get("help", "")}
                if tp is bool:
                    kwargs.update(type=_strbool, nargs="?", const=True, default=default)
                else:
                    kwargs.update(type=tp, default=default)
                self.add_argument("--" + f.name, **kwargs)

    def parse_args_into_dataclasses(self, args=None, return_remaining_strings=False) -> Tuple[Any, ...]:
        import sys as _sys
        args = list(_sys.argv[1:] if args is None else args)
        for i, tok in enumerate(args):                       # renamed upstream flags
            if tok.startswith("--"):
                name, eq, val = tok[2:].partition("=")
                if name in HF_ALIASES:
                    args[i] = "--" + HF_ALIASES[name] + (eq + val if eq else "")
        ns, remaining = self.parse_known_args(args)
        remaining, ignored = _strip_accepted(remaining)
        if ignored:
            print("[args] accepted for transformers.TrainingArguments compatibility, not used by this trainer: " + ", ".join(sorted(set(ignored))),
                  flush=True)
        outs = []
        for dt in self.dataclass_types:
            keys = {f.name for f in fields(dt) if f.init}
            outs.append(dt(**{k: v for k, v in vars(ns).items() if k in keys}))
        if return_remaining_strings:
            return (*outs, remaining)
        if remaining:
            raise ValueError(f"Some specified arguments are not used by the HfArgumentParser: {remaining}")
        return tuple(outs)
