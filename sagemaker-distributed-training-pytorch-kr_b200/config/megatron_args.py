"""Megatron-style argument system: flag-compatible with the reference's
3_training_megatron-lm/megatron/arguments.py (17 argument groups, :18-56 / :448-1305) and its
``validate_args`` derivations (:58-398), re-expressed as a declarative flag table plus a list of
derivation / consistency rules.

Differences on purpose:
  * ``--fp16`` works both as a bare flag and with a string boolean (``--fp16 true``, the SageMaker edit).
  * interleaved pipelining is accepted for pipeline size 2 (the reference asserts > 2).
  * features that are non-goals here (retro, biencoder, vision pre-training, TransformerEngine/fp8,
    experts) are parsed and accepted, and rejected only when switched on.
"""
from __future__ import annotations

import argparse
import os
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Sequence

import torch

T, F_ = "store_true", "store_false"


def _strbool(v):
    if isinstance(v, bool):
        return v
    return str(v).lower() in ("1", "true", "yes", "y", "t")


@dataclass
class Flag:
    names: Sequence[str]
    type: Any = None
    default: Any = None
    action: Optional[str] = None
    choices: Optional[list] = None
    dest: Optional[str] = None
    nargs: Optional[str] = None
    const: Any = None
    help: str = ""


def _f(*names, **kw) -> Flag:
    return Flag(names=names, **kw)


# group title -> flags.  Group names follow the reference's _add_*_args functions.
GROUPS: Dict[str, List[Flag]] = {
    "network size": [
        _f("--num-layers", type=int), _f("--encoder-num-layers", type=int), _f("--decoder-num-layers", type=int),
        _f("--hidden-size", type=int), _f("--ffn-hidden-size", type=int), _f("--num-attention-heads", type=int),
        _f("--kv-channels", type=int), _f("--group-query-attention", action=T), _f("--num-query-groups", type=int, default=1),
        _f("--max-position-embeddings", type=int),
        _f("--position-embedding-type", type=str, default="learned_absolute", choices=["learned_absolute", "rope"]),
        _f("--use-rotary-position-embeddings", action=T), _f("--rotary-percent", type=float, default=1.0),
        _f("--no-position-embedding", action=F_, dest="add_position_embedding"),
        _f("--make-vocab-size-divisible-by", type=int, default=128), _f("--layernorm-epsilon", type=float, default=1e-5),
        _f("--apply-layernorm-1p", action=T), _f("--apply-residual-connection-post-layernorm", action=T),
        _f("--openai-gelu", action=T), _f("--squared-relu", action=T), _f("--swiglu", action=T),
        _f("--onnx-safe", type=_strbool), _f("--bert-no-binary-head", action=F_, dest="bert_binary_head"),
        _f("--num-experts", type=int), _f("--untie-embeddings-and-output-weights", action=T),
        _f("--embedding-weights-in-fp32", action=T),
    ],
    "logging": [
        _f("--log-params-norm", action=T), _f("--log-num-zeros-in-grad", action=T), _f("--timing-log-level", type=int, default=0, choices=[0, 1, 2]),
        _f("--no-barrier-with-level-1-timing", action=F_, dest="barrier_with_L1_time"),
        _f("--timing-log-option", type=str, default="minmax", choices=["max", "minmax", "all"]),
        _f("--tensorboard-log-interval", type=int, default=1), _f("--tensorboard-queue-size", type=int, default=1000),
        _f("--log-timers-to-tensorboard", action=T), _f("--log-batch-size-to-tensorboard", action=T),
        _f("--no-log-learnig-rate-to-tensorboard", action=F_, dest="log_learning_rate_to_tensorboard"),
        _f("--no-log-loss-scale-to-tensorboard", action=F_, dest="log_loss_scale_to_tensorboard"),
        _f("--log-validation-ppl-to-tensorboard", action=T), _f("--log-memory-to-tensorboard", action=T),
        _f("--log-world-size-to-tensorboard", action=T),
    ],
    "regularization": [
        _f("--attention-dropout", type=float, default=0.1), _f("--hidden-dropout", type=float, default=0.1),
        _f("--weight-decay", type=float, default=0.01), _f("--start-weight-decay", type=float), _f("--end-weight-decay", type=float),
        _f("--weight-decay-incr-style", type=str, default="constant", choices=["constant", "linear", "cosine"]),
        _f("--clip-grad", type=float, default=1.0), _f("--adam-beta1", type=float, default=0.9), _f("--adam-beta2", type=float, default=0.999),
        _f("--adam-eps", type=float, default=1e-8), _f("--sgd-momentum", type=float, default=0.9),
    ],
    "training": [
        _f("--micro-batch-size", type=int), _f("--batch-size", type=int), _f("--global-batch-size", type=int),
        _f("--rampup-batch-size", nargs="*"), _f("--recompute-activations", action=T),
        _f("--recompute-granularity", type=str, choices=["full", "selective"]), _f("--distribute-saved-activations", action=T),
        _f("--recompute-method", type=str, choices=["uniform", "block"]), _f("--recompute-num-layers", type=int, default=1),
        _f("--profile", action=T), _f("--profile-step-start", type=int, default=10), _f("--profile-step-end", type=int, default=12),
        _f("--profile-ranks", type=int, nargs="+", default=[0]), _f("--checkpoint-activations", action=T),
        _f("--train-iters", type=int), _f("--train-samples", type=int), _f("--log-interval", type=int, default=100),
        _f("--exit-interval", type=int), _f("--exit-duration-in-mins", type=int), _f("--exit-signal-handler", action=T),
        _f("--tensorboard-dir", type=str), _f("--no-masked-softmax-fusion", action=F_, dest="masked_softmax_fusion"),
        _f("--no-bias-gelu-fusion", action=F_, dest="bias_gelu_fusion"), _f("--no-bias-dropout-fusion", action=F_, dest="bias_dropout_fusion"),
        _f("--use-flash-attn", action=T), _f("--disable-bias-linear", action=F_, dest="add_bias_linear"),
        _f("--optimizer", type=str, default="adam", choices=["adam", "sgd"]), _f("--dataloader-type", type=str, choices=["single", "cyclic"]),
        _f("--no-async-tensor-model-parallel-allreduce", action=F_, dest="async_tensor_model_parallel_allreduce"),
        _f("--no-persist-layer-norm", action=T), _f("--sequence-parallel", action=T),
        _f("--no-gradient-accumulation-fusion", action=F_, dest="gradient_accumulation_fusion"),
        _f("--cuda-graph", action=T, help="(addition) replay the micro-batch forward+backward as a CUDA graph (PP=1)"),
    ],
    "initialization": [
        _f("--seed", type=int, default=1234), _f("--data-parallel-random-init", action=T), _f("--init-method-std", type=float, default=0.02),
        _f("--init-method-xavier-uniform", action=T),
    ],
    "learning rate": [
        _f("--lr", type=float), _f("--lr-decay-style", type=str, default="linear", choices=["constant", "linear", "cosine", "inverse-square-root"]),
        _f("--lr-decay-iters", type=int), _f("--lr-decay-samples", type=int), _f("--lr-warmup-fraction", type=float),
        _f("--lr-warmup-iters", type=int, default=0), _f("--lr-warmup-samples", type=int, default=0), _f("--warmup", type=int),
        _f("--min-lr", type=float, default=0.0), _f("--override-opt_param-scheduler", action=T), _f("--use-checkpoint-opt_param-scheduler", action=T),
    ],
    "checkpointing": [
        _f("--save", type=str), _f("--save-interval", type=int), _f("--no-save-optim", action=T, default=None), _f("--no-save-rng", action=T, default=None),
        _f("--load", type=str), _f("--no-load-optim", action=T, default=None), _f("--no-load-rng", action=T, default=None), _f("--finetune", action=T),
        _f("--no-initialization", action=F_, dest="perform_initialization"), _f("--use-checkpoint-args", action=T), _f("--exit-on-missing-checkpoint", action=T),
    ],
    "mixed precision": [
        _f("--fp16", type=_strbool, nargs="?", const=True, default=False), _f("--bf16", type=_strbool, nargs="?", const=True, default=False),
        _f("--loss-scale", type=float), _f("--initial-loss-scale", type=float, default=2 ** 32), _f("--min-loss-scale", type=float, default=1.0),
        _f("--loss-scale-window", type=float, default=1000), _f("--hysteresis", type=int, default=2), _f("--fp32-residual-connection", action=T),
        _f("--no-query-key-layer-scaling", action=F_, dest="apply_query_key_layer_scaling"), _f("--attention-softmax-in-fp32", action=T),
        _f("--accumulate-allreduce-grads-in-fp32", action=T), _f("--fp16-lm-cross-entropy", action=T),
    ],
    "distributed": [
        _f("--tensor-model-parallel-size", type=int, default=1), _f("--pipeline-model-parallel-size", type=int, default=1),
        _f("--pipeline-model-parallel-split-rank", type=int), _f("--model-parallel-size", type=int), _f("--num-layers-per-virtual-pipeline-stage", type=int),
        _f("--overlap-p2p-communication", action=T, dest="overlap_p2p_comm"),
        _f("--distributed-backend", default="nccl", choices=["nccl", "gloo"]), _f("--distributed-timeout-minutes", type=int, default=10),
        _f("--DDP-impl", default="local", choices=["local", "torch"]),
        _f("--no-contiguous-buffers-in-local-ddp", action=F_, dest="use_contiguous_buffers_in_local_ddp"),
        _f("--no-scatter-gather-tensors-in-pipeline", action=F_, dest="scatter_gather_tensors_in_pipeline"),
        _f("--use-ring-exchange-p2p", action=T), _f("--local_rank", type=int), _f("--lazy-mpu-init", type=_strbool),
        _f("--use-cpu-initialization", action=T, default=None), _f("--empty-unused-memory-level", type=int, default=0, choices=[0, 1, 2]),
        _f("--standalone-embedding-stage", action=T), _f("--use-distributed-optimizer", action=T),
        # additions: communication path of the tensor/data-parallel collectives
        _f("--comm-backend", type=str, default="native", choices=["native", "nccl"],
           help="native = fused NVLink/NVSwitch kernels (ag_gemm, gemm_rs, fused bucket all-reduce); nccl = baseline semantics"),
    ],
    "validation": [
        _f("--eval-iters", type=int, default=100), _f("--eval-interval", type=int, default=1000), _f("--skip-train", action=T),
    ],
    "data and dataloader": [
        _f("--data-path", nargs="*"), _f("--split", type=str, default="969, 30, 1"), _f("--train-data-path", nargs="*"),
        _f("--valid-data-path", nargs="*"), _f("--test-data-path", nargs="*"), _f("--data-cache-path"),
        _f("--vocab-size", type=int), _f("--vocab-file", type=str), _f("--merge-file", type=str), _f("--vocab-extra-ids", type=int, default=0),
        _f("--seq-length", type=int), _f("--encoder-seq-length", type=int), _f("--decoder-seq-length", type=int),
        _f("--retriever-seq-length", type=int, default=256), _f("--sample-rate", type=float, default=1.0), _f("--mask-prob", type=float, default=0.15),
        _f("--short-seq-prob", type=float, default=0.1), _f("--mmap-warmup", action=T), _f("--num-workers", type=int, default=2),
        _f("--tokenizer-type", type=str, choices=["BertWordPieceLowerCase", "BertWordPieceCase", "GPT2BPETokenizer", "SentencePieceTokenizer",
                                                  "GPTSentencePieceTokenizer", "NullTokenizer"]),
        _f("--tokenizer-model", type=str), _f("--data-impl", type=str, default="infer", choices=["lazy", "cached", "mmap", "infer"]),
        _f("--reset-position-ids", action=T), _f("--reset-attention-mask", action=T), _f("--eod-mask-loss", action=T),
        _f("--mock-data", action=T, help="(addition) synthetic token stream of the configured shape"),
        _f("--attention-impl", type=str, default="flash", choices=["flash", "unfused"],
           help="(addition) flash = tcgen05 flash attention (default, what --use-flash-attn selects); unfused = the reference's CoreAttention: "
                "batched GEMM, scale+mask+softmax kernel, dropout, batched GEMM"),
        _f("--metrics-jsonl", type=str, help="(addition) append one JSON record per log point: loss, lr, step ms, tokens/s, MFU"),
        _f("--inject-fault", type=str, help="(addition) RANK:ITERATION - raise on that rank after that iteration (recovery tests)"),
    ],
    "autoresume": [_f("--adlr-autoresume", action=T), _f("--adlr-autoresume-interval", type=int, default=1000)],
    "biencoder": [
        _f("--ict-head-size", type=int), _f("--biencoder-projection-dim", type=int, default=0), _f("--biencoder-shared-query-context-model", action=T),
        _f("--ict-load", type=str), _f("--bert-load", type=str), _f("--titles-data-path", type=str), _f("--query-in-block-prob", type=float, default=0.1),
        _f("--use-one-sent-docs", action=T), _f("--evidence-data-path", type=str), _f("--retriever-report-topk-accuracies", type=int, nargs="+", default=[]),
        _f("--retriever-score-scaling", action=T), _f("--block-data-path", type=str), _f("--embedding-path", type=str),
        _f("--indexer-batch-size", type=int, default=128), _f("--indexer-log-interval", type=int, default=1000),
    ],
    "vision": [
        _f("--num-classes", type=int, default=1000), _f("--img-h", type=int, default=224), _f("--img-w", type=int, default=224),
        _f("--num-channels", type=int, default=3), _f("--patch-dim", type=int, default=16), _f("--classes-fraction", type=float, default=1.0),
        _f("--data-per-class-fraction", type=float, default=1.0), _f("--no-data-sharding", action=F_, dest="data_sharding"),
        _f("--head-lr-mult", type=float, default=1.0), _f("--vision-pretraining", action=T),
        _f("--vision-pretraining-type", type=str, default="classify", choices=["classify", "inpaint", "dino"]),
        _f("--vision-backbone-type", type=str, default="vit", choices=["vit", "mit", "swin"]),
        _f("--swin-backbone-type", type=str, default="tiny", choices=["tiny", "base", "h3"]),
        _f("--mask-type", type=str, default="random", choices=["random", "row"]), _f("--mask-factor", type=float, default=1.0),
        _f("--iter-per-epoch", type=int, default=1250), _f("--dino-local-img-size", type=int, default=96),
        _f("--dino-local-crops-number", type=int, default=10), _f("--dino-head-hidden-size", type=int, default=2048),
        _f("--dino-bottleneck-size", type=int, default=256), _f("--dino-freeze-last-layer", type=float, default=1),
        _f("--dino-norm-last-layer", action=T), _f("--dino-warmup-teacher-temp", type=float, default=0.04),
        _f("--dino-teacher-temp", type=float, default=0.07), _f("--dino-warmup-teacher-temp-epochs", type=int, default=30),
    ],
    "inference": [
        _f("--inference-batch-times-seqlen-threshold", type=int, default=512), _f("--max-tokens-to-oom", type=int, default=12000),
        _f("--output-bert-embeddings", action=T), _f("--bert-embedder-type", default="megatron", choices=["megatron", "huggingface"]),
    ],
    "transformer-engine": [
        _f("--fp8-e4m3", action=T), _f("--fp8-hybrid", action=T), _f("--no-fp8-wgrad", action=F_, dest="fp8_wgrad"),
        _f("--fp8-margin", type=int, default=0), _f("--fp8-interval", type=int, default=1),
        _f("--transformer-impl", default="local", choices=["local", "transformer_engine"]),
        _f("--fp8-amax-history-len", type=int, default=1), _f("--fp8-amax-compute-algo", default="most_recent", choices=["most_recent", "max"]),
    ],
    "retro": [
        _f("--retro-workdir"), _f("--retro-add-retriever", action=T), _f("--retro-cyclic-train-iters", type=int),
        _f("--retro-encoder-layers", type=int, default=2), _f("--retro-encoder-hidden-dropout", type=float, default=0.1),
        _f("--retro-encoder-attention-dropout", type=float, default=0.1), _f("--retro-num-neighbors", type=int, default=2),
        _f("--retro-num-retrieved-chunks", type=int, default=2), _f("--retro-return-doc-ids", action=T),
    ],
}


def build_parser(extra_args_provider: Optional[Callable] = None, ignore_unknown_args: bool = False) -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="b200dist Megatron-style arguments", allow_abbrev=False)
    for title, flags in GROUPS.items():
        g = parser.add_argument_group(title=title)
        for fl in flags:
            kw: Dict[str, Any] = {}
            if fl.action:
                kw["action"] = fl.action
                if fl.default is not None or fl.action == T and fl.names[0] in ("--no-save-optim", "--no-save-rng", "--no-load-optim", "--no-load-rng", "--use-cpu-initialization"):
                    kw["default"] = fl.default
            else:
                if fl.type is not None:
                    kw["type"] = fl.type
                kw["default"] = fl.default
                if fl.choices is not None:
                    kw["choices"] = fl.choices
                if fl.nargs is not None:
                    kw["nargs"] = fl.nargs
                if fl.const is not None:
                    kw["const"] = fl.const
            if fl.dest:
                kw["dest"] = fl.dest
            if fl.help:
                kw["help"] = fl.help
            g.add_argument(*fl.names, **kw)
    if extra_args_provider is not None:
        parser = extra_args_provider(parser)
    return parser


def all_flag_names() -> List[str]:
    return [n for flags in GROUPS.values() for fl in flags for n in fl.names]


def parse_args(extra_args_provider=None, ignore_unknown_args=False, argv=None):
    """reference: arguments.py:18-56 (also reads RANK / WORLD_SIZE from the environment, :53-54)."""
    parser = build_parser(extra_args_provider)
    if ignore_unknown_args:
        args, _ = parser.parse_known_args(argv)
    else:
        args = parser.parse_args(argv)
    args.rank = int(os.getenv("RANK", "0"))
    args.world_size = int(os.getenv("WORLD_SIZE", "1"))
    return args


class ArgumentError(ValueError):
    pass


def _require(cond: bool, msg: str):
    if not cond:
        raise ArgumentError(msg)


def validate_args(args, defaults: Optional[dict] = None, verbose: bool = True):
    """Derived fields + consistency checks (reference: arguments.py:58-398)."""
    defaults = defaults or {}
    log = (lambda *a: print(*a, flush=True)) if (verbose and args.rank == 0) else (lambda *a: None)

    # ---- parallel sizes
    args.tensor_model_parallel_size = min(args.tensor_model_parallel_size, args.world_size)
    _require(args.world_size % args.tensor_model_parallel_size == 0,
             f"world size ({args.world_size}) is not divisible by tensor model parallel size ({args.tensor_model_parallel_size})")
    args.pipeline_model_parallel_size = min(args.pipeline_model_parallel_size, args.world_size // args.tensor_model_parallel_size)
    args.transformer_pipeline_model_parallel_size = args.pipeline_model_parallel_size - (1 if args.standalone_embedding_stage else 0)
    mp = args.pipeline_model_parallel_size * args.tensor_model_parallel_size
    _require(not getattr(args, "reset_attention_mask", False) or getattr(args, "attention_impl", "flash") == "unfused",
             "--reset-attention-mask (attention within a document only) needs --attention-impl unfused: the flash kernel is causal over "
             "the whole sequence, the scale/mask/softmax kernel of the unfused path takes the document mask")
    _require(args.world_size % mp == 0, f"world size ({args.world_size}) is not divisible by TP ({args.tensor_model_parallel_size}) x PP ({args.pipeline_model_parallel_size})")
    args.data_parallel_size = args.world_size // mp
    log(f"using world size: {args.world_size}, data-parallel-size: {args.data_parallel_size}, tensor-model-parallel size: "
        f"{args.tensor_model_parallel_size}, pipeline-model-parallel size: {args.pipeline_model_parallel_size} ")
    if args.pipeline_model_parallel_size > 1 and args.pipeline_model_parallel_split_rank is not None:
        _require(args.pipeline_model_parallel_split_rank < args.pipeline_model_parallel_size, "split rank needs to be less than pipeline model parallel size")

    # ---- retired flags
    for dead, hint in (("batch_size", "--micro-batch-size"), ("warmup", "--lr-warmup-fraction"), ("model_parallel_size", "--tensor-model-parallel-size")):
        _require(getattr(args, dead, None) is None, f"--{dead.replace('_', '-')} is no longer valid, use {hint} instead")
        if hasattr(args, dead):
            delattr(args, dead)
    _require(not args.checkpoint_activations, "--checkpoint-activations is no longer valid, use --recompute-activations")
    del args.checkpoint_activations
    if args.recompute_activations:
        args.recompute_granularity = "selective"
    del args.recompute_activations

    # ---- injected defaults (pretrain_gpt.py:149 passes tokenizer_type)
    for k, v in defaults.items():
        if getattr(args, k, None) is not None:
            log(f"WARNING: overriding default arguments for {k}:{v} with {k}:{getattr(args, k)}")
        else:
            setattr(args, k, v)

    # ---- batch sizes
    _require(args.micro_batch_size is not None and args.micro_batch_size > 0, "--micro-batch-size must be given and positive")
    if args.global_batch_size is None:
        args.global_batch_size = args.micro_batch_size * args.data_parallel_size
        log(f"setting global batch size to {args.global_batch_size}")
    _require(args.global_batch_size > 0, "global batch size must be positive")
    _require(args.global_batch_size % (args.micro_batch_size * args.data_parallel_size) == 0,
             f"global batch size ({args.global_batch_size}) is not divisible by micro batch size ({args.micro_batch_size}) "
             f"times data parallel size ({args.data_parallel_size})")

    # ---- layers (needed by the virtual-pipeline derivation)
    if args.num_layers is not None:
        _require(args.encoder_num_layers is None, "cannot have both num-layers and encoder-num-layers specified")
        args.encoder_num_layers = args.num_layers
    else:
        _require(args.encoder_num_layers is not None, "either num-layers or encoder-num-layers should be specified")
        args.num_layers = args.encoder_num_layers

    # ---- interleaved schedule: PP >= 2 is enough here (the reference demands > 2)
    if args.num_layers_per_virtual_pipeline_stage is not None:
        _require(args.pipeline_model_parallel_size >= 2, "interleaved schedule needs pipeline-model-parallel size >= 2")
        _require(args.num_layers % args.num_layers_per_virtual_pipeline_stage == 0,
                 "number of layers is not divisible by number of layers per virtual pipeline stage")
        per_stage = args.num_layers // args.transformer_pipeline_model_parallel_size
        _require(per_stage % args.num_layers_per_virtual_pipeline_stage == 0, "layers per pipeline stage not divisible by layers per virtual stage")
        args.virtual_pipeline_model_parallel_size = per_stage // args.num_layers_per_virtual_pipeline_stage
        if args.virtual_pipeline_model_parallel_size <= 1:
            args.virtual_pipeline_model_parallel_size = None
    else:
        args.virtual_pipeline_model_parallel_size = None

    # ---- precision
    args.params_dtype = torch.float
    if args.fp16:
        _require(not args.bf16, "--fp16 and --bf16 are exclusive")
        args.params_dtype = torch.half
    if args.bf16:
        args.params_dtype = torch.bfloat16
        if not args.accumulate_allreduce_grads_in_fp32:
            args.accumulate_allreduce_grads_in_fp32 = True
            log("accumulate and all-reduce gradients in fp32 for bfloat16 data type.")
    log(f"using {args.params_dtype} for parameters ...")
    if args.accumulate_allreduce_grads_in_fp32 or args.use_distributed_optimizer:
        _require(args.DDP_impl == "local" and args.use_contiguous_buffers_in_local_ddp, "fp32 grad accumulation / distributed optimizer need the local DDP with contiguous buffers")
    if args.DDP_impl == "torch":
        args.use_contiguous_buffers_in_local_ddp = False
    if args.dataloader_type is None:
        args.dataloader_type = "single"
    args.consumed_train_samples = 0
    args.consumed_valid_samples = 0
    args.variable_seq_lengths = False

    # ---- iteration- vs sample-based training
    if args.train_iters:
        _require(args.train_samples is None, "expected iteration-based training")
        _require(args.lr_decay_samples is None, "expected iteration-based learning rate decay")
        _require(args.lr_warmup_samples == 0, "expected iteration-based learning rate warmup")
        _require(args.rampup_batch_size is None, "expected no batch-size rampup for iteration-based training")
        if args.lr_warmup_fraction is not None:
            _require(args.lr_warmup_iters == 0, "can only specify one of lr-warmup-fraction and lr-warmup-iters")
    if args.train_samples:
        _require(args.train_iters is None, "expected sample-based training")
        _require(args.lr_decay_iters is None, "expected sample-based learning rate decay")
        _require(args.lr_warmup_iters == 0, "expected sample-based learning rate warmup")
        if args.lr_warmup_fraction is not None:
            _require(args.lr_warmup_samples == 0, "can only specify one of lr-warmup-fraction and lr-warmup-samples")

    # ---- model shape
    for req in ("num_layers", "hidden_size", "num_attention_heads", "max_position_embeddings"):
        _require(getattr(args, req) is not None, f"{req} argument is None")
    if args.ffn_hidden_size is None:
        args.ffn_hidden_size = 4 * args.hidden_size
        if args.swiglu:
            args.ffn_hidden_size = int((4 * args.hidden_size * 2 / 3) / 64) * 64
    if args.kv_channels is None:
        _require(args.hidden_size % args.num_attention_heads == 0, "hidden size must be divisible by the number of heads")
        args.kv_channels = args.hidden_size // args.num_attention_heads
    _require(args.num_attention_heads % args.tensor_model_parallel_size == 0,
             f"num attention heads ({args.num_attention_heads}) not divisible by tensor parallel size ({args.tensor_model_parallel_size})")
    if args.seq_length is not None:
        _require(args.encoder_seq_length is None, "cannot give both seq-length and encoder-seq-length")
        args.encoder_seq_length = args.seq_length
    else:
        _require(args.encoder_seq_length is not None, "seq-length is required")
        args.seq_length = args.encoder_seq_length
    _require(args.max_position_embeddings >= args.seq_length, "max-position-embeddings must be >= seq-length")
    if args.decoder_seq_length is not None:
        _require(args.max_position_embeddings >= args.decoder_seq_length, "max-position-embeddings must be >= decoder-seq-length")
    if args.lr is not None:
        _require(args.min_lr <= args.lr, "min-lr must be <= lr")
    if args.save is not None:
        _require(args.save_interval is not None, "--save needs --save-interval")
    if args.fp16_lm_cross_entropy:
        _require(args.fp16, "lm cross entropy in fp16 only support in fp16 mode.")
    if args.fp32_residual_connection:
        _require(args.fp16 or args.bf16, "residual connection in fp32 only supported when using fp16 or bf16.")

    # ---- weight decay schedule
    if args.weight_decay_incr_style == "constant":
        _require(args.start_weight_decay is None and args.end_weight_decay is None, "constant weight decay takes no start/end values")
        args.start_weight_decay = args.end_weight_decay = args.weight_decay
    else:
        _require(args.start_weight_decay is not None and args.end_weight_decay is not None, "weight decay ramp needs start and end values")

    # ---- recompute
    if args.distribute_saved_activations:
        _require(args.tensor_model_parallel_size > 1, "can distribute recomputed activations only across tensor model parallel groups")
        _require(args.recompute_granularity == "full", "distributed recompute activations is only applicable to full recompute granularity")
        _require(args.recompute_method is not None, "distributed recompute needs a recompute method")
    if args.recompute_granularity == "selective":
        _require(args.recompute_method is None, "recompute method is not supported for selective recomputing granularity")

    # ---- sequence parallel / async TP all-reduce interplay
    if args.tensor_model_parallel_size == 1:
        args.sequence_parallel = False
    if args.sequence_parallel:
        args.async_tensor_model_parallel_allreduce = False
    # The reference requires CUDA_DEVICE_MAX_CONNECTIONS=1 so the async collective is queued before the
    # wgrad GEMM (arguments.py:347-355).  The fused in-kernel collectives do not depend on launch order,
    # so the variable is only enforced for the nccl baseline path.
    if args.comm_backend == "nccl" and torch.cuda.is_available() and os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS") != "1":
        if args.sequence_parallel or (args.async_tensor_model_parallel_allreduce and args.tensor_model_parallel_size > 1):
            raise ArgumentError("--comm-backend nccl with sequence parallelism / async TP all-reduce requires CUDA_DEVICE_MAX_CONNECTIONS=1")
    if not args.add_bias_linear:
        args.bias_gelu_fusion = False

    # ---- non-goals: accepted on the command line, rejected when enabled
    _require(not (args.fp8_e4m3 or args.fp8_hybrid) and args.transformer_impl == "local", "TransformerEngine / fp8 paths are not part of this framework")
    _require(not args.retro_add_retriever, "retro is not part of this framework")
    if args.num_experts is not None:
        _require(args.num_experts > 1, "--num-experts must be greater than 1")
    if args.group_query_attention:
        _require(args.num_attention_heads % args.num_query_groups == 0, "num-attention-heads must be a multiple of num-query-groups")
        _require(args.num_query_groups % args.tensor_model_parallel_size == 0, "num-query-groups must be divisible by the tensor parallel size")
    _require(sum(bool(x) for x in (args.swiglu, args.squared_relu, args.openai_gelu)) <= 1, "choose one of --swiglu, --squared-relu, --openai-gelu")
    if args.use_rotary_position_embeddings:
        args.position_embedding_type = "rope"
    if not args.add_position_embedding and args.position_embedding_type != "rope":
        raise ArgumentError("--no-position-embedding is deprecated, use --position-embedding-type")
    # ---- flags that parse but do not change what runs here: say so once instead of silently ignoring them
    inert = []
    if args.fp32_residual_connection:
        inert.append("--fp32-residual-connection (the residual stream stays in the parameter dtype; LayerNorm / softmax / loss accumulate in fp32)")
    if not args.apply_query_key_layer_scaling:
        inert.append("--no-query-key-layer-scaling (scores are scaled by 1/sqrt(d) inside the attention kernel with an fp32 softmax: the layer-number trick is unnecessary)")
    if getattr(args, "overlap_p2p_comm", False):
        inert.append("--overlap-p2p-communication (pipeline p2p is one fused send+recv group per schedule boundary; B2_PP_P2P=links posts receives ahead)")
    if not getattr(args, "scatter_gather_tensors_in_pipeline", True):
        inert.append("--no-scatter-gather-tensors-in-pipeline (stage-to-stage tensors are sent whole; under sequence parallelism they already are 1/t)")
    if getattr(args, "use_ring_exchange_p2p", False):
        inert.append("--use-ring-exchange-p2p")
    if not getattr(args, "bias_dropout_fusion", True):
        inert.append("--no-bias-dropout-fusion (bias + dropout + residual is always one kernel or a GEMM epilogue)")
    if getattr(args, "attention_impl", "flash") == "flash" and not args.masked_softmax_fusion:
        inert.append("--no-masked-softmax-fusion (only affects --attention-impl unfused)")
    args.inert_flags = inert
    if inert and args.rank == 0:
        print("> accepted for CLI compatibility, no effect in this runtime:\n    " + "\n    ".join(inert), flush=True)
    if verbose:
        print_args("arguments", args)
    return args


def print_args(title, args):
    """Sorted `name ....... value` dump (reference: arguments.py:401-413)."""
    if args.rank != 0:
        return
    print(f"------------------------ {title} ------------------------", flush=True)
    lines = []
    for k in vars(args):
        dots = "." * (48 - len(k))
        lines.append(f"  {k} {dots} {getattr(args, k)}")
    for ln in sorted(lines, key=lambda x: x.lower()):
        print(ln, flush=True)
    print(f"-------------------- end of {title} ---------------------", flush=True)


@dataclass
class TransformerConfig:
    """Model-facing subset of the arguments (reference: core_transformer_config_from_args, arguments.py:419-446)."""
    num_layers: int = 12
    hidden_size: int = 768
    num_attention_heads: int = 12
    ffn_hidden_size: int = 3072
    kv_channels: int = 64
    seq_length: int = 1024
    max_position_embeddings: int = 1024
    padded_vocab_size: int = 32768
    hidden_dropout: float = 0.1
    attention_dropout: float = 0.1
    layernorm_epsilon: float = 1e-5
    init_method_std: float = 0.02
    init_method_xavier_uniform: bool = False
    params_dtype: torch.dtype = torch.float32
    add_bias_linear: bool = True
    swiglu: bool = False
    squared_relu: bool = False
    openai_gelu: bool = False
    bias_gelu_fusion: bool = True
    num_query_groups: Optional[int] = None          # None / == num_attention_heads: multi-head attention
    apply_layernorm_1p: bool = False
    apply_residual_connection_post_layernorm: bool = False
    num_experts: Optional[int] = None
    rotary_percent: float = 1.0
    rotary_seq_len_interpolation_factor: Optional[int] = None
    position_embedding_type: str = "learned_absolute"
    attention_impl: str = "flash"                   # "unfused": scores -> scale/mask/softmax kernel -> context (CoreAttention)
    masked_softmax_fusion: bool = True
    untie_embeddings_and_output_weights: bool = False
    apply_query_key_layer_scaling: bool = True
    tensor_model_parallel_size: int = 1
    pipeline_model_parallel_size: int = 1
    virtual_pipeline_model_parallel_size: Optional[int] = None
    sequence_parallel: bool = False
    gradient_accumulation_fusion: bool = True
    recompute_granularity: Optional[str] = None
    recompute_method: Optional[str] = None
    recompute_num_layers: int = 1
    distribute_saved_activations: bool = False
    fp16: bool = False
    bf16: bool = False
    fp32_residual_connection: bool = False
    comm_backend: str = "native"
    pipeline_dtype: torch.dtype = torch.float32
    batch_p2p_comm: bool = True
    deallocate_pipeline_outputs: bool = True
    use_cpu_initialization: bool = False
    perform_initialization: bool = True


def core_transformer_config_from_args(args) -> TransformerConfig:
    kw = {}
    for f in TransformerConfig.__dataclass_fields__:
        if hasattr(args, f) and getattr(args, f) is not None:
            kw[f] = getattr(args, f)
    cfg = TransformerConfig(**kw)
    if not getattr(args, "group_query_attention", False):
        cfg.num_query_groups = None                 # the flag, not the group count, switches GQA on (arguments.py:566-567)
    cfg.pipeline_dtype = args.params_dtype
    cfg.batch_p2p_comm = not args.overlap_p2p_comm
    cfg.deallocate_pipeline_outputs = True
    cfg.use_cpu_initialization = bool(args.use_cpu_initialization)
    cfg.standalone_embedding_stage = bool(getattr(args, "standalone_embedding_stage", False))      # read by models.gpt.get_layer_range
    return cfg


def pad_vocab_size(orig_vocab_size: int, divisible_by: int, tp: int, verbose: bool = False) -> int:
    """Pad so each tensor-parallel shard is a multiple of `divisible_by` (log line: 3_...ipynb:1212)."""
    mult = divisible_by * tp
    after = (orig_vocab_size + mult - 1) // mult * mult
    if verbose:
        print(f" > padded vocab (size: {orig_vocab_size}) with {after - orig_vocab_size} dummy tokens (new size: {after})", flush=True)
    return after
