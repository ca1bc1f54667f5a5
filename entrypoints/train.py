"""Alpaca supervised fine-tuning with ZeRO-3 - same CLI as the reference's 4_training_alpaca_deepspeed/train.py
(ModelArguments / DataArguments / TrainingArguments via HfArgumentParser :72-89,210-211; special-token handling
:226-240; Trainer(...).train(); save_state(); save_model() :243-246), on the b200dist runtime.

    torchrun --nproc-per-node 8 entrypoints/train.py --model_name_or_path llama-7b --bf16 True --output_dir out \
        --num_train_epochs 1 --per_device_train_batch_size 4 --gradient_accumulation_steps 8 --learning_rate 2e-5 \
        --warmup_ratio 0.03 --deepspeed configs/zero3_b200_resident.json --tf32 True --synthetic_data True
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

# SageMaker / mpirun shim (reference train.py:21-25)
from b200dist.launch import env as _env
_env.export_torch_env()

import torch

from b200dist.config.hf_args import DataArguments, HfArgumentParser, ModelArguments, TrainingArguments
from b200dist.data import sft
from b200dist.engine.hf_trainer import Trainer
from b200dist.models.llama import build_causal_lm


def load_tokenizer(name_or_path, training_args):
    """HF tokenizer when the files are local (no network here), else the dependency-free byte tokenizer."""
    if name_or_path and os.path.isdir(name_or_path):
        try:
            import transformers
            return transformers.AutoTokenizer.from_pretrained(name_or_path, cache_dir=training_args.cache_dir,
                                                              model_max_length=training_args.model_max_length, padding_side="right", use_fast=False)
        except Exception as e:
            print(f"[train] could not load a HF tokenizer from {name_or_path} ({type(e).__name__}); using the byte tokenizer")
    return sft.ByteTokenizer(model_max_length=training_args.model_max_length)


def train(argv=None):
    parser = HfArgumentParser((ModelArguments, DataArguments, TrainingArguments))
    model_args, data_args, training_args = parser.parse_args_into_dataclasses(argv)
    info = _env.init_distributed("nccl")
    dev = _env.device_for(info)
    if training_args.tf32 and dev.type == "cuda":
        torch.backends.cuda.matmul.allow_tf32 = True
    dtype = torch.bfloat16 if (training_args.bf16 and dev.type == "cuda") else torch.float32
    torch.manual_seed(training_args.seed)
    model = build_causal_lm(model_args.model_name_or_path, device=dev, dtype=dtype)
    if training_args.model_max_length > model.config.max_position_embeddings:
        print(f"[train] model_max_length {training_args.model_max_length} > max_position_embeddings "
              f"{model.config.max_position_embeddings}: clamping")
        training_args.model_max_length = model.config.max_position_embeddings
    tokenizer = load_tokenizer(model_args.model_name_or_path, training_args)
    special = {}
    if tokenizer.pad_token is None:
        special["pad_token"] = sft.DEFAULT_PAD_TOKEN
    if tokenizer.eos_token is None:
        special["eos_token"] = sft.DEFAULT_EOS_TOKEN
    if tokenizer.bos_token is None:
        special["bos_token"] = sft.DEFAULT_BOS_TOKEN
    if tokenizer.unk_token is None:
        special["unk_token"] = sft.DEFAULT_UNK_TOKEN
    if isinstance(tokenizer, sft.ByteTokenizer):
        tokenizer.add_special_tokens(special)          # synthetic / byte path: vocabulary stays the model's
    else:
        sft.smart_tokenizer_and_embedding_resize(special, tokenizer, model)
    data_module = sft.make_supervised_data_module(tokenizer, data_args, training_args, vocab_size=model.config.vocab_size)
    trainer = Trainer(model=model, tokenizer=tokenizer, args=training_args, **data_module)
    resume = training_args.resume_from_checkpoint
    metrics = trainer.train(resume_from_checkpoint=(True if str(resume).lower() == "true" else resume) if resume else None)
    trainer.save_state()
    trainer.save_model(output_dir=training_args.output_dir)
    if torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()
    return metrics


if __name__ == "__main__":
    train()
