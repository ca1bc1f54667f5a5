"""Pretrain GPT - same structure and CLI as the reference's 3_training_megatron-lm/pretrain_gpt.py
(model_provider :46-58, get_batch :61-90, loss_func :92-100, forward_step :103-117,
train_valid_test_datasets_provider :120-140, pretrain(...) :143-149), on the b200dist runtime.

    torchrun --nproc-per-node 8 entrypoints/pretrain_gpt.py --num-layers 12 --hidden-size 768 \
        --num-attention-heads 12 --seq-length 1024 --max-position-embeddings 1024 --micro-batch-size 12 \
        --global-batch-size 192 --tensor-model-parallel-size 2 --sequence-parallel --bf16 \
        --tokenizer-type NullTokenizer --vocab-size 32767 --mock-data --train-iters 100 --lr 5e-4
"""
import os
import sys
from functools import partial

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

# SageMaker / mpirun shim (reference pretrain_gpt.py:10-14): publish torch-style rank variables
from b200dist.launch import env as _env
_env.export_torch_env()

import torch

from b200dist.config.megatron_args import core_transformer_config_from_args
from b200dist.data.gpt_dataset import build_train_valid_test_datasets
from b200dist.engine.megatron import get_args, get_timers, get_tokenizer, pretrain, print_rank_0
from b200dist.models.gpt import GPTModel
from b200dist.parallel import state as mpu
from b200dist.parallel import tensor as tensor_parallel


def model_provider(pre_process=True, post_process=True):
    """Build the model."""
    print_rank_0("building GPT model ...")
    args = get_args()
    config = core_transformer_config_from_args(args)
    dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    return GPTModel(config, num_tokentypes=0, parallel_output=True, pre_process=pre_process, post_process=post_process,
                    device=dev, seed=args.seed)


def get_batch(data_iterator):
    """Generate a batch: tokens / labels shifted by one, loss mask, position ids.  The [1,1,s,s] causal mask
    of the reference (get_ltor_masks_and_position_ids, :83-88) is implicit in the attention kernel."""
    args, tokenizer = get_args(), get_tokenizer()
    data = next(data_iterator) if data_iterator is not None else None
    if getattr(args, "_static_batch", False) and data is not None and data["text"].is_cuda:
        data_b = data        # CUDA-graph mode: the batch already sits in the static device buffer on every TP rank
    else:
        data_b = tensor_parallel.broadcast_data(["text"], data, torch.int64)
    tokens_ = data_b["text"].long()
    labels = tokens_[:, 1:].contiguous()
    tokens = tokens_[:, :-1].contiguous()
    loss_mask = torch.ones(tokens.shape, dtype=torch.float, device=tokens.device)
    if args.eod_mask_loss:
        loss_mask[tokens == tokenizer.eod] = 0.0
    position_ids = torch.arange(tokens.shape[1], dtype=torch.long, device=tokens.device).unsqueeze(0).expand_as(tokens)
    if args.reset_position_ids:
        position_ids = reset_position_ids_after_eod(tokens, tokenizer.eod)
    attention_mask = None
    if args.reset_attention_mask:
        # the reference materialises a [b, 1, s, s] mask; the document id of every token says the same in [b, s]: a token attends to
        # earlier tokens of its own document only (an EOD closes the document it ends)
        is_eod = (tokens == tokenizer.eod).long()
        attention_mask = torch.cumsum(is_eod, dim=1) - is_eod
    return tokens, labels, loss_mask, attention_mask, position_ids


def reset_position_ids_after_eod(tokens, eod):
    """--reset-position-ids (get_ltor_masks_and_position_ids, megatron/utils.py): positions restart at 0 after every end-of-document
    token.  position[i] = i - (index of the last EOD before i, + 1), computed with a running maximum instead of the reference's
    per-document Python loop."""
    S = tokens.shape[1]
    idx = torch.arange(S, dtype=torch.long, device=tokens.device).unsqueeze(0).expand_as(tokens)
    starts = torch.where(tokens == eod, idx + 1, torch.zeros_like(idx))          # a document starts right after an EOD
    starts = torch.cat([torch.zeros_like(starts[:, :1]), starts[:, :-1]], 1)     # ... which affects the tokens AFTER it
    return idx - torch.cummax(starts, dim=1).values


def loss_func(loss_mask, output_tensor):
    losses = output_tensor.float()
    loss_mask = loss_mask.view(-1).float()
    loss = torch.sum(losses.view(-1) * loss_mask) / loss_mask.sum()
    # data-parallel average for logging (reference :98); kept on the device, read only at log time
    averaged = loss.detach().clone()
    if (torch.distributed.is_initialized() and mpu.get_data_parallel_world_size() > 1
            and not getattr(get_args(), "_static_batch", False)):      # CUDA-graph mode averages once per step outside the graph
        torch.distributed.all_reduce(averaged, group=mpu.get_data_parallel_group())
        averaged = averaged / mpu.get_data_parallel_world_size()
    return loss, {"lm loss": averaged}


def forward_step(data_iterator, model):
    args, timers = get_args(), get_timers()
    timers("batch-generator", log_level=2).start()
    tokens, labels, loss_mask, attention_mask, position_ids = get_batch(data_iterator)
    timers("batch-generator").stop()
    output_tensor = model(tokens, position_ids, attention_mask, labels=labels)
    return output_tensor, partial(loss_func, loss_mask)


def train_valid_test_datasets_provider(train_val_test_num_samples):
    args = get_args()
    print_rank_0("> building train, validation, and test datasets for GPT ...")
    train_ds, valid_ds, test_ds = build_train_valid_test_datasets(
        data_prefix=args.data_path, data_impl=args.data_impl, splits_string=args.split,
        train_valid_test_num_samples=train_val_test_num_samples, seq_length=args.seq_length, seed=args.seed,
        skip_warmup=(not args.mmap_warmup), train_data_prefix=args.train_data_path, valid_data_prefix=args.valid_data_path,
        test_data_prefix=args.test_data_path, data_cache_path=args.data_cache_path, mock=args.mock_data,
        vocab_size=args.padded_vocab_size)
    print_rank_0("> finished creating GPT datasets ...")
    return train_ds, valid_ds, test_ds


if __name__ == "__main__":
    pretrain(train_valid_test_datasets_provider, model_provider, "encoder_or_decoder", forward_step,
             args_defaults={"tokenizer_type": "GPT2BPETokenizer"})
    # orderly teardown: leaving the process with live communicator threads can abort at interpreter exit
    # ("terminate called without an active exception") when the peers are already gone
    if torch.distributed.is_initialized():
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
