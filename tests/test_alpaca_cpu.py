"""Alpaca / ZeRO-3 workload on CPU: HF-style args, DeepSpeed "auto" resolution (golden values from the reference's
launch, 4_...ipynb cell 24 + log :1955-2006), SFT preprocessing, HF checkpoint round trip, DeepSpeed-layout shards and
zero_to_fp32 consolidation, and the train.py entry point on 2 gloo ranks."""
import json
import os
import subprocess
import sys

import pytest
import torch

from b200dist.ckpt import deepspeed as ds_ckpt
from b200dist.ckpt import hf as hf_ckpt
from b200dist.config.ds_config import load_ds_config, resolve_auto
from b200dist.config.hf_args import DataArguments, HfArgumentParser, ModelArguments, TrainingArguments
from b200dist.data import sft
from b200dist.models.llama import build_causal_lm


def _free_port():
    import socket
    with socket.socket() as so:
        so.bind(("127.0.0.1", 0))
        return so.getsockname()[1]


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _args(extra=()):
    p = HfArgumentParser((ModelArguments, DataArguments, TrainingArguments))
    return p.parse_args_into_dataclasses(["--output_dir", "/tmp/x", *extra])


def test_hf_args_reference_launch_line():
    m, d, t = _args(["--model_name_or_path", "facebook/opt-125m", "--data_path", "alpaca_data.json", "--bf16", "True",
                     "--num_train_epochs", "1", "--per_device_train_batch_size", "1", "--per_device_eval_batch_size", "1",
                     "--gradient_accumulation_steps", "8", "--evaluation_strategy", "no", "--save_strategy", "steps",
                     "--save_steps", "2000", "--save_total_limit", "1", "--learning_rate", "2e-5", "--weight_decay", "0.",
                     "--warmup_ratio", "0.03", "--deepspeed", "configs/default_offload_opt_param.json", "--tf32", "True"])
    assert m.model_name_or_path == "facebook/opt-125m" and d.data_path == "alpaca_data.json"
    assert t.bf16 is True and t.tf32 is True and t.gradient_accumulation_steps == 8 and t.learning_rate == 2e-5
    assert t.model_max_length == 512 and t.optim == "adamw_torch"          # train.py:84-88 defaults
    assert t.get_warmup_steps(1000) == 30


def test_ds_auto_resolution_golden():
    _, _, t = _args(["--bf16", "True", "--per_device_train_batch_size", "1", "--gradient_accumulation_steps", "8",
                     "--learning_rate", "2e-5", "--warmup_ratio", "0.03"])
    cfg = load_ds_config(os.path.join(ROOT, "configs", "default_offload_opt_param.json"))
    ds = resolve_auto(cfg, t, hidden_size=768, world_size=8, num_training_steps=812)
    # the values DeepSpeed printed for this exact launch (SURVEY 6.4 / notebook log)
    assert ds.train_batch_size == 64 and ds.train_micro_batch_size_per_gpu == 1 and ds.gradient_accumulation_steps == 8
    assert ds.zero.stage == 3 and ds.zero.reduce_bucket_size == 589824 and ds.zero.stage3_prefetch_bucket_size == 530841
    assert ds.zero.stage3_param_persistence_threshold == 7680
    # T14 drops the optimizer-offload device (only pin_memory is left); T15 (-original) keeps both on the CPU
    assert ds.zero.offload_optimizer_device == "none" and ds.zero.offload_param_device == "cpu"
    orig = resolve_auto(load_ds_config(os.path.join(ROOT, "configs", "default_offload_opt_param-original.json")), t, 768, 8, 812)
    assert orig.zero.offload_optimizer_device == "cpu" and orig.zero.offload_param_device == "cpu"
    assert ds.bf16 and not ds.fp16 and ds.optimizer_type == "AdamW" and ds.optimizer_params["lr"] == 2e-5
    assert ds.scheduler_type == "WarmupDecayLR" and ds.scheduler_params["warmup_num_steps"] == 25   # ceil(812*0.03)
    assert ds.scheduler_params["total_num_steps"] == 812 and ds.gradient_clipping == 1.0
    assert "auto" not in json.dumps(cfg)


def test_ds_unresolved_auto_raises():
    _, _, t = _args()
    with pytest.raises(ValueError):
        resolve_auto({"something_new": {"x": "auto"}}, t, 64, 1, 10)


def test_sft_preprocess_masks_prompt():
    tok = sft.ByteTokenizer(model_max_length=512)
    tok.add_special_tokens(dict(pad_token=sft.DEFAULT_PAD_TOKEN, eos_token=sft.DEFAULT_EOS_TOKEN, bos_token=sft.DEFAULT_BOS_TOKEN,
                                unk_token=sft.DEFAULT_UNK_TOKEN))
    src = sft.PROMPT_DICT["prompt_no_input"].format_map({"instruction": "Say hi."})
    d = sft.preprocess([src], ["hi</s>"], tok)
    ids, lab = d["input_ids"][0], d["labels"][0]
    n = len(tok.encode(src))
    assert (lab[:n] == sft.IGNORE_INDEX).all() and (lab[n:] == ids[n:]).all()
    assert ids[-1].item() == tok.eos_token_id and len(ids) == n + 3
    batch = sft.DataCollatorForSupervisedDataset(tok)([dict(input_ids=ids, labels=lab), dict(input_ids=ids[:10], labels=lab[:10])])
    assert batch["input_ids"].shape == (2, len(ids)) and batch["input_ids"][1, 10:].eq(tok.pad_token_id).all()
    assert batch["labels"][1, 10:].eq(sft.IGNORE_INDEX).all() and batch["attention_mask"][1].sum() == 10


def test_smart_resize_mean_init():
    model = build_causal_lm("llama-tiny")
    old = model.embed_tokens.shape[0]
    tok = sft.ByteTokenizer(vocab_size=old + 3)
    sft.smart_tokenizer_and_embedding_resize(dict(pad_token="[PAD]"), tok, model)
    assert model.embed_tokens.shape[0] == old + 3 and model.config.vocab_size == old + 3
    assert torch.allclose(model.embed_tokens[old:], model.embed_tokens[:old].mean(0, keepdim=True).expand(3, -1), atol=1e-6)


@pytest.mark.parametrize("name,safe", [("llama-tiny", False), ("opt-tiny", False), ("llama-tiny", True)])
def test_hf_checkpoint_round_trip(tmp_path, name, safe):
    if safe:
        pytest.importorskip("safetensors")
    torch.manual_seed(0)
    model = build_causal_lm(name)
    hf_ckpt.save_pretrained(model, str(tmp_path), max_shard_size=200_000, safe_serialization=safe)     # forces several shards
    stem = "model.safetensors" if safe else "pytorch_model.bin"
    idx = json.load(open(tmp_path / (stem + ".index.json")))
    assert set(idx["weight_map"].values()) == {f for f in os.listdir(tmp_path) if f.startswith(stem.split(".")[0] + "-")}
    model2 = hf_ckpt.load_pretrained(str(tmp_path))
    ids = torch.randint(0, model.config.vocab_size, (2, 16))
    a, b = model(ids), model2(ids)
    assert torch.equal(a, b)
    sd = hf_ckpt.read_state_dict(str(tmp_path))
    key = "model.layers.0.self_attn.q_proj.weight" if name.startswith("llama") else "model.decoder.layers.0.self_attn.q_proj.weight"
    assert key in sd and sd[key].shape == (model.config.hidden_size, model.config.hidden_size)


def _run_train(tmp_path, nproc, extra):
    data = tmp_path / "alpaca.json"
    data.write_text(json.dumps([{"instruction": "Add.", "input": "2 and 3", "output": "5"}, {"instruction": "Say hi.", "input": "", "output": "hi"},
                                {"instruction": "Name a color.", "input": "", "output": "blue"}, {"instruction": "Double it.", "input": "4", "output": "8"}]))
    out = tmp_path / "out"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "entrypoints", "train.py"),
           "--model_name_or_path", "opt-tiny", "--data_path", str(data), "--output_dir", str(out), "--max_steps", "5",
           "--per_device_train_batch_size", "1", "--gradient_accumulation_steps", "2", "--learning_rate", "1e-3", "--logging_steps", "1",
           "--model_max_length", "512", "--bf16", "True", *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env={**os.environ, "OMP_NUM_THREADS": "1"})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    losses = [float(l.split("'loss': ")[1].split(",")[0]) for l in r.stdout.splitlines() if l.startswith("{'loss'")]
    return out, losses


def test_train_entrypoint_zero3_shards_and_consolidation(tmp_path):
    out, losses = _run_train(tmp_path, 2, ["--deepspeed", os.path.join(ROOT, "configs", "default_offload_opt_param.json")])
    assert len(losses) == 5 and losses[-1] < losses[0]
    assert (out / "latest").read_text() == "global_step5" and (out / "trainer_state.json").exists()
    files = sorted(os.listdir(out / "global_step5"))
    assert len(files) == 4 and all("zero_pp_rank_" in f for f in files)
    sd = ds_ckpt.zero_to_fp32(str(out))
    model = build_causal_lm("opt-tiny")
    n_model = sum(p.numel() for p in model.parameters())
    assert sum(v.numel() for v in sd.values()) == n_model
    # the consolidation script dropped next to the shards works stand-alone (torch only)
    assert (out / "zero_to_fp32.py").exists()
    r = subprocess.run([sys.executable, str(out / "zero_to_fp32.py"), str(out), str(out / "fp32.pt"), "--strip-unit-prefix"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    flat = torch.load(out / "fp32.pt")
    assert sum(v.numel() for v in flat.values()) == n_model and all(torch.equal(flat[k], v) for k, v in hf_ckpt.rename_unit_keys(sd).items() if k in flat)
    assert all(v.dtype == torch.float32 and torch.isfinite(v).all() for v in sd.values())


def test_train_entrypoint_gathered_save(tmp_path):
    out, losses = _run_train(tmp_path, 2, ["--deepspeed", os.path.join(ROOT, "configs", "zero3_b200_resident.json")])
    assert losses[-1] < losses[0]
    model = hf_ckpt.load_pretrained(str(out))
    assert torch.isfinite(model(torch.randint(0, 256, (1, 8)))).all()


def test_train_entrypoint_original_config_offloads_the_optimizer_to_the_cpu(tmp_path):
    """configs/default_offload_opt_param-original.json (offload_optimizer.device = cpu): ZeRO-Offload with the C++ CPU Adam;
    same loss curve as the config that keeps the optimizer on the device"""
    out, losses = _run_train(tmp_path, 2, ["--deepspeed", os.path.join(ROOT, "configs", "default_offload_opt_param-original.json")])
    assert len(losses) == 5 and losses[-1] < losses[0]
    (tmp_path / "b").mkdir()
    _, ref = _run_train(tmp_path / "b", 2, ["--deepspeed", os.path.join(ROOT, "configs", "default_offload_opt_param.json")])
    assert all(abs(a - b) < 5e-3 for a, b in zip(losses, ref)), (losses, ref)
    sd = ds_ckpt.zero_to_fp32(str(out))
    assert all(torch.isfinite(v).all() for v in sd.values())


def test_trainer_checkpoint_rotation_resume_and_evaluation(tmp_path):
    """save_steps / save_total_limit (4_...ipynb:474-476) produce resumable checkpoints; --resume_from_checkpoint continues
    with the losses of the uninterrupted run; evaluation_strategy=steps reports eval_loss"""
    out, losses = _run_train(tmp_path, 2, ["--max_steps", "6", "--save_steps", "2", "--save_total_limit", "2", "--weight_decay", "0.01"])
    assert len(losses) == 6
    assert sorted(n for n in os.listdir(out) if n.startswith("checkpoint-")) == ["checkpoint-4", "checkpoint-6"]
    assert json.load(open(out / "checkpoint-4" / "trainer_state.json"))["global_step"] == 4
    import shutil
    shutil.rmtree(out / "checkpoint-6")
    out2, resumed = _run_train(tmp_path, 2, ["--max_steps", "6", "--save_steps", "2", "--save_total_limit", "2", "--weight_decay", "0.01",
                                            "--resume_from_checkpoint", "True"])
    assert len(resumed) == 2 and all(abs(a - b) < 2e-4 for a, b in zip(resumed, losses[4:])), (resumed, losses)
    # evaluation in-process (single rank)
    from b200dist.config.hf_args import DataArguments, TrainingArguments
    from b200dist.data import sft
    from b200dist.engine.hf_trainer import Trainer
    targs = TrainingArguments(output_dir=str(tmp_path / "ev"), max_steps=4, per_device_train_batch_size=2, logging_steps=2, evaluation_strategy="steps",
                              eval_steps=2, save_strategy="no", learning_rate=1e-3, synthetic_data=True, synthetic_samples=16, model_max_length=32,
                              comm_backend="nccl")
    model = build_causal_lm("opt-tiny")
    tok = sft.ByteTokenizer(model_max_length=32)
    dm = sft.make_supervised_data_module(tok, DataArguments(data_path=None), targs, vocab_size=model.config.vocab_size)
    dm["eval_dataset"] = dm["train_dataset"]
    tr = Trainer(model=model, tokenizer=tok, args=targs, **dm)
    tr.train()
    evals = [h["eval_loss"] for h in tr.state.log_history if "eval_loss" in h]
    assert len(evals) == 2 and evals[1] < evals[0]


def test_hf_training_arguments_compatibility_flags(capsys):
    """flags of transformers.TrainingArguments that the trainer does not act on are accepted with a notice; unknown flags fail"""
    p = HfArgumentParser((ModelArguments, DataArguments, TrainingArguments))
    m, d, t = p.parse_args_into_dataclasses(["--model_name_or_path", "x", "--output_dir", "o", "--logging_dir", "l", "--fsdp", "full_shard auto_wrap",
                                             "--eval_strategy", "steps", "--overwrite_output_dir", "--do_train", "True", "--report_to", "none",
                                             "--gradient_checkpointing", "True", "--ddp_timeout=1800"])
    assert t.evaluation_strategy == "steps" and t.gradient_checkpointing is True and t.output_dir == "o"
    assert "fsdp" in capsys.readouterr().out
    with pytest.raises(ValueError):
        p.parse_args_into_dataclasses(["--output_dir", "o", "--definitely_not_a_flag", "1"])
