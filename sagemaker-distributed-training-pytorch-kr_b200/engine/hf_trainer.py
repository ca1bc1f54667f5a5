"""HF-Trainer-style loop on the ZeRO-3 engine (reference: train.py:243-246 `Trainer(...).train(); save_state();
save_model()` -> transformers.Trainer + DeepSpeedEngine).

  Zero3CausalLM  - splits a causal LM into gather units (embeddings | each decoder layer | final norm + head) and runs
                   them through Zero3Engine.wrap (fetch / prefetch / reduce-scatter(+AdamW) around every unit).
  Trainer        - gradient accumulation, WarmupDecayLR / linear schedule, logging every `logging_steps`
                   ({'loss', 'learning_rate', 'epoch'}), final metrics line {'train_runtime', 'train_samples_per_second',
                   'train_steps_per_second', 'train_loss', 'epoch'} (4_...ipynb:2306), save_state -> trainer_state.json,
                   save_model -> ZeRO-3 shard files (or gathered 16-bit HF weights when the config asks for them).
"""
from __future__ import annotations

import json
import math
import os
import time
from typing import Dict, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from ..config import ds_config as dsc
from ..optim.schedulers import WarmupDecayLR, WarmupLR
from ..parallel.zero3 import Zero3Engine


class _EmbedUnit(nn.Module):
    def __init__(self, model):
        super().__init__()
        self.embed_tokens = model.embed_tokens
        if hasattr(model, "embed_positions"):
            self.embed_positions = model.embed_positions
        self._m = [model]

    def forward(self, ids):
        return self._m[0].embed(ids)


class _HeadUnit(nn.Module):
    def __init__(self, model):
        super().__init__()
        for n in ("norm", "final_layer_norm_w", "final_layer_norm_b", "lm_head"):
            if getattr(model, n, None) is not None:
                setattr(self, n, getattr(model, n))
        self._m = [model]

    def forward(self, h, labels):
        return self._m[0].head_loss(h, labels)


class Zero3CausalLM(nn.Module):
    def __init__(self, model: nn.Module, *, lr=1e-5, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_grad_norm=1.0,
                 gradient_accumulation_steps=1, persistence_threshold=0, comm_backend="native", offload_param="none", group=None,
                 offload_optimizer="none", gradient_checkpointing=False, max_live_parameters=0, prefetch_bucket_size=0, max_reuse_distance=0,
                 fuse_head_gather=True, materialize=None, device=None):
        super().__init__()
        self.model = model
        self.tied = getattr(model, "lm_head", None) is None
        if self.tied:      # a matrix used by two units (embedding + head) stays replicated; its two gradients add up in .grad
            model.embed_tokens._zero3_persistent = True
        units: List[nn.Module] = [_EmbedUnit(model)] + list(model.layers) + [_HeadUnit(model)]
        names = ["embed"] + [f"layers.{i}" for i in range(len(model.layers))] + ["head"]
        self.engine = Zero3Engine(units, names, group, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, max_grad_norm=max_grad_norm,
                                  gradient_accumulation_steps=gradient_accumulation_steps, persistence_threshold=persistence_threshold,
                                  comm_backend=comm_backend, offload_param=offload_param, offload_optimizer=offload_optimizer,
                                  checkpoint_units=range(1, len(units) - 1) if gradient_checkpointing else (),      # the decoder layers
                                  max_live_parameters=max_live_parameters, prefetch_bucket_size=prefetch_bucket_size,
                                  max_reuse_distance=max_reuse_distance,
                                  # the untied LM head reaches its GEMM through ops.functional.linear: gather it inside that GEMM
                                  fuse_gather_gemm_units=() if (self.tied or not fuse_head_gather) else (len(units) - 1,),
                                  materialize=materialize, device=device)
        self._units = units

    def forward(self, input_ids, attention_mask=None, labels=None):
        m, eng = self.model, self.engine
        B, S = input_ids.shape
        extra = ()
        if hasattr(m, "rope_tables"):
            cos_t, sin_t = m.rope_tables(input_ids.device)
            pos = torch.arange(S, device=input_ids.device, dtype=torch.int32).repeat_interleave(B)
            extra = (cos_t, sin_t, pos)
        h = eng.wrap(0, self._units[0], input_ids)
        for i, layer in enumerate(m.layers):
            h = eng.wrap(i + 1, layer, h, *extra)
        last = len(self._units) - 1
        return eng.wrap(last, lambda hh: self._units[last](hh, labels), h)


class TrainerState:
    def __init__(self):
        self.global_step, self.epoch, self.log_history, self.max_steps = 0, 0.0, [], 0
        self.total_flos, self.best_metric = 0, None

    def to_json(self):
        return {"global_step": self.global_step, "epoch": self.epoch, "max_steps": self.max_steps, "log_history": self.log_history,
                "total_flos": self.total_flos, "best_metric": self.best_metric}


class Trainer:
    def __init__(self, model: nn.Module, tokenizer=None, args=None, train_dataset=None, eval_dataset=None, data_collator=None):
        self.raw_model, self.tokenizer, self.args = model, tokenizer, args
        self.train_dataset, self.eval_dataset, self.data_collator = train_dataset, eval_dataset, data_collator
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        args.world_size = self.world
        self.state = TrainerState()
        self.ds: Optional[dsc.DSConfig] = None
        self.zmodel: Optional[Zero3CausalLM] = None
        self.scheduler = None

    # ------------------------------------------------------------------ setup
    def _num_steps(self) -> int:
        a = self.args
        per_step = a.per_device_train_batch_size * a.gradient_accumulation_steps * self.world
        if a.max_steps and a.max_steps > 0:
            return a.max_steps
        return max(1, math.ceil(len(self.train_dataset) * a.num_train_epochs / per_step))

    def _setup(self):
        a = self.args
        total = self._num_steps()
        self.state.max_steps = total
        hidden = self.raw_model.config.hidden_size
        if a.deepspeed:
            raw = dsc.load_ds_config(a.deepspeed)
        else:       # no JSON given: plain ZeRO-3 with everything derived from the arguments
            raw = {"bf16": {"enabled": "auto"}, "optimizer": {"type": "AdamW", "params": {"lr": "auto", "betas": "auto", "eps": "auto", "weight_decay": "auto"}},
                   "scheduler": {"type": "WarmupDecayLR", "params": {"total_num_steps": "auto", "warmup_min_lr": "auto", "warmup_max_lr": "auto", "warmup_num_steps": "auto"}},
                   "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": "auto"}, "gradient_accumulation_steps": "auto",
                   "gradient_clipping": "auto", "train_batch_size": "auto", "train_micro_batch_size_per_gpu": "auto"}
        self.ds = dsc.resolve_auto(raw, a, hidden, self.world, total)
        if self.ds.zero.stage != 3:
            raise ValueError("this trainer implements ZeRO stage 3 (the reference config); got stage %d" % self.ds.zero.stage)
        if self.ds.zero.offload_optimizer_device == "cpu" and self.rank == 0:
            print("[zero3] offload_optimizer.device=cpu: fp32 master weights and Adam moments in pinned host memory, step on the CPU "
                  "(csrc/host/cpu_adam.cpp); a B200 holds them in HBM comfortably - configs/zero3_b200_resident.json keeps them there")
        op = self.ds.optimizer_params           # the JSON optimizer block wins over --optim
        self.zmodel = Zero3CausalLM(self.raw_model, lr=op.get("lr", a.learning_rate), betas=tuple(op.get("betas", (a.adam_beta1, a.adam_beta2))),
                                    eps=op.get("eps", a.adam_epsilon), weight_decay=op.get("weight_decay", a.weight_decay),
                                    max_grad_norm=self.ds.gradient_clipping, gradient_accumulation_steps=self.ds.gradient_accumulation_steps,
                                    persistence_threshold=self.ds.zero.stage3_param_persistence_threshold, comm_backend=a.comm_backend,
                                    offload_param=self.ds.zero.offload_param_device,
                                    offload_optimizer=self.ds.zero.offload_optimizer_device,
                                    gradient_checkpointing=bool(a.gradient_checkpointing),
                                    max_live_parameters=self.ds.zero.stage3_max_live_parameters,
                                    prefetch_bucket_size=self.ds.zero.stage3_prefetch_bucket_size,
                                    max_reuse_distance=self.ds.zero.stage3_max_reuse_distance)
        eng = self.zmodel.engine
        sp = self.ds.scheduler_params
        if self.ds.scheduler_type == "WarmupDecayLR":
            self.scheduler = WarmupDecayLR(eng, sp["total_num_steps"], sp.get("warmup_min_lr", 0), sp.get("warmup_max_lr", a.learning_rate),
                                           sp.get("warmup_num_steps", 0), sp.get("warmup_type", "log"))
        elif self.ds.scheduler_type == "WarmupLR":
            self.scheduler = WarmupLR(eng, sp.get("warmup_min_lr", 0), sp.get("warmup_max_lr", a.learning_rate), sp.get("warmup_num_steps", 0))
        elif self.ds.scheduler_type is None:      # no scheduler block in the JSON: the HF Trainer's own schedule (--lr_scheduler_type)
            from ..optim.schedulers import HFScheduler
            self.scheduler = HFScheduler(eng, a.lr_scheduler_type, a.get_warmup_steps(total), total)
        if self.rank == 0:
            n_persist = eng.num_persistent
            print(f"[zero3] partitioned {eng.total_params() - n_persist} parameters in {len(eng.units)} units over {self.world} ranks; "
                  f"Parameter Offload: Total persistent parameters: {n_persist} in {len(eng.persistent)} params", flush=True)

    def _loader(self):
        a = self.args
        sampler = None
        if self.world > 1:
            sampler = torch.utils.data.distributed.DistributedSampler(self.train_dataset, self.world, self.rank, shuffle=True, seed=a.seed)
        return torch.utils.data.DataLoader(self.train_dataset, batch_size=a.per_device_train_batch_size, sampler=sampler, shuffle=sampler is None,
                                           collate_fn=self.data_collator, num_workers=a.dataloader_num_workers, pin_memory=torch.cuda.is_available(),
                                           drop_last=True), sampler

    # ------------------------------------------------------------------ training
    def training_step(self, batch) -> torch.Tensor:
        dev = self.zmodel.engine.dev
        ids = batch["input_ids"].to(dev, non_blocking=True)
        labels = batch["labels"].to(dev, non_blocking=True)
        loss = self.zmodel(ids, attention_mask=None, labels=labels)
        self.zmodel.engine.backward(loss)
        return loss.detach()

    def train(self, resume_from_checkpoint=None):
        """`resume_from_checkpoint`: a `checkpoint-N` directory, or True for the newest one under output_dir (the
        transformers.Trainer contract; the reference never passes it, train.py:244, but its save_steps/save_total_limit
        settings - 4_...ipynb:474-476 - exist for exactly this)."""
        from ..utils.metrics import JsonlSink
        torch.manual_seed(self.args.seed)
        self._setup()
        a, eng = self.args, self.zmodel.engine
        loader, sampler = self._loader()
        total = self.state.max_steps
        steps_per_epoch = max(1, len(loader) // a.gradient_accumulation_steps)
        sink = JsonlSink(rank=self.rank)
        skip_batches, epoch = 0, 0
        if resume_from_checkpoint:
            ck = self._latest_checkpoint() if resume_from_checkpoint is True else str(resume_from_checkpoint)
            if ck is None:
                raise ValueError(f"no checkpoint-* directory under {a.output_dir} to resume from")
            self._load_checkpoint(ck)
            consumed = self.state.global_step * a.gradient_accumulation_steps
            epoch, skip_batches = consumed // max(1, len(loader)), consumed % max(1, len(loader))
            if self.rank == 0:
                print(f"[trainer] resumed from {ck}: global step {self.state.global_step}, epoch {epoch}, skipping {skip_batches} batches", flush=True)
        t0 = time.time()
        tr_loss = torch.zeros((), device=eng.dev)
        logged_loss, samples = 0.0, 0
        done = self.state.global_step >= total
        while not done:
            if sampler is not None:
                sampler.set_epoch(epoch)
            else:
                torch.manual_seed(a.seed + epoch)          # single process: the shuffle order is a function of the epoch (resumable)
            for i, batch in enumerate(loader):
                if skip_batches > 0:
                    skip_batches -= 1
                    continue
                tr_loss += self.training_step(batch) / a.gradient_accumulation_steps
                samples += batch["input_ids"].shape[0] * self.world
                if eng.step():
                    if self.scheduler is not None:
                        self.scheduler.step()
                    self.state.global_step += 1
                    self.state.epoch = self.state.global_step / steps_per_epoch
                    if a.logging_steps and self.state.global_step % a.logging_steps == 0:
                        cur = float(tr_loss)
                        log = {"loss": round((cur - logged_loss) / a.logging_steps, 4), "learning_rate": eng.param_groups[0]["lr"],
                               "epoch": round(self.state.epoch, 2)}
                        logged_loss = cur
                        self.state.log_history.append({**log, "step": self.state.global_step})
                        if self.rank == 0:
                            print(log, flush=True)
                            sink.write(workload="sft", step=self.state.global_step, **log)
                    ev = a.evaluation_strategy
                    if self.eval_dataset is not None and ((ev == "steps" and self.state.global_step % (a.eval_steps or a.logging_steps) == 0)
                                                          or (ev == "epoch" and self.state.global_step % steps_per_epoch == 0)):
                        self.evaluate()
                    if a.save_strategy == "steps" and a.save_steps and self.state.global_step % a.save_steps == 0:
                        self._save_checkpoint()
                    elif a.save_strategy == "epoch" and self.state.global_step % steps_per_epoch == 0:
                        self._save_checkpoint()
                    if self.state.global_step >= total:
                        done = True
                        break
            epoch += 1
        if eng.dev.type == "cuda":
            torch.cuda.synchronize()
        runtime = time.time() - t0
        metrics = {"train_runtime": round(runtime, 4), "train_samples_per_second": round(samples / runtime, 3),
                   "train_steps_per_second": round(self.state.global_step / runtime, 3),
                   "train_loss": float(tr_loss) / max(self.state.global_step, 1), "epoch": round(self.state.epoch, 2)}
        self.state.log_history.append({**metrics, "step": self.state.global_step})
        if self.rank == 0:
            print(metrics, flush=True)
        return metrics

    # ------------------------------------------------------------------ evaluation
    def evaluate(self, eval_dataset=None) -> dict:
        """mean per-token loss over the (rank-sharded) evaluation set; logs {'eval_loss': ...} like transformers.Trainer"""
        ds = eval_dataset if eval_dataset is not None else self.eval_dataset
        if ds is None:
            return {}
        a, eng = self.args, self.zmodel.engine
        sampler = torch.utils.data.distributed.DistributedSampler(ds, self.world, self.rank, shuffle=False) if self.world > 1 else None
        loader = torch.utils.data.DataLoader(ds, batch_size=a.per_device_eval_batch_size, sampler=sampler, shuffle=False,
                                             collate_fn=self.data_collator, drop_last=self.world > 1)
        tot = torch.zeros(2, dtype=torch.float64, device=eng.dev)
        was_training = self.zmodel.training
        self.zmodel.eval()
        t0 = time.time()
        with torch.no_grad():
            for batch in loader:
                ids, labels = batch["input_ids"].to(eng.dev), batch["labels"].to(eng.dev)
                n = (labels[:, 1:] != -100).sum()
                tot[0] += self.zmodel(ids, attention_mask=None, labels=labels).double() * n
                tot[1] += n
        if self.world > 1:
            t32 = tot.float()
            dist.all_reduce(t32)
            tot = t32.double()
        self.zmodel.train(was_training)
        metrics = {"eval_loss": float(tot[0] / tot[1].clamp(min=1)), "eval_runtime": round(time.time() - t0, 4), "epoch": round(self.state.epoch, 2)}
        self.state.log_history.append({**metrics, "step": self.state.global_step})
        if self.rank == 0:
            print(metrics, flush=True)
        return metrics

    # ------------------------------------------------------------------ persistence
    def _checkpoints(self):
        out = self.args.output_dir
        if not os.path.isdir(out):
            return []
        found = []
        for n in os.listdir(out):
            if n.startswith("checkpoint-") and n[len("checkpoint-"):].isdigit() and os.path.isfile(os.path.join(out, n, "latest")):
                found.append((int(n[len("checkpoint-"):]), os.path.join(out, n)))
        return [p for _, p in sorted(found)]

    def _latest_checkpoint(self):
        c = self._checkpoints()
        return c[-1] if c else None

    def _save_checkpoint(self):
        """resumable intermediate checkpoint: ZeRO-3 shards of every rank (fp32 master + Adam moments + scheduler) in
        `output_dir/checkpoint-N/`, trainer_state.json next to them; `save_total_limit` keeps the newest N"""
        import shutil
        from ..ckpt import deepspeed as ds_ckpt
        d = os.path.join(self.args.output_dir, f"checkpoint-{self.state.global_step}")
        os.makedirs(d, exist_ok=True)
        ds_ckpt.save_zero3_checkpoint(self.zmodel.engine, d, self.state.global_step, self.ds.raw if self.ds else {},
                                      scheduler=self.scheduler, persistent=self._persistent_named())
        if self.rank == 0:
            with open(os.path.join(d, "trainer_state.json"), "w") as f:
                json.dump(self.state.to_json(), f, indent=2)
            limit = self.args.save_total_limit
            if limit is not None and limit > 0:
                for old in self._checkpoints()[:-limit]:
                    shutil.rmtree(old, ignore_errors=True)
        if dist.is_initialized():
            dist.barrier()

    def _load_checkpoint(self, ck: str):
        from ..ckpt import deepspeed as ds_ckpt
        step = ds_ckpt.load_zero3_checkpoint(self.zmodel.engine, ck, scheduler=self.scheduler)
        self.state.global_step = int(step)
        ts = os.path.join(ck, "trainer_state.json")
        if os.path.isfile(ts):
            with open(ts) as f:
                self.state.log_history = json.load(f).get("log_history", [])

    def save_state(self):
        if self.rank == 0:
            os.makedirs(self.args.output_dir, exist_ok=True)
            with open(os.path.join(self.args.output_dir, "trainer_state.json"), "w") as f:
                json.dump(self.state.to_json(), f, indent=2)

    def save_model(self, output_dir: Optional[str] = None):
        from ..ckpt import deepspeed as ds_ckpt
        from ..ckpt import hf as hf_ckpt
        out = output_dir or self.args.output_dir
        os.makedirs(out, exist_ok=True)
        if self.ds is not None and self.ds.zero.stage3_gather_16bit_weights_on_model_save:
            sd = self.zmodel.engine.gather_full_state_dict()
            for p, n in self._persistent_named():
                sd[n] = p.detach().cpu()
            if self.rank == 0:
                hf_ckpt.save_pretrained(self.raw_model, out, state_dict_override=hf_ckpt.rename_unit_keys(sd))
        else:
            if self.rank == 0:
                print("[zero3] stage3_gather_16bit_weights_on_model_save=false: saving ZeRO-3 shards "
                      "(use ckpt.deepspeed.zero_to_fp32 to consolidate)", flush=True)
            ds_ckpt.save_zero3_checkpoint(self.zmodel.engine, out, self.state.global_step, self.ds.raw if self.ds else {},
                                          scheduler=self.scheduler, persistent=self._persistent_named())
        if dist.is_initialized():
            dist.barrier()

    def _persistent_named(self):
        ids = {id(p): n for n, p in self.raw_model.named_parameters()}
        return [(p, ids.get(id(p), f"persistent.{i}")) for i, p in enumerate(self.zmodel.engine.persistent)]
