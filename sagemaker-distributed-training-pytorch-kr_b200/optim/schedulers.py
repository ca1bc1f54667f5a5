"""Learning-rate / weight-decay schedulers.

* ``OptimizerParamScheduler`` - Megatron's: linear warmup then {constant, linear, cosine,
  inverse-square-root} decay to ``min_lr``, weight-decay ramp {constant, linear, cosine}; counted in
  iterations or samples; checkpointable (megatron/optimizer_param_scheduler.py; flags
  arguments.py:876-919, 693-699; LR trace 3_...ipynb:4347-4360).
* ``WarmupDecayLR`` / ``WarmupLR`` - DeepSpeed's (configs/default_offload_opt_param.json:14-22).
"""
from __future__ import annotations

import math


class OptimizerParamScheduler:
    def __init__(self, optimizer, max_lr, min_lr, lr_warmup_steps, lr_decay_steps, lr_decay_style, start_wd, end_wd,
                 wd_incr_steps, wd_incr_style, use_checkpoint_opt_param_scheduler=True, override_opt_param_scheduler=False):
        self.optimizer = optimizer
        self.max_lr, self.min_lr = float(max_lr), float(min_lr)
        assert 0.0 <= self.min_lr <= self.max_lr
        self.lr_warmup_steps, self.lr_decay_steps, self.lr_decay_style = lr_warmup_steps, lr_decay_steps, lr_decay_style
        assert self.lr_decay_steps > 0 and self.lr_warmup_steps < self.lr_decay_steps
        self.start_wd, self.end_wd, self.wd_incr_steps, self.wd_incr_style = start_wd, end_wd, wd_incr_steps, wd_incr_style
        assert 0.0 <= self.start_wd <= self.end_wd
        self.override, self.use_ckpt = override_opt_param_scheduler, use_checkpoint_opt_param_scheduler
        assert not (self.override and self.use_ckpt), "both override and use-checkpoint are set."
        self.num_steps = 0
        self.step(0)

    def get_wd(self) -> float:
        if self.num_steps > self.wd_incr_steps:
            return self.end_wd
        if self.wd_incr_style == "constant":
            assert self.start_wd == self.end_wd
            return self.end_wd
        ratio = self.num_steps / self.wd_incr_steps
        if self.wd_incr_style == "linear":
            coeff = ratio
        elif self.wd_incr_style == "cosine":
            coeff = 0.5 * (math.cos(math.pi * (1 - ratio)) + 1.0)
        else:
            raise ValueError(f"{self.wd_incr_style} weight decay increment style is not supported.")
        return self.start_wd + coeff * (self.end_wd - self.start_wd)

    def get_lr(self) -> float:
        if self.lr_warmup_steps > 0 and self.num_steps <= self.lr_warmup_steps:
            return self.max_lr * self.num_steps / self.lr_warmup_steps
        if self.lr_decay_style == "constant":
            return self.max_lr
        if self.num_steps > self.lr_decay_steps:
            return self.min_lr
        if self.lr_decay_style == "inverse-square-root":
            w = max(self.lr_warmup_steps, 1)
            n = max(self.num_steps, 1)
            return max(self.min_lr, self.max_lr * w ** 0.5 / n ** 0.5)
        ratio = (self.num_steps - self.lr_warmup_steps) / (self.lr_decay_steps - self.lr_warmup_steps)
        delta = self.max_lr - self.min_lr
        if self.lr_decay_style == "linear":
            coeff = 1.0 - ratio
        elif self.lr_decay_style == "cosine":
            coeff = 0.5 * (math.cos(math.pi * ratio) + 1.0)
        else:
            raise ValueError(f"{self.lr_decay_style} decay style is not supported.")
        return self.min_lr + coeff * delta

    def step(self, increment: int):
        self.num_steps += increment
        lr, wd = self.get_lr(), self.get_wd()
        for g in self.optimizer.param_groups:
            g["lr"] = lr * g.get("lr_mult", 1.0)
            g["weight_decay"] = wd * g.get("wd_mult", 1.0)

    def state_dict(self):
        return {k: getattr(self, k) for k in ("max_lr", "lr_warmup_steps", "num_steps", "lr_decay_style", "lr_decay_steps", "min_lr",
                                              "start_wd", "end_wd", "wd_incr_style", "wd_incr_steps")}

    def _pick(self, cls_value, sd_value, name):
        if self.override:
            return cls_value
        if not self.use_ckpt:
            assert cls_value == sd_value, f"OptimizerParamScheduler: class input value {cls_value} and checkpoint value {sd_value} for {name} do not match"
        return sd_value

    def load_state_dict(self, sd):
        self.max_lr = self._pick(self.max_lr, sd.get("max_lr", sd.get("start_lr")), "learning rate")
        self.min_lr = self._pick(self.min_lr, sd["min_lr"], "minimum learning rate")
        self.lr_warmup_steps = self._pick(self.lr_warmup_steps, sd.get("lr_warmup_steps", sd.get("warmup_iter", sd.get("warmup_steps"))), "warmup iterations")
        self.lr_decay_steps = self._pick(self.lr_decay_steps, sd.get("lr_decay_steps", sd.get("end_iter", sd.get("decay_steps"))), "total number of iterations")
        self.lr_decay_style = self._pick(self.lr_decay_style, sd.get("lr_decay_style", sd.get("decay_style")), "learning rate decay style")
        if "start_wd" in sd:
            self.start_wd = self._pick(self.start_wd, sd["start_wd"], "start weight decay")
            self.end_wd = self._pick(self.end_wd, sd["end_wd"], "end weight decay")
            self.wd_incr_steps = self._pick(self.wd_incr_steps, sd["wd_incr_steps"], "total number of weight decay iterations")
            self.wd_incr_style = self._pick(self.wd_incr_style, sd["wd_incr_style"], "weight decay incr style")
        self.num_steps = 0
        self.step(sd.get("num_steps", sd.get("num_iters", 0)))


class WarmupLR:
    """DeepSpeed WarmupLR: log or linear warmup from warmup_min_lr to warmup_max_lr, then constant."""

    def __init__(self, optimizer, warmup_min_lr=0.0, warmup_max_lr=1e-3, warmup_num_steps=1000, warmup_type="log", last_batch_iteration=-1):
        self.optimizer = optimizer
        self.min_lr, self.max_lr = warmup_min_lr, warmup_max_lr
        self.warmup_num_steps = max(2, int(warmup_num_steps))
        self.warmup_type = warmup_type
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        self.last_batch_iteration = last_batch_iteration
        if last_batch_iteration == -1:
            self._apply(self.get_lr())

    def _gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            if self.warmup_type == "log":
                return self.inverse_log_warm_up * math.log(self.last_batch_iteration + 1)
            return self.last_batch_iteration / self.warmup_num_steps
        return 1.0

    def get_lr(self):
        if self.last_batch_iteration < 0:
            return self.min_lr
        return self.min_lr + (self.max_lr - self.min_lr) * self._gamma()

    def _apply(self, lr):
        for g in self.optimizer.param_groups:
            g["lr"] = lr

    def step(self, last_batch_iteration=None):
        self.last_batch_iteration = self.last_batch_iteration + 1 if last_batch_iteration is None else last_batch_iteration
        self._apply(self.get_lr())

    def get_last_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups]

    def state_dict(self):
        return {"last_batch_iteration": self.last_batch_iteration}

    def load_state_dict(self, sd):
        self.last_batch_iteration = sd["last_batch_iteration"]
        self._apply(self.get_lr())


class WarmupDecayLR(WarmupLR):
    """DeepSpeed WarmupDecayLR: warmup as above, then linear decay to warmup_min_lr at total_num_steps."""

    def __init__(self, optimizer, total_num_steps, warmup_min_lr=0.0, warmup_max_lr=1e-3, warmup_num_steps=1000, warmup_type="log", last_batch_iteration=-1):
        self.total_num_steps = total_num_steps
        super().__init__(optimizer, warmup_min_lr, warmup_max_lr, warmup_num_steps, warmup_type, last_batch_iteration)

    def _gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            return super()._gamma()
        return max(0.0, (self.total_num_steps - self.last_batch_iteration) / max(1.0, self.total_num_steps - self.warmup_num_steps))


class HFScheduler:
    """transformers.get_scheduler(name, optimizer, num_warmup_steps, num_training_steps) for the case the DeepSpeed JSON has no
    `scheduler` block (then the HF Trainer builds its own from --lr_scheduler_type / --warmup_ratio / --warmup_steps):
    linear | cosine | constant | constant_with_warmup | polynomial (power 1 = linear) | inverse_sqrt.  Multiplies the base
    learning rate of every param group; step() after each optimizer step."""

    def __init__(self, optimizer, name: str, num_warmup_steps: int, num_training_steps: int, last_step: int = 0):
        assert name in ("linear", "cosine", "constant", "constant_with_warmup", "polynomial", "inverse_sqrt"), f"unsupported lr_scheduler_type {name}"
        self.optimizer, self.name, self.warmup, self.total = optimizer, name, int(num_warmup_steps), int(num_training_steps)
        self.base_lrs = [g["lr"] for g in optimizer.param_groups]
        self.last_step = last_step
        self._apply()

    def factor(self, step: int) -> float:
        w, t = self.warmup, self.total
        if self.name != "constant" and step < w:
            return step / max(1, w)
        if self.name in ("constant", "constant_with_warmup"):
            return 1.0
        if self.name in ("linear", "polynomial"):
            return max(0.0, (t - step) / max(1, t - w))
        if self.name == "cosine":
            prog = (step - w) / max(1, t - w)
            return max(0.0, 0.5 * (1.0 + math.cos(math.pi * prog)))
        shift = max(w, 1)                                   # inverse_sqrt
        return math.sqrt(shift) / math.sqrt(max(step, shift))

    def _apply(self):
        f = self.factor(self.last_step)
        for g, b in zip(self.optimizer.param_groups, self.base_lrs):
            g["lr"] = b * f

    def step(self):
        self.last_step += 1
        self._apply()

    def get_last_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups]

    def state_dict(self):
        return {"last_step": self.last_step, "base_lrs": list(self.base_lrs)}

    def load_state_dict(self, sd):
        self.last_step, self.base_lrs = sd["last_step"], list(sd["base_lrs"])
        self._apply()
