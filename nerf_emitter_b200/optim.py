"""Optimizer step of the emitter's two parameter groups on the GPU: the host-side mirror of
``Optimizers`` (NS/engine/optimizers.py:62-160) for the groups of NS/models/nerfacto.py:313-317 with the
sdf-nerfacto hyper-parameters (NS/configs/method_configs.py:331-342: Adam lr 1e-2, eps 1e-15, max_norm 0.1,
max_value 0.1, ExponentialDecayStepScheduler lr_final 1e-4, max_steps 20000, step_pretrain 2000, lr_lambda 0.01).  All arithmetic is
``nemit_adam_step`` (two multi-tensor kernel launches); this file only keeps step counters, the learning-rate
schedule (a scalar per step) and the Adam state tensors."""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import Dict, Optional

import torch

from . import lib as L
from .emitter import NerfEmitter


@dataclass
class AdamOptimizerConfig:
    """NS/engine/optimizers.py:33-59 (torch.optim.Adam defaults for the betas)."""
    lr: float = 1e-2
    eps: float = 1e-15
    betas: tuple = (0.9, 0.999)
    max_norm: Optional[float] = 0.1
    max_value: Optional[float] = 0.1


@dataclass
class ExponentialDecayStepSchedulerConfig:
    """NS/engine/schedulers.py:96-110,146-153."""
    lr_final: Optional[float] = 1e-4
    max_steps: int = 20000
    warmup_steps: int = 0
    lr_pre_warmup: float = 1e-8
    ramp: str = "cosine"
    step_pretrain: int = 0
    lr_lambda: float = 1.0

    @classmethod
    def sdf_nerfacto(cls):
        """The scheduler of both groups in the reference's sdf-nerfacto method (NS/configs/method_configs.py:331-342):
        after PRETRAIN_ITER = 2000 steps - i.e. in the phase Mitsuba has taken over, which is the phase this emitter path
        serves - the learning rate is cut to 1 % (lr_lambda)."""
        return cls(lr_final=1e-4, max_steps=20000, step_pretrain=PRETRAIN_ITER, lr_lambda=0.01)

    def multiplier(self, step: int, lr_init: float) -> float:
        """ExponentialDecayStepScheduler.get_scheduler.func (schedulers.py:169-190)."""
        lr_final = lr_init if self.lr_final is None else self.lr_final
        if lr_init == 0:
            return 1.0
        if step < self.warmup_steps:
            if self.ramp == "cosine":
                lr = self.lr_pre_warmup + (lr_init - self.lr_pre_warmup) * math.sin(
                    0.5 * math.pi * min(max(step / self.warmup_steps, 0.0), 1.0))
            else:
                lr = self.lr_pre_warmup + (lr_init - self.lr_pre_warmup) * step / self.warmup_steps
        else:
            t = min(max((step - self.warmup_steps) / (self.max_steps - self.warmup_steps), 0.0), 1.0)
            lr = math.exp(math.log(lr_init) * (1 - t) + math.log(lr_final) * t)
        res = lr / lr_init
        if step + 1 >= self.step_pretrain:
            res *= self.lr_lambda
        return res


PRETRAIN_ITER = 2000          # NS/configs/method_configs.py:288
GROUPS = ("fields", "proposal_networks")


def sdf_nerfacto_optimizers() -> Dict[str, Dict]:
    """method_configs["sdf-nerfacto"].optimizers (NS/configs/method_configs.py:331-342)."""
    return {g: {"optimizer": AdamOptimizerConfig(lr=1e-2, eps=1e-15, max_norm=0.1, max_value=0.1),
                "scheduler": ExponentialDecayStepSchedulerConfig.sdf_nerfacto()} for g in GROUPS}


class EmitterOptimizers:
    """``Optimizers`` for the emitter's parameters: ``optimizer_scaler_step_all`` / ``optimizer_step_all`` /
    ``scheduler_step_all`` / ``zero_grad_all`` with the reference's semantics, executed by the library on the
    tensors the emitter has bound (parameters and gradients are updated in place)."""

    def __init__(self, emitter: NerfEmitter, config: Optional[Dict[str, Dict]] = None):
        self.E = emitter
        # default: the reference's sdf-nerfacto hyper-parameters, INCLUDING the 100x learning-rate cut after the
        # pre-training phase (step_pretrain = 2000, lr_lambda = 0.01)
        config = config or sdf_nerfacto_optimizers()
        self.config = config
        if not emitter.grads:
            emitter.bind_grads()
        # Adam state in the parameter layouts, one flat buffer each (zero-initialised like torch's lazy state)
        self.state = {}
        for key in ("exp_avg", "exp_avg_sq"):
            flat = torch.zeros(sum((p.numel() + 63) // 64 * 64 for p in emitter.params.values()), device=emitter.device)
            views, off = {}, 0
            for k, p in emitter.params.items():
                views[k] = flat[off:off + p.numel()].view_as(p)
                off += (p.numel() + 63) // 64 * 64
            self.state[key] = views
            self.state[key + "_flat"] = flat
        self._m = emitter._pack(self.state["exp_avg"])
        self._v = emitter._pack(self.state["exp_avg_sq"])
        # applied optimizer steps per group (Adam's `step`), kept ON THE DEVICE: a step skipped by the non-finite check
        # or for lack of gradients must not advance it (torch Adam only counts optimizer.step() calls), and finding
        # that out on the host would cost a synchronisation per iteration
        self.step_state = torch.zeros(2, dtype=torch.int64, device=emitter.device)
        self.sched_step = 0                      # LambdaLR's last_epoch
        self.stats = torch.zeros(4, device=emitter.device)

    @property
    def step_count(self):
        """Applied steps of (fields, proposal_networks); reads the device counters (synchronises)."""
        return tuple(int(x) for x in self.step_state.tolist())

    def lr(self, group: str) -> float:
        cfg = self.config[group]
        lr0 = cfg["optimizer"].lr
        return lr0 * cfg["scheduler"].multiplier(self.sched_step, lr0) if "scheduler" in cfg else lr0

    def _group(self, name: str, enabled: bool = True) -> L.NemitAdamGroup:
        o = self.config[name]["optimizer"]
        g = L.NemitAdamGroup()
        g.enabled = int(bool(enabled))
        g.lr, g.beta1, g.beta2, g.eps = self.lr(name), o.betas[0], o.betas[1], o.eps
        g.max_norm = o.max_norm if o.max_norm is not None else 0.0
        g.max_value = o.max_value if o.max_value is not None else 0.0
        return g

    def optimizer_scaler_step_all(self, grad_scale: float = 1.0, zero_grads: bool = False,
                                  proposal_has_grads: bool = True) -> None:
        """optimizers.py:122-137 with ``grad_scale`` = GradScaler.get_scale().  Per group, as the reference keeps one
        optimizer and one GradScaler ``found_inf`` record per group: a group with a non-finite unscaled gradient skips ITS
        step on the device (``self.stats[2 + g]`` = 1) without advancing its step count; ``proposal_has_grads=False``
        (no backward chunk of this iteration ran with ``proposal_requires_grad`` - with sdf-nerfacto's
        ``proposal_update_every=1`` that is every other iteration, the ``.grad``s being ``None`` after ``zero_grad``)
        leaves the proposal networks' parameters, moments and step count untouched (optimizers.py:136)."""
        E = self.E
        gf, gp = self._group("fields"), self._group("proposal_networks", proposal_has_grads)
        L.check(E.lib.nemit_adam_step(E._h, C.byref(self._m), C.byref(self._v), C.byref(gf), C.byref(gp),
                                      C.c_void_p(self.step_state.data_ptr()), float(grad_scale), int(zero_grads),
                                      C.c_void_p(self.stats.data_ptr()), E._stream()),
                "nemit_adam_step")

    def optimizer_step_all(self, zero_grads: bool = False, proposal_has_grads: bool = True) -> None:
        """optimizers.py:139-149 (no GradScaler)."""
        self.optimizer_scaler_step_all(1.0, zero_grads, proposal_has_grads)

    def scheduler_step_all(self, step: int = 0) -> None:
        self.sched_step += 1

    def zero_grad_all(self) -> None:
        self.E.zero_grads()

    def found_inf(self, group: Optional[str] = None) -> bool:
        """Host read (synchronises): what GradScaler.update consumes - any group by default, or one group's record."""
        flags = self.stats[2:4].tolist()
        if group is None:
            return bool(flags[0] != 0.0 or flags[1] != 0.0)
        return bool(flags[GROUPS.index(group)] != 0.0)

    def state_dict(self):
        return {"step": list(self.step_count), "sched_step": self.sched_step,
                "exp_avg": {k: v.clone() for k, v in self.state["exp_avg"].items()},
                "exp_avg_sq": {k: v.clone() for k, v in self.state["exp_avg_sq"].items()}}

    def load_state_dict(self, sd) -> None:
        steps = sd["step"] if isinstance(sd["step"], (list, tuple)) else [sd["step"]] * 2
        self.step_state.copy_(torch.tensor([int(x) for x in steps], dtype=torch.int64))
        self.sched_step = int(sd["sched_step"])
        for key in ("exp_avg", "exp_avg_sq"):
            for k, v in sd[key].items():
                self.state[key][k].copy_(v)
