"""TEST INFRASTRUCTURE ONLY (oracle). The reference's optimizer step for the emitter's parameter groups, CPU:
Optimizers.optimizer_scaler_step_all (nerfstudio/engine/optimizers.py:122-137) = GradScaler.unscale_ ->
torch.nn.utils.clip_grad_norm_ -> clip_grad_value_ -> torch.optim.Adam.step, with the groups of
nerfstudio/models/nerfacto.py:313-317 and the hyper-parameters of nerfstudio/configs/method_configs.py:331-342.
The reference calls torch for every one of these, so the oracle does too (torch is part of the oracle);
GradScaler itself is CUDA-only, so unscale_/step are restated: grads / scale, skip the step when any is non-finite.
The schedule is the reference's own ExponentialDecayStepScheduler when /root/reference is importable (test only)."""
from __future__ import annotations

from typing import Dict

import torch


def group_of(name: str) -> str:
    return "proposal_networks" if name.startswith("proposal_networks.") else "fields"


class OracleOptimizers:
    def __init__(self, params: Dict[str, torch.Tensor], lr=1e-2, eps=1e-15, max_norm=0.1, max_value=0.1):
        self.params = {k: torch.nn.Parameter(v.detach().clone().float()) for k, v in params.items()}
        self.groups = {g: [p for k, p in self.params.items() if group_of(k) == g] for g in ("fields", "proposal_networks")}
        self.opt = {g: torch.optim.Adam(ps, lr=lr, eps=eps) for g, ps in self.groups.items()}
        self.max_norm, self.max_value = max_norm, max_value

    def set_lr(self, group: str, lr: float):
        for pg in self.opt[group].param_groups:
            pg["lr"] = lr

    def step(self, grads: Dict[str, torch.Tensor], grad_scale: float = 1.0, has_grads=None) -> Dict[str, bool]:
        """One optimizer_scaler_step_all.  ``has_grads`` = {group: bool}: a group none of whose parameters has a .grad is
        not stepped at all (optimizers.py:136).  Returns {group: stepped}: GradScaler keeps found_inf PER OPTIMIZER
        (torch/amp/grad_scaler.py, _per_optimizer_states), so a non-finite gradient skips only the step of its own group."""
        has_grads = has_grads or {}
        stepped = {}
        for g, ps in self.groups.items():
            if not has_grads.get(g, True):
                for p in ps:
                    p.grad = None
                stepped[g] = False
                continue
            for k, p in self.params.items():
                if group_of(k) == g:
                    p.grad = grads[k].detach().clone().float() / grad_scale          # unscale_
            finite = all(bool(torch.isfinite(p.grad).all()) for p in ps)
            if self.max_norm is not None:
                torch.nn.utils.clip_grad_norm_(ps, self.max_norm)
            if self.max_value is not None:
                torch.nn.utils.clip_grad_value_(ps, self.max_value)
            if finite:                                                       # GradScaler.step
                self.opt[g].step()
            stepped[g] = finite
        return stepped
