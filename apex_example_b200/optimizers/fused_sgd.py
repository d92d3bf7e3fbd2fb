"""apex.optimizers.FusedSGD on the sm_100a fused step kernel.

Signature and semantics follow apex/optimizers/fused_sgd.py:
``wd_after_momentum``, ``materialize_master_grads``, first-run momentum initialisation.  With
``materialize_master_grads=False`` under amp O2 the kernel reads the fp16 model gradients,
applies 1/loss_scale itself and writes the fp16 model copy — apex's "N=4" case — in one launch
for the whole list (b200_fused_sgd).
"""
from __future__ import annotations

import torch
from torch.optim.optimizer import required

from .. import ops
from ._fused_base import FusedOptimizerBase


class FusedSGD(FusedOptimizerBase):
    _name = "FusedSGD"

    def __init__(self, params, lr=required, momentum=0, dampening=0, weight_decay=0, nesterov=False,
                 wd_after_momentum=False, materialize_master_grads=True, set_grad_none=False):
        if lr is not required and lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if momentum < 0.0:
            raise ValueError("Invalid momentum value: {}".format(momentum))
        if weight_decay < 0.0:
            raise ValueError("Invalid weight_decay value: {}".format(weight_decay))
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov)
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, defaults)
        self.wd_after_momentum = wd_after_momentum
        self.materialize_master_grads = materialize_master_grads
        self.most_recent_scale = 1.0
        self.scale_set_by_backward = False
        self.set_grad_none = set_grad_none

    def __setstate__(self, state):
        super().__setstate__(state)
        for group in self.param_groups:
            group.setdefault("nesterov", False)

    def load_state_dict(self, state_dict):
        """Checkpoints written by torch.optim.SGD or apex's FusedSGD carry ``momentum_buffer`` in the
        state but no step count in the param groups.  apex keys ``first_run`` on whether the buffers
        exist; here it is keyed on the group's step, so a group that arrives with buffers and without
        a step is marked as having stepped (otherwise the first step would overwrite the loaded
        momentum with the gradient)."""
        super().load_state_dict(state_dict)
        for group in self.param_groups:
            if "step" not in group and any("momentum_buffer" in self.state.get(p, {}) for p in group["params"]):
                group["step"] = 1

    def _state_columns(self, p, group):
        if group["momentum"] == 0:
            return (None,)
        state = self.state[p]
        if "momentum_buffer" not in state:
            # contents are ignored on the group's first successful step (first_run: mom = grad)
            state["momentum_buffer"] = torch.zeros_like(p.data)
        return (state["momentum_buffer"],)

    def _launch(self, table, key, group_ids, steps, scale_args, sync):
        gdt, sdt, cdt, _ = key
        hyper = []
        for local in range(max(group_ids) + 1):
            gi = group_ids.get(local)
            if gi is None:
                hyper.append(ops.SgdGroup(0.0, 0.0, 0.0, 0.0, 2, 0, 0))
                continue
            g = self.param_groups[gi]
            hyper.append(ops.SgdGroup(float(g["lr"]), float(g["momentum"]), float(g["dampening"]),
                                      float(g["weight_decay"]), int(steps[gi]), 1 if g["nesterov"] else 0,
                                      1 if self.wd_after_momentum else 0))
        inv_scale, dev_inv, dev_found = scale_args
        ops.fused_sgd(table, gdt, sdt, cdt, hyper, inv_scale=inv_scale, dev_inv_scale=dev_inv,
                      dev_found_inf=dev_found, sync=sync)

    def step(self, closure=None):
        loss = None
        if closure is not None:
            loss = closure()
        self._fused_step()
        return loss
