"""apex.optimizers.FusedAdam on the sm_100a fused step kernel.

Signature and error behaviour follow apex/optimizers/fused_adam.py; the
reference reaches an optimizer's ``step`` at /root/reference/test_apex_distributed_spawn.py:164.
Under amp O2 one launch per dtype set does unscale + Adam/AdamW + fp32-master -> fp16-model copy
and is skipped on the device when the loss scaler saw an overflow (b200_fused_adam).
"""
from __future__ import annotations

import torch

from .. import ops
from ._fused_base import FusedOptimizerBase


class FusedAdam(FusedOptimizerBase):
    _name = "FusedAdam"

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, adam_w_mode=True,
                 weight_decay=0., amsgrad=False, set_grad_none=True):
        if amsgrad:
            raise RuntimeError("FusedAdam does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none

    def _state_columns(self, p, group):
        state = self.state[p]
        if len(state) == 0:
            # apex: zeros_like(p.data) -> fp16 moments for fp16 params (O3)
            state["exp_avg"] = torch.zeros_like(p.data)
            state["exp_avg_sq"] = torch.zeros_like(p.data)
        return state["exp_avg"], state["exp_avg_sq"]

    def _launch(self, table, key, group_ids, steps, scale_args, sync):
        gdt, sdt, cdt, _ = key
        hyper = []
        for local in range(max(group_ids) + 1):
            gi = group_ids.get(local)
            if gi is None:
                hyper.append(ops.AdamGroup(0.0, 0.9, 0.999, 1e-8, 0.0, 1, 1, self.adam_w_mode))
                continue
            g = self.param_groups[gi]
            beta1, beta2 = g["betas"]
            hyper.append(ops.AdamGroup(float(g["lr"]), float(beta1), float(beta2), float(g["eps"]),
                                       float(g["weight_decay"]), int(steps[gi]), 1 if g.get("bias_correction", True) else 0,
                                       self.adam_w_mode))
        inv_scale, dev_inv, dev_found = scale_args
        ops.fused_adam(table, gdt, sdt, cdt, hyper, inv_scale=inv_scale, dev_inv_scale=dev_inv,
                       dev_found_inf=dev_found, sync=sync)

    def step(self, closure=None, grads=None, output_params=None, scale=None, grad_norms=None):
        if any(a is not None for a in (grads, output_params, scale, grad_norms)):
            raise RuntimeError("FusedAdam has been updated.  Simply initialize it identically to torch.optim.Adam, "
                               "and call step() with no arguments.")
        loss = None
        if closure is not None:
            loss = closure()
        self._fused_step()
        return loss
