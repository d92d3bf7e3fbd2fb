"""apex.optimizers.FusedNovoGrad on sm_100a kernels.

Signature and defaults follow apex/optimizers/fused_novograd.py.  apex runs a per-tensor multi_tensor_l2norm over the
gradients, blends the norms into the layer-wise second moments (one float per tensor) and launches multi_tensor_novograd;
here: b200_l2norm_per_tensor (deterministic) -> b200_novograd_blend -> b200_fused_novograd over the device-resident
tensor tables, with the loss-scale unscale folded into the blend and the update (no fp32 master gradients under O2).
``state[p]["exp_avg_sq"]`` is a 0-d view of the set's moment array, so ``state_dict()`` carries it per parameter.
As for FusedLAMB, a dynamic loss scale's overflow flag is read on the host before the step (one synchronisation).
"""
from __future__ import annotations

import torch

from .. import ops
from ._fused_base import FusedOptimizerBase


class FusedNovoGrad(FusedOptimizerBase):
    _name = "FusedNovoGrad"

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.,
                 amsgrad=False, reg_inside_moment=False, grad_averaging=True, norm_type=2, init_zero=False,
                 set_grad_none=True):
        if amsgrad:
            raise RuntimeError("FusedNovoGrad does not support the AMSGrad variant.")
        if norm_type != 2:
            raise RuntimeError("FusedNovoGrad only support l2/inf norm now.  (libapex_b200: l2 only)")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        grad_averaging=grad_averaging, norm_type=norm_type, init_zero=init_zero)
        super().__init__(params, defaults)
        self.moment_mode = 0 if reg_inside_moment else 1
        self.set_grad_none = set_grad_none
        self._aux = {}

    def _state_columns(self, p, group):
        state = self.state[p]
        if "exp_avg" not in state:
            state["exp_avg"] = torch.zeros_like(p.data)
        if "exp_avg_sq" not in state:
            state["exp_avg_sq"] = torch.zeros((), dtype=torch.float32, device=p.device)
        return (state["exp_avg"],)

    def _table_built(self, key, s, table):
        dev = table.dev.device
        n = table.n_tensors
        # one contiguous moment array per launch set; the per-parameter state entries become views of it
        v = torch.zeros(max(n, 1), dtype=torch.float32, device=dev)
        by_ptr = {q.data_ptr(): q for g in self.param_groups for q in g["params"]}
        for i, t in enumerate(s["p"]):
            q = by_ptr[t.data_ptr()]
            v[i] = self.state[q]["exp_avg_sq"].to(dev)
            self.state[q]["exp_avg_sq"] = v[i]
        self._aux[key] = {"v": v, "first": table.first_tiles(),
                          "tile_ss": torch.empty(max(table.n_tiles, 1), dtype=torch.float32, device=dev),
                          "ss": torch.empty(max(n, 1), dtype=torch.float32, device=dev)}

    def _hyper(self, group_ids, steps):
        out = []
        for local in range(max(group_ids) + 1):
            gi = group_ids.get(local)
            if gi is None:
                out.append(ops.NovoGradGroup(0.0, 0.9, 0.999, 1e-8, 0.0, 2, 1, 0, 1, 1))
                continue
            g = self.param_groups[gi]
            beta1, beta2 = g["betas"]
            out.append(ops.NovoGradGroup(float(g["lr"]), float(beta1), float(beta2), float(g["eps"]), float(g["weight_decay"]),
                                         int(steps[gi]), 1 if g.get("bias_correction", True) else 0,
                                         1 if self.moment_mode == 0 else 0, 1 if g.get("grad_averaging", True) else 0,
                                         1 if g.get("init_zero", False) else 0))
        return out

    def _fused_step(self):
        plan = self._ensure_plan()
        st = self._amp_stash
        if st is not None and st.scaler is not None:
            st.scaler.note_step()
        if plan is None:
            return
        device = plan["device"]
        inv_scale, dev_inv, dev_found = self._scale_args()
        if dev_found is not None and int(dev_found.reshape(-1)[0].item()) != 0:
            return                                           # skipped step: state and step count untouched
        steps, _ = self._next_steps(device, False)
        with torch.cuda.device(device):
            for key, table, gids in plan["launches"]:
                gdt, sdt, cdt, _ = key
                a = self._aux[key]
                hyper = self._hyper(gids, steps)
                ops.l2norm_per_tensor(table, 0, gdt, a["first"], a["tile_ss"], a["ss"])
                ops.novograd_blend(table, hyper, a["v"], a["ss"], inv_scale=inv_scale, dev_inv_scale=dev_inv)
                ops.fused_novograd(table, gdt, sdt, cdt, hyper, a["v"], inv_scale=inv_scale, dev_inv_scale=dev_inv)

    def step(self, closure=None):
        loss = None
        if closure is not None:
            loss = closure()
        self._fused_step()
        return loss
