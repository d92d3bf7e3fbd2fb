"""apex.optimizers.FusedLAMB on sm_100a kernels.

Signature and defaults follow apex/optimizers/fused_lamb.py.  apex runs: two multi_tensor_l2norm launches (fp32 and fp16
gradients) blended into a global gradient norm, then multi_tensor_lamb = stage 1 (moments + update), per-tensor norms
of the parameters and the updates, stage 2 (trust ratio applied).  Here the same stages run over the device-resident
tensor tables (b200_l2norm_per_tensor, b200_lamb_stage1, b200_lamb_stage2), all norms deterministic (no floating-point
atomics), and the loss-scale unscale is folded into stage 1, so under amp O2 no fp32 master gradients are made.

Difference from FusedAdam/FusedSGD here: with a dynamic loss scale the overflow flag is read on the host before the step
(one stream synchronisation per step — what apex's patched ``step`` does too), because the bias corrections are host
scalars and LAMB has no device-side step window.
"""
from __future__ import annotations

import torch

from .. import ops
from ..tables import TensorTable
from ._fused_base import FusedOptimizerBase


class FusedLAMB(FusedOptimizerBase):
    _name = "FusedLAMB"

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01,
                 amsgrad=False, adam_w_mode=True, grad_averaging=True, set_grad_none=True, max_grad_norm=1.0,
                 use_nvlamb=False):
        if amsgrad:
            raise RuntimeError("FusedLAMB does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        grad_averaging=grad_averaging, max_grad_norm=max_grad_norm)
        super().__init__(params, defaults)
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none
        self.use_nvlamb = use_nvlamb
        self._update = {}           # parameter -> fp32 update buffer (scratch between the stages, not state)
        self._aux = {}              # dtype-set key -> stage-2 table and norm buffers
        self._gsumsq = None

    def _state_columns(self, p, group):
        state = self.state[p]
        if len(state) == 0:
            state["exp_avg"] = torch.zeros_like(p.data)
            state["exp_avg_sq"] = torch.zeros_like(p.data)
        u = self._update.get(p)
        if u is None or u.shape != p.shape or u.device != p.device:
            u = self._update[p] = torch.zeros_like(p.data, dtype=torch.float32)
        return state["exp_avg"], state["exp_avg_sq"], u

    def _table_columns(self, s):
        return [s["g"], s["p"]] + s["state"]                 # g, p, m, v, u — the model copy belongs to stage 2

    def _table_built(self, key, s, table):
        u = s["state"][2]
        dev = table.dev.device
        # same tile size as the stage-1 table: the trust ratio's summation order is part of the contract
        t2 = TensorTable([u, s["p"]] + ([s["c"]] if s["c"] else []), groups=s["grp"], tile_elems=table.tile_elems)
        n = table.n_tensors
        self._aux[key] = {"table2": t2, "first": table.first_tiles(),
                          "tile_ss": torch.empty(max(table.n_tiles, 1), dtype=torch.float32, device=dev),
                          "ss": torch.empty(max(n, 1), dtype=torch.float32, device=dev),
                          "pn": torch.empty(max(n, 1), dtype=torch.float32, device=dev),
                          "un": torch.empty(max(n, 1), dtype=torch.float32, device=dev)}

    def _hyper(self, group_ids, steps):
        out = []
        for local in range(max(group_ids) + 1):
            gi = group_ids.get(local)
            if gi is None:
                out.append(ops.LambGroup(0.0, 0.9, 0.999, 1e-6, 0.0, 1, 1, self.adam_w_mode, 1, 0))
                continue
            g = self.param_groups[gi]
            beta1, beta2 = g["betas"]
            out.append(ops.LambGroup(float(g["lr"]), float(beta1), float(beta2), float(g["eps"]), float(g["weight_decay"]),
                                     int(steps[gi]), 1 if g.get("bias_correction", True) else 0, self.adam_w_mode,
                                     1 if g.get("grad_averaging", True) else 0, 1 if self.use_nvlamb else 0))
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
        max_norm = float(self.defaults["max_grad_norm"] or 0.0)
        with torch.cuda.device(device):
            if self._gsumsq is None or self._gsumsq.device != device:
                self._gsumsq = torch.zeros(1, dtype=torch.float32, device=device)
            self._gsumsq.zero_()
            for key, table, _gids in plan["launches"]:
                a = self._aux[key]
                ops.l2norm_per_tensor(table, 0, key[0], a["first"], a["tile_ss"], a["ss"], total=self._gsumsq)
            for key, table, gids in plan["launches"]:
                gdt, sdt, cdt, _ = key
                a = self._aux[key]
                hyper = self._hyper(gids, steps)
                ops.l2norm_per_tensor(table, 1, sdt, a["first"], a["tile_ss"], a["ss"], norms=a["pn"])
                ops.lamb_stage1(table, gdt, sdt, hyper, self._gsumsq, max_norm, inv_scale=inv_scale, dev_inv_scale=dev_inv)
                ops.l2norm_per_tensor(table, 4, torch.float32, a["first"], a["tile_ss"], a["ss"], norms=a["un"])
                ops.lamb_stage2(a["table2"], sdt, cdt, hyper, a["pn"], a["un"])

    def step(self, closure=None):
        loss = None
        if closure is not None:
            loss = closure()
        self._fused_step()
        return loss
