"""Shared machinery of FusedAdam / FusedSGD: launch sets, table caching, amp hand-shake and the
exact step counter.

A *launch set* is the sub-list of parameters that one kernel launch can cover: same
(grad, state, copy) element types and at most ``B200_MAX_GROUPS`` param groups.  apex launches
once per (group, dtype) and then ceil(chunks/320) times inside each; here a
typical O2 model needs one launch for the fp16 parameters and one for the fp32 BatchNorm ones.
"""
from __future__ import annotations

import torch

from .. import _lib, ops
from ..tables import TensorTable

_SUPPORTED = (torch.float16, torch.bfloat16, torch.float32)


class AmpStash:
    """What amp._process_optimizer leaves on an optimizer (apex calls it ``_amp_stash``)."""

    def __init__(self):
        self.model_of = {}              # fp32 master param -> fp16/bf16 model param (O2)
        self.scaler = None              # LossScaler whose device record the step kernels read
        self.grads_scaled = False       # True: .grad still carries the loss scale (fused unscale)
        self.lazy_master_grads = True   # False: materialise fp32 master grads at scale_loss exit
        self.all_fp16_params = []
        self.all_fp32_from_fp16_params = []
        self.all_fp32_params = []


class FusedOptimizerBase(torch.optim.Optimizer):
    _name = "FusedOptimizer"

    def __init__(self, params, defaults):
        super().__init__(params, defaults)
        self._amp_stash = None
        self._tables = {}               # launch-set key -> TensorTable
        self._counter = None            # ops.StepCounter once device-side skipping is in play
        self._step_offsets = {}         # group index -> offset of the group's step from the counter
        self._plain_state = None

    # ---- subclass interface ----------------------------------------------------------------------
    def _state_columns(self, p, group):
        """Return the state tensors of ``p`` in kernel column order (after grad and param)."""
        raise NotImplementedError

    def _launch(self, table, key, group_ids, steps, scale_args, sync):
        raise NotImplementedError

    # ---- helpers ------------------------------------------------------------------------------------
    def _grad_and_copy(self, p):
        st = self._amp_stash
        if st is not None and p in st.model_of:
            model_p = st.model_of[p]
            if p.grad is not None:              # materialised (already unscaled) fp32 master grad
                return p.grad, model_p
            return model_p.grad, model_p
        return p.grad, None

    def _scale_args(self):
        """(inv_scale, dev_inv_scale, dev_found_inf) for the step kernels."""
        st = self._amp_stash
        if st is None or st.scaler is None:
            return 1.0, None, None
        sc = st.scaler
        found = sc.found_inf_ref() if sc.dynamic else None
        if st.grads_scaled:
            return 1.0, sc.inv_scale_ref(), found
        return 1.0, None, found

    def _next_steps(self, device, skip_on_device):
        """Advance every group's step.  Returns (steps per group, first StepSync or None)."""
        groups = self.param_groups
        if not skip_on_device and self._counter is None:
            for g in groups:
                g["step"] = g.get("step", 0) + 1
            return [g["step"] for g in groups], None
        if self._counter is None:
            self._counter = ops.StepCounter(device)
        step_t, sync = self._counter.next()
        steps = []
        for gi, g in enumerate(groups):
            off = self._step_offsets.get(gi)
            if off is None:
                off = g.get("step", 0) - (step_t - 1)
                self._step_offsets[gi] = off
            g["step"] = max(step_t + off, 1)
            steps.append(g["step"])
        return steps, sync

    def _sync_steps(self):
        """Make group['step'] exact (one stream sync); used by state_dict()."""
        if self._counter is not None and self._counter.calls:
            t = self._counter.exact_steps()
            for gi, g in enumerate(self.param_groups):
                if gi in self._step_offsets:
                    g["step"] = t + self._step_offsets[gi]

    def state_dict(self):
        self._sync_steps()
        return super().state_dict()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._counter = None
        self._step_offsets = {}
        self._tables = {}

    def zero_grad(self, set_to_none=None):
        if set_to_none is None:
            set_to_none = getattr(self, "set_grad_none", False)
        if set_to_none:
            for group in self.param_groups:
                for p in group["params"]:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    # ---- the step ------------------------------------------------------------------------------------
    def _fused_step(self):
        sets = {}
        device = None
        for gi, group in enumerate(self.param_groups):
            for p in group["params"]:
                g, copy = self._grad_and_copy(p)
                if g is None:
                    continue
                if g.is_sparse:
                    raise RuntimeError(f"{self._name} does not support sparse gradients, "
                                       "please consider SparseAdam instead")
                if p.dtype not in _SUPPORTED or g.dtype not in _SUPPORTED:
                    raise RuntimeError(f"{self._name} only support fp16 and fp32.")
                if p.device.type != "cuda":
                    raise RuntimeError(f"{self._name} needs CUDA parameters: libapex_b200 has no CPU fallback")
                device = p.device
                key = (g.dtype, p.dtype, None if copy is None else copy.dtype, gi // _lib.MAX_GROUPS)
                s = sets.get(key)
                if s is None:
                    s = sets[key] = {"g": [], "p": [], "state": None, "c": [], "grp": [], "gids": {}}
                cols = self._state_columns(p, group)
                if s["state"] is None:
                    s["state"] = [[] for _ in cols]
                s["g"].append(g if g.is_contiguous() else g.contiguous())
                s["p"].append(p.data)
                for lst, t in zip(s["state"], cols):
                    lst.append(t)
                if copy is not None:
                    s["c"].append(copy.data)
                s["grp"].append(gi % _lib.MAX_GROUPS)
                s["gids"][gi % _lib.MAX_GROUPS] = gi
        if not sets:
            return
        inv_scale, dev_inv, dev_found = self._scale_args()
        steps, sync = self._next_steps(device, dev_found is not None)
        first = True
        with torch.cuda.device(device):
            for key, s in sets.items():
                columns = [s["g"], s["p"]] + s["state"] + ([s["c"]] if s["c"] else [])
                # m / v / momentum are owned by this object (stable); grads, params and model copies
                # can be re-allocated behind our back, so their addresses form the cache key
                sig = (tuple(t.data_ptr() for t in s["g"]), tuple(t.data_ptr() for t in s["p"]),
                       tuple(t.data_ptr() for t in s["c"]))
                table = self._tables.get(key)
                if table is None or table.signature != sig:
                    table = TensorTable(columns, groups=s["grp"], signature=sig)
                    self._tables[key] = table
                this_sync = None
                if sync is not None:
                    this_sync = sync if first else ops.StepSync(sync.dev_counters, sync.mirror, sync.known_skips, 1)
                self._launch(table, key, s["gids"], steps, (inv_scale, dev_inv, dev_found), this_sync)
                first = False
