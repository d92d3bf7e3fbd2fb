"""Shared machinery of FusedAdam / FusedSGD: launch sets, table caching, amp hand-shake and the
exact step counter.

A *launch set* is the sub-list of parameters that one kernel launch can cover: same
(grad, state, copy) element types and at most ``B200_MAX_GROUPS`` param groups.  apex launches
once per (group, dtype) and then ceil(chunks/320) times inside each; here a
typical O2 model needs one launch for the fp16 parameters and one for the fp32 BatchNorm ones.
"""
from __future__ import annotations

import operator

import torch

from .. import _lib, ops
from ..tables import TensorTable, is_dense, layout_key

_SUPPORTED = (torch.float16, torch.bfloat16, torch.float32)
_GRAD_OF = operator.attrgetter("grad")
_DATA_PTR = torch.Tensor.data_ptr


class ListWatch:
    """Re-validation of a cached launch plan at a bounded host cost.

    A plan is right for as long as (a) every gradient is still the tensor OBJECT it was built
    from and (b) no watched storage (parameters, model copies) has moved.  Both are checked with
    C-level passes (``map`` over ``is`` / ``data_ptr``) — completely, every step, for lists of up
    to ``WINDOW`` tensors (every real model: GPT-2 medium has 292), and in rotating slices of
    ``WINDOW`` for larger ones, so a 34,000-tensor list costs the step ~0.1 ms instead of ~5 ms
    and a change the owner did not announce (``invalidate()``) is caught within
    ceil(n / WINDOW) steps.  Everything inside this package that replaces gradients or
    parameters announces it."""

    WINDOW = 1024
    __slots__ = ("holders", "grads", "tensors", "ptrs", "_g", "_t")

    def __init__(self, holders, grads, tensors=()):
        self.holders, self.grads = holders, grads
        self.tensors = list(tensors)
        self.ptrs = tuple(map(_DATA_PTR, self.tensors))
        self._g = self._t = 0

    def ok(self) -> bool:
        n, w = len(self.holders), self.WINDOW
        if n <= w:
            if not all(map(operator.is_, map(_GRAD_OF, self.holders), self.grads)):
                return False
        else:
            a = self._g
            b = min(n, a + w)
            if not all(map(operator.is_, map(_GRAD_OF, self.holders[a:b]), self.grads[a:b])):
                return False
            self._g = b if b < n else 0
        n = len(self.tensors)
        if n <= w:
            return self.ptrs == tuple(map(_DATA_PTR, self.tensors))
        a = self._t
        b = min(n, a + w)
        if self.ptrs[a:b] != tuple(map(_DATA_PTR, self.tensors[a:b])):
            return False
        self._t = b if b < n else 0
        return True


class _Never:
    """Holder whose .grad never matches: marks plans that must not be replayed."""
    grad = object()


_NEVER = _Never()


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
        self.accumulating = False       # a second backward arrived before step(): materialised path until the step
        self.version = 0                # bumped whenever amp itself replaces gradient objects


class FusedOptimizerBase(torch.optim.Optimizer):
    _name = "FusedOptimizer"

    def __init__(self, params, defaults):
        super().__init__(params, defaults)
        self._amp_stash = None
        self._tables = {}               # launch-set key -> TensorTable
        self._counter = None            # ops.StepCounter once device-side skipping is in play
        self._step_offsets = {}         # group index -> offset of the group's step from the counter
        self._plan = None               # cached launch plan (see _plan_is_current)
        self._version = 0               # bumped by everything in here that replaces gradients / parameters
        self._clip = None               # pending (unscale x clip) factor from clip_grad_norm_
        self._norm_buf = None

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
        self.invalidate()

    def invalidate(self):
        """Forget the cached launch plan (gradients or parameters were replaced)."""
        self._plan = None
        self._version += 1

    def add_param_group(self, param_group):
        super().add_param_group(param_group)
        if hasattr(self, "_plan"):
            self.invalidate()

    def zero_grad(self, set_to_none=None):
        if set_to_none is None:
            set_to_none = getattr(self, "set_grad_none", False)
        if set_to_none:
            for group in self.param_groups:
                for p in group["params"]:
                    p.grad = None
            self.invalidate()
        else:
            super().zero_grad(set_to_none=False)

    # ---- the step ------------------------------------------------------------------------------------
    def _grads_scaled_now(self):
        st = self._amp_stash
        return bool(st is not None and st.grads_scaled)

    def _plan_is_current(self):
        """Steady-state fast path.  When every gradient is the very tensor object the plan was built
        from (gradients accumulate in place, or live in DDP buckets) and no parameter storage moved,
        the cached tables are still right and the step is launched without touching the parameters
        one by one again (see ListWatch for what is checked and at what cost)."""
        plan = self._plan
        if plan is None:
            return False
        st = self._amp_stash
        if plan["version"] != (self._version, st.version if st is not None else 0):
            return False
        if plan["n_groups"] != len(self.param_groups) or plan["grads_scaled"] != self._grads_scaled_now():
            return False
        if plan["n_params"] != sum(map(len, map(operator.itemgetter("params"), self.param_groups))):
            return False
        return plan["watch"].ok()

    def _table_columns(self, s):
        """Columns of the step table of one dtype set: gradient, parameter, state..., model copy."""
        return [s["g"], s["p"]] + s["state"] + ([s["c"]] if s["c"] else [])

    def _table_built(self, key, s, table):
        """Hook: a dtype set got a new table (subclasses that launch more than one kernel per set build theirs here)."""

    def _ensure_plan(self):
        if self._plan_is_current():
            return self._plan
        self._plan = None
        sets = {}
        device = None
        holders, grad_objs, watch = [], [], []
        for gi, group in enumerate(self.param_groups):
            for p in group["params"]:
                g, copy = self._grad_and_copy(p)
                holder = p
                st_ = self._amp_stash
                if st_ is not None and p in st_.model_of and p.grad is None:
                    holder = st_.model_of[p]                 # the gradient lives on the fp16 model parameter
                holders.append(holder)
                grad_objs.append(g)
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
                if layout_key(g) != layout_key(p) or not is_dense(g):
                    # the gradient was produced in another memory layout than the parameter (e.g. a strided
                    # view, or NCHW for a channels_last weight): bring it into the parameter's element order
                    g = torch.empty_like(p.data, dtype=g.dtype).copy_(g)
                    grad_objs[-1] = None                     # a temporary copy: never replay this plan
                    holders[-1] = _NEVER
                s["g"].append(g)
                s["p"].append(p.data)
                watch.append(p)
                if copy is not None:
                    watch.append(copy)
                for lst, t in zip(s["state"], cols):
                    lst.append(t)
                if copy is not None:
                    s["c"].append(copy.data)
                s["grp"].append(gi % _lib.MAX_GROUPS)
                s["gids"][gi % _lib.MAX_GROUPS] = gi
        if not sets:
            return None
        launches = []
        for key, s in sets.items():
            columns = self._table_columns(s)
            # m / v / momentum are owned by this object (stable); grads, params and model copies
            # can be re-allocated behind our back, so their addresses form the cache key
            sig = (tuple(t.data_ptr() for t in s["g"]), tuple(t.data_ptr() for t in s["p"]),
                   tuple(t.data_ptr() for t in s["c"]))
            table = self._tables.get(key)
            if table is None or table.signature != sig:
                table = TensorTable(columns, groups=s["grp"], signature=sig)
                self._tables[key] = table
                self._table_built(key, s, table)
            launches.append((key, table, s["gids"]))
        st_ = self._amp_stash
        self._plan = {"launches": launches, "device": device,
                      "n_groups": len(self.param_groups), "n_params": sum(len(g["params"]) for g in self.param_groups),
                      "grads_scaled": self._grads_scaled_now(), "watch": ListWatch(holders, grad_objs, watch),
                      "version": (self._version, st_.version if st_ is not None else 0)}
        return self._plan

    def _fused_step(self):
        plan = self._ensure_plan()
        st = self._amp_stash
        if st is not None and st.scaler is not None:
            st.scaler.note_step()           # the pending skip decision is being consumed by this step
        if plan is None:
            return
        device = plan["device"]
        inv_scale, dev_inv, dev_found = self._scale_args()
        if self._clip is not None:                      # clip_grad_norm_ left a combined (unscale x clip) factor
            dev_inv = self._clip["scale"]
            self._clip = None
        steps, sync = self._next_steps(device, dev_found is not None)
        first = True
        with torch.cuda.device(device):
            for key, table, gids in plan["launches"]:
                this_sync = None
                if sync is not None:
                    this_sync = sync if first else ops.StepSync(sync.dev_counters, sync.mirror, sync.known_skips, 1)
                self._launch(table, key, gids, steps, (inv_scale, dev_inv, dev_found), this_sync)
                first = False

    # ---- gradient clipping fused into the step ---------------------------------------
    def clip_grad_norm_(self, max_norm: float):
        """Clip the global gradient norm like ``torch.nn.utils.clip_grad_norm_(amp.master_params(opt), max_norm)``
        — without materialising fp32 master gradients and without a scaling pass: ONE read-only pass
        over the (still loss-scaled) gradients computes the norm on the device, and the clip
        coefficient is folded into the factor the next ``step()`` kernel multiplies every gradient by.
        Returns the total (unscaled) norm as a 0-d device tensor; nothing synchronises."""
        plan = self._ensure_plan()
        if plan is None:
            return torch.zeros((), device="cuda")
        device = plan["device"]
        if self._norm_buf is None or self._norm_buf["sumsq"].device != device:
            self._norm_buf = {"sumsq": torch.zeros(1, device=device), "scratch": ops.new_norm_scratch(device),
                              "scale": torch.ones(1, device=device), "norm": torch.zeros(1, device=device)}
        nb = self._norm_buf
        inv_scale, dev_inv, _ = self._scale_args()
        with torch.cuda.device(device):
            for key, table, _gids in plan["launches"]:
                ops.multi_tensor_l2norm(table, 0, key[0], nb["sumsq"], nb["scratch"])
            ops.clip_scale(nb["sumsq"], max_norm, nb["scale"], nb["norm"], inv_scale=inv_scale, dev_inv_scale=dev_inv)
        self._clip = {"scale": nb["scale"], "max_norm": float(max_norm)}
        return nb["norm"][0]
