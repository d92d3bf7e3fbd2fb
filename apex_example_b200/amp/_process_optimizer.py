"""Optimizer patching for amp (apex/amp/_process_optimizer.py).

Two paths:

* fused (``apex.optimizers.FusedAdam`` / ``FusedSGD`` from this package): at ``scale_loss`` exit
  the gradients are only CHECKED (unless DDP's bucket epilogue already did); the unscale, the
  update and the fp32-master -> fp16-model copy happen in the optimizer's single kernel, which
  reads ``found_inf`` / ``inv_scale`` from the scaler's device record.  No host sync.
* generic (any torch optimizer, e.g. the reference's ``torch.optim.SGD`` at
  /root/reference/test_apex_distributed_spawn.py:119): apex's sequence — fp32 master grads are
  materialised by one multi_tensor_scale launch, the host reads the skip decision once, the
  optimizer steps on the masters, one more launch copies masters to the model.
"""
from __future__ import annotations

import operator
import types

import torch

from .. import ops
from ..optimizers._fused_base import AmpStash, FusedOptimizerBase
from ..tables import TensorTable, signature_of
from ._amp_state import _amp_state

_HALF_TYPES = (torch.float16, torch.bfloat16)
_GRAD_OF = operator.attrgetter("grad")


def _is_fused(optimizer) -> bool:
    return isinstance(optimizer, FusedOptimizerBase)


def _lazy(optimizer) -> bool:
    """True when master grads are NOT materialised at scale_loss exit (fused single-pass path)."""
    if not _is_fused(optimizer):
        return False
    if getattr(optimizer, "materialize_master_grads", False) and type(optimizer).__name__ == "FusedSGD":
        return False
    return optimizer._amp_stash.lazy_master_grads


def _master_params_to_model_params(optimizer):
    """fp32 master -> fp16/bf16 model weights, one launch (apex: multi_tensor_scale with scale 1.0)."""
    stash = optimizer._amp_stash
    if not stash.all_fp32_from_fp16_params:
        return
    masters = [p.data for p in stash.all_fp32_from_fp16_params]
    models = [p.data for p in stash.all_fp16_params]
    sig = signature_of([masters, models])
    tbl = getattr(stash, "_copy_table", None)
    if tbl is None or tbl.signature != sig:
        tbl = TensorTable([masters, models], signature=sig)
        stash._copy_table = tbl
    ops.multi_tensor_scale(tbl, torch.float32, models[0].dtype, scale=1.0)


def _prepare_amp_backward(self):
    stash = self._amp_stash
    if _lazy(self):
        return
    # apex: existing master grads are stashed so a second backward can be merged with axpby
    stash.stashed_master_grads = [p.grad for p in stash.all_fp32_from_fp16_params]
    for p in stash.all_fp32_from_fp16_params:
        p.grad = None
    if any(g is not None for g in stash.stashed_master_grads):
        return
    stash.stashed_master_grads = None


def _post_amp_backward(self, scaler):
    stash = self._amp_stash
    stash.scaler = scaler

    if _lazy(self):
        # Only an overflow CHECK happens here (the unscale is fused into the step kernel).  The list of
        # gradient tensors is cached and re-validated with one C-level identity pass, so the steady
        # state costs no per-tensor Python work (matters for lists of tens of thousands of tensors).
        plan = stash.__dict__.get("_check_plan")
        if plan is None or not all(map(operator.is_, map(_GRAD_OF, plan["holders"]), plan["grads"])):
            holders = stash.all_fp16_params + stash.all_fp32_params
            grads = [p.grad for p in holders]
            by_dtype = {}
            for g in grads:
                if g is not None:
                    by_dtype.setdefault(g.dtype, []).append(g)
            scaler._ensure_state()
            plan = {"holders": holders, "grads": grads,
                    "calls": [(TensorTable([gs]), dt) for dt, gs in by_dtype.items()],
                    "ids": frozenset(id(p) for p, g in zip(holders, grads) if g is not None)}
            stash._check_plan = plan
        checked = _amp_state.ddp_checked_backward
        if checked:
            # DDP's bucket epilogue raised the flag already — provided it saw every gradient we own
            checked = plan["ids"] <= getattr(_amp_state, "ddp_checked_ids", frozenset())
        if not checked:
            flag = scaler.overflow_ref()
            for tbl, dt in plan["calls"]:
                ops.multi_tensor_check(tbl, 0, dt, flag)
        stash.grads_scaled = True
        return

    model16 = [p for p in stash.all_fp16_params if p.grad is not None]
    fp32 = [p for p in stash.all_fp32_params if p.grad is not None]

    # materialised path -------------------------------------------------------------------------
    stash.grads_scaled = False
    if stash.all_fp32_from_fp16_params:
        pairs = [(m, p) for m, p in zip(stash.all_fp32_from_fp16_params, stash.all_fp16_params) if p.grad is not None]
        stashed = getattr(stash, "stashed_master_grads", None)
        if pairs:
            bufs = stash.__dict__.setdefault("_master_grad_bufs", {})
            new_master = []
            for m, p in pairs:
                b = bufs.get(m)
                if b is None or b.shape != m.shape:
                    b = bufs[m] = torch.empty_like(m.data)
                new_master.append(b)
            if stashed is None:
                scaler.unscale([p.grad for _, p in pairs], new_master)
                for (m, _), b in zip(pairs, new_master):
                    m.grad = b
            else:
                # gradient accumulation: master = stashed + model/scale  (apex multi_tensor_axpby)
                idx = {m: i for i, m in enumerate(stash.all_fp32_from_fp16_params)}
                have = [(m, p, stashed[idx[m]]) for m, p in pairs if stashed[idx[m]] is not None]
                fresh = [(m, p) for m, p in pairs if stashed[idx[m]] is None]
                if fresh:
                    fb = [bufs[m] for m, _ in fresh]
                    scaler.unscale([p.grad for _, p in fresh], fb)
                    for (m, _), b in zip(fresh, fb):
                        m.grad = b
                if have:
                    xs = [p.grad for _, p, _ in have]
                    ys = [s for _, _, s in have]
                    tbl = TensorTable([xs, ys, ys])
                    scaler._ensure_state()
                    inv = 1.0 / scaler.loss_scale()
                    ops.multi_tensor_axpby(tbl, xs[0].dtype, torch.float32, torch.float32, inv, 1.0, 0,
                                           scaler.overflow_ref())
                    for (m, _, s) in have:
                        m.grad = s
        stash.stashed_master_grads = None
    for dt in (torch.float32,) + _HALF_TYPES:
        grads = [p.grad for p in fp32 if p.grad.dtype == dt]
        scaler.unscale(grads, grads)            # in place


def _materialize_master_grads(self):
    """amp.master_params() hook: turn still-scaled fp16 model grads into unscaled fp32 master grads
    (what apex always does at scale_loss exit) so callers such as clip_grad_norm_ see them."""
    stash = self._amp_stash
    if stash is None or not stash.grads_scaled or stash.scaler is None:
        return
    scaler = stash.scaler
    pairs = [(m, p) for m, p in zip(stash.all_fp32_from_fp16_params, stash.all_fp16_params) if p.grad is not None]
    bufs = stash.__dict__.setdefault("_master_grad_bufs", {})
    outs = []
    for m, p in pairs:
        b = bufs.get(m)
        if b is None:
            b = bufs[m] = torch.empty_like(m.data)
        outs.append(b)
    if pairs:
        tbl = TensorTable([[p.grad for _, p in pairs], outs])
        ops.multi_tensor_scale(tbl, pairs[0][1].grad.dtype, torch.float32, dev_scale=scaler.inv_scale_ref())
        for (m, _), b in zip(pairs, outs):
            m.grad = b
    fp32 = [p.grad for p in stash.all_fp32_params if p.grad is not None]
    for dt in (torch.float32,) + _HALF_TYPES:
        gs = [g for g in fp32 if g.dtype == dt]
        if gs:
            tbl = TensorTable([gs, gs])
            ops.multi_tensor_scale(tbl, dt, dt, dev_scale=scaler.inv_scale_ref())
    stash.grads_scaled = False


def _process_optimizer(optimizer, properties):
    if getattr(optimizer, "_amp_stash", None) is not None:
        raise RuntimeError("A given optimizer should only be passed through amp.initialize once.")
    stash = AmpStash()
    optimizer._amp_stash = stash
    stash.params_have_scaled_gradients = False
    fused = _is_fused(optimizer)

    if properties.master_weights:
        for group in optimizer.param_groups:
            for i, param in enumerate(group["params"]):
                if not param.requires_grad:
                    continue
                if param.dtype in _HALF_TYPES:
                    master = param.detach().clone().float()
                    master.requires_grad = True
                    group["params"][i] = master
                    if param in optimizer.state:
                        optimizer.state[master] = optimizer.state.pop(param)
                    stash.model_of[master] = param
                    stash.all_fp16_params.append(param)
                    stash.all_fp32_from_fp16_params.append(master)
                elif param.dtype == torch.float32:
                    stash.all_fp32_params.append(param)
                else:
                    raise TypeError("Optimizer's parameters must be either torch.cuda.FloatTensor or "
                                    "torch.cuda.HalfTensor. Received {}".format(param.type()))
    else:
        for group in optimizer.param_groups:
            for param in group["params"]:
                if param.requires_grad:
                    stash.all_fp32_params.append(param)

    optimizer._prepare_amp_backward = types.MethodType(_prepare_amp_backward, optimizer)
    optimizer._post_amp_backward = types.MethodType(_post_amp_backward, optimizer)
    optimizer._amp_materialize_master_grads = types.MethodType(_materialize_master_grads, optimizer)

    old_step = optimizer.step
    old_zero_grad = optimizer.zero_grad

    def new_step(self, closure=None, *args, **kwargs):
        if closure is not None:
            raise RuntimeError("Currently, Amp does not support closure use with optimizers.")
        retval = old_step(*args, **kwargs)
        if not fused and properties.master_weights:
            _master_params_to_model_params(self)
        if properties.master_weights:
            # apex clears the master grads once they are consumed
            for p in stash.all_fp32_from_fp16_params:
                p.grad = None
        return retval

    def new_zero_grad(self, *args, **kwargs):
        set_none = bool(getattr(self, "set_grad_none", False))
        params = stash.all_fp16_params + stash.all_fp32_params
        if set_none:
            for p in params:
                p.grad = None
        else:
            grads = []
            for p in params:
                if p.grad is not None:
                    if p.grad.grad_fn is not None:
                        p.grad.detach_()
                    else:
                        p.grad.requires_grad_(False)      # views of the DDP buckets cannot be detached in place
                    grads.append(p.grad)
            if grads:
                torch._foreach_zero_(grads)
        for p in stash.all_fp32_from_fp16_params:
            p.grad = None
        stash.grads_scaled = False

    optimizer.step = types.MethodType(new_step, optimizer)
    optimizer.zero_grad = types.MethodType(new_zero_grad, optimizer)
    optimizer._amp_original_step = old_step
    optimizer._amp_original_zero_grad = old_zero_grad
    return optimizer
