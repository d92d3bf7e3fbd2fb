"""Optimizer patching for amp (apex/amp/_process_optimizer.py).

Two paths:

* fused: ``apex.optimizers.FusedAdam`` / ``FusedSGD`` from this package, and — new — a stock
  ``torch.optim.SGD`` / ``Adam`` / ``AdamW`` (the reference constructs ``torch.optim.SGD`` at
  /root/reference/test_apex_distributed_spawn.py:119) whose options the kernels cover: such an
  optimizer is *adopted* (see ``_adopt_torch_optimizer``), i.e. its ``step`` is served by the same
  single-launch kernel out of its own ``param_groups`` / ``state``.  At ``scale_loss`` exit the
  gradients are only CHECKED (unless DDP's bucket epilogue already did); the unscale, the update
  and the fp32-master -> fp16-model copy happen in the optimizer's kernel, which reads
  ``found_inf`` / ``inv_scale`` from the scaler's device record.  No host sync.
* generic (any other optimizer): apex's sequence — fp32 master grads are materialised by one
  multi_tensor_scale launch, the host reads the skip decision once, the optimizer steps on the
  masters, one more launch copies masters to the model.

Gradient accumulation (several ``scale_loss`` blocks before one ``step()``) follows apex's
``prepare_backward_*`` / ``post_backward_*``: at entry the model gradients are dropped and the
fp32 ones stashed, at exit the fresh gradients are unscaled and merged into the stash with
``multi_tensor_axpby``.  The fused path switches to this materialised sequence from the second
backward of a cycle on (the first one's gradients are folded into fp32 master grads first).
"""
from __future__ import annotations

import os
import types

import torch

from .. import ops
from ..optimizers._fused_base import AmpStash, FusedOptimizerBase, ListWatch
from ..tables import TensorTable, signature_of
from ._amp_state import _amp_state

_HALF_TYPES = (torch.float16, torch.bfloat16)


def _engine(optimizer):
    """The object whose fused kernels serve ``optimizer.step`` (the optimizer itself, or the engine
    that adopted a stock torch optimizer), else None."""
    if isinstance(optimizer, FusedOptimizerBase):
        return optimizer
    return getattr(optimizer, "_b200_engine", None)


def _is_fused(optimizer) -> bool:
    return _engine(optimizer) is not None


def _lazy(optimizer) -> bool:
    """True when master grads are NOT materialised at scale_loss exit (fused single-pass path)."""
    eng = _engine(optimizer)
    if eng is None:
        return False
    if getattr(eng, "materialize_master_grads", False) and type(eng).__name__ == "FusedSGD":
        return False
    stash = optimizer._amp_stash
    return stash.lazy_master_grads and not stash.accumulating


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
    """scale_loss entry (apex prepare_backward_with_master_weights / _no_master_weights)."""
    stash = self._amp_stash
    if _lazy(self):
        if not stash.grads_scaled:
            return
        # A backward is about to ADD to gradients that still carry the previous block's loss scale
        # (no step() / zero_grad() in between, no delay_unscale).  Fold what is there into unscaled
        # fp32 master grads — what apex does at every exit — and stay on the materialised sequence
        # until the step, so each micro-batch is unscaled with the scale it was produced under.
        self._amp_materialize_master_grads()
        stash.accumulating = True
    # apex: fp16 model grads are dropped (the master grads carry the running sum); fp32 grads are
    # stashed and dropped, so this backward's contribution can be unscaled on its own and merged
    for p in stash.all_fp16_params:
        p.grad = None
    stash.stashed_fp32_grads = [p.grad for p in stash.all_fp32_params]
    for p in stash.all_fp32_params:
        p.grad = None
    stash.version += 1
    if not any(g is not None for g in stash.stashed_fp32_grads):
        stash.stashed_fp32_grads = None


def _bufs_for(stash, masters):
    bufs = stash.__dict__.setdefault("_master_grad_bufs", {})
    out = []
    for m in masters:
        b = bufs.get(m)
        if b is None or b.shape != m.shape:
            b = bufs[m] = torch.empty_like(m.data)
        out.append(b)
    return out


def _axpby_into(scaler, xs, ys):
    """ys += xs / current_scale, raising the overflow flag on non-finite xs (apex unscale_with_stashed)."""
    scaler._ensure_state()
    inv = 1.0 / scaler.loss_scale()
    for dt in {x.dtype for x in xs}:
        for ydt in {y.dtype for y in ys}:
            pair = [(x, y) for x, y in zip(xs, ys) if x.dtype == dt and y.dtype == ydt]
            if not pair:
                continue
            xx, yy = [x for x, _ in pair], [y for _, y in pair]
            ops.multi_tensor_axpby(TensorTable([xx, yy, yy]), dt, ydt, ydt, inv, 1.0, 0, scaler.overflow_ref())


def _post_amp_backward(self, scaler):
    stash = self._amp_stash
    stash.scaler = scaler

    if _lazy(self):
        # Only an overflow CHECK happens here (the unscale is fused into the step kernel).  The list of
        # gradient tensors is cached and re-validated at a bounded cost (ListWatch), so the steady
        # state costs no per-tensor Python work (matters for lists of tens of thousands of tensors).
        plan = stash.__dict__.get("_check_plan")
        eng = _engine(self)
        key = (stash.version, eng._version)
        if plan is None or plan["key"] != key or not plan["watch"].ok():
            holders = stash.all_fp16_params + stash.all_fp32_params
            grads = [p.grad for p in holders]
            by_dtype = {}
            for g in grads:
                if g is not None:
                    by_dtype.setdefault(g.dtype, []).append(g)
            scaler._ensure_state()
            plan = {"watch": ListWatch(holders, grads), "key": key,
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

    # materialised path (apex post_backward_with_master_weights / post_backward_models_are_masters) ----
    stash.grads_scaled = False
    if stash.all_fp32_from_fp16_params:
        fresh_x, fresh_m, merge_x, merge_y = [], [], [], []
        for m, p in zip(stash.all_fp32_from_fp16_params, stash.all_fp16_params):
            if p.grad is None:
                continue                        # unused in this backward; a running sum (if any) stays on m.grad
            if m.grad is None:
                fresh_x.append(p.grad)
                fresh_m.append(m)
            else:
                merge_x.append(p.grad)
                merge_y.append(m.grad)
        if fresh_x:
            outs = _bufs_for(stash, fresh_m)
            scaler.unscale(fresh_x, outs)
            for m, b in zip(fresh_m, outs):
                m.grad = b
        if merge_x:
            _axpby_into(scaler, merge_x, merge_y)

    stashed = getattr(stash, "stashed_fp32_grads", None) or [None] * len(stash.all_fp32_params)
    plain, merges = [], []
    for p, s in zip(stash.all_fp32_params, stashed):
        if p.grad is None:
            if s is not None:
                p.grad = s                      # unused in this backward: the running sum stays
        elif s is None:
            plain.append(p.grad)
        else:
            merges.append((p, s))
    for dt in (torch.float32,) + _HALF_TYPES:
        grads = [g for g in plain if g.dtype == dt]
        scaler.unscale(grads, grads)            # in place
    if merges:
        _axpby_into(scaler, [p.grad for p, _ in merges], [s for _, s in merges])
        for p, s in merges:
            p.grad = s                          # the running sum lives in the stashed tensor
    stash.stashed_fp32_grads = None
    stash.version += 1


def _materialize_master_grads(self):
    """amp.master_params() hook: turn still-scaled fp16 model grads into unscaled fp32 master grads
    (what apex always does at scale_loss exit) so callers such as clip_grad_norm_ see them."""
    stash = self._amp_stash
    if stash is None or not stash.grads_scaled or stash.scaler is None:
        return
    scaler = stash.scaler
    pairs = [(m, p) for m, p in zip(stash.all_fp32_from_fp16_params, stash.all_fp16_params) if p.grad is not None]
    if pairs:
        outs = _bufs_for(stash, [m for m, _ in pairs])
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
    stash.version += 1


# ---------------------------------------------------------------------------------------------------
# stock torch optimizers on the fused kernels
# ---------------------------------------------------------------------------------------------------
def _adoptable(optimizer) -> str | None:
    """'sgd' / 'adam' when ``optimizer`` is EXACTLY torch.optim.SGD / Adam / AdamW with options the
    fused kernels implement, else None (subclasses may override step(): they keep the generic path)."""
    if os.environ.get("APEX_B200_ADOPT_TORCH_OPTIM", "1") == "0":
        return None
    kind = {torch.optim.SGD: "sgd", torch.optim.Adam: "adam", torch.optim.AdamW: "adam"}.get(type(optimizer))
    if kind is None:
        return None
    for g in optimizer.param_groups:
        if g.get("maximize") or g.get("differentiable") or g.get("capturable") or g.get("amsgrad"):
            return None
        if any(isinstance(g.get(k), torch.Tensor) for k in ("lr", "weight_decay", "momentum", "eps")):
            return None
        if kind == "adam" and any(isinstance(b, torch.Tensor) for b in g["betas"]):
            return None
    return kind


def _adopt_torch_optimizer(optimizer, kind):
    """Serve a stock torch optimizer's step() with b200_fused_sgd / b200_fused_adam.

    The engine is a FusedSGD / FusedAdam object that SHARES the torch optimizer's ``param_groups``
    list and ``state`` dict: hyper-parameters (and schedulers that edit them) are read from the
    groups at every step, and the state keys are torch's own (``momentum_buffer``, ``exp_avg``,
    ``exp_avg_sq``), so ``state_dict()`` can be loaded into an unpatched torch optimizer and back.
    Adam's per-parameter ``step`` tensors are filled in when a state_dict is taken."""
    from ..optimizers import FusedAdam, FusedSGD
    dummy = [torch.zeros(1)]
    if kind == "sgd":
        eng = FusedSGD(dummy, lr=0.0, materialize_master_grads=False, set_grad_none=False)
    else:
        g0 = optimizer.param_groups[0]
        decoupled = isinstance(optimizer, torch.optim.AdamW) or bool(g0.get("decoupled_weight_decay", False))
        eng = FusedAdam(dummy, adam_w_mode=decoupled, set_grad_none=False)
    eng.param_groups = optimizer.param_groups
    eng.state = optimizer.state
    eng._amp_stash = optimizer._amp_stash
    eng._adopted_kind = kind
    optimizer._b200_engine = eng

    def _before_state_dict(opt):
        eng._sync_steps()
        if kind == "adam":
            for g in opt.param_groups:
                for p in g["params"]:
                    st = opt.state.get(p)
                    if st:
                        st["step"] = torch.tensor(float(g.get("step", 0)), dtype=torch.float32)

    def _after_load(opt):
        eng.param_groups, eng.state = opt.param_groups, opt.state
        eng._counter, eng._step_offsets, eng._tables = None, {}, {}
        eng.invalidate()
        for g in opt.param_groups:
            steps = [opt.state[p]["step"] for p in g["params"] if p in opt.state and "step" in opt.state[p]]
            if steps:
                g["step"] = int(max(float(s) for s in steps))
            elif "step" not in g and any("momentum_buffer" in opt.state.get(p, {}) for p in g["params"]):
                g["step"] = 1

    optimizer.register_state_dict_pre_hook(_before_state_dict)
    optimizer.register_load_state_dict_post_hook(_after_load)
    if optimizer.state:                       # loaded from a checkpoint before amp.initialize: pick up its step counts
        _after_load(optimizer)
    return eng


def _process_optimizer(optimizer, properties):
    if getattr(optimizer, "_amp_stash", None) is not None:
        raise RuntimeError("A given optimizer should only be passed through amp.initialize once.")
    stash = AmpStash()
    optimizer._amp_stash = stash
    stash.params_have_scaled_gradients = False

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

    on_cuda = all(p.is_cuda for g in optimizer.param_groups for p in g["params"])
    kind = _adoptable(optimizer) if (on_cuda and not isinstance(optimizer, FusedOptimizerBase)) else None
    if kind is not None:
        _adopt_torch_optimizer(optimizer, kind)
    engine = _engine(optimizer)
    fused = engine is not None
    if isinstance(optimizer, FusedOptimizerBase):
        optimizer.invalidate()

    optimizer._prepare_amp_backward = types.MethodType(_prepare_amp_backward, optimizer)
    optimizer._post_amp_backward = types.MethodType(_post_amp_backward, optimizer)
    optimizer._amp_materialize_master_grads = types.MethodType(_materialize_master_grads, optimizer)

    old_step = optimizer.step
    old_zero_grad = optimizer.zero_grad

    def new_step(self, closure=None, *args, **kwargs):
        if closure is not None:
            raise RuntimeError("Currently, Amp does not support closure use with optimizers.")
        if engine is not None and engine is not self:
            retval = engine.step()                       # adopted torch optimizer: one fused launch per dtype set
        else:
            retval = old_step(*args, **kwargs)
            if not fused and stash.scaler is not None:
                stash.scaler.note_step()
        if not fused and properties.master_weights:
            _master_params_to_model_params(self)
        if properties.master_weights:
            # apex clears the master grads once they are consumed
            for p in stash.all_fp32_from_fp16_params:
                p.grad = None
        stash.grads_scaled = False                       # consumed: a later backward starts a new cycle
        if stash.accumulating:
            stash.accumulating = False
            stash.version += 1
        return retval

    def new_zero_grad(self, *args, **kwargs):
        set_none = bool(getattr(self, "set_grad_none", False))
        params = stash.all_fp16_params + stash.all_fp32_params
        if set_none:
            for p in params:
                p.grad = None
            stash.version += 1
        else:
            _zero_grads(params)
        for p in stash.all_fp32_from_fp16_params:
            p.grad = None
        stash.grads_scaled = False
        if stash.accumulating:
            stash.accumulating = False
            stash.version += 1

    optimizer.step = types.MethodType(new_step, optimizer)
    optimizer.zero_grad = types.MethodType(new_zero_grad, optimizer)
    optimizer._amp_original_step = old_step
    optimizer._amp_original_zero_grad = old_zero_grad
    return optimizer


def _zero_grads(params):
    """Patched zero_grad without set-to-none (apex a9).  Gradients that are views of a DDP bucket are
    cleared with ONE memset per bucket (the wrapper knows which buckets currently back every gradient
    of theirs); whatever is left is zeroed with one foreach launch per dtype."""
    left = params
    for ddp in list(_amp_state.ddp_wrappers):
        left = ddp.zero_bucket_grads(left)
    grads = []
    for p in left:
        g = p.grad
        if g is not None:
            if g.grad_fn is not None:
                g.detach_()
            else:
                g.requires_grad_(False)          # views of the DDP buckets cannot be detached in place
            grads.append(g)
    if grads:
        torch._foreach_zero_(grads)
