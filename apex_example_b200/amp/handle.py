"""amp.scale_loss — the context manager around backward().

Mirrors apex/amp/handle.py::scale_loss, entered by the reference at
/root/reference/test_apex_distributed_spawn.py:160.  The loss is multiplied by a DEVICE scalar
(the scaler's state record), so the scale can change on the device without the host knowing.
"""
from __future__ import annotations

import contextlib
import types

import torch

from ._amp_state import _amp_state, maybe_print
from ._process_optimizer import _is_fused


@contextlib.contextmanager
def scale_loss(loss, optimizers, loss_id=0, model=None, delay_unscale=False, delay_overflow_check=False):
    if not _amp_state.initialized:
        raise RuntimeError("Invoked 'with amp.scale_loss`, but internal Amp state has not been initialized.  "
                           "model, optimizer = amp.initialize(model, optimizer, opt_level=...) must be called "
                           "before `with amp.scale_loss`.")
    if not _amp_state.enabled or not _amp_state.opt_properties.enabled:
        yield loss
        return

    if isinstance(optimizers, torch.optim.Optimizer) or type(optimizers).__name__ == "LARC":
        optimizers = [optimizers]

    loss_scaler = _amp_state.loss_scalers[loss_id]

    # plain O0 / O3: static scale 1.0 and no masters -> nothing to unscale or check
    if (not _amp_state.opt_properties.master_weights and not loss_scaler.dynamic
            and loss_scaler._loss_scale == 1.0):
        yield loss.float()
        return

    if not delay_unscale:
        for optimizer in optimizers:
            if not optimizer._amp_stash.params_have_scaled_gradients:
                optimizer._prepare_amp_backward()

    loss_scaler._ensure_state(loss.device if loss.is_cuda else None)
    loss_scaler.poll_overflow_message()
    yield loss.float() * loss_scaler.loss_scale_tensor()

    if delay_unscale:
        for optimizer in optimizers:
            optimizer._amp_stash.params_have_scaled_gradients = True
    else:
        loss_scaler.clear_overflow_state()
        for optimizer in optimizers:
            optimizer._post_amp_backward(loss_scaler)
            optimizer._amp_stash.params_have_scaled_gradients = False
        # generic optimizers need the decision on the host (one sync, as in apex); fused ones
        # read found_inf on the device
        need_host = any(not _is_fused(o) for o in optimizers)
        should_skip = False if delay_overflow_check else loss_scaler.update_scale(need_host_decision=need_host)
        if should_skip:
            for optimizer in optimizers:
                if _is_fused(optimizer):
                    continue
                if not hasattr(optimizer, "_amp_stash") or getattr(optimizer._amp_stash, "already_patched", False):
                    continue
                _patch_skip_step(optimizer, loss_scaler, loss_id)
    _amp_state.ddp_checked_backward = False


def _patch_skip_step(optimizer, loss_scaler, loss_id):
    old_step = optimizer.step
    stash = optimizer._amp_stash

    def skip_step(self, closure=None):
        if closure is not None:
            raise RuntimeError("Currently, Amp does not support closure use with optimizers.")
        maybe_print("Gradient overflow.  Skipping step, loss scaler {} reducing loss scale to {}".format(
            loss_id, loss_scaler.loss_scale()))
        for p in stash.all_fp32_from_fp16_params:
            p.grad = None
        loss_scaler.note_step()                  # the skip consumed the decision: the next update starts afresh
        stash.grads_scaled = False
        if stash.accumulating:
            stash.accumulating = False
            stash.version += 1
        self.step = old_step
        stash.already_patched = False

    optimizer.step = types.MethodType(skip_step, optimizer)
    stash.already_patched = True


@contextlib.contextmanager
def disable_casts():
    """O1 helper: run a region without autocast."""
    with torch.autocast(device_type="cuda", enabled=False):
        yield
