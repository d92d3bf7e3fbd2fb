"""LossScaler with its state resident on the device.

Restates apex/amp/scaler.py (reached from /root/reference/test_apex_distributed_spawn.py:160,
``amp.scale_loss`` exit) — SURVEY.md App. B.2 — but the overflow flag, the scale and the growth
counter live in one 32-byte device record (b200_scaler_state, include/apex_b200.h) that the check
/ epilogue kernels raise and the optimizer kernels read.  No ``.item()`` per step.
"""
from __future__ import annotations

import torch

from .. import ops
from ..tables import TensorTable, signature_of
from ._amp_state import maybe_print


class LossScaler:
    warned_no_fused_kernel = False
    warned_unscaling_non_fp32_grad = False
    has_fused_kernel = True

    def __init__(self, loss_scale, init_scale=2.0 ** 16, scale_factor=2.0, scale_window=2000,
                 min_loss_scale=None, max_loss_scale=2.0 ** 24, device=None):
        if loss_scale == "dynamic":
            self.dynamic = True
            self._loss_scale = min(max_loss_scale, init_scale)
        else:
            self.dynamic = False
            self._loss_scale = float(loss_scale)
        self._max_loss_scale = max_loss_scale
        self._min_loss_scale = min_loss_scale
        self._scale_seq_len = scale_window
        self._scale_factor = scale_factor
        self._unskipped = 0
        self._device = device
        self._state = None          # int32[8] device tensor == b200_scaler_state
        self._mirror = None         # pinned host copy written by the update kernel
        self._seen_seq = 0
        self._seen_skips = 0
        self._loss_id = 0
        self._tables = {}
        self._updates_since_step = 0    # update_scale() calls no optimizer step has consumed yet

    # ---- device record -----------------------------------------------------------------------
    def _ensure_state(self, device=None):
        if self._state is None:
            device = device or self._device or torch.device("cuda", torch.cuda.current_device())
            self._state = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32, device=device)
            self._mirror = torch.zeros(ops.SCALER_WORDS, dtype=torch.int32).pin_memory()
            ops.scaler_init(self._state, float(self._loss_scale))
            if self._unskipped:
                self._state[ops.IDX_UNSKIPPED] = self._unskipped
        return self._state

    @property
    def state(self) -> torch.Tensor:
        return self._ensure_state()

    def loss_scale_tensor(self) -> torch.Tensor:
        """0-d fp32 view of the current scale (what scale_loss multiplies the loss with)."""
        return self.state.view(torch.float32)[ops.IDX_LOSS_SCALE]

    def inv_scale_ref(self) -> torch.Tensor:
        return self.state.view(torch.float32)[ops.IDX_INV_SCALE:]

    def found_inf_ref(self) -> torch.Tensor:
        return self.state[ops.IDX_FOUND_INF:]

    def overflow_ref(self) -> torch.Tensor:
        return self.state[ops.IDX_OVERFLOW_ACC:]

    # ---- apex API -----------------------------------------------------------------------------
    def loss_scale(self) -> float:
        """Host value of the scale.  Reads the pinned mirror; only synchronises when the device is
        ahead of the last mirrored update."""
        self._refresh(sync=True)
        return self._loss_scale

    def clear_overflow_state(self):
        # the update kernel clears overflow_acc itself; nothing to launch
        pass

    def unscale(self, model_grads, master_grads, unused_scale=None, models_are_masters=False, scale_override=None):
        """apex LossScaler.unscale: master = model * (1/scale), raise the flag on inf/NaN input.
        One launch for the whole list (apex: ceil(chunks/320) launches)."""
        if not model_grads:
            return
        cols = [model_grads, master_grads]
        key = ("unscale", model_grads[0].dtype, master_grads[0].dtype, len(model_grads))
        sig = signature_of(cols)
        tbl = self._tables.get(key)
        if tbl is None or tbl.signature != sig:
            tbl = TensorTable(cols, signature=sig)
            self._tables[key] = tbl
        self._ensure_state()
        if scale_override is not None:
            ops.multi_tensor_scale(tbl, model_grads[0].dtype, master_grads[0].dtype, scale=float(scale_override),
                                   dev_flag=self.overflow_ref())
        else:
            # the backward that produced these grads was scaled by the CURRENT loss_scale (the update
            # has not run yet), whose reciprocal the device record keeps in inv_loss_scale
            ops.multi_tensor_scale(tbl, model_grads[0].dtype, master_grads[0].dtype,
                                   dev_scale=self.state.view(torch.float32)[ops.IDX_INV_LOSS_SCALE:],
                                   dev_flag=self.overflow_ref())

    def check(self, grads):
        """Overflow check without the arithmetic (the unscale itself is fused into the optimizer step)."""
        if not grads:
            return
        key = ("check", grads[0].dtype, len(grads))
        sig = signature_of([grads])
        tbl = self._tables.get(key)
        if tbl is None or tbl.signature != sig:
            tbl = TensorTable([grads], signature=sig)
            self._tables[key] = tbl
        ops.multi_tensor_check(tbl, 0, grads[0].dtype, self.overflow_ref())

    def update_scale(self, need_host_decision: bool = True):
        """apex LossScaler.update_scale.  Runs the state machine on the device.  Returns
        ``should_skip`` when the caller needs it on the host (generic optimizers — one sync, as in
        apex), else None (fused optimizers skip on the device)."""
        self._ensure_state()
        # A second update before any step() (gradient accumulation without delay_unscale) must not
        # forget an overflow the first one saw: apex's patched skip_step stays until step() runs.
        mode = (ops.SCALER_DYNAMIC if self.dynamic else 0) | (ops.SCALER_KEEP_SKIP if self._updates_since_step else 0)
        self._updates_since_step += 1
        ops.scaler_update(self._state, mode, self._scale_seq_len, self._scale_factor,
                          float(self._min_loss_scale or 0.0), float(self._max_loss_scale), self._mirror)
        if need_host_decision and self.dynamic:
            self._refresh(sync=True, quiet=True)        # the patched skip_step prints the message (as apex)
            return bool(self._mirror[ops.IDX_FOUND_INF])
        if not self.dynamic:
            return False
        return None

    def note_step(self):
        """An optimizer step has been issued: the next update_scale starts a fresh decision."""
        self._updates_since_step = 0

    def _refresh(self, sync: bool, quiet: bool = False):
        """Fold what the device has mirrored so far into the host copy; prints apex's overflow line
        once per skip the host had not seen yet."""
        if self._state is None:
            return
        if sync:
            torch.cuda.current_stream(self._state.device).synchronize()
        m = self._mirror
        seq = int(m[ops.IDX_SEQ])
        if seq == self._seen_seq:
            return
        self._seen_seq = seq
        self._loss_scale = float(m.view(torch.float32)[ops.IDX_LOSS_SCALE])
        self._unskipped = int(m[ops.IDX_UNSKIPPED])
        skipped = int(m[ops.IDX_SKIPPED_TOTAL])
        for _ in range(0 if quiet else skipped - self._seen_skips):
            maybe_print("Gradient overflow.  Skipping step, loss scaler {} reducing loss scale to {}".format(
                self._loss_id, self._loss_scale))
        self._seen_skips = skipped

    def poll_overflow_message(self):
        """No-sync variant used at scale_loss entry: report whatever has been mirrored by now."""
        self._refresh(sync=False)

    # ---- checkpoint surface (amp.state_dict) -----------------------------------------------------
    def state_dict(self):
        self._refresh(sync=True)
        return {"loss_scale": self._loss_scale, "unskipped": self._unskipped}

    def load_state_dict(self, sd):
        self._loss_scale = float(sd["loss_scale"])
        self._unskipped = int(sd["unskipped"])
        if self._state is not None:
            ops.scaler_init(self._state, self._loss_scale)
            self._state[ops.IDX_UNSKIPPED] = self._unskipped
