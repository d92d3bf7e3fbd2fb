"""apex.multi_tensor_apply — API kept, scheduler replaced.

``multi_tensor_applier(op, noop_flag_buffer, tensor_lists, *args)`` still forwards to
``op(chunk_size, noop_flag_buffer, tensor_lists, *args)``; the ops in :mod:`amp_C` build (and
cache) a device-resident table for ``tensor_lists`` and cover the list in ONE launch, so
``chunk_size`` is accepted and ignored.
"""
from __future__ import annotations

from collections import OrderedDict

import torch

from .. import ops
from ..tables import TensorTable, signature_of


class MultiTensorApply:
    available = True
    warned = False

    def __init__(self, chunk_size):
        self.chunk_size = chunk_size

    def check_avail(self):
        from .. import _lib
        _lib.load()

    def __call__(self, op, noop_flag_buffer, tensor_lists, *args):
        return op(self.chunk_size, noop_flag_buffer, tensor_lists, *args)


multi_tensor_applier = MultiTensorApply(2048 * 32)

# ---- table cache for the amp_C-compatible entry points -------------------------------------------
_CACHE: "OrderedDict[tuple, TensorTable]" = OrderedDict()
_CACHE_MAX = 64


def table_for(tensor_lists, groups=None) -> TensorTable:
    sig = signature_of(tensor_lists)
    tbl = _CACHE.get(sig)
    if tbl is None:
        tbl = TensorTable(tensor_lists, groups=groups, signature=sig)
        _CACHE[sig] = tbl
        if len(_CACHE) > _CACHE_MAX:
            _CACHE.popitem(last=False)
    else:
        _CACHE.move_to_end(sig)
    return tbl


def _check_flag(noop_flag):
    if noop_flag.dtype != torch.int32 or noop_flag.numel() < 1:
        raise RuntimeError("noop_flag must be an int32 tensor with at least one element")


def multi_tensor_scale(chunk_size, noop_flag, tensor_lists, scale):
    """amp_C.multi_tensor_scale(chunk_size, noop_flag, [in, out], scale)."""
    if len(tensor_lists) != 2:
        raise RuntimeError("multi_tensor_scale expects tensor_lists = [inputs, outputs]")
    if not tensor_lists[0]:
        return
    _check_flag(noop_flag)
    ops.multi_tensor_scale(table_for(tensor_lists), tensor_lists[0][0].dtype, tensor_lists[1][0].dtype,
                           scale=float(scale), dev_flag=noop_flag)


def multi_tensor_axpby(chunk_size, noop_flag, tensor_lists, a, b, arg_to_check):
    """amp_C.multi_tensor_axpby(chunk_size, noop_flag, [x, y, out], a, b, arg_to_check)."""
    if len(tensor_lists) != 3:
        raise RuntimeError("multi_tensor_axpby expects tensor_lists = [x, y, out]")
    if not tensor_lists[0]:
        return
    _check_flag(noop_flag)
    ops.multi_tensor_axpby(table_for(tensor_lists), tensor_lists[0][0].dtype, tensor_lists[1][0].dtype,
                           tensor_lists[2][0].dtype, float(a), float(b), int(arg_to_check), noop_flag)


def multi_tensor_sgd(chunk_size, noop_flag, tensor_lists, wd, momentum, dampening, lr, nesterov, first_run,
                     wd_after_momentum, scale):
    """amp_C.multi_tensor_sgd: lists [g, w, mom] or [g, w, mom, model_copy]; early-out on *noop_flag."""
    if len(tensor_lists) not in (3, 4):
        raise RuntimeError("multi_tensor_sgd expects 3 or 4 tensor lists")
    if not tensor_lists[0]:
        return
    _check_flag(noop_flag)
    cdt = tensor_lists[3][0].dtype if len(tensor_lists) == 4 else None
    grp = ops.SgdGroup(float(lr), float(momentum), float(dampening), float(wd), 1 if first_run else 2,
                       int(bool(nesterov)), int(bool(wd_after_momentum)))
    ops.fused_sgd(table_for(tensor_lists), tensor_lists[0][0].dtype, tensor_lists[1][0].dtype, cdt, [grp],
                  inv_scale=float(scale), dev_found_inf=noop_flag)


def multi_tensor_adam(chunk_size, noop_flag, tensor_lists, lr, beta1, beta2, eps, step, mode, bias_correction,
                      weight_decay):
    """amp_C.multi_tensor_adam: lists [g, p, m, v] (+ optional model copy); mode 0 = L2, 1 = AdamW."""
    if len(tensor_lists) not in (4, 5):
        raise RuntimeError("multi_tensor_adam expects 4 (or 5) tensor lists")
    if not tensor_lists[0]:
        return
    _check_flag(noop_flag)
    cdt = tensor_lists[4][0].dtype if len(tensor_lists) == 5 else None
    grp = ops.AdamGroup(float(lr), float(beta1), float(beta2), float(eps), float(weight_decay), int(step),
                        int(bool(bias_correction)), int(mode))
    ops.fused_adam(table_for(tensor_lists), tensor_lists[0][0].dtype, tensor_lists[1][0].dtype, cdt, [grp],
                   dev_found_inf=noop_flag)


_NORM_SCRATCH = {}


def multi_tensor_l2norm(chunk_size, noop_flag, tensor_lists, per_tensor=False):
    """amp_C.multi_tensor_l2norm(chunk_size, noop_flag, [tensors], per_tensor) -> (norm, per_tensor_norms).
    One read-only pass; the flag is raised on inf/NaN like apex's.  ``per_tensor=True`` (LAMB's trust ratios) returns
    the norm of every tensor too, summed in a fixed order (b200_l2norm_per_tensor)."""
    if per_tensor:
        if len(tensor_lists) != 1:
            raise RuntimeError("multi_tensor_l2norm expects tensor_lists = [tensors]")
        tensors = tensor_lists[0]
        device = tensors[0].device if tensors else noop_flag.device
        total = torch.zeros(1, device=device)
        norms = torch.zeros(len(tensors), device=device)
        if tensors:
            _check_flag(noop_flag)
            table = table_for(tensor_lists)
            ss = torch.empty(len(tensors), device=device)
            tile_ss = torch.empty(max(table.n_tiles, 1), device=device)
            ops.l2norm_per_tensor(table, 0, tensors[0].dtype, table.first_tiles(), tile_ss, ss, norms=norms, total=total,
                                  dev_flag=noop_flag)
        return total.sqrt_(), norms
    if len(tensor_lists) != 1:
        raise RuntimeError("multi_tensor_l2norm expects tensor_lists = [tensors]")
    tensors = tensor_lists[0]
    device = tensors[0].device if tensors else noop_flag.device
    sumsq = torch.zeros(1, device=device)
    if tensors:
        _check_flag(noop_flag)
        scratch = _NORM_SCRATCH.get(device)
        if scratch is None:
            scratch = _NORM_SCRATCH[device] = ops.new_norm_scratch(device)
        ops.multi_tensor_l2norm(table_for(tensor_lists), 0, tensors[0].dtype, sumsq, scratch, noop_flag)
    return sumsq.sqrt_(), torch.empty(0, device=device)
