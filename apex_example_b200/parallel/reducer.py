"""apex.parallel.Reducer — the manual-trigger sibling of DistributedDataParallel.

apex's Reducer (apex/parallel/distributed.py) all-reduces a module's gradients only when the user
calls ``reduce()`` — e.g. every few iterations — and averages over the process group.  apex
implements it with ``flat_dist_call``: flatten copy, ``all_reduce``, ``mul_(1/world)``, unflatten
copy.  Here the gradients are reduced through persistent buckets and the same epilogue kernel as
the DDP wrapper (``b200_bucket_epilogue``: x 1/world in the gradient dtype + amp's overflow
check), and when the gradients already sit in one flat buffer (views of a flat tensor, of DDP-style
buckets, or of a flat gradient arena) they are reduced IN PLACE: no flatten, no unflatten, host work
O(buckets) per call however many tensors there are.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch.nn.modules import Module

from .. import ops
from ..tables import TensorTable, flat_view_like
from .distributed import flat_dist_call


def _extract_tensors(maybe_tensor, out):
    if torch.is_tensor(maybe_tensor):
        out.append(maybe_tensor)
        return
    try:
        for item in maybe_tensor:
            _extract_tensors(item, out)
    except TypeError:
        return


class Reducer:
    """``Reducer(module_or_grads_list)``; ``reduce()`` averages the gradients across the group.

    A module's parameters are broadcast from rank 0 at construction (as apex does); with a list of
    gradient tensors the caller is responsible for starting from identical parameters.
    ``message_size`` (elements) bounds one all-reduce, as in DistributedDataParallel."""

    def __init__(self, module_or_grads_list, message_size=10000000, gradient_average=True, process_group=None):
        if not dist.is_available() or not dist.is_initialized():
            raise RuntimeError("apex.parallel.Reducer requires torch.distributed to be initialized")
        self.group = process_group
        self.world_size = dist.get_world_size(process_group)
        self.message_size = int(message_size)
        self.gradient_average = gradient_average
        if isinstance(module_or_grads_list, Module):
            self.module = module_or_grads_list
            flat_dist_call([p.data for p in self.module.parameters()], dist.broadcast, (0,))
            self.grads = None
        else:
            self.module = None
            self.grads = []
            _extract_tensors(module_or_grads_list, self.grads)
        self._plan = None

    # ---- planning ----------------------------------------------------------------------------------
    def _current_grads(self):
        if self.module is not None:
            return [p.grad for p in self.module.parameters() if p.grad is not None]
        return self.grads

    @staticmethod
    def _flat_span(tensors):
        """If every tensor is a contiguous view into ONE storage, return a flat tensor over the span
        they cover (gaps are alignment padding that the caller keeps zero or does not care about)."""
        first = tensors[0]
        store = first.untyped_storage()
        base = store.data_ptr()
        lo, hi = None, None
        for t in tensors:
            if not t.is_contiguous() or t.untyped_storage().data_ptr() != base or t.dtype != first.dtype:
                return None
            a = t.data_ptr()
            b = a + t.numel() * t.element_size()
            lo = a if lo is None or a < lo else lo
            hi = b if hi is None or b > hi else hi
        covered = sum(t.numel() for t in tensors) * first.element_size()
        if covered < 0.9 * (hi - lo):            # mostly foreign bytes in between: not a gradient arena
            return None
        es = first.element_size()
        return torch.empty(0, dtype=first.dtype, device=first.device).set_(store, (lo - base) // es, ((hi - lo) // es,))

    def _build_plan(self, grads):
        by_dtype = {}
        for g in grads:
            by_dtype.setdefault(g.dtype, []).append(g)
        plan = {"sig": tuple(g.data_ptr() for g in grads), "sets": []}
        for dt, gs in by_dtype.items():
            flat = self._flat_span(gs) if gs[0].is_cuda else None
            entry = {"dtype": dt, "grads": gs, "in_place": flat is not None}
            if flat is None:
                total = sum(g.numel() for g in gs)
                flat = torch.zeros(total, dtype=dt, device=gs[0].device)
                views, off = [], 0
                for g in gs:
                    views.append(flat_view_like(flat, off, g))
                    off += g.numel()
                entry["views"] = views
                if gs[0].is_cuda:
                    entry["gather"] = TensorTable([gs, views])
                    entry["scatter"] = TensorTable([views, gs])
            n = flat.numel()
            entry["chunks"] = [flat[o:min(n, o + self.message_size)] for o in range(0, n, self.message_size)]
            plan["sets"].append(entry)
        return plan

    def _overflow_flag(self, device):
        try:
            from ..amp._amp_state import _amp_state
        except Exception:
            return None, None
        if not (_amp_state.initialized and _amp_state.enabled and len(_amp_state.loss_scalers) == 1):
            return None, None
        scaler = _amp_state.loss_scalers[0]
        scaler._ensure_state(device)
        return scaler.overflow_ref(), _amp_state

    # ---- the call -----------------------------------------------------------------------------------
    def reduce(self):
        grads = self._current_grads()
        if not grads:
            return
        plan = self._plan
        n = len(grads)
        if plan is None or len(plan["sig"]) != n or (
                n <= 4096 and plan["sig"] != tuple(g.data_ptr() for g in grads)):
            plan = self._plan = self._build_plan(grads)
        factor = 1.0 / self.world_size if self.gradient_average else 1.0
        for s in plan["sets"]:
            cuda = s["grads"][0].is_cuda
            flag, amp_state = self._overflow_flag(s["grads"][0].device) if cuda else (None, None)
            if not s["in_place"]:
                if cuda:
                    ops.multi_tensor_scale(s["gather"], s["dtype"], s["dtype"], scale=1.0)
                else:
                    for g, v in zip(s["grads"], s["views"]):
                        v.copy_(g)
            for c in s["chunks"]:
                if self.world_size > 1:
                    dist.all_reduce(c, group=self.group)
                if cuda:
                    ops.bucket_epilogue(c, factor, flag)
                elif factor != 1.0:
                    c.mul_(factor)
            if not s["in_place"]:
                if cuda:
                    ops.multi_tensor_scale(s["scatter"], s["dtype"], s["dtype"], scale=1.0)
                else:
                    for g, v in zip(s["grads"], s["views"]):
                        g.copy_(v)
            if flag is not None:
                # the epilogue inspected every reduced gradient: amp's scale_loss exit need not do it again
                amp_state.ddp_checked_backward = True
                amp_state.ddp_checked_ids = getattr(amp_state, "ddp_checked_ids", frozenset()) | self._ids(grads)

    def _ids(self, grads):
        if self.module is not None:
            return frozenset(id(p) for p in self.module.parameters() if p.grad is not None)
        ids = getattr(self, "_param_ids", None)
        return ids if ids is not None else frozenset()

    def covers(self, params):
        """Tell the Reducer which parameters own the gradient tensors it was given, so amp can rely on
        its overflow check for them (only needed with a list of gradients; a module's are known)."""
        self._param_ids = frozenset(id(p) for p in params)
        return self
