"""Thin Python launchers over the C ABI (include/apex_b200.h).  Every function enqueues one
kernel of libapex_b200.so on the current CUDA stream and returns; none synchronises."""
from __future__ import annotations

import ctypes
from typing import Sequence

import torch

from . import _lib
from .tables import TensorTable, dtype_code

AdamGroup, SgdGroup, StepSync = _lib.AdamGroup, _lib.SgdGroup, _lib.StepSync
STEP_WINDOW = _lib.STEP_WINDOW


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _addr(t) -> int:
    """Device address of a tensor (or 0)."""
    if t is None:
        return 0
    if isinstance(t, int):
        return t
    return t.data_ptr()


def _require_cuda(t: TensorTable):
    if t.dev.device.type != "cuda":
        raise RuntimeError("libapex_b200 kernels need CUDA tensors; there is no CPU fallback")


def multi_tensor_scale(table: TensorTable, in_dtype, out_dtype, scale: float = 1.0, dev_scale=None, dev_flag=None):
    _require_cuda(table)
    _lib.check(_lib.load().b200_multi_tensor_scale(table.ptr, table.n_tiles, dtype_code(in_dtype), dtype_code(out_dtype),
                                                   scale, _addr(dev_scale), _addr(dev_flag), _stream()),
               "multi_tensor_scale")


def multi_tensor_axpby(table: TensorTable, x_dtype, y_dtype, out_dtype, a: float, b: float, arg_to_check: int = -1,
                       dev_flag=None):
    _require_cuda(table)
    _lib.check(_lib.load().b200_multi_tensor_axpby(table.ptr, table.n_tiles, dtype_code(x_dtype), dtype_code(y_dtype),
                                                   dtype_code(out_dtype), a, b, arg_to_check, _addr(dev_flag), _stream()),
               "multi_tensor_axpby")


def multi_tensor_check(table: TensorTable, col: int, dtype, dev_flag):
    _require_cuda(table)
    _lib.check(_lib.load().b200_multi_tensor_check(table.ptr, table.n_tiles, col, dtype_code(dtype), _addr(dev_flag),
                                                   _stream()), "multi_tensor_check")


def fused_adam(table: TensorTable, grad_dtype, state_dtype, copy_dtype, groups: Sequence[AdamGroup],
               inv_scale: float = 1.0, dev_inv_scale=None, dev_found_inf=None, sync: StepSync | None = None,
               pull=None):
    _require_cuda(table)
    arr = (AdamGroup * len(groups))(*groups)
    _lib.check(_lib.load().b200_fused_adam_peer(table.ptr, table.n_tiles, dtype_code(grad_dtype), dtype_code(state_dtype),
                                                dtype_code(copy_dtype), arr, len(groups), inv_scale, _addr(dev_inv_scale),
                                                _addr(dev_found_inf), ctypes.byref(sync) if sync is not None else None,
                                                ctypes.byref(pull) if pull is not None else None, _stream()),
               "fused_adam")


def fused_sgd(table: TensorTable, grad_dtype, state_dtype, copy_dtype, groups: Sequence[SgdGroup],
              inv_scale: float = 1.0, dev_inv_scale=None, dev_found_inf=None, sync: StepSync | None = None,
              pull=None):
    _require_cuda(table)
    arr = (SgdGroup * len(groups))(*groups)
    _lib.check(_lib.load().b200_fused_sgd_peer(table.ptr, table.n_tiles, dtype_code(grad_dtype), dtype_code(state_dtype),
                                               dtype_code(copy_dtype), arr, len(groups), inv_scale, _addr(dev_inv_scale),
                                               _addr(dev_found_inf), ctypes.byref(sync) if sync is not None else None,
                                               ctypes.byref(pull) if pull is not None else None, _stream()),
               "fused_sgd")


PeerPull = _lib.PeerPull


def peer_reduce_shard(peer_ptrs, reduced: torch.Tensor, begin: int, end: int, factor: float, dev_flag=None, sync=None):
    """Reduce-scatter + mean + overflow check for this rank's shard: peer_ptrs[r] is rank r's flat
    gradient buffer as addressable from this process (device address, int).  ``sync`` (a PeerSync)
    adds the in-kernel hand-shake; the flag is then delivered by :func:`peer_wait_flags`."""
    arr = (ctypes.c_void_p * len(peer_ptrs))(*peer_ptrs)
    _lib.check(_lib.load().b200_peer_reduce_shard(arr, len(peer_ptrs), reduced.data_ptr(), begin, end,
                                                  dtype_code(reduced.dtype), factor, _addr(dev_flag),
                                                  ctypes.byref(sync) if sync is not None else None, _stream()),
               "peer_reduce_shard")


def peer_wait_flags(sync, dev_flag):
    _lib.check(_lib.load().b200_peer_wait_flags(ctypes.byref(sync), _addr(dev_flag), _stream()), "peer_wait_flags")


def make_peer_sync(pad_ptrs, rank: int, epoch: int):
    """b200_peer_sync for this step; keeps the pointer array alive on the returned struct."""
    arr = (ctypes.c_void_p * len(pad_ptrs))(*pad_ptrs)
    ps = _lib.PeerSync(ctypes.cast(arr, ctypes.POINTER(ctypes.c_void_p)), len(pad_ptrs), rank, epoch, 0)
    ps._keep = arr
    return ps


def make_peer_pull(local_base_ptr: int, bounds, reduced_ptrs, elem_size: int, first_tile: int = 0) -> "PeerPull":
    world = len(reduced_ptrs)
    pp = PeerPull()
    pp.local_base = local_base_ptr
    for r in range(world + 1):
        pp.bounds[r] = int(bounds[r])
    for r in range(world):
        pp.delta[r] = int(reduced_ptrs[r]) - int(local_base_ptr)
    pp.world = world
    pp.elem_size = elem_size
    pp.first_tile = int(first_tile)
    return pp


class StepCounter:
    """Host half of b200_step_sync for one launch set: device {calls, skips}, a pinned mirror the
    kernel writes, and the bookkeeping that turns them into an exact step number."""

    def __init__(self, device):
        self.dev = torch.zeros(2, dtype=torch.int32, device=device)
        self.mirror = torch.zeros(2, dtype=torch.int32).pin_memory()
        self.calls = 0                      # launches issued by the host
        self.base_step = 0                  # successful steps before this counter existed (load_state_dict)

    def next(self):
        """Account for one more launch.  Returns (step, StepSync) where step is the 1-based step
        number assuming no skip the host has not seen yet."""
        self.calls += 1
        seen_calls = int(self.mirror[0])
        known_skips = int(self.mirror[1])
        if self.calls - 1 - seen_calls > STEP_WINDOW - 1:
            # the device is more than a window behind what the host has enqueued: wait for it
            torch.cuda.current_stream(self.dev.device).synchronize()
            known_skips = int(self.mirror[1])
        step = self.base_step + self.calls - known_skips
        return max(step, 1), StepSync(self.dev.data_ptr(), self.mirror.data_ptr(), known_skips, 0)

    def exact_steps(self) -> int:
        """Number of steps that really ran (synchronises)."""
        torch.cuda.current_stream(self.dev.device).synchronize()
        return self.base_step + int(self.mirror[0]) - int(self.mirror[1]) if self.calls else self.base_step


def bucket_epilogue(bucket: torch.Tensor, factor: float, dev_flag=None):
    if bucket.device.type != "cuda":
        raise RuntimeError("libapex_b200 kernels need CUDA tensors; there is no CPU fallback")
    if not bucket.is_contiguous():
        raise RuntimeError("bucket must be contiguous")
    _lib.check(_lib.load().b200_bucket_epilogue(bucket.data_ptr(), bucket.numel(), dtype_code(bucket.dtype), factor,
                                                _addr(dev_flag), _stream()), "bucket_epilogue")


# ---- loss-scaler state ------------------------------------------------------------------------
SCALER_WORDS = 12           # sizeof(b200_scaler_state) / 4
(IDX_LOSS_SCALE, IDX_INV_SCALE, IDX_FOUND_INF, IDX_UNSKIPPED, IDX_OVERFLOW_ACC, IDX_SEQ, IDX_HAD_OVERFLOW,
 IDX_INV_LOSS_SCALE, IDX_SKIPPED_TOTAL) = range(9)


def scaler_init(state: torch.Tensor, init_scale: float):
    _lib.check(_lib.load().b200_scaler_init(state.data_ptr(), init_scale, _stream()), "scaler_init")


SCALER_DYNAMIC, SCALER_KEEP_SKIP = 1, 2      # b200_scaler_update mode bits


def scaler_update(state: torch.Tensor, dynamic, scale_window: int, scale_factor: float, min_scale: float,
                  max_scale: float, mirror: torch.Tensor | None = None):
    _lib.check(_lib.load().b200_scaler_update(state.data_ptr(), int(dynamic), scale_window, scale_factor, min_scale,
                                              max_scale, _addr(mirror), _stream()), "scaler_update")


def set_launch_config(ctas_per_sm: int = 0, threads: int = 256, unroll: int = 2):
    _lib.check(_lib.load().b200_set_launch_config(ctas_per_sm, threads, unroll), "set_launch_config")


def launch_count() -> int:
    return int(_lib.load().b200_launch_count())


def multi_tensor_l2norm(table: TensorTable, col: int, dtype, sumsq: torch.Tensor, scratch: torch.Tensor, dev_flag=None):
    """Adds the list's sum of squares to ``sumsq`` (fp32[1]) in one read-only pass; also the overflow check."""
    _require_cuda(table)
    _lib.check(_lib.load().b200_multi_tensor_l2norm(table.ptr, table.n_tiles, col, dtype_code(dtype), sumsq.data_ptr(),
                                                    scratch.data_ptr(), scratch.numel() * scratch.element_size(),
                                                    _addr(dev_flag), _stream()), "multi_tensor_l2norm")


LambGroup = _lib.LambGroup


def l2norm_per_tensor(table: TensorTable, col: int, dtype, first_tiles: torch.Tensor, tile_ss: torch.Tensor,
                      sumsq: torch.Tensor, norms: torch.Tensor | None = None, total: torch.Tensor | None = None,
                      dev_flag=None):
    """Per-tensor sums of squares (``sumsq``: fp32[n_tensors]) and norms of column ``col``; ADDS the list's total to
    ``total`` (fp32[1]).  ``first_tiles`` = table.first_tiles(); ``tile_ss`` = fp32 scratch of n_tiles elements."""
    _require_cuda(table)
    if tile_ss.numel() < table.n_tiles or sumsq.numel() < table.n_tensors:
        raise RuntimeError("l2norm_per_tensor: scratch or output too small")
    _lib.check(_lib.load().b200_l2norm_per_tensor(table.ptr, table.n_tiles, col, dtype_code(dtype), first_tiles.data_ptr(),
                                                  table.n_tensors, tile_ss.data_ptr(), sumsq.data_ptr(), _addr(norms),
                                                  _addr(total), _addr(dev_flag), _stream()), "l2norm_per_tensor")


def lamb_stage1(table: TensorTable, grad_dtype, state_dtype, groups: Sequence[LambGroup], grad_sumsq: torch.Tensor,
                max_grad_norm: float, inv_scale: float = 1.0, dev_inv_scale=None, dev_found_inf=None):
    _require_cuda(table)
    arr = (LambGroup * len(groups))(*groups)
    _lib.check(_lib.load().b200_lamb_stage1(table.ptr, table.n_tiles, dtype_code(grad_dtype), dtype_code(state_dtype), arr,
                                            len(groups), inv_scale, _addr(dev_inv_scale), _addr(dev_found_inf),
                                            grad_sumsq.data_ptr(), float(max_grad_norm), _stream()), "lamb_stage1")


def lamb_stage2(table: TensorTable, state_dtype, copy_dtype, groups: Sequence[LambGroup], p_norm: torch.Tensor,
                u_norm: torch.Tensor, dev_found_inf=None):
    _require_cuda(table)
    arr = (LambGroup * len(groups))(*groups)
    _lib.check(_lib.load().b200_lamb_stage2(table.ptr, table.n_tiles, dtype_code(state_dtype), dtype_code(copy_dtype), arr,
                                            len(groups), _addr(dev_found_inf), p_norm.data_ptr(), u_norm.data_ptr(),
                                            _stream()), "lamb_stage2")


NovoGradGroup = _lib.NovoGradGroup


def novograd_blend(table: TensorTable, groups: Sequence[NovoGradGroup], v: torch.Tensor, grad_sumsq: torch.Tensor,
                   inv_scale: float = 1.0, dev_inv_scale=None, dev_found_inf=None):
    """v[t] <- blend(v[t], norm of the unscaled gradient of tensor t); see b200_novograd_blend."""
    _require_cuda(table)
    arr = (NovoGradGroup * len(groups))(*groups)
    _lib.check(_lib.load().b200_novograd_blend(table.ptr, table.n_tensors, arr, len(groups), inv_scale, _addr(dev_inv_scale),
                                               _addr(dev_found_inf), v.data_ptr(), grad_sumsq.data_ptr(), _stream()),
               "novograd_blend")


def fused_novograd(table: TensorTable, grad_dtype, state_dtype, copy_dtype, groups: Sequence[NovoGradGroup], v: torch.Tensor,
                   inv_scale: float = 1.0, dev_inv_scale=None, dev_found_inf=None):
    _require_cuda(table)
    arr = (NovoGradGroup * len(groups))(*groups)
    _lib.check(_lib.load().b200_fused_novograd(table.ptr, table.n_tiles, dtype_code(grad_dtype), dtype_code(state_dtype),
                                               dtype_code(copy_dtype), arr, len(groups), inv_scale, _addr(dev_inv_scale),
                                               _addr(dev_found_inf), v.data_ptr(), _stream()), "fused_novograd")


def new_norm_scratch(device) -> torch.Tensor:
    return torch.zeros(_lib.NORM_SCRATCH_BYTES // 4, dtype=torch.float32, device=device)


def clip_scale(sumsq: torch.Tensor, max_norm: float, scale_out: torch.Tensor, norm_out=None, inv_scale: float = 1.0,
               dev_inv_scale=None):
    _lib.check(_lib.load().b200_clip_scale(sumsq.data_ptr(), _addr(dev_inv_scale), inv_scale, float(max_norm),
                                           scale_out.data_ptr(), _addr(norm_out), _stream()), "clip_scale")


def peer_gather(reduced_ptrs, local_out: torch.Tensor, bounds, dev_flag=None, sync=None):
    """All-gather half: pull every shard of the mean from its owner into ``local_out``."""
    arr = (ctypes.c_void_p * len(reduced_ptrs))(*reduced_ptrs)
    b = (ctypes.c_int64 * len(bounds))(*bounds)
    _lib.check(_lib.load().b200_peer_gather(arr, len(reduced_ptrs), local_out.data_ptr(), b, dtype_code(local_out.dtype),
                                            _addr(dev_flag), ctypes.byref(sync) if sync is not None else None, _stream()),
               "peer_gather")


def set_peer_grid(max_ctas: int = 0):
    _lib.check(_lib.load().b200_set_peer_grid(max_ctas), "set_peer_grid")


# ---- speculative single-pass step (b200_fused_adam_spec / b200_spec_commit / b200_spec_repair) -------------------
def spec_shadow_delta(numel: int, elem_size: int) -> int:
    """Bytes from a state tensor to its shadow copy."""
    return int(_lib.check(_lib.load().b200_spec_shadow_delta(numel, elem_size), "spec_shadow_delta", count=True))


def spec_state_like(t: torch.Tensor, fill=None):
    """Allocate a double-buffered state tensor shaped like ``t``: returns (primary view, whole buffer).
    The shadow copy starts spec_shadow_delta bytes after the primary one, in the same allocation."""
    es = t.element_size()
    delta = spec_shadow_delta(t.numel(), es) // es
    buf = torch.zeros(delta + t.numel(), dtype=t.dtype, device=t.device)
    prim = buf[:t.numel()].view(t.shape)
    if fill is not None:
        prim.copy_(fill)
    return prim, buf


def spec_current(prim: torch.Tensor, buf: torch.Tensor, parity: int) -> torch.Tensor:
    """The copy of a double-buffered state tensor that ``parity`` (0/1) makes current."""
    if not parity:
        return prim
    es = prim.element_size()
    delta = spec_shadow_delta(prim.numel(), es) // es
    return buf[delta:delta + prim.numel()].view(prim.shape)


def fused_adam_spec(table: TensorTable, grad_dtype, state_dtype, copy_dtype, groups: Sequence[AdamGroup], dev_flag,
                    dev_parity, inv_scale: float = 1.0, dev_inv_scale=None, sync: StepSync | None = None, pull=None):
    _require_cuda(table)
    arr = (AdamGroup * len(groups))(*groups)
    _lib.check(_lib.load().b200_fused_adam_spec(table.ptr, table.n_tiles, dtype_code(grad_dtype), dtype_code(state_dtype),
                                                dtype_code(copy_dtype), arr, len(groups), inv_scale, _addr(dev_inv_scale),
                                                _addr(dev_flag), _addr(dev_parity),
                                                ctypes.byref(sync) if sync is not None else None,
                                                ctypes.byref(pull) if pull is not None else None, _stream()),
               "fused_adam_spec")


def spec_commit(state: torch.Tensor, dynamic, scale_window: int, scale_factor: float, min_scale: float, max_scale: float,
                dev_parity, mirror: torch.Tensor | None = None, sync: StepSync | None = None):
    _lib.check(_lib.load().b200_spec_commit(state.data_ptr(), int(dynamic), scale_window, scale_factor, min_scale, max_scale,
                                            _addr(mirror), _addr(dev_parity),
                                            ctypes.byref(sync) if sync is not None else None, _stream()), "spec_commit")


def spec_repair(table: TensorTable, state_dtype, copy_dtype, dev_found_inf, dev_parity):
    _require_cuda(table)
    _lib.check(_lib.load().b200_spec_repair(table.ptr, table.n_tiles, dtype_code(state_dtype), dtype_code(copy_dtype),
                                            _addr(dev_found_inf), _addr(dev_parity), _stream()), "spec_repair")


def peer_spec_step(table: TensorTable, grad_dtype, state_dtype, copy_dtype, groups: Sequence[AdamGroup], dev_flag, dev_parity,
                   pull, peer_grad_ptrs, reduced: torch.Tensor, factor: float, psync, n_reducers: int, chunk_elems: int,
                   scratch: torch.Tensor, inv_scale: float = 1.0, dev_inv_scale=None, sync: StepSync | None = None):
    """Gradient mean across ranks + speculative Adam step in ONE kernel (b200_peer_spec_step)."""
    _require_cuda(table)
    arr = (AdamGroup * len(groups))(*groups)
    ptrs = (ctypes.c_void_p * len(peer_grad_ptrs))(*peer_grad_ptrs)
    _lib.check(_lib.load().b200_peer_spec_step(table.ptr, table.n_tiles, dtype_code(grad_dtype), dtype_code(state_dtype),
                                               dtype_code(copy_dtype), arr, len(groups), inv_scale, _addr(dev_inv_scale),
                                               _addr(dev_flag), _addr(dev_parity),
                                               ctypes.byref(sync) if sync is not None else None, ctypes.byref(pull), ptrs,
                                               reduced.data_ptr(), factor, ctypes.byref(psync), n_reducers, chunk_elems,
                                               scratch.data_ptr(), scratch.numel() * scratch.element_size(), _stream()),
               "peer_spec_step")


PeerFan = _lib.PeerFan


def make_peer_fan(grad_ptrs, model_ptrs, rank: int, factor: float, model_mc: int | None = None,
                  grads_mc: int | None = None) -> "PeerFan":
    """b200_peer_fan: every rank's flat gradient and model-weight buffer as addressable from this process."""
    fan = PeerFan()
    for r, (g, m) in enumerate(zip(grad_ptrs, model_ptrs)):
        fan.grads[r] = g
        fan.models[r] = m
    fan.world, fan.rank, fan.factor = len(grad_ptrs), rank, factor
    fan.model_mc = model_mc or None
    fan.grads_mc = grads_mc or None
    return fan


def peer_shard_step(table: TensorTable, grad_dtype, state_dtype, copy_dtype, groups: Sequence[AdamGroup], dev_parity, fan,
                    psync=None, dev_flag=None, inv_scale: float = 1.0, dev_inv_scale=None, sync: StepSync | None = None):
    """Reduce-scatter + speculative Adam on this rank's shard + all-gather of the new weights, one kernel
    (b200_peer_shard_step).  ``table`` lists the shard only."""
    _require_cuda(table)
    arr = (AdamGroup * len(groups))(*groups)
    _lib.check(_lib.load().b200_peer_shard_step(table.ptr, table.n_tiles, dtype_code(grad_dtype), dtype_code(state_dtype),
                                                dtype_code(copy_dtype), arr, len(groups), inv_scale, _addr(dev_inv_scale),
                                                _addr(dev_flag), _addr(dev_parity),
                                                ctypes.byref(sync) if sync is not None else None, ctypes.byref(fan),
                                                ctypes.byref(psync) if psync is not None else None, _stream()),
               "peer_shard_step")


def peer_shard_repair(table: TensorTable, state_dtype, copy_dtype, dev_found_inf, dev_parity, fan, psync=None):
    _require_cuda(table)
    _lib.check(_lib.load().b200_peer_shard_repair(table.ptr, table.n_tiles, dtype_code(state_dtype), dtype_code(copy_dtype),
                                                  _addr(dev_found_inf), _addr(dev_parity), ctypes.byref(fan),
                                                  ctypes.byref(psync) if psync is not None else None, _stream()),
               "peer_shard_repair")


def new_peer_step_scratch(device) -> torch.Tensor:
    return torch.zeros(2048 + 20 * 512, dtype=torch.int32, device=device)     # tickets, chunk counters (+ room for profile builds)


PEER_STEP_PAD_WORDS = _lib.PEER_STEP_PAD_BYTES // 4      # int32 words of a signal pad + progress board
