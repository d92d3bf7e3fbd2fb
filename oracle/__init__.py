"""CPU oracle for the mixed-precision step hot path — TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package; nothing under ``apex_example_b200/`` does.

Two layers:

* ``liboracle.so`` (``amp_step_oracle.c``, plain scalar C) — the restatement the CUDA kernels
  are compared with, bit for bit.  The functions here are thin ctypes wrappers that take
  **CPU** ``torch`` tensors.
* ``torch_*`` functions — the same steps written with torch 2.11's own CPU ops
  (``torch.optim.Adam/AdamW/SGD(foreach=False)``, ``torch._amp_*``).  They exist to PIN the C
  restatement (tests/test_oracle.py) and to generate tests/golden/ (make_golden.py).

PARITY STATUS: unpinned against apex (not vendored, not installable offline; the reference
repository holds no golden vectors for this path — SURVEY.md §8c); pinned against
torch.optim.Adam/SGD, which BASELINE.json's north_star names as the numerical contract.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass, field

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")

F32, F16, BF16 = 0, 1, 2
_DT = {torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16}


def build(force: bool = False) -> str:
    """Compile amp_step_oracle.c with the committed Makefile (gcc is in the image)."""
    src = os.path.join(_HERE, "amp_step_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-s", "-C", _HERE], check=True, capture_output=True)
    return _SO


class _AdamGroup(ctypes.Structure):
    _fields_ = [("lr", ctypes.c_double), ("beta1", ctypes.c_double), ("beta2", ctypes.c_double),
                ("eps", ctypes.c_double), ("weight_decay", ctypes.c_double),
                ("step", ctypes.c_int64), ("bias_correction", ctypes.c_int32),
                ("adam_w_mode", ctypes.c_int32)]


class _LambGroup(ctypes.Structure):
    _fields_ = [("lr", ctypes.c_double), ("beta1", ctypes.c_double), ("beta2", ctypes.c_double),
                ("eps", ctypes.c_double), ("weight_decay", ctypes.c_double),
                ("step", ctypes.c_int64), ("bias_correction", ctypes.c_int32),
                ("adam_w_mode", ctypes.c_int32), ("grad_averaging", ctypes.c_int32),
                ("use_nvlamb", ctypes.c_int32)]


class _NovoGroup(ctypes.Structure):
    _fields_ = [("lr", ctypes.c_double), ("beta1", ctypes.c_double), ("beta2", ctypes.c_double),
                ("eps", ctypes.c_double), ("weight_decay", ctypes.c_double),
                ("step", ctypes.c_int64), ("bias_correction", ctypes.c_int32),
                ("reg_inside_moment", ctypes.c_int32), ("grad_averaging", ctypes.c_int32),
                ("init_zero", ctypes.c_int32)]


class _SgdGroup(ctypes.Structure):
    _fields_ = [("lr", ctypes.c_double), ("momentum", ctypes.c_double),
                ("dampening", ctypes.c_double), ("weight_decay", ctypes.c_double),
                ("nesterov", ctypes.c_int32), ("first_run", ctypes.c_int32),
                ("wd_after_momentum", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class _ScalerState(ctypes.Structure):
    _fields_ = [("loss_scale", ctypes.c_float), ("inv_scale", ctypes.c_float),
                ("found_inf", ctypes.c_int32), ("unskipped", ctypes.c_int32),
                ("overflow_acc", ctypes.c_int32), ("seq", ctypes.c_int32),
                ("had_overflow", ctypes.c_int32), ("inv_loss_scale", ctypes.c_float),
                ("skipped_total", ctypes.c_int32), ("reserved", ctypes.c_int32 * 3)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.oracle_check.restype = ctypes.c_int
        if os.environ.get("ORACLE_FORCE_GENERIC"):
            ctypes.c_int.in_dll(_lib, "oracle_force_generic").value = 1
    return _lib


def _p(t):
    if t is None:
        return ctypes.c_void_p(0)
    assert t.device.type == "cpu" and t.is_contiguous(), "oracle takes contiguous CPU tensors"
    return ctypes.c_void_p(t.data_ptr())


def _dt(t):
    return _DT[t.dtype] if t is not None else F32


# --------------------------------------------------------------------------- C oracle wrappers
def convert(src: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    out = torch.empty(src.shape, dtype=dtype)
    lib().oracle_convert(_p(src), _dt(src), _p(out), _DT[dtype], ctypes.c_int64(src.numel()))
    return out


def check(g: torch.Tensor) -> int:
    return int(lib().oracle_check(_p(g), _dt(g), ctypes.c_int64(g.numel())))


def scale(src, dst, factor: float) -> int:
    flag = ctypes.c_int(0)
    lib().oracle_scale(_p(src), _dt(src), _p(dst), _dt(dst), ctypes.c_int64(src.numel()),
                       ctypes.c_float(factor), ctypes.byref(flag))
    return flag.value


def axpby(x, y, out, a: float, b: float, arg_to_check: int = -1) -> int:
    flag = ctypes.c_int(0)
    lib().oracle_axpby(_p(x), _dt(x), _p(y), _dt(y), _p(out), _dt(out), ctypes.c_int64(x.numel()),
                       ctypes.c_float(a), ctypes.c_float(b), ctypes.c_int(arg_to_check),
                       ctypes.byref(flag))
    return flag.value


def adam(g, p, m, v, copy=None, *, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0,
         step=1, bias_correction=True, adam_w_mode=True, inv_scale=1.0) -> None:
    """In place on p, m, v (and copy).  See oracle_adam in amp_step_oracle.c."""
    grp = _AdamGroup(lr, beta1, beta2, eps, weight_decay, step, int(bias_correction), int(adam_w_mode))
    assert p.dtype == m.dtype == v.dtype
    lib().oracle_adam(_p(g), _dt(g), _p(p), _p(m), _p(v), _dt(p), _p(copy), _dt(copy),
                      ctypes.c_int64(p.numel()), ctypes.byref(grp), ctypes.c_float(inv_scale))


def sgd(g, p, mom, copy=None, *, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False,
        first_run=False, wd_after_momentum=False, inv_scale=1.0) -> None:
    grp = _SgdGroup(lr, momentum, dampening, weight_decay, int(nesterov), int(first_run),
                    int(wd_after_momentum), 0)
    lib().oracle_sgd(_p(g), _dt(g), _p(p), _p(mom), _dt(p), _p(copy), _dt(copy),
                     ctypes.c_int64(p.numel()), ctypes.byref(grp), ctypes.c_float(inv_scale))


def bucket_epilogue(x, factor: float) -> int:
    flag = ctypes.c_int(0)
    lib().oracle_bucket_epilogue(_p(x), _dt(x), ctypes.c_int64(x.numel()), ctypes.c_float(factor),
                                 ctypes.byref(flag))
    return flag.value


@dataclass
class LossScaler:
    """apex.amp.scaler.LossScaler restated on top of oracle_scaler_update."""
    loss_scale: float = 2.0 ** 16
    dynamic: bool = True
    scale_window: int = 2000
    scale_factor: float = 2.0
    min_loss_scale: float | None = None
    max_loss_scale: float = 2.0 ** 24
    unskipped: int = 0
    _s: _ScalerState = field(default=None, repr=False)

    def __post_init__(self):
        if self.dynamic:
            self.loss_scale = min(self.max_loss_scale, self.loss_scale)
        self._s = _ScalerState(self.loss_scale, 1.0 / self.loss_scale, 0, self.unskipped, 0, 0, 0, 1.0 / self.loss_scale, 0)

    def update_scale(self, overflow: bool, keep_skip: bool = False) -> bool:
        """Returns should_skip.  ``keep_skip``: no optimizer step ran since the previous update."""
        self._s.overflow_acc = int(bool(overflow))
        lib().oracle_scaler_update(ctypes.byref(self._s), ctypes.c_int(int(self.dynamic) | (2 if keep_skip else 0)),
                                   ctypes.c_int(self.scale_window), ctypes.c_float(self.scale_factor),
                                   ctypes.c_float(self.min_loss_scale or 0.0),
                                   ctypes.c_float(self.max_loss_scale))
        self.loss_scale = float(self._s.loss_scale)
        self.unskipped = int(self._s.unskipped)
        return bool(self._s.found_inf)

    @property
    def inv_scale(self) -> float:
        return float(self._s.inv_scale)


def _ptr_array(tensors):
    arr = (ctypes.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


def adam_list(gs, ps, ms, vs, copies=None, *, check_overflow=True, inv_scale=1.0, **hyper) -> int:
    """Whole-list O2 step (check, then unscale+Adam+copy unless overflow).  The timed CPU baseline."""
    grp = _AdamGroup(hyper.get("lr", 1e-3), hyper.get("beta1", 0.9), hyper.get("beta2", 0.999),
                     hyper.get("eps", 1e-8), hyper.get("weight_decay", 0.0), hyper.get("step", 1),
                     int(hyper.get("bias_correction", True)), int(hyper.get("adam_w_mode", True)))
    n = len(gs)
    numel = (ctypes.c_int64 * n)(*[t.numel() for t in ps])
    fn = lib().oracle_adam_list
    fn.restype = ctypes.c_int
    return int(fn(ctypes.c_int(n), _ptr_array(gs), _dt(gs[0]), _ptr_array(ps), _ptr_array(ms),
                  _ptr_array(vs), _dt(ps[0]),
                  _ptr_array(copies) if copies else None, _dt(copies[0]) if copies else F32,
                  numel, ctypes.byref(grp), ctypes.c_float(inv_scale), ctypes.c_int(int(check_overflow))))


def sgd_list(gs, ps, moms, copies=None, *, check_overflow=True, inv_scale=1.0, **hyper) -> int:
    grp = _SgdGroup(hyper["lr"], hyper.get("momentum", 0.0), hyper.get("dampening", 0.0),
                    hyper.get("weight_decay", 0.0), int(hyper.get("nesterov", False)),
                    int(hyper.get("first_run", False)), int(hyper.get("wd_after_momentum", False)), 0)
    n = len(gs)
    numel = (ctypes.c_int64 * n)(*[t.numel() for t in ps])
    fn = lib().oracle_sgd_list
    fn.restype = ctypes.c_int
    return int(fn(ctypes.c_int(n), _ptr_array(gs), _dt(gs[0]), _ptr_array(ps),
                  _ptr_array(moms) if moms else None, _dt(ps[0]),
                  _ptr_array(copies) if copies else None, _dt(copies[0]) if copies else F32,
                  numel, ctypes.byref(grp), ctypes.c_float(inv_scale), ctypes.c_int(int(check_overflow))))


def l2norm_per_tensor(xs, tile_elems: int):
    """(per-tensor sums of squares as a float32 tensor, total as a python float, inf/NaN flag), summed in the order
    the contract fixes (see amp_step_oracle.c)."""
    n = len(xs)
    numel = (ctypes.c_int64 * n)(*[t.numel() for t in xs])
    out = torch.zeros(max(n, 1), dtype=torch.float32)
    total = ctypes.c_float(0.0)
    fn = lib().oracle_l2norm_per_tensor
    fn.restype = ctypes.c_int
    flag = int(fn(ctypes.c_int(n), _ptr_array(xs), _dt(xs[0]) if n else F32, numel, ctypes.c_int(tile_elems),
                  ctypes.c_void_p(out.data_ptr()), ctypes.byref(total)))
    return out[:n], total.value, flag


def lamb_list(gs, ps, ms, vs, copies=None, *, tile_elems: int, grad_sumsq: float, inv_scale=1.0, max_grad_norm=0.0,
              **hyper) -> None:
    """One LAMB step over a list with one hyper-parameter group (state updated in place)."""
    grp = _LambGroup(hyper.get("lr", 1e-3), hyper.get("beta1", 0.9), hyper.get("beta2", 0.999),
                     hyper.get("eps", 1e-6), hyper.get("weight_decay", 0.01), hyper.get("step", 1),
                     int(hyper.get("bias_correction", True)), int(hyper.get("adam_w_mode", True)),
                     int(hyper.get("grad_averaging", True)), int(hyper.get("use_nvlamb", False)))
    n = len(gs)
    numel = (ctypes.c_int64 * n)(*[t.numel() for t in ps])
    us = [torch.zeros(t.numel(), dtype=torch.float32) for t in ps]
    lib().oracle_lamb_list(ctypes.c_int(n), _ptr_array(gs), _dt(gs[0]), _ptr_array(ps), _ptr_array(ms), _ptr_array(vs),
                           _dt(ps[0]), _ptr_array(copies) if copies else None, _dt(copies[0]) if copies else F32,
                           _ptr_array(us), numel, ctypes.c_int(tile_elems), ctypes.byref(grp),
                           ctypes.c_float(inv_scale), ctypes.c_float(grad_sumsq), ctypes.c_float(max_grad_norm))


def novograd_list(gs, ps, ms, v, grad_sumsq, copies=None, *, inv_scale=1.0, **hyper) -> None:
    """One NovoGrad step over a list with one hyper-parameter group.  ``v``: float32 tensor, one entry per tensor
    (updated in place); ``grad_sumsq``: per-tensor sums of squares of the still scaled gradients."""
    grp = _NovoGroup(hyper.get("lr", 1e-3), hyper.get("beta1", 0.9), hyper.get("beta2", 0.999),
                     hyper.get("eps", 1e-8), hyper.get("weight_decay", 0.0), hyper.get("step", 1),
                     int(hyper.get("bias_correction", True)), int(hyper.get("reg_inside_moment", False)),
                     int(hyper.get("grad_averaging", True)), int(hyper.get("init_zero", False)))
    n = len(gs)
    numel = (ctypes.c_int64 * n)(*[t.numel() for t in ps])
    assert v.dtype == torch.float32 and grad_sumsq.dtype == torch.float32 and v.numel() >= n
    lib().oracle_novograd_list(ctypes.c_int(n), _ptr_array(gs), _dt(gs[0]), _ptr_array(ps), _ptr_array(ms), _dt(ps[0]),
                               _ptr_array(copies) if copies else None, _dt(copies[0]) if copies else F32,
                               ctypes.c_void_p(v.data_ptr()), ctypes.c_void_p(grad_sumsq.data_ptr()), numel,
                               ctypes.byref(grp), ctypes.c_float(inv_scale))


# ------------------------------------------------------------------- torch-op restatement (pin)
def torch_unscale_check(grads, inv_scale: float):
    """fp32 master grads + flag via torch's own op (TORCH/amp/grad_scaler.py:292-298)."""
    master = [g.float() for g in grads]
    found = torch.zeros(1)
    torch._amp_foreach_non_finite_check_and_unscale_(master, found, torch.tensor(inv_scale))
    return master, int(found.item())


def torch_adam_steps(p0, grads_per_step, *, adamw: bool, inv_scale=1.0, **kw):
    """Run torch.optim.Adam/AdamW(foreach=False) on a clone of p0 (fp32) over a list of per-step
    gradients (any dtype; unscaled here like LossScaler.unscale).  Returns p, exp_avg, exp_avg_sq."""
    p = torch.nn.Parameter(p0.clone().float())
    cls = torch.optim.AdamW if adamw else torch.optim.Adam
    opt = cls([p], foreach=False, fused=False, **kw)
    for g in grads_per_step:
        p.grad = g.float() * inv_scale
        opt.step()
    st = opt.state[p]
    return p.detach(), st["exp_avg"], st["exp_avg_sq"]


def torch_sgd_steps(p0, grads_per_step, *, inv_scale=1.0, **kw):
    p = torch.nn.Parameter(p0.clone().float())
    opt = torch.optim.SGD([p], foreach=False, fused=False, **kw)
    for g in grads_per_step:
        p.grad = g.float() * inv_scale
        opt.step()
    return p.detach(), opt.state[p].get("momentum_buffer")


def torch_update_scale(scale: float, growth_tracker: int, found_inf: int, window: int):
    """torch._amp_update_scale_ (TORCH/amp/grad_scaler.py:549-556): no max/min caps."""
    s = torch.tensor(scale, dtype=torch.float32)
    t = torch.tensor(growth_tracker, dtype=torch.int32)
    torch._amp_update_scale_(s, t, torch.tensor(float(found_inf)), 2.0, 0.5, window)
    return float(s), int(t)


def peer_reduce(bufs, out, begin: int, end: int, factor: float) -> int:
    """Mean of one shard over per-rank flat buffers (oracle_peer_reduce)."""
    flag = ctypes.c_int(0)
    lib().oracle_peer_reduce(_ptr_array(bufs), ctypes.c_int(len(bufs)), _p(out), _dt(out), ctypes.c_int64(begin),
                             ctypes.c_int64(end), ctypes.c_float(factor), ctypes.byref(flag))
    return flag.value
