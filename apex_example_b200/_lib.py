"""ctypes binding of libapex_b200.so (the C ABI declared in include/apex_b200.h).

There is deliberately no fallback: if the shared library is missing or a call fails, a
RuntimeError is raised.  Nothing here (or anywhere under apex_example_b200/) imports oracle/.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("APEX_B200_SO") or os.path.join(_HERE, "libapex_b200.so")   # override: A/B experiments
CSRC = os.path.join(_HERE, "csrc")

F32, F16, BF16, NONE = 0, 1, 2, -1
MAX_DEPTH, MAX_GROUPS, TILE_ELEMS = 5, 8, 4096

c_void_p, c_int32, c_int64, c_float, c_double = (ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64,
                                                 ctypes.c_float, ctypes.c_double)


class AdamGroup(ctypes.Structure):           # b200_adam_group
    _fields_ = [("lr", c_double), ("beta1", c_double), ("beta2", c_double), ("eps", c_double),
                ("weight_decay", c_double), ("step", c_int64), ("bias_correction", c_int32),
                ("adam_w_mode", c_int32)]


class LambGroup(ctypes.Structure):           # b200_lamb_group
    _fields_ = [("lr", c_double), ("beta1", c_double), ("beta2", c_double), ("eps", c_double),
                ("weight_decay", c_double), ("step", c_int64), ("bias_correction", c_int32),
                ("adam_w_mode", c_int32), ("grad_averaging", c_int32), ("use_nvlamb", c_int32)]


class NovoGradGroup(ctypes.Structure):       # b200_novograd_group
    _fields_ = [("lr", c_double), ("beta1", c_double), ("beta2", c_double), ("eps", c_double),
                ("weight_decay", c_double), ("step", c_int64), ("bias_correction", c_int32),
                ("reg_inside_moment", c_int32), ("grad_averaging", c_int32), ("init_zero", c_int32)]


class SgdGroup(ctypes.Structure):            # b200_sgd_group
    _fields_ = [("lr", c_double), ("momentum", c_double), ("dampening", c_double),
                ("weight_decay", c_double), ("step", c_int64), ("nesterov", c_int32),
                ("wd_after_momentum", c_int32)]


class StepSync(ctypes.Structure):            # b200_step_sync
    _fields_ = [("dev_counters", c_void_p), ("mirror", c_void_p), ("known_skips", c_int32),
                ("read_only", c_int32)]

STEP_WINDOW = 4
MAX_PEERS = 8


class PeerSync(ctypes.Structure):            # b200_peer_sync
    _fields_ = [("pads", ctypes.POINTER(c_void_p)), ("world", c_int32), ("rank", c_int32), ("epoch", c_int32),
                ("reserved", c_int32)]


SIGNAL_PAD_BYTES = 128
PEER_STEP_MAX_CHUNKS = 1024
PEER_STEP_PAD_BYTES = 128 + MAX_PEERS * PEER_STEP_MAX_CHUNKS * 4     # signal pad followed by the progress board


class PeerPull(ctypes.Structure):            # b200_peer_pull
    _fields_ = [("local_base", c_void_p), ("bounds", c_int64 * (MAX_PEERS + 1)), ("delta", c_int64 * MAX_PEERS),
                ("world", c_int32), ("elem_size", c_int32), ("first_tile", ctypes.c_uint32), ("reserved", ctypes.c_uint32)]

class PeerFan(ctypes.Structure):             # b200_peer_fan
    _fields_ = [("grads", c_void_p * MAX_PEERS), ("models", c_void_p * MAX_PEERS), ("world", c_int32), ("rank", c_int32),
                ("factor", c_float), ("reserved", c_int32), ("model_mc", c_void_p), ("grads_mc", c_void_p)]


NOALLOC_TABLE_BYTES = 256 * 1024
NORM_SCRATCH_BYTES = 16384


class TableHeader(ctypes.Structure):         # b200_table_header
    _fields_ = [("magic", ctypes.c_uint32), ("version", ctypes.c_uint32), ("n_tensors", ctypes.c_uint32),
                ("depth", ctypes.c_uint32), ("tile_elems", ctypes.c_uint32), ("n_tiles", ctypes.c_uint32),
                ("rows_off", ctypes.c_uint64), ("tiles_off", ctypes.c_uint64), ("total_elems", ctypes.c_uint64),
                ("image_bytes", ctypes.c_uint64), ("reserved", ctypes.c_uint64)]


class TableRow(ctypes.Structure):            # b200_table_row
    _fields_ = [("ptr", c_void_p * MAX_DEPTH), ("numel", c_int64), ("group", ctypes.c_uint32),
                ("flags", ctypes.c_uint32), ("reserved", ctypes.c_uint64)]


class TableTile(ctypes.Structure):           # b200_table_tile
    _fields_ = [("tensor", ctypes.c_uint32), ("tile", ctypes.c_uint32)]


# name -> (restype, argtypes); must list every symbol include/apex_b200.h declares
SIGNATURES = {
    "b200_abi_version": (c_int32, []),
    "b200_last_error": (ctypes.c_char_p, []),
    "b200_table_bytes": (c_int64, [c_int32, ctypes.POINTER(c_int64), c_int32]),
    "b200_table_build": (c_int64, [c_void_p, c_int64, c_int32, c_int32, ctypes.POINTER(c_void_p),
                                   ctypes.POINTER(c_int32), ctypes.POINTER(c_int64),
                                   ctypes.POINTER(c_int32), c_int32]),
    "b200_table_upload": (c_int32, [c_void_p, c_void_p, c_int64, c_void_p]),
    "b200_scaler_init": (c_int32, [c_void_p, c_float, c_void_p]),
    "b200_scaler_update": (c_int32, [c_void_p, c_int32, c_int32, c_float, c_float, c_float, c_void_p, c_void_p]),
    "b200_multi_tensor_scale": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_float, c_void_p, c_void_p, c_void_p]),
    "b200_multi_tensor_axpby": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_int32, c_float, c_float,
                                          c_int32, c_void_p, c_void_p]),
    "b200_multi_tensor_check": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_void_p, c_void_p]),
    "b200_multi_tensor_l2norm": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_void_p, c_void_p, c_int64, c_void_p,
                                           c_void_p]),
    "b200_l2norm_per_tensor": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_void_p, c_int32, c_void_p, c_void_p,
                                         c_void_p, c_void_p, c_void_p, c_void_p]),
    "b200_lamb_stage1": (c_int32, [c_void_p, c_int64, c_int32, c_int32, ctypes.POINTER(LambGroup), c_int32, c_float,
                                   c_void_p, c_void_p, c_void_p, c_float, c_void_p]),
    "b200_lamb_stage2": (c_int32, [c_void_p, c_int64, c_int32, c_int32, ctypes.POINTER(LambGroup), c_int32, c_void_p,
                                   c_void_p, c_void_p, c_void_p]),
    "b200_novograd_blend": (c_int32, [c_void_p, c_int32, ctypes.POINTER(NovoGradGroup), c_int32, c_float, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_void_p]),
    "b200_fused_novograd": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_int32, ctypes.POINTER(NovoGradGroup), c_int32,
                                      c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b200_clip_scale": (c_int32, [c_void_p, c_void_p, c_float, c_float, c_void_p, c_void_p, c_void_p]),
    "b200_fused_adam": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_int32, ctypes.POINTER(AdamGroup),
                                  c_int32, c_float, c_void_p, c_void_p, ctypes.POINTER(StepSync), c_void_p]),
    "b200_fused_sgd": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_int32, ctypes.POINTER(SgdGroup),
                                 c_int32, c_float, c_void_p, c_void_p, ctypes.POINTER(StepSync), c_void_p]),
    "b200_peer_alloc": (c_int32, [c_int64, ctypes.POINTER(c_void_p), c_void_p]),
    "b200_peer_open": (c_int32, [c_void_p, ctypes.POINTER(c_void_p)]),
    "b200_peer_close": (c_int32, [c_void_p]),
    "b200_peer_free": (c_int32, [c_void_p]),
    "b200_peer_reduce_shard": (c_int32, [ctypes.POINTER(c_void_p), c_int32, c_void_p, c_int64, c_int64, c_int32, c_float,
                                         c_void_p, ctypes.POINTER(PeerSync), c_void_p]),
    "b200_peer_wait_flags": (c_int32, [ctypes.POINTER(PeerSync), c_void_p, c_void_p]),
    "b200_peer_gather": (c_int32, [ctypes.POINTER(c_void_p), c_int32, c_void_p, ctypes.POINTER(c_int64), c_int32, c_void_p,
                                   ctypes.POINTER(PeerSync), c_void_p]),
    "b200_set_peer_grid": (c_int32, [c_int32]),
    "b200_fused_adam_peer": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_int32, ctypes.POINTER(AdamGroup),
                                       c_int32, c_float, c_void_p, c_void_p, ctypes.POINTER(StepSync),
                                       ctypes.POINTER(PeerPull), c_void_p]),
    "b200_fused_sgd_peer": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_int32, ctypes.POINTER(SgdGroup),
                                      c_int32, c_float, c_void_p, c_void_p, ctypes.POINTER(StepSync),
                                      ctypes.POINTER(PeerPull), c_void_p]),
    "b200_spec_shadow_delta": (c_int64, [c_int64, c_int32]),
    "b200_fused_adam_spec": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_int32, ctypes.POINTER(AdamGroup), c_int32,
                                       c_float, c_void_p, c_void_p, c_void_p, ctypes.POINTER(StepSync),
                                       ctypes.POINTER(PeerPull), c_void_p]),
    "b200_spec_commit": (c_int32, [c_void_p, c_int32, c_int32, c_float, c_float, c_float, c_void_p, c_void_p,
                                   ctypes.POINTER(StepSync), c_void_p]),
    "b200_spec_repair": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_void_p, c_void_p, c_void_p]),
    "b200_table_order_for_peer_step": (c_int32, [c_void_p, c_void_p, c_int32, ctypes.POINTER(c_int64), c_int32, c_int32,
                                                 c_int64]),
    "b200_peer_spec_step": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_int32, ctypes.POINTER(AdamGroup), c_int32,
                                      c_float, c_void_p, c_void_p, c_void_p, ctypes.POINTER(StepSync),
                                      ctypes.POINTER(PeerPull), ctypes.POINTER(c_void_p), c_void_p, c_float,
                                      ctypes.POINTER(PeerSync), c_int32, c_int64, c_void_p, c_int64, c_void_p]),
    "b200_peer_shard_step": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_int32, ctypes.POINTER(AdamGroup), c_int32,
                                       c_float, c_void_p, c_void_p, c_void_p, ctypes.POINTER(StepSync),
                                       ctypes.POINTER(PeerFan), ctypes.POINTER(PeerSync), c_void_p]),
    "b200_peer_shard_repair": (c_int32, [c_void_p, c_int64, c_int32, c_int32, c_void_p, c_void_p, ctypes.POINTER(PeerFan),
                                         ctypes.POINTER(PeerSync), c_void_p]),
    "b200_fused_adam_noalloc": (c_int32, [c_int32, ctypes.POINTER(c_void_p), ctypes.POINTER(c_int64), c_int32, c_int32,
                                          c_int32, ctypes.POINTER(AdamGroup), c_float, c_void_p, c_void_p, c_void_p]),
    "b200_bucket_epilogue": (c_int32, [c_void_p, c_int64, c_int32, c_float, c_void_p, c_void_p]),
    "b200_set_launch_config": (c_int32, [c_int32, c_int32, c_int32]),
    "b200_launch_count": (c_int64, []),
}

_lib = None


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile libapex_b200.so in-tree with nvcc for sm_100a (csrc/Makefile)."""
    cmd = ["make", "-C", CSRC] + (["-B"] if force else [])
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose:
        print(r.stdout, r.stderr)
    if r.returncode != 0:
        raise RuntimeError("building libapex_b200.so failed:\n" + r.stdout + r.stderr)
    return SO_PATH


def load():
    """Return the loaded library; raises RuntimeError when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise RuntimeError(
                f"{SO_PATH} is missing: the sm_100a CUDA library has not been built "
                "(run `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
        lib = ctypes.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)          # AttributeError here = header/library mismatch
            fn.restype, fn.argtypes = res, args
        if lib.b200_abi_version() != 1:
            raise RuntimeError("libapex_b200.so ABI version mismatch")
        _lib = lib
    return _lib


def check(rc: int, what: str = "", count: bool = False) -> int:
    """Map the C ABI's status codes to RuntimeError (SURVEY.md §8b error conventions).
    `count=True` is for the entry points that return a non-negative size on success."""
    if rc < 0 or (rc > 0 and not count):
        msg = load().b200_last_error().decode(errors="replace")
        raise RuntimeError(f"libapex_b200 {what or 'call'} failed (code {rc}): {msg}")
    return rc
