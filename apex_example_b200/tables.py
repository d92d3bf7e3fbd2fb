"""S1 — tensor-list tables: the host side of the multi_tensor_apply chunk scheduler.

apex re-packs tensor addresses into kernel arguments on every ``multi_tensor_applier`` call
(SURVEY.md §2.4 K1).  Here a list is described once by a table image (include/apex_b200.h),
uploaded to device memory, and reused for as long as the tensors' addresses do not change.
"""
from __future__ import annotations

import ctypes
from typing import Sequence

import torch

from . import _lib

_DT = {torch.float32: _lib.F32, torch.float16: _lib.F16, torch.bfloat16: _lib.BF16}


def dtype_code(dt) -> int:
    if dt is None:
        return _lib.NONE
    try:
        return _DT[dt]
    except KeyError:
        raise RuntimeError(f"libapex_b200 only supports fp16, bf16 and fp32 tensors, got {dt}") from None


def layout_key(x: torch.Tensor):
    """None for the default (row-major contiguous) layout, else (shape, strides).  The kernels treat every
    tensor as ``numel`` consecutive elements in MEMORY order, so the columns of one table row may use any
    dense layout (e.g. channels_last convolution weights) as long as they all use the same one."""
    if x.is_contiguous():
        return None
    return (tuple(x.shape), tuple(x.stride()))


def is_dense(x: torch.Tensor) -> bool:
    """True when the tensor's elements occupy ``numel`` consecutive slots of memory, in any dimension order."""
    if x.is_contiguous():
        return True
    expect = 1
    for stride, size in sorted((st, sz) for sz, st in zip(x.shape, x.stride()) if sz != 1):
        if stride != expect:
            return False
        expect *= size
    return True


def flat_view_like(flat: torch.Tensor, offset: int, like: torch.Tensor) -> torch.Tensor:
    """View of ``flat[offset : offset + like.numel()]`` with ``like``'s shape AND memory layout."""
    v = flat[offset:offset + like.numel()]
    if like.is_contiguous() or not is_dense(like):
        return v.view(like.shape)
    return v.as_strided(like.size(), like.stride())


def auto_tile_elems(columns) -> int:
    """Tile size by measurement (profiles/r1_tuning.md): a tile should be exactly one trip of a
    CTA — 256 threads x 8 elements x the kernel's groups in flight.  The 4-5 column optimizer
    kernels keep 2 groups in flight (4096 elements, ~114 KB of traffic); the one-column overflow
    check keeps 8 and wants at least 32 KB per tile, so it gets 32768 elements."""
    width = 0
    for col in columns:
        t = next((x for x in col if x is not None), None)
        if t is not None:
            width += t.element_size()
    if width <= 4:
        return 32768
    # all-16-bit lists (amp O3: g, p, m, v in fp16/bf16; 8 bytes per element over the columns): the step kernels
    # keep 4 groups in flight there, i.e. 8192 elements per trip (measured on the 34k-tensor config-5 list:
    # 4.57 ms with 4096-element tiles, 3.99 ms with 8192, profiles/r02_summary.md)
    return 8192 if width <= 8 else _lib.TILE_ELEMS


def build_image(columns: Sequence[Sequence[torch.Tensor | None]], groups: Sequence[int] | None = None,
                tile_elems: int | None = None, pin: bool = False):
    """Build the host image for ``columns`` (column-major: columns[d][t]).

    Returns (uint8 host tensor, n_tiles).  Pure host code — works without a GPU.
    """
    lib = _lib.load()
    if tile_elems is None:
        tile_elems = auto_tile_elems(columns)
    depth = len(columns)
    n = len(columns[0]) if depth else 0
    if depth < 1 or depth > _lib.MAX_DEPTH:
        raise RuntimeError(f"tensor_lists depth must be 1..{_lib.MAX_DEPTH}, got {depth}")
    ref = None
    for col in columns:
        if len(col) != n:
            raise RuntimeError("tensor_lists must all have the same length")
        if ref is None and all(t is not None for t in col):
            ref = col
    if ref is None:
        raise RuntimeError("at least one column must be fully present")
    numel = (ctypes.c_int64 * max(n, 1))()
    ptrs = (ctypes.c_void_p * max(n * depth, 1))()
    esize = (ctypes.c_int32 * depth)()
    for t in range(n):
        numel[t] = ref[t].numel()
    layouts = [layout_key(x) for x in ref]
    for d, col in enumerate(columns):
        es = 4
        for t, x in enumerate(col):
            if x is None:
                ptrs[d * n + t] = None
                continue
            if x.numel() != numel[t]:
                raise RuntimeError(f"size mismatch in tensor_lists[{d}][{t}]: {x.numel()} vs {numel[t]}")
            if layout_key(x) != layouts[t]:
                raise RuntimeError(f"tensor_lists[{d}][{t}] has another memory layout than the other columns of "
                                   f"tensor {t} (shape {tuple(x.shape)}, strides {tuple(x.stride())})")
            if layouts[t] is not None and not is_dense(x):
                raise RuntimeError(f"tensor_lists[{d}][{t}] is not contiguous")
            es = x.element_size()
            ptrs[d * n + t] = x.data_ptr()
        esize[d] = es if es in (2, 4) else 4
    grp = None
    if groups is not None:
        grp = (ctypes.c_int32 * max(n, 1))(*groups)
    nbytes = _lib.check(lib.b200_table_bytes(n, numel, tile_elems), "table_bytes", count=True)
    image = torch.empty(nbytes, dtype=torch.uint8, pin_memory=pin)
    n_tiles = _lib.check(lib.b200_table_build(image.data_ptr(), nbytes, n, depth, ptrs, esize, numel, grp,
                                              tile_elems), "table_build", count=True)
    return image, int(n_tiles)


class TensorTable:
    """A table image resident in device memory plus what the launchers need on the host."""

    __slots__ = ("image", "dev", "n_tiles", "n_tensors", "total_elems", "signature")

    def __init__(self, columns, groups=None, tile_elems: int | None = None, device=None, signature=None):
        ref = next(c for c in columns if all(t is not None for t in c)) if columns and columns[0] else []
        if device is None:
            device = ref[0].device if len(ref) else torch.device("cuda")
        self.image, self.n_tiles = build_image(columns, groups, tile_elems, pin=torch.cuda.is_available())
        self.dev = self.image.to(device, non_blocking=True)
        self.n_tensors = len(ref)
        self.total_elems = sum(t.numel() for t in ref)
        self.signature = signature

    @property
    def ptr(self) -> int:
        return self.dev.data_ptr()

    @property
    def tile_elems(self) -> int:
        return int(self.image[16:20].view(torch.int32)[0])

    def first_tiles(self) -> torch.Tensor:
        """int32[n_tensors + 1] on the table's device: exclusive prefix sum of the tensors' tile counts (what
        b200_l2norm_per_tensor takes as dev_first_tile), read back from the image's own rows."""
        rows_off = int(self.image[24:32].view(torch.int64)[0])
        rows = self.image[rows_off:rows_off + 64 * self.n_tensors].view(torch.int64).view(self.n_tensors, 8)
        numel = rows[:, _lib.MAX_DEPTH]
        te = self.tile_elems
        first = torch.zeros(self.n_tensors + 1, dtype=torch.int64)
        torch.cumsum((numel + te - 1) // te, 0, out=first[1:])
        return first.to(torch.int32).to(self.dev.device)

    def order_for_peer_step(self, local_base_ptr: int, grad_elem_size: int, bounds, rank: int, chunk_elems: int):
        """Re-order the tile list for b200_peer_spec_step (chunk-major, this rank's shard first) and upload
        the image again.  ``bounds`` are the world+1 tile-aligned shard offsets of the flat gradient buffer."""
        b = (ctypes.c_int64 * len(bounds))(*[int(x) for x in bounds])
        _lib.check(_lib.load().b200_table_order_for_peer_step(self.image.data_ptr(), local_base_ptr, grad_elem_size, b,
                                                              len(bounds) - 1, rank, chunk_elems),
                   "table_order_for_peer_step")
        self.dev.copy_(self.image, non_blocking=True)
        return self


def signature_of(columns) -> tuple:
    """Cheap identity of a list of columns: the device addresses (sizes are fixed per parameter)."""
    return tuple(tuple(0 if t is None else t.data_ptr() for t in col) for col in columns)


def tile_aligned_shards(offsets: Sequence[int], sizes: Sequence[int], total: int, world: int,
                        tile_elems: int = _lib.TILE_ELEMS, return_first_tiles: bool = False):
    """Split a flat buffer (tensor t at element offset offsets[t], sizes[t] long) into `world`
    contiguous shards of roughly equal work whose boundaries never cut a tile of the optimizer's
    table — the precondition of b200_fused_*_peer.  Returns world+1 element offsets and, on
    request, each rank's first tile index (valid when the table lists the tensors in flat order)."""
    starts = []
    for off, n in zip(offsets, sizes):
        starts.extend(range(off, off + n, tile_elems))
    if starts != sorted(starts):
        if return_first_tiles:
            raise RuntimeError("tile_aligned_shards: tensors must be listed in flat-buffer order")
        starts.sort()
    bounds, first = [0], [0]
    for r in range(1, world):
        k = round(r * len(starts) / world)
        bounds.append(starts[k] if k < len(starts) else total)
        first.append(min(k, max(len(starts) - 1, 0)))
    bounds.append(total)
    for a_, b_ in zip(bounds, bounds[1:]):
        assert a_ <= b_
    return (bounds, first) if return_first_tiles else bounds
