"""Peer-addressable gradient buffers for the fused reduce-scatter / optimizer path (S3').

Plumbing only: every rank allocates its flat buffers through the library (cudaMalloc + a CUDA IPC
handle), the 64-byte handles are gathered over the process group, and every rank maps its peers'
buffers into its own address space with peer access enabled.  The arithmetic is in libapex_b200
(b200_peer_reduce_shard, b200_fused_*_peer); NCCL carries two 4-byte all-reduces per step that act
as cross-rank barriers and OR the overflow flags.

Single node only (NVLink / NVSwitch): every pair of GPUs must be peer-accessible.
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from .. import _lib, ops

_TYPESTR = {torch.float16: "<f2", torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1"}


class _RawDeviceArray:
    """Lets torch wrap a raw device pointer without copying (``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, numel: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False), "version": 2,
                                         "strides": None}


def _wrap(ptr: int, numel: int, dtype: torch.dtype, device) -> torch.Tensor:
    if dtype == torch.bfloat16:       # no numpy typestr for bf16: wrap as int16 and reinterpret
        t = torch.as_tensor(_RawDeviceArray(ptr, numel, "<i2"), device=device)
        return t.view(torch.bfloat16)
    return torch.as_tensor(_RawDeviceArray(ptr, numel, _TYPESTR[dtype]), device=device)


class PeerBuffers:
    """``names`` flat buffers of ``numel`` elements per rank, each addressable from every rank.

    ``local[name]`` is this rank's buffer as a torch tensor (gradient views are carved out of it);
    ``ptrs(name)[r]`` is rank r's buffer as a device address valid in THIS process.
    Call :meth:`close` on every rank before the process group goes away."""

    def __init__(self, numel: int, dtype: torch.dtype, names=("grads", "reduced"), group=None, device=None,
                 multicast=()):
        if not dist.is_initialized():
            raise RuntimeError("PeerBuffers needs an initialised process group")
        lib = _lib.load()
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world > _lib.MAX_PEERS:
            raise RuntimeError(f"the peer-memory path supports up to {_lib.MAX_PEERS} ranks (one NVSwitch domain)")
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.numel, self.dtype = numel, dtype
        self.names = tuple(names)
        esize = torch.empty(0, dtype=dtype).element_size()
        nbytes = (numel * esize + 255) // 256 * 256
        self._own, self._opened = {}, []
        handles = {}
        # buffers named in `multicast` come from torch's symmetric memory, which also binds them to an NVSwitch
        # multicast object (set-up only: no torch collective runs in the step); mc_ptr(name) is then the address
        # through which ONE store reaches every rank's buffer.  Anything missing -> plain IPC buffers, mc_ptr = None.
        self._symm, self._mc, self.multicast_note = {}, {}, None
        for n in multicast:
            if n not in self.names:
                raise RuntimeError(f"multicast buffer {n!r} is not one of {self.names}")
            try:
                self._symm[n] = self._symm_alloc(nbytes // esize, dtype, group)
            except Exception as exc:                      # noqa: BLE001 — every rank must take the same branch below
                self._symm[n] = None
                self.multicast_note = f"{type(exc).__name__}: {exc}"[:200]
        if multicast:
            ok = torch.tensor([int(all(self._symm[n] is not None and self._symm[n][1].multicast_ptr for n in multicast))],
                              device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if not int(ok.item()):
                if self.multicast_note is None:
                    self.multicast_note = "no multicast address (has_multicast_support is false on some rank)"
                self._symm = {}
        with torch.cuda.device(self.device):
            torch.cuda.current_stream().synchronize()
            for n in self.names:
                if n in self._symm:
                    continue
                ptr = ctypes.c_void_p()
                h = (ctypes.c_ubyte * 64)()
                _lib.check(lib.b200_peer_alloc(nbytes, ctypes.byref(ptr), h), "peer_alloc")
                self._own[n] = ptr.value
                handles[n] = bytes(h)
            gathered = [None] * self.world
            dist.all_gather_object(gathered, handles, group=group)
            self._ptrs = {}
            for n in self.names:
                if n in self._symm:
                    t, hdl = self._symm[n]
                    self._ptrs[n] = [int(x) for x in hdl.buffer_ptrs]
                    self._mc[n] = int(hdl.multicast_ptr)
                    continue
                row = []
                for r in range(self.world):
                    if r == self.rank:
                        row.append(self._own[n])
                        continue
                    ptr = ctypes.c_void_p()
                    hb = (ctypes.c_ubyte * 64).from_buffer_copy(gathered[r][n])
                    _lib.check(lib.b200_peer_open(hb, ctypes.byref(ptr)), "peer_open")
                    self._opened.append(ptr.value)
                    row.append(ptr.value)
                self._ptrs[n] = row
        self.local = {n: (self._symm[n][0][:numel] if n in self._symm else _wrap(self._own[n], numel, dtype, self.device))
                      for n in self.names}
        self.signal_ptrs = None
        self._bar = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._closed = False
        dist.barrier(group=group)

    def _symm_alloc(self, numel, dtype, group):
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(numel, dtype=dtype, device=self.device)
        hdl = symm.rendezvous(t, group if group is not None else dist.group.WORLD)
        t.zero_()
        return t, hdl

    def ptrs(self, name):
        return list(self._ptrs[name])

    def mc_ptr(self, name):
        """Multicast address of buffer ``name`` (None unless it was requested and the fabric provides one)."""
        return self._mc.get(name) or None

    def signal_pads(self, count: int = 1, index: int = 0):
        """Allocate (once) ``count`` 128-byte signal pads per rank and return pad ``index`` of every rank
        as addresses valid in this process (b200_peer_signal_pad, zero-initialised by the allocator)."""
        if self.signal_ptrs is None:
            self._n_pads = count
            self._real_pads = count          # how many of them are signal pads (the rest may be a progress board)
            lib = _lib.load()
            with torch.cuda.device(self.device):
                ptr = ctypes.c_void_p()
                h = (ctypes.c_ubyte * 64)()
                _lib.check(lib.b200_peer_alloc(max(256, 128 * count), ctypes.byref(ptr), h), "peer_alloc")
                self._own["__signal__"] = ptr.value
                gathered = [None] * self.world
                dist.all_gather_object(gathered, bytes(h), group=self.group)
                row = []
                for r in range(self.world):
                    if r == self.rank:
                        row.append(ptr.value)
                        continue
                    q = ctypes.c_void_p()
                    hb = (ctypes.c_ubyte * 64).from_buffer_copy(gathered[r])
                    _lib.check(lib.b200_peer_open(hb, ctypes.byref(q)), "peer_open")
                    self._opened.append(q.value)
                    row.append(q.value)
                self.signal_ptrs = row
                self._signal_view = _wrap(ptr.value, 32 * count, torch.int32, self.device)
            dist.barrier(group=self.group)
        if index >= self._n_pads:
            raise RuntimeError("signal pad index out of range")
        return [p + 128 * index for p in self.signal_ptrs]

    def check_signals(self):
        """Raise if a bounded in-kernel wait gave up (synchronises)."""
        if self.signal_ptrs is not None and bool(self._signal_view.view(-1, 32)[:self._real_pads, 16].any().item()):
            raise RuntimeError("libapex_b200: a cross-GPU wait timed out (a rank did not reach the step)")

    def barrier(self, flag: torch.Tensor | None = None):
        """Stream-ordered cross-rank barrier.  With ``flag`` (int32) it also ORs the ranks' overflow
        flags in place (all-reduce MAX of one word) — 4 bytes over NCCL, no host sync."""
        t = flag[:1] if flag is not None else self._bar
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)

    def close(self):
        """Unmap the peers' buffers, wait for everybody, then free our own.  The tensors in
        ``local`` (and every view of them) must not be used afterwards."""
        if self._closed:
            return
        self._closed = True
        lib = _lib.load()
        with torch.cuda.device(self.device):
            torch.cuda.synchronize()
            timed_out = self.signal_ptrs is not None and bool(self._signal_view.view(-1, 32)[:self._real_pads, 16].any().item())
            for p in self._opened:
                lib.b200_peer_close(p)
            self._opened = []
            try:
                dist.barrier(group=self.group)
            except Exception:
                pass
            self.local = {}
            self._symm, self._mc = {}, {}
            self._signal_view = None
            for p in self._own.values():
                lib.b200_peer_free(p)
            self._own = {}
        if timed_out:
            raise RuntimeError("libapex_b200: a cross-GPU wait timed out during this run (a rank did not reach a step "
                               "within ~30 s); results after that point are not valid")


class PeerGradientReducer:
    """reduce-scatter + mean + check (one kernel) followed by a flag-OR barrier; hands the optimizer a
    b200_peer_pull so its step kernel gathers the reduced shards on the fly."""

    def __init__(self, buffers: PeerBuffers, bounds, gradient_average: bool = True, first_tiles=None,
                 barrier: str = "signal"):
        self.buf = buffers
        self.barrier = barrier
        self.epoch = 0
        self._pads = buffers.signal_pads() if barrier == "signal" else None
        self.bounds = list(bounds)
        assert len(self.bounds) == buffers.world + 1
        self.factor = (1.0 / buffers.world) if gradient_average else 1.0
        self._grad_ptrs = buffers.ptrs("grads")
        self._red_ptrs = buffers.ptrs("reduced")
        self.pull = ops.make_peer_pull(buffers.local["grads"].data_ptr(), self.bounds, self._red_ptrs,
                                       buffers.local["grads"].element_size(),
                                       first_tile=first_tiles[buffers.rank] if first_tiles else 0)

    def reduce(self, overflow_flag: torch.Tensor):
        """Enqueue on the current stream: barrier -> shard reduction -> barrier + flag OR.

        barrier="signal": both hand-shakes happen inside the kernels through signal pads in peer memory
        (no NCCL call in the step); barrier="nccl": two 4-byte NCCL all-reduces."""
        b = self.buf
        if self.barrier == "signal":
            self.epoch += 1
            sync = ops.make_peer_sync(self._pads, b.rank, self.epoch)
            ops.peer_reduce_shard(self._grad_ptrs, b.local["reduced"], self.bounds[b.rank], self.bounds[b.rank + 1],
                                  self.factor, None, sync)
            ops.peer_wait_flags(sync, overflow_flag)
            return
        b.barrier()                                     # every rank's gradients are in its `grads` buffer,
        #                                                 and every rank is done reading last step's `reduced`
        ops.peer_reduce_shard(self._grad_ptrs, b.local["reduced"], self.bounds[b.rank], self.bounds[b.rank + 1],
                              self.factor, overflow_flag)
        b.barrier(overflow_flag)                        # all shards written; flags OR-ed across ranks


class PeerSpecStep:
    """Gradient mean + speculative Adam step as ONE kernel per rank (b200_peer_spec_step): every CTA keeps one
    reduce piece of this rank's shard in flight (cp.async into shared memory) while it updates a tile, and a
    tile is updated as soon as the owner of its gradient has announced the chunk it lies in — NVLink time and
    HBM time overlap instead of adding up.  ``table`` must list its gradient column as views of
    ``buffers.local["grads"]``; it is re-ordered in place.  Must be the first user of the buffers' signal pads
    (the progress board lives right behind the pad)."""

    def __init__(self, buffers: PeerBuffers, bounds, table, gradient_average: bool = True, chunk_elems: int | None = None,
                 n_reducers: int | None = None):
        self.buf = buffers
        self.bounds = list(bounds)
        assert len(self.bounds) == buffers.world + 1
        self.factor = (1.0 / buffers.world) if gradient_average else 1.0
        shard = max(b - a for a, b in zip(self.bounds, self.bounds[1:]))
        if chunk_elems is None:
            chunk_elems = 1 << 19
        min_chunk = -(-shard // _lib.PEER_STEP_MAX_CHUNKS)
        chunk_elems = max(int(chunk_elems), min_chunk)
        self.chunk_elems = -(-chunk_elems // 131072) * 131072
        # measured on B200 (GPT-2-small list): 32 reducer CTAs are best at 2 ranks, 16 at 8 (profiles/r02_summary.md)
        self.n_reducers = int(n_reducers) if n_reducers else (32 if buffers.world <= 2 else 24 if buffers.world <= 4 else 16)
        self.epoch = 0
        if buffers.signal_ptrs is not None and buffers._n_pads < _lib.PEER_STEP_PAD_BYTES // 128:
            raise RuntimeError("PeerSpecStep must allocate the signal pads itself (pad + progress board)")
        self._pads = buffers.signal_pads(count=_lib.PEER_STEP_PAD_BYTES // 128)
        buffers._real_pads = 1               # one signal pad; the other 128-byte units are the progress board
        self._grad_ptrs = buffers.ptrs("grads")
        g = buffers.local["grads"]
        self.pull = ops.make_peer_pull(g.data_ptr(), self.bounds, buffers.ptrs("reduced"), g.element_size())
        self.table = table.order_for_peer_step(g.data_ptr(), g.element_size(), self.bounds, buffers.rank, self.chunk_elems)
        self.scratch = ops.new_peer_step_scratch(buffers.device)

    def step(self, grad_dtype, state_dtype, copy_dtype, groups, dev_flag, dev_parity, dev_inv_scale=None, inv_scale=1.0,
             sync=None):
        b = self.buf
        self.epoch += 1
        psync = ops.make_peer_sync(self._pads, b.rank, self.epoch)
        ops.peer_spec_step(self.table, grad_dtype, state_dtype, copy_dtype, groups, dev_flag, dev_parity, self.pull,
                           self._grad_ptrs, b.local["reduced"], self.factor, psync, self.n_reducers, self.chunk_elems,
                           self.scratch, inv_scale=inv_scale, dev_inv_scale=dev_inv_scale, sync=sync)


def even_shards(total: int, world: int, align: int = 1024):
    """world+1 element offsets splitting [0, total) evenly at multiples of ``align`` (pure host code)."""
    bounds = [0]
    for r in range(1, world):
        b = min(total, (r * total // world + align - 1) // align * align)
        bounds.append(max(b, bounds[-1]))
    bounds.append(total)
    return bounds


def shard_segments(offsets, sizes, begin: int, end: int):
    """The pieces of the tensors (tensor t at flat offset offsets[t], sizes[t] long) that fall inside [begin, end):
    a list of (tensor index, first element inside the tensor, length, flat offset).  Pure host code."""
    out = []
    for t, (o, n) in enumerate(zip(offsets, sizes)):
        a, b = max(o, begin), min(o + n, end)
        if a < b:
            out.append((t, a - o, b - a, a))
    return out


class PeerShardedStep:
    """ZeRO-1 style data-parallel step as ONE kernel per rank (b200_peer_shard_step): each rank keeps the fp32
    master / m / v of its shard only, averages every rank's gradients for that shard in rank order, runs the
    speculative Adam update and stores the new 16-bit weights into EVERY rank's model buffer — gradient reads
    (NVLink inbound) and weight writes (outbound) overlap, and no tile waits for another.  Followed here by the
    flag exchange, the commit (loss-scaler update + parity flip) and the repair that re-sends the old weights
    after an overflow.

    ``buffers`` must have been built with names containing "grads" and "model" (same 16-bit dtype): every rank
    produces its gradients into ``buffers.local["grads"]`` and reads its weights from ``buffers.local["model"]``
    (tensor t at element offset offsets[t]).  ``masters``: optional fp32 flat tensor (the whole model) to take
    the shard's initial master weights from; default = the local model buffer widened to fp32.
    ``use_multicast``: "none" | "stores" | "reduce" | "both" — which of the two NVSwitch multicast paths to use when
    the buffers have a multicast address (PeerBuffers(multicast=("grads", "model"))): "stores" sends every 16 bytes
    of new weights out once (multimem.st), "reduce" sums the gradients inside the switch (multimem.ld_reduce: fp32
    accumulation, addend order and partial-sum precision unspecified -> the mean may differ from the rank-ordered one by a few
    16-bit ulps)."""

    def __init__(self, buffers: PeerBuffers, offsets, sizes, group_ids=None, gradient_average: bool = True,
                 bounds=None, masters: torch.Tensor | None = None, tile_elems: int | None = None, use_multicast="both"):
        from ..tables import TensorTable
        if "grads" not in buffers.names or "model" not in buffers.names:
            raise RuntimeError('PeerShardedStep needs PeerBuffers(names=("grads", "model"))')
        self.buf = buffers
        self.bounds = list(bounds) if bounds is not None else even_shards(buffers.numel, buffers.world)
        assert len(self.bounds) == buffers.world + 1
        self.begin, self.end = self.bounds[buffers.rank], self.bounds[buffers.rank + 1]
        if self.end <= self.begin:
            raise RuntimeError("PeerShardedStep: this rank's shard is empty (fewer elements than ranks x 1024?)")
        self.segments = shard_segments(offsets, sizes, self.begin, self.end)
        g, w = buffers.local["grads"], buffers.local["model"]
        src = masters if masters is not None else w
        self.state = {"p": [], "m": [], "v": []}
        gcol, wcol, grp = [], [], []
        for t, first, n, flat in self.segments:
            gcol.append(g[flat:flat + n])
            wcol.append(w[flat:flat + n])
            init = src[flat:flat + n].to(device=buffers.device, dtype=torch.float32)
            self.state["p"].append(ops.spec_state_like(init, fill=init))
            zero = torch.zeros(n, dtype=torch.float32, device=buffers.device)
            self.state["m"].append(ops.spec_state_like(zero))
            self.state["v"].append(ops.spec_state_like(zero))
            grp.append(int(group_ids[t]) if group_ids is not None else 0)
        self.table = TensorTable([gcol] + [[a for a, _ in self.state[k]] for k in "pmv"] + [wcol], groups=grp,
                                 tile_elems=tile_elems)
        self.factor = (1.0 / buffers.world) if gradient_average else 1.0
        want = {"none": (False, False), "stores": (True, False), "reduce": (False, True), "both": (True, True),
                True: (True, False), False: (False, False)}[use_multicast]
        self.multicast_stores = bool(buffers.mc_ptr("model")) and want[0]
        self.multicast_reduce = bool(buffers.mc_ptr("grads")) and want[1]
        self.multicast = self.multicast_stores
        self.fan = ops.make_peer_fan(buffers.ptrs("grads"), buffers.ptrs("model"), buffers.rank, self.factor,
                                     model_mc=buffers.mc_ptr("model") if self.multicast_stores else None,
                                     grads_mc=buffers.mc_ptr("grads") if self.multicast_reduce else None)
        self._pads = buffers.signal_pads()
        self.parity = torch.zeros(1, dtype=torch.int32, device=buffers.device)
        self.epoch = 0
        self.dtype = buffers.dtype

    def current(self, name: str, parity: int):
        """The current copies of the shard's state ('p', 'm' or 'v'), one tensor per segment."""
        return [ops.spec_current(a, b, parity) for a, b in self.state[name]]

    def step(self, groups, scaler_state: torch.Tensor, dynamic: bool = True, scale_window: int = 2000,
             scale_factor: float = 2.0, min_scale: float = 0.0, max_scale: float = 2.0 ** 24, sync=None,
             ev_after_kernel=None):
        """One data-parallel step.  ``scaler_state``: the device-side b200_scaler_state words (its inverse scale
        unscales the gradients, its overflow word collects every rank's flag, the commit updates it)."""
        b = self.buf
        self.epoch += 2
        fs = scaler_state.view(torch.float32)
        psync = ops.make_peer_sync(self._pads, b.rank, self.epoch)
        ops.peer_shard_step(self.table, self.dtype, torch.float32, self.dtype, groups, self.parity, self.fan, psync,
                            dev_inv_scale=fs[ops.IDX_INV_LOSS_SCALE:], sync=sync)
        if ev_after_kernel is not None:
            ev_after_kernel.record()
        ops.peer_wait_flags(psync, scaler_state[ops.IDX_OVERFLOW_ACC:])
        ops.spec_commit(scaler_state, dynamic, scale_window, scale_factor, min_scale, max_scale, self.parity, sync=sync)
        ops.peer_shard_repair(self.table, torch.float32, self.dtype, scaler_state[ops.IDX_FOUND_INF:], self.parity, self.fan,
                              ops.make_peer_sync(self._pads, b.rank, self.epoch + 1))
