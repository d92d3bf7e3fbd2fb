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

    def __init__(self, numel: int, dtype: torch.dtype, names=("grads", "reduced"), group=None, device=None):
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
        with torch.cuda.device(self.device):
            torch.cuda.current_stream().synchronize()
            for n in self.names:
                ptr = ctypes.c_void_p()
                h = (ctypes.c_ubyte * 64)()
                _lib.check(lib.b200_peer_alloc(nbytes, ctypes.byref(ptr), h), "peer_alloc")
                self._own[n] = ptr.value
                handles[n] = bytes(h)
            gathered = [None] * self.world
            dist.all_gather_object(gathered, handles, group=group)
            self._ptrs = {}
            for n in self.names:
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
        self.local = {n: _wrap(self._own[n], numel, dtype, self.device) for n in self.names}
        self.signal_ptrs = None
        self._bar = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._closed = False
        dist.barrier(group=group)

    def ptrs(self, name):
        return list(self._ptrs[name])

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
