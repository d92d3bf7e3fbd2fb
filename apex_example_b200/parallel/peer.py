"""Peer-addressable gradient buffers for the fused reduce-scatter / optimizer path (S3').

Plumbing only: every rank allocates its flat buffers with torch, exports them with CUDA IPC
(torch's own storage-sharing mechanism), gathers the handles over the process group and opens
its peers' buffers into its own address space with peer access enabled.  The arithmetic is in
libapex_b200 (b200_peer_reduce_shard, b200_fused_*_peer); NCCL is used for two 4-byte
all-reduces per step that act as cross-rank barriers and OR the overflow flags.

Single node only (NVLink / NVSwitch): every pair of GPUs must be peer-accessible.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .. import ops


class PeerBuffers:
    """``names`` flat buffers of ``numel`` elements per rank, each addressable from every rank."""

    def __init__(self, numel: int, dtype: torch.dtype, names=("grads", "reduced"), group=None, device=None):
        if not dist.is_initialized():
            raise RuntimeError("PeerBuffers needs an initialised process group")
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.numel, self.dtype = numel, dtype
        self.local = {n: torch.zeros(numel, dtype=dtype, device=self.device) for n in names}
        self.peers = {}
        self._keep = []
        exported = {}
        for n, t in self.local.items():
            st = t.untyped_storage()
            exported[n] = (st._share_cuda_(), t.storage_offset() * t.element_size())
        gathered = [None] * self.world
        dist.all_gather_object(gathered, exported, group=group)
        for n in names:
            views = []
            for r in range(self.world):
                if r == self.rank:
                    views.append(self.local[n])
                    continue
                handle, extra_off = gathered[r][n]
                # open in THIS device's address space: the storage then lives on the peer GPU but is
                # addressable from kernels running here (peer access is enabled lazily by the runtime)
                args = (self.device.index,) + tuple(handle[1:])
                storage = torch.UntypedStorage._new_shared_cuda(*args)
                t = torch.empty(0, dtype=dtype, device=self.device).set_(storage, extra_off // self.local[n].element_size(),
                                                                          (numel,))
                self._keep.append(storage)
                views.append(t)
            self.peers[n] = views
        # tiny tensors for the two per-step barriers
        self._bar = torch.zeros(1, dtype=torch.int32, device=self.device)
        dist.barrier(group=group)

    def close(self):
        """Drop the mappings of the peers' memory (call on every rank before the process group is
        destroyed; the owners' buffers must outlive everybody's mappings, hence the barrier)."""
        if self.peers is None:
            return
        torch.cuda.synchronize(self.device)
        try:
            dist.barrier(group=self.group)
        except Exception:
            pass
        self.peers = None
        self._keep.clear()
        torch.cuda.synchronize(self.device)
        try:
            dist.barrier(group=self.group)
        except Exception:
            pass

    def ptrs(self, name):
        return [t.data_ptr() for t in self.peers[name]]

    def barrier(self, flag: torch.Tensor | None = None):
        """Stream-ordered cross-rank barrier.  With ``flag`` (int32[1]) it also ORs the ranks' overflow
        flags in place (all-reduce MAX) — 4 bytes over NCCL, no host sync."""
        t = flag if flag is not None else self._bar
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)


class PeerGradientReducer:
    """reduce-scatter + mean + check (one kernel) followed by a flag-OR barrier; hands the optimizer a
    b200_peer_pull so its step kernel gathers the reduced shards on the fly."""

    def __init__(self, buffers: PeerBuffers, bounds, gradient_average: bool = True, predivide: float = 1.0):
        self.buf = buffers
        self.bounds = list(bounds)
        assert len(self.bounds) == buffers.world + 1
        self.factor = (1.0 / buffers.world) if gradient_average else 1.0
        self._grad_ptrs = buffers.ptrs("grads")
        self._red_ptrs = buffers.ptrs("reduced")
        self.pull = ops.make_peer_pull(buffers.local["grads"].data_ptr(), self.bounds, self._red_ptrs,
                                       buffers.local["grads"].element_size())

    def reduce(self, overflow_flag: torch.Tensor):
        """Enqueue on the current stream: barrier -> shard reduction -> barrier + flag OR."""
        b = self.buf
        b.barrier()                                     # every rank's gradients are in its `grads` buffer,
        #                                                 and every rank is done reading last step's `reduced`
        ops.peer_reduce_shard(self._grad_ptrs, b.local["reduced"], self.bounds[b.rank], self.bounds[b.rank + 1],
                              self.factor, overflow_flag)
        b.barrier(overflow_flag)                        # all shards written; flags OR-ed across ranks
