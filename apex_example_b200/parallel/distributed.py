"""apex.parallel.DistributedDataParallel, B200-first.

The reference wraps its model at /root/reference/test_apex_distributed_spawn.py:122 and drives
this code from ``scaled_loss.backward()`` (:161).  apex's version (apex/parallel/distributed.py,
SURVEY.md App. B.7) learns bucket order in iteration 1, then per bucket: ``apex_C.flatten``
(copy), ``all_reduce``, ``mul_(1/world)``, ``unflatten`` + copy back (another copy).

Here:
  * buckets are PERSISTENT flat buffers laid out once (reverse parameter order ~ backward arrival
    order, one dtype per bucket, closed at ``message_size`` elements); rank 0's layout is broadcast
    and every rank checks it matches (apex's bucket-structure broadcast, C2);
  * when the last gradient of a bucket arrives, ONE multi-tensor launch gathers the bucket's
    gradients into the buffer on a side stream, NCCL all-reduces it in place over NVLink, and ONE
    epilogue launch applies the averaging factor and raises the loss scaler's overflow flag
    (b200_bucket_epilogue) — so amp does not need its own pass over the gradients;
  * ``param.grad`` is then re-pointed at the bucket view: no unflatten copy, and the fused
    optimizer reads gradients straight from the bucket.
The main stream waits for the side streams at the end of backward (engine callback), exactly where
apex does.
"""
from __future__ import annotations

import operator

import torch
import torch.distributed as dist
from torch.autograd import Variable
from torch.nn.modules import Module

from .. import ops
from ..tables import TensorTable, flat_view_like, layout_key

_SUPPORTED = (torch.float16, torch.bfloat16, torch.float32)
_GRAD_OF = operator.attrgetter("grad")


def split_by_dtype(tensors):
    """apex's split_half_float_double, plus bfloat16: one list per dtype, in first-seen order."""
    buckets = {}
    for t in tensors:
        buckets.setdefault(t.dtype, []).append(t)
    return list(buckets.values())


def flat_dist_call(tensors, call, extra_args=None):
    """Run a collective on dtype-grouped, flattened copies of ``tensors`` and copy the result back
    (apex flat_dist_call; used once, for the rank-0 parameter broadcast at construction)."""
    from torch._utils import _flatten_dense_tensors, _unflatten_dense_tensors
    for group in split_by_dtype(tensors):
        flat = _flatten_dense_tensors(group)
        if extra_args is not None:
            call(flat, *extra_args)
        else:
            call(flat)
        for buf, synced in zip(group, _unflatten_dense_tensors(flat, group)):
            buf.copy_(synced)


class _CudaBucketOps:
    """Device work of one bucket, on libapex_b200 kernels."""

    def gather(self, srcs, dsts, scale, table_cache, key):
        sig = (tuple(t.data_ptr() for t in srcs), tuple(t.data_ptr() for t in dsts))
        tbl = table_cache.get(key)
        if tbl is None or tbl.signature != sig:
            tbl = TensorTable([srcs, dsts], signature=sig)
            table_cache[key] = tbl
        ops.multi_tensor_scale(tbl, srcs[0].dtype, dsts[0].dtype, scale=scale)

    def epilogue(self, flat, factor, flag):
        ops.bucket_epilogue(flat, factor, flag)


class _Bucket:
    __slots__ = ("index", "dtype", "params", "offsets", "flat", "flat32", "views", "ready", "launched", "stream",
                 "ready_event", "done_event", "numel", "reduced", "grad_ptrs", "red_ptrs", "pads", "epoch", "bounds")

    def __init__(self, index, dtype):
        self.index = index
        self.dtype = dtype
        self.params = []
        self.offsets = []
        self.flat = None
        self.flat32 = None
        self.views = []
        self.ready = 0
        self.launched = False
        self.stream = None
        self.ready_event = None
        self.done_event = None


class DistributedDataParallel(Module):
    """Data-parallel wrapper with apex's constructor.

    Args (as apex): ``message_size`` elements per bucket; ``delay_allreduce`` reduce everything at
    the end of backward; ``allreduce_trigger_params`` close a bucket at each of these parameters;
    ``retain_allreduce_buffers`` expose the flat buffers as ``self.allreduce_buffers``;
    ``allreduce_always_fp32`` reduce half buckets in fp32; ``num_allreduce_streams`` side streams;
    ``allreduce_communicators`` = (list of process groups, list of streams);
    ``gradient_average`` / ``gradient_predivide_factor`` the averaging rule
    ``g = sum(g_r / predivide) * predivide / world``; ``prof`` NVTX ranges.
    """

    def __init__(self, module, message_size=10000000, delay_allreduce=False, shared_param=None,
                 allreduce_trigger_params=None, retain_allreduce_buffers=False, allreduce_always_fp32=False,
                 num_allreduce_streams=1, allreduce_communicators=None, gradient_average=True,
                 gradient_predivide_factor=1.0, gradient_average_split_factor=None, prof=False,
                 allreduce_backend=None, _bucket_ops=None):
        super().__init__()
        if not dist.is_available() or not dist.is_initialized():
            raise RuntimeError("apex.parallel.DistributedDataParallel requires torch.distributed to be initialized "
                               "(call torch.distributed.init_process_group first).")
        if shared_param is not None:
            raise ValueError("shared_param is no longer supported as an option.  It was misleadingly named from "
                             "the start.  It turns out overlapping communication with computation should work "
                             "fine with shared parameters.  If you still wish to delay communication to the end "
                             "of the backward pass, use delay_allreduce=True|False instead.")
        if gradient_average_split_factor is not None:
            print("Warning:  gradient_average_split_factor has been renamed to gradient_predivide_factor.  "
                  "For now, gradient_average_split_factor will also work, but please update to "
                  "gradient_predivide_factor instead.")
            gradient_predivide_factor = gradient_average_split_factor

        self.world_size = float(dist.get_world_size())
        self.module = module
        self.message_size = int(message_size)
        self.delay_allreduce = bool(delay_allreduce)
        self.retain_allreduce_buffers = retain_allreduce_buffers
        self.allreduce_always_fp32 = allreduce_always_fp32
        self.gradient_average = gradient_average
        self.gradient_predivide_factor = float(gradient_predivide_factor)
        self.prof = prof
        self.custom_allreduce_triggers = allreduce_trigger_params is not None
        self._trigger_ids = set(id(p) for p in allreduce_trigger_params) if allreduce_trigger_params else set()
        if self.custom_allreduce_triggers and self.delay_allreduce:
            raise ValueError("Setting allreduce_trigger_params is only valid if delay_allreduce=False.")
        self._ops = _bucket_ops
        # "nccl" (default, as apex) or "peer": the bucket all-reduce runs as two kernels of this library
        # over NVLink peer memory (reduce-scatter + mean + overflow check, then gather), no NCCL call
        import os
        self.allreduce_backend = (allreduce_backend or os.environ.get("APEX_B200_DDP_BACKEND") or "nccl").lower()
        if self.allreduce_backend not in ("nccl", "peer"):
            raise ValueError("allreduce_backend must be 'nccl' or 'peer'")
        self._peer = {}
        self._enabled = True
        self._armed = False
        self._callback_queued = False
        self._table_cache = {}
        self.allreduce_buffers = []

        params = [p for p in module.parameters() if p.requires_grad]
        self._on_cuda = bool(params) and params[0].is_cuda
        for p in params:
            if p.dtype not in _SUPPORTED:
                raise TypeError("apex.parallel.DistributedDataParallel only supports fp16, bf16 and fp32 "
                                "parameters, found {}".format(p.type()))
            if not p.is_cuda and self._ops is None:
                raise TypeError("apex.parallel.DistributedDataParallel only supports CUDA parameters "
                                "(found {}); its bucket kernels have no CPU fallback.".format(p.type()))
        if self._ops is None:
            self._ops = _CudaBucketOps()

        if allreduce_communicators:
            self.bucket_pgs, self.bucket_streams = allreduce_communicators
            self.num_allreduce_streams = len(self.bucket_streams)
            assert len(self.bucket_pgs) == self.num_allreduce_streams
        else:
            self.num_allreduce_streams = max(1, int(num_allreduce_streams))
            self.bucket_pgs = [None] * self.num_allreduce_streams
            self.bucket_streams = ([torch.cuda.Stream() for _ in range(self.num_allreduce_streams)]
                                   if self._on_cuda else [None] * self.num_allreduce_streams)

        # C1: every rank starts from rank 0's parameters (and buffers)
        tensors = [p.data for p in module.parameters()] + [b.data for b in module.buffers()
                                                            if b.is_floating_point()]
        if tensors:
            flat_dist_call(tensors, dist.broadcast, (0,))
            self._resync_amp_masters()

        self._build_buckets(params)
        self._sync_bucket_structure()
        self._next = 0
        if self.world_size > 1 or not self._on_cuda:
            for p in params:
                p.register_post_accumulate_grad_hook(self._hook)
        try:
            from ..amp._amp_state import _amp_state
            _amp_state.ddp_wrappers.add(self)
        except Exception:
            pass

    def zero_bucket_grads(self, params):
        """amp's patched ``zero_grad`` (a9): every bucket whose parameters' gradients currently ARE its
        views is cleared with one memset of the flat buffer; returns the parameters not covered."""
        if not self._on_cuda:
            return params
        handled = set()
        for b in self.buckets:
            if b.flat is None or not b.views:
                continue
            if all(map(operator.is_, map(_GRAD_OF, b.params), b.views)):
                b.flat.zero_()
                handled.update(map(id, b.params))
        if not handled:
            return params
        return [p for p in params if id(p) not in handled]

    def _resync_amp_masters(self):
        """amp.initialize runs BEFORE this wrapper (reference :121 then :122), so at O2 the fp32
        masters were cloned from each rank's own initial weights.  apex leaves them as they are
        (ranks must seed identically); here the masters of the broadcast parameters are re-derived
        from rank 0's weights so the replicas cannot start apart."""
        try:
            from ..amp._amp_state import _amp_state
        except Exception:
            return
        mine = set(id(p) for p in self.module.parameters())
        with torch.no_grad():
            for opt in getattr(_amp_state, "optimizers", []):
                stash = getattr(opt, "_amp_stash", None)
                if stash is None:
                    continue
                for master, model_p in stash.model_of.items():
                    if id(model_p) in mine:
                        master.copy_(model_p.float())

    # ---- layout --------------------------------------------------------------------------------------
    def _build_buckets(self, params):
        self.buckets = []
        self._bucket_of = {}
        open_by_dtype = {}
        for p in reversed(params):
            b = open_by_dtype.get(p.dtype)
            if b is None:
                b = _Bucket(len(self.buckets), p.dtype)
                self.buckets.append(b)
                open_by_dtype[p.dtype] = b
            off = sum(q.numel() for q in b.params)
            # keep every view 32-byte aligned so the gather kernel takes its 128/256-bit path
            pad = (-off * p.element_size()) % 32 // p.element_size()
            b.offsets.append(off + pad)
            b.params.append(p)
            self._bucket_of[id(p)] = b
            b_numel = b.offsets[-1] + p.numel()
            close = id(p) in self._trigger_ids if self.custom_allreduce_triggers else b_numel >= self.message_size
            if close:
                del open_by_dtype[p.dtype]
        for b in self.buckets:
            # offsets were computed incrementally without earlier padding; recompute exactly
            off = 0
            b.offsets = []
            for p in b.params:
                off += (-off * p.element_size()) % 32 // p.element_size()
                b.offsets.append(off)
                off += p.numel()
            device = b.params[0].device
            b.numel = off
            b.flat = None if self._use_peer() else torch.zeros(off, dtype=b.dtype, device=device)
            if b.flat is not None:
                b.views = [flat_view_like(b.flat, o, p) for o, p in zip(b.offsets, b.params)]
            if self.allreduce_always_fp32 and b.dtype != torch.float32 and not self._use_peer():
                b.flat32 = torch.zeros(off, dtype=torch.float32, device=device)
            slot = b.index % self.num_allreduce_streams
            b.stream = self.bucket_streams[slot]
            if self._on_cuda:
                b.ready_event = torch.cuda.Event()
                b.done_event = torch.cuda.Event()
        if self._use_peer():
            self._build_peer_buckets()
            # the peer kernels overlap with backward and spin while slower ranks catch up: keep them a
            # small "communication" grid (process-wide knob of the library)
            import os
            ops.set_peer_grid(int(os.environ.get("APEX_B200_PEER_CTAS", "128")))   # measured: 8 -> 45.4, 32 -> 40.6, 128 -> 39.5 ms/step
        if self.retain_allreduce_buffers:
            self.allreduce_buffers = [b.flat for b in self.buckets]
        self._checked_ids = frozenset(self._bucket_of)

    def _use_peer(self):
        return self.allreduce_backend == "peer" and self._on_cuda and self.world_size > 1

    def _build_peer_buckets(self):
        """Peer backend: all buckets of one dtype are slices of ONE peer-addressable buffer pair (gradients,
        reduced shards); every bucket has its own signal pad and its own epoch counter."""
        from .peer import PeerBuffers
        world = int(self.world_size)
        by_dtype = {}
        for b in self.buckets:
            by_dtype.setdefault(b.dtype, []).append(b)
        for dtype, group in by_dtype.items():
            starts, total = [], 0
            for b in group:
                total += (-total) % 64                         # 128-byte aligned bucket starts
                starts.append(total)
                total += b.numel
            bufs = PeerBuffers(max(total, 64), dtype, names=("grads", "reduced"))
            bufs.signal_pads(count=len(group))
            self._peer[dtype] = bufs
            esize = bufs.local["grads"].element_size()
            for i, (b, st) in enumerate(zip(group, starts)):
                b.flat = bufs.local["grads"][st:st + b.numel]
                b.reduced = bufs.local["reduced"][st:st + b.numel]
                b.views = [flat_view_like(b.flat, o, p) for o, p in zip(b.offsets, b.params)]
                b.grad_ptrs = [ptr + st * esize for ptr in bufs.ptrs("grads")]
                b.red_ptrs = [ptr + st * esize for ptr in bufs.ptrs("reduced")]
                b.pads = bufs.signal_pads(index=i)
                b.epoch = 0
                shard = -(-b.numel // world)
                shard += (-shard) % 8
                b.bounds = [min(r * shard, b.numel) for r in range(world)] + [b.numel]
        self._rank = dist.get_rank()

    def close(self):
        """Release the peer-addressable buffers (peer backend; call on every rank before the process
        group is destroyed).  Gradients that are views of the buckets must not be used afterwards."""
        for p in self.module.parameters():
            p.grad = None
        for b in self.buckets:
            b.flat = b.reduced = None
            b.views = []
        for bufs in self._peer.values():
            bufs.close()
        self._peer = {}

    def _sync_bucket_structure(self):
        """C2: rank 0 broadcasts [num_buckets] + bucket sizes + parameter sizes; all ranks must agree."""
        info = [len(self.buckets)] + [b.flat.numel() for b in self.buckets] + \
               [p.numel() for b in self.buckets for p in b.params]
        device = self.buckets[0].flat.device if self.buckets else torch.device("cpu")
        mine = torch.tensor(info, dtype=torch.int64, device=device)
        ref = mine.clone()
        dist.broadcast(ref, 0)
        if ref.numel() != mine.numel() or not torch.equal(ref, mine):
            raise RuntimeError("apex.parallel.DistributedDataParallel: this rank's bucket structure differs from "
                               "rank 0's (the models are not identical across ranks).")
        self.bucket_sizes = [len(b.params) for b in self.buckets]

    # ---- public switches (apex API) ---------------------------------------------------------------------
    def enable_allreduce(self):
        self._enabled = True

    def disable_allreduce(self):
        self._enabled = False

    def forward(self, *inputs, **kwargs):
        self._armed = self._enabled
        self._callback_queued = False
        self._next = 0
        for b in self.buckets:
            b.ready = 0
            b.launched = False
        return self.module(*inputs, **kwargs)

    # ---- backward-time machinery ---------------------------------------------------------------------------
    def _hook(self, param):
        if not self._armed:
            return
        if not self._callback_queued:
            Variable._execution_engine.queue_callback(self._finalize)
            self._callback_queued = True
        b = self._bucket_of[id(param)]
        b.ready += 1
        if b.ready == len(b.params) and not self.delay_allreduce:
            # Collectives must be issued in the same order on every rank, whatever order the gradients
            # arrive in (a data-dependent graph can differ between ranks): buckets go out in index order,
            # a bucket that is ready early waits for its predecessors (apex: next_bucket).
            while self._next < len(self.buckets):
                nb = self.buckets[self._next]
                if not nb.launched:
                    if nb.ready != len(nb.params):
                        break
                    self._reduce_bucket(nb)
                self._next += 1

    def _overflow_flag(self):
        """Device flag of amp's loss scaler (if amp is active): the epilogue doubles as the overflow check."""
        try:
            from ..amp._amp_state import _amp_state
        except Exception:
            return None
        if not (_amp_state.initialized and _amp_state.enabled and len(_amp_state.loss_scalers) == 1):
            return None
        scaler = _amp_state.loss_scalers[0]
        if not self._on_cuda:
            return None
        scaler._ensure_state(self.buckets[0].flat.device)
        return scaler.overflow_ref()

    def _reduce_bucket(self, b):
        b.launched = True
        if self.prof:
            torch.cuda.nvtx.range_push("allreduce_bucket_%d" % b.index)
        side = b.stream
        if self._on_cuda:
            b.ready_event.record(torch.cuda.current_stream())
            side.wait_event(b.ready_event)
            ctx = torch.cuda.stream(side)
        else:
            import contextlib
            ctx = contextlib.nullcontext()
        pre = 1.0 / self.gradient_predivide_factor
        with ctx:
            srcs, dsts = [], []
            in_place = 0
            for p, view in zip(b.params, b.views):
                g = p.grad
                if g is None:
                    view.zero_()                      # parameter unused in this backward
                elif g.data_ptr() == view.data_ptr():
                    in_place += 1
                else:
                    if self._on_cuda:
                        g.record_stream(side)         # the gather reads it on the side stream
                    if layout_key(g) == layout_key(view):
                        srcs.append(g)
                        dsts.append(view)
                    else:                             # produced in another memory layout than the parameter's
                        view.copy_(g)
                        if pre != 1.0:
                            view.mul_(pre)
            if srcs:
                self._ops.gather(srcs, dsts, pre, self._table_cache, b.index)
            if in_place and pre != 1.0:
                # gradients that already live in the bucket (accumulated in place) still need the predivide
                raise RuntimeError("gradient_predivide_factor != 1 cannot be combined with accumulating into "
                                   "bucket views across iterations; call zero_grad() between reductions.")
            if self._use_peer():
                self._peer_allreduce(b)
            else:
                self._nccl_allreduce(b)
            if self._on_cuda:
                b.done_event.record(side)
        for p, view in zip(b.params, b.views):
            p.grad = view
        if self.prof:
            torch.cuda.nvtx.range_pop()

    def _averaging_factor(self):
        if self.gradient_average:
            return self.gradient_predivide_factor / self.world_size
        if self.gradient_predivide_factor != 1.0:
            return self.gradient_predivide_factor
        return 1.0

    def _nccl_allreduce(self, b):
        buf = b.flat
        if b.flat32 is not None:
            self._ops.gather([b.flat], [b.flat32], 1.0, self._table_cache, ("up", b.index))
            buf = b.flat32
        pg = self.bucket_pgs[b.index % self.num_allreduce_streams]
        if self.world_size > 1:
            if pg is not None:
                dist.all_reduce(buf, group=pg)
            else:
                dist.all_reduce(buf)
        factor = self._averaging_factor()
        flag = self._overflow_flag()
        if b.flat32 is not None:
            self._ops.epilogue(buf, factor, None)
            self._ops.gather([b.flat32], [b.flat], 1.0, self._table_cache, ("down", b.index))
            self._ops.epilogue(b.flat, 1.0, flag)
        else:
            self._ops.epilogue(buf, factor, flag)

    def _peer_allreduce(self, b):
        """All-reduce of one bucket in two launches over NVLink peer memory: (1) reduce-scatter + mean +
        overflow bit of this rank's shard, with the cross-GPU hand-shake inside the kernel; (2) gather of
        every shard from its owner into the local bucket, which first waits for all shards and ORs the
        overflow bits into the loss scaler's flag.  The sum is accumulated in fp32 in rank order."""
        b.epoch += 1
        sync = ops.make_peer_sync(b.pads, self._rank, b.epoch)
        ops.peer_reduce_shard(b.grad_ptrs, b.reduced, b.bounds[self._rank], b.bounds[self._rank + 1],
                              self._averaging_factor(), None, sync)
        ops.peer_gather(b.red_ptrs, b.flat, b.bounds, self._overflow_flag(), sync)

    def _finalize(self):
        """End of backward: flush what has not been reduced yet and make the main stream wait."""
        if not self._armed:
            return
        for b in self.buckets:
            if not b.launched:
                self._reduce_bucket(b)
        if self._on_cuda:
            main = torch.cuda.current_stream()
            for b in self.buckets:
                main.wait_event(b.done_event)
            try:
                from ..amp._amp_state import _amp_state
                if self._overflow_flag() is not None:
                    _amp_state.ddp_checked_backward = True
                    _amp_state.ddp_checked_ids = self._checked_ids
            except Exception:
                pass
        self._callback_queued = False
        self._armed = False
