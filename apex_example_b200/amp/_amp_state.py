"""Process-wide amp state (mirrors apex/amp/_amp_state.py: one object shared by frontend, handle
and the patched optimizers)."""
from __future__ import annotations


import weakref


class AmpState:
    def __init__(self):
        # live apex.parallel.DistributedDataParallel wrappers (the patched zero_grad asks them to clear
        # gradients that are views of their buckets with one memset per bucket); survives reset()
        self.ddp_wrappers = weakref.WeakSet()
        self.reset()

    def reset(self):
        self.initialized = False
        self.enabled = False
        self.opt_properties = None
        self.loss_scalers = []
        self.verbosity = 1
        self.min_loss_scale = None
        self.max_loss_scale = 2.0 ** 24
        self.optimizers = []
        # set by apex.parallel.DistributedDataParallel when its bucket epilogue has already run
        # the overflow check over every gradient of this backward
        self.ddp_checked_backward = False


_amp_state = AmpState()


def maybe_print(msg: str, rank0: bool = False):
    """apex's maybe_print: honours verbosity; optionally rank-0 only."""
    if _amp_state.verbosity <= 0:
        return
    if rank0:
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_rank() != 0:
                return
        except Exception:
            pass
    print(msg)


def master_params(optimizer):
    """Generator over the parameters the optimizer actually steps (fp32 masters at O2).

    apex materialises fp32 master gradients at ``scale_loss`` exit; the fused optimizers here do
    not (the unscale is folded into the step kernel), so asking for the master params is taken as
    the signal that the caller wants ``master.grad`` (e.g. for clip_grad_norm_) and they are
    materialised on demand."""
    hook = getattr(optimizer, "_amp_materialize_master_grads", None)
    if hook is not None:
        hook()
    for group in optimizer.param_groups:
        for p in group["params"]:
            yield p
