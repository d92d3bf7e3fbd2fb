"""apex.parallel surface: DistributedDataParallel (what the reference wraps its model in) and its
manual-trigger sibling Reducer.  SyncBatchNorm and LARC are out of scope (the
reference never reaches them)."""
from .distributed import DistributedDataParallel, flat_dist_call, split_by_dtype
from .reducer import Reducer

__all__ = ["DistributedDataParallel", "Reducer", "flat_dist_call", "split_by_dtype"]
