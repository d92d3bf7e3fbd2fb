"""apex.parallel surface used by the reference (DistributedDataParallel).  SyncBatchNorm, LARC
and Reducer are out of scope (the reference never reaches them)."""
from .distributed import DistributedDataParallel, flat_dist_call, split_by_dtype

__all__ = ["DistributedDataParallel", "flat_dist_call", "split_by_dtype"]
