"""apex_example_b200 — B200-native (sm_100a) mixed-precision training step behind apex's API.

The hot path of enijkamp/apex_example (/root/reference/test_apex_distributed_spawn.py:121-164):
``amp.initialize`` / ``amp.scale_loss`` / ``apex.parallel.DistributedDataParallel`` /
``apex.optimizers.FusedAdam|FusedSGD``, re-implemented over hand-written CUDA kernels that are
reached through a plain C ABI (include/apex_b200.h, libapex_b200.so).  The top-level ``apex``
package in this repository re-exports these modules under apex's names so the reference script
runs unmodified.
"""
__version__ = "0.1.0"

from . import _lib  # noqa: F401
