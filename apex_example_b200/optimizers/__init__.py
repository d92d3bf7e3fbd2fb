from .fused_adam import FusedAdam
from .fused_sgd import FusedSGD

__all__ = ["FusedAdam", "FusedSGD"]
