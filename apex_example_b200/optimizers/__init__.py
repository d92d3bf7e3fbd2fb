from .fused_adam import FusedAdam
from .fused_lamb import FusedLAMB
from .fused_novograd import FusedNovoGrad
from .fused_sgd import FusedSGD

__all__ = ["FusedAdam", "FusedLAMB", "FusedNovoGrad", "FusedSGD"]
