from apex_example_b200.optimizers import FusedAdam, FusedLAMB, FusedNovoGrad, FusedSGD  # noqa: F401
