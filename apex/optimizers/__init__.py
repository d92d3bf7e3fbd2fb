from apex_example_b200.optimizers import FusedAdam, FusedLAMB, FusedSGD  # noqa: F401
