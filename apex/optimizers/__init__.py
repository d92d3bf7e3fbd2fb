from apex_example_b200.optimizers import FusedAdam, FusedSGD  # noqa: F401
