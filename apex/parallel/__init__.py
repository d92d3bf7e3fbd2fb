from apex_example_b200.parallel import DistributedDataParallel, Reducer, flat_dist_call  # noqa: F401
