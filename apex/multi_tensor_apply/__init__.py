from apex_example_b200.multi_tensor_apply import MultiTensorApply, multi_tensor_applier  # noqa: F401
