"""Drop-in for apex's ``amp_C`` extension module: same call signatures, served by
libapex_b200.so through its C ABI.  Third-party code that does
``multi_tensor_applier(amp_C.multi_tensor_scale, flag, [src, dst], s)`` keeps working."""
from apex_example_b200.multi_tensor_apply import (multi_tensor_adam, multi_tensor_axpby,  # noqa: F401
                                                  multi_tensor_l2norm, multi_tensor_scale, multi_tensor_sgd)
