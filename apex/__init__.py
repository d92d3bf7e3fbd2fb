"""Drop-in ``apex`` package: the names the reference imports
(/root/reference/test_apex_distributed_spawn.py:31-32 — ``import apex.parallel`` and
``from apex import amp``) resolve to the B200-native implementation in apex_example_b200."""
from apex_example_b200 import __version__  # noqa: F401

from . import amp, multi_tensor_apply, optimizers, parallel  # noqa: F401
