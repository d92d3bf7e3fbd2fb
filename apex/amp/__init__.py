from apex_example_b200.amp import *  # noqa: F401,F403
from apex_example_b200.amp import __all__, _amp_state  # noqa: F401
from apex_example_b200.amp import frontend, handle, scaler  # noqa: F401
