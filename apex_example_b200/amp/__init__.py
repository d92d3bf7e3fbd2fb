"""apex.amp surface: initialize, scale_loss, master_params, state_dict,
load_state_dict and the O1 registration helpers."""
from ._amp_state import _amp_state, master_params
from .frontend import initialize, load_state_dict, state_dict
from .handle import disable_casts, scale_loss
from .scaler import LossScaler


def _identity_decorator(fn):
    return fn


# O1 function-registration API.  apex uses these to extend its monkey-patch lists; O1 here is
# torch.autocast, whose policy is fixed, so the decorators only keep user code
# importable and the register_* calls are accepted and ignored.
def half_function(fn):
    import functools
    import torch

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        with torch.autocast(device_type="cuda", enabled=False):
            cast = lambda t: t.half() if isinstance(t, torch.Tensor) and t.is_floating_point() else t  # noqa: E731
            return fn(*[cast(a) for a in args], **{k: cast(v) for k, v in kwargs.items()})
    return wrapper


def float_function(fn):
    import functools
    import torch

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        with torch.autocast(device_type="cuda", enabled=False):
            cast = lambda t: t.float() if isinstance(t, torch.Tensor) and t.is_floating_point() else t  # noqa: E731
            return fn(*[cast(a) for a in args], **{k: cast(v) for k, v in kwargs.items()})
    return wrapper


promote_function = _identity_decorator


def register_half_function(module, name):
    setattr(module, name, half_function(getattr(module, name)))


def register_float_function(module, name):
    setattr(module, name, float_function(getattr(module, name)))


def register_promote_function(module, name):
    pass


__all__ = ["initialize", "scale_loss", "master_params", "state_dict", "load_state_dict", "disable_casts",
           "half_function", "float_function", "promote_function", "register_half_function",
           "register_float_function", "register_promote_function", "LossScaler"]
