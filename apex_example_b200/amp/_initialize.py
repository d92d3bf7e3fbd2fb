"""Model casting, forward patching and optimizer processing behind amp.initialize
(apex/amp/_initialize.py; SURVEY.md §3.2 lists what the reference's O2 call does)."""
from __future__ import annotations

import functools
from collections.abc import Mapping, Sequence

import torch
from torch.nn.modules.batchnorm import _BatchNorm

from ._amp_state import _amp_state, maybe_print
from ._process_optimizer import _process_optimizer
from .scaler import LossScaler

_HALF_TYPES = (torch.float16, torch.bfloat16)


def to_type(dtype, t):
    if isinstance(t, torch.Tensor):
        if not t.is_cuda:
            maybe_print("Warning:  An input tensor was not cuda.")
        if t.is_floating_point() and t.dtype != dtype:
            return t.to(dtype)
        return t
    return t


def applier(value, fn):
    """Apply fn to every tensor in a nested structure of lists / tuples / dicts."""
    if isinstance(value, torch.Tensor):
        return fn(value)
    if isinstance(value, (str, bytes)):
        return value
    if isinstance(value, Mapping):
        mapped = {applier(k, fn): applier(v, fn) for k, v in value.items()}
        if type(value) is dict:
            return mapped
        try:                                    # dict subclasses (e.g. dataclass-style model outputs) keep their type
            return type(value)(mapped)
        except Exception:
            try:
                return type(value)(**mapped)
            except Exception:
                return mapped
    if isinstance(value, tuple) and hasattr(value, "_fields"):       # namedtuple
        return type(value)(*(applier(v, fn) for v in value))
    if isinstance(value, Sequence):
        return type(value)(applier(v, fn) for v in value)
    if hasattr(value, "to") and not isinstance(value, torch.nn.Module):
        return value
    return value


def check_models(models):
    for model in models:
        parallel_type = None
        if isinstance(model, torch.nn.parallel.DistributedDataParallel):
            parallel_type = "torch.nn.parallel.DistributedDataParallel"
        if type(model).__name__ == "DistributedDataParallel" and type(model).__module__.startswith(
                ("apex", "apex_example_b200")):
            parallel_type = "apex.parallel.DistributedDataParallel"
        if isinstance(model, torch.nn.parallel.DataParallel):
            parallel_type = "torch.nn.parallel.DataParallel"
        if parallel_type is not None:
            raise RuntimeError("Incoming model is an instance of {}. ".format(parallel_type) +
                               "Parallel wrappers should only be applied to the model(s) AFTER \n"
                               "the model(s) have been returned from amp.initialize.")


def check_params_fp32(models):
    for model in models:
        for name, param in model.named_parameters():
            if param.is_floating_point():
                if not param.is_cuda:
                    maybe_print("Warning:  Found param {} with type {}, expected torch.cuda.FloatTensor.\n"
                                "When using amp.initialize, you need to provide a model with parameters\n"
                                "located on a CUDA device before passing it no matter what optimization level\n"
                                "you chose. Use model.to('cuda') to use the default device.".format(name, param.type()))
                elif param.dtype != torch.float32:
                    maybe_print("Warning:  Found param {} with type {}, expected torch.cuda.FloatTensor.\n"
                                "When using amp.initialize, you do not need to call .half() on your model\n"
                                "before passing it, no matter what optimization level you choose.".format(
                                    name, param.type()))


def check_optimizers(optimizers):
    for optim in optimizers:
        bad = None
        if type(optim).__name__ in ("FP16_Optimizer", "FP16_Optimizer_general", "FP16_Optimizer_for_fused"):
            bad = type(optim).__name__
        if bad is not None:
            raise RuntimeError("An incoming optimizer is an instance of {}. ".format(bad) +
                               "The optimizer(s) passed to amp.initialize() must be bare \n"
                               "instances of either ordinary Pytorch optimizers, or Apex fused \n"
                               "optimizers.\n")


def convert_module(module, dtype):
    for param in module.parameters(recurse=False):
        if param is not None and param.data.is_floating_point():
            param.data = param.data.to(dtype=dtype)
            if param._grad is not None and param._grad.data.is_floating_point():
                param._grad.data = param._grad.data.to(dtype=dtype)
    for name, buf in module.named_buffers(recurse=False):
        if buf is not None and buf.data.is_floating_point():
            module._buffers[name] = buf.data.to(dtype=dtype)


def convert_network(network, dtype, keep_batchnorm_fp32):
    """Cast the network's floating-point tensors to ``dtype``; affine batch norms stay fp32 when asked
    (O2: keep_batchnorm_fp32=True)."""
    for module in network.modules():
        if keep_batchnorm_fp32 and isinstance(module, _BatchNorm) and module.affine is True:
            continue
        convert_module(module, dtype)
        if isinstance(module, (torch.nn.RNNBase, torch.nn.modules.rnn.RNNBase)):
            module.flatten_parameters()
    return network


def _patch_forward(model, input_caster, output_caster, autocast_dtype=None):
    old_fwd = model.forward

    @functools.wraps(old_fwd)
    def new_fwd(*args, **kwargs):
        if input_caster is not None:
            args = applier(args, input_caster)
            kwargs = applier(kwargs, input_caster)
        if autocast_dtype is not None:
            with torch.autocast(device_type="cuda", dtype=autocast_dtype):
                output = old_fwd(*args, **kwargs)
        else:
            output = old_fwd(*args, **kwargs)
        if output_caster is not None:
            output = applier(output, output_caster)
        return output

    model.forward = new_fwd


class _O2StateDictHook:
    """model.state_dict() hands out fp32 weights so checkpoints do not depend on the opt_level (apex
    O2StateDictHook).  apex casts the fp16 weights up, which drops the masters' extra bits; here a
    parameter that has an fp32 master is saved AS its master, so a restored run continues exactly."""

    def __init__(self, master_of):
        self.master_of = master_of          # id(model param) -> fp32 master, filled by _initialize

    def __call__(self, module, state_dict, prefix, local_metadata):
        for name, param in module._parameters.items():
            key = prefix + name
            if param is None or key not in state_dict:
                continue
            master = self.master_of.get(id(param))
            if master is not None:
                state_dict[key] = master.detach().clone()
        for key in state_dict:
            param = state_dict[key]
            if param.dtype in _HALF_TYPES:
                state_dict[key] = param.to(torch.float32)


def _initialize(models, optimizers, properties, num_losses=1, cast_model_outputs=None):
    from torch.optim import Optimizer

    if isinstance(optimizers, Optimizer) or type(optimizers).__name__ == "LARC":
        optimizers_was_list = False
        optimizers = [optimizers]
    elif optimizers is None:
        optimizers_was_list = False
        optimizers = []
    elif isinstance(optimizers, list):
        optimizers_was_list = True
    else:
        check_optimizers([optimizers])
        raise TypeError("optimizers must be either a single optimizer or a list of optimizers.")

    if isinstance(models, torch.nn.Module):
        models_was_list = False
        models = [models]
    elif isinstance(models, list):
        models_was_list = True
    else:
        raise TypeError("models must be either a single model or a list of models.")

    master_of = {}                          # id(model param) -> fp32 master (O2), shared with the state_dict hooks
    check_models(models)
    if not getattr(_amp_state, "allow_incoming_model_not_fp32", False):
        check_params_fp32(models)
    check_optimizers(optimizers)

    if properties.cast_model_type:
        dtype = properties.cast_model_type
        for model in models:
            convert_network(model, dtype, bool(properties.keep_batchnorm_fp32))
        input_caster = functools.partial(to_type, dtype)
        out_dtype = cast_model_outputs if cast_model_outputs is not None else torch.float32
        output_caster = functools.partial(to_type, out_dtype)
        for model in models:
            _patch_forward(model, input_caster, output_caster)
        # state_dict() hands out fp32 weights so checkpoints are O-level independent
        if dtype in _HALF_TYPES:
            for model in models:
                for module in model.modules():
                    module._register_state_dict_hook(_O2StateDictHook(master_of))
    elif cast_model_outputs is not None:
        output_caster = functools.partial(to_type, cast_model_outputs)
        for model in models:
            _patch_forward(model, None, output_caster)

    if properties.patch_torch_functions:
        # O1.  apex monkey-patches torch.* / Tensor.* / F.* with white/black lists; the same casting
        # policy ships in torch as autocast, so the forward is wrapped in it.
        for model in models:
            _patch_forward(model, None, None, autocast_dtype=torch.float16)

    for i, optimizer in enumerate(optimizers):
        optimizers[i] = _process_optimizer(optimizer, properties)

    if properties.master_weights:
        # model.load_state_dict(fp32 checkpoint) would only reach the fp16 weights (rounded); hand the
        # exact fp32 values to the masters as well, so a restored run continues bit-exactly
        for optimizer in optimizers:
            for master, model_p in optimizer._amp_stash.model_of.items():
                master_of[id(model_p)] = master

        def _masters_from_checkpoint(module, state_dict, prefix, *unused):
            with torch.no_grad():
                for name, param in module._parameters.items():
                    if param is None:
                        continue
                    master = master_of.get(id(param))
                    src = state_dict.get(prefix + name)
                    if master is not None and src is not None and src.shape == master.shape:
                        master.copy_(src.to(device=master.device, dtype=torch.float32))

        for model in models:
            for module in model.modules():
                module._register_load_state_dict_pre_hook(_masters_from_checkpoint, with_module=True)

    _amp_state.loss_scalers = []
    for idx in range(num_losses):
        scaler = LossScaler(properties.loss_scale, min_loss_scale=_amp_state.min_loss_scale,
                            max_loss_scale=_amp_state.max_loss_scale)
        scaler._loss_id = idx
        _amp_state.loss_scalers.append(scaler)
    _amp_state.optimizers = list(optimizers)
    _amp_state.initialized = True
    _amp_state.enabled = True

    if optimizers_was_list:
        return (models if models_was_list else models[0]), optimizers
    if len(optimizers) == 0:
        return models if models_was_list else models[0]
    return (models if models_was_list else models[0]), optimizers[0]
