"""amp.initialize / state_dict / load_state_dict for the loss-scaling runtime (reference:
apex/amp/frontend.py).  The opt-level table is the reference's (frontend.py:118-256).  Function-level
cast patching of O1/O4 (apex/amp/amp.py, wrap.py, lists/) is out of scope here; those levels run the
model's forward under torch.autocast instead, which is the stock-PyTorch equivalent."""
from collections import OrderedDict

import torch

from ._amp_state import _amp_state, maybe_print
from ._process_optimizer import _process_optimizer
from .scaler import LossScaler


class Properties(object):
    def __init__(self):
        self.enabled = False
        self.opt_level = None
        self.cast_model_type = None
        self.patch_torch_functions = False
        self.patch_torch_functions_type = None
        self.keep_batchnorm_fp32 = None
        self.master_weights = None
        self.loss_scale = 1.0

    @property
    def options(self):
        return {k: v for k, v in self.__dict__.items()}


#            cast_model_type, autocast dtype, keep_bn_fp32, master_weights, loss_scale
OPT_LEVELS = {
    "O0": (torch.float32, None, None, False, 1.0),
    "O1": (None, torch.float16, None, None, "dynamic"),
    "O2": (torch.float16, None, True, True, "dynamic"),
    "O3": (torch.float16, None, False, False, 1.0),
    "O4": (None, torch.bfloat16, None, None, 1),
    "O5": (torch.bfloat16, None, True, True, 1),
}

_BN = (torch.nn.modules.batchnorm._BatchNorm,)


def _cast_model(model, dtype, keep_batchnorm_fp32):
    for module in model.modules():
        if keep_batchnorm_fp32 and isinstance(module, _BN):
            module.float()
            continue
        for t in list(module.parameters(recurse=False)) + list(module.buffers(recurse=False)):
            if t.is_floating_point():
                t.data = t.data.to(dtype)
    return model


def _wrap_forward(model, in_dtype, out_dtype, autocast_dtype):
    inner = model.forward

    def cast(x, dtype):
        if isinstance(x, torch.Tensor):
            return x.to(dtype) if x.is_floating_point() else x
        if isinstance(x, (list, tuple)):
            return type(x)(cast(v, dtype) for v in x)
        if isinstance(x, dict):
            return {k: cast(v, dtype) for k, v in x.items()}
        return x

    def forward(*args, **kwargs):
        if autocast_dtype is not None:
            with torch.autocast("cuda", dtype=autocast_dtype):
                return inner(*args, **kwargs)
        out = inner(*cast(args, in_dtype), **cast(kwargs, in_dtype))
        return cast(out, out_dtype) if out_dtype is not None else out

    model.forward = forward


def initialize(models, optimizers=None, enabled=True, opt_level="O1", cast_model_type=None,
               patch_torch_functions=None, patch_torch_functions_type=None, keep_batchnorm_fp32=None,
               master_weights=None, loss_scale=None, cast_model_outputs=None, num_losses=1, verbosity=1,
               min_loss_scale=None, max_loss_scale=2. ** 24):
    """Same call as the reference (frontend.py:259-425): returns (models, optimizers) in the shapes given."""
    _amp_state.opt_properties = props = Properties()
    _amp_state.verbosity = verbosity
    if not enabled:
        return models if optimizers is None else (models, optimizers)
    if opt_level not in OPT_LEVELS:
        raise RuntimeError("Unexpected optimization level {}. Options are 'O0', 'O1', 'O2', 'O3', 'O4', 'O5'.  "
                           "Note that in `O0`, `O1`, etc., the prefix O is the letter O, not the number zero.".format(opt_level))
    cmt, ac, kbn, mw, ls = OPT_LEVELS[opt_level]
    props.enabled, props.opt_level = True, opt_level
    props.cast_model_type, props.patch_torch_functions_type = cmt, ac
    props.patch_torch_functions = ac is not None
    props.keep_batchnorm_fp32, props.master_weights, props.loss_scale = kbn, mw, ls
    if cast_model_type is not None:
        props.cast_model_type = cast_model_type or None      # False = "leave the model alone" (reference frontend.py:65-77)
    if keep_batchnorm_fp32 is not None:
        props.keep_batchnorm_fp32 = keep_batchnorm_fp32 if not isinstance(keep_batchnorm_fp32, str) else keep_batchnorm_fp32 == "True"
    if master_weights is not None:
        props.master_weights = master_weights
    if loss_scale is not None:
        props.loss_scale = loss_scale if loss_scale == "dynamic" else float(loss_scale)
    if patch_torch_functions is not None:
        props.patch_torch_functions = patch_torch_functions
    if patch_torch_functions_type is not None:
        props.patch_torch_functions_type = patch_torch_functions_type

    models_was_list = isinstance(models, list)
    model_list = models if models_was_list else [models]
    opt_was_list = isinstance(optimizers, list)
    opt_list = [] if optimizers is None else (optimizers if opt_was_list else [optimizers])

    for model in model_list:
        for name, p in model.named_parameters():
            if not p.is_cuda:
                raise RuntimeError("Found param {} with type {}, expected torch.cuda.FloatTensor.\nWhen using amp.initialize, "
                                   "you need to provide a model with parameters located on a CUDA device before passing it "
                                   "no matter what optimization level you chose.".format(name, p.type()))
        if props.cast_model_type is not None and props.cast_model_type != torch.float32:
            _cast_model(model, props.cast_model_type, props.keep_batchnorm_fp32)
            _wrap_forward(model, props.cast_model_type, cast_model_outputs or torch.float32, None)
        elif props.patch_torch_functions and props.patch_torch_functions_type is not None:
            _wrap_forward(model, None, cast_model_outputs, props.patch_torch_functions_type)

    for opt in opt_list:
        _process_optimizer(opt, props)

    _amp_state.min_loss_scale, _amp_state.max_loss_scale = min_loss_scale, max_loss_scale
    _amp_state.loss_scalers = [LossScaler(props.loss_scale, min_loss_scale=min_loss_scale, max_loss_scale=max_loss_scale)
                               for _ in range(num_losses)]
    if optimizers is None:
        return models
    return models, (opt_list if opt_was_list else opt_list[0])


def state_dict(destination=None):
    """{'loss_scaler%d': {'loss_scale', 'unskipped'}} (reference frontend.py:428-437)."""
    destination = OrderedDict() if destination is None else destination
    for idx, ls in enumerate(_amp_state.loss_scalers):
        destination["loss_scaler%d" % idx] = {"loss_scale": ls.loss_scale(), "unskipped": ls._unskipped}
    return destination


def load_state_dict(state_dict):
    if len(state_dict) != len(_amp_state.loss_scalers):
        print("Warning: state_dict contains {} entries, while {} loss_scalers are used".format(
            len(state_dict), len(_amp_state.loss_scalers)))
    unexpected = [k for k in state_dict if "loss_scaler" not in k]
    if unexpected:
        raise RuntimeError("Error(s) in loading state_dict. Unexpected key(s) in state_dict: {}. ".format(
            ", ".join('"{}"'.format(k) for k in unexpected)))
    for idx, key in enumerate(state_dict):
        if idx >= len(_amp_state.loss_scalers):
            print("Skipping loss_scaler[{}], since num_losses was set to {}".format(idx, len(_amp_state.loss_scalers)))
            break
        _amp_state.loss_scalers[idx]._loss_scale = state_dict[key]["loss_scale"]
        _amp_state.loss_scalers[idx]._unskipped = state_dict[key]["unskipped"]
