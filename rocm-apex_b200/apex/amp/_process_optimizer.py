"""Optimizer patching for amp (reference: apex/amp/_process_optimizer.py:1-489).

amp.initialize() hands every optimizer through process_optimizer(), which attaches an `_amp_stash`
and the hooks scale_loss() drives:
    _prepare_amp_backward()   before backward: stash or drop existing grads so that backward writes fresh ones
    _post_amp_backward(scaler) after backward: unscale (and, with master weights, copy fp16 grads into fp32
                               master grads) through LossScaler -> multi_tensor_scale / multi_tensor_axpby
and, with master weights, wraps step() so that the updated fp32 masters are written back into the
fp16/bf16 model parameters with one multi_tensor_scale(…, 1.0) launch.
Attribute names on `_amp_stash` are the reference's: FusedSGD reads them directly.
"""
import types

import torch

from ..multi_tensor_apply import multi_tensor_applier
from ..optimizers import FusedSGD

LOW_PRECISION = (torch.float16, torch.bfloat16)


class AmpOptimizerState(object):
    pass


def _classify(param):
    if not param.is_cuda:
        raise TypeError("Optimizer's parameters must be one of torch.cuda.FloatTensor, torch.cuda.HalfTensor, "
                        "torch.cuda.BFloat16Tensor. Received {}".format(param.type()))
    if param.dtype in LOW_PRECISION:
        return "low"
    if param.dtype == torch.float32:
        return "fp32"
    raise TypeError("Optimizer's parameters must be one of torch.cuda.FloatTensor, torch.cuda.HalfTensor, "
                    "torch.cuda.BFloat16Tensor. Received {}".format(param.type()))


def _flatten(groups):
    return [p for g in groups for p in g]


# ---------------------------------------------------------------------------------------------------
# shared: unscale grads that live on the very tensors the optimizer steps
# ---------------------------------------------------------------------------------------------------
def post_backward_models_are_masters(scaler, params, stashed_grads, scale_override=None):
    """reference :93-139.  A stashed grad is one that existed before this backward (accumulation)."""
    grads_have_scale, stashed_have_scale, out_scale = scaler.loss_scale(), 1.0, 1.0
    if scaler.loss_scale() == 1.0 and not scaler.dynamic:
        stashed_grads[:] = [None] * len(stashed_grads)   # backward accumulated directly; nothing to undo
        return
    if scale_override is not None:
        grads_have_scale, stashed_have_scale, out_scale = scale_override
    fresh, accum, stash_for_accum = [], [], []
    for param, stashed in zip(params, stashed_grads):
        if param.grad is None and stashed is not None:
            param.grad = stashed
        elif param.grad is not None and stashed is None:
            fresh.append(param.grad)
        elif param.grad is not None and stashed is not None:
            accum.append(param.grad)
            stash_for_accum.append(stashed)
    if fresh:
        scaler.unscale(fresh, fresh, None, models_are_masters=True, scale_override=grads_have_scale / out_scale)
    if accum:
        scaler.unscale_with_stashed(accum, stash_for_accum, accum,
                                    scale_override=(grads_have_scale, stashed_have_scale, out_scale))
    stashed_grads[:] = [None] * len(stashed_grads)


# ---------------------------------------------------------------------------------------------------
# master-weight mode (O2 / O5)
# ---------------------------------------------------------------------------------------------------
def _split_group_with_masters(optimizer, params):
    """Replaces low-precision params of one group by fp32 master copies, in place in `params`."""
    low, masters, fp32 = [], [], []
    for i, param in enumerate(params):
        if not param.requires_grad:
            continue
        kind = _classify(param)
        if kind == "low":
            master = param.detach().clone().float()
            master.requires_grad = True
            params[i] = master
            if optimizer is not None and param in optimizer.state:
                optimizer.state[master] = optimizer.state.pop(param)
            low.append(param)
            masters.append(master)
        else:
            fp32.append(param)
    return low, masters, fp32


def lazy_init_with_master_weights(self):
    stash = self._amp_stash
    stash.fp16_groups, stash.fp32_from_fp16_groups, stash.fp32_from_fp32_groups = [], [], []
    for group in self.param_groups:
        low, masters, fp32 = _split_group_with_masters(self, group["params"])
        stash.fp16_groups.append(low)
        stash.fp32_from_fp16_groups.append(masters)
        stash.fp32_from_fp32_groups.append(fp32)
    stash.all_fp16_params = _flatten(stash.fp16_groups)
    stash.all_fp32_from_fp16_params = _flatten(stash.fp32_from_fp16_groups)
    stash.all_fp32_from_fp32_params = _flatten(stash.fp32_from_fp32_groups)
    stash.all_fp16_grad_stash = [None] * len(stash.all_fp16_params)
    stash.all_fp32_from_fp32_grad_stash = [None] * len(stash.all_fp32_from_fp32_params)
    for p in stash.all_fp32_from_fp16_params + stash.all_fp32_from_fp32_params:
        p.grad = None
    self.load_state_dict(self.state_dict())   # re-casts optimizer state to the new master params


def master_params_to_model_params(self):
    """fp32 masters -> low-precision model params, one launch (reference :14-25)."""
    stash = self._amp_stash
    if len(stash.all_fp16_params) > 0:
        multi_tensor_applier(stash.multi_tensor_scale, stash.dummy_overflow_buf,
                             [stash.all_fp32_from_fp16_params, stash.all_fp16_params], 1.0)


def prepare_backward_with_master_weights(self):
    stash = self._amp_stash
    self._amp_lazy_init()
    for p in stash.all_fp16_params:
        p.grad = None     # model grads are write-only scratch: backward recreates them
    for i, p in enumerate(stash.all_fp32_from_fp32_params):
        stash.all_fp32_from_fp32_grad_stash[i] = p.grad
        p.grad = None


def post_backward_with_master_weights(self, scaler):
    stash = self._amp_stash
    self._amp_lazy_init()
    fresh_model, fresh_master, accum_model, accum_master = [], [], [], []
    for model_p, master_p in zip(stash.all_fp16_params, stash.all_fp32_from_fp16_params):
        if model_p.grad is None:
            continue
        if master_p.grad is None:
            master_p.grad = torch.empty_like(master_p)
            fresh_model.append(model_p.grad)
            fresh_master.append(master_p.grad)
        else:
            accum_model.append(model_p.grad)
            accum_master.append(master_p.grad)
    if fresh_model:
        scaler.unscale(fresh_model, fresh_master, scaler.loss_scale(), models_are_masters=False)
    if accum_model:
        scaler.unscale_with_stashed(accum_model, accum_master, accum_master)
    post_backward_models_are_masters(scaler, stash.all_fp32_from_fp32_params, stash.all_fp32_from_fp32_grad_stash)


# ---------------------------------------------------------------------------------------------------
# no master weights (O0 / O1 / O3 / O4)
# ---------------------------------------------------------------------------------------------------
def lazy_init_no_master_weights(self):
    stash = self._amp_stash
    stash.all_fp16_params, stash.all_fp32_params = [], []
    for group in self.param_groups:
        for p in group["params"]:
            (stash.all_fp16_params if _classify(p) == "low" else stash.all_fp32_params).append(p)
    stash.all_fp16_grad_stash = [None] * len(stash.all_fp16_params)
    stash.all_fp32_grad_stash = [None] * len(stash.all_fp32_params)


def prepare_backward_no_master_weights(self):
    stash = self._amp_stash
    self._amp_lazy_init()
    for params, slots in ((stash.all_fp16_params, stash.all_fp16_grad_stash),
                          (stash.all_fp32_params, stash.all_fp32_grad_stash)):
        for i, p in enumerate(params):
            slots[i] = p.grad
            p.grad = None


def post_backward_no_master_weights(self, scaler):
    stash = self._amp_stash
    self._amp_lazy_init()
    post_backward_models_are_masters(scaler, stash.all_fp16_params, stash.all_fp16_grad_stash)
    post_backward_models_are_masters(scaler, stash.all_fp32_params, stash.all_fp32_grad_stash)


# ---------------------------------------------------------------------------------------------------
# FusedSGD: can consume scaled grads directly (the unscale rides in the kernel), reference :258-310
# ---------------------------------------------------------------------------------------------------
def prepare_backward_with_master_weights_FusedSGD(self):
    if self.materialize_master_grads:
        return prepare_backward_with_master_weights(self)
    stash = self._amp_stash
    self._amp_lazy_init()
    for params, slots in ((stash.all_fp16_params, stash.all_fp16_grad_stash),
                          (stash.all_fp32_from_fp32_params, stash.all_fp32_from_fp32_grad_stash)):
        for i, p in enumerate(params):
            slots[i] = p.grad
            p.grad = None


def post_backward_with_master_weights_FusedSGD(self, scaler):
    if self.materialize_master_grads:
        return post_backward_with_master_weights(self, scaler)
    stash = self._amp_stash
    self._amp_lazy_init()
    grads_have_scale = scaler.loss_scale()
    stashed_have_scale = self.most_recent_scale
    out_scale = min(grads_have_scale, self.most_recent_scale) if self.scale_set_by_backward else grads_have_scale
    for params, slots in ((stash.all_fp16_params, stash.all_fp16_grad_stash),
                          (stash.all_fp32_from_fp32_params, stash.all_fp32_from_fp32_grad_stash)):
        post_backward_models_are_masters(scaler, params, slots, (grads_have_scale, stashed_have_scale, out_scale))
    self.most_recent_scale = out_scale
    self.scale_set_by_backward = True


def _amp_lazy_init(self):
    stash = self._amp_stash
    if not stash.lazy_init_called:
        self._lazy_init_maybe_master_weights()
        stash.lazy_init_called = True


HOOK_NAMES = ("_lazy_init_maybe_master_weights", "_master_params_to_model_params", "_prepare_amp_backward",
              "_post_amp_backward", "_amp_lazy_init")


def _process_optimizer(optimizer, properties):
    if hasattr(optimizer, "_amp_stash"):
        raise RuntimeError("A given optimizer should only be passed through amp.initialize once.")
    for name in HOOK_NAMES:
        if hasattr(optimizer, name):
            raise RuntimeError("Incoming optimizer already has {} defined.".format(name))
    stash = optimizer._amp_stash = AmpOptimizerState()
    stash.lazy_init_called = False
    stash.already_patched = False
    stash.params_have_scaled_gradients = False
    if not multi_tensor_applier.available:
        raise RuntimeError("amp requires the multi-tensor kernels (libb200mt.so); there is no python fallback")
    import amp_C
    stash.multi_tensor_scale = amp_C.multi_tensor_scale
    stash.multi_tensor_l2norm = amp_C.multi_tensor_l2norm
    stash.dummy_overflow_buf = torch.zeros(1, dtype=torch.int32, device="cuda")

    def bind(name, fn):
        setattr(optimizer, name, types.MethodType(fn, optimizer))

    sgd = isinstance(optimizer, FusedSGD)
    if properties.master_weights:
        bind("_lazy_init_maybe_master_weights", lazy_init_with_master_weights)
        bind("_master_params_to_model_params", master_params_to_model_params)
        bind("_prepare_amp_backward", prepare_backward_with_master_weights_FusedSGD if sgd else prepare_backward_with_master_weights)
        bind("_post_amp_backward", post_backward_with_master_weights_FusedSGD if sgd else post_backward_with_master_weights)
        inner_step = optimizer.step

        def step_then_copy_back(self, closure=None):
            if closure is not None:
                raise RuntimeError("Currently, Amp does not support closure use with optimizers.")
            ret = inner_step()
            if not sgd:                       # FusedSGD writes the model copy inside its own kernel
                self._master_params_to_model_params()
            for p in self._amp_stash.all_fp32_from_fp16_params:
                p.grad = None                 # master grads are per-iteration scratch
            return ret

        def zero_grad_with_masters(self, set_to_none=False):
            st = self._amp_stash
            self._amp_lazy_init()
            for p in st.all_fp16_params + st.all_fp32_from_fp32_params:
                if p.grad is not None:
                    p.grad.detach_()
                    p.grad.zero_()
            for p in st.all_fp32_from_fp16_params:
                p.grad = None

        bind("step", step_then_copy_back)
        bind("zero_grad", zero_grad_with_masters)
    else:
        bind("_lazy_init_maybe_master_weights", lazy_init_no_master_weights)
        bind("_prepare_amp_backward", prepare_backward_no_master_weights)
        bind("_post_amp_backward", post_backward_no_master_weights)
    bind("_amp_lazy_init", _amp_lazy_init)

    inner_add_param_group = optimizer.add_param_group

    def add_param_group(self, new_group):
        """Keeps the stash in step with param groups added after amp.initialize (reference :411-487)."""
        st = self._amp_stash
        self._amp_lazy_init()
        assert isinstance(new_group, dict), "param group must be a dict"
        params = new_group["params"]
        if isinstance(params, torch.Tensor):
            new_group["params"] = [params]
        elif isinstance(params, set):
            raise TypeError("optimizer parameters need to be organized in ordered collections, but the ordering of "
                            "tensors in sets will change between runs. Please use a list instead.")
        else:
            new_group["params"] = list(params)
        if properties.master_weights:
            low, masters, fp32 = _split_group_with_masters(None, new_group["params"])
            st.fp16_groups.append(low)
            st.fp32_from_fp16_groups.append(masters)
            st.fp32_from_fp32_groups.append(fp32)
            st.all_fp16_params += low
            st.all_fp32_from_fp16_params += masters
            st.all_fp32_from_fp32_params += fp32
            st.all_fp16_grad_stash += [None] * len(low)
            st.all_fp32_from_fp32_grad_stash += [None] * len(fp32)
        else:
            for p in new_group["params"]:
                if _classify(p) == "low":
                    st.all_fp16_params.append(p)
                    st.all_fp16_grad_stash.append(None)
                else:
                    st.all_fp32_params.append(p)
                    st.all_fp32_grad_stash.append(None)
        inner_add_param_group(new_group)

    bind("add_param_group", add_param_group)
    return optimizer
