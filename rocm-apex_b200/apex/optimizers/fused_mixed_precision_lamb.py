"""FusedMixedPrecisionLamb (reference: apex/optimizers/fused_mixed_precision_lamb.py:9-256): LAMB with
fp32 master weights for reduced-precision model parameters, and no host synchronisation anywhere:
learning rate and step are device tensors, the loss scale and found-inf flag come straight from a
torch GradScaler, and every kernel skips itself when the overflow flag is set."""
from collections import defaultdict
from copy import deepcopy
from itertools import chain

import torch

from ..multi_tensor_apply import multi_tensor_applier
from . import _common


class FusedMixedPrecisionLamb(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, step=0, bias_correction=True, betas=(0.9, 0.999), eps=1e-6,
                 weight_decay=0.01, amsgrad=False, adam_w_mode=True, grad_averaging=True,
                 max_grad_norm=1.0, use_nvlamb=False, reduced_precision_dtype=None):
        if amsgrad:
            raise RuntimeError("FusedLAMB does not support the AMSGrad variant.")
        defaults = dict(lr=torch.tensor(lr, dtype=torch.float32), step=torch.tensor([step], dtype=torch.int),
                        bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        grad_averaging=grad_averaging, max_grad_norm=max_grad_norm)
        super().__init__(params, defaults)
        device = self.param_groups[0]["params"][0].device
        for group in self.param_groups:           # lr and step live on the device (sync-free stepping)
            for key in ("lr", "step"):
                group[key] = group[key].to(device=device)
        amp_C = _common.require_extension("FusedMixedPrecisionLamb")
        self.multi_tensor_l2norm = amp_C.multi_tensor_l2norm_mp
        self.multi_tensor_lamb = amp_C.multi_tensor_lamb_mp
        self._dummy_overflow_buf = _common.overflow_buf(device)
        self.reduced_precision_dtype = reduced_precision_dtype
        self.param_groups_full_precision = []
        self._step_supports_amp_scaling = True
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.use_nvlamb = use_nvlamb

    def load_state_dict(self, state_dict):
        """Like Optimizer.load_state_dict but without casting state to the parameter dtype: the
        moments of reduced-precision parameters must stay fp32 (reference :55-115)."""
        state_dict = deepcopy(state_dict)
        groups, saved = self.param_groups, state_dict["param_groups"]
        if len(groups) != len(saved):
            raise ValueError("loaded state dict has a different number of parameter groups")
        if any(len(g["params"]) != len(s["params"]) for g, s in zip(groups, saved)):
            raise ValueError("loaded state dict contains a parameter group that doesn't match the size of optimizer's group")
        id_map = dict(zip(chain.from_iterable(s["params"] for s in saved),
                          chain.from_iterable(g["params"] for g in groups)))
        state = defaultdict(dict)
        for k, v in state_dict["state"].items():
            state[id_map.get(k, k)] = v
        new_groups = []
        for g, s in zip(groups, saved):
            s["params"] = g["params"]
            new_groups.append(s)
        self.__setstate__({"state": state, "param_groups": new_groups})

    def _setup_full_precision_params(self):
        rp = self.reduced_precision_dtype
        for pg in self.param_groups:
            self.param_groups_full_precision.append({"params": [
                p.detach().clone().to(dtype=torch.float32) if (rp is not None and p.dtype == rp) else None
                for p in pg["params"]]})

    def add_param_group(self, param_group):
        super().add_param_group(param_group)
        for name, default in self.defaults.items():
            if isinstance(default, torch.Tensor):
                self.param_groups[-1][name] = default.clone()

    @torch.no_grad()
    def step(self, closure=None, grad_scaler=None):
        loss = closure() if closure is not None else None
        # master copies are made at the first step, not in the constructor, so that a DDP
        # broadcast of the model parameters after construction is picked up (reference :146-150)
        if len(self.param_groups_full_precision) == 0:
            self._setup_full_precision_params()

        grad_list = []
        for group in self.param_groups:
            for p in group["params"]:
                assert group["params"][0].dtype == p.dtype, \
                    "Error: Parameters are not of the identical type: {} != {}".format(group["params"][0].dtype, p.dtype)
                if p.grad is not None:
                    grad_list.append(p.grad)

        device = self.param_groups[0]["params"][0].device
        if grad_scaler is not None:
            found_inf = grad_scaler._check_inf_per_device(self)[device]
            scale = grad_scaler._get_scale_async()
            inv_scale = scale.double().reciprocal().float()
        else:
            found_inf = torch.zeros((1,), device=device)
            scale = torch.ones((1,), device=device)
            inv_scale = torch.ones((1,), device=device)
        self._dummy_overflow_buf.copy_(found_inf)

        # the norm is taken over the still-scaled gradients, so the clip threshold is scaled too
        max_grad_norm = self.defaults["max_grad_norm"] * scale
        grad_norm = multi_tensor_applier(self.multi_tensor_l2norm, self._dummy_overflow_buf, [grad_list], False)[0]

        for group, group_full in zip(self.param_groups, self.param_groups_full_precision):
            beta1, beta2 = group["betas"]
            group["step"] += (self._dummy_overflow_buf != 1).to(torch.int)   # advance only on a clean step
            lists = [[], [], [], []]
            if self.reduced_precision_dtype is not None:
                lists.append([])
            for p, p_full in zip(group["params"], group_full["params"]):
                if p.grad is None:
                    continue
                assert not p.grad.is_sparse
                state = self.state[p]
                if len(state) == 0:
                    dt = torch.float32 if (self.reduced_precision_dtype is not None
                                           and p.dtype == self.reduced_precision_dtype) else p.dtype
                    state["exp_avg"] = torch.zeros_like(p, dtype=dt)
                    state["exp_avg_sq"] = torch.zeros_like(p, dtype=dt)
                lists[0].append(p.grad)
                lists[1].append(p_full if (self.reduced_precision_dtype is not None and p_full is not None) else p)
                lists[2].append(state["exp_avg"])
                lists[3].append(state["exp_avg_sq"])
                if self.reduced_precision_dtype is not None:
                    lists[4].append(p)
            if len(lists[0]) == 0:
                continue
            multi_tensor_applier(self.multi_tensor_lamb, self._dummy_overflow_buf, lists, group["lr"], beta1, beta2,
                                 group["eps"], group["step"], 1 if group["bias_correction"] else 0,
                                 group["weight_decay"], 1 if group["grad_averaging"] else 0, self.adam_w_mode,
                                 grad_norm, max_grad_norm, self.use_nvlamb, found_inf, inv_scale)
        return loss
