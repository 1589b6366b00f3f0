"""FusedAdam (reference: apex/optimizers/fused_adam.py:4-173): Adam / AdamW over all parameters of a
group in one multi-tensor launch per dtype."""
import torch

from ..multi_tensor_apply import multi_tensor_applier
from . import _common


class FusedAdam(torch.optim.Optimizer):
    """Same constructor and state (`exp_avg`, `exp_avg_sq`, per-group `step`) as the reference.
    `adam_w_mode=True` is decoupled weight decay (AdamW), False is L2 regularisation."""

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 adam_w_mode=True, weight_decay=0., amsgrad=False, set_grad_none=True):
        if amsgrad:
            raise RuntimeError("FusedAdam does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay))
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none
        amp_C = _common.require_extension("FusedAdam")
        self._dummy_overflow_buf = _common.overflow_buf()
        self.multi_tensor_adam = amp_C.multi_tensor_adam
        self._lists = _common.GroupCache("FusedAdam", ("exp_avg", "exp_avg_sq"))

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group["params"]:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    @torch.no_grad()
    def step(self, closure=None, grads=None, output_params=None, scale=None, grad_norms=None):
        if any(a is not None for a in (grads, output_params, scale, grad_norms)):
            raise RuntimeError("FusedAdam has been updated.  Simply initialize it identically to "
                               "torch.optim.Adam, and call step() with no arguments.")
        loss = closure() if closure is not None else None
        for gi, group in enumerate(self.param_groups):
            group["step"] = group.get("step", 0) + 1      # one counter per group, as in the reference
            beta1, beta2 = group["betas"]
            # [g, p, exp_avg, exp_avg_sq] per dtype: the reference's per-parameter loop (:118-160), run once and memoised
            for lists in self._lists.launch_sets(self, gi, group, self._init_state):
                multi_tensor_applier(self.multi_tensor_adam, self._dummy_overflow_buf, lists, group["lr"],
                                     beta1, beta2, group["eps"], group["step"], self.adam_w_mode,
                                     1 if group["bias_correction"] else 0, group["weight_decay"])
        return loss

    @staticmethod
    def _init_state(p, state):
        state["exp_avg"] = torch.zeros_like(p)
        state["exp_avg_sq"] = torch.zeros_like(p)
