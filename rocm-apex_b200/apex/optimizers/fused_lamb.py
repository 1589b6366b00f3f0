"""FusedLAMB (reference: apex/optimizers/fused_lamb.py:4-215): global gradient-norm clipping, then the
LAMB update with a per-tensor trust ratio, one fused call per dtype and parameter group."""
import torch

from ..multi_tensor_apply import multi_tensor_applier
from . import _common


class FusedLAMB(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-6,
                 weight_decay=0.01, amsgrad=False, adam_w_mode=True, grad_averaging=True,
                 set_grad_none=True, max_grad_norm=1.0, use_nvlamb=False):
        if amsgrad:
            raise RuntimeError("FusedLAMB does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, grad_averaging=grad_averaging,
                                      max_grad_norm=max_grad_norm))
        amp_C = _common.require_extension("FusedLAMB")
        self.multi_tensor_l2norm = amp_C.multi_tensor_l2norm
        self.multi_tensor_lamb = amp_C.multi_tensor_lamb
        self._dummy_overflow_buf = _common.overflow_buf(self.param_groups[0]["params"][0].device)
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none
        self.use_nvlamb = use_nvlamb

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group["params"]:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    def _global_grad_norm(self):
        """sqrt(||g32||^2 + ||g16||^2 ...): one l2norm per dtype, then the l2norm of those norms
        (apex/optimizers/fused_lamb.py:107-136)."""
        device = self.param_groups[0]["params"][0].device
        grads = {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.dtype not in _common.SUPPORTED:
                    raise RuntimeError("FusedLAMB only support fp16, bfloat16 and fp32.")
                grads.setdefault(p.dtype, []).append(p.grad)
        norms = [multi_tensor_applier(self.multi_tensor_l2norm, self._dummy_overflow_buf, [g], False)[0]
                 for g in grads.values()]
        while len(norms) < 2:
            norms.append(torch.zeros(1, device=device))
        return multi_tensor_applier(self.multi_tensor_l2norm, self._dummy_overflow_buf, [norms], False)[0]

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        global_grad_norm = self._global_grad_norm()
        max_grad_norm = self.defaults["max_grad_norm"]
        for group in self.param_groups:
            group["step"] = group.get("step", 0) + 1
            beta1, beta2 = group["betas"]
            entries = []
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedLAMB does not support sparse gradients, please consider SparseAdam instead")
                state = self.state[p]
                if len(state) == 0:
                    state["exp_avg"] = torch.zeros_like(p)
                    state["exp_avg_sq"] = torch.zeros_like(p)
                entries.append((p.grad, p, state["exp_avg"], state["exp_avg_sq"]))
            for bucket in _common.by_dtype(entries):
                lists = [list(col) for col in zip(*bucket)]
                multi_tensor_applier(self.multi_tensor_lamb, self._dummy_overflow_buf, lists, group["lr"],
                                     beta1, beta2, group["eps"], group["step"],
                                     1 if group["bias_correction"] else 0, group["weight_decay"],
                                     1 if group["grad_averaging"] else 0, self.adam_w_mode,
                                     global_grad_norm, max_grad_norm, self.use_nvlamb)
        return loss
