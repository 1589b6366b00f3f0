"""FusedNovoGrad (reference: apex/optimizers/fused_novograd.py:4-214): Adam-like first moment per element,
second moment one scalar per tensor (the blended gradient norm), kept per param group in `exp_avg_sq`
as [norms of the 16-bit tensors, norms of the fp32 tensors]."""
import torch

from ..multi_tensor_apply import multi_tensor_applier
from . import _common


class FusedNovoGrad(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.,
                 amsgrad=False, reg_inside_moment=False, grad_averaging=True, norm_type=2, init_zero=False,
                 set_grad_none=True):
        if amsgrad:
            raise RuntimeError("FusedNovoGrad does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, grad_averaging=grad_averaging, norm_type=norm_type,
                                      init_zero=init_zero))
        amp_C = _common.require_extension("FusedNovoGrad")
        self._dummy_overflow_buf = _common.overflow_buf(self.param_groups[0]["params"][0].device)
        self.multi_tensor_novograd = amp_C.multi_tensor_novograd
        self.moment_mode = 0 if reg_inside_moment else 1
        self.set_grad_none = set_grad_none

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group["params"]:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        for group in self.param_groups:   # the per-group norm vectors must follow the parameters' device
            if len(group["params"]) > 0 and "exp_avg_sq" in group:
                dev = group["params"][0].device
                group["exp_avg_sq"] = [t.to(dev) for t in group["exp_avg_sq"]]

    @staticmethod
    def _initial_norms(grads, norm_type, device):
        if norm_type == 0:
            vals = [g.float().abs().max() for g in grads]
        elif norm_type == 2:
            vals = [g.float().pow(2).sum().sqrt() for g in grads]
        else:
            raise RuntimeError("FusedNovoGrad only support l2/inf norm now.")
        return torch.stack(vals).float() if vals else torch.zeros(0, device=device)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            group["step"] = group.get("step", 0) + 1
            beta1, beta2 = group["betas"]
            low, full = [], []       # (g, p, m) of the 16-bit and of the fp32 parameters
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedNovoGrad does not support sparse gradients, please consider SparseAdam instead")
                state = self.state[p]
                if len(state) == 0:
                    state["exp_avg"] = torch.zeros_like(p)
                if p.dtype in (torch.float16, torch.bfloat16):
                    low.append((p.grad, p, state["exp_avg"]))
                elif p.dtype == torch.float32:
                    full.append((p.grad, p, state["exp_avg"]))
                else:
                    raise RuntimeError("FusedNovoGrad only support fp16, bfloat16 and fp32.")
            dev = group["params"][0].device
            if "exp_avg_sq" not in group:
                if group["init_zero"]:
                    group["exp_avg_sq"] = [torch.zeros(len(low), device=dev), torch.zeros(len(full), device=dev)]
                else:   # start from this step's norm, so that the first blend changes nothing
                    group["exp_avg_sq"] = [self._initial_norms([e[0] for e in low], group["norm_type"], dev),
                                           self._initial_norms([e[0] for e in full], group["norm_type"], dev)]
            else:
                assert len(low) == group["exp_avg_sq"][0].numel()
                assert len(full) == group["exp_avg_sq"][1].numel()
            for entries, norms in ((low, group["exp_avg_sq"][0]), (full, group["exp_avg_sq"][1])):
                if not entries:
                    continue
                lists = [list(col) for col in zip(*entries)]
                multi_tensor_applier(self.multi_tensor_novograd, self._dummy_overflow_buf, lists, norms, group["lr"],
                                     beta1, beta2, group["eps"], group["step"], 1 if group["bias_correction"] else 0,
                                     group["weight_decay"], 1 if group["grad_averaging"] else 0, self.moment_mode,
                                     group["norm_type"])
        return loss
