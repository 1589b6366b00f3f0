"""FusedAdagrad (reference: apex/optimizers/fused_adagrad.py:5-121)."""
import torch

from ..multi_tensor_apply import multi_tensor_applier
from . import _common


class FusedAdagrad(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-2, eps=1e-10, weight_decay=0., set_grad_none=True, adagrad_w_mode=False):
        super().__init__(params, dict(lr=lr, eps=eps, weight_decay=weight_decay))
        self.adagrad_w_mode = 1 if adagrad_w_mode else 0
        self.set_grad_none = set_grad_none
        amp_C = _common.require_extension("FusedAdagrad")
        self._dummy_overflow_buf = _common.overflow_buf()
        self.multi_tensor_adagrad = amp_C.multi_tensor_adagrad

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group["params"]:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            entries = []
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdagrad does not support sparse gradients")
                if p.dtype not in _common.SUPPORTED:
                    raise RuntimeError("FusedAdagrad only support fp16, bfloat16 and fp32.")
                state = self.state[p]
                if len(state) == 0:
                    state["sum"] = torch.zeros_like(p)
                entries.append((p.grad, p, state["sum"]))
            for bucket in _common.by_dtype(entries):
                lists = [list(col) for col in zip(*bucket)]
                multi_tensor_applier(self.multi_tensor_adagrad, self._dummy_overflow_buf, lists, group["lr"],
                                     group["eps"], self.adagrad_w_mode, group["weight_decay"])
        return loss
