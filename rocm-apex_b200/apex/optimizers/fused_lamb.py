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
        self._lists = _common.GroupCache("FusedLAMB", ("exp_avg", "exp_avg_sq"), grad_handles=True)
        self._zero_norm = torch.zeros(1, device=self.param_groups[0]["params"][0].device)   # pads the list of norms to two

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group["params"]:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    @staticmethod
    def _init_state(p, state):
        state["exp_avg"] = torch.zeros_like(p)
        state["exp_avg_sq"] = torch.zeros_like(p)

    def _global_grad_norm(self, per_group):
        """sqrt(||g32||^2 + ||g16||^2 ...): one l2norm per dtype, then the l2norm of those norms
        (apex/optimizers/fused_lamb.py:107-136).  A dtype that lives in one bucket -- one parameter group, the usual
        case -- is normed through the bucket's prepared gradient handle; a dtype spread over several groups is
        normed as one concatenated list, as the reference does."""
        by_dtype = {}
        for gi, sets in enumerate(per_group):
            for (_, grads), bucket in zip(sets, self._lists.groups[gi].buckets):
                by_dtype.setdefault(bucket.params[0].dtype, []).append((grads, bucket))
        norms = []
        for entries in by_dtype.values():
            if len(entries) == 1:
                grads = entries[0][0]
            else:
                grads = []
                for g, bucket in entries:
                    grads.extend(g if isinstance(g, list) else [r[0].grad for r in bucket.rows])
            norms.append(multi_tensor_applier(self.multi_tensor_l2norm, self._dummy_overflow_buf,
                                              grads if not isinstance(grads, list) else [grads], False)[0])
        while len(norms) < 2:
            norms.append(self._zero_norm)
        return multi_tensor_applier(self.multi_tensor_l2norm, self._dummy_overflow_buf, [norms], False)[0]

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        # [g, p, exp_avg, exp_avg_sq] per group and dtype: the reference's per-parameter loops (:107-136,149-190), memoised,
        # and handed over as prepared handles while the gradients keep their addresses (_common.GroupCache.launch_sets)
        per_group = [self._lists.launch_sets_with_grads(self, gi, group, self._init_state)
                     for gi, group in enumerate(self.param_groups)]
        global_grad_norm = self._global_grad_norm(per_group)
        max_grad_norm = self.defaults["max_grad_norm"]
        for group, sets in zip(self.param_groups, per_group):
            group["step"] = group.get("step", 0) + 1
            beta1, beta2 = group["betas"]
            for lists, _ in sets:
                multi_tensor_applier(self.multi_tensor_lamb, self._dummy_overflow_buf, lists, group["lr"],
                                     beta1, beta2, group["eps"], group["step"],
                                     1 if group["bias_correction"] else 0, group["weight_decay"],
                                     1 if group["grad_averaging"] else 0, self.adam_w_mode,
                                     global_grad_norm, max_grad_norm, self.use_nvlamb)
        return loss
