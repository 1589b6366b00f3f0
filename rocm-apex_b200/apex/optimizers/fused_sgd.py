"""FusedSGD (reference: apex/optimizers/fused_sgd.py:6-227).  With amp master weights it consumes the
still-scaled gradients directly -- the unscale rides in the kernel as `1/most_recent_scale` -- and
writes the fp16 model weights in the same pass."""
import collections

import torch
from torch.optim.optimizer import required

from ..multi_tensor_apply import multi_tensor_applier
from . import _common

# one parameter dtype of a memoised group: rows of (param, state_dict, momentum, momentum) as csrc/packer.cpp: check_rows walks them
_Bucket = collections.namedtuple("_Bucket", "rows params momentums handles")


class FusedSGD(torch.optim.Optimizer):
    def __init__(self, params, lr=required, momentum=0, dampening=0, weight_decay=0, nesterov=False,
                 wd_after_momentum=False, materialize_master_grads=True, set_grad_none=False):
        if lr is not required and lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if momentum < 0.0:
            raise ValueError("Invalid momentum value: {}".format(momentum))
        if weight_decay < 0.0:
            raise ValueError("Invalid weight_decay value: {}".format(weight_decay))
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening,
                                      weight_decay=weight_decay, nesterov=nesterov))
        self.wd_after_momentum = wd_after_momentum
        self.materialize_master_grads = materialize_master_grads
        self.most_recent_scale = 1.0
        self.scale_set_by_backward = False
        self.set_grad_none = set_grad_none
        amp_C = _common.require_extension("FusedSGD")
        self._dummy_overflow_buf = _common.overflow_buf(self.param_groups[0]["params"][0].device)
        self.multi_tensor_sgd = amp_C.multi_tensor_sgd
        self._sets = {}

    def __setstate__(self, state):
        super().__setstate__(state)
        self._sets = {}
        for group in self.param_groups:
            group.setdefault("nesterov", False)

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group["params"]:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    def get_momentums(self, params):
        """Momentum buffers for `params`, created on first use.  first_run is decided by the last
        parameter, like the reference (apex/optimizers/fused_sgd.py:121-136)."""
        bufs, first_run = [], True
        for p in params:
            st = self.state[p]
            first_run = "momentum_buffer" not in st
            if first_run:
                st["momentum_buffer"] = torch.zeros_like(p)
            bufs.append(st["momentum_buffer"])
        return bufs, first_run

    def _launch_sets(self, gid, group):
        stash = getattr(self, "_amp_stash", None)
        if stash is not None and hasattr(stash, "fp32_from_fp16_groups"):
            # amp O2: explicit fp32 master copies of fp16 model params
            masters, models = stash.fp32_from_fp16_groups[gid], stash.fp16_groups[gid]
            fp32_params = [p for p in stash.fp32_from_fp32_groups[gid] if p.grad is not None]
            fp32_set = [[p.grad for p in fp32_params], fp32_params]
            if self.materialize_master_grads:
                keep = [i for i, mp in enumerate(masters) if mp.grad is not None]
                grads = [masters[i].grad for i in keep]          # fp32 master grads
            else:
                keep = [i for i, mp in enumerate(models) if mp.grad is not None]
                grads = [models[i].grad for i in keep]           # fp16 model grads, still scaled
            mparams = [masters[i] for i in keep]
            moms16, first16 = self.get_momentums(mparams)
            moms32, first32 = self.get_momentums(fp32_params)
            return [([grads, mparams, moms16, [models[i] for i in keep]], first16),
                    (fp32_set + [moms32], first32)]
        # The reference walks every parameter three times per step here (:182-199); once the momentum buffers exist
        # the lists are memoised and a step only re-reads p.grad and checks by identity that nothing moved
        # (see _common.GroupCache for the rules).
        c = self._sets.get(gid)
        if c is not None and c.state is self.state and c.params is group["params"] and c.nparams == len(group["params"]):
            sets = self._memoised_sets(c)
            if sets is not None:
                return sets
        sets, rows_by_dt, all_warm = [], [], True
        for dt in (torch.float16, torch.float32):
            ps = [p for p in group["params"] if p.dtype == dt and p.grad is not None]
            moms, first = self.get_momentums(ps)
            all_warm = all_warm and not (first and ps)
            sets.append(([[p.grad for p in ps], ps, moms], first))
            rows_by_dt.append(_Bucket([(p, self.state[p], m, m) for p, m in zip(ps, moms)], ps, moms, {}))
        if all_warm:
            idle = [p for p in group["params"] if p.grad is None]
            self._sets[gid] = _common.Group(self.state, group["params"], len(group["params"]), idle, rows_by_dt)
        else:
            self._sets.pop(gid, None)
        return sets

    @staticmethod
    def _memoised_sets(c):
        """The cached [g, p, momentum] sets of one group, each as a prepared list handle when its gradients still sit
        where they sat last step (one native loop checks addresses and the identity of the momentum buffers; see
        _common.GroupCache.launch_sets), else as plain lists."""
        import _b200mt as mt
        for p in c.idle:
            if p.grad is not None:
                return None
        sets = []
        for rows, ps, moms, h in c.buckets:
            if not rows:
                sets.append(([[], ps, moms], False))
                continue
            if mt.check_rows is not None and h.get("prep") is not None:
                rc = mt.check_rows(rows, "momentum_buffer", "momentum_buffer", h["prep"].ptrs.ctypes.data, h["code"])
                if rc == 0:
                    h["misses"] = 0
                    sets.append((h["prep"], False))
                    continue
                if rc == 2:
                    return None
            gs = []
            add = gs.append
            for p, st, mom, _ in rows:
                g = p.grad
                if g is None or st.get("momentum_buffer") is not mom:
                    return None
                add(g)
            lists = [gs, ps, moms]
            if mt.check_rows is not None:
                if h.get("prep") is not None:
                    h["prep"].release()
                    h["prep"] = None
                    h["misses"] = h.get("misses", 0) + 1
                if h.get("misses", 0) < 3 and not any(g.is_sparse for g in gs):
                    try:
                        h["prep"] = mt.prepare(lists, keep=False)     # gradient addresses are re-checked before every use
                        h["code"] = mt.dtype_code(gs[0])
                        sets.append((h["prep"], False))
                        continue
                    except RuntimeError:
                        pass                                          # the plain call reports what is wrong with the lists
            sets.append((lists, False))
        return sets

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        for gid, group in enumerate(self.param_groups):
            for lists, first_run in self._launch_sets(gid, group):
                if not hasattr(lists, "handle"):                      # plain lists (a prepared handle was validated when made)
                    assert len(lists[0]) == len(lists[1]) == len(lists[2])
                    if len(lists[0]) == 0:
                        continue
                multi_tensor_applier(self.multi_tensor_sgd, self._dummy_overflow_buf, lists,
                                     group["weight_decay"], group["momentum"], group["dampening"], group["lr"],
                                     group["nesterov"], first_run, self.wd_after_momentum,
                                     1.0 / self.most_recent_scale)
        self.most_recent_scale = 1.0
        self.scale_set_by_backward = False
        return loss
