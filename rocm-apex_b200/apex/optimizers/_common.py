"""Helpers shared by the fused optimizers of this mirror."""
import torch

from ..multi_tensor_apply import multi_tensor_applier

SUPPORTED = (torch.float32, torch.float16, torch.bfloat16)


def require_extension(name):
    if not multi_tensor_applier.available:
        raise RuntimeError("apex.optimizers.%s requires cuda extensions" % name)
    import amp_C
    return amp_C


def overflow_buf(device=None):
    return torch.zeros(1, dtype=torch.int32, device=device if device is not None else "cuda")


class GroupCache(object):
    """Per-param-group memo of the tensor lists a fused optimizer hands to its kernel.

    The reference rebuilds [grads, params, state...] with a Python loop over every parameter on every step
    (apex/optimizers/fused_adam.py:118-160): ~0.7 us per parameter, which on a ResNet-50-sized list is as long as the
    GPU work itself.  Here the slow pass runs once; afterwards a step only re-reads `p.grad` and checks, by identity,
    that nothing it relied on has moved: the optimizer's `state` mapping, the group's `params` list, every cached state
    tensor (`state[p][name] is cached`), and which parameters have a gradient at all.  Any mismatch falls back to the
    slow pass, which is the reference's loop with its checks and error messages."""

    def __init__(self, optimizer_name, state_names):
        self.name = optimizer_name
        self.names = tuple(state_names)
        assert len(self.names) == 2, "specialised for two state tensors per parameter"
        self.groups = {}

    def clear(self):
        self.groups = {}

    def buckets(self, opt, gi, group, init_state):
        """-> list of (grads, params, state0, state1) lists, one entry per parameter dtype (fp32 first)."""
        c = self.groups.get(gi)
        if c is not None and c[0] is opt.state and c[1] is group["params"] and c[2] == len(group["params"]):
            out = self._fast(c)
            if out is not None:
                return out
        return self._slow(opt, gi, group, init_state)

    def _fast(self, c):
        n0, n1 = self.names
        for p in c[3]:                       # parameters that had no gradient when the cache was built
            if p.grad is not None:
                return None
        out = []
        for rows, ps, s0, s1 in c[4]:
            gs = []
            add = gs.append
            for p, st, t0, t1 in rows:
                g = p.grad
                if g is None or st[n0] is not t0 or st[n1] is not t1:
                    return None
                add(g)
            out.append((gs, ps, s0, s1))
        return out

    def _slow(self, opt, gi, group, init_state):
        n0, n1 = self.names
        idle, by = [], {}
        for p in group["params"]:
            if p.grad is None:
                idle.append(p)
                continue
            if p.grad.is_sparse:
                raise RuntimeError("%s does not support sparse gradients, please consider SparseAdam instead" % self.name)
            if p.dtype not in SUPPORTED:
                raise RuntimeError("%s only support fp16, bfloat16 and fp32." % self.name)
            st = opt.state[p]
            if len(st) == 0:
                init_state(p, st)
            by.setdefault(p.dtype, []).append((p, st, st[n0], st[n1]))
        cached, out = [], []
        for dt in SUPPORTED:
            rows = by.get(dt)
            if not rows:
                continue
            ps, s0, s1 = [r[0] for r in rows], [r[2] for r in rows], [r[3] for r in rows]
            cached.append((rows, ps, s0, s1))
            out.append(([r[0].grad for r in rows], ps, s0, s1))
        self.groups[gi] = (opt.state, group["params"], len(group["params"]), idle, cached)
        return out


def by_dtype(entries, key=1):
    """Groups (grad, param, ...) tuples by the dtype of member `key` (the parameter), fp32 first.  One launch
    per dtype: the reference lumps fp16 and bf16 into one list and dispatches on its first tensor
    (apex/optimizers/fused_adam.py:134-138), which mistypes mixed lists; keeping them apart does not."""
    out = {}
    for e in entries:
        out.setdefault(e[key].dtype, []).append(e)
    return [out[d] for d in SUPPORTED if d in out]
