"""Helpers shared by the fused optimizers of this mirror."""
import collections

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


# What GroupCache remembers of one param group, and of one parameter dtype inside it.  `rows` is a list of
# (param, state_dict, state0_tensor, state1_tensor) tuples -- the layout csrc/packer.cpp: check_rows walks -- and `handles` a dict
# with the prepared list handles ("prep": [g, p, s0, s1]; "prep_g": [g] alone), the gradient dtype code they were made for and
# the number of consecutive steps the gradients had moved.
Group = collections.namedtuple("Group", "state params nparams idle buckets")
Bucket = collections.namedtuple("Bucket", "rows params state0 state1 handles")


class GroupCache(object):
    """Per-param-group memo of the tensor lists a fused optimizer hands to its kernel.

    The reference rebuilds [grads, params, state...] with a Python loop over every parameter on every step
    (apex/optimizers/fused_adam.py:118-160): ~0.7 us per parameter, which on a ResNet-50-sized list is as long as the
    GPU work itself.  Here the slow pass runs once; afterwards a step only re-reads `p.grad` and checks, by identity,
    that nothing it relied on has moved: the optimizer's `state` mapping, the group's `params` list, every cached state
    tensor (`state[p][name] is cached`), and which parameters have a gradient at all.  Any mismatch falls back to the
    slow pass, which is the reference's loop with its checks and error messages."""

    def __init__(self, optimizer_name, state_names, grad_handles=False):
        self.name = optimizer_name
        self.names = tuple(state_names)
        assert len(self.names) == 2, "specialised for two state tensors per parameter"
        self.groups = {}
        self.grad_handles = grad_handles      # also keep a depth-1 handle over the gradients alone (FusedLAMB's norm)

    def clear(self):
        for c in self.groups.values():
            for b in c.buckets:
                for k in ("prep", "prep_g"):
                    if b.handles.get(k) is not None:
                        b.handles[k].release()
        self.groups = {}

    def buckets(self, opt, gi, group, init_state):
        """-> list of (grads, params, state0, state1) lists, one entry per parameter dtype (fp32 first)."""
        c = self.groups.get(gi)
        if c is not None and c.state is opt.state and c.params is group["params"] and c.nparams == len(group["params"]):
            out = self._fast(c)
            if out is not None:
                return out
        return self._slow(opt, gi, group, init_state)

    def launch_sets(self, opt, gi, group, init_state):
        """What step() hands to the kernel, one entry per parameter dtype: the prepared handle (amp_C.prepare) of the
        bucket's [grads, params, state0, state1] when every gradient still sits at the address it had when the handle was
        made -- one C loop checks that, together with the identity of the state tensors -- else the four lists.
        Gradients re-allocated every iteration usually come back at the same addresses (caching allocator), so the
        handle survives `zero_grad(set_to_none=True)`; a bucket that keeps moving stops being prepared after three misses."""
        import _b200mt as mt
        if mt.check_rows is None:
            return [list(b) for b in self.buckets(opt, gi, group, init_state)]
        n0, n1 = self.names
        c = self.groups.get(gi)
        if c is not None and c.state is opt.state and c.params is group["params"] and c.nparams == len(group["params"]) \
                and all(p.grad is None for p in c.idle):
            out = []
            for bucket in c.buckets:
                h = bucket.handles
                if h.get("prep") is not None:
                    rc = mt.check_rows(bucket.rows, n0, n1, h["prep"].ptrs.ctypes.data, h["code"])
                    if rc == 0:
                        h["misses"] = 0
                        out.append(h["prep"])
                        continue
                    if rc == 2:
                        out = None
                        break
                gs = []
                for p, st, t0, t1 in bucket.rows:
                    g = p.grad
                    if g is None or st[n0] is not t0 or st[n1] is not t1:
                        gs = None
                        break
                    gs.append(g)
                if gs is None:
                    out = None
                    break
                out.append(self._maybe_prepare(bucket, gs))
            if out is not None:
                return out
        self._slow(opt, gi, group, init_state)
        return [self._maybe_prepare(b, [r[0].grad for r in b.rows]) for b in self.groups[gi].buckets]

    def _maybe_prepare(self, bucket, gs):
        import _b200mt as mt
        rows, ps, s0, s1, h = bucket
        lists = [gs, ps, s0, s1]
        if h.get("prep") is not None:
            h["prep"].release()
            h["prep"] = None
            h["misses"] = h.get("misses", 0) + 1
        if h.get("prep_g") is not None:
            h["prep_g"].release()
            h["prep_g"] = None
        if h.get("misses", 0) >= 3 or any(g.is_sparse for g in gs):
            return lists
        try:
            h["prep"] = mt.prepare(lists, keep=False)      # the gradients are not held: their addresses are re-checked
            h["code"] = mt.dtype_code(gs[0])
            if self.grad_handles:
                h["prep_g"] = mt.prepare([gs], keep=False)  # same addresses, so the same check covers it
            return h["prep"]
        except RuntimeError:
            return lists                                   # let the plain call report whatever is wrong with the lists

    def launch_sets_with_grads(self, opt, gi, group, init_state):
        """launch_sets() plus, per entry, what a norm over the bucket's gradients alone should be handed: the depth-1
        handle that was prepared together with the four-list one, else the list of gradients.  -> [(set, grads)]"""
        sets = self.launch_sets(opt, gi, group, init_state)
        out = []
        for s, bucket in zip(sets, self.groups[gi].buckets):
            if isinstance(s, list):
                out.append((s, s[0]))
            else:
                g = bucket.handles.get("prep_g")
                out.append((s, g if g is not None else [r[0].grad for r in bucket.rows]))
        return out

    def _fast(self, c):
        n0, n1 = self.names
        for p in c.idle:                     # parameters that had no gradient when the cache was built
            if p.grad is not None:
                return None
        out = []
        for rows, ps, s0, s1, _ in c.buckets:
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
            cached.append(Bucket(rows, ps, s0, s1, {}))
            out.append(([r[0].grad for r in rows], ps, s0, s1))
        self.groups[gi] = Group(opt.state, group["params"], len(group["params"]), idle, cached)
        return out


def by_dtype(entries, key=1):
    """Groups (grad, param, ...) tuples by the dtype of member `key` (the parameter), fp32 first.  One launch
    per dtype: the reference lumps fp16 and bf16 into one list and dispatches on its first tensor
    (apex/optimizers/fused_adam.py:134-138), which mistypes mixed lists; keeping them apart does not."""
    out = {}
    for e in entries:
        out.setdefault(e[key].dtype, []).append(e)
    return [out[d] for d in SUPPORTED if d in out]
