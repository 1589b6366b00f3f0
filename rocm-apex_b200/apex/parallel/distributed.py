"""apex.parallel.DistributedDataParallel / Reducer: the data-parallel gradient all-reduce that runs right
before the optimizer hot path (reference: apex/parallel/distributed.py:1-640).

Same constructor keywords and semantics as the reference (buckets of >= message_size elements per dtype in
gradient-arrival order, all-reduced on side streams while backward is still running; delay_allreduce;
gradient_average / gradient_predivide_factor; allreduce_always_fp32; retain_allreduce_buffers;
disable_allreduce()/enable_allreduce(); parameters broadcast from rank 0 at construction).

What is different is the bucket data path.  The reference does, per bucket (:426-476)
    flatten (cat) -> mul_(1/predivide) -> all_reduce -> mul_(predivide/world) -> unflatten copy
i.e. four extra read+write passes over the gradients and five or more launches.  Here a bucket owns a
persistent flat buffer and the path is
    ONE multi_tensor_scale launch  [grads] -> [views of the flat buffer], scale 1/predivide      (gather + predivide)
    all_reduce(flat buffer)
    ONE multi_tensor_scale launch  [views] -> [grads], scale predivide/world                     (average + scatter)
two passes, two launches, no allocation per iteration; both kernels take their tensor lists from the
device-resident descriptor table, which is uploaded once because the buffers never move.
"""
from collections import OrderedDict

import torch
import torch.distributed as dist
from torch.nn.modules import Module

from ..multi_tensor_apply import multi_tensor_applier


def _scale_op():
    import amp_C
    return amp_C.multi_tensor_scale


def split_by_type(tensors):
    buckets = OrderedDict()
    for t in tensors:
        buckets.setdefault((t.dtype, t.device), []).append(t)
    return buckets


def apply_flat_dist_call(bucket, call, extra_args=None):
    """Runs a collective on the concatenation of `bucket` (one dtype) and copies the result back
    (reference :36-47); all_reduce results are averaged over the world size."""
    flat = torch.cat([t.reshape(-1) for t in bucket])
    call(flat, *extra_args) if extra_args is not None else call(flat)
    if call is dist.all_reduce:
        flat /= dist.get_world_size()
    off = 0
    for t in bucket:
        t.copy_(flat[off:off + t.numel()].view_as(t))
        off += t.numel()


def flat_dist_call(tensors, call, extra_args=None):
    for bucket in split_by_type(tensors).values():
        apply_flat_dist_call(bucket, call, extra_args)


def extract_tensors(maybe_tensor, tensor_list):
    if torch.is_tensor(maybe_tensor):
        tensor_list.append(maybe_tensor)
    else:
        try:
            for item in maybe_tensor:
                extract_tensors(item, tensor_list)
        except TypeError:
            return


class Reducer(object):
    """Manual variant: nothing happens during backward, the user calls reduce() (reference :89-127).
    Constructing it from a module broadcasts the parameters from rank 0."""

    def __init__(self, module_or_grads_list):
        if isinstance(module_or_grads_list, Module):
            self.module = module_or_grads_list
            flat_dist_call([p.data for p in self.module.parameters()], dist.broadcast, (0,))
        else:
            self.module = None
            self.grads = []
            extract_tensors(module_or_grads_list, self.grads)

    def reduce(self):
        if self.module:
            grads = [p.grad.data for p in self.module.parameters() if p.grad is not None]
            flat_dist_call(grads, dist.all_reduce)
        else:
            flat_dist_call(self.grads, dist.all_reduce)


class _Bucket(object):
    """A fixed set of parameters of one dtype with a persistent flat buffer."""

    def __init__(self, params, comm_dtype):
        self.params = params
        self.numel = sum(p.numel() for p in params)
        dev = params[0].device
        self.flat = torch.zeros(self.numel, dtype=comm_dtype, device=dev)
        self.views, off = [], 0
        for p in params:
            self.views.append(self.flat[off:off + p.numel()].view_as(p))
            off += p.numel()
        self.pending = len(params)
        self.flag = torch.zeros(1, dtype=torch.int32, device=dev)


class DistributedDataParallel(Module):
    """Wraps a module; during backward the gradients are averaged over the ranks bucket by bucket,
    overlapped with the rest of backward (reference :129-640).  Iteration 1 records the order in which
    gradients arrive (rank 0's order is broadcast so every rank builds the same buckets) and reduces
    everything at the end of backward; from iteration 2 on a bucket is reduced as soon as its last gradient
    has been produced AND every earlier bucket has been launched: all ranks issue their all_reduce calls in
    bucket-index order whatever order their own gradients arrive in (reference comm_ready_buckets, :520-554).
    With allreduce_trigger_params the buckets are cut at those parameters in iteration 1 (reference :373-376)."""

    def __init__(self, module, message_size=10000000, delay_allreduce=False, shared_param=None,
                 allreduce_trigger_params=None, retain_allreduce_buffers=False, allreduce_always_fp32=False,
                 num_allreduce_streams=1, allreduce_communicators=None, gradient_average=True,
                 gradient_predivide_factor=1.0, gradient_average_split_factor=None, prof=False):
        super().__init__()
        if shared_param is not None:
            raise ValueError("shared_param is no longer supported as an option.  It was misleadingly named from the start.  "
                             "It turns out overlapping communication with computation should work fine with shared parameters.  "
                             "If you still wish to delay communication to the end of the backward pass, use delay_allreduce=True|False instead.")
        if num_allreduce_streams > 1 and delay_allreduce:
            raise ValueError("self.allreduce_different_streams may only be used if delay_allreduce=False.")
        if allreduce_trigger_params is not None and delay_allreduce:
            raise ValueError("Setting allreduce_trigger_params is only valid if delay_allreduce=False.")
        if not multi_tensor_applier.available:
            raise RuntimeError("apex.parallel.DistributedDataParallel requires the multi-tensor kernels (libb200mt.so)")
        self.module = module
        self.world_size = float(dist.get_world_size())
        self.message_size = message_size
        self.delay_allreduce = delay_allreduce
        self.retain_allreduce_buffers = retain_allreduce_buffers
        self.allreduce_always_fp32 = allreduce_always_fp32
        self.gradient_average = gradient_average
        self.gradient_predivide_factor = gradient_predivide_factor
        self.num_allreduce_streams = max(1, num_allreduce_streams)
        self.allreduce_communicators = allreduce_communicators
        self.trigger_ids = set(id(p) for p in allreduce_trigger_params) if allreduce_trigger_params is not None else None
        self.prof = prof
        self._disable_allreduce = False
        self.allreduce_buffers = []
        self._params = [p for p in self.module.parameters() if p.requires_grad]
        if dist.get_backend() == "nccl":
            for p in self.module.parameters():
                assert p.is_cuda, "NCCL backend only supports model parameters to be on GPU."
        self._cuda = len(self._params) > 0 and self._params[0].is_cuda
        self._index = {id(p): i for i, p in enumerate(self._params)}
        self._multi_tensor_scale = _scale_op()
        self._buckets = None            # fixed after the first iteration
        self._next_bucket = 0           # index of the next bucket that may be launched (in-order rule)
        self._ready = set()             # filled buckets waiting for an earlier one
        self._num_regular = 0           # buckets of parameters whose gradients showed up in iteration 1
        self._bucket_of = {}
        self._arrival = []
        self._callback_queued = False
        self._streams = [torch.cuda.Stream() for _ in range(self.num_allreduce_streams)] if self._cuda else []
        self._groups = None
        if self.num_allreduce_streams > 1:
            self._groups = (allreduce_communicators[0] if allreduce_communicators
                            else [dist.new_group() for _ in range(self.num_allreduce_streams)])
        for p in self._params:
            p.register_post_accumulate_grad_hook(self._make_hook(p))
        flat_dist_call([p.data for p in self.module.parameters()], dist.broadcast, (0,))

    # ---- user switches (reference :262-268) ----------------------------------------------------------
    def disable_allreduce(self):
        self._disable_allreduce = True

    def enable_allreduce(self):
        self._disable_allreduce = False

    def forward(self, *inputs, **kwargs):
        if not self._disable_allreduce and self._buckets is not None:
            for b in self._buckets:
                b.pending = len(b.params)
        self._callback_queued = False
        return self.module(*inputs, **kwargs)

    # ---- backward-time machinery -------------------------------------------------------------------------
    def _make_hook(self, param):
        def hook(*unused):
            if self._disable_allreduce:
                return
            if not self._callback_queued:
                torch.autograd.Variable._execution_engine.queue_callback(self._end_of_backward)
                self._callback_queued = True
            if self._buckets is None:
                self._arrival.append(self._index[id(param)])
                return
            if self.delay_allreduce:
                return
            b = self._bucket_of[id(param)]
            b.pending -= 1
            if b.pending == 0 and not b.launched and not b.straggler:
                # launch strictly in bucket order: a bucket that fills early waits for its predecessors
                self._ready.add(b.index)
                while self._next_bucket in self._ready:
                    self._ready.discard(self._next_bucket)
                    self._reduce_bucket(self._buckets[self._next_bucket])
                    self._next_bucket += 1
        return hook

    def _build_buckets(self):
        """Buckets in rank 0's gradient-arrival order, one dtype each, closed once they reach message_size
        elements (reference :340-404); parameters whose gradient did not show up in iteration 1 go last."""
        order = list(dict.fromkeys(self._arrival))
        narrived = len(order)
        order += [i for i in range(len(self._params)) if i not in set(order)]
        t = torch.tensor([narrived] + order, dtype=torch.int64, device=self._params[0].device)
        dist.broadcast(t, 0)
        narrived, order = int(t[0]), t[1:].tolist()
        trigger_idx = None
        if self.trigger_ids is not None:
            trigger_idx = set(i for i, p in enumerate(self._params) if id(p) in self.trigger_ids)

        def cut(indices):
            open_by_type, out = {}, []
            for i in indices:
                p = self._params[i]
                key = (p.dtype, p.device)
                cur = open_by_type.setdefault(key, [])
                cur.append(p)
                if self.delay_allreduce:
                    continue
                full = (i in trigger_idx) if trigger_idx is not None else (sum(q.numel() for q in cur) >= self.message_size)
                if full:
                    out.append(cur)
                    open_by_type[key] = []
            return out + [cur for cur in open_by_type.values() if cur]

        regular = cut(order[:narrived])
        # parameters whose gradient never showed up on rank 0 in iteration 1 (unused parameters) live in buckets of
        # their own at the end: those cannot fill during backward and are reduced, in order, by the epilogue
        stragglers = cut(order[narrived:])
        self._num_regular = len(regular)
        self._buckets = []
        for k, params in enumerate(regular + stragglers):
            comm_dtype = torch.float32 if self.allreduce_always_fp32 else params[0].dtype
            b = _Bucket(params, comm_dtype)
            b.index, b.launched, b.straggler = k, False, k >= len(regular)
            self._buckets.append(b)
            for p in params:
                self._bucket_of[id(p)] = b
        self.allreduce_buffers = [None] * len(self._buckets)

    def _reduce_bucket(self, b, on_main_stream=False):
        """gather+predivide (one launch) -> all_reduce -> average+scatter (one launch)."""
        b.launched = True
        live = [(p.grad, v, p) for p, v in zip(b.params, b.views) if p.grad is not None]
        if len(live) != len(b.params):
            for p, v in zip(b.params, b.views):
                if p.grad is None:
                    v.zero_()           # a parameter unused on this rank still takes part in the average
        stream = None
        if self._cuda and not on_main_stream:
            stream = self._streams[b.index % self.num_allreduce_streams]
            stream.wait_stream(torch.cuda.current_stream())
        ctx = torch.cuda.stream(stream) if stream is not None else _NullContext()
        with ctx:
            if live:
                multi_tensor_applier(self._multi_tensor_scale, b.flag, [[g for g, _, _ in live], [v for _, v, _ in live]],
                                     1.0 / self.gradient_predivide_factor)
            group = self._groups[b.index % self.num_allreduce_streams] if (self._groups and not on_main_stream) else None
            dist.all_reduce(b.flat, group=group)
            post = (self.gradient_predivide_factor / self.world_size) if self.gradient_average else 1.0
            if self.retain_allreduce_buffers:
                if post != 1.0:
                    b.flat.mul_(post)
                self.allreduce_buffers[b.index] = b.flat
            else:
                for p, v in zip(b.params, b.views):
                    if p.grad is None:
                        p.grad = torch.empty_like(p)
                multi_tensor_applier(self._multi_tensor_scale, b.flag, [b.views, [p.grad for p in b.params]], post)

    def _end_of_backward(self):
        """Queued on the autograd engine by the first gradient hook of a backward pass (reference :405-424)."""
        self._callback_queued = False
        if self._disable_allreduce:
            return
        first = self._buckets is None
        if first:
            self._build_buckets()
        if not first and not self.delay_allreduce and self._next_bucket != self._num_regular:
            # an overlapped bucket never filled on this rank: the other ranks have (or will) run its all_reduce, so
            # going on would hang or mix buckets -- same refusal as the reference's epilogue (:339-349)
            nb, self._next_bucket = self._next_bucket, 0
            self._ready.clear()
            for b in self._buckets:
                b.launched, b.pending = False, len(b.params)
            raise RuntimeError("In epilogue, next_bucket ({}) != num_buckets ({}).  This probably indicates some buckets "
                               "were not allreduced.".format(nb, self._num_regular))
        for b in self._buckets:       # iteration 1 / delayed mode: everything; otherwise the straggler buckets, in order
            if not b.launched:
                self._reduce_bucket(b, on_main_stream=self.delay_allreduce)
        if self._cuda:
            cur = torch.cuda.current_stream()
            for s in self._streams:
                cur.wait_stream(s)
        for b in self._buckets:
            b.launched = False
            b.pending = len(b.params)
        self._next_bucket = 0
        self._ready.clear()


class _NullContext(object):
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
