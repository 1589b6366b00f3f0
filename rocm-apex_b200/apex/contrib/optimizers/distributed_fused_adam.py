"""DistributedFusedAdam: ZeRO-style sharded AdamW for one NVSwitch domain.

Same public surface as the reference (apex/contrib/optimizers/distributed_fused_adam.py:8-636):
constructor keywords, set_global_scale / global_scale, complete_reductions(), step(), L2_grad_norm,
has_overflow / peek_overflow, strided_check_finite(), state_dict() = {step, fp32_p, fp32_m, fp32_v}
(this rank's shards).  The pipeline is the reference's --

    gradients --(predivide 1/W, flatten)--> flat grads --reduce-scatter--> grad shard
        --> shard-local fused Adam (distributed_adam_cuda.multi_tensor_fused_adam, writes fp32 p/m/v and the
            low-precision p_copy in one pass) --> all-gather of p_copy --> model parameters

-- but laid out for 8 B200 behind one NVSwitch instead of the reference's multi-node tree:
  * one flat index space, cut into K pipeline chunks x W rank shards: element i of shard r of chunk k
    sits at ((k*W + r)*shard + i) in BOTH the flat gradient buffer and the gathered-parameter buffer, so
    reduce_scatter_tensor / all_gather_into_tensor run zero-copy on contiguous slices (the reference needs
    a patched c10d with `no_copy=` for that, :370-371) and a parameter has the same offset in both;
  * one communicator (NVSwitch gives every peer full bandwidth; the reference's 4+ process groups per
    collective only help on a network with several rails) and one side stream: RS(k+1..) overlaps
    Adam(k), AG(k) overlaps Adam(k+1..);
  * gradients are flattened with ONE multi_tensor_scale launch and gathered parameters are scattered
    back with ONE maybe_cast_mt launch, whatever the number of parameters.
`dwu_group_size` keeps its meaning (shard inside groups of that many ranks, all-reduce the shard across groups: the
multi-node layout).  The kwargs that only steer the reference's process-group fan-out (dwu_num_rs_pg, dwu_num_ar_pg,
dwu_num_ag_pg, num_process_groups, ...) are accepted and ignored.
"""
import importlib
import math

import torch
import torch.distributed as dist

from ...multi_tensor_apply import multi_tensor_applier

ALIGN = 64          # parameter offsets are multiples of 64 elements (reference :165-173)
PAD_UNIT = 256      # total length is a multiple of 256*K*W (reference :182-183)


def build_layout(numels, world_size, num_chunks):
    """Offsets of every parameter in the flat space and the chunk/shard geometry.  Pure arithmetic
    (shared by the optimizer, the tests and the benchmark)."""
    offsets, off = [], 0
    for n in numels:
        offsets.append(off)
        off = (off + n + ALIGN - 1) // ALIGN * ALIGN
    unit = PAD_UNIT * num_chunks * world_size
    total = max(unit, (off + unit - 1) // unit * unit)
    chunk = total // num_chunks
    shard = chunk // world_size
    return {"offsets": offsets, "total": total, "chunk": chunk, "shard": shard,
            "num_chunks": num_chunks, "world_size": world_size}


def shard_fragments(layout, numels, rank):
    """For every pipeline chunk k: the list of (param index, offset inside the param, offset inside this
    rank's shard of chunk k, length) where a parameter overlaps the shard (reference :260-301)."""
    W, shard, chunk = layout["world_size"], layout["shard"], layout["chunk"]
    out = []
    for k in range(layout["num_chunks"]):
        lo = k * chunk + rank * shard
        hi = lo + shard
        frags = []
        for i, (off, n) in enumerate(zip(layout["offsets"], numels)):
            a, b = max(lo, off), min(hi, off + n)
            if a < b:
                frags.append((i, a - off, a - lo, b - a))
        out.append(frags)
    return out


def reshard_state_dicts(state_dicts, numels, new_world_size, new_num_chunks, old_num_chunks=None):
    """Checkpoint shards written by `state_dict()` on every rank of one job (list indexed by rank, each
    {step, fp32_p, fp32_m, fp32_v}: reference :598-616) -> the shards of a job with another world size
    and / or pipeline-chunk count.  The reference can only resume on the layout that wrote the files;
    here a parameter has the same flat offset in every layout (build_layout), so resharding is a
    re-slicing of the flat space.  Host-side tensor indexing only; one of p/m/v is in flight at a time.
    `numels`: parameter sizes in optimizer order; `old_num_chunks` defaults to what the shard sizes imply."""
    old_world = len(state_dicts)
    steps = {int(sd["step"]) for sd in state_dicts}
    if len(steps) != 1:
        raise ValueError("shards come from different steps: %s" % sorted(steps))
    shard_len = state_dicts[0]["fp32_p"].numel()
    if old_num_chunks is None and "layout" in state_dicts[0]:
        old_num_chunks = int(state_dicts[0]["layout"]["num_chunks"])
        if int(state_dicts[0]["layout"]["world_size"]) != old_world:
            raise ValueError("%d shards given, written by a job of %d ranks" % (old_world, state_dicts[0]["layout"]["world_size"]))
    if old_num_chunks is None:
        # K*shard == total/W and total is the smallest multiple of 256*K*W covering the parameters
        fits = [k for k in range(1, 257) if build_layout(numels, old_world, k)["shard"] * k == shard_len]
        if len(fits) != 1:
            raise ValueError("shard length %d fits pipeline-chunk counts %s for these parameters: pass old_num_chunks "
                             "(dwu_num_blocks * dwu_num_chunks of the job that wrote the shards)" % (shard_len, fits))
        old_num_chunks = fits[0]
    old = build_layout(numels, old_world, old_num_chunks)
    if old["shard"] * old_num_chunks != shard_len:
        raise ValueError("shard length %d does not match world size %d x %d chunks" % (shard_len, old_world, old_num_chunks))
    new = build_layout(numels, new_world_size, new_num_chunks)
    used = min(old["total"], new["total"])        # everything past the last parameter is padding (zeros)
    out = [{"step": min(steps), "layout": {"world_size": new_world_size, "num_chunks": new_num_chunks}}
           for _ in range(new_world_size)]
    for key in ("fp32_p", "fp32_m", "fp32_v"):
        full = torch.zeros(max(old["total"], new["total"]), dtype=torch.float32)
        for r, sd in enumerate(state_dicts):
            t = sd[key].detach().to("cpu", torch.float32).reshape(old_num_chunks, old["shard"])
            for k in range(old_num_chunks):
                lo = k * old["chunk"] + r * old["shard"]
                full[lo:lo + old["shard"]] = t[k]
        full[used:] = 0
        for r in range(new_world_size):
            pieces = [full[k * new["chunk"] + r * new["shard"]:k * new["chunk"] + (r + 1) * new["shard"]]
                      for k in range(new_num_chunks)]
            out[r][key] = torch.cat(pieces).clone()
        del full
    return out


class DistributedFusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 eps_inside_sqrt=False, weight_decay=0., max_grad_norm=0., amsgrad=False, flat_mt=False,
                 overlap_reductions=True, compute_L2_grad_norm=False, dwu_group_size=0, dwu_num_blocks=4,
                 dwu_num_chunks=4, dwu_num_rs_pg=1, dwu_num_ar_pg=4, dwu_num_ag_pg=0, predivide=True,
                 e5m2_allgather=False, do_not_flatten_model=False, step_supports_amp_scaling=True,
                 num_process_groups=1, current_process_group=None, process_group_id=0, process_group_size=0,
                 clip_grad_norm=True, model_parallel=False, zero_copy_params=False, fused_collectives=False):
        # fused_collectives (not in the reference): the flat gradient and gathered-parameter buffers are
        # symmetric-memory allocations bound to NVSwitch multicast objects, and the whole step is ONE kernel:
        # gradients are summed over the ranks inside the switch while they are loaded (multimem.ld_reduce) and
        # the updated low-precision parameters are replicated to every rank by the switch as they are stored
        # (multimem.st).  No reduce_scatter, no all_gather, no pipeline chunks; two cross-GPU barriers per step.
        # True picks the form by world size: with TWO ranks the kernel reads the peer's gradients and writes the peer's
        # parameters directly through peer-mapped memory ("p2p": 4 bytes per shard element and direction on the link,
        # where the multicast form moves 6 because the switch pulls the requester's own half too); from three ranks on the
        # in-switch reduction wins ("multimem").  Either name forces that form.
        if fused_collectives not in (False, True, "multimem", "p2p"):
            raise ValueError("fused_collectives must be False, True, 'multimem' or 'p2p'")
        self._fused = bool(fused_collectives)
        self._fused_form = fused_collectives if isinstance(fused_collectives, str) else None
        # zero_copy_params (not in the reference): re-point every model parameter's storage at its slot of
        # the all-gather output buffer, so the gathered parameters ARE the model parameters and the final
        # scatter-back pass (maybe_cast_mt, 4 B/param of HBM traffic) disappears.
        self._zero_copy_params = zero_copy_params
        if amsgrad:
            raise RuntimeError("DistributedFusedAdam does not support the AMSGrad variant.")
        # e5m2_allgather (reference :64,198,206): the step kernel writes the parameter copy as e5m2 bytes,
        # so the all-gather (or, fused, the multicast store) moves 1 byte per parameter instead of 2; the
        # final maybe_cast_mt launch unpacks the bytes into the model parameters.
        self._e5m2 = e5m2_allgather
        if e5m2_allgather and zero_copy_params:
            raise ValueError("zero_copy_params needs the gathered buffer in the parameter dtype; "
                             "e5m2_allgather gathers bytes")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, max_grad_norm=max_grad_norm))
        # kernels are looked up by module name, exactly as the reference does (:78-81)
        self._dist_adam = importlib.import_module("distributed_adam_cuda")
        self._contrib = importlib.import_module("fused_adam_cuda")
        amp_C = importlib.import_module("amp_C")
        self.multi_tensor_l2norm = amp_C.multi_tensor_l2norm
        self._multi_tensor_scale = amp_C.multi_tensor_scale

        self.eps_mode = 0 if eps_inside_sqrt else 1
        self._step_supports_amp_scaling = step_supports_amp_scaling
        self._predivide = predivide
        self._compute_L2_grad_norm = compute_L2_grad_norm
        self._clip_grad_norm = clip_grad_norm
        self._overlap_reductions = overlap_reductions
        self._num_chunks_total = max(1, dwu_num_blocks * dwu_num_chunks)   # K pipeline chunks
        if self._fused:
            self._num_chunks_total = 1          # one kernel, nothing to pipeline
            if compute_L2_grad_norm:
                raise NotImplementedError("fused_collectives never materialises the reduced gradients; "
                                          "compute the gradient norm before the step instead")
        self._global_scale = None
        self._has_overflow = False
        self._L2_grad_norm = None
        self._l2_sq = None
        self._step = 0
        self._init_done = False
        self._resume_state = None
        self._reduced = False
        self._last_step = False
        self._iter_started = False       # the overflow flag has been cleared for the current iteration

        self._pg = current_process_group if current_process_group is not None else dist.group.WORLD
        self._world_size = dist.get_world_size(self._pg)
        self._rank = dist.get_rank(self._pg)
        # dwu_group_size (reference :115-117,328-345,397-418): shard over groups of that many ranks (one node) and keep
        # the optimizer state replicated across groups; gradients are reduce-scattered inside a group, then the
        # shard is all-reduced across groups.  0 or >= the world: one group, the plain ZeRO layout.
        self._bcast_pg, self._grad_world, self._ar_pg = self._pg, self._world_size, None
        if 0 < dwu_group_size < self._world_size:
            if self._world_size % dwu_group_size != 0:
                raise ValueError("dwu_group_size %d does not divide the %d ranks" % (dwu_group_size, self._world_size))
            if self._fused:
                raise NotImplementedError("fused_collectives spans one NVSwitch domain: no rank groups")
            if current_process_group is not None:
                raise NotImplementedError("rank groups inside a caller-provided process group")   # new_group() is world-collective
            gs, ngroups, me = dwu_group_size, self._world_size // dwu_group_size, self._rank
            for j in range(ngroups):                       # every rank creates every group, in the same order
                g = dist.new_group(ranks=[j * gs + k for k in range(gs)])
                if me // gs == j:
                    self._pg = g
            for j in range(gs):
                g = dist.new_group(ranks=[j + k * gs for k in range(ngroups)])
                if me % gs == j:
                    self._ar_pg = g
            self._world_size, self._rank = gs, me % gs

    # ------------------------------------------------------------------------------------------
    # lazy construction of the flat buffers (first complete_reductions / step)
    # ------------------------------------------------------------------------------------------
    def _init_everything(self):
        if self._init_done:
            return
        self._model_params, self._group_of = [], []
        for gi, group in enumerate(self.param_groups):
            for p in group["params"]:
                if p.requires_grad:
                    self._model_params.append(p)
                    self._group_of.append(gi)
        assert len(self._model_params) > 0, "no parameters to optimise"
        if self._grad_world > 1:
            # every rank starts from the first rank's parameters (reference :144-145)
            src = dist.get_global_rank(self._bcast_pg, 0)
            for group in self.param_groups:
                for p in group["params"]:
                    dist.broadcast(p.data, src=src, group=self._bcast_pg)
        p0 = self._model_params[0]
        dev, low = p0.device, p0.dtype
        assert all(p.dtype == low and p.device == dev for p in self._model_params), \
            "all parameters must share one dtype and device"
        self._device, self._low = dev, low
        self._cuda = dev.type == "cuda"
        W, K, r = self._world_size, self._num_chunks_total, self._rank
        numels = [p.numel() for p in self._model_params]
        L = self._layout = build_layout(numels, W, K)
        total, chunk, shard = L["total"], L["chunk"], L["shard"]

        self._overflow_buf = torch.zeros(1, dtype=torch.int32, device=dev)
        gathered = torch.uint8 if self._e5m2 else low      # dtype that travels in the all-gather
        if self._fused:
            import torch.distributed._symmetric_memory as symm_mem
            assert self._cuda, "fused_collectives needs CUDA devices on one NVSwitch domain"
            self._flat_grads = symm_mem.empty(total, dtype=low, device=dev)
            self._new_params = symm_mem.empty(total, dtype=gathered, device=dev)
            self._flat_grads.zero_(); self._new_params.zero_()
            self._hdl_g = symm_mem.rendezvous(self._flat_grads, self._pg.group_name)
            self._hdl_p = symm_mem.rendezvous(self._new_params, self._pg.group_name)
            if self._fused_form is None:
                self._fused_form = "p2p" if (W == 2 and not self._e5m2 and low == torch.bfloat16) else "multimem"
            if self._fused_form == "p2p":
                if W != 2 or self._e5m2 or low != torch.bfloat16:
                    raise ValueError("fused_collectives='p2p' is the two-rank bf16 form")
                peer = 1 - r
                # byte distance from this rank's buffer to the peer's copy of it, both mapped into this process
                self._p2p_delta = (int(self._hdl_g.buffer_ptrs[peer]) - int(self._hdl_g.buffer_ptrs[r]),
                                   int(self._hdl_p.buffer_ptrs[peer]) - int(self._hdl_p.buffer_ptrs[r]))
            elif not (self._hdl_g.multicast_ptr and self._hdl_p.multicast_ptr):
                raise RuntimeError("fused_collectives: no multicast (NVLS) address for the symmetric buffers; "
                                   "this needs an NVSwitch domain with multicast support")
        else:
            self._flat_grads = torch.zeros(total, dtype=low, device=dev)      # RS input  [k][r][shard]
            self._new_params = torch.zeros(total, dtype=gathered, device=dev) # AG output [k][r][shard]
        nlow = 0 if self._fused else K * shard                            # not materialised when fused
        self._fp16_g = torch.zeros(nlow, dtype=low, device=dev)           # RS output, this rank
        self._fp16_p = torch.zeros(nlow, dtype=gathered, device=dev)      # Adam p_copy, AG input
        self._fp32_p = torch.zeros(K * shard, dtype=torch.float32, device=dev)
        self._fp32_m = torch.zeros(K * shard, dtype=torch.float32, device=dev)
        self._fp32_v = torch.zeros(K * shard, dtype=torch.float32, device=dev)

        # views of each parameter's slot in the two full-size buffers
        self._grad_views = [self._flat_grads[o:o + n].view_as(p) for o, n, p in zip(L["offsets"], numels, self._model_params)]
        self._param_views = [self._new_params[o:o + n].view_as(p) for o, n, p in zip(L["offsets"], numels, self._model_params)]

        # per-chunk kernel lists over the fragments of this rank's shard + per-fragment hyper-parameters
        self._chunk_lists, self._chunk_hyper = [], []
        for k, frags in enumerate(shard_fragments(L, numels, r)):
            base = k * shard
            lists = [[], [], [], [], []]
            b1, b2, bc, ep, wd = [], [], [], [], []
            for (pi, poff, soff, n) in frags:
                if self._fused:
                    # multicast accesses are whole 16-byte vectors of 8 elements: extend the fragment into the (zero, unowned)
                    # alignment padding that follows every parameter; g and p_copy are multicast addresses of
                    # this fragment's position in the flat space
                    from _b200mt import RawBuffer
                    n = (n + 7) // 8 * 8
                    flat_pos = k * chunk + r * shard + soff
                    esize = self._flat_grads.element_size()
                    for lst, buf in zip(lists[:3], (self._fp32_p, self._fp32_m, self._fp32_v)):
                        lst.append(buf[base + soff:base + soff + n])
                    if self._fused_form == "p2p":      # this rank's own views; the kernel adds the distance to the peer's copies
                        lists[3].append(self._flat_grads[flat_pos:flat_pos + n])
                        lists[4].append(self._new_params[flat_pos:flat_pos + n])
                    else:
                        lists[3].append(RawBuffer(self._hdl_g.multicast_ptr + flat_pos * esize, n, low, dev))
                        lists[4].append(RawBuffer(self._hdl_p.multicast_ptr + flat_pos * self._new_params.element_size(),
                                                  n, gathered, dev))
                else:
                    sl = slice(base + soff, base + soff + n)
                    for lst, buf in zip(lists, (self._fp32_p, self._fp32_m, self._fp32_v, self._fp16_g, self._fp16_p)):
                        lst.append(buf[sl])
                g = self.param_groups[self._group_of[pi]]
                b1.append(g["betas"][0]); b2.append(g["betas"][1]); bc.append(1 if g["bias_correction"] else 0)
                ep.append(g["eps"]); wd.append(g["weight_decay"])
            self._chunk_lists.append(lists)
            self._chunk_hyper.append((torch.tensor(b1, dtype=torch.float32, device=dev),
                                      torch.tensor(b2, dtype=torch.float32, device=dev),
                                      torch.tensor(bc, dtype=torch.int32, device=dev),
                                      torch.tensor(ep, dtype=torch.float32, device=dev),
                                      torch.tensor(wd, dtype=torch.float32, device=dev)))

        # master weights: this rank's slice of the model parameters (or a loaded checkpoint)
        if self._resume_state is not None:
            st = self._resume_state
            self._step = st["step"]
            self._fp32_p.copy_(st["fp32_p"]); self._fp32_m.copy_(st["fp32_m"]); self._fp32_v.copy_(st["fp32_v"])
            self._resume_state = None
        else:
            with torch.no_grad():
                for k, frags in enumerate(shard_fragments(L, numels, r)):
                    for (pi, poff, soff, n) in frags:
                        self._fp32_p[k * shard + soff:k * shard + soff + n].copy_(
                            self._model_params[pi].detach().reshape(-1)[poff:poff + n])
        if self._zero_copy_params:
            with torch.no_grad():
                for p, v in zip(self._model_params, self._param_views):
                    v.copy_(p.detach())
                    p.data = v
        self._ag_pg = self._pg
        if self._cuda:
            # high priority: a collective started from a gradient hook must get SMs while backward's kernels keep the GPU
            # full, otherwise it only advances in the gaps and nothing is hidden (tools/bench_overlap.py)
            self._comm_stream = torch.cuda.Stream(device=dev, priority=-1)
            self._ag_stream = self._comm_stream
            self._rs_done = [torch.cuda.Event() for _ in range(K)]
            self._adam_done = [torch.cuda.Event() for _ in range(K)]
            if (W > 1 and K > 1 and not self._fused
                    and dist.get_process_group_ranks(self._pg) == list(range(dist.get_world_size()))):
                # The all-gathers get their own communicator and stream (the reference does the same with its
                # dwu_num_ag_pg groups, :345-364): AG(k) then runs while RS(k-1..0) are still in flight -- the
                # reduce-scatter loads a GPU's outgoing links, the all-gather its incoming ones -- instead of queueing
                # behind them on one communicator.  Only when the optimizer spans the whole job: new_group() is
                # collective over the default group.
                self._ag_pg = dist.new_group(ranks=list(range(W)))
                self._ag_stream = torch.cuda.Stream(device=dev, priority=-1)
        # overlapped reductions (reference :158-164,381-395,489-502): a hook per parameter marks its gradient
        # as produced; pipeline chunks are flushed (flattened + reduce-scattered on the side stream) from the
        # LAST chunk down, each as soon as every parameter that touches it has its gradient -- backward
        # produces gradients roughly last parameter first, so the reductions hide behind the rest of backward.
        # Every rank flushes chunks in the same order, which keeps the collectives matched.
        self._chunk_params = []
        for k in range(K):
            lo, hi = k * chunk, (k + 1) * chunk
            self._chunk_params.append([i for i, (o, n) in enumerate(zip(L["offsets"], numels)) if o < hi and o + n > lo])
        self._grad_ready = [False] * len(numels)
        self._grad_flat = [False] * len(numels)       # already scaled into the flat buffer this iteration
        self._next_chunk = K - 1
        self._hooks = []
        if self._overlap_reductions and not self._fused:
            for i, p in enumerate(self._model_params):
                self._hooks.append(p.register_post_accumulate_grad_hook(lambda _p, i=i: self._on_grad(i)))
        self._init_done = True

    # ------------------------------------------------------------------------------------------
    # reference API
    # ------------------------------------------------------------------------------------------
    def set_global_scale(self, global_scale):
        """The loss scale the incoming gradients carry; the step kernel divides it out."""
        self._global_scale = global_scale

    @property
    def global_scale(self):
        return self._global_scale

    @property
    def has_overflow(self):
        """Overflow seen by strided_check_finite(); reading clears it (reference :509-516)."""
        has = self._has_overflow
        self._has_overflow = False
        return has

    @property
    def peek_overflow(self):
        return self._has_overflow

    def strided_check_finite(self, output_params, stride=1, start=-1, end=-1, clear=True):
        out_p = output_params[start:end] if (0 <= start < end) else output_params
        self._contrib.strided_check_finite(self._overflow_buf, out_p, stride, 1 if clear else 0)
        self._has_overflow = self._overflow_buf.item() != 0
        return self._has_overflow

    @property
    def L2_grad_norm(self):
        """Global gradient norm (of the still-scaled, averaged gradients); host value, syncs once."""
        if not self._compute_L2_grad_norm:
            return None
        if self._L2_grad_norm is None and self._l2_sq is not None:
            self._L2_grad_norm = math.sqrt(float(self._l2_sq.item()))
        return self._L2_grad_norm

    def attach_flat_grads(self):
        """Points every parameter's .grad at its slot of the flat gradient buffer, so that backward
        accumulates straight into the reduce-scatter input and complete_reductions() has nothing to
        flatten (the reference keeps such views in _individual_flat_grads, :209-211)."""
        self._init_everything()
        for p, v in zip(self._model_params, self._grad_views):
            p.grad = v
        return self._flat_grads

    def set_last_step(self, last_step):
        """While set, gradient hooks do not start reductions; complete_reductions() does all of them
        (reference :378-379,496,556-569)."""
        self._last_step = last_step

    def _reduce_scatter_chunk(self, k):
        shard, chunk = self._layout["shard"], self._layout["chunk"]
        dist.reduce_scatter_tensor(self._fp16_g[k * shard:(k + 1) * shard],
                                   self._flat_grads[k * chunk:(k + 1) * chunk], group=self._pg)
        if self._ar_pg is not None:         # the same shard of the other rank groups (reference :411-418)
            dist.all_reduce(self._fp16_g[k * shard:(k + 1) * shard], group=self._ar_pg)
        if self._cuda:
            self._rs_done[k].record(torch.cuda.current_stream(self._device))

    def _flush_chunks(self, chunks, final):
        """Flatten (predivide + copy, ONE launch) the not yet flattened gradients of every parameter touching
        `chunks`, then reduce-scatter those chunks on the communication stream.  `final`: parameters without a
        gradient are treated as unused (zero), as complete_reductions() does in the reference (:556-564)."""
        scale = 1.0 / self._grad_world if self._predivide else 1.0      # every rank of every group contributes
        if not self._iter_started:
            # first flush of this iteration: the flag of the flatten pass starts from zero, as the reference's
            # _flatten_grad_mt does (:481) -- otherwise one non-finite gradient would gate maybe_cast_mt in step()
            # for the rest of the run and the model parameters would silently stop following the masters
            self._overflow_buf.zero_()
            self._iter_started = True
        src, dst, inplace = [], [], []
        for k in chunks:
            for i in self._chunk_params[k]:
                if self._grad_flat[i]:
                    continue
                p, v = self._model_params[i], self._grad_views[i]
                if p.grad is None:
                    assert final
                    v.zero_()
                elif p.grad.data_ptr() != v.data_ptr():
                    src.append(p.grad); dst.append(v)
                elif scale != 1.0:
                    inplace.append(v)               # produced directly into the flat buffer (attach_flat_grads)
                self._grad_flat[i] = True
        if src or inplace:
            multi_tensor_applier(self._multi_tensor_scale, self._overflow_buf, [src + inplace, dst + inplace], scale)
        if self._fused:
            return                                  # the reduction happens inside the step kernel
        if self._cuda:
            self._comm_stream.wait_stream(torch.cuda.current_stream(self._device))
            with torch.cuda.stream(self._comm_stream):
                for k in chunks:
                    self._reduce_scatter_chunk(k)
        else:
            for k in chunks:
                self._reduce_scatter_chunk(k)

    def _on_grad(self, i):
        """Gradient hook of parameter i (reference _do_overlapped_reduction, :489-502)."""
        if getattr(self, "_is_accumulation_step", False):   # DistributedFusedLAMB: nothing leaves until the last micro-batch
            return
        self._grad_ready[i] = True
        if self._last_step or not self._overlap_reductions:
            return
        while self._next_chunk >= 0 and all(self._grad_ready[j] for j in self._chunk_params[self._next_chunk]):
            with torch.no_grad():
                self._flush_chunks([self._next_chunk], final=False)
            self._next_chunk -= 1

    def complete_reductions(self):
        """Flatten the gradients (predivided by the world size) and start the reduce-scatters of every
        pipeline chunk the gradient hooks have not already flushed (reference :552-575)."""
        self._init_everything()
        with torch.no_grad():
            self._flush_chunks(list(range(self._next_chunk, -1, -1)), final=True)
        self._next_chunk = self._num_chunks_total - 1
        self._iter_started = False
        self._grad_ready = [False] * len(self._grad_ready)
        self._grad_flat = [False] * len(self._grad_flat)
        self._L2_grad_norm, self._l2_sq = None, None
        if self._compute_L2_grad_norm:
            # the shard buffer is contiguous after the reductions: one norm + one scalar all-reduce (:421-441)
            if self._cuda:
                torch.cuda.current_stream(self._device).wait_stream(self._comm_stream)
            n = multi_tensor_applier(self.multi_tensor_l2norm, self._overflow_buf, [[self._fp16_g]], False)[0]
            self._l2_sq = n * n
            dist.all_reduce(self._l2_sq, group=self._pg)
        self._reduced = True

    def _combined_scale(self):
        """global_scale, enlarged when the (unscaled) gradient norm exceeds max_grad_norm (reference :443-453)."""
        gs = self._global_scale if self._global_scale is not None else 1.0
        mx = self.defaults["max_grad_norm"]
        if self._clip_grad_norm and mx > 0 and self._compute_L2_grad_norm:
            l2 = self.L2_grad_norm
            if l2 is not None and math.isfinite(l2):
                c = mx / (l2 / gs + 1e-6)
                return gs / min(1, c)
        return gs

    def _adam_chunk(self, k, combined_scale):
        lists = self._chunk_lists[k]
        if len(lists[0]) == 0:
            return
        b1, b2, bc, ep, wd = self._chunk_hyper[k]
        if self._fused and self._fused_form == "p2p":
            multi_tensor_applier(self._dist_adam.multi_tensor_fused_adam_p2p, self._overflow_buf, lists, b1, b2, bc, ep, wd,
                                 self.param_groups[0]["lr"], combined_scale, self._step, self.eps_mode, *self._p2p_delta)
            return
        op = self._dist_adam.multi_tensor_fused_adam_nvls if self._fused else self._dist_adam.multi_tensor_fused_adam
        multi_tensor_applier(op, self._overflow_buf, lists, b1, b2, bc, ep, wd,
                             self.param_groups[0]["lr"], combined_scale, self._step, self.eps_mode)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        if not self._reduced:
            self.complete_reductions()
        self._reduced = False
        K, shard, chunk = self._num_chunks_total, self._layout["shard"], self._layout["chunk"]
        combined_scale = self._combined_scale()
        self._step += 1
        if self._fused:
            self._hdl_g.barrier(channel=0)      # every rank's gradients are in its flat buffer
            self._adam_chunk(0, combined_scale) # load = in-switch reduce, store = in-switch broadcast
            self._hdl_p.barrier(channel=1)      # every rank's kernel is done: parameters complete everywhere,
                                                # and nobody is still reading this rank's gradients
        elif self._cuda:
            cur = torch.cuda.current_stream(self._device)
            order = range(K - 1, -1, -1)                        # the order the reduce-scatters were issued in
            for k in order:                                     # Adam(k) as soon as RS(k) has landed
                cur.wait_event(self._rs_done[k])
                self._adam_chunk(k, combined_scale)
                self._adam_done[k].record(cur)
            with torch.cuda.stream(self._ag_stream):            # AG(k) behind Adam(k), beside the remaining RS
                for k in order:
                    self._ag_stream.wait_event(self._adam_done[k])
                    dist.all_gather_into_tensor(self._new_params[k * chunk:(k + 1) * chunk],
                                                self._fp16_p[k * shard:(k + 1) * shard], group=self._ag_pg)
            cur.wait_stream(self._ag_stream)
            cur.wait_stream(self._comm_stream)
        else:
            for k in range(K):
                self._adam_chunk(k, combined_scale)
                dist.all_gather_into_tensor(self._new_params[k * chunk:(k + 1) * chunk],
                                            self._fp16_p[k * shard:(k + 1) * shard], group=self._pg)
        if not self._zero_copy_params:
            # gathered parameters -> model parameters, one launch (reference :584-591)
            multi_tensor_applier(self._contrib.maybe_cast_mt, self._overflow_buf,
                                 [self._param_views, [p.data for p in self._model_params]])
        return loss

    def state_dict(self):
        """This rank's shard: {step, fp32_p, fp32_m, fp32_v} (reference :598-616), plus a `layout` entry
        (world size, pipeline chunks) that load_state_dict ignores and reshard_state_dicts() reads."""
        self._init_everything()
        return {"step": self._step, "fp32_p": self._fp32_p, "fp32_m": self._fp32_m, "fp32_v": self._fp32_v,
                "layout": {"world_size": self._world_size, "num_chunks": self._num_chunks_total}}

    def load_state_dict(self, state_dict):
        if self._init_done:
            self._step = state_dict["step"]
            self._fp32_p.copy_(state_dict["fp32_p"]); self._fp32_m.copy_(state_dict["fp32_m"])
            self._fp32_v.copy_(state_dict["fp32_v"])
        else:
            self._resume_state = state_dict
