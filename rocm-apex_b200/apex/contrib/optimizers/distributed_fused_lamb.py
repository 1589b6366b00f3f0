"""DistributedFusedLAMB: ZeRO-style sharded LAMB for one NVSwitch domain.

Public surface of the reference (apex/contrib/optimizers/distributed_fused_lamb.py:8-910): constructor keywords,
set_global_scale / global_scale, set_is_accumulation_step, set_last_step, complete_reductions(), L2_grad_norm,
step(closure, grad_scaler), state_dict() = {step, fp32_p, fp32_m, fp32_v}.  The step is the reference's (:694-748) --

    grads --(1/W, flatten)--> reduce-scatter --> global ||g|| (scalar all-reduce) --> overflow flag
      --> compute_update_term on the shard (u, m, v; also leaves sum u^2 per fragment)
      --> per-parameter ||u||^2: all-reduce of one float per parameter --> update_weights (trust ratio) --> all-gather

-- and, like it, never synchronises the host: learning rate, step counter, loss scale, gradient norm and the
overflow flag are device tensors, every kernel returns at once when the flag is set, and the step counter only
advances on a clean step.  Layout, buffers and collectives are those of DistributedFusedAdam (one flat index
space, zero-copy reduce_scatter_tensor / all_gather_into_tensor); there is no chunk pipeline because the update
needs the global norms first.  Differences from the reference: the per-fragment update norms come out of the
update kernel itself instead of a separate pass over u, and a parameter split over several fragments of one
rank is summed (index_add_) where the reference's scatter_ would keep only the last fragment.
"""
import importlib

import torch
import torch.distributed as dist

from ...multi_tensor_apply import multi_tensor_applier
from .distributed_fused_adam import DistributedFusedAdam, shard_fragments


class DistributedFusedLAMB(DistributedFusedAdam):
    def __init__(self, params, lr=1e-3, bias_correction=True, grad_averaging=True, betas=(0.9, 0.999), eps=1e-8,
                 weight_decay=0., max_grad_norm=0., adam_w_mode=True, use_nvlamb=False, step_supports_amp_scaling=True,
                 overlap_reductions=True, dwu_group_size=0, dwu_num_blocks=4, dwu_num_chunks=4, dwu_num_rs_pg=1,
                 dwu_num_ar_pg=4, dwu_num_ag_pg=0, e5m2_allgather=False, verbose=False, clip_after_ar=True,
                 zero_copy_params=False):
        if not clip_after_ar:
            raise NotImplementedError("clip_after_ar=False (clipping before the reduction) is not part of this path")
        super().__init__(params, lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                         max_grad_norm=max_grad_norm, overlap_reductions=overlap_reductions,
                         dwu_group_size=dwu_group_size, dwu_num_blocks=1,
                         dwu_num_chunks=1, predivide=True, e5m2_allgather=e5m2_allgather,
                         step_supports_amp_scaling=step_supports_amp_scaling, zero_copy_params=zero_copy_params)
        for g in self.param_groups:
            g.setdefault("grad_averaging", grad_averaging)
        lamb = importlib.import_module("distributed_lamb_cuda")
        self.multi_tensor_lamb_compute_update_term = lamb.multi_tensor_lamb_compute_update_term
        self.multi_tensor_lamb_update_weights = lamb.multi_tensor_lamb_update_weights
        self._adam_w_mode = 1 if adam_w_mode else 0
        self._use_nvlamb = use_nvlamb
        self._is_accumulation_step = False
        self._last_step = False
        self._verbose = verbose
        self._lamb_ready = False

    # ---- reference API ------------------------------------------------------------------------
    def set_is_accumulation_step(self, is_accumulation_step):
        self._is_accumulation_step = is_accumulation_step

    def set_last_step(self, last_step):
        self._last_step = last_step

    def set_global_scale(self, global_scale):
        self._init_lamb()
        if torch.is_tensor(global_scale):
            self._global_scale_t.copy_(global_scale.reshape(1))
        else:
            self._global_scale_t.fill_(float(global_scale))
        self._global_scale = global_scale

    @property
    def L2_grad_norm(self):
        """Device tensor (fp32[1]): norm of the averaged, still-scaled gradients (reference :809-811)."""
        return self._L2_grad_norm

    # ---- set-up ---------------------------------------------------------------------------------
    def _init_lamb(self):
        if self._lamb_ready:
            return
        self._init_everything()
        dev, shard = self._device, self._layout["shard"]
        numels = [p.numel() for p in self._model_params]
        self._fp32_u = torch.zeros(shard, dtype=torch.float32, device=dev)
        frags = shard_fragments(self._layout, numels, self._rank)[0]
        self._frag_param = torch.tensor([pi for (pi, _, _, _) in frags], dtype=torch.int64, device=dev)
        lists1, lists2 = [[], [], [], [], []], [[], [], []]
        for (pi, poff, soff, n) in frags:
            sl = slice(soff, soff + n)
            for lst, buf in zip(lists1, (self._fp16_g, self._fp32_p, self._fp32_m, self._fp32_v, self._fp32_u)):
                lst.append(buf[sl])
            for lst, buf in zip(lists2, (self._fp32_u, self._fp32_p, self._fp16_p)):
                lst.append(buf[sl])
        self._stage1_lists, self._stage2_lists = lists1, lists2
        b1, b2, bc, ep, wd = self._chunk_hyper[0]
        self._hyper = dict(beta1=b1, beta2=b2, beta3=1.0 - b1, bias_correction=bc, eps=ep, decay=wd)
        self._frag_sumsq = torch.zeros(max(1, len(frags)), dtype=torch.float32, device=dev)
        self._lr = torch.tensor([float(self.param_groups[0]["lr"])], dtype=torch.float32, device=dev)
        self._step_t = torch.zeros(1, dtype=torch.int32, device=dev) + int(self._step)
        self._global_scale_t = torch.ones(1, dtype=torch.float32, device=dev)
        self._one = torch.ones(1, dtype=torch.int32, device=dev)
        # the all-gather input must hold the current parameters from the start: a first step that overflows
        # skips update_weights and gathers whatever is in this buffer
        if self._e5m2:      # bytes: pack, do not convert numerically
            multi_tensor_applier(self._contrib.maybe_cast_mt, torch.zeros_like(self._overflow_buf),
                                 [[self._fp32_p], [self._fp16_p]])
        else:
            self._fp16_p.copy_(self._fp32_p)
        self._L2_grad_norm = None
        self._lamb_ready = True

    # ---- reductions: RS + global gradient norm, all on the device ---------------------------------
    def complete_reductions(self):
        self._init_lamb()
        if self._is_accumulation_step:
            return
        compute = self._compute_L2_grad_norm
        self._compute_L2_grad_norm = False
        super().complete_reductions()           # flatten (1/W) + reduce-scatter
        self._compute_L2_grad_norm = compute
        if self._cuda:
            torch.cuda.current_stream(self._device).wait_stream(self._comm_stream)
        local = multi_tensor_applier(self.multi_tensor_l2norm, self._overflow_buf, [[self._fp16_g]], False)[0]
        sq = local * local
        dist.all_reduce(sq, group=self._pg)
        self._L2_grad_norm = sq.sqrt()

    def _adam_chunk(self, k, combined_scale):    # the Adam kernel of the parent class is not used here
        raise RuntimeError("DistributedFusedLAMB steps through compute_update_term / update_weights")

    @torch.no_grad()
    def step(self, closure=None, grad_scaler=None):
        loss = closure() if closure is not None else None
        if not self._reduced:
            self.complete_reductions()
        self._reduced = False
        chunk, shard = self._layout["chunk"], self._layout["shard"]
        gn = self._L2_grad_norm
        # overflow <=> the global norm is not finite (reference :700-708); identical on every rank after the all-reduce
        is_finite = (gn + 1 > gn).int()
        self._overflow_buf.copy_(self._one - is_finite)
        dist.all_reduce(self._overflow_buf, op=dist.ReduceOp.MAX, group=self._pg)
        self._step_t += (1 - self._overflow_buf)
        self._lr.fill_(float(self.param_groups[0]["lr"]))

        h = self._hyper
        max_grad_norm = self.defaults["max_grad_norm"]
        # per-parameter ||p|| from the model parameters every rank holds in full (reference :673-684)
        pnorm = multi_tensor_applier(self.multi_tensor_l2norm, self._overflow_buf,
                                     [[p.data for p in self._model_params]], True)[1]
        pnorm_frag = pnorm.index_select(0, self._frag_param) if self._frag_param.numel() else pnorm[:0]
        nparams = len(self._model_params)
        unorm_sq = torch.zeros(nparams, dtype=torch.float32, device=self._device)
        if len(self._stage1_lists[0]) > 0:
            self._frag_sumsq.zero_()
            multi_tensor_applier(self.multi_tensor_lamb_compute_update_term, self._overflow_buf, self._stage1_lists,
                                 h["beta1"], h["beta2"], h["beta3"], h["bias_correction"], self._step_t, h["eps"],
                                 self._adam_w_mode, h["decay"], self._global_scale_t, gn, max_grad_norm, self._frag_sumsq)
            unorm_sq.index_add_(0, self._frag_param, self._frag_sumsq[:self._frag_param.numel()])
        dist.all_reduce(unorm_sq, group=self._pg)       # one float per parameter (reference :686-692)
        unorm = unorm_sq.sqrt()
        if len(self._stage2_lists[0]) > 0:
            multi_tensor_applier(self.multi_tensor_lamb_update_weights, self._overflow_buf, self._stage2_lists,
                                 pnorm_frag, unorm, self._frag_param, self._lr, h["decay"], gn, self._use_nvlamb)
        dist.all_gather_into_tensor(self._new_params[0:chunk], self._fp16_p[0:shard], group=self._pg)
        if grad_scaler is not None:     # hand the flag to torch's GradScaler (reference :842-846)
            state = grad_scaler._per_optimizer_states[id(self)]
            state["found_inf_per_device"][self._device] = self._overflow_buf.float()
        if not self._zero_copy_params:
            # skipped on overflow: maybe_cast_mt is gated on the flag, like the reference's (:850-861)
            multi_tensor_applier(self._contrib.maybe_cast_mt, self._overflow_buf,
                                 [self._param_views, [p.data for p in self._model_params]])
        return loss

    def state_dict(self):
        self._init_lamb()
        return {"step": self._step_t.clone(), "fp32_p": self._fp32_p, "fp32_m": self._fp32_m, "fp32_v": self._fp32_v}

    def load_state_dict(self, state_dict):
        self._init_lamb()
        st = state_dict["step"]
        self._step_t.fill_(int(st.item()) if torch.is_tensor(st) else int(st))
        self._fp32_p.copy_(state_dict["fp32_p"]); self._fp32_m.copy_(state_dict["fp32_m"])
        self._fp32_v.copy_(state_dict["fp32_v"])
