"""LossScaler (reference: apex/amp/scaler.py:42-226): unscales gradients with the fused
multi_tensor_scale / multi_tensor_axpby kernels, which raise a device-side overflow flag; the flag is
read back once per iteration in update_scale() -- the only host<->device sync on the path."""
import torch

from ..multi_tensor_apply import multi_tensor_applier
from ._amp_state import _amp_state, maybe_print


class LossScaler(object):
    warned_no_fused_kernel = False
    warned_unscaling_non_fp32_grad = False
    has_fused_kernel = False

    def __init__(self, loss_scale, init_scale=2. ** 16, scale_factor=2., scale_window=2000,
                 min_loss_scale=None, max_loss_scale=2. ** 24):
        self.dynamic = (loss_scale == "dynamic")
        self._loss_scale = min(max_loss_scale, init_scale) if self.dynamic else loss_scale
        self._max_loss_scale = max_loss_scale
        self._min_loss_scale = min_loss_scale
        self._scale_seq_len = scale_window
        self._unskipped = 0
        self._has_overflow = False
        # no python fallback here: this build has no CPU path, the fused kernels are mandatory
        if not multi_tensor_applier.available:
            raise RuntimeError("apex.amp.LossScaler requires the multi-tensor kernels (libb200mt.so): "
                               + repr(getattr(multi_tensor_applier, "import_err", None)))
        import amp_C
        self._overflow_buf = torch.zeros(1, dtype=torch.int32, device="cuda")
        LossScaler.has_fused_kernel = True
        LossScaler.multi_tensor_scale_cuda = amp_C.multi_tensor_scale
        LossScaler.multi_tensor_axpby_cuda = amp_C.multi_tensor_axpby

    def loss_scale(self):
        return self._loss_scale

    def unscale(self, model_grads, master_grads, unused_scale, models_are_masters=False, scale_override=None):
        """master = model / scale, overflow flag raised for non-finite model grads (reference :103-128)."""
        if self._has_overflow:
            return
        scale = self._loss_scale if scale_override is None else scale_override
        if scale == 1.0 and models_are_masters and not self.dynamic:
            return
        multi_tensor_applier(LossScaler.multi_tensor_scale_cuda, self._overflow_buf,
                             [model_grads, master_grads], 1. / scale)

    def unscale_with_stashed(self, model_grads, stashed_master_grads, master_grads, scale_override=None):
        """master = model*(out/grads_scale) + stashed*(out/stashed_scale): gradient accumulation across
        backward passes; only the incoming model grads are checked for inf/NaN (reference :161-193)."""
        if self._has_overflow:
            return
        grads_have_scale, stashed_have_scale, out_scale = self._loss_scale, 1.0, 1.0
        if scale_override is not None:
            grads_have_scale, stashed_have_scale, out_scale = scale_override
        if not LossScaler.warned_unscaling_non_fp32_grad and master_grads[0].dtype == torch.float16:
            print("Warning:  unscaling grads that are not FP32. Unscaling non-fp32 grads may indicate an error. "
                  "When using Amp, you don't need to call .half() on your model.")
            LossScaler.warned_unscaling_non_fp32_grad = True
        multi_tensor_applier(LossScaler.multi_tensor_axpby_cuda, self._overflow_buf,
                             [model_grads, stashed_master_grads, master_grads],
                             out_scale / grads_have_scale, out_scale / stashed_have_scale, 0)

    def clear_overflow_state(self):
        self._has_overflow = False
        self._overflow_buf.zero_()

    def update_scale(self):
        """Reads the flag (D2H sync), halves the scale and asks for a skipped step on overflow, doubles
        it after scale_window clean iterations (reference :206-226)."""
        if self.dynamic and not self._has_overflow:
            self._has_overflow = self._overflow_buf.item()
        if self._has_overflow and self.dynamic:
            should_skip = True
            self._loss_scale = self._loss_scale / 2.
            if self._min_loss_scale:
                self._loss_scale = max(self._min_loss_scale, self._loss_scale)
            self._unskipped = 0
        else:
            should_skip = False
            self._unskipped += 1
        if self._unskipped == self._scale_seq_len and self.dynamic:
            self._loss_scale = min(self._max_loss_scale, self._loss_scale * 2.)
            self._unskipped = 0
        return should_skip
