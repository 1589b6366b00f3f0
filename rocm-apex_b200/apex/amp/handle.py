"""amp.scale_loss (reference: apex/amp/handle.py:16-158): the context manager around backward()."""
import contextlib

import torch

from ._amp_state import _amp_state, maybe_print


def _skip_step_once(opt, loss_scaler, loss_id):
    """Replaces opt.step by a one-shot no-op that reports the overflow and restores the real step."""
    real_step = opt.step

    def skip_step(closure=None):
        if closure is not None:
            raise RuntimeError("Currently, Amp does not support closure use with optimizers.")
        maybe_print("Gradient overflow.  Skipping step, loss scaler {} reducing loss scale to {}".format(
            loss_id, loss_scaler.loss_scale()))
        for p in getattr(opt._amp_stash, "all_fp32_from_fp16_params", ()):
            p.grad = None     # drop the master grads the skipped step would have consumed
        if hasattr(opt, "most_recent_scale"):
            opt.most_recent_scale = 1.0
            opt.scale_set_by_backward = False
        opt.step = real_step
        opt._amp_stash.already_patched = False

    opt.step = skip_step
    opt._amp_stash.already_patched = True


@contextlib.contextmanager
def scale_loss(loss, optimizers, loss_id=0, model=None, delay_unscale=False, delay_overflow_check=False):
    """with amp.scale_loss(loss, optimizer) as scaled_loss: scaled_loss.backward()

    Yields loss.float()*loss_scale.  On exit (unless delay_unscale) gradients are unscaled and checked
    for inf/NaN by the fused kernels; on overflow the dynamic scale is halved and the next
    optimizer.step() becomes a no-op."""
    if not hasattr(_amp_state, "opt_properties"):
        raise RuntimeError("Invoked 'with amp.scale_loss`, but internal Amp state has not been initialized.  "
                           "model, optimizer = amp.initialize(model, optimizer, opt_level=...) must be called "
                           "before `with amp.scale_loss`.")
    props = _amp_state.opt_properties
    if not props.enabled:
        yield loss
        return
    if isinstance(optimizers, torch.optim.Optimizer):
        optimizers = [optimizers]
    loss_scaler = _amp_state.loss_scalers[loss_id]
    loss_scale = loss_scaler.loss_scale()
    if (not props.master_weights) and (not loss_scaler.dynamic) and loss_scale == 1.0:
        yield loss.float()      # nothing to scale, nothing to unscale
        return
    if not delay_unscale:
        for opt in optimizers:
            if not opt._amp_stash.params_have_scaled_gradients:
                opt._prepare_amp_backward()
    yield loss.float() * loss_scale
    if delay_unscale:
        for opt in optimizers:
            opt._amp_stash.params_have_scaled_gradients = True
        return
    loss_scaler.clear_overflow_state()
    for opt in optimizers:
        opt._post_amp_backward(loss_scaler)
        opt._amp_stash.params_have_scaled_gradients = False
    should_skip = False if delay_overflow_check else loss_scaler.update_scale()
    if should_skip:
        for opt in optimizers:
            if not opt._amp_stash.already_patched:
                _skip_step_once(opt, loss_scaler, loss_id)
