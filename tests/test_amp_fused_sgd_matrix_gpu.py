"""GPU: the amp x FusedSGD overflow matrix of the reference, tests/L0/run_amp/test_fused_sgd.py:79-183
(test_2models2losses1optimizer), over this repo's amp runtime and kernels:

  materialize_master_grads in {False, True} x opt_level O0..O3 x how gradients are zeroed x one or two loss scalers
  x an inf injected into an fp16 or an fp32 gradient, at iteration 0 or 1, after the first or the second backward.

Two models share one FusedSGD (two param groups); each model has one fp32 and one fp16 parameter; two losses per
iteration, each through amp.scale_loss.  As in the reference: the iteration that saw the inf is skipped (parameters
untouched), the gradients of every other iteration equal those of a plain torch.optim.SGD run, and the final
parameters equal that run's -- i.e. a SGD that skipped exactly the same iterations.
"""
import itertools

import pytest
import torch

pytestmark = pytest.mark.gpu


class MyModel(torch.nn.Module):
    def __init__(self, unique):
        super().__init__()
        self.weight0 = torch.nn.Parameter(unique + torch.arange(2, device="cuda", dtype=torch.float32))
        self.weight1 = torch.nn.Parameter(1. + unique + torch.arange(2, device="cuda", dtype=torch.float16))

    def forward(self, x):
        return ((x * self.weight0.float()) * self.weight1.float()).sum()


@pytest.fixture(scope="module")
def sgd_reference():
    """Plain torch.optim.SGD on the same two models: gradients of its two iterations and the final parameters."""
    x = torch.ones(2, device="cuda")
    m0, m1 = MyModel(1), MyModel(2)
    opt = torch.optim.SGD([{"params": m0.parameters(), "lr": 0.25}, {"params": m1.parameters(), "lr": 0.5}], momentum=0.125)
    grads = []
    for _ in range(2):
        opt.zero_grad()
        m0(x).backward()
        m1(x).backward()
        grads.append([p.grad.detach().clone() for p in itertools.chain(m0.parameters(), m1.parameters())])
        opt.step()
    return grads, [p.detach().clone() for p in itertools.chain(m0.parameters(), m1.parameters())]


def _cases():
    out = []
    for mmg, opt_level, zero, two_scalers in itertools.product((False, True), ("O0", "O1", "O2", "O3"),
                                                               ("none", "model", "optimizer"), (False, True)):
        out.append((mmg, opt_level, zero, two_scalers, -1, None, None))
        if opt_level in ("O1", "O2"):                        # the levels with a dynamic loss scale
            for inject, loc, which in itertools.product((0, 1), ("fp16", "fp32"), (0, 1)):
                out.append((mmg, opt_level, zero, two_scalers, inject, loc, which))
    return out


@pytest.mark.parametrize("mmg,opt_level,zero,two_scalers,inject,loc,which", _cases())
def test_amp_fused_sgd_overflow_matrix(sgd_reference, mmg, opt_level, zero, two_scalers, inject, loc, which):
    from apex import amp
    from apex.amp import _amp_state
    from apex.optimizers import FusedSGD
    ref_grads, final = sgd_reference
    x = torch.ones(2, device="cuda")
    m0, m1 = MyModel(1), MyModel(2)
    models = [m0, m1]
    opt = FusedSGD([{"params": m0.parameters(), "lr": 0.25}, {"params": m1.parameters(), "lr": 0.5}], momentum=0.125,
                   materialize_master_grads=mmg)
    _amp_state.allow_incoming_model_not_fp32 = True
    [m0, m1], opt = amp.initialize([m0, m1], opt, opt_level=opt_level, verbosity=0, cast_model_type=False,
                                   num_losses=2 if two_scalers else 1)
    _amp_state.allow_incoming_model_not_fp32 = False
    _amp_state.loss_scalers[0]._loss_scale = 4.0
    if two_scalers:
        _amp_state.loss_scalers[1]._loss_scale = 16.0
    loss_ids = [0, 1] if two_scalers else [0, 0]
    unskipped, skipped = 0, 0
    for i in range(3 if inject >= 0 else 2):
        if zero == "none":
            for m in models:
                for p in m.parameters():
                    p.grad = None
        elif zero == "model":
            for m in models:
                m.zero_grad()
        else:
            opt.zero_grad()
        l0, l1 = m0(x), m1(x)
        with amp.scale_loss(l0, opt, loss_id=loss_ids[0]) as s:
            s.backward()
            if i == inject and which == 0:
                (m0.weight0 if loc == "fp32" else m0.weight1).grad[0] = float("inf")
        with amp.scale_loss(l1, opt, loss_id=loss_ids[1]) as s:
            s.backward()
            if i == inject and which == 1:
                (m1.weight0 if loc == "fp32" else m1.weight1).grad[0] = float("inf")
        if i != inject:
            if not (opt_level == "O2" and not mmg):      # without materialised master grads there is nothing to look at
                for p, want in zip(amp.master_params(opt), ref_grads[unskipped]):
                    assert torch.allclose(p.grad.float(), want.float()), (i, p.grad, want)
            unskipped += 1
            opt.step()
        else:
            before = [p.detach().clone() for p in amp.master_params(opt)]
            model_before = [p.detach().clone() for m in models for p in m.parameters()]
            opt.step()                                   # the one-shot skipped step
            assert all(torch.equal(a, b.detach()) for a, b in zip(before, amp.master_params(opt)))
            assert all(torch.equal(a, b.detach()) for a, b in zip(model_before, (p for m in models for p in m.parameters())))
            skipped += 1
    assert unskipped == 2 and skipped == (1 if inject >= 0 else 0)
    model_params = [p for m in models for p in m.parameters()]
    for model_p, master_p, want in zip(model_params, amp.master_params(opt), final):
        assert torch.allclose(model_p, want), (model_p, want)
        assert torch.allclose(model_p, master_p.to(model_p.dtype))
    if inject >= 0:       # the scaler that saw the inf halved its scale, the other one did not move
        hit = loss_ids[which]
        assert _amp_state.loss_scalers[hit].loss_scale() == (4.0 if hit == 0 else 16.0) / 2.0
