"""GPU: further behaviours the reference's own tests pin for this path --
channels-last tensors through the multi-tensor ops (tests/L0/run_amp/test_multi_tensor_axpby.py:143-172),
several losses / loss scalers (test_multiple_models_optimizers_losses.py), param groups added after
amp.initialize (test_add_param_group.py), and delayed unscale across two backward passes."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu
CHUNK = 2048 * 32


def flag_tensor(v=0):
    return torch.full((1,), v, dtype=torch.int32, device="cuda")


def test_channels_last_tensors_are_accepted_and_elementwise_correct():
    import amp_C
    shape = (8, 32, 17, 19)
    x = torch.full(shape, 4.0, device="cuda").to(memory_format=torch.channels_last)
    y = torch.full(shape, 16.0, device="cuda", dtype=torch.float16).to(memory_format=torch.channels_last)
    out = torch.zeros(shape, device="cuda").to(memory_format=torch.channels_last)
    assert not x.is_contiguous() and x.is_contiguous(memory_format=torch.channels_last)
    f = flag_tensor()
    amp_C.multi_tensor_axpby(CHUNK, f, [[x], [y], [out]], 2.0, 8.0, -1)
    assert int(f) == 0 and torch.equal(out, torch.full_like(out, 136.0))
    # same memory format on every list: element k of one is element k of the others, whatever the strides
    r = torch.randn(shape, device="cuda").to(memory_format=torch.channels_last)
    o = torch.empty_like(r)
    amp_C.multi_tensor_scale(CHUNK, f, [[r], [o]], 0.5)
    assert torch.equal(o, r * 0.5) and o.is_contiguous(memory_format=torch.channels_last)
    n, _ = amp_C.multi_tensor_l2norm(CHUNK, f, [[r]], False)
    assert abs(float(n) - float(r.norm())) <= 1e-5 * float(r.norm())
    r.view(-1)[0] if False else None
    r[3, 5, 7, 11] = float("nan")
    amp_C.multi_tensor_scale(CHUNK, f, [[r], [o]], 0.5)
    assert int(f) == 1


def _two_models():
    torch.manual_seed(0)
    m0 = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.ReLU(), torch.nn.Linear(32, 4)).cuda()
    m1 = torch.nn.Sequential(torch.nn.Linear(16, 8), torch.nn.Tanh(), torch.nn.Linear(8, 4)).cuda()
    return m0, m1


def test_two_models_two_optimizers_two_losses_with_separate_scalers():
    from apex import amp
    from apex.optimizers import FusedSGD
    m0, m1 = _two_models()
    r0, r1 = copy.deepcopy(m0), copy.deepcopy(m1)
    o0, o1 = FusedSGD(m0.parameters(), lr=0.1, momentum=0.5), FusedSGD(m1.parameters(), lr=0.05, momentum=0.0)
    ro0, ro1 = torch.optim.SGD(r0.parameters(), lr=0.1, momentum=0.5), torch.optim.SGD(r1.parameters(), lr=0.05)
    (m0, m1), (o0, o1) = amp.initialize([m0, m1], [o0, o1], opt_level="O2", num_losses=2, verbosity=0)
    scalers = amp._amp_state.loss_scalers
    assert len(scalers) == 2
    x = torch.randn(8, 16, device="cuda")
    for it in range(4):
        l0 = m0(x).float().pow(2).mean()
        l1 = m1(x).float().pow(2).mean()
        with amp.scale_loss(l0, o0, loss_id=0) as s0:
            s0.backward()
        with amp.scale_loss(l1, o1, loss_id=1) as s1:
            s1.backward()
            if it == 1:
                m1[0].weight.grad[0, 0] = float("inf")      # only loss 1 / optimizer 1 overflows
        before1 = [p.detach().clone() for p in amp.master_params(o1)]
        o0.step(); o1.step()
        o0.zero_grad(); o1.zero_grad()
        ro0.zero_grad(); r0(x).pow(2).mean().backward(); ro0.step()
        if it == 1:
            for a, b in zip(before1, amp.master_params(o1)):
                assert torch.equal(a, b.detach())            # skipped
        else:
            ro1.zero_grad(); r1(x).pow(2).mean().backward(); ro1.step()
    assert scalers[0].loss_scale() == 2. ** 16 and scalers[1].loss_scale() == 2. ** 15
    for p, r in zip(amp.master_params(o0), r0.parameters()):
        assert float((p.detach() - r.detach()).abs().max()) <= 5e-3
    for p, r in zip(amp.master_params(o1), r1.parameters()):
        assert float((p.detach() - r.detach()).abs().max()) <= 5e-3
    sd = amp.state_dict()
    assert set(sd) == {"loss_scaler0", "loss_scaler1"} and sd["loss_scaler1"]["unskipped"] == 2


@pytest.mark.parametrize("opt_level", ["O0", "O2"])
def test_add_param_group_after_initialize(opt_level):
    from apex import amp
    from apex.optimizers import FusedAdam
    m0, m1 = _two_models()
    opt = FusedAdam(m0.parameters(), lr=1e-2)
    (m0, m1), opt = amp.initialize([m0, m1], opt, opt_level=opt_level, verbosity=0)
    x = torch.randn(8, 16, device="cuda")
    with amp.scale_loss((m0(x).float() + m1(x).float()).pow(2).mean(), opt) as s:
        s.backward()
    opt.step(); opt.zero_grad()
    w1 = [p.detach().clone() for p in m1.parameters()]
    opt.add_param_group({"params": list(m1.parameters()), "lr": 5e-3})
    assert len(opt.param_groups) == 2
    if opt_level == "O2":
        assert all(p.dtype == torch.float32 for p in opt.param_groups[1]["params"])      # masters, not the fp16 model params
        assert len(opt._amp_stash.fp16_groups) == 2
    for _ in range(2):
        with amp.scale_loss((m0(x).float() + m1(x).float()).pow(2).mean(), opt) as s:
            s.backward()
        opt.step(); opt.zero_grad()
    assert any(not torch.equal(a, b.detach()) for a, b in zip(w1, m1.parameters()))      # the new group is being stepped


def test_delay_unscale_accumulates_scaled_gradients_then_unscales_once():
    from apex import amp
    from apex.optimizers import FusedAdam
    torch.manual_seed(2)
    model = torch.nn.Linear(16, 4).cuda()
    ref = copy.deepcopy(model)
    opt = FusedAdam(model.parameters(), lr=1e-2)
    model, opt = amp.initialize(model, opt, opt_level="O0", loss_scale=64.0, verbosity=0)
    xs = [torch.randn(8, 16, device="cuda") for _ in range(2)]
    with amp.scale_loss(model(xs[0]).pow(2).mean(), opt, delay_unscale=True) as s:
        s.backward()
    with amp.scale_loss(model(xs[1]).pow(2).mean(), opt) as s:
        s.backward()
    for x in xs:
        ref(x).pow(2).mean().backward()
    for p, r in zip(model.parameters(), ref.parameters()):
        assert torch.allclose(p.grad, r.grad, rtol=1e-5, atol=1e-7)
