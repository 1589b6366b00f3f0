"""GPU: the host-side mirror of the reference's Python (apex.optimizers, apex.amp) on top of the sm_100a
kernels, tested the way the reference tests its own (tests/L0/run_optimizers/test_fused_optimizer.py,
test_lamb.py, tests/L0/run_amp/test_fused_sgd.py, test_checkpointing.py): against torch.optim on the
same seeded parameters and gradients, plus the amp overflow / skipped-step protocol."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu

SIZES = [[4096, 1024], [4096], [4096, 2048], [32320, 1024], [1]]   # test_fused_optimizer.py:113


def gen_params(sizes, dtype=torch.float32, seed=9876):
    torch.manual_seed(seed)
    ref = [torch.nn.Parameter(torch.rand(s, device="cuda", dtype=dtype)) for s in sizes]
    tst = [torch.nn.Parameter(p.detach().clone()) for p in ref]
    return ref, tst


def gen_grads(ref, tst):
    for r, t in zip(ref, tst):
        r.grad = torch.rand_like(r)
        t.grad = r.grad.clone()


def max_abs_diff(ref, tst):
    return max(float((r.detach().float() - t.detach().float()).abs().max()) for r, t in zip(ref, tst))


@pytest.mark.parametrize("options", [dict(lr=5e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, amsgrad=False),
                                     dict(lr=1e-3, betas=(0.8, 0.99), eps=1e-6, weight_decay=0.1, amsgrad=False)])
@pytest.mark.parametrize("adam_w", [True, False])
def test_fused_adam_vs_torch(options, adam_w):
    from apex.optimizers import FusedAdam
    ref_p, tst_p = gen_params([[278011]] + SIZES)
    ref = (torch.optim.AdamW if adam_w else torch.optim.Adam)(ref_p, **options)
    tst = FusedAdam(tst_p, adam_w_mode=adam_w, **options)
    for _ in range(7):
        gen_grads(ref_p, tst_p)
        ref.step(); tst.step()
        assert max_abs_diff(ref_p, tst_p) <= 1e-6     # reference's own bound is 1e-3 (test_fused_optimizer.py:12-16)
    assert tst.param_groups[0]["step"] == 7
    st = tst.state[tst_p[0]]
    assert set(st) == {"exp_avg", "exp_avg_sq"}
    with pytest.raises(RuntimeError, match="AMSGrad"):
        FusedAdam(tst_p, amsgrad=True)
    with pytest.raises(RuntimeError, match="FusedAdam has been updated"):
        tst.step(grads=[1])


def test_fused_adam_mixed_dtype_groups_and_zero_grad():
    from apex.optimizers import FusedAdam
    torch.manual_seed(1)
    p32 = torch.nn.Parameter(torch.rand(1000, device="cuda"))
    p16 = torch.nn.Parameter(torch.rand(1000, device="cuda").half())
    pbf = torch.nn.Parameter(torch.rand(1000, device="cuda").bfloat16())
    opt = FusedAdam([p32, p16, pbf], lr=1e-2)
    before = [p.detach().clone() for p in (p32, p16, pbf)]
    for p in (p32, p16, pbf):
        p.grad = torch.ones_like(p)
    opt.step()
    for p, b in zip((p32, p16, pbf), before):
        assert float((p.detach().float() - b.float()).abs().max()) > 1e-3    # every dtype bucket was stepped
    opt.zero_grad()
    assert p32.grad is None and p16.grad is None


def test_fused_adam_state_dict_roundtrip():
    from apex.optimizers import FusedAdam
    ref_p, tst_p = gen_params(SIZES[:3])
    a, b = FusedAdam(ref_p, lr=1e-3), FusedAdam(tst_p, lr=1e-3)
    for _ in range(3):
        gen_grads(ref_p, tst_p)
        a.step(); b.step()
    c_p = [torch.nn.Parameter(p.detach().clone()) for p in tst_p]
    c = FusedAdam(c_p, lr=1e-3)
    c.load_state_dict(copy.deepcopy(b.state_dict()))
    gen_grads(ref_p, tst_p)
    for cp, tp in zip(c_p, tst_p):
        cp.grad = tp.grad.clone()
    a.step(); b.step(); c.step()
    assert max_abs_diff(tst_p, c_p) == 0.0 and c.param_groups[0]["step"] == 4


@pytest.mark.parametrize("options", [dict(lr=0.25, momentum=0.125, dampening=0.0, weight_decay=0.0, nesterov=False),
                                     dict(lr=0.1, momentum=0.9, dampening=0.0, weight_decay=1e-4, nesterov=True),
                                     dict(lr=0.1, momentum=0.0, dampening=0.0, weight_decay=1e-4, nesterov=False)])
def test_fused_sgd_vs_torch(options):
    from apex.optimizers import FusedSGD
    ref_p, tst_p = gen_params([[278011]] + SIZES)
    ref, tst = torch.optim.SGD(ref_p, **options), FusedSGD(tst_p, **options)
    for _ in range(7):
        gen_grads(ref_p, tst_p)
        ref.step(); tst.step()
        assert max_abs_diff(ref_p, tst_p) <= 1e-5
    if options["momentum"]:
        assert "momentum_buffer" in tst.state[tst_p[0]]


class RefLAMB(torch.optim.Optimizer):
    """Plain-PyTorch LAMB with global-norm clipping, written from the update rule the reference kernels
    implement (csrc/multi_tensor_lamb.cu:110-142,255-263): the check for FusedLAMB."""

    def __init__(self, params, lr, betas, eps, weight_decay, max_grad_norm=1.0, use_nvlamb=False, grad_averaging=True):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self.max_grad_norm, self.use_nvlamb, self.grad_averaging = max_grad_norm, use_nvlamb, grad_averaging

    @torch.no_grad()
    def step(self):
        gn = torch.sqrt(sum((p.grad.double() ** 2).sum() for g in self.param_groups for p in g["params"]))
        clip = max(float(gn) / self.max_grad_norm, 1.0)
        for group in self.param_groups:
            b1, b2 = group["betas"]
            group["step"] = group.get("step", 0) + 1
            for p in group["params"]:
                st = self.state[p]
                if not st:
                    st["m"], st["v"] = torch.zeros_like(p, dtype=torch.float64), torch.zeros_like(p, dtype=torch.float64)
                g = p.grad.double() / clip
                st["m"].mul_(b1).add_(g, alpha=(1 - b1) if self.grad_averaging else 1.0)
                st["v"].mul_(b2).addcmul_(g, g, value=1 - b2)
                mh, vh = st["m"] / (1 - b1 ** group["step"]), st["v"] / (1 - b2 ** group["step"])
                u = mh / (vh.sqrt() + group["eps"]) + group["weight_decay"] * p.double()
                ratio = 1.0
                if self.use_nvlamb or group["weight_decay"] != 0:
                    pn, un = p.double().norm(), u.norm()
                    ratio = float(pn / un) if (pn != 0 and un != 0) else 1.0
                p.copy_((p.double() - group["lr"] * ratio * u).to(p.dtype))


@pytest.mark.parametrize("wd", [0.0, 0.01])
@pytest.mark.parametrize("nvlamb", [False, True])
def test_fused_lamb_vs_reference_rule(wd, nvlamb):
    from apex.optimizers import FusedLAMB
    ref_p, tst_p = gen_params([[4096, 64], [4096], [333, 77], [1]])
    opts = dict(lr=1e-2, betas=(0.9, 0.999), eps=1e-6, weight_decay=wd)
    ref = RefLAMB(ref_p, max_grad_norm=1.0, use_nvlamb=nvlamb, **opts)
    tst = FusedLAMB(tst_p, max_grad_norm=1.0, use_nvlamb=nvlamb, **opts)
    for _ in range(5):
        gen_grads(ref_p, tst_p)
        ref.step(); tst.step()
        assert max_abs_diff(ref_p, tst_p) <= 2e-6


def test_fused_mixed_precision_lamb_with_grad_scaler():
    from apex.optimizers import FusedLAMB, FusedMixedPrecisionLamb
    torch.manual_seed(3)
    sizes = [[512, 64], [512], [100, 7]]
    master = [torch.rand(s, device="cuda") for s in sizes]
    p32 = [torch.nn.Parameter(m.half().float()) for m in master]            # fp32 run on the fp16-representable start
    p16 = [torch.nn.Parameter(m.half()) for m in master]
    opts = dict(lr=1e-2, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01, max_grad_norm=1.0)
    ref = FusedLAMB(p32, **opts)
    tst = FusedMixedPrecisionLamb(p16, reduced_precision_dtype=torch.float16, **opts)
    scaler = torch.amp.GradScaler("cuda", init_scale=128.0, growth_interval=1000)
    for it in range(4):
        g = [torch.rand_like(p).half() for p in p32]
        for a, b, gg in zip(p32, p16, g):
            a.grad = gg.float()
            b.grad = gg * 128.0       # what a scaled backward would leave
        ref.step()
        scaler._lazy_init_scale_growth_tracker(torch.device("cuda")) if scaler._scale is None else None
        tst.step(grad_scaler=scaler)
        for a, b in zip(p32, p16):
            assert float((a.detach() - b.detach().float()).abs().max()) <= 2e-3      # fp16 model copy of the same masters
        for a, mfull in zip(p32, tst.param_groups_full_precision[0]["params"]):
            assert float((a.detach() - mfull).abs().max()) <= 1e-5
    assert int(tst.param_groups[0]["step"]) == 4
    # an inf gradient: found_inf -> every kernel skips, step counter does not advance
    before = [p.detach().clone() for p in p16]
    for b in p16:
        b.grad = torch.full_like(b, float("inf"))
    tst.step(grad_scaler=scaler)
    assert int(tst.param_groups[0]["step"]) == 4
    for a, b in zip(before, p16):
        assert torch.equal(a, b.detach())


# ---------------------------------------------------------------------------------------------------
# amp: loss scaling protocol on top of the fused kernels
# ---------------------------------------------------------------------------------------------------
class Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        # no bias in front of the BatchNorm: its gradient is pure rounding noise (analytically zero), and
        # Adam turns noise of either sign into +-lr steps, which no two implementations reproduce
        self.l1 = torch.nn.Linear(64, 128, bias=False)
        self.bn = torch.nn.BatchNorm1d(128)
        self.l2 = torch.nn.Linear(128, 10)

    def forward(self, x):
        return self.l2(torch.relu(self.bn(self.l1(x))))


@pytest.mark.parametrize("opt_level", ["O0", "O2", "O5"])
@pytest.mark.parametrize("which", ["adam", "sgd"])
def test_amp_training_matches_fp32_reference_and_skips_on_overflow(opt_level, which):
    from apex import amp
    from apex.optimizers import FusedAdam, FusedSGD
    torch.manual_seed(0)
    model = Net().cuda()
    ref_model = copy.deepcopy(model)
    mk = (lambda ps: FusedAdam(ps, lr=1e-3)) if which == "adam" else (lambda ps: FusedSGD(ps, lr=0.05, momentum=0.9))
    mk_ref = (lambda ps: torch.optim.AdamW(ps, lr=1e-3, weight_decay=0.0)) if which == "adam" else \
        (lambda ps: torch.optim.SGD(ps, lr=0.05, momentum=0.9))
    opt, ref_opt = mk(model.parameters()), mk_ref(ref_model.parameters())
    model, opt = amp.initialize(model, opt, opt_level=opt_level, verbosity=0)
    if opt_level == "O5" and which == "sgd":
        # FusedSGD's fused model-weight write-out is Half only, in the reference as well
        # (csrc/multi_tensor_sgd_kernel.cu:160-161): bf16 model params + master weights are refused
        with amp.scale_loss(model(torch.randn(32, 64, device="cuda")).float().sum(), opt) as scaled:
            scaled.backward()
        with pytest.raises(RuntimeError, match="Additional output tensors should always be fp16"):
            opt.step()
        return
    if opt_level in ("O2", "O5"):
        low = torch.float16 if opt_level == "O2" else torch.bfloat16
        assert model.l1.weight.dtype == low and model.bn.weight.dtype == torch.float32
    x = torch.randn(32, 64, device="cuda")
    y = torch.randint(0, 10, (32,), device="cuda")
    scaler = amp._amp_state.loss_scalers[0]
    start_scale = scaler.loss_scale()
    skipped = 0
    for it in range(6):
        loss = torch.nn.functional.cross_entropy(model(x), y)
        with amp.scale_loss(loss, opt) as scaled:
            scaled.backward()
            if it == 2 and opt_level == "O2":      # poison one gradient the way test_fused_sgd.py:152-163 does
                model.l2.weight.grad[0, 0] = float("inf")
        if it == 2 and opt_level == "O2":
            assert scaler.loss_scale() == start_scale / 2.0
            before = [p.detach().clone() for p in amp.master_params(opt)]
            opt.step()                              # must be the one-shot no-op
            for a, b in zip(before, amp.master_params(opt)):
                assert torch.equal(a, b.detach())
            skipped += 1
        else:
            opt.step()
            ref_loss = torch.nn.functional.cross_entropy(ref_model(x), y)
            ref_opt.zero_grad()
            ref_loss.backward()
            ref_opt.step()
        opt.zero_grad()
    tol = {"O0": 1e-5, "O2": 3e-2, "O5": 1e-1}[opt_level]
    for p, r in zip(model.parameters(), ref_model.parameters()):
        assert float((p.detach().float() - r.detach()).abs().max()) <= tol
    if opt_level in ("O2", "O5"):
        # model weights are the rounded master weights
        stash = opt._amp_stash
        for m16, m32 in zip(stash.all_fp16_params, stash.all_fp32_from_fp16_params):
            assert torch.equal(m16.detach(), m32.detach().to(m16.dtype))
    sd = amp.state_dict()
    assert set(sd) == {"loss_scaler0"} and set(sd["loss_scaler0"]) == {"loss_scale", "unskipped"}
    # O0 (static scale 1.0, no master weights) takes scale_loss's early exit and never touches the scaler
    assert sd["loss_scaler0"]["unskipped"] == {"O0": 0, "O2": 3, "O5": 6}[opt_level]
    amp.load_state_dict({"loss_scaler0": {"loss_scale": 4.0, "unskipped": 1}})
    assert scaler.loss_scale() == 4.0 and scaler._unskipped == 1


def test_amp_gradient_accumulation_uses_axpby():
    from apex import amp
    from apex.optimizers import FusedSGD
    torch.manual_seed(1)
    model = torch.nn.Linear(16, 4).cuda()
    ref_model = copy.deepcopy(model)
    opt = FusedSGD(model.parameters(), lr=0.1, momentum=0.0, materialize_master_grads=True)
    model, opt = amp.initialize(model, opt, opt_level="O2", loss_scale=128.0, verbosity=0)
    xs = [torch.randn(8, 16, device="cuda") for _ in range(2)]
    for x in xs:      # two backward passes, one step: the second unscale goes through multi_tensor_axpby
        with amp.scale_loss(model(x).float().pow(2).mean(), opt) as scaled:
            scaled.backward()
    opt.step()
    ref_opt = torch.optim.SGD(ref_model.parameters(), lr=0.1)
    for x in xs:
        ref_model(x).pow(2).mean().backward()
    ref_opt.step()
    for p, r in zip(amp.master_params(opt), ref_model.parameters()):
        assert float((p.detach() - r.detach()).abs().max()) <= 5e-3


def test_multi_tensor_applier_contract():
    from apex.multi_tensor_apply import multi_tensor_applier, MultiTensorApply
    import amp_C
    assert multi_tensor_applier.available and multi_tensor_applier.chunk_size == 2048 * 32
    a = [torch.full((100,), 4.0, device="cuda")]
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    multi_tensor_applier(amp_C.multi_tensor_scale, f, [a, a], 0.25)
    assert torch.equal(a[0], torch.ones(100, device="cuda"))
    n, _ = MultiTensorApply(333)(amp_C.multi_tensor_l2norm, f, [a], False)
    assert abs(float(n) - 10.0) < 1e-5


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_tensors_on_another_device_than_current():
    # tests/L0/run_optimizers/test_fused_optimizer.py:104-109: current device != tensor device
    from apex.optimizers import FusedAdam
    ref_p, tst_p = gen_params([[4096]])
    ref_p = [torch.nn.Parameter(p.detach().to("cuda:1")) for p in ref_p]
    tst_p = [torch.nn.Parameter(p.detach().to("cuda:1")) for p in tst_p]
    with torch.cuda.device(0):
        ref, tst = torch.optim.AdamW(ref_p, lr=1e-3, weight_decay=0.0), FusedAdam(tst_p, lr=1e-3)
        tst._dummy_overflow_buf = tst._dummy_overflow_buf.to("cuda:1")
        gen_grads(ref_p, tst_p)
        ref.step(); tst.step()
    assert max_abs_diff(ref_p, tst_p) <= 1e-6
    mixed = [torch.nn.Parameter(torch.rand(8, device="cuda:0")), torch.nn.Parameter(torch.rand(8, device="cuda:1"))]
    opt = FusedAdam(mixed, lr=1e-3)
    for p in mixed:
        p.grad = torch.rand_like(p)
    with pytest.raises(RuntimeError, match="not on the same device"):
        opt.step()
