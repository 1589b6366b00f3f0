"""Subprocess body of tests/test_reference_python_gpu.py: runs the REFERENCE's own Python (oracle/_ref/pyref.zip: the
unmodified package, compiled to bytecode and archived by oracle/build_ref.sh) over one of two back ends and saves what it computed.

  python tests/_refpy_worker.py ours <out.pt>    amp_C = this repo's shim over libb200mt.so (rocm-apex_b200/amp_C.py)
  python tests/_refpy_worker.py ref  <out.pt>    amp_C = oracle/_ref/amp_C_ref.so, the reference's own kernels

The scenarios follow the reference's tests: tests/L0/run_optimizers/test_fused_optimizer.py:63-113 (7 iterations, the
278011-element tensor and the multi-parameter sizes), test_lamb.py:147-333, and the LossScaler protocol of apex/amp/scaler.py.
"""
import collections.abc
import importlib.util
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PYREF = os.path.join(ROOT, "oracle", "_ref", "pyref.zip")      # bytecode archive of the reference's apex package


def setup(backend):
    six = types.ModuleType("torch._six")            # removed from torch 2.x; apex/amp/_initialize.py:2 imports it
    six.string_classes = (str, bytes)
    six.container_abcs = collections.abc
    sys.modules["torch._six"] = six
    # the reference's package first; this repo's mirror `apex` (rocm-apex_b200/apex) must NOT be the one imported
    sys.path[:] = [p for p in sys.path if not p.rstrip("/").endswith("rocm-apex_b200")]
    sys.path.insert(0, PYREF)
    if backend == "ours":
        sys.path.insert(1, os.path.join(ROOT, "rocm-apex_b200"))
    else:
        path = os.path.join(ROOT, "oracle", "_ref", "amp_C_ref.so")
        import torch  # noqa: F401  (libtorch must be loaded before the extension)
        spec = importlib.util.spec_from_file_location("amp_C_ref", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        sys.modules["amp_C"] = mod
    import apex
    import amp_C
    assert os.path.realpath(apex.__file__).startswith(os.path.realpath(PYREF)), apex.__file__
    origin = os.path.realpath(getattr(amp_C, "__file__", ""))
    want = "rocm-apex_b200/amp_C.py" if backend == "ours" else "oracle/_ref/amp_C_ref.so"
    assert origin.endswith(want), (backend, origin)
    from apex.multi_tensor_apply import multi_tensor_applier
    assert multi_tensor_applier.available
    return apex


def main():
    backend, out = sys.argv[1], sys.argv[2]
    setup(backend)
    import torch
    from apex.optimizers import FusedAdam, FusedLAMB, FusedSGD, FusedMixedPrecisionLamb
    from apex.amp.scaler import LossScaler
    dev = torch.device("cuda:0")
    res = {}
    SIZES = {"one": [[278011]], "multi": [[4096, 1024], [4096], [4096, 2048], [32320, 1024], [1]]}

    def run(tag, make_opt, sizes, iters=7, dtype=torch.float32, grad_scale=1.0, step_kwargs=None):
        g = torch.Generator().manual_seed(9876)
        params = [torch.nn.Parameter((torch.rand(s, generator=g)).to(dev).to(dtype)) for s in sizes]
        opt = make_opt(params)
        for it in range(iters):
            for p in params:
                p.grad = (torch.rand(p.shape, generator=g) * grad_scale).to(dev).to(dtype)
            opt.step(**(step_kwargs or {}))
        torch.cuda.synchronize()
        res[tag] = [p.detach().float().cpu() for p in params]
        return opt

    for name, sizes in SIZES.items():
        run("adam_w/" + name, lambda ps: FusedAdam(ps, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01), sizes)
        run("adam_l2/" + name, lambda ps: FusedAdam(ps, lr=5e-4, weight_decay=0.01, adam_w_mode=False, amsgrad=False), sizes)
        run("lamb_nv/" + name, lambda ps: FusedLAMB(ps, lr=5e-4, weight_decay=0.01, use_nvlamb=True, max_grad_norm=1.0), sizes)
        run("lamb_wd0/" + name, lambda ps: FusedLAMB(ps, lr=5e-4, weight_decay=0.0, max_grad_norm=1.0), sizes)
        run("sgd_mom/" + name, lambda ps: FusedSGD(ps, lr=0.25, momentum=0.125, weight_decay=0.01), sizes)
        run("sgd_nesterov/" + name, lambda ps: FusedSGD(ps, lr=0.1, momentum=0.9, nesterov=True, wd_after_momentum=True, weight_decay=0.01), sizes)
        run("mplamb_fp32/" + name, lambda ps: FusedMixedPrecisionLamb(ps, lr=5e-4, weight_decay=0.01, max_grad_norm=1.0), sizes)
    opt = run("mplamb_fp16/multi", lambda ps: FusedMixedPrecisionLamb(ps, lr=5e-4, weight_decay=0.01, max_grad_norm=1.0,
                                                                     reduced_precision_dtype=torch.float16), SIZES["multi"],
              dtype=torch.float16)
    res["mplamb_fp16/multi/masters"] = [p.detach().float().cpu() for g in opt.param_groups_full_precision for p in g["params"]]
    res["mplamb_fp16/multi/step"] = int(opt.param_groups[0]["step"])

    # ---- LossScaler: unscale (multi_tensor_scale), unscale_with_stashed (multi_tensor_axpby), overflow -> halved scale
    g = torch.Generator().manual_seed(5)
    model = [(torch.randn(s, generator=g) * 100.0).to(dev).half() for s in ([300, 7], [64], [70001])]
    master = [torch.zeros(t.shape, device=dev) for t in model]
    stashed = [torch.randn(t.shape, generator=g).to(dev) for t in model]
    sc = LossScaler("dynamic", init_scale=2.0 ** 10)
    sc.clear_overflow_state()
    sc.unscale(model, master, None)
    res["scaler/unscale"] = [t.cpu() for t in master]
    over0 = sc.update_scale()
    sc.clear_overflow_state()
    sc.unscale_with_stashed(model, stashed, master)
    res["scaler/unscale_with_stashed"] = [t.cpu() for t in master]
    over1 = sc.update_scale()
    model[2][12345] = float("inf")
    sc.clear_overflow_state()
    sc.unscale(model, master, None)
    over2 = sc.update_scale()
    res["scaler/flags"] = [bool(over0), bool(over1), bool(over2), float(sc.loss_scale())]
    res["backend"] = backend
    torch.save(res, out)
    print("refpy worker ok:", backend, len(res), "entries")


if __name__ == "__main__":
    main()
