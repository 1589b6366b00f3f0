"""Shared helpers for the parity tests."""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def load_ref(name):
    """Import one of the reference's own kernels compiled unmodified into oracle/_ref
    (oracle/build_ref.sh).  Returns None when it was not built."""
    path = os.path.join(REF_DIR, name + ".so")
    if not os.path.exists(path):
        return None
    import importlib.util
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def rel_err(a, b, floor=1e-30):
    """max |a-b| / max(|b|, floor) over elements, both widened to fp64; nan/inf must coincide."""
    a = a.detach().double().cpu().reshape(-1)
    b = b.detach().double().cpu().reshape(-1)
    fa, fb = torch.isfinite(a), torch.isfinite(b)
    if not torch.equal(fa, fb):
        return float("inf")
    if not torch.equal(a[~fa].isnan(), b[~fb].isnan()):
        return float("inf")
    a, b = a[fa], b[fb]
    if a.numel() == 0:
        return 0.0
    return float(((a - b).abs() / b.abs().clamp_min(floor)).max())


def assert_close_rel(a, b, rtol, atol=0.0, what="", scale=None):
    """Element-wise |a-b| <= atol + rtol*|b| + rtol*scale, compared in fp64.

    `scale` is the magnitude of the terms that were added to produce b: fp32 results that come out of
    a cancellation (m = b1*m + (1-b1)*g landing near zero, p - lr*u ...) differ between two correct
    implementations by an ulp of the *operands* (FMA contraction, approximate division in the
    --use_fast_math reference), which is not small relative to the tiny result.  scale="self" uses
    max|b| of the tensor; a float gives it explicitly (e.g. the update magnitude for p)."""
    a64 = a.detach().double().cpu().reshape(-1)
    b64 = b.detach().double().cpu().reshape(-1)
    assert a64.shape == b64.shape, what
    fin = torch.isfinite(b64)
    assert torch.equal(torch.isfinite(a64), fin), "%s: finiteness differs" % what
    if scale == "self":
        scale = float(b64[fin].abs().max()) if fin.any() else 0.0
    err = (a64[fin] - b64[fin]).abs()
    bound = atol + rtol * b64[fin].abs() + (rtol * scale if scale else 0.0)
    bad = err > bound
    if bad.any():
        i = int(torch.nonzero(bad)[0])
        raise AssertionError("%s: %d/%d elements outside rtol=%g atol=%g scale=%r; first: got %r want %r" % (
            what, int(bad.sum()), err.numel(), rtol, atol, scale, float(a64[fin][i]), float(b64[fin][i])))


def resnet50_shapes():
    import param_shapes
    return param_shapes.resnet50()


def rel_err_stats(a, b, floor=0.0):
    """Element-wise relative error |a-b|/|b| in fp64 over the elements where b is finite and |b| > floor:
    (max, 99.9th percentile, number of elements).  Works on lists of tensors (concatenated)."""
    if isinstance(a, (list, tuple)):
        a = torch.cat([t.detach().reshape(-1).double() for t in a])
        b = torch.cat([t.detach().reshape(-1).double() for t in b])
    else:
        a, b = a.detach().reshape(-1).double(), b.detach().reshape(-1).double()
    keep = torch.isfinite(b) & (b.abs() > floor)
    if not bool(keep.any()):
        return 0.0, 0.0, 0
    r = (a[keep] - b[keep]).abs() / b[keep].abs()
    n = r.numel()
    sub = r if n <= (1 << 24) else r[:: (n >> 24) + 1]      # the percentile of very long lists comes from a strided sample
    p999 = float(torch.kthvalue(sub, max(1, int(math.ceil(0.999 * sub.numel()))))[0])
    return float(r.max()), p999, n


def assert_rel_pure(a, b, rtol, what="", tiny=0.0):
    """Pure element-wise relative bound |a-b| <= rtol*|b| (+ tiny absolute, for flushed denormals only):
    for quantities that are sums of non-negative terms (second moments, norms, trust ratios) there is no
    cancellation to excuse anything looser."""
    if isinstance(a, (list, tuple)):
        for i, (x, y) in enumerate(zip(a, b)):
            assert_rel_pure(x, y, rtol, "%s[%d]" % (what, i), tiny)
        return
    a64, b64 = a.detach().double().reshape(-1), b.detach().double().reshape(-1)
    assert a64.shape == b64.shape, what
    fin = torch.isfinite(b64)
    assert torch.equal(torch.isfinite(a64), fin), "%s: finiteness differs" % what
    err = (a64[fin] - b64[fin]).abs()
    bad = err > (rtol * b64[fin].abs() + tiny)
    if bool(bad.any()):
        i = int(torch.nonzero(bad)[0])
        raise AssertionError("%s: %d/%d elements outside pure rtol=%g; first: got %r want %r (rel %.3g)" % (
            what, int(bad.sum()), err.numel(), rtol, float(a64[fin][i]), float(b64[fin][i]),
            float(err[i] / b64[fin][i].abs().clamp_min(1e-300))))
