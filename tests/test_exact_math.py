"""ve_exact_math.h against the host libm: the 'oracle vs libm' deviation DESIGN.md quotes."""
import math

import numpy as np
import pytest

from util import ROOT  # noqa: F401
from oracle import pyoracle as po


def ulp(x):
    return math.ulp(x) if x != 0 else 5e-324


def _err(fn_id, ref, xs, ys=None):
    worst, mism = 0.0, 0
    for i, x in enumerate(xs):
        y = ys[i] if ys is not None else 0.0
        got = po.vm(fn_id, float(x), float(y))
        want = ref(float(x), float(y)) if ys is not None else ref(float(x))
        if got != want:
            mism += 1
            worst = max(worst, abs(got - want) / ulp(want))
    return worst, mism / len(xs)


def test_log_exp_pow_close_to_libm():
    rng = np.random.default_rng(1)
    n = 20000
    w, m = _err(0, math.log, np.exp((rng.random(n) - 0.5) * 60))
    assert w <= 1.0 and m < 1e-3
    w, m = _err(1, math.exp, (rng.random(n) - 0.5) * 80)
    assert w <= 1.0 and m < 5e-3
    w, m = _err(2, math.pow, np.exp((rng.random(n) - 0.3) * 30), (rng.random(n) - 0.2) * 3)
    assert w <= 1.0 and m < 5e-3
    w, m = _err(3, math.log1p, -(1e-7 + rng.random(n) * (1 - 2e-7)))
    assert w <= 1.0   # glibc's log1p itself is not correctly rounded
    w, m = _err(4, math.tanh, (rng.random(n) - 0.5) * 40)
    assert w <= 2.0


def test_special_values():
    assert po.vm(2, 1024.0, 0.2) == 4.0          # chunk_area.powf(q), erosion.rs:926
    assert po.vm(2, 7.5, 0.0) == 1.0 and po.vm(2, 0.0, 0.0) == 1.0
    assert po.vm(2, 0.0, 0.7) == 0.0             # wrec_k.powf(p) on empty slots, erosion.rs:2468
    assert po.vm(2, 45.254833995939045, 1.0) == 45.254833995939045
    assert po.vm(5, 0.0, 0.0) == 1.0             # powf(dh, n - 1) with n = 1, erosion.rs:1345
    assert po.vm(5, 2.25, 0.5) == 1.5
    assert po.vm(0, 1.0) == 0.0 and po.vm(1, 0.0) == 1.0
    assert po.vm(4, 0.0) == 0.0 and po.vm(4, 50.0) == 1.0 and po.vm(4, -50.0) == -1.0


def _rel(a, b):
    s_ = np.maximum(np.abs(a), np.abs(b))
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.where(s_ > 0, np.abs(a - b) / s_, 0.0)


def test_oracle_libm_vs_exact_math_one_step():
    """Sensitivity of one erode() step (128x128, from identical state) to the two substitutions, each alone, in pure
    relative terms (profiles/r02_libm_drift_*.json has the same over the full 100-step schedule):
      * f64 pow/exp/ln/ln_1p/tanh from glibc instead of ve_exact_math.h: heights agree to ~1e-13, far inside
        north_star's 1e-9, and every integer array is identical;
      * the Newton loop's f32 powf(dh, 0.5) from glibc's powf instead of sqrtf: up to ~1e-5 in single cells (the
        loop stops at tolp = 1e-3, so a last-place change of a term can change the iteration count).  The oracle
        and the kernels use sqrtf: q_ is the constant 1.5 (erosion.rs:749), so the exponent is the literal 0.5 and
        LLVM's libcall simplifier rewrites llvm.pow(x, 0.5) to sqrt(x) (with fabs / -inf handling) without
        fast-math flags."""
    from veloren_b200 import synth
    f = synth.generate(7, 7)
    base = po.Oracle(f)
    f64 = po.Oracle(f, use_libm=True, libm_f32_powf=False)
    f32 = po.Oracle(f, use_libm=False, libm_f32_powf=True)
    for o in (base, f64, f32):
        o.step()
    hb = base.get("h")
    for name in ("dh", "indirection", "newh", "mstack", "mrec"):
        assert np.array_equal(base.get(name), f64.get(name)), name
    assert int(base.scalar("n_gs")) == int(f64.scalar("n_gs"))
    r64 = _rel(hb, f64.get("h"))
    assert r64.max() < 1e-11, r64.max()
    r32 = _rel(hb, f32.get("h"))
    assert r32.max() < 1e-4, r32.max()
