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


@pytest.mark.parametrize("lg", [5])
def test_oracle_libm_vs_exact_math_one_step(lg):
    """Sensitivity of a whole erode() step to the libm substitution.  A last-place difference in an
    f64 pow can flip the f32 rounding of a k_fs/k_df factor (erosion.rs:905, 998), i.e. a 6e-8
    relative change of one coefficient; the resulting height change stays around 1e-9."""
    from veloren_b200 import synth
    f = synth.generate(lg, lg)
    a, b = po.Oracle(f, use_libm=False), po.Oracle(f, use_libm=True)
    a.step(); b.step()
    ha, hb = a.get("h"), b.get("h")
    assert np.array_equal(a.get("dh"), b.get("dh")) and np.array_equal(a.get("mstack"), b.get("mstack"))
    assert np.max(np.abs(ha - hb) / np.maximum(1.0, np.abs(ha))) < 1e-7
