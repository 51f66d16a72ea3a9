"""Shared helpers for the parity tests."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

INTERMEDIATES = ["dh0", "dh", "indirection", "newh", "wh_mrec", "mrec", "mwrec", "mstack", "area", "max_slope", "kfs",
                 "kdf", "h_spl", "lake_sill", "lake_silt", "lake_water_volume", "b_spl", "h_pre_diff", "kd_eff", "wh",
                 "h", "b"]
INTEGER_ARRAYS = ["dh0", "dh", "indirection", "newh", "mrec", "mstack", "lake_sill"]


def bit_diff(a, b):
    """number of entries whose bit patterns differ (floats compared bitwise: -0.0 != +0.0)."""
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    assert a.dtype == b.dtype and a.shape == b.shape, (a.dtype, b.dtype, a.shape, b.shape)
    if a.dtype.kind == "f":
        v = np.uint64 if a.dtype.itemsize == 8 else np.uint32
        return int((a.view(v) != b.view(v)).sum())
    return int((a != b).sum())


def max_rel(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    d = np.abs(a - b)
    s = np.maximum(np.abs(a), np.abs(b))
    with np.errstate(invalid="ignore", divide="ignore"):
        r = np.where(s > 0, d / s, 0.0)
    return float(np.nanmax(r)) if r.size else 0.0


def compare_all(oracle, gpu, names=INTERMEDIATES):
    """dict name -> (bit differences, max relative difference) for arrays that differ."""
    bad = {}
    for name in names:
        a, b = oracle.get(name), gpu.get(name)
        nd = bit_diff(a, b)
        if nd:
            bad[name] = (nd, max_rel(a, b) if a.dtype.kind == "f" else None)
    return bad


def neighbors(i, x_lg, y_lg):
    nx, ny = 1 << x_lg, 1 << y_lg
    x, y = i & (nx - 1), i >> x_lg
    for dx, dy in ((-1, -1), (0, -1), (1, -1), (-1, 0), (1, 0), (-1, 1), (0, 1), (1, 1)):
        xx, yy = x + dx, y + dy
        if 0 <= xx < nx and 0 <= yy < ny:
            yield (yy << x_lg) | xx


def check_order_invariants(x_lg, y_lg, dh, newh, mrec, mstack):
    """The reference's own asserts, as checks: newh/mstack are permutations (erosion.rs:1903, 2291,
    2510), no -1 receiver survives get_lakes (1250-1252), receivers precede donors."""
    n = dh.size
    assert np.array_equal(np.sort(newh), np.arange(n, dtype=newh.dtype))
    assert np.array_equal(np.sort(mstack), np.arange(n, dtype=mstack.dtype))
    assert not (dh == -1).any()
    pos = np.empty(n, np.int64)
    pos[newh] = np.arange(n)
    has = dh >= 0
    assert (pos[dh[has]] < pos[has]).all(), "a receiver appears after its donor in newh"
    spos = np.empty(n, np.int64)
    spos[mstack] = np.arange(n)
    nx = 1 << x_lg
    xs, ys = np.arange(n) & (nx - 1), np.arange(n) >> x_lg
    for k, (dx, dy) in enumerate(((-1, -1), (0, -1), (1, -1), (-1, 0), (1, 0), (-1, 1), (0, 1), (1, 1))):
        m = ((mrec >> k) & 1).astype(bool)
        r = ((ys[m] + dy) << x_lg) | (xs[m] + dx)
        assert (spos[r] > spos[np.flatnonzero(m)]).all(), "a multi-receiver appears before its donor in mstack"
