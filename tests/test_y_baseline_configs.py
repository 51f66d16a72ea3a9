"""-m gpu, collected after the other parity tests: BASELINE.json's configurations at their full sizes.

C1 (256^2 x 10) lives in test_gpu_parity.py.  Here: C2 the default 1024^2 world free-running for 10 steps plus
the zero-step second do_erosion call (f32 round trip, mod.rs:1399-1421); C5 the depression-heavy plateau at
1024^2 for 10 steps; C3 2048^2 for 2 steps; C4 8192^2 for one step (about 90 s of oracle time); and 20 seeded
random cases (sizes 16..512, both map kinds, four uplift scales).  Every intermediate of every step is compared
bit for bit with the CPU oracle (receivers, lake ids, orders exactly; heights and flux within 1e-9 relative is
the contract, bit-identical is what is asserted)."""
import os

import numpy as np
import pytest

from util import INTEGER_ARRAYS, bit_diff, compare_all
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
REL_TOL = 1e-9


def _ctx(fields):
    c = ErosionContext(fields["x_lg"], fields["y_lg"])
    c.configure(fields)
    return c


def _check(o, c, where, names=None):
    bad = compare_all(o, c) if names is None else compare_all(o, c, names)
    ints = {k: v for k, v in bad.items() if k in INTEGER_ARRAYS}
    assert not ints, "%s: integer arrays differ: %s" % (where, ints)
    flts = {k: v for k, v in bad.items() if v[1] is not None and v[1] > REL_TOL}
    assert not flts, "%s: floats beyond %g: %s" % (where, REL_TOL, flts)
    assert not bad, "%s: within tolerance but not bit-identical: %s" % (where, bad)
    assert int(o.scalar("n_gs")) == int(c.scalar("n_gs")), where


def test_c2_default_1024_ten_steps_free_running_and_f32_round_trip():
    f = synth.generate(10, 10)
    o, c = po.Oracle(f, n_threads=os.cpu_count()), _ctx(f)
    for it in range(10):
        o.step(); c.step()
        _check(o, c, "C2 step %d" % it)
    # second do_erosion call of WorldSim::generate: n_small_steps = 0, inputs are the outputs cast to f32
    h, b = c.get("h"), c.get("b")
    f2 = dict(f); f2["oldh"], f2["oldb"] = h.astype(np.float32), b.astype(np.float32)
    c.set_inputs(f2)
    h2, b2 = c.run(0)
    assert np.array_equal(h2, h.astype(np.float32).astype(np.float64))
    assert np.array_equal(b2, b.astype(np.float32).astype(np.float64))


def test_c5_plateau_1024_ten_steps():
    f = synth.generate(10, 10, kind=1)
    o, c = po.Oracle(f, n_threads=os.cpu_count()), _ctx(f)
    for it in range(10):
        o.step(); c.step()
        if it == 0:
            assert o.scalar("n_lakes") > 10000          # the lake-heavy case it is meant to be
        _check(o, c, "C5 step %d" % it)


def test_c3_2048_two_steps():
    f = synth.generate(11, 11)
    o, c = po.Oracle(f, n_threads=os.cpu_count()), _ctx(f)
    for it in range(2):
        o.step(); c.step()
        _check(o, c, "C3 step %d" % it)


def test_stress_twenty_random_cases():
    rng = np.random.RandomState(20261107)
    for case in range(20):
        x_lg, y_lg = int(rng.randint(4, 10)), int(rng.randint(4, 10))
        kind = int(rng.randint(0, 2))
        seed = int(rng.randint(1, 2 ** 31 - 1))
        scale = float(rng.choice([0.5, 1.0, 2.0, 4.0]))
        f = synth.generate(x_lg, y_lg, seed=seed, kind=kind, scale=scale)
        o, c = po.Oracle(f, n_threads=os.cpu_count()), _ctx(f)
        for it in range(3 if x_lg + y_lg <= 16 else 2):
            o.step(); c.step()
            _check(o, c, "case %d (%dx%d kind %d seed %d scale %.1f) step %d" % (case, 1 << x_lg, 1 << y_lg, kind, seed, scale, it))
        c.close()


def test_c4_8192_one_step():
    """67 M cells: ~50 GB of device state, ~90 s of oracle time.  Skipped when the box lacks the memory."""
    import torch
    free, total = torch.cuda.mem_get_info(0)
    if total < 120 * 2 ** 30:
        pytest.skip("needs a 180 GB device")
    try:
        import psutil
        if psutil.virtual_memory().available < 150 * 2 ** 30:
            pytest.skip("needs ~150 GB of host memory for the oracle")
    except ImportError:
        pass
    f = synth.generate(13, 13)
    o, c = po.Oracle(f, n_threads=os.cpu_count()), _ctx(f)
    o.step(); c.step()
    _check(o, c, "C4 step 0", names=["dh", "indirection", "newh", "mstack", "mrec", "area", "h", "b", "wh"])
