"""-m gpu: the CUDA path (through the C ABI) against the CPU oracle and the golden vectors.

Bar: bit-exact for receivers, orders and lake ids; heights/flux within 1e-9 relative
(BASELINE.json north_star) — the tests below actually require bit-identical floats too, since
both sides evaluate the same IEEE operation sequence (include/ve_exact_math.h)."""
import os

import numpy as np
import pytest

from util import ROOT, INTERMEDIATES, INTEGER_ARRAYS, bit_diff, max_rel, compare_all, check_order_invariants
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext, do_erosion, VeError
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
GOLD = os.path.join(ROOT, "tests", "golden")
REL_TOL = 1e-9


def _ctx(fields):
    c = ErosionContext(fields["x_lg"], fields["y_lg"])
    c.configure(fields)
    return c


def _assert_step_equal(o, c, where):
    bad = compare_all(o, c)
    ints = {k: v for k, v in bad.items() if k in INTEGER_ARRAYS}
    assert not ints, "%s: integer arrays differ: %s" % (where, ints)
    flts = {k: v for k, v in bad.items() if v[1] is not None and v[1] > REL_TOL}
    assert not flts, "%s: floats beyond %g: %s" % (where, REL_TOL, flts)
    assert not bad, "%s: not bit-identical (within tolerance though): %s" % (where, bad)
    assert int(o.scalar("n_gs")) == int(c.scalar("n_gs"))
    assert o.scalar("maxh") == c.scalar("maxh")
    assert int(o.scalar("boundary_len")) == int(c.scalar("boundary_len"))


@pytest.mark.parametrize("name", ["erode_fbm_32", "erode_plateau_32", "erode_fbm_64"])
def test_gpu_matches_golden(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    lg, kind, steps = int(z["lg"][0]), int(z["kind"][0]), int(z["steps"][0])
    c = _ctx(synth.generate(lg, lg, kind=kind))
    for s in range(steps):
        c.step()
        for key in z.files:
            if key.startswith("s%d_" % s) and not key.endswith("n_gs"):
                assert bit_diff(z[key], c.get(key[len("s%d_" % s):])) == 0, key
        assert int(z["s%d_n_gs" % s][0]) == int(c.scalar("n_gs"))


def test_c1_256_ten_iterations_every_intermediate():
    """configs[0]: 256x256, default seed, 10 iterations; each step starts from the oracle's state
    and every intermediate is compared; a free-running GPU copy is compared at the end."""
    f = synth.generate(8, 8)
    o, c, free = po.Oracle(f), _ctx(f), _ctx(f)
    for it in range(10):
        c.set_state(o.get("h"), o.get("b"), o.get("wh"))
        o.step(); c.step(); free.step()
        _assert_step_equal(o, c, "C1 step %d" % it)
    assert bit_diff(o.get("h"), free.get("h")) == 0 and bit_diff(o.get("b"), free.get("b")) == 0


def test_c5_plateau_256_lake_heavy():
    f = synth.generate(8, 8, kind=1)
    o, c = po.Oracle(f), _ctx(f)
    for it in range(4):
        o.step(); c.step()
        assert o.scalar("n_lakes") > 100 or it > 0
        _assert_step_equal(o, c, "C5 step %d" % it)
        assert 1 <= int(c.scalar("lake_rounds")) <= 6      # the parallel lake growth ran and found its fixed point


def test_thread_per_node_sweeps_match_too():
    """The sweeps can also run several nodes per warp (one thread per node) instead of one warp per node
    (kept for comparison); force that path on a small map through the debug setter."""
    f = synth.generate(7, 7)
    o, c = po.Oracle(f), _ctx(f)
    c.set_debug(lanes_up=8, lanes_down=8, lanes_other=8)
    for it in range(3):
        o.step(); c.step()
        _assert_step_equal(o, c, "lanes=8 step %d" % it)
    g = synth.generate(7, 7, kind=1)
    o, c = po.Oracle(g), _ctx(g)
    c.set_debug(lanes_up=8, lanes_down=8, lanes_other=8)
    for it in range(2):
        o.step(); c.step()
        _assert_step_equal(o, c, "lanes=8 plateau step %d" % it)


@pytest.mark.parametrize("knob", [dict(newton=4), dict(newton_variant=1), dict(serial_err=1), dict(thermal_sweep=1),
                                  dict(lakes_serial=1), dict(bps_up=1, bps_down=1, poll_sleep_ns=50), dict(level_order=1), dict(rank_jump=1), dict(quad=1), dict(wh_full_sweep=1), dict(rank_launches=1), dict(no_walk_cache=1)])
def test_alternative_code_paths_match(knob):
    """The up-sweep solves a node with explicit in-range sqrt/division sequences and re-solves it with the
    plain operators if a range test ever failed: newton=4 forces the second solve for every node,
    newton_variant selects another formulation of the Newton loop; serial_err=1 sums the Gauss-Seidel error in
    the reference's serial order every iteration (normally only when it lands within rounding of the
    tolerance); thermal_sweep=1 the ordered sweep of erosion.rs:1620 instead of pointer jumping; lakes_serial=1
    grows the lake forest with the one-warp heap (the path degenerate inputs take) instead of the parallel
    fixed-point formulation; bps_*/poll_sleep_ns change the sweeps' occupancy and poll pacing; quad=1 runs the
    sediment sweep with four nodes per warp over the level order padded to quads (the default from 2048^2 on);
    no_walk_cache=1 makes the stack build walk the parent forest for every pair of donors it cannot decide by their keys
    (by default pairs that hang below the same two block anchors share one walk); rank_launches=1 runs every pointer-jumping round as a launch of its own instead of one cooperative launch;
    wh_full_sweep=1 walks every cell in the water-surface sweep instead of only those a running maximum leaves open."""
    for kind in (0, 1):
        f = synth.generate(7, 7, kind=kind)
        o, c = po.Oracle(f), _ctx(f)
        c.set_debug(**knob)
        for it in range(3 if kind == 0 else 2):
            o.step(); c.step()
            _assert_step_equal(o, c, "%s kind %d step %d" % (knob, kind, it))
        if "lakes_serial" in knob:
            assert int(c.scalar("lake_rounds")) == 0
        if "wh_full_sweep" not in knob:
            assert 0 < int(c.scalar("wh_undecided")) < (1 << 14)        # the pre-pass decided some cells, not all


def test_host_noise_callback_matches_device_tables():
    """ve_set_noise_cb: rock strength from a host NoiseFn.  With the callback evaluating the same fBm as
    the device tables the steps must be bit-identical to the oracle (and hence to the table mode)."""
    f = synth.generate(7, 7)
    calls = []

    def noise(points):
        calls.append(points.shape)
        return po.noise_eval(f["noise"], points)

    o, c = po.Oracle(f), _ctx(f)
    c.set_noise_callback(noise)
    for it in range(2):
        o.step(); c.step()
        _assert_step_equal(o, c, "noise callback step %d" % it)
    assert calls == [(1 << 14, 3)] * 2
    c.set_noise_callback(None)
    o.step(); c.step()
    _assert_step_equal(o, c, "back to the tables")
    assert len(calls) == 2
    # callback only: no device table was ever uploaded (ve_set_noise not called)
    c2 = ErosionContext(7, 7)
    c2.set_params(k_d_scale=f["k_d_scale"], k_da=f["k_da"])
    c2.set_inputs(f)
    c2.set_noise_callback(noise)
    o2 = po.Oracle(f)
    o2.step(); c2.step()
    _assert_step_equal(o2, c2, "callback-only context")


def test_lake_water_volume_replay_mid_loop():
    """lake_water_volume updates are logged and replayed once after the Gauss-Seidel loop; a full log is
    is first compacted (adds that can no longer stand are dropped) and replayed early only if that does not make
    room.  lwv_small_log forces either before every iteration (plateau: many lakes)."""
    for mode in (1, 2):     # 1: replay (sort + walk) before every iteration; 2: compact the log before every iteration
        f = synth.generate(7, 7, kind=1)
        o, c = po.Oracle(f), _ctx(f)
        c.set_debug(lwv_small_log=mode)
        for it in range(3):
            o.step(); c.step()
            _assert_step_equal(o, c, "small lwv log (mode %d) step %d" % (mode, it))


@pytest.mark.parametrize("quad", [0, 1])
def test_non_square_and_small_maps(quad):
    for x_lg, y_lg in ((5, 7), (7, 5), (3, 3), (2, 2)):
        f = synth.generate(x_lg, y_lg)
        o, c = po.Oracle(f), _ctx(f)
        c.set_debug(quad=quad)
        for it in range(2):
            o.step(); c.step()
            _assert_step_equal(o, c, "%dx%d step %d" % (1 << x_lg, 1 << y_lg, it))


def test_general_exponent_and_no_deposition():
    """n != 1 takes the general f32 powf path of the Newton loop (erosion.rs:1343-1347);
    g == 0 everywhere makes err = 0 after one Gauss-Seidel pass (erosion.rs:1458)."""
    f = synth.generate(6, 6)
    f["n"] = np.full_like(f["n"], 1.25); f["m"] = (np.float32(0.4) * f["n"]).astype(np.float32)
    o, c = po.Oracle(f), _ctx(f)
    o.step(); c.step()
    _assert_step_equal(o, c, "n=1.25")
    f = synth.generate(6, 6)
    f["g"] = np.zeros_like(f["g"])
    o, c = po.Oracle(f), _ctx(f)
    o.step(); c.step()
    assert int(c.scalar("n_gs")) == 1
    _assert_step_equal(o, c, "g=0")


def test_default_1024_one_step_and_invariants():
    """configs[1] size: one erode() at 1024x1024 against the oracle, plus the reference's asserts."""
    f = synth.generate(10, 10)
    o, c = po.Oracle(f), _ctx(f)
    o.step(); c.step()
    _assert_step_equal(o, c, "1024^2 step 0")
    check_order_invariants(10, 10, c.get("dh"), c.get("newh"), c.get("mrec"), c.get("mstack"))
    area = c.get("area")
    assert abs(area[c.get("mrec") == 0].sum() - c.n) <= 1e-6 * c.n


def test_do_erosion_mirror_and_progress():
    """The do_erosion mirror (erosion.rs:2524) with closures, progress cadence (2648-2651), and
    the zero-step call that only round-trips the state (mod.rs:1399-1421)."""
    f = synth.generate(6, 6)
    nz = f["noise"]
    seen = []
    h, b = do_erosion((6, 6), 32.0, 3, None, nz, lambda i: f["oldh"][i], f["oldb"], f["is_ocean"],
                      lambda i: f["uplift"][i].astype(np.float64), f["n"], lambda i: np.full(np.shape(i), 0.4, np.float32),
                      f["kf"], f["kd"], f["g"], f["epsilon_0"], f["alpha"], lambda n: 0.5 ** n, f["k_d_scale"],
                      lambda q: 0.5 ** (-2.0 * q), report_progress=seen.append)
    o = po.Oracle(f)
    ho, bo = o.run(3)
    assert seen == [(i / 3.0) * 100.0 for i in range(3)]   # (i as f64 / n_steps as f64) * 100.0
    assert max_rel(h, ho) <= REL_TOL and max_rel(b, bo) <= REL_TOL
    h0, b0 = do_erosion((6, 6), 1.0, 0, None, nz, h.astype(np.float32), b.astype(np.float32), f["is_ocean"], f["uplift"],
                        f["n"], np.full(f["n"].size, 0.4, np.float32), f["kf"], f["kd"], f["g"], f["epsilon_0"], f["alpha"],
                        lambda n: 0.5 ** n, f["k_d_scale"], lambda q: 0.5 ** (-2.0 * q))
    assert np.array_equal(h0, h.astype(np.float32).astype(np.float64))


def test_hydrology_entry_points():
    """mod.rs:1474-1550: downhill (plain is_ocean predicate), get_lakes, get_multi_rec,
    get_multi_drainage, fill_sinks, plus get_drainage and the standalone diffusion."""
    lg = 7
    f = synth.generate(lg, lg)
    o = po.Oracle(f)
    for _ in range(2):
        o.step()
    alt = o.get("h")
    oc = f["is_ocean"]
    c = ErosionContext(lg, lg)
    dh_o = po.downhill(lg, lg, alt, oc)
    dh_g = c.downhill(alt, oc)
    assert np.array_equal(dh_o, dh_g)
    # ocean cells above zero would leave -1 lakes next to the border untouched in both paths
    r_o = po.get_lakes(lg, lg, alt, dh_o)
    r_g = c.get_lakes(alt, dh_g)
    for a, b in zip(r_o[:3], r_g[:3]):
        assert np.array_equal(a, b)
    assert r_o[3] == r_g[3] and r_o[4] == r_g[4]
    m_o = po.get_multi_rec(lg, lg, alt, r_o[0], r_o[2])
    m_g = c.get_multi_rec(alt, r_g[0], r_g[2])
    for a, b in zip(m_o, m_g):
        assert bit_diff(a, b) == 0
    assert bit_diff(po.get_multi_drainage(lg, lg, m_o[2], m_o[1], m_o[3]), c.get_multi_drainage(m_g[2], m_g[1], m_g[3])) == 0
    assert bit_diff(po.get_drainage(lg, lg, r_o[2], r_o[0]), c.get_drainage(r_g[2], r_g[0])) == 0
    w_o, _ = po.fill_sinks(lg, lg, alt, oc)
    assert bit_diff(w_o, c.fill_sinks(alt, oc)) == 0
    # 8(f) rank 1: get_oceans on the eroded heights and the water_alt chain (mod.rs:1474, 1502-1592)
    oc2_o = po.get_oceans(lg, lg, alt)
    oc2_g = c.get_oceans(alt)
    assert np.array_equal(oc2_o, oc2_g) and oc2_o.any()
    dh2 = po.downhill(lg, lg, alt, oc2_o)
    lk = po.get_lakes(lg, lg, alt, dh2)
    f_o, w_o = po.water_alt(lg, lg, alt, oc2_o, lk[0], lk[1])
    f_g, w_g = c.water_alt(alt, oc2_o, lk[0], lk[1])
    assert bit_diff(f_o, f_g) == 0 and bit_diff(w_o, w_g) == 0
    # 8(f) rank 2: get_rivers on the same fields (mod.rs:1556-1564)
    mr = po.get_multi_rec(lg, lg, alt, lk[0], lk[2])
    fl = po.get_multi_drainage(lg, lg, mr[2], mr[1], mr[3])
    r_o = po.get_rivers(lg, lg, 2.0, lk[2], f_o, lk[0], lk[1], fl)
    r_g = c.get_rivers(2.0, lk[2], f_o, lk[0], lk[1], fl)
    assert r_o.tobytes() == r_g.tobytes()
    assert (r_o["kind"] == 3).any() and (r_o["kind"] == 1).any()
    kd = o.get("kd_eff")
    h_o, b_o = po.diffusion(1 << lg, 1 << lg, 32000.0, alt, o.get("b"), kd)
    h_g, b_g = c.diffusion(32000.0, alt, o.get("b"), kd)
    assert bit_diff(h_o, h_g) == 0 and bit_diff(b_o, b_g) == 0


def test_fill_sinks_full_size_properties():
    """1024x1024 (BASELINE size): fixed-point property instead of an oracle run."""
    lg = 10
    f = synth.generate(lg, lg, kind=1)
    h = f["oldh"].astype(np.float64)
    oc = f["is_ocean"].astype(bool)
    c = ErosionContext(lg, lg)
    w = c.fill_sinks(h, f["is_ocean"])
    assert (w >= h).all() and np.array_equal(w[oc], h[oc])
    W = w.reshape(1 << lg, 1 << lg)
    P = np.pad(W, 1, constant_values=np.inf)
    mn = np.full_like(W, np.inf)
    for dy in (0, 1, 2):
        for dx in (0, 1, 2):
            if dy != 1 or dx != 1:
                mn = np.minimum(mn, P[dy:dy + W.shape[0], dx:dx + W.shape[1]])
    fixed = np.maximum(h.reshape(W.shape), mn)
    assert np.array_equal(W[~oc.reshape(W.shape)], fixed[~oc.reshape(W.shape)])
    assert np.array_equal(c.fill_sinks(w, f["is_ocean"]), w)   # idempotent


def test_errors_are_reported_not_swallowed():
    c = ErosionContext(4, 4)
    with pytest.raises(VeError):
        c.step()            # no inputs/noise configured
    with pytest.raises(Exception):
        ErosionContext(1, 1)
    with pytest.raises(VeError):
        c.set_debug(no_such_knob=1)


def test_error_codes_mirror_the_reference_panics():
    """assert_ne!(bet, 0.0) (diffusion.rs:445) -> VE_ERR_TRIDAG_SINGULAR; panic!("Disconnected lake!")
    (erosion.rs:1251) -> VE_ERR_DISCONNECTED_LAKE; an order that is not topological (the reference's
    assert_eq!(stack.len(), nn), erosion.rs:2510, can never pass such an order on) -> the sweep gives up
    with VE_ERR_STACK_INCOMPLETE instead of hanging the device."""
    lg = 5
    n = 1 << (2 * lg)
    c = ErosionContext(lg, lg)
    rng = np.random.default_rng(3)
    h = rng.random(n) * 10.0
    # dt/2/dx^2 = 1 and kd = -0.5 everywhere: diag = 1 + (-0.5) + (-0.5) == 0 at the first interior element
    with pytest.raises(VeError) as ei:
        c.diffusion(2048.0, h, h - 1.0, np.full(n, -0.5))
    assert ei.value.code == -4
    # no ocean cell at all: every local minimum is a -1 lake and none can reach the boundary
    dh = c.downhill(h, np.zeros(n, np.uint8))
    assert (dh == -1).any() and not (dh == -2).any()
    with pytest.raises(VeError) as ei:
        c.get_lakes(h, dh)
    assert ei.value.code == -3
    # get_multi_drainage is a pure function of its inputs ...
    f = synth.generate(lg, lg)
    o = po.Oracle(f); o.step()
    mstack, mrec, mwrec = o.get("mstack"), o.get("mrec"), o.get("mwrec")
    assert bit_diff(c.get_multi_drainage(mstack, mrec, mwrec), o.get("area")) == 0
    # ... and an mstack in which a node depends on one that never appears (its receiver's entry replaced by a
    # duplicate) stalls the gather, which gives up instead of hanging
    bad = mstack.copy()
    has_don = np.flatnonzero(mrec != 0)
    victim = int(has_don[0])                             # a node with receivers: it is some receiver's donor
    pos = int(np.flatnonzero(mstack == victim)[0])
    bad[pos] = mstack[0] if pos != 0 else mstack[1]      # victim never published
    c.set_debug(spin_limit_ms=300)
    with pytest.raises(VeError) as ei:
        c.get_multi_drainage(bad, mrec, mwrec)
    assert ei.value.code == -5
    c.set_debug(spin_limit_ms=0)
    assert bit_diff(c.get_multi_drainage(mstack, mrec, mwrec), o.get("area")) == 0      # the context is usable again


def test_two_contexts_on_two_devices_in_one_process():
    """Function attributes (the >48 KB shared-memory opt-ins) and occupancy are per device: stepping contexts on
    two GPUs from one process must work (skipped on a single-GPU box)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    f = synth.generate(7, 7, kind=1)
    o = po.Oracle(f)
    cs = [ErosionContext(7, 7, device=d) for d in (0, 1)]
    for c in cs:
        c.configure(f)
    for it in range(2):
        o.step()
        for c in cs:
            c.step()
            _assert_step_equal(o, c, "device ctx step %d" % it)


def test_map_edge_factor():
    for x_lg, y_lg in ((7, 7), (6, 9), (10, 10)):
        c = ErosionContext(x_lg, y_lg)
        assert bit_diff(po.map_edge_factor(x_lg, y_lg), c.map_edge_factor()) == 0


def test_uniform_noise_and_horizon_map():
    """rank CDF (util.rs:171-207) and the horizon ray march (util.rs:371-449)."""
    lg = 6
    f = synth.generate(lg, lg)
    c = ErosionContext(lg, lg)
    v = f["oldh"].astype(np.float32).copy()
    v[f["is_ocean"].astype(bool)] = np.nan          # None on ocean chunks, as at mod.rs:1118-1124
    v[5:50] = v[5]                                  # ties -> broken by index
    assert bit_diff(po.uniform_noise(v), c.uniform_noise(v)) == 0
    h = (f["oldh"].astype(np.float32) * np.float32(0.25) + np.float32(140.0))
    got = c.horizon_map(140.0, 140.0 + 300.0, h)
    want = po.horizon_map(lg, lg, 140.0, 140.0 + 300.0, h)
    for a, b in zip(want, got):
        assert bit_diff(a, b) == 0
    with pytest.raises(VeError):
        c.horizon_map(10.0, 5.0, h)


def test_parameter_closures_on_device():
    """ve_set_inputs_from_fields against a numpy restatement of mod.rs:849-893 and 1102-1337 (f32/f64
    casts as in the reference; log / ln_1p / tanh / pow through the oracle's exact-math probes), then a step from the
    device-built inputs against the oracle fed with the downloaded arrays."""
    lg = 6
    n = 1 << (2 * lg)
    f = synth.generate(lg, lg)
    rng = np.random.default_rng(5)
    # alt_old in the generator's range with an ocean border, noise samples in [0, 1]
    nx = 1 << lg
    yy, xx = np.divmod(np.arange(n), nx)
    r2 = ((xx + 0.5 - nx / 2) / (nx / 2)) ** 2 + ((yy + 0.5 - nx / 2) / (nx / 2)) ** 2
    alt_old = (0.6 - 1.1 * r2 + 0.2 * rng.standard_normal(n)).astype(np.float32)
    uheight = rng.random(n)
    rock = rng.random(n)
    scale = 2.0
    c = ErosionContext(lg, lg)
    got = c.set_inputs_from_fields(scale, alt_old, uheight, rock)
    # --- restatement
    oc = po.get_oceans(lg, lg, alt_old.astype(np.float64)).astype(bool)          # mod.rs:1092
    assert np.array_equal(got["is_ocean"].astype(bool), oc) and oc.any() and (~oc).any()
    gs = 1.0 / (4.0 / scale)
    hs = po.vm(2, gs, 1.0); ts = hs
    old_height = alt_old * np.float32(2048.0) * np.float32(hs)
    alt = np.where(oc, old_height, (old_height.astype(np.float64) / 2048.0).astype(np.float32) - np.float32(0.5)).astype(np.float32)
    assert bit_diff(alt, got["oldh"]) == 0
    v = old_height.copy(); v[oc] = np.nan
    cdf = po.uniform_noise(v)
    mn = 1.0 / max(float(n), 2.220446049250313e-16 * 0.5); mx = min(1.0 - 1.0 / n, 1.0 - 2.220446049250313e-16 * 0.5)
    hgt = (cdf.astype(np.float64) - 0.5 + 0.5) ** 2
    hgt = hgt * (mx - mn) + mn
    hgt = (hgt - mn) / (mx - mn)
    up = np.where(oc, 0.0, hgt * (64.0 * ts)).astype(np.float32)
    assert bit_diff(up, got["uplift"]) == 0
    kfs = po.vm(2, gs, -2.0 * float(np.float32(0.4) * np.float32(1.0)))
    assert bit_diff(np.where(oc, 1.0e-4, 1.0e-6) * kfs, got["kf"]) == 0
    assert bit_diff(np.full(n, (1.0e-2 / 4.0) * (gs * gs / ts)), got["kd"]) == 0
    assert bit_diff(np.where(po.map_edge_factor(lg, lg) == 0, np.float32(0), np.float32(1)).astype(np.float32), got["g"]) == 0
    l2b = (3.0 ** 0.5) * 0.63661977236758134308
    logit = lambda x: po.vm(0, x) - po.vm(3, -x)
    us = np.array([po.vm(4, (logit(min(1 - 1e-7, max(1e-7, float(rk)))) + (logit(min(0.45, max(0.35, float(uh)))) - logit(0.4))) / l2b) * 0.5 + 0.5
                   for rk, uh in zip(rock, uheight)])
    eps_scale = np.float32(hs / ts); a_scale = np.float32(1.0 / hs)
    eps_land = ((1.0 - us) * (2.078e-3 - 5.3e-5) + 5.3e-5).astype(np.float32) / np.float32(4.0) * eps_scale
    eps = np.where(oc, np.float32(2.078e-3) / np.float32(4.0) * eps_scale, eps_land).astype(np.float32)
    al = np.where(oc, np.float32(3.7e-2) * a_scale, (us * (4.2e-2 - 1.6e-2) + 1.6e-2).astype(np.float32) * a_scale).astype(np.float32)
    assert bit_diff(eps, got["epsilon_0"]) == 0 and bit_diff(al, got["alpha"]) == 0
    assert got["max_erosion_per_delta_t"] == 64.0 * ts
    assert np.all(got["n"] == 1.0) and bit_diff(got["m"], np.full(n, np.float32(0.4), np.float32)) == 0
    # --- the context is ready to step from these inputs: same result as an oracle given the downloaded arrays
    c.set_params(k_d_scale=f["k_d_scale"], k_da=f["k_da"]); c.set_noise(f["noise"])
    f2 = dict(f); f2.update({k: got[k] for k in ("oldh", "oldb", "is_ocean", "uplift", "n", "m", "g", "epsilon_0", "alpha", "kf", "kd",
                                                  "height_scale_cell")})
    o = po.Oracle(f2)
    o.step(); c.step()
    _assert_step_equal(o, c, "step from device-built inputs")
