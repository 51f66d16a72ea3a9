"""CPU tests of the oracle: the reference's own asserts as invariants, analytic checks, and the
committed golden vectors (manufactured by tests/golden/make_golden.py)."""
import heapq
import os

import numpy as np
import pytest

from util import ROOT, check_order_invariants, bit_diff, neighbors
from veloren_b200 import synth
from oracle import pyoracle as po

GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.mark.parametrize("name", ["erode_fbm_32", "erode_plateau_32", "erode_fbm_64"])
def test_oracle_matches_golden(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    lg, kind, steps = int(z["lg"][0]), int(z["kind"][0]), int(z["steps"][0])
    o = po.Oracle(synth.generate(lg, lg, kind=kind))
    for s in range(steps):
        o.step()
        for key in z.files:
            if key.startswith("s%d_" % s) and not key.endswith("n_gs"):
                assert bit_diff(z[key], o.get(key[len("s%d_" % s):])) == 0, key
        assert int(z["s%d_n_gs" % s][0]) == int(o.scalar("n_gs"))


@pytest.mark.parametrize("lg,kind", [(6, 0), (6, 1), (7, 0)])
def test_step_invariants(lg, kind):
    f = synth.generate(lg, lg, kind=kind)
    o = po.Oracle(f)
    n = o.n
    for _ in range(2):
        o.step()
        dh, newh, mrec, mstack = o.get("dh"), o.get("newh"), o.get("mrec"), o.get("mstack")
        check_order_invariants(lg, lg, dh, newh, mrec, mstack)
        # area conservation: every node with receivers hands on all of its area (sum w = 1)
        area = o.get("area")
        sinks = mrec == 0
        assert abs(area[sinks].sum() - n) <= 1e-6 * n
        mw = o.get("mwrec").reshape(n, 8)
        s = mw.sum(axis=1)
        assert np.all((np.abs(s - 1) < 1e-12) | (mrec == 0))
        assert np.all(mw[sinks] == 0)
        h, b = o.get("h"), o.get("b")
        assert np.isfinite(h).all() and np.isfinite(b).all()
        assert (b <= h).all()                                   # diffusion.rs:384-386
        assert 1 <= o.scalar("n_gs") <= 99


def _minimax_fill(h, ocean, x_lg, y_lg):
    """Independent statement of fill_sinks' fixed point: W(c) = min over paths to the ocean of the
    max height along the path (priority flood)."""
    n = h.size
    w = np.full(n, np.inf)
    pq = []
    for i in np.flatnonzero(ocean):
        w[i] = h[i]
        heapq.heappush(pq, (h[i], int(i)))
    done = np.zeros(n, bool)
    while pq:
        lvl, i = heapq.heappop(pq)
        if done[i]:
            continue
        done[i] = True
        for j in neighbors(i, x_lg, y_lg):
            if not ocean[j]:
                c = max(h[j], lvl)
                if c < w[j]:
                    w[j] = c
                    heapq.heappush(pq, (c, j))
    return w


@pytest.mark.parametrize("kind", [0, 1])
def test_fill_sinks_is_the_minimax_level(kind):
    lg = 6
    f = synth.generate(lg, lg, kind=kind)
    h = f["oldh"].astype(np.float64)
    oc = f["is_ocean"]
    out, sweeps = po.fill_sinks(lg, lg, h, oc)
    assert sweeps >= 1
    ref = _minimax_fill(h, oc.astype(bool), lg, lg)
    # ocean cells are not relaxed by the reference (newh == oldh there)
    assert bit_diff(out, np.where(oc.astype(bool), h, ref)) == 0
    assert (out >= h).all()


def test_diffusion_boundaries_and_residual():
    nx = ny = 32
    rng = np.random.default_rng(3)
    h = rng.normal(size=nx * ny) * 10
    b = h - rng.random(nx * ny)
    kd = 1e-3 + rng.random(nx * ny) * 1e-2
    dt = 32000.0
    h2, b2 = po.diffusion(nx, ny, dt, h, b, kd)
    H, H2 = h.reshape(ny, nx), h2.reshape(ny, nx)
    # Dirichlet frame (diffusion.rs:177-180, 207-210, 296-299, 330-333)
    assert np.array_equal(H[0], H2[0]) and np.array_equal(H[-1], H2[-1])
    assert np.array_equal(H[:, 0], H2[:, 0]) and np.array_equal(H[:, -1], H2[:, -1])
    assert (b2 <= h2).all() and (b2 <= b + 1e-300).all()
    # diffusion smooths: interior variance does not grow
    assert H2[1:-1, 1:-1].var() <= H[1:-1, 1:-1].var() * (1 + 1e-12)


def test_get_drainage_counts_upstream_cells():
    lg = 6
    f = synth.generate(lg, lg)
    o = po.Oracle(f)
    o.step()
    dh, newh = o.get("dh"), o.get("newh")
    flux = po.get_drainage(lg, lg, newh, dh)
    # exact in f32 while counts stay below 2^24: flux = 1 + number of upstream cells
    cnt = np.ones(dh.size)
    for c in newh[::-1]:
        if dh[c] >= 0:
            cnt[dh[c]] += cnt[c]
    assert np.array_equal(flux.astype(np.float64), cnt)
    assert flux[dh < 0].sum() == dh.size


def test_standalone_pieces_agree_with_step():
    lg = 6
    f = synth.generate(lg, lg)
    o = po.Oracle(f)
    h0 = o.get("h").copy()
    o.step()
    pred = (f["is_ocean"].astype(bool) & (h0 <= 0)).astype(np.uint8)
    dh0 = po.downhill(lg, lg, h0, pred)
    assert np.array_equal(dh0, o.get("dh0"))
    dh, ind, newh, maxh, bl = po.get_lakes(lg, lg, h0, dh0)
    assert np.array_equal(dh, o.get("dh")) and np.array_equal(ind, o.get("indirection"))
    assert np.array_equal(newh, o.get("newh")) and maxh == o.scalar("maxh") and bl == o.scalar("boundary_len")
    wh, mrec, mstack, mwrec = po.get_multi_rec(lg, lg, h0, dh, newh)
    assert np.array_equal(mstack, o.get("mstack")) and bit_diff(mwrec, o.get("mwrec")) == 0
    area = po.get_multi_drainage(lg, lg, mstack, mrec, mwrec)
    assert bit_diff(area, o.get("area")) == 0


def test_get_oceans_is_the_border_connected_component():
    from scipy import ndimage
    lg = 6
    f = synth.generate(lg, lg)
    o = po.Oracle(f)
    o.step()
    h = o.get("h")
    oc = po.get_oceans(lg, lg, h).reshape(1 << lg, 1 << lg).astype(bool)
    low = (h <= 0).reshape(oc.shape)
    lab, _ = ndimage.label(low, structure=np.ones((3, 3)))
    border = np.zeros_like(low); border[0, :] = border[-1, :] = border[:, 0] = border[:, -1] = True
    keep = np.unique(lab[border & low])
    assert np.array_equal(oc, np.isin(lab, keep[keep > 0]))


def test_water_alt_properties():
    lg = 6
    f = synth.generate(lg, lg, kind=1)
    o = po.Oracle(f)
    o.step()
    alt = o.get("h")
    oc = po.get_oceans(lg, lg, alt)
    dh = po.downhill(lg, lg, alt, oc)
    dh, ind, newh, _, _ = po.get_lakes(lg, lg, alt, dh)
    filled, final = po.water_alt(lg, lg, alt, oc, dh, ind)
    assert (filled >= alt).all()
    lake = np.where(ind < 0, np.arange(ind.size), ind)
    assert (final[dh[lake] < 0] == 0).all()


def test_get_rivers_invariants():
    lg = 7
    f = synth.generate(lg, lg)
    o = po.Oracle(f)
    for _ in range(3):
        o.step()
    alt = o.get("h")
    oc = po.get_oceans(lg, lg, alt)
    dh = po.downhill(lg, lg, alt, oc)
    dh, ind, newh, _, _ = po.get_lakes(lg, lg, alt, dh)
    wh, mrec, mstack, mwrec = po.get_multi_rec(lg, lg, alt, dh, newh)
    area = po.get_multi_drainage(lg, lg, mstack, mrec, mwrec)
    filled, _ = po.water_alt(lg, lg, alt, oc, dh, ind)
    r = po.get_rivers(lg, lg, 2.0, newh, filled, dh, ind, area)
    kind = r["kind"]
    assert ((kind == 1) == (dh < 0)).all()                     # ocean <=> no receiver (erosion.rs:253-258)
    riv = np.flatnonzero(kind == 3)
    assert riv.size > 0
    # "unbroken stream": a river's receiver is a river, a lake or the ocean (erosion.rs:489-493)
    assert (kind[dh[riv]] != 0).all()
    cs = r["aux"][riv].view(np.float32)
    assert (cs[:, 0] <= 64.0 + 1e-6).all()                     # width capped at RECT_SIZE.x * river_max_width
    assert np.isfinite(r["velocity"]).all() and np.isfinite(r["spline_derivative"]).all()


def test_world_file_codec_round_trip(tmp_path):
    """bincode-legacy layout of WorldFile::Veloren0_7_0 (mod.rs:498-551)."""
    import struct
    from veloren_b200 import worldfile
    n = 16 * 8
    alt = np.arange(n, dtype=np.float64) * 0.5 - 3
    bas = alt - 1.25
    p = tmp_path / "map.bin"
    worldfile.save(p, alt, bas, 4, 3, 2.157574498096227)
    raw = p.read_bytes()
    assert len(raw) == 4 + 8 + 8 + 8 + n * 8 + 8 + n * 8
    assert struct.unpack_from("<IIIdQ", raw, 0) == (1, 4, 3, 2.157574498096227, n)
    m = worldfile.load(p)
    assert m["version"] == "0.7.0" and (m["x_lg"], m["y_lg"]) == (4, 3)
    assert np.array_equal(m["alt"], alt) and np.array_equal(m["basement"], bas)
    # a 0.5.0 file: variant 0, no size header
    p2 = tmp_path / "old.bin"
    a16 = np.arange(16, dtype=np.float64)
    p2.write_bytes(struct.pack("<IQ", 0, 16) + a16.tobytes() + struct.pack("<Q", 16) + (a16 - 0.5).tobytes())
    m2 = worldfile.load(p2)
    assert m2["version"] == "0.5.0" and np.array_equal(m2["basement"], a16 - 0.5)
    # WorldMap_0_5_0::into_modern: pow_size = trailing_zeros(16) / 2 = 2, scale = 2^(2 - 10)
    assert (m2["x_lg"], m2["y_lg"]) == (2, 2) and m2["continent_scale_hack"] == 2.0 ** -8
    p3 = tmp_path / "bad.bin"          # 8 entries is not 4^k: WorldSizeInvalid
    p3.write_bytes(struct.pack("<IQ", 0, 8) + a16[:8].tobytes() + struct.pack("<Q", 8) + a16[:8].tobytes())
    with pytest.raises(ValueError):
        worldfile.load(p3)


def test_uniform_noise_restated():
    v = np.array([0.5, np.nan, -1.0, 0.5, 2.0], np.float32)
    out = po.uniform_noise(v)
    assert np.array_equal(out, np.array([2, 0, 1, 3, 4], np.float32) / np.float32(4))


def test_map_edge_factor_restatement():
    """util.rs:22-29: hand-checked values (1024 chunks: 16 chunks = 1.0; the far edge is one chunk short)."""
    m = po.map_edge_factor(10, 10).reshape(1024, 1024)
    assert m[0, 0] == 0.0 and m[8, 8] == np.float32(0.5) and m[16, 500] == 1.0 and m[512, 512] == 1.0
    assert m[1023, 3] == np.float32(1.0 / 16.0) and m[3, 1023] == np.float32(1.0 / 16.0)
    r = po.map_edge_factor(6, 9).reshape(512, 64)          # non-square: x decides near the x edges
    assert r[256, 0] == 0.0 and r[256, 1] == np.float32(1.0) and r[4, 32] == np.float32(0.5)
