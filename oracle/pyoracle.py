"""ctypes wrapper of the CPU oracle (TEST INFRASTRUCTURE — see erosion_oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  PARITY UNPINNED (no reference golden vectors exist for this path).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

ARRAYS = ["h", "b", "wh", "dh", "indirection", "newh", "mrec", "mstack", "mwrec", "area", "max_slope", "kfs", "kdf",
          "h_spl", "lake_silt", "lake_sill", "lake_water_volume", "kd_eff", "dh0", "wh_mrec", "b_spl", "h_pre_diff"]
_DTYPES = dict(h=np.float64, b=np.float64, wh=np.float64, dh=np.int32, indirection=np.int32, newh=np.uint32,
               mrec=np.uint8, mstack=np.uint32, mwrec=np.float64, area=np.float64, max_slope=np.float64,
               kfs=np.float32, kdf=np.float32, h_spl=np.float64, lake_silt=np.float64, lake_sill=np.int32,
               lake_water_volume=np.float64, kd_eff=np.float64, dh0=np.int32, wh_mrec=np.float64, b_spl=np.float64,
               h_pre_diff=np.float64)
SCALARS = ["n_gs", "err", "maxh", "boundary_len", "n_lakes", "dag_levels", "n_underwater", "n_sources", "max_uplift",
           "max_g", "newton_iters", "newton_nodes", "newton_max"]


class Params(ctypes.Structure):
    _fields_ = [("k_d_scale", ctypes.c_double), ("k_da", ctypes.c_double), ("chunk_area_pow_q", ctypes.c_double),
                ("min_max_angle", ctypes.c_double), ("max_max_angle", ctypes.c_double), ("mid_slope", ctypes.c_double),
                ("mountain_scale", ctypes.c_double), ("use_libm", ctypes.c_int), ("n_threads", ctypes.c_int)]


class Inputs(ctypes.Structure):
    _fields_ = [(k, ctypes.c_void_p) for k in ("oldh", "oldb", "is_ocean", "uplift", "n", "m", "g", "epsilon_0",
                                               "alpha", "kf", "kd", "height_scale_cell")]


class Noise(ctypes.Structure):
    _fields_ = [("octaves", ctypes.c_uint32), ("frequency", ctypes.c_double), ("lacunarity", ctypes.c_double),
                ("persistence", ctypes.c_double), ("scale_factor", ctypes.c_double), ("perm", ctypes.c_void_p)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liberosion_oracle.so")
        if not os.path.exists(path):
            build()
        L = ctypes.CDLL(path)
        L.orc_create.restype = ctypes.c_void_p
        L.orc_array_bytes.restype = ctypes.c_size_t
        L.orc_get_scalar.restype = ctypes.c_double
        L.orc_timer_name.restype = ctypes.c_char_p
        L.orc_vm.restype = ctypes.c_double
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class Oracle:
    """One map's erosion state on the CPU; mirrors the ve_* C-ABI so tests read the same."""

    def __init__(self, fields, use_libm=False, n_threads=0, libm_f32_powf=None):
        # use_libm: f64 transcendentals from glibc; libm_f32_powf: the Newton loop's powf(dh, 0.5) from glibc's
        # powf (default: follows use_libm, the round-1 meaning of the switch)
        L = lib()
        self.x_lg, self.y_lg = fields["x_lg"], fields["y_lg"]
        self.n = (1 << self.x_lg) * (1 << self.y_lg)
        self.ctx = ctypes.c_void_p(L.orc_create(ctypes.c_uint32(self.x_lg), ctypes.c_uint32(self.y_lg)))
        self._keep = fields
        inp = Inputs(*[_p(np.ascontiguousarray(fields[k])) for k, _ in Inputs._fields_])
        L.orc_set_inputs(self.ctx, ctypes.byref(inp))
        prm = Params()
        L.orc_default_params(ctypes.byref(prm))
        prm.k_d_scale = fields["k_d_scale"]
        prm.k_da = fields["k_da"]
        prm.use_libm = (1 if use_libm else 0) | (2 if (use_libm if libm_f32_powf is None else libm_f32_powf) else 0)
        prm.n_threads = int(n_threads)
        L.orc_set_params(self.ctx, ctypes.byref(prm))
        nz = fields["noise"]
        self._perm = np.ascontiguousarray(nz["perm"])
        cn = Noise(nz["octaves"], nz["frequency"], nz["lacunarity"], nz["persistence"], nz["scale_factor"],
                   _p(self._perm))
        L.orc_set_noise(self.ctx, ctypes.byref(cn))

    def close(self):
        if self.ctx:
            lib().orc_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_state(self, h=None, b=None, wh=None):
        f = lambda a: _p(np.ascontiguousarray(a, np.float64)) if a is not None else None
        lib().orc_set_state(self.ctx, f(h), f(b), f(wh))

    def step(self):
        rc = lib().orc_step(self.ctx)
        if rc != 0:
            raise RuntimeError("oracle step failed: %d" % rc)

    def run(self, n_steps):
        h = np.empty(self.n, np.float64)
        b = np.empty(self.n, np.float64)
        rc = lib().orc_run(self.ctx, ctypes.c_uint32(n_steps), _p(h), _p(b))
        if rc != 0:
            raise RuntimeError("oracle run failed: %d" % rc)
        return h, b

    def get(self, name):
        which = ARRAYS.index(name)
        nb = lib().orc_array_bytes(self.ctx, which)
        dt = np.dtype(_DTYPES[name])
        out = np.empty(nb // dt.itemsize, dt)
        rc = lib().orc_get_array(self.ctx, which, _p(out))
        if rc != 0:
            raise RuntimeError("oracle array %s unavailable" % name)
        return out

    def scalar(self, name):
        return lib().orc_get_scalar(self.ctx, SCALARS.index(name))

    def timers(self):
        buf = (ctypes.c_double * 32)()
        n = lib().orc_get_timers(self.ctx, buf, 32)
        return {lib().orc_timer_name(i).decode(): buf[i] for i in range(n)}


def downhill(x_lg, y_lg, h, ocean):
    n = h.size
    dh = np.empty(n, np.int32)
    lib().orc_downhill(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), _p(np.ascontiguousarray(h, np.float64)),
                       _p(np.ascontiguousarray(ocean, np.uint8)), _p(dh))
    return dh


def get_lakes(x_lg, y_lg, h, dh):
    n = h.size
    dh = np.ascontiguousarray(dh, np.int32).copy()
    ind = np.empty(n, np.int32)
    newh = np.empty(n, np.uint32)
    maxh = ctypes.c_double()
    bl = ctypes.c_uint64()
    rc = lib().orc_get_lakes(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), _p(np.ascontiguousarray(h, np.float64)),
                             _p(dh), _p(ind), _p(newh), ctypes.byref(maxh), ctypes.byref(bl))
    if rc != 0:
        raise RuntimeError("oracle get_lakes failed: %d" % rc)
    return dh, ind, newh, maxh.value, bl.value


def get_multi_rec(x_lg, y_lg, h, dh, newh, wh=None, use_libm=False):
    n = h.size
    wh = np.zeros(n, np.float64) if wh is None else np.ascontiguousarray(wh, np.float64).copy()
    mrec = np.empty(n, np.uint8)
    mstack = np.empty(n, np.uint32)
    mwrec = np.empty(n * 8, np.float64)
    lib().orc_get_multi_rec(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), _p(np.ascontiguousarray(h, np.float64)),
                            _p(np.ascontiguousarray(dh, np.int32)), _p(np.ascontiguousarray(newh, np.uint32)), _p(wh),
                            _p(mrec), _p(mstack), _p(mwrec), ctypes.c_int(int(use_libm)))
    return wh, mrec, mstack, mwrec


def get_multi_drainage(x_lg, y_lg, mstack, mrec, mwrec):
    area = np.empty(mrec.size, np.float64)
    lib().orc_get_multi_drainage(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), _p(mstack), _p(mrec), _p(mwrec),
                                 _p(area))
    return area


def get_drainage(x_lg, y_lg, newh, dh):
    flux = np.empty(dh.size, np.float32)
    lib().orc_get_drainage(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), _p(np.ascontiguousarray(newh, np.uint32)),
                           _p(np.ascontiguousarray(dh, np.int32)), _p(flux))
    return flux


def fill_sinks(x_lg, y_lg, h, ocean):
    out = np.empty(h.size, np.float64)
    sweeps = lib().orc_fill_sinks(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg),
                                  _p(np.ascontiguousarray(h, np.float64)), _p(np.ascontiguousarray(ocean, np.uint8)),
                                  _p(out))
    return out, sweeps


def diffusion(nx, ny, dt, h, b, kd, kdsed=-1.0):
    h = np.ascontiguousarray(h, np.float64).copy()
    b = np.ascontiguousarray(b, np.float64).copy()
    lib().orc_diffusion(ctypes.c_uint32(nx), ctypes.c_uint32(ny), ctypes.c_double(nx * 32.0),
                        ctypes.c_double(ny * 32.0), ctypes.c_double(dt), _p(h), _p(b),
                        _p(np.ascontiguousarray(kd, np.float64)), ctypes.c_double(kdsed))
    return h, b


def get_oceans(x_lg, y_lg, h):
    out = np.empty(h.size, np.uint8)
    lib().orc_get_oceans(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), _p(np.ascontiguousarray(h, np.float64)), _p(out))
    return out


def water_alt(x_lg, y_lg, alt, is_ocean, dh, indirection):
    filled = np.empty(alt.size, np.float64)
    final = np.empty(alt.size, np.float32)
    lib().orc_water_alt(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), _p(np.ascontiguousarray(alt, np.float64)),
                        _p(np.ascontiguousarray(is_ocean, np.uint8)), _p(np.ascontiguousarray(dh, np.int32)),
                        _p(np.ascontiguousarray(indirection, np.int32)), _p(filled), _p(final))
    return filled, final


RIVER_DTYPE = np.dtype([("velocity", np.float32, 3), ("spline_derivative", np.float32, 2), ("kind", np.int32),
                        ("aux", np.uint32, 2)])


def get_rivers(x_lg, y_lg, scale, newh, water_alt_filled, dh, indirection, drainage, use_libm=False):
    out = np.zeros(dh.size, RIVER_DTYPE)
    rc = lib().orc_get_rivers(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), ctypes.c_double(scale),
                              _p(np.ascontiguousarray(newh, np.uint32)), _p(np.ascontiguousarray(water_alt_filled, np.float64)),
                              _p(np.ascontiguousarray(dh, np.int32)), _p(np.ascontiguousarray(indirection, np.int32)),
                              _p(np.ascontiguousarray(drainage, np.float64)), _p(out), ctypes.c_int(int(use_libm)))
    if rc != 0:
        raise RuntimeError("oracle get_rivers failed: %d" % rc)
    return out


def noise_eval(nz, xyz):
    """Fbm<Perlin>::get at the rows of xyz (n x 3): the value a host noise callback hands to the library."""
    xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
    perm = np.ascontiguousarray(nz["perm"])
    cn = Noise(nz["octaves"], nz["frequency"], nz["lacunarity"], nz["persistence"], nz["scale_factor"], _p(perm))
    out = np.empty(xyz.shape[0], dtype=np.float64)
    L = lib()
    L.orc_noise_eval.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]
    L.orc_noise_eval.restype = None
    L.orc_noise_eval(ctypes.byref(cn), _p(xyz), _p(out), xyz.shape[0])
    return out


def map_edge_factor(x_lg, y_lg):
    """util.rs:22-29 restated in numpy (f32 arithmetic)."""
    sx, sy = 1 << x_lg, 1 << y_lg
    ex = np.arange(sx, dtype=np.int64)[None, :].repeat(sy, 0)
    ey = np.arange(sy, dtype=np.int64)[:, None].repeat(sx, 1)
    fx = (sx // 2 - np.abs(ex - sx // 2)).astype(np.float32) / (np.float32(16.0) / np.float32(1024.0) * np.float32(sx))
    fy = (sy // 2 - np.abs(ey - sy // 2)).astype(np.float32) / (np.float32(16.0) / np.float32(1024.0) * np.float32(sy))
    return np.clip(np.minimum(fx, fy), np.float32(0), np.float32(1)).astype(np.float32).ravel()


def uniform_noise(values):
    """util.rs:171-207 restated in numpy: NaN = None; (1 + rank) as f32 / count as f32, ties by index."""
    v = np.asarray(values, np.float32)
    out = np.zeros(v.size, np.float32)
    idx = np.flatnonzero(~np.isnan(v))
    order = idx[np.lexsort((idx, v[idx]))]
    total = np.float32(order.size)
    out[order] = (np.arange(1, order.size + 1).astype(np.float32)) / total
    return out


def horizon_map(x_lg, y_lg, minh, maxh, h):
    """util.rs:371-449 restated literally (pure Python loops: small maps only), F = f32."""
    nx, ny = 1 << x_lg, 1 << y_lg
    h = np.asarray(h, np.float32)
    f = np.float32
    res = []
    for step in (-1, 1):
        ang = np.zeros(h.size, np.float32); hei = np.zeros(h.size, np.float32)
        for posi in range(h.size):
            x = posi & (nx - 1)
            maxdx = x if step < 0 else nx - x
            slope = f(0); h0 = h[posi]
            if h0 < f(minh):
                hv = f(0)
            else:
                max_height = f(0); maxdz = f(maxh) - h0
                for d in range(1, maxdx):
                    deltax = f(32) * f(d)
                    est = slope * deltax
                    if est > maxdz:
                        break
                    act = h[posi + d * step] - h0
                    if est <= act:
                        slope = act / deltax; max_height = act
                hv = h0 - f(minh) + max_height
            ang[posi] = slope; hei[posi] = hv
        res += [ang, hei]
    return res


def vm(fn, x, y=0.0):
    return lib().orc_vm(ctypes.c_int(fn), ctypes.c_double(x), ctypes.c_double(y))
