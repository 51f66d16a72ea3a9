"""Host side of the erosion path: a thin ctypes layer over libvelerode.so (include/velerode.h).

`do_erosion` keeps the argument list and meaning of the reference's
`world/src/sim/erosion.rs:2524-2547` (closures become callables or arrays); everything that
computes runs in the CUDA library.  There is no CPU fallback: if the library or a CUDA device is
missing, this module raises.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "csrc", "libvelerode.so")
_LIB = None

ARRAYS = ["h", "b", "wh", "dh", "indirection", "newh", "mrec", "mstack", "mwrec", "area", "max_slope", "kfs", "kdf",
          "h_spl", "lake_silt", "lake_sill", "lake_water_volume", "kd_eff", "dh0", "wh_mrec", "b_spl", "h_pre_diff"]
_DTYPES = dict(h=np.float64, b=np.float64, wh=np.float64, dh=np.int32, indirection=np.int32, newh=np.uint32,
               mrec=np.uint8, mstack=np.uint32, mwrec=np.float64, area=np.float64, max_slope=np.float64,
               kfs=np.float32, kdf=np.float32, h_spl=np.float64, lake_silt=np.float64, lake_sill=np.int32,
               lake_water_volume=np.float64, kd_eff=np.float64, dh0=np.int32, wh_mrec=np.float64, b_spl=np.float64,
               h_pre_diff=np.float64)
SCALARS = ["n_gs", "err", "maxh", "boundary_len", "n_lakes", "max_uplift", "max_g", "n_underwater", "up_ms_total",
           "up_launches_total", "lake_rounds", "wh_undecided"]

EXPORTS = ["ve_default_params", "ve_create", "ve_destroy", "ve_last_error", "ve_set_params", "ve_set_noise",
           "ve_set_noise_cb", "ve_mg_attach", "ve_set_debug", "ve_set_inputs", "ve_set_inputs_from_fields", "ve_set_state", "ve_run", "ve_step", "ve_run_device", "ve_run_device_timed", "ve_get_array", "ve_array_bytes",
           "ve_get_scalar", "ve_launch_count", "ve_get_history", "ve_set_timing", "ve_get_timers", "ve_timer_name", "ve_downhill",
           "ve_get_lakes", "ve_get_multi_rec", "ve_get_multi_drainage", "ve_fill_sinks", "ve_get_drainage",
           "ve_diffusion", "ve_diffusion_static", "ve_get_oceans", "ve_water_alt", "ve_get_rivers", "ve_uniform_noise",
           "ve_map_edge_factor", "ve_horizon_map"]


class VeError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libvelerode error %d: %s" % (code, msg))
        self.code = code


class Params(ctypes.Structure):
    _fields_ = [(k, ctypes.c_double) for k in ("k_d_scale", "k_da", "chunk_area_pow_q", "min_max_angle",
                                               "max_max_angle", "mid_slope", "mountain_scale")]


class Inputs(ctypes.Structure):
    _fields_ = [(k, ctypes.c_void_p) for k in ("oldh", "oldb", "is_ocean", "uplift", "n", "m", "g", "epsilon_0",
                                               "alpha", "kf", "kd", "height_scale_cell")]


class InputsOut(ctypes.Structure):
    _fields_ = [(k, ctypes.c_void_p) for k in ("oldh", "is_ocean", "uplift", "n", "m", "g", "epsilon_0", "alpha", "kf", "kd",
                                               "height_scale_cell")] + [("max_erosion_per_delta_t", ctypes.c_double)]


class Noise(ctypes.Structure):
    _fields_ = [("octaves", ctypes.c_uint32), ("frequency", ctypes.c_double), ("lacunarity", ctypes.c_double),
                ("persistence", ctypes.c_double), ("scale_factor", ctypes.c_double), ("perm", ctypes.c_void_p)]


RIVER_DTYPE = np.dtype([("velocity", np.float32, 3), ("spline_derivative", np.float32, 2), ("kind", np.int32),
                        ("aux", np.uint32, 2)])   # ve_river

PROGRESS_CB = ctypes.CFUNCTYPE(None, ctypes.c_void_p, ctypes.c_double)
ALLGATHER_CB = ctypes.CFUNCTYPE(None, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)   # ve_allgather_cb


def build():
    """nvcc build for sm_100a (cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-s", "-j4", "-C", os.path.join(_HERE, "csrc")])


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError("libvelerode.so is not built (run veloren_b200.erosion.build()); no CPU fallback exists")
        L = ctypes.CDLL(_LIB_PATH)
        L.ve_last_error.restype = ctypes.c_char_p
        L.ve_array_bytes.restype = ctypes.c_size_t
        L.ve_get_scalar.restype = ctypes.c_double
        L.ve_launch_count.restype = ctypes.c_uint64
        L.ve_timer_name.restype = ctypes.c_char_p
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class ErosionContext:
    """Device-resident erosion state of one map (one ve_ctx)."""

    def __init__(self, x_lg, y_lg, device=None):
        L = lib()
        self.x_lg, self.y_lg = int(x_lg), int(y_lg)
        self.n = (1 << self.x_lg) * (1 << self.y_lg)
        self.ctx = ctypes.c_void_p()
        if device is None:
            rc = L.ve_create(ctypes.byref(self.ctx), ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), None, 0)
        else:
            ids = (ctypes.c_int * 1)(int(device))
            rc = L.ve_create(ctypes.byref(self.ctx), ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), ids, 1)
        if rc != 0:
            msg = L.ve_last_error(self.ctx).decode() if self.ctx else "ve_create failed"
            if self.ctx:
                L.ve_destroy(self.ctx)
                self.ctx = None
            raise VeError(rc, msg)
        self._keep = []

    def _chk(self, rc):
        if rc != 0:
            raise VeError(rc, lib().ve_last_error(self.ctx).decode())

    def close(self):
        if getattr(self, "ctx", None):
            lib().ve_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- configuration -------------------------------------------------------------------------
    def set_params(self, **kw):
        prm = Params()
        lib().ve_default_params(ctypes.byref(prm))
        for k, v in kw.items():
            setattr(prm, k, float(v))
        self._chk(lib().ve_set_params(self.ctx, ctypes.byref(prm)))

    def attach_multi(self, rank, world, allgather=None, barrier=None, group=None):
        """ve_mg_attach: shard this map's Gauss-Seidel sweeps as row bands over `world` GPUs (one process per GPU;
        every rank holds a context for the same map with the same inputs and makes the same calls from here on).
        allgather(bytes) -> list of every rank's bytes and barrier() default to torch.distributed on `group`
        (a gloo group is the natural choice: this is the control plane, the data moves over NVLink)."""
        if allgather is None or barrier is None:
            import torch.distributed as dist

            def allgather(b, _g=group):
                out = [None] * world
                dist.all_gather_object(out, b, group=_g)
                return out

            def barrier(_g=group):
                dist.barrier(group=_g)

        def tramp(_user, send, recv, nbytes):
            if nbytes == 0:
                barrier()
                return
            parts = allgather(ctypes.string_at(send, nbytes))
            for i, part in enumerate(parts):
                ctypes.memmove(recv + i * nbytes, part, nbytes)

        self._mg_cb = ALLGATHER_CB(tramp)                  # keep the trampoline alive
        self._chk(lib().ve_mg_attach(self.ctx, int(rank), int(world), self._mg_cb, None))

    def set_debug(self, **kw):
        """Test / tuning knobs of ve_set_debug (lanes_up=8, newton=4, spin_limit_ms=200, ...)."""
        for k, v in kw.items():
            self._chk(lib().ve_set_debug(self.ctx, k.encode(), ctypes.c_longlong(int(v))))

    def set_noise(self, noise):
        perm = np.ascontiguousarray(noise["perm"], np.uint8)
        nz = Noise(int(noise["octaves"]), float(noise["frequency"]), float(noise["lacunarity"]),
                   float(noise["persistence"]), float(noise["scale_factor"]), _p(perm))
        self._chk(lib().ve_set_noise(self.ctx, ctypes.byref(nz)))

    def set_noise_callback(self, fn):
        """rock_strength_nz as a host function (ve_set_noise_cb): fn(points) with points an (N, 3) float64
        array of the `.get()` arguments (erosion.rs:585-589), returning N float64 values.  None switches
        back to the table set with set_noise."""
        if fn is None:
            self._noise_cb = None
            self._chk(lib().ve_set_noise_cb(self.ctx, None, None))
            return
        proto = ctypes.CFUNCTYPE(None, ctypes.c_void_p, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double),
                                 ctypes.c_size_t)

        def tramp(_user, xyz, out, n):
            pts = np.ctypeslib.as_array(xyz, shape=(n, 3))
            res = np.ctypeslib.as_array(out, shape=(n,))
            res[:] = np.asarray(fn(pts), dtype=np.float64).reshape(n)

        self._noise_cb = proto(tramp)                      # keep the trampoline alive
        self._chk(lib().ve_set_noise_cb(self.ctx, self._noise_cb, None))

    def set_inputs(self, fields):
        spec = dict(oldh=np.float32, oldb=np.float32, is_ocean=np.uint8, uplift=np.float32, n=np.float32,
                    m=np.float32, g=np.float32, epsilon_0=np.float32, alpha=np.float32, kf=np.float64,
                    kd=np.float64, height_scale_cell=np.float64)
        arrs = {k: np.ascontiguousarray(fields[k], dt) for k, dt in spec.items()}
        for k, a in arrs.items():
            if a.size != self.n:
                raise ValueError("%s has %d entries, expected %d" % (k, a.size, self.n))
        inp = Inputs(*[_p(arrs[k]) for k, _ in Inputs._fields_])
        self._chk(lib().ve_set_inputs(self.ctx, ctypes.byref(inp)))

    def set_inputs_from_fields(self, gen_scale, alt_old, uheight, rock_strength, is_ocean=None):
        """ve_set_inputs_from_fields: the parameter closures of WorldSim::generate (mod.rs:1102-1337) evaluated on the
        device from the host's noise samples; returns the twelve input arrays (as ve_set_inputs would take them) plus
        max_erosion_per_delta_t, for inspection."""
        n = self.n
        spec = dict(oldh=np.float32, is_ocean=np.uint8, uplift=np.float32, n=np.float32, m=np.float32, g=np.float32,
                    epsilon_0=np.float32, alpha=np.float32, kf=np.float64, kd=np.float64, height_scale_cell=np.float64)
        arrs = {k: np.empty(n, dt) for k, dt in spec.items()}
        out = InputsOut(*[_p(arrs[k]) for k, _ in InputsOut._fields_[:-1]], 0.0)
        a = np.ascontiguousarray(alt_old, np.float32); u = np.ascontiguousarray(uheight, np.float64)
        r = np.ascontiguousarray(rock_strength, np.float64)
        oc = np.ascontiguousarray(is_ocean, np.uint8) if is_ocean is not None else None
        self._chk(lib().ve_set_inputs_from_fields(self.ctx, ctypes.c_double(gen_scale), _p(a), _p(oc) if oc is not None else None,
                                                  _p(u), _p(r), ctypes.byref(out)))
        arrs["oldb"] = arrs["oldh"].copy()
        arrs["max_erosion_per_delta_t"] = out.max_erosion_per_delta_t
        return arrs

    def configure(self, fields):
        """fields: the dict veloren_b200.synth.generate returns (arrays + scalars + noise)."""
        self.set_params(k_d_scale=fields["k_d_scale"], k_da=fields["k_da"])
        self.set_noise(fields["noise"])
        self.set_inputs(fields)

    def set_state(self, h=None, b=None, wh=None):
        f = lambda a: _p(np.ascontiguousarray(a, np.float64)) if a is not None else None
        hh, bb, ww = (np.ascontiguousarray(a, np.float64) if a is not None else None for a in (h, b, wh))
        self._chk(lib().ve_set_state(self.ctx, _p(hh) if hh is not None else None, _p(bb) if bb is not None else None,
                                     _p(ww) if ww is not None else None))

    # -- running -------------------------------------------------------------------------------
    def step(self):
        self._chk(lib().ve_step(self.ctx))

    def run_device(self, n_steps):
        self._chk(lib().ve_run_device(self.ctx, ctypes.c_uint32(n_steps)))

    def run_device_timed(self, n_steps):
        """n_steps erode() calls; returns device milliseconds (CUDA events on the library's stream)."""
        ms = ctypes.c_double()
        self._chk(lib().ve_run_device_timed(self.ctx, ctypes.c_uint32(n_steps), ctypes.byref(ms)))
        return ms.value

    def run(self, n_steps, report_progress=None, out_h=None, out_b=None):
        h = out_h if out_h is not None else np.empty(self.n, np.float64)
        b = out_b if out_b is not None else np.empty(self.n, np.float64)
        cb = PROGRESS_CB(lambda user, pct: report_progress(pct)) if report_progress else PROGRESS_CB()
        self._chk(lib().ve_run(self.ctx, ctypes.c_uint32(n_steps), cb, None, _p(h), _p(b)))
        return h, b

    def get(self, name):
        which = ARRAYS.index(name)
        nb = lib().ve_array_bytes(self.ctx, which)
        dt = np.dtype(_DTYPES[name])
        out = np.empty(nb // dt.itemsize, dt)
        self._chk(lib().ve_get_array(self.ctx, which, _p(out)))
        return out

    def scalar(self, name):
        return lib().ve_get_scalar(self.ctx, SCALARS.index(name))

    def launch_count(self):
        return int(lib().ve_launch_count(self.ctx))

    def history(self, last):
        """(n_gs, n_lakes) lists of the last `last` erode() steps."""
        g = (ctypes.c_int32 * last)()
        l = (ctypes.c_uint32 * last)()
        k = lib().ve_get_history(self.ctx, g, l, int(last))
        return list(g[:k]), list(l[:k])

    def set_timing(self, on=True):
        self._chk(lib().ve_set_timing(self.ctx, int(on)))

    def timers(self):
        buf = (ctypes.c_double * 32)()
        n = lib().ve_get_timers(self.ctx, buf, 32)
        return {lib().ve_timer_name(i).decode(): buf[i] for i in range(n)}

    # -- post-erosion hydrology (world/src/sim/mod.rs:1474-1550) -------------------------------
    def downhill(self, h, is_ocean):
        h = np.ascontiguousarray(h, np.float64)
        oc = np.ascontiguousarray(is_ocean, np.uint8)
        dh = np.empty(self.n, np.int32)
        self._chk(lib().ve_downhill(self.ctx, _p(h), _p(oc), _p(dh)))
        return dh

    def get_lakes(self, h, dh):
        h = np.ascontiguousarray(h, np.float64)
        dh = np.ascontiguousarray(dh, np.int32).copy()
        ind = np.empty(self.n, np.int32)
        newh = np.empty(self.n, np.uint32)
        maxh = ctypes.c_double()
        bl = ctypes.c_uint64()
        self._chk(lib().ve_get_lakes(self.ctx, _p(h), _p(dh), _p(ind), _p(newh), ctypes.byref(maxh), ctypes.byref(bl)))
        return dh, ind, newh, maxh.value, bl.value

    def get_multi_rec(self, h, dh, newh, wh=None):
        h = np.ascontiguousarray(h, np.float64)
        dh = np.ascontiguousarray(dh, np.int32)
        newh = np.ascontiguousarray(newh, np.uint32)
        wh = np.zeros(self.n, np.float64) if wh is None else np.ascontiguousarray(wh, np.float64).copy()
        mrec = np.empty(self.n, np.uint8)
        mstack = np.empty(self.n, np.uint32)
        mwrec = np.empty(self.n * 8, np.float64)
        self._chk(lib().ve_get_multi_rec(self.ctx, _p(h), _p(dh), _p(newh), _p(wh), _p(mrec), _p(mstack), _p(mwrec)))
        return wh, mrec, mstack, mwrec

    def get_multi_drainage(self, mstack, mrec, mwrec):
        area = np.empty(self.n, np.float64)
        self._chk(lib().ve_get_multi_drainage(self.ctx, _p(np.ascontiguousarray(mstack, np.uint32)),
                                              _p(np.ascontiguousarray(mrec, np.uint8)),
                                              _p(np.ascontiguousarray(mwrec, np.float64)), _p(area)))
        return area

    def fill_sinks(self, h, is_ocean):
        out = np.empty(self.n, np.float64)
        self._chk(lib().ve_fill_sinks(self.ctx, _p(np.ascontiguousarray(h, np.float64)),
                                      _p(np.ascontiguousarray(is_ocean, np.uint8)), _p(out)))
        return out

    def get_oceans(self, h):
        out = np.empty(self.n, np.uint8)
        self._chk(lib().ve_get_oceans(self.ctx, _p(np.ascontiguousarray(h, np.float64)), _p(out)))
        return out

    def water_alt(self, alt, is_ocean, dh, indirection):
        filled = np.empty(self.n, np.float64)
        final = np.empty(self.n, np.float32)
        self._chk(lib().ve_water_alt(self.ctx, _p(np.ascontiguousarray(alt, np.float64)),
                                     _p(np.ascontiguousarray(is_ocean, np.uint8)), _p(np.ascontiguousarray(dh, np.int32)),
                                     _p(np.ascontiguousarray(indirection, np.int32)), _p(filled), _p(final)))
        return filled, final

    def get_rivers(self, scale, newh, water_alt_filled, dh, indirection, drainage):
        out = np.zeros(self.n, RIVER_DTYPE)
        self._chk(lib().ve_get_rivers(self.ctx, ctypes.c_double(scale), _p(np.ascontiguousarray(newh, np.uint32)),
                                      _p(np.ascontiguousarray(water_alt_filled, np.float64)),
                                      _p(np.ascontiguousarray(dh, np.int32)), _p(np.ascontiguousarray(indirection, np.int32)),
                                      _p(np.ascontiguousarray(drainage, np.float64)), _p(out)))
        return out

    def uniform_noise(self, values):
        v = np.ascontiguousarray(values, np.float32)
        out = np.empty(self.n, np.float32)
        self._chk(lib().ve_uniform_noise(self.ctx, _p(v), _p(out)))
        return out

    def map_edge_factor(self):
        out = np.empty(self.n, np.float32)
        self._chk(lib().ve_map_edge_factor(self.ctx, _p(out)))
        return out

    def horizon_map(self, minh, maxh, h):
        hh = np.ascontiguousarray(h, np.float32)
        outs = [np.empty(self.n, np.float32) for _ in range(4)]
        self._chk(lib().ve_horizon_map(self.ctx, ctypes.c_float(minh), ctypes.c_float(maxh), _p(hh), *[_p(o) for o in outs]))
        return outs

    def get_drainage(self, newh, dh):
        flux = np.empty(self.n, np.float32)
        self._chk(lib().ve_get_drainage(self.ctx, _p(np.ascontiguousarray(newh, np.uint32)),
                                        _p(np.ascontiguousarray(dh, np.int32)), _p(flux)))
        return flux

    def diffusion(self, dt, h, b, kd):
        h = np.ascontiguousarray(h, np.float64).copy()
        b = np.ascontiguousarray(b, np.float64).copy()
        self._chk(lib().ve_diffusion(self.ctx, ctypes.c_double(dt), _p(h), _p(b), _p(np.ascontiguousarray(kd, np.float64))))
        return h, b


def diffusion_static(x_lg, y_lg, dt, d_h, d_b, d_kd_eff, stream=0):
    """ve_diffusion_static: in-place ADI diffusion on caller-owned DEVICE buffers (integer device
    addresses), using the work space built into the library instead of allocating.  For hosts
    that cannot allocate GPU memory; limited to 512x512 cells.  Raises VeError otherwise."""
    buf = ctypes.create_string_buffer(256)
    rc = lib().ve_diffusion_static(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), ctypes.c_double(dt),
                                   ctypes.c_void_p(d_h), ctypes.c_void_p(d_b), ctypes.c_void_p(d_kd_eff),
                                   ctypes.c_void_p(stream), buf, ctypes.c_size_t(256))
    if rc != 0:
        raise VeError(rc, buf.value.decode(errors="replace"))


def _materialise(f, n, dtype):
    """erosion.rs:2549-2609: evaluate a per-cell closure once into a flat array."""
    if callable(f):
        idx = np.arange(n)
        try:
            out = np.asarray(f(idx))
            if out.shape != (n,):
                raise TypeError
        except Exception:
            out = np.fromiter((f(int(i)) for i in range(n)), dtype=np.float64, count=n)
        return np.ascontiguousarray(out.astype(dtype))
    return np.ascontiguousarray(np.asarray(f).reshape(-1).astype(dtype))


def do_erosion(map_size_lg, _max_uplift, n_steps, seed, rock_strength_nz, oldh, oldb, is_ocean, uplift, n, theta, kf,
               kd, g, epsilon_0, alpha, height_scale, k_d_scale, k_da_scale, threadpool=None, report_progress=None,
               device=None):
    """Drop-in mirror of `do_erosion` (world/src/sim/erosion.rs:2524-2547).

    map_size_lg: (x_lg, y_lg).  Per-cell closures may be callables (index -> value, vectorised or
    scalar) or arrays.  rock_strength_nz: dict with the Fbm<Perlin> tables (ve_noise_fbm_perlin3).
    height_scale: callable n -> f64; k_da_scale: callable q -> f64.  `seed` and `threadpool` are
    accepted and unused, as in the reference (`_seed`, erosion.rs:711).  Returns (h, b) as f64.
    """
    x_lg, y_lg = map_size_lg
    N = (1 << x_lg) * (1 << y_lg)
    n_arr = _materialise(n, N, np.float32)
    theta_arr = _materialise(theta, N, np.float32)
    fields = dict(
        oldh=_materialise(oldh, N, np.float32), oldb=_materialise(oldb, N, np.float32),
        is_ocean=_materialise(is_ocean, N, np.uint8),
        uplift=_materialise(uplift, N, np.float64).astype(np.float32),   # `uplift(posi) as f32`, erosion.rs:2607
        n=n_arr, m=(theta_arr * n_arr).astype(np.float32),               # f32 product, erosion.rs:2570
        g=_materialise(g, N, np.float32), epsilon_0=_materialise(epsilon_0, N, np.float32),
        alpha=_materialise(alpha, N, np.float32), kf=_materialise(kf, N, np.float64),
        kd=_materialise(kd, N, np.float64),
        height_scale_cell=np.asarray([height_scale(float(v)) for v in np.unique(n_arr)], np.float64)[
            np.searchsorted(np.unique(n_arr), n_arr)],
    )
    ctx = ErosionContext(x_lg, y_lg, device=device)
    try:
        ctx.set_params(k_d_scale=k_d_scale, k_da=2.5 * k_da_scale(0.2))   # erosion.rs:745-750
        ctx.set_noise(rock_strength_nz)
        ctx.set_inputs(fields)
        return ctx.run(n_steps, report_progress=report_progress)
    finally:
        ctx.close()
