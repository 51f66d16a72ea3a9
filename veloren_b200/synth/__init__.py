"""Synthetic pre-erosion fields (host side, deterministic).

Follows the recipes of the reference's WorldSim::generate (world/src/sim/mod.rs:849-893,
1092-1337); see gen_inputs.cpp.  Returns plain numpy arrays in the layout ve_set_inputs takes.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

DEFAULT_WORLD_SEED = 130626853  # world/src/sim/mod.rs:586


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libvegen.so")
        if not os.path.exists(path):
            build()
        _LIB = ctypes.CDLL(path)
        _LIB.vegen_generate.restype = ctypes.c_int
    return _LIB


def generate(x_lg, y_lg, seed=DEFAULT_WORLD_SEED, kind=0, scale=2.0):
    """kind 0: fBm continent (C1-C4); kind 1: depression-heavy plateau (C5)."""
    n = (1 << x_lg) * (1 << y_lg)
    f32 = lambda: np.empty(n, np.float32)
    f64 = lambda: np.empty(n, np.float64)
    out = dict(oldh=f32(), oldb=f32(), is_ocean=np.empty(n, np.uint8), uplift=f32(), n=f32(), m=f32(),
               g=f32(), epsilon_0=f32(), alpha=f32(), kf=f64(), kd=f64(), height_scale_cell=f64())
    perm = np.empty(10 * 256, np.uint8)
    scalars = np.zeros(8, np.float64)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    rc = _lib().vegen_generate(ctypes.c_uint32(x_lg), ctypes.c_uint32(y_lg), ctypes.c_uint64(seed), ctypes.c_int(kind),
                               ctypes.c_double(scale), p(out["oldh"]), p(out["oldb"]), p(out["is_ocean"]),
                               p(out["uplift"]), p(out["n"]), p(out["m"]), p(out["g"]), p(out["epsilon_0"]),
                               p(out["alpha"]), p(out["kf"]), p(out["kd"]), p(out["height_scale_cell"]), p(perm),
                               p(scalars))
    if rc != 0:
        raise RuntimeError("vegen_generate failed: %d" % rc)
    out["x_lg"], out["y_lg"] = x_lg, y_lg
    out["k_d_scale"] = float(scalars[0])
    out["k_da"] = float(scalars[1])
    out["max_erosion_per_delta_t"] = float(scalars[2])
    out["noise"] = dict(octaves=10, frequency=float(scalars[3]), lacunarity=float(scalars[4]),
                        persistence=float(scalars[5]), scale_factor=float(scalars[6]), perm=perm)
    out["n_land"] = int(scalars[7])
    return out
