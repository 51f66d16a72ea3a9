"""World-file codec: the on-disk form of do_erosion's output.

`WorldFile` (world/src/sim/mod.rs:498-551) is a bincode-"legacy" stream (fixed-width little-endian
integers, u64 sequence lengths; `encode_into_std_write(map, .., legacy())`, mod.rs:424-455):

    u32 variant                      0 = Veloren0_5_0, 1 = Veloren0_7_0        (#[repr(u32)] enum)
    variant 1 only: u32 x_lg, u32 y_lg (map_size_lg: Vec2<u32>), f64 continent_scale_hack
    u64 len, len * f64  alt
    u64 len, len * f64  basement

Host-side only (no GPU work): it lets a map eroded by libvelerode be loaded by the reference
(`FileOpts::Load`) and vice versa.
"""
import struct

import numpy as np


def save(path, alt, basement, x_lg, y_lg, continent_scale_hack=2.0):
    alt = np.ascontiguousarray(alt, "<f8").reshape(-1)
    basement = np.ascontiguousarray(basement, "<f8").reshape(-1)
    n = (1 << x_lg) * (1 << y_lg)
    if alt.size != n or basement.size != n:
        raise ValueError("alt/basement must have 2^x_lg * 2^y_lg entries")
    with open(path, "wb") as f:
        f.write(struct.pack("<IIId", 1, x_lg, y_lg, float(continent_scale_hack)))
        f.write(struct.pack("<Q", alt.size)); f.write(alt.tobytes())
        f.write(struct.pack("<Q", basement.size)); f.write(basement.tobytes())


def load(path):
    """-> dict(version, x_lg, y_lg, continent_scale_hack, alt, basement); 0.5.0 files have no size
    (the reference assumes 1024x1024 for them, mod.rs:553-580)."""
    buf = open(path, "rb").read()
    (variant,) = struct.unpack_from("<I", buf, 0)
    off = 4
    out = {"version": {0: "0.5.0", 1: "0.7.0"}.get(variant)}
    if variant == 1:
        x_lg, y_lg, scale = struct.unpack_from("<IId", buf, off); off += 16
        out.update(x_lg=x_lg, y_lg=y_lg, continent_scale_hack=scale)
    elif variant == 0:
        out.update(x_lg=10, y_lg=10, continent_scale_hack=2.0)
    else:
        raise ValueError("unknown WorldFile variant %d" % variant)
    arrays = []
    for _ in range(2):
        (ln,) = struct.unpack_from("<Q", buf, off); off += 8
        arrays.append(np.frombuffer(buf, "<f8", ln, off).copy()); off += 8 * ln
    if off != len(buf):
        raise ValueError("trailing bytes in world file")
    out["alt"], out["basement"] = arrays
    return out
