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
    """-> dict(version, x_lg, y_lg, continent_scale_hack, alt, basement).  0.5.0 files carry no size: as in
    WorldMap_0_5_0::into_modern (mod.rs:616-635) the side is derived from the length, pow_size =
    trailing_zeros(len) / 2 (the length must be 4^pow_size), and continent_scale_hack = 2^(pow_size - 10).
    Inconsistent lengths raise, like WorldFileError::WorldSizeInvalid (mod.rs:619-621, 641-646)."""
    buf = open(path, "rb").read()
    (variant,) = struct.unpack_from("<I", buf, 0)
    off = 4
    out = {"version": {0: "0.5.0", 1: "0.7.0"}.get(variant)}
    if variant == 1:
        x_lg, y_lg, scale = struct.unpack_from("<IId", buf, off); off += 16
        out.update(x_lg=x_lg, y_lg=y_lg, continent_scale_hack=scale)
    elif variant == 0:
        pass                                  # size derived from the array length below
    else:
        raise ValueError("unknown WorldFile variant %d" % variant)
    arrays = []
    for _ in range(2):
        (ln,) = struct.unpack_from("<Q", buf, off); off += 8
        arrays.append(np.frombuffer(buf, "<f8", ln, off).copy()); off += 8 * ln
    if off != len(buf):
        raise ValueError("trailing bytes in world file")
    out["alt"], out["basement"] = arrays
    n = arrays[0].size
    if variant == 0:
        pow_size = ((n & -n).bit_length() - 1) // 2 if n else 0
        if n != 1 << (2 * pow_size):
            raise ValueError("world size of map is invalid: %d is not 4^k" % n)
        out.update(x_lg=pow_size, y_lg=pow_size, continent_scale_hack=2.0 ** (pow_size - 10))
    if arrays[1].size != n or n != 1 << (out["x_lg"] + out["y_lg"]) or not out["continent_scale_hack"] > 0.0:
        raise ValueError("world size of map is invalid")
    return out
