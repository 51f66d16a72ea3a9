"""libm sensitivity of the oracle over the full schedule (CPU only; SURVEY hard part 7).

The CUDA path and the oracle evaluate pow/exp/ln/ln_1p/tanh through include/ve_exact_math.h and the Newton
loop's f32 powf(dh, 0.5) as sqrtf; the Rust reference calls the host libm.  Nothing pins which bits rustc+glibc
produce, so this measures how far the two choices move the result: two oracles run free from the same initial
state, one with the exact-math header (baseline), one with a switch flipped, compared after every step.

usage: python tools/libm_drift.py LG STEPS OUT.json
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from veloren_b200 import synth          # noqa: E402
from oracle import pyoracle as po       # noqa: E402


def main():
    lg, steps, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
    f = synth.generate(lg, lg)
    nt = os.cpu_count()
    base = po.Oracle(f, n_threads=nt)
    variants = {"f64_transcendentals_from_glibc": po.Oracle(f, use_libm=True, libm_f32_powf=False, n_threads=nt),
                "f32_powf_half_from_glibc_powf": po.Oracle(f, use_libm=False, libm_f32_powf=True, n_threads=nt),
                "both": po.Oracle(f, use_libm=True, libm_f32_powf=True, n_threads=nt)}
    res = {"map": "%dx%d" % (1 << lg, 1 << lg), "steps": steps, "metric": "pure relative |dh|/max(|h_a|,|h_b|), free-running",
           "variants": {k: {"max_rel_h": [], "cells_beyond_1e-9": [], "first_step_dh_differs": None,
                            "first_step_mstack_differs": None, "first_step_newh_differs": None,
                            "first_step_n_gs_differs": None} for k in variants}}
    for it in range(steps):
        base.step()
        hb = base.get("h")
        ints = {k: base.get(k) for k in ("dh", "mstack", "newh")}
        for name, o in variants.items():
            o.step()
            h = o.get("h")
            s = np.maximum(np.abs(h), np.abs(hb))
            with np.errstate(invalid="ignore", divide="ignore"):
                r = np.where(s > 0, np.abs(h - hb) / s, 0.0)
            v = res["variants"][name]
            v["max_rel_h"].append(float(r.max()))
            v["cells_beyond_1e-9"].append(int((r > 1e-9).sum()))
            for k in ("dh", "mstack", "newh"):
                key = "first_step_%s_differs" % k
                if v[key] is None and not np.array_equal(ints[k], o.get(k)):
                    v[key] = it
            if v["first_step_n_gs_differs"] is None and int(o.scalar("n_gs")) != int(base.scalar("n_gs")):
                v["first_step_n_gs_differs"] = it
        print(it, {k: "%.2e" % v["max_rel_h"][-1] for k, v in res["variants"].items()}, flush=True)
        json.dump(res, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
