"""Manufactures the golden vectors under tests/golden/ from the CPU oracle.

The reference has NO golden vectors, known-answer tests or fixtures for the erosion path
 and cannot be built here, so these fixtures pin the *oracle* (and, through the
-m gpu tests, the CUDA path) against regressions; they do not pin the oracle to the reference.
Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from veloren_b200 import synth  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

KEEP = ["dh0", "dh", "indirection", "newh", "mrec", "mstack", "area", "wh_mrec", "h_spl", "lake_sill",
        "lake_water_volume", "kd_eff", "h", "b", "wh"]


def make(name, lg, kind, steps):
    f = synth.generate(lg, lg, kind=kind)
    o = po.Oracle(f)
    out = {}
    for s in range(steps):
        o.step()
        for k in KEEP:
            out["s%d_%s" % (s, k)] = o.get(k)
        out["s%d_n_gs" % s] = np.array([o.scalar("n_gs")])
    np.savez_compressed(os.path.join(HERE, name + ".npz"), lg=np.array([lg]), kind=np.array([kind]),
                        steps=np.array([steps]), **out)
    print(name, "written")


if __name__ == "__main__":
    make("erode_fbm_32", 5, 0, 3)
    make("erode_plateau_32", 5, 1, 2)
    make("erode_fbm_64", 6, 0, 2)
