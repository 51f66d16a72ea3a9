"""Debug helper: run the oracle and the CUDA library for a few steps from the same state and
report, per intermediate array, how many entries differ.  (Needs a GPU.)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext
from oracle import pyoracle as po

ORDER = ["dh0", "dh", "indirection", "newh", "wh_mrec", "mrec", "mwrec", "mstack", "area", "max_slope", "kfs", "kdf",
         "h_spl", "lake_sill", "lake_silt", "lake_water_volume", "b_spl", "h_pre_diff", "kd_eff", "wh", "h", "b"]

def compare(o, c, verbose=True):
    bad = {}
    for name in ORDER:
        a, b = o.get(name), c.get(name)
        if a.dtype.kind == "f":
            neq = ~((a == b) | (np.isnan(a) & np.isnan(b)))
            # bitwise too (sign of zero)
            bits = (a.view(np.uint64 if a.dtype == np.float64 else np.uint32) != b.view(np.uint64 if a.dtype == np.float64 else np.uint32))
            nbad, nbits = int(neq.sum()), int(bits.sum())
        else:
            neq = a != b
            nbad = nbits = int(neq.sum())
        if nbits:
            bad[name] = (nbad, nbits)
            if verbose:
                idx = np.flatnonzero(neq if nbad else bits)[:5]
                print("  MISMATCH %-18s value-diff %d bit-diff %d first idx %s oracle %s gpu %s" % (name, nbad, nbits, idx, a[idx], b[idx]))
    for s in ("n_gs", "maxh", "boundary_len", "n_lakes"):
        if o.scalar(s) != c.scalar(s):
            bad[s] = (o.scalar(s), c.scalar(s))
            if verbose: print("  MISMATCH scalar", s, o.scalar(s), c.scalar(s))
    return bad

if __name__ == "__main__":
    lg = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    kind = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    f = synth.generate(lg, lg, kind=kind)
    o = po.Oracle(f)
    c = ErosionContext(lg, lg)
    c.configure(f)
    c.set_timing(True)
    for it in range(steps):
        # same starting state for both: take the oracle's
        c.set_state(o.get("h"), o.get("b"), o.get("wh"))
        t0 = time.time(); o.step(); t1 = time.time(); c.step(); t2 = time.time()
        print("step %d: oracle %.3fs gpu %.3fs n_gs %d/%d lakes %d" % (it, t1 - t0, t2 - t1, o.scalar("n_gs"), c.scalar("n_gs"), o.scalar("n_lakes")), flush=True)
        bad = compare(o, c)
        print("   ", "ALL EQUAL" if not bad else "differences: %s" % sorted(bad), flush=True)
        print("    gpu timers ms:", {k: round(v * 1e3, 2) for k, v in c.timers().items()}, flush=True)
