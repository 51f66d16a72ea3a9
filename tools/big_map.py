"""Timing (and optional oracle parity) on larger maps: python tools/big_map.py LG STEPS [check]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext
lg = int(sys.argv[1]); steps = int(sys.argv[2]); check = len(sys.argv) > 3
t0 = time.time(); f = synth.generate(lg, lg); print("generate %.1fs" % (time.time() - t0), flush=True)
c = ErosionContext(lg, lg); c.configure(f); c.set_timing(True)
if check:
    from oracle import pyoracle as po
    o = po.Oracle(f)
for it in range(steps):
    t0 = time.time(); c.step(); tg = time.time() - t0
    msg = "step %d gpu %.3fs n_gs %d lakes %d" % (it, tg, int(c.scalar("n_gs")), int(c.scalar("n_lakes")))
    if check:
        t0 = time.time(); o.step(); tc = time.time() - t0
        bad = [k for k in ("dh", "newh", "mstack", "indirection", "h", "b", "wh", "area") if not np.array_equal(o.get(k).view(np.uint8), c.get(k).view(np.uint8))]
        msg += " cpu %.3fs speedup %.1fx %s" % (tc, tc / tg, "BIT-EXACT" if not bad else "DIFF %s" % bad)
    print(msg, {k: round(v * 1e3, 1) for k, v in c.timers().items()}, flush=True)
