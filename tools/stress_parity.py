"""Randomised parity sweep: seeds x sizes x kinds x scales, a few erode() steps each, every intermediate compared
bit for bit with the oracle (needs a GPU).  usage: python tools/stress_parity.py [n_cases] [rng_seed] [knob=value ...]
(knobs: ve_set_debug keys applied to every context, e.g. quad=1 lwv_small_log=2)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from util import compare_all
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext
from oracle import pyoracle as po

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 24
rng = np.random.RandomState(int(sys.argv[2]) if len(sys.argv) > 2 else 12345)
knobs = {k: int(v) for k, v in (a.split("=") for a in sys.argv[3:])}
bad_total = 0
t0 = time.time()
for case in range(n_cases):
    x_lg, y_lg = int(rng.randint(4, 10)), int(rng.randint(4, 10))
    kind = int(rng.randint(0, 2))
    seed = int(rng.randint(1, 2 ** 31 - 1))
    scale = float(rng.choice([0.5, 1.0, 2.0, 4.0]))
    steps = 3 if x_lg + y_lg <= 16 else 2
    f = synth.generate(x_lg, y_lg, seed=seed, kind=kind, scale=scale)
    o = po.Oracle(f, n_threads=os.cpu_count())
    c = ErosionContext(x_lg, y_lg); c.configure(f)
    if knobs:
        c.set_debug(**knobs)
    status = "ok"
    for it in range(steps):
        o.step(); c.step()
        bad = compare_all(o, c)
        if bad or int(o.scalar("n_gs")) != int(c.scalar("n_gs")):
            status = "DIFF at step %d: %s" % (it, sorted(bad)[:6]); bad_total += 1
            break
    print("case %2d: %4dx%-4d kind %d seed %10d scale %.1f lakes %6d n_gs %2d: %s" %
          (case, 1 << x_lg, 1 << y_lg, kind, seed, scale, int(c.scalar("n_lakes")), int(c.scalar("n_gs")), status), flush=True)
print("%d cases, %d with differences, %.0f s" % (n_cases, bad_total, time.time() - t0))
sys.exit(1 if bad_total else 0)
