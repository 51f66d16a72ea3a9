"""Per-stage GPU timers for a few steps (needs a GPU)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 10
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
kind = int(sys.argv[3]) if len(sys.argv) > 3 else 0
c = ErosionContext(lg, lg); c.configure(synth.generate(lg, lg, kind=kind)); c.set_timing(True)
for kv in sys.argv[4:]:
    k, v = kv.split("="); c.set_debug(**{k: int(v)})
for it in range(steps):
    c.step()
    print(it, "n_gs", int(c.scalar("n_gs")), "lakes", int(c.scalar("n_lakes")), "lake_rounds", int(c.scalar("lake_rounds")), {k: round(v * 1e3, 2) for k, v in c.timers().items()}, flush=True)
