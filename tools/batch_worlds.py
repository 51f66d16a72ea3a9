"""k independent worlds eroded concurrently on ONE GPU (the batch_generate use-case, world/examples/batch_generate.rs):
the sweeps are latency-bound and leave most of the SMs' issue slots idle, so several contexts, each limited to a
share of the resident blocks (ve_set_debug "df_bps"), overlap.  usage: python tools/batch_worlds.py K BPS [LG] [STEPS]"""
import sys, os, time, threading
k = int(sys.argv[1]); bps = sys.argv[2]
lg = int(sys.argv[3]) if len(sys.argv) > 3 else 10
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 10
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext
ctxs = []
for i in range(k):
    c = ErosionContext(lg, lg); c.configure(synth.generate(lg, lg, seed=130626853 + 7919 * i))
    if bps != "0":
        c.set_debug(df_bps=int(bps))
    c.run_device(3)
    ctxs.append(c)
def work(c):
    c.run_device(steps)
t0 = time.perf_counter()
ths = [threading.Thread(target=work, args=(c,)) for c in ctxs]
for t in ths: t.start()
for t in ths: t.join()
dt = time.perf_counter() - t0
n = (1 << lg) ** 2
print("worlds %d, VE_DF_BPS %s: %d steps each in %.3f s -> %.1f ms per step per world, aggregate %.2f Mcell-iters/s" %
      (k, bps, steps, dt, dt / steps * 1e3, k * n * steps / dt / 1e6))
