"""Full do_erosion schedule (WorldSim::generate: n_steps = 100, then the zero-step call that round-trips the
state through f32, mod.rs:1370-1421) through the public host API, wall-clock, GPU and optionally the CPU oracle.
usage: python tools/full_run.py LG STEPS KIND [cpu]"""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext
lg = int(sys.argv[1]); steps = int(sys.argv[2]); kind = int(sys.argv[3]); cpu = len(sys.argv) > 4
f = synth.generate(lg, lg, kind=kind)
n = (1 << lg) ** 2
t0 = time.perf_counter()
c = ErosionContext(lg, lg); c.configure(f)
h, b = c.run(steps)
h32, b32 = h.astype(np.float32), b.astype(np.float32)          # |posi| alt[posi] as f32 (mod.rs:1405-1406)
f2 = dict(f); f2["oldh"], f2["oldb"] = h32, b32
c.set_inputs(f2); h2, b2 = c.run(0)                             # second do_erosion call, n_small_steps = 0
tg = time.perf_counter() - t0
res = {"map": "%dx%d" % (1 << lg, 1 << lg), "kind": kind, "steps": steps, "gpu_wall_s": tg, "gpu_mcell_iters_per_s": n * steps / tg / 1e6,
       "f32_round_trip_ok": bool(np.array_equal(h2, h32.astype(np.float64)))}
if cpu:
    from oracle import pyoracle as po
    t0 = time.perf_counter()
    o = po.Oracle(f, n_threads=os.cpu_count()); ho, bo = o.run(steps)
    tc = time.perf_counter() - t0
    rel = np.abs(h - ho) / np.maximum(np.abs(ho), 1e-300)
    res.update(cpu_wall_s=tc, cpu_cores=os.cpu_count(), speedup=tc / tg, h_bit_identical=bool(np.array_equal(h, ho)),
               b_bit_identical=bool(np.array_equal(b, bo)), h_max_rel=float(np.nanmax(np.where(ho != 0, rel, 0))))
print(json.dumps(res), flush=True)
