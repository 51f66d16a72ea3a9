"""Per-stage GPU timers for several debug-knob variants in one process (needs a GPU).
usage: python tools/variants.py LG WARMUP STEPS "name:key=val,key=val" ...   (a bare "base" = defaults)"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext
lg, warm, steps = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
f = synth.generate(lg, lg)
for spec in sys.argv[4:]:
    name, _, kv = spec.partition(":")
    knobs = {k: int(v) for k, v in (p.split("=") for p in kv.split(",") if p)}
    c = ErosionContext(lg, lg); c.configure(f)
    if knobs:
        c.set_debug(**knobs)
    c.run_device(warm)
    ms = c.run_device_timed(steps)
    c.set_timing(True)
    acc = {}
    for _ in range(3):
        c.step()
        for k, v in c.timers().items():
            acc[k] = acc.get(k, 0.0) + v * 1e3 / 3
    print(json.dumps({"variant": name, "knobs": knobs, "ms_per_step": ms / steps, "n_gs_last": int(c.scalar("n_gs")), "lake_rounds": int(c.scalar("lake_rounds")),
                      "stage_ms_next3": {k: round(v, 2) for k, v in acc.items()}}), flush=True)
    c.close()
