"""One map sharded as row bands over the ranks of a torchrun job (ve_mg_attach): per-step time and per-stage timers.
usage: torchrun --nproc-per-node N tools/mg_scale.py LG STEPS [check]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from veloren_b200 import synth
from veloren_b200.erosion import ErosionContext
lg, steps = int(sys.argv[1]), int(sys.argv[2])
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("gloo")
f = synth.generate(lg, lg)
c = ErosionContext(lg, lg, device=local); c.configure(f)
if world > 1:
    c.attach_multi(rank, world)
c.step()                                     # warm-up (allocations, first-touch)
c.set_timing(True)
acc, tot = {}, 0.0
for it in range(steps):
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter(); c.step(); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    tot += dt
    for k, v in c.timers().items():
        acc[k] = acc.get(k, 0.0) + v * 1e3 / steps
up = (c.scalar("up_ms_total"), c.scalar("up_launches_total"))
ups = [up]
if world > 1:
    ups = [None] * world; dist.all_gather_object(ups, up)
if rank == 0:
    print("up-sweep kernel ms per launch, per rank:", [round(a / max(b, 1), 2) for a, b in ups], flush=True)
    n = 1 << (2 * lg)
    print(json.dumps({"map": "%dx%d" % (1 << lg, 1 << lg), "n_gpus": world, "steps": steps, "ms_per_step": tot / steps * 1e3,
                      "mcell_iters_per_s": n * steps / tot / 1e6, "n_gs_last": int(c.scalar("n_gs")),
                      "stage_ms": {k: round(v, 1) for k, v in acc.items()}, "checksum_h": float(c.get("h").sum())}), flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
