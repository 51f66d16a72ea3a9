#!/usr/bin/env python
"""bench.py — erosion throughput (Mcell-iters/s) of the B200 path, per the driver's contract.

  python bench.py --gpus N --steps K --warmup W            # our CUDA path
  python bench.py --impl reference --gpus N --steps K ...   # the CPU arm (oracle on host cores)

A "step" is one erode() (world/src/sim/erosion.rs:699) over the whole map.  Workload =
BASELINE.json configs[1]: the default 1024x1024 world (MapSizeLg 10x10) with the synthetic fields
of veloren_b200.synth (reference recipes, in-repo noise).  `value` = N_cells * K / device time,
state resident in HBM; `e2e` = the same through the public call with host buffers (inputs H2D,
results D2H inside the timed region).  With --gpus N>1 every rank erodes its own copy of the SAME
world (replicas: the reference scales the same way, world/examples/batch_generate.rs runs k
independent generations); no data-path collective, weak scaling, value = sum over ranks / max time.
The reference arm at N>1 runs N oracle instances concurrently on the host cores (cores/N OpenMP
threads each), i.e. the same N-world job.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

X_LG = Y_LG = 10
# algorithmic bytes per cell of the dominant kernel (G4, the implicit up-sweep)
G4_BYTES_PER_CELL = 145
FIXED_BYTES_PER_CELL, GS_BYTES_PER_CELL = 684, 374


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per up-sweep launch from the committed ncu --set full capture
    (profiles/r02_ncu_traffic.json); None when no capture of the current kernel exists."""
    p = os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")
    try:
        d = json.load(open(p))
        return d["OpGsUp"]["dram_bytes_per_launch"], d["OpGsUp"].get("source", p)
    except Exception:
        return None, None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled during the timed region (recipe's clocks line)."""

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def run_reference(args):
    """CPU arm: the oracle (restatement of the reference; the reference itself is Rust and cannot be
    built in this image) on all host cores, same config/metric.  Rank 0 only.  At N>1 the job is N
    copies of the world, so N oracle instances run concurrently, cores/N OpenMP threads each."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    from veloren_b200 import synth
    from oracle import pyoracle as po
    nw = max(1, args.gpus)
    cores = os.cpu_count()
    per = max(1, cores // nw)
    fields = synth.generate(X_LG, Y_LG)
    oracles = [po.Oracle(fields, n_threads=per) for _ in range(nw)]   # torchrun exports OMP_NUM_THREADS=1: ask explicitly
    n = oracles[0].n

    def steps(k):
        if nw == 1:
            for _ in range(k):
                oracles[0].step()
            return
        ths = [threading.Thread(target=lambda o=o: [o.step() for _ in range(k)]) for o in oracles]
        for t in ths:
            t.start()
        for t in ths:
            t.join()

    steps(args.warmup)
    t0 = time.perf_counter()
    steps(args.steps)
    dt = time.perf_counter() - t0
    v = n * nw * args.steps / dt / 1e6
    line = {"impl": "reference", "metric": "erosion_mcell_iters_per_s", "value": v, "unit": "Mcell-iters/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(nw),
            "cpu_baseline": {"value": v, "unit": "Mcell-iters/s", "cores": cores, "kind": "port",
                             "sample": "%d erode() steps of %d copy(ies) of the 1024x1024 world after %d warm-up steps, C++ oracle "
                                       "(OpenMP where the reference uses rayon; %d threads per world); parity with the Rust "
                                       "reference unpinned" % (args.steps, nw, args.warmup, per)},
            "e2e": {"value": v, "unit": "Mcell-iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(n_gpus):
    return {"workload": "default 1024x1024 world (MapSizeLg 10x10), erode() steps of the WorldSim::generate schedule",
            "x_lg": X_LG, "y_lg": Y_LG, "seed": 130626853, "scale": 2.0,
            "parallelism": "replicas: %d independent copies of the same world, one per GPU, no data-path collective" % n_gpus,
            "l2": "working set ~0.7 GB per world > 126 MB L2 (no flush needed)"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from veloren_b200 import synth
    from veloren_b200.erosion import ErosionContext
    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ["NCCL_DEBUG"] = "WARN"            # keep NCCL's version banner off stdout: one JSON line only
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # every rank erodes its own copy of the same world (same seed: efficiency then measures the machine,
    # not the world), inputs resident before timing
    from veloren_b200 import multi
    fields = synth.generate(X_LG, Y_LG, seed=multi.seed_for_rank(rank))
    n = (1 << X_LG) * (1 << Y_LG)
    ctx = ErosionContext(X_LG, Y_LG, device=local)
    ctx.configure(fields)
    ctx.run_device(args.warmup)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    l0 = ctx.launch_count()
    up_ms0, up_n0 = ctx.scalar("up_ms_total"), ctx.scalar("up_launches_total")
    ms = ctx.run_device_timed(args.steps)        # CUDA events on the library's stream around exactly K steps
    gs_counts, lake_counts = ctx.history(args.steps)
    launches = ctx.launch_count() - l0
    up_ms = ctx.scalar("up_ms_total") - up_ms0
    up_n = int(ctx.scalar("up_launches_total") - up_n0)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    value, ms_max = multi.aggregate_mcell_iters_per_s(n, args.steps, ms, device="cuda")
    per_rank = multi.gather_objects({"rank": rank, "ms_per_step": ms / args.steps, "n_gs": gs_counts, "n_lakes": lake_counts})

    # ---- roofline of the dominant kernel (G4 up-sweep): CUDA events recorded around every launch of the
    # timed steps themselves, on the stream it is launched on ----------------------------------------
    peak, how = peaks()
    per_launch = up_ms * 1e-3 / max(up_n, 1)
    achieved = G4_BYTES_PER_CELL * n / per_launch / 1e9
    traffic, traffic_src = ncu_traffic()
    roofline = {"bound": "hbm", "kernel": "k_sweep_warp<OpGsUp> (implicit stream-power up-sweep, erosion.rs:1237-1450)",
                "achieved": achieved, "peak": peak, "peak_source": how, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_src, "avg_launch_ms": per_launch * 1e3,
                "launches_timed": up_n, "share_of_step": up_ms / ms if ms > 0 else None,
                "algorithmic_bytes_per_launch": G4_BYTES_PER_CELL * n,
                "whole_step_bytes": (FIXED_BYTES_PER_CELL + GS_BYTES_PER_CELL * (sum(gs_counts) / max(len(gs_counts), 1))) * n,
                "note": "latency-bound by the dependency depth of the receiver DAG at 1024^2 (DESIGN.md section 7)"}

    # ---- e2e: public call with host buffers (pinned), H2D inputs + K steps + D2H results -----
    pinned = {}
    for k in ("oldh", "oldb", "is_ocean", "uplift", "n", "m", "g", "epsilon_0", "alpha", "kf", "kd", "height_scale_cell"):
        tns = torch.from_numpy(np.ascontiguousarray(fields[k])).pin_memory()
        pinned[k] = tns.numpy()
    out_h = torch.empty(n, dtype=torch.float64).pin_memory().numpy()
    out_b = torch.empty(n, dtype=torch.float64).pin_memory().numpy()
    h2d = sum(a.nbytes for a in pinned.values())
    d2h = out_h.nbytes + out_b.nbytes
    e2e_steps = args.steps
    barrier()
    t0 = time.perf_counter()
    ctx.set_inputs(pinned)
    ctx.run(e2e_steps, out_h=out_h, out_b=out_b)
    torch.cuda.synchronize()
    te = time.perf_counter() - t0
    e2e_v, _ = multi.aggregate_mcell_iters_per_s(n, e2e_steps, te * 1e3, device="cuda")
    e2e = {"value": e2e_v, "unit": "Mcell-iters/s", "h2d_bytes_per_step": h2d / e2e_steps, "d2h_bytes_per_step": d2h / e2e_steps,
           "note": "ErosionContext.set_inputs + run(K) from the initial state: inputs H2D once per call, (h,b) D2H at the end",
           "checksum_h": float(out_h.sum())}

    # ---- the metric's second half: full do_erosion wall time (100 steps + the zero-step f32 round trip of the
    # second call, mod.rs:1370-1421) through the public host API, copies and context creation included ----------
    full = None
    if not args.no_full:
        barrier()
        t0 = time.perf_counter()
        c2 = ErosionContext(X_LG, Y_LG, device=local)
        c2.configure(pinned | {k: fields[k] for k in ("k_d_scale", "k_da", "noise")})
        hf, bf = c2.run(100)
        f2 = dict(pinned); f2["oldh"], f2["oldb"] = hf.astype(np.float32), bf.astype(np.float32)
        c2.set_inputs(f2)
        h2, _ = c2.run(0)
        torch.cuda.synchronize()
        tfull = time.perf_counter() - t0
        c2.close()
        tfull_max = multi.max_over_ranks(tfull, device="cuda")
        full = {"wall_s": tfull_max, "steps": 100, "mcell_iters_per_s": n * 100 * world / tfull_max / 1e6,
                "f32_round_trip_ok": bool(np.array_equal(h2, hf.astype(np.float32).astype(np.float64))),
                "checksum_h": float(hf.sum())}

    # ---- N > 1: BASELINE configs[2], ONE 2048x2048 world sharded as row bands over the N ranks (ve_mg_attach; strong
    # scaling), next to the same world on one GPU measured by rank 0 in the same run ---------------------------------
    row_bands = None
    if world > 1 and not args.no_row_bands:
        ctl = dist.new_group(backend="gloo")              # control plane (IPC handles, barriers); data moves over NVLink
        f3 = synth.generate(11, 11)
        c3 = ErosionContext(11, 11, device=local)
        c3.configure(f3)
        c3.attach_multi(rank, world, group=ctl)
        c3.run_device(3)
        barrier()
        ms3 = multi.max_over_ranks(c3.run_device_timed(5), device="cuda")
        chk3 = float(c3.get("h").sum())
        c3.close()
        barrier()
        ms31 = None
        if rank == 0:
            c31 = ErosionContext(11, 11, device=local)
            c31.configure(f3)
            c31.run_device(3)
            ms31 = c31.run_device_timed(5)
            same = float(c31.get("h").sum()) == chk3
            c31.close()
            row_bands = {"workload": "BASELINE configs[2]: one 2048x2048 world, row bands over %d GPUs" % world, "scaling": "strong",
                         "steps": 5, "warmup": 3, "ms_per_step": ms3 / 5, "mcell_iters_per_s": (1 << 22) * 5 / (ms3 * 1e-3) / 1e6,
                         "one_gpu_ms_per_step": ms31 / 5, "speedup_vs_one_gpu": ms31 / ms3, "identical_to_one_gpu": same,
                         "sharded": "Gauss-Seidel sweeps (erosion.rs:1127-1151, 1237-1450): owner-computes per band, other bands' "
                                    "words polled over NVLink (CUDA IPC peer memory)",
                         "replicated": "every other stage (dependency-depth bound at this size: DESIGN.md section 8)"}
        barrier()

    # ---- CPU baseline on the box's host cores (rank 0, N=1 only): bounded sample ---------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import pyoracle as po
        o = po.Oracle(fields, n_threads=os.cpu_count())
        o.step()
        t0 = time.perf_counter()
        cs = 3
        for _ in range(cs):
            o.step()
        dt = time.perf_counter() - t0
        cpu = {"value": n * cs / dt / 1e6, "unit": "Mcell-iters/s", "cores": os.cpu_count(), "kind": "port",
               "sample": "%d erode() steps (after 1 warm-up) of the same 1024x1024 world, C++ oracle, OpenMP on all host "
                         "cores where the reference uses rayon, serial elsewhere" % cs,
               "ms_per_step": dt / cs * 1e3}
    if rank == 0:
        line = {"metric": "erosion_mcell_iters_per_s", "value": value, "unit": "Mcell-iters/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(world), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                "gpu_launches": int(launches), "clocks": clocks, "per_rank": per_rank,
                "full_do_erosion": full, "row_bands": row_bands,
                "full_do_erosion_estimate_s": ms_max / args.steps * 100 / 1e3}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-full", action="store_true", help="skip the measured 100-step do_erosion run")
    ap.add_argument("--no-row-bands", action="store_true", help="N > 1: skip the sharded 2048x2048 measurement")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
