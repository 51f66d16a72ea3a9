"""world_size-2 gloo test of the multi-GPU host logic (replicas of one world, max-over-ranks
timing, whole-job throughput), on CPU."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from util import ROOT  # noqa: F401


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from veloren_b200 import multi, synth
    seed = multi.seed_for_rank(rank)
    f = synth.generate(4, 4, seed=seed)
    local_ms = 10.0 + 5.0 * rank                     # rank 1 is the slow one
    v, ms = multi.aggregate_mcell_iters_per_s(256, 2, local_ms)
    objs = multi.gather_objects({"rank": rank, "ms": local_ms})
    q.put((rank, seed, float(f["oldh"].sum()), v, ms, [o["ms"] for o in objs]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_replicas():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, s0, sum0, v0, ms0, g0), (r1, s1, sum1, v1, ms1, g1) = res
    assert s1 == s0 and sum0 == sum1                  # replicas of the same world on every rank
    assert g0 == g1 == [10.0, 15.0]                   # per-rank times reach every rank
    assert ms0 == ms1 == 15.0                         # max over ranks
    assert v0 == v1 == 256 * 2 * 2 / 15e-3 / 1e6      # whole-job aggregate


def test_single_process_identity():
    from veloren_b200 import multi
    v, ms = multi.aggregate_mcell_iters_per_s(1000, 3, 2.0)
    assert ms == 2.0 and v == 1000 * 3 / 2e-3 / 1e6


def test_row_band_ownership():
    from veloren_b200 import multi
    n, ny = 1 << 20, 1 << 10
    for world in (1, 2, 4, 8):
        per = n // world
        assert multi.band_of_cell(0, n, world) == 0 and multi.band_of_cell(n - 1, n, world) == world - 1
        for r in range(world):
            lo, hi = multi.band_rows(r, ny, world)
            assert multi.band_of_cell(lo * 1024, n, world) == r and multi.band_of_cell(hi * 1024 - 1, n, world) == r
            assert (hi - lo) * 1024 == per
