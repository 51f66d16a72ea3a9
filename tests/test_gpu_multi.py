"""-m gpu, needs two GPUs (skipped otherwise): one map sharded as row bands over two ranks (ve_mg_attach) must end
every step bit-identical to the single-GPU result, on every rank."""
import os
import socket

import numpy as np
import pytest

from util import ROOT  # noqa: F401

pytestmark = pytest.mark.gpu
ARRS = ["h", "b", "wh", "dh", "newh", "mstack", "area", "h_spl", "lake_sill", "lake_silt", "lake_water_volume", "kd_eff"]


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, lg, kind, steps, q):
    import torch
    import torch.distributed as dist
    from veloren_b200 import synth
    from veloren_b200.erosion import ErosionContext
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    f = synth.generate(lg, lg, kind=kind)
    c = ErosionContext(lg, lg, device=rank)
    c.configure(f)
    c.attach_multi(rank, world)
    single = None
    if rank == 0:
        single = ErosionContext(lg, lg, device=0)
        single.configure(f)
    bad = []
    for it in range(steps):
        c.step()
        if rank == 0:
            single.step()
            for name in ARRS:
                a, b = single.get(name), c.get(name)
                if a.tobytes() != b.tobytes():
                    bad.append("step %d %s" % (it, name))
            if int(single.scalar("n_gs")) != int(c.scalar("n_gs")):
                bad.append("step %d n_gs" % it)
    h = c.get("h")
    sums = [None] * world
    dist.all_gather_object(sums, float(h.sum()))
    q.put((rank, bad, sums))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("lg,kind,steps", [(8, 0, 3), (9, 1, 2)])
def test_two_gpu_row_bands_bit_identical(lg, kind, steps):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, lg, kind, steps, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=600) for _ in range(2))
    for p in ps:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res[0][1] == [], res[0][1]
    assert res[0][2][0] == res[0][2][1]          # every rank ends with the same complete state
