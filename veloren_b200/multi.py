"""Multi-GPU host logic (round 1: replicas only).

The erosion of ONE map does not shard without cross-GPU dataflow, so N GPUs erode
N independent worlds — the reference scales the same way (`batch_generate --threads k`,
world/examples/batch_generate.rs:287-352).  There is no data-path collective; the only
communication is a barrier and a max-reduce of the per-rank device time.
"""
import torch
import torch.distributed as dist



def band_of_cell(cell, n_cells, world):
    """Row-band ownership used by ve_mg_attach: cells are row-major, rank r owns the contiguous range
    [r * n/world, (r+1) * n/world), i.e. rows [r * ny/world, (r+1) * ny/world)."""
    per = n_cells // world
    return cell // per


def band_rows(rank, ny, world):
    per = ny // world
    return rank * per, (rank + 1) * per


BASE_SEED = 130626853  # DEFAULT_WORLD_SEED, world/src/sim/mod.rs:586


def seed_for_rank(rank):
    """Every rank erodes its own copy of the SAME world (BASE_SEED): with a seed per rank the step time
    varied 2x from world to world and the scaling figure measured the worlds, not the machine."""
    return BASE_SEED


def gather_objects(obj):
    """list of every rank's object on every rank (identity list when not initialised)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [obj]
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, obj)
    return out


def max_over_ranks(value, device="cpu"):
    """max of a python float over all ranks (identity when not initialised)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def aggregate_mcell_iters_per_s(cells_per_rank, steps, local_ms, device="cpu"):
    """Whole-job throughput: every rank did cells_per_rank*steps cell-iterations; the job took as
    long as its slowest rank."""
    world = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
    ms = max_over_ranks(local_ms, device)
    return cells_per_rank * steps * world / (ms * 1e-3) / 1e6, ms
