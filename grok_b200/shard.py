"""Host-side sharding of independent units (tiles of an image, frames of a batch) over the GPUs of
one box.  Tiles and frames are coded independently (reference: one TileProcessor per tile,
src/lib/jp2/codestream/CodeStreamDecompress.cpp:945-955 composites finished tiles into the output
image), so there is no data-path collective: a rank decodes its own units and writes its own
rectangles.  torch.distributed is used only for the barrier and the max-over-ranks of timings.
"""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple


def dist_env() -> Tuple[int, int, int]:
    """(rank, local_rank, world_size) from the torchrun environment (1 process per GPU)."""
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return rank, local, world


def bind_to_device(local_rank: int) -> bool:
    """grkb_bind_host_to_device: keep this rank's threads (and the page-locked buffers it is about
    to allocate) on the NUMA node of its GPU.  GRKB_NO_NUMA_BIND=1 leaves the affinity alone."""
    if os.environ.get("GRKB_NO_NUMA_BIND") == "1":
        return False
    from ._lib import lib
    return lib.grkb_bind_host_to_device(int(local_rank)) == 0


def shard_units(n_units: int, rank: int, world: int) -> List[int]:
    """Unit u goes to rank u mod world (tile t -> GPU t mod G)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world %d/%d" % (rank, world))
    return list(range(rank, n_units, world))


def shard_by_cost(costs: Sequence[int], world: int) -> List[List[int]]:
    """Longest-processing-time-first assignment of units with known cost (compressed size of a
    tile) to `world` ranks; deterministic, so every rank computes the same table."""
    order = sorted(range(len(costs)), key=lambda i: (-int(costs[i]), i))
    load = [0] * world
    out: List[List[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += int(costs[i])
    for lst in out:
        lst.sort()
    return out


def tile_grid(x0: int, y0: int, x1: int, y1: int, tx0: int, ty0: int, tw: int, th: int):
    """Tile rectangles of an image area on the reference grid (ISO 15444-1 B.3; reference
    CodingParams::getTileBounds, src/lib/jp2/codestream/CodingParams.cpp): tile (p, q) covers
    [max(tx0 + p*tw, x0), min(tx0 + (p+1)*tw, x1)) x likewise, row-major tile index."""
    if tw <= 0 or th <= 0:
        raise ValueError("empty tile size")
    nx = -(-(x1 - tx0) // tw)
    ny = -(-(y1 - ty0) // th)
    tiles = []
    for q in range(ny):
        for p in range(nx):
            rx0, ry0 = max(tx0 + p * tw, x0), max(ty0 + q * th, y0)
            rx1, ry1 = min(tx0 + (p + 1) * tw, x1), min(ty0 + (q + 1) * th, y1)
            tiles.append((rx0, ry0, rx1, ry1))
    return tiles


def composite(dst, tile_planes, rect, origin=(0, 0)):
    """Write one decoded tile (list of 2-D arrays, one per component) into the image planes
    `dst` (list of 2-D arrays) at its canvas rectangle."""
    x0, y0, x1, y1 = rect
    ox, oy = origin
    for c, p in enumerate(tile_planes):
        dst[c][y0 - oy:y1 - oy, x0 - ox:x1 - ox] = p[:y1 - y0, :x1 - x0]


def barrier(world: int):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def max_over_ranks(value: float, world: int, device=None) -> float:
    """Timing rule of the bench contract: the slowest rank's time."""
    if world <= 1:
        return float(value)
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float, world: int, device=None) -> float:
    if world <= 1:
        return float(value)
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def reduce_bench_figures(world: int, device, times_ms: Sequence[float], link_gbs: Sequence[float],
                         leg_values: dict, leg_names: Sequence[str], extra_ms: dict, extra_names: Sequence[str]):
    """The reductions of a bench run over the ranks, with shapes that do not depend on what happened on a
    rank (a leg that failed somewhere must not leave the ranks reducing vectors of different lengths):
    times_ms -> max, link_gbs -> min, leg_values[name] (throughputs; missing or None: 0) -> sum with the
    number of ranks that have one, extra_ms[name] (times; missing or None: 0) -> max with that count.
    Returns (times, link, {name: (sum, ranks)}, {name: (max, ranks)})."""
    lv = [float(leg_values.get(k) or 0.0) for k in leg_names]
    ev = [float(extra_ms.get(k) or 0.0) for k in extra_names]
    have = [1.0 if v > 0 else 0.0 for v in lv + ev]
    times, link = [float(x) for x in times_ms], [float(x) for x in link_gbs]
    if world > 1:
        import torch
        import torch.distributed as dist
        t = torch.tensor(times + ev, dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        got = [float(x) for x in t]
        times, ev = got[:len(times)], got[len(times):]
        t = torch.tensor(link, dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        link = [float(x) for x in t]
        t = torch.tensor(lv + have, dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        got = [float(x) for x in t]
        lv, have = got[:len(lv)], got[len(lv):]
    legs = {k: (lv[i], int(have[i])) for i, k in enumerate(leg_names)}
    extras = {k: (ev[i], int(have[len(lv) + i])) for i, k in enumerate(extra_names)}
    return times, link, legs, extras
