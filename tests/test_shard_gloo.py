"""N > 1 host path on CPU: two gloo ranks shard tiles/frames, composite their rectangles and agree
on the max-over-ranks timing.  No compute kernels run here (there is no CPU fallback)."""
import os
import socket
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from grok_b200 import shard  # noqa: E402


def test_shard_units_partition():
    for n in (0, 1, 7, 64):
        for world in (1, 2, 3, 8):
            got = [shard.shard_units(n, r, world) for r in range(world)]
            flat = sorted(u for g in got for u in g)
            assert flat == list(range(n))
            assert max(len(g) for g in got) - min(len(g) for g in got) <= 1
    with pytest.raises(ValueError):
        shard.shard_units(4, 2, 2)


def test_shard_by_cost_balanced():
    rng = np.random.default_rng(5)
    costs = rng.integers(1, 1000, size=64).tolist()
    parts = shard.shard_by_cost(costs, 8)
    assert sorted(u for p in parts for u in p) == list(range(64))
    loads = [sum(costs[u] for u in p) for p in parts]
    assert max(loads) - min(loads) <= max(costs)


def test_tile_grid_covers_image():
    # 8192x8192 in 1024x1024 tiles (BASELINE configs[3]) and a ragged grid with offsets
    t = shard.tile_grid(0, 0, 8192, 8192, 0, 0, 1024, 1024)
    assert len(t) == 64 and t[0] == (0, 0, 1024, 1024) and t[-1] == (7168, 7168, 8192, 8192)
    t = shard.tile_grid(5, 3, 1000, 700, 2, 1, 256, 300)
    area = np.zeros((700, 1000), np.int32)
    for (x0, y0, x1, y1) in t:
        area[y0:y1, x0:x1] += 1
    assert np.all(area[3:, 5:] == 1) and area[:3].sum() == 0 and area[:, :5].sum() == 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert shard.dist_env() == (rank, rank, world)
        tiles = shard.tile_grid(0, 0, 300, 200, 0, 0, 128, 64)
        mine = shard.shard_units(len(tiles), rank, world)
        # stand-in for the decode of a tile: planes whose value names the tile
        img = [np.zeros((200, 300), np.int32) for _ in range(3)]
        for t in mine:
            x0, y0, x1, y1 = tiles[t]
            planes = [np.full((y1 - y0, x1 - x0), 100 * c + t + 1, np.int32) for c in range(3)]
            shard.composite(img, planes, tiles[t])
        np.save(os.path.join(out_dir, "img%d.npy" % rank), np.stack(img))
        shard.barrier(world)
        ms = shard.max_over_ranks(10.0 + rank, world)
        assert ms == 10.0 + world - 1
        total = shard.sum_over_ranks(float(len(mine)), world)
        assert total == len(tiles)
        # the bench's reductions keep their shape when a leg failed on one rank only (rank 1 has no second leg
        # and no tiled figure): no rank waits for a vector of another length
        legs = {"a": 100.0 + rank, "b": 50.0} if rank == 0 else {"a": 100.0 + rank, "b": None}
        extra = {"tiled": 7.0 + rank, "inflight": 3.0} if rank == 0 else {"tiled": None}
        times, link, lg, ex = shard.reduce_bench_figures(world, None, [10.0 + rank, 20.0 - rank], [5.0 + rank], legs,
                                                         ("a", "b"), extra, ("tiled", "inflight"))
        assert times == [10.0 + world - 1, 20.0] and link == [5.0]
        assert lg["a"] == (sum(100.0 + r for r in range(world)), world) and lg["b"] == (50.0, 1)
        assert ex["tiled"] == (7.0, 1) and ex["inflight"] == (3.0, 1)
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo(tmp_path):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    a = np.load(tmp_path / "img0.npy")
    b = np.load(tmp_path / "img1.npy")
    assert np.all((a == 0) | (b == 0))  # disjoint rectangles
    full = a + b
    tiles = shard.tile_grid(0, 0, 300, 200, 0, 0, 128, 64)
    for t, (x0, y0, x1, y1) in enumerate(tiles):
        for c in range(3):
            assert np.all(full[c, y0:y1, x0:x1] == 100 * c + t + 1)
