"""world_size-2 (and 3) gloo tests on CPU of the tile split + gather-to-rank-0 logic
(gpu_raytracer_b200/tiles.py).  The device copies are replaced by their numpy
statements; the frame content comes from the CPU oracle."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import tiles


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, w, h, what, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(1234)  # same "rendered frame" on every rank
    bpp = tiles.BYTES_PER_PIXEL[what]
    full = rng.integers(0, 255, (h, w, bpp), dtype=np.uint8)
    # this rank only "renders" its own tiles
    mine = np.zeros_like(full)
    for t in tiles.tiles_of_rank(w, h, world, rank):
        ox, oy, tw, th = tiles.tile_rect(w, h, t)
        mine[oy:oy + th, ox:ox + tw] = full[oy:oy + th, ox:ox + tw]
    tg = tiles.TileGather(dist, torch, w, h, what, "cpu")

    def pack(dst):
        dst.copy_(torch.from_numpy(tiles.pack_tiles_host(mine, world, rank).reshape(-1)))

    def unpack(src, r):
        comp = src.numpy().reshape(tiles.tiles_per_rank_padded(w, h, world), tiles.TILE, tiles.TILE, bpp)
        tiles.unpack_tiles_host(mine, comp, world, r)

    tg.run(pack, unpack)
    if rank == 0:
        q.put(bool(np.array_equal(mine, full)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h,what", [(2, 1280, 720, tiles.WHAT_COMPOSITE), (3, 300, 200, tiles.WHAT_ACCUM),
                                            (2, 128, 128, tiles.WHAT_COMPOSITE)])
def test_tile_gather_reassembles_frame(world, w, h, what):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, w, h, what, q)) for r in range(world)]
    for p in procs:
        p.start()
    ok = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok


def test_tile_partition_covers_frame_once():
    for (w, h) in ((1280, 720), (1920, 1080), (3840, 2160), (100, 50)):
        tx, ty = tiles.tile_grid(w, h)
        for world in (1, 2, 4, 8):
            seen = sorted(t for r in range(world) for t in tiles.tiles_of_rank(w, h, world, r))
            assert seen == list(range(tx * ty))
            assert sum(tiles.pixels_of_rank(w, h, world, r) for r in range(world)) == w * h
            assert max(len(tiles.tiles_of_rank(w, h, world, r)) for r in range(world)) == tiles.tiles_per_rank_padded(w, h, world)
    assert tiles.tile_grid(1920, 1080) == (30, 17) and tiles.tile_rect(1920, 1080, 509) == (1856, 1024, 64, 56)


def _cost_model(w, h, tile_list):
    """A synthetic per-rank frame time: fixed part + a per-tile cost that grows down the image."""
    tx, _ = tiles.tile_grid(w, h)
    return 0.7 + sum(0.015 + 0.02 * (t // tx) / 17.0 + (0.01 if t % 8 == 5 else 0.0) for t in tile_list)


def _balance_worker(rank, world, port, w, h, q):
    """The bench's balancing loop on gloo: every rank all-gathers the per-rank times, runs tiles.rebalance on the
    same numbers and must end up with the same lists; then the list-based gather reassembles the frame."""
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lists = tiles.initial_lists(w, h, world)
    spread = []
    for _ in range(4):
        t = torch.tensor([_cost_model(w, h, lists[rank])], dtype=torch.float64)
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        times = [float(x.item()) for x in allt]
        spread.append(max(times) / min(times))
        lists = tiles.rebalance(lists, times)
    full = np.random.default_rng(7).integers(0, 255, (h, w, 4), dtype=np.uint8)
    mine = np.zeros_like(full)
    tiles.unpack_tile_list_host(mine, tiles.pack_tile_list_host(full, lists[rank]), lists[rank])
    tg = tiles.TileGather(dist, torch, w, h, tiles.WHAT_COMPOSITE, "cpu", lists=lists)
    n_pad = max(len(l) for l in lists)

    def pack(dst):
        dst.copy_(torch.from_numpy(tiles.pack_tile_list_host(mine, lists[rank], n_pad).reshape(-1)))

    def unpack(src, r):
        tiles.unpack_tile_list_host(mine, src.numpy().reshape(n_pad, tiles.TILE, tiles.TILE, 4), lists[r])

    tg.run(pack, unpack)
    agree = [None] * world
    dist.all_gather_object(agree, lists)
    if rank == 0:
        q.put((bool(np.array_equal(mine, full)), all(a == lists for a in agree), spread,
               sorted(t for l in lists for t in l) == list(range(len(full[::tiles.TILE, ::tiles.TILE].reshape(-1, 4))))))
    dist.barrier()
    dist.destroy_process_group()


def test_tile_lists_rebalance_and_gather():
    world, w, h = 2, 1920, 1080
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_balance_worker, args=(r, world, port, w, h, q)) for r in range(world)]
    for p in procs:
        p.start()
    ok, agree, spread, partition = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok and agree and partition
    assert spread[-1] <= spread[0] + 1e-9


def test_rebalance_is_a_partition_and_converges():
    w, h, world = 1920, 1080, 8
    lists = tiles.initial_lists(w, h, world)
    first = None
    for it in range(6):
        times = [_cost_model(w, h, l) for l in lists]
        if first is None:
            first = max(times)
        lists = tiles.rebalance(lists, times)
        assert sorted(t for l in lists for t in l) == list(range(510))
    times = [_cost_model(w, h, l) for l in lists]
    assert max(times) < first and max(times) / min(times) < 1.03
    assert sum(tiles.pixels_of_tiles(w, h, l) for l in lists) == w * h
