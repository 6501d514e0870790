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
