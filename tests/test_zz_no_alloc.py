"""rt_pack_tile_list_static — the entry point that allocates nothing (static device scratch; no counterpart in the
reference).  Collected last on purpose (file name).
Checks it against the existing rt_pack_tile_list through the C ABI on two calls in a row (the second with a
shorter, different list: the scratch keeps the first call's entries), for both buffers, bit for bit."""
import numpy as np
import pytest

import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi, tiles

pytestmark = pytest.mark.gpu


def test_pack_tile_list_static_equals_pack_tile_list_twice_in_a_row():
    import torch
    s = g.default_scene()
    w, h = 1000, 600  # 16 x 10 split tiles, ragged right and bottom edges
    ctx = g.Context(0, w, h)
    g.BufferManager(ctx).upload_scene(s)
    g.ComputeRenderer(ctx).render_frame(s.camera, mode=abi.MODE_PATH, spp=2, max_depth=3, base_seed=4,
                                        outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE)
    ctx.synchronize()
    n_tiles = 16 * 10
    rng = np.random.default_rng(5)
    first = rng.permutation(n_tiles).astype(np.uint32)             # every tile, shuffled (edge tiles included)
    second = np.array([159, 3, 15, 144, 77, 3], np.uint32)         # shorter, with a repeated tile
    for lst in (first, second):                                    # two calls in a row on the shared scratch
        for what, bpp in ((tiles.WHAT_COMPOSITE, 4), (tiles.WHAT_ACCUM, 16)):
            nbytes = len(lst) * tiles.TILE * tiles.TILE * bpp
            want = torch.full((nbytes,), 0xAB, dtype=torch.uint8, device="cuda")
            got = torch.full((nbytes,), 0xCD, dtype=torch.uint8, device="cuda")
            ctx.pack_tile_list(what, lst, want.data_ptr())          # existing entry point (allocating scratch)
            ctx.pack_tile_list_static(what, lst, got.data_ptr())    # static scratch, no allocation
            ctx.synchronize()
            assert torch.equal(got, want), (len(lst), what)
            assert int(got.to(torch.int64).sum()) > 0
    # and the packed tiles are the frame's: un-tile the first list on the host and compare with the composite
    comp = ctx.read_composite()
    buf = torch.zeros(len(first) * tiles.TILE * tiles.TILE * 4, dtype=torch.uint8, device="cuda")
    ctx.pack_tile_list_static(tiles.WHAT_COMPOSITE, first, buf.data_ptr())
    ctx.synchronize()
    frame = np.zeros_like(comp)
    tiles.unpack_tile_list_host(frame, buf.cpu().numpy().reshape(len(first), tiles.TILE, tiles.TILE, 4), list(first))
    assert np.array_equal(frame, comp)
    # limits and errors: beyond the static scratch -> a clear error, nothing launched
    launches = ctx.launch_count
    too_many = np.zeros(abi.STATIC_TILE_LIST_MAX + 1, np.uint32)
    with pytest.raises(g.RtError) as e:
        ctx.pack_tile_list_static(tiles.WHAT_COMPOSITE, too_many, buf.data_ptr())
    assert e.value.code == -1 and "static scratch" in str(e.value)
    with pytest.raises(g.RtError):
        ctx.pack_tile_list_static(tiles.WHAT_COMPOSITE, np.array([n_tiles], np.uint32), buf.data_ptr())  # not a tile of the frame
    with pytest.raises(g.RtError):
        ctx.pack_tile_list_static(2, second, buf.data_ptr())
    assert ctx.launch_count == launches
    ctx.pack_tile_list_static(tiles.WHAT_COMPOSITE, np.zeros(0, np.uint32), buf.data_ptr())  # empty list: no-op
    ctx.close()
