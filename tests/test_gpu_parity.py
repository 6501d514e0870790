"""GPU parity tests proper: the CUDA path, called through the C ABI, against the
CPU oracle on the same inputs.  Bit-exact for hit records (kind, primitive id, t
bits, material) and rgba8 images; radiance tolerance stated below (target and
observed: bit-exact)."""
import numpy as np
import pytest

import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi
import oracle_binding as ob

pytestmark = pytest.mark.gpu

# north_star tolerance for fp32 radiance (depth 0): max-abs <= 1e-6, RMSE <= 1e-7 per channel.
MAX_ABS_TOL = 1e-6
RMSE_TOL = 1e-7


def check_radiance(got, want):
    d = got.astype(np.float64) - want.astype(np.float64)
    assert np.isnan(got).sum() == np.isnan(want).sum()
    d = np.nan_to_num(d)
    assert np.abs(d).max() <= MAX_ABS_TOL, np.abs(d).max()
    for c in range(3):
        assert np.sqrt((d[..., c] ** 2).mean()) <= RMSE_TOL


def check_hits(got, want):
    for f in ("kind", "prim", "t_bits", "material_id"):
        bad = got[f] != want[f]
        assert not bad.any(), (f, int(bad.sum()), np.argwhere(bad)[:5], got[bad][:5], want[bad][:5])


def render_gpu_fused(scene, w, h, mode=abi.MODE_LEGACY, traversal=abi.TRAVERSAL_AUTO):
    ctx = g.Context(0, w, h)
    g.BufferManager(ctx).upload_scene(scene)
    cr = g.ComputeRenderer(ctx)
    cr.render_frame(scene.camera, mode=mode, traversal=traversal,
                    outputs=abi.OUT_ACCUM | abi.OUT_HITS | abi.OUT_CHANNELS | abi.OUT_COMPOSITE)
    out = dict(channels=[ctx.read_channel(i) for i in range(3)], composite=ctx.read_composite(),
               accum=ctx.read_accum(), hits=ctx.read_hits(), traversal=ctx.last_traversal)
    ctx.close()
    return out


def compare(gpu, ref):
    check_hits(gpu["hits"], ref["hits"])
    for c in range(3):
        assert np.array_equal(gpu["channels"][c], ref["channels"][c]), c
    assert np.array_equal(gpu["composite"], ref["composite"])
    check_radiance(gpu["accum"][..., :3], ref["radiance"])
    assert np.array_equal(gpu["accum"][..., :3].view(np.uint32), ref["radiance"].view(np.uint32)), "not bit-exact"


@pytest.fixture(scope="module")
def default_ref():
    s = g.default_scene()
    return s, ob.render(s, 1280, 720)


@pytest.mark.parametrize("traversal", [abi.TRAVERSAL_LITERAL, abi.TRAVERSAL_ORDERED])
def test_c1_default_scene_fused(default_ref, traversal):
    s, ref = default_ref
    gpu = render_gpu_fused(s, 1280, 720, traversal=traversal)
    assert gpu["traversal"] == traversal
    compare(gpu, ref)


def test_c1_default_scene_run_compute_progressive(default_ref):
    """The reference's own loop: run_compute until every tile is done (9 frames at 720p)."""
    s, ref = default_ref
    ctx = g.Context(0, 1280, 720)
    g.BufferManager(ctx).upload_scene(s)
    cr = g.ComputeRenderer(ctx)
    frames, pending = 0, cr.run_compute(s.camera, True)
    frames += 1
    while pending > 0:
        pending = cr.run_compute(s.camera, False)
        frames += 1
    assert frames == 9  # 60 tiles, 7 per frame (shared/src/lib.rs:1195-1203)
    for c in range(3):
        assert np.array_equal(ctx.read_channel(c), ref["channels"][c])
    check_hits(ctx.read_hits(), ref["hits"])
    assert np.array_equal(ctx.read_accum()[..., :3].view(np.uint32), ref["radiance"].view(np.uint32))
    ctx.close()


def test_c1_wavefront_depth0(default_ref):
    s, _ = default_ref
    ref = ob.render(s, 640, 360, wavefront=(0, 4, 123))
    gpu = render_gpu_fused(s, 640, 360, mode=abi.MODE_WAVEFRONT0)
    compare(gpu, ref)


@pytest.mark.parametrize("traversal", [abi.TRAVERSAL_LITERAL, abi.TRAVERSAL_ORDERED])
def test_c2_terrain_100k_1080p(traversal):
    s = g.terrain_scene(250, 200)
    assert len(s.triangles) == 100000 and len(s.bvh_nodes) == 199999
    ref = ob.render(s, 1920, 1080)
    gpu = render_gpu_fused(s, 1920, 1080, traversal=traversal)
    assert gpu["traversal"] == traversal
    compare(gpu, ref)
    assert (ref["hits"]["kind"] == 2).mean() > 0.9


@pytest.fixture(scope="module")
def c3_scene_and_ref():
    s = g.terrain_scene(1000, 500)
    assert len(s.triangles) == 1000000 and len(s.bvh_nodes) == 20005
    return s, ob.render(s, 480, 270)


@pytest.mark.parametrize("traversal", [abi.TRAVERSAL_LITERAL, abi.TRAVERSAL_ORDERED, abi.TRAVERSAL_REBUILT,
                                       abi.TRAVERSAL_AUTO])
def test_c3_terrain_1m_chunked_bvh(c3_scene_and_ref, traversal):
    """1M-triangle terrain on the reference's chunked BVH (100 triangles per leaf): every
    traversal policy returns the oracle's hit records bit for bit."""
    s, ref = c3_scene_and_ref
    gpu = render_gpu_fused(s, 480, 270, traversal=traversal)
    assert gpu["traversal"] == (abi.TRAVERSAL_REBUILT if traversal == abi.TRAVERSAL_AUTO else traversal)
    compare(gpu, ref)


def test_c3_full_size_policies_agree_1080p():
    """At BASELINE.json's full size the oracle is too slow for a unit test; the
    size-independent property is that the three policies agree with each other
    (ORDERED and LITERAL are pinned to the oracle above)."""
    s = g.terrain_scene(1000, 500)
    outs = [render_gpu_fused(s, 1920, 1080, traversal=t) for t in (abi.TRAVERSAL_ORDERED, abi.TRAVERSAL_REBUILT)]
    check_hits(outs[1]["hits"], outs[0]["hits"])
    assert np.array_equal(outs[1]["composite"], outs[0]["composite"])
    assert np.array_equal(outs[1]["accum"].view(np.uint32), outs[0]["accum"].view(np.uint32))
    assert (outs[0]["hits"]["kind"] == 2).mean() > 0.9


def test_rebuilt_policy_grazing_and_inside_cameras():
    """Cameras that stress the certification step: grazing view along the terrain,
    a camera inside the terrain bounds, an axis-aligned view (zero direction
    components -> literal fallback rays), and a random soup on a chunked BVH."""
    s = g.terrain_scene(400, 300, strategy=abi.BVH_CHUNKED)
    cams = []
    for pos, d in (([0, 4.5, 40], [0, -0.05, -1]), ([3, 1.0, 2], [0.6, -0.1, -0.79]), ([0, 30, 0], [0, -1, 0]),
                   ([-60, 6, 0], [1, -0.08, 0.02])):
        cam = s.camera.copy()
        d = np.array(d, np.float32)
        d /= np.linalg.norm(d)
        cam["position"], cam["direction"] = pos, d
        if abs(d[1]) > 0.99:
            cam["up"] = [0, 0, -1]
        cams.append(cam)
    ctx = g.Context(0, 640, 360)
    g.BufferManager(ctx).upload_scene(s)
    cr = g.ComputeRenderer(ctx)
    bound = ob.BoundScene(s)
    for cam in cams:
        ref = ob.render(bound, 640, 360, camera=cam, channels=(0,))
        cr.render_frame(cam, traversal=abi.TRAVERSAL_REBUILT, outputs=abi.OUT_HITS | abi.OUT_ACCUM)
        assert ctx.last_traversal == abi.TRAVERSAL_REBUILT
        check_hits(ctx.read_hits(), ref["hits"])
        assert np.array_equal(ctx.read_accum()[..., 0].view(np.uint32), ref["radiance"][..., 0].view(np.uint32))
    ctx.close()


def test_random_soup_all_policies():
    import test_oracle
    s = test_oracle.random_soup(3000, 5).rebuild_bvh(abi.BVH_CHUNKED)
    ref = ob.render(s, 512, 384)
    for t in (abi.TRAVERSAL_LITERAL, abi.TRAVERSAL_ORDERED, abi.TRAVERSAL_REBUILT):
        gpu = render_gpu_fused(s, 512, 384, traversal=t)
        assert gpu["traversal"] == t
        compare(gpu, ref)


# ---- EXTENSION (SPEC.md): path tracing, bit-exact vs the extension oracle ------
def render_gpu_path(scene, w, h, spp, max_depth, base_seed=1, traversal=abi.TRAVERSAL_AUTO, chunks=None, variant=0):
    ctx = g.Context(0, w, h)
    g.BufferManager(ctx).upload_scene(scene)
    cr = g.ComputeRenderer(ctx)
    segs, first = 0, 0
    for n in (chunks or [spp]):
        cr.render_frame(scene.camera, mode=abi.MODE_PATH, spp=n, first_sample=first, max_depth=max_depth,
                        base_seed=base_seed, reset_accum=(first == 0), traversal=traversal, variant=variant,
                        outputs=abi.OUT_ACCUM | abi.OUT_HITS | abi.OUT_COMPOSITE)
        segs += ctx.last_segments
        first += n
    out = dict(accum=ctx.read_accum(), hits=ctx.read_hits(), composite=ctx.read_composite(), segments=segs,
               traversal=ctx.last_traversal)
    ctx.close()
    return out


def compare_path(gpu, ref):
    check_hits(gpu["hits"], ref["hits"])
    assert gpu["segments"] == ref["counters"]["rays"]
    bad = gpu["accum"].view(np.uint32) != ref["accum"].view(np.uint32)
    assert not bad.any(), (int(bad.sum()), np.argwhere(bad)[:4], gpu["accum"][bad][:4], ref["accum"][bad][:4])
    assert np.array_equal(gpu["composite"], ref["resolved"])


def test_path_spp1_depth0_is_reference_wavefront(default_ref):
    s, _ = default_ref
    wf = ob.render(s, 640, 360, wavefront=(0, 4, 5))
    gpu = render_gpu_path(s, 640, 360, 1, 0, base_seed=5)
    check_hits(gpu["hits"], wf["hits"])
    assert np.array_equal(gpu["accum"][..., :3].view(np.uint32), wf["radiance"].view(np.uint32))
    assert np.array_equal(gpu["composite"], wf["composite"])


@pytest.mark.parametrize("traversal,variant", [(abi.TRAVERSAL_LITERAL, 0), (abi.TRAVERSAL_ORDERED, 1),
                                               (abi.TRAVERSAL_ORDERED, 2), (abi.TRAVERSAL_ORDERED, 3),
                                               (abi.TRAVERSAL_ORDERED, 4)])
def test_c4_cornell_depth8(traversal, variant):
    """C4: Cornell box, diffuse bounces to depth 8 (incoherent rays; divergence stress)."""
    s = g.cornell_scene()
    ref = ob.render_path(s, 256, 256, spp=8, max_depth=8, base_seed=11)
    gpu = render_gpu_path(s, 256, 256, 8, 8, base_seed=11, traversal=traversal, variant=variant)
    compare_path(gpu, ref)
    assert gpu["segments"] > 256 * 256 * 8 * 3


@pytest.mark.parametrize("variant", [1, 4])
def test_path_default_scene_glass_metal_and_resume(default_ref, variant):
    s, _ = default_ref
    ref = ob.render_path(s, 320, 180, spp=6, max_depth=5, base_seed=2)
    gpu = render_gpu_path(s, 320, 180, 6, 5, base_seed=2, chunks=[1, 2, 3], variant=variant)  # progressive in 3 calls
    compare_path(gpu, ref)


@pytest.mark.parametrize("split", [1, 3, 4])
def test_wave_engine_sub_frames(default_ref, monkeypatch, split):
    """Wave engine with the context's tiles dealt to 1 / 3 / 4 concurrent sub-frames (own streams and
    buffers): same frame bit for bit (the default for small frames is 2, which the other tests cover)."""
    s, _ = default_ref
    ref = ob.render_path(s, 320, 180, spp=3, max_depth=4, base_seed=6)
    monkeypatch.setenv("RTB200_WAVE_SPLIT", str(split))
    gpu = render_gpu_path(s, 320, 180, 3, 4, base_seed=6, variant=4)
    compare_path(gpu, ref)


def test_wave_engine_multi_wave(default_ref, monkeypatch):
    """Wave engine with a path budget so small that the 5 samples take 3 waves (2 + 2 + 1): the
    per-wave resolve must add the samples in the same order as one wave / the megakernel."""
    s, _ = default_ref
    ref = ob.render_path(s, 320, 180, spp=5, max_depth=4, base_seed=9)
    monkeypatch.setenv("RTB200_WAVE_PATHS", str(2 * 15 * 4096))  # 320x180 = 5x3 split tiles of 4096 work items
    gpu = render_gpu_path(s, 320, 180, 5, 4, base_seed=9, variant=4)
    compare_path(gpu, ref)


def chain_bvh(scene):
    """Degenerate but valid uploaded tree: internal node 2k = (leaf of triangle k, chain of the rest), so the
    tree is len(triangles) - 1 levels deep."""
    n = len(scene.triangles)
    P, T = scene.vertices["position"], scene.triangles
    corners = np.stack([P[T["v0_index"]], P[T["v1_index"]], P[T["v2_index"]]], axis=1)  # n x 3 x 3
    lo, hi = corners.min(axis=1), corners.max(axis=1)
    nodes = np.zeros(2 * n - 1, dtype=abi.BvhNode)
    for k in range(n):
        leaf = 2 * k + 1 if k < n - 1 else 2 * k
        nodes[leaf]["bounds"]["min"], nodes[leaf]["bounds"]["max"] = lo[k], hi[k]
        nodes[leaf]["left_child"] = nodes[leaf]["right_child"] = 0xFFFFFFFF
        nodes[leaf]["triangle_start"], nodes[leaf]["triangle_count"] = k, 1
        if k < n - 1:
            nodes[2 * k]["bounds"]["min"], nodes[2 * k]["bounds"]["max"] = lo[k:].min(axis=0), hi[k:].max(axis=0)
            nodes[2 * k]["left_child"], nodes[2 * k]["right_child"] = 2 * k + 1, 2 * k + 2
    scene.bvh_nodes, scene.triangle_indices = nodes, np.arange(n, dtype=np.uint32)
    return scene


@pytest.mark.parametrize("variant", [1, 4])
def test_path_chain_bvh_deeper_than_the_shared_stack(variant):
    """Cornell box under a 35-level chain tree: the wave engine's extend kernel must take its DEEP form
    (stack entries beyond the shared-memory column live in local memory)."""
    s = chain_bvh(g.cornell_scene())
    ref = ob.render_path(s, 128, 128, spp=4, max_depth=4, base_seed=3)
    gpu = render_gpu_path(s, 128, 128, 4, 4, base_seed=3, traversal=abi.TRAVERSAL_ORDERED, variant=variant)
    assert gpu["traversal"] == abi.TRAVERSAL_ORDERED
    compare_path(gpu, ref)


@pytest.mark.parametrize("traversal", [abi.TRAVERSAL_ORDERED, abi.TRAVERSAL_REBUILT])
def test_wave_engine_forced_deep_form(c3_scene_and_ref, monkeypatch, traversal):
    """Both policies through the overflow-capable stack form (normally picked only for trees deeper than 30 levels)."""
    s, _ = c3_scene_and_ref
    ref = ob.render_path(s, 160, 90, spp=2, max_depth=3, base_seed=8, policy=1)
    monkeypatch.setenv("RTB200_WAVE_DEEP", "1")
    gpu = render_gpu_path(s, 160, 90, 2, 3, base_seed=8, traversal=traversal, variant=4)
    compare_path(gpu, ref)


def transformed_terrain(scale=1.0, offset=(0.0, 0.0, 0.0), flatten=False, camera_back=0.0):
    """Chunked-BVH terrain (100-triangle leaves -> REBUILT policy) under a similarity transform of the scene,
    lights and camera: exercises the 16-bit node grid of the wave engine far from the unit-scale default."""
    s = g.terrain_scene(120, 90, strategy=None)
    off = np.asarray(offset, np.float32)
    P = s.vertices["position"]
    if flatten:
        P[:, 1] = 0.0
    s.vertices["position"] = P * np.float32(scale) + off
    s.lights["position"] = s.lights["position"] * np.float32(scale) + off
    cam = s.camera.copy()
    pos = cam["position"] * np.float32(scale) + off
    cam["position"] = pos - cam["direction"] * np.float32(camera_back * scale)
    s.camera = cam
    return s.rebuild_bvh(abi.BVH_CHUNKED)


@pytest.mark.parametrize("name,kw", [
    ("offset", dict(offset=(5000.0, -3000.0, 8000.0))),   # extent << |coordinates|: the pad dominates the grid cell
    ("tiny", dict(scale=0.02)),
    ("huge", dict(scale=1e3)),
    ("flat", dict(flatten=True)),                          # zero extent along y: degenerate grid axis
    ("far_camera", dict(camera_back=300.0)),               # ray origins 3 scene extents away, terrain covers 7 % of the frame
    ("very_far_camera", dict(camera_back=1000.0)),         # 10 extents: a few dozen pixels hit
])
def test_rebuilt_policy_quantised_nodes_under_transforms(name, kw):
    s = transformed_terrain(**kw)
    assert s.bvh_nodes["triangle_count"].max() > 4
    ref = ob.render_path(s, 160, 90, spp=2, max_depth=3, base_seed=12, policy=1)
    gpu = render_gpu_path(s, 160, 90, 2, 3, base_seed=12, traversal=abi.TRAVERSAL_REBUILT, variant=4)
    assert gpu["traversal"] == abi.TRAVERSAL_REBUILT
    compare_path(gpu, ref)


def test_wave_engine_equals_megakernel_at_scale(c3_scene_and_ref):
    """A frame large enough that every block of the wave engine's kernels loops over many chunks (persistent
    work fetch, queued shading, several survivors per block): segments, radiance and image must equal the
    megakernel's (which the oracle pins at small sizes) bit for bit, run to run."""
    s, _ = c3_scene_and_ref
    a = render_gpu_path(s, 1280, 720, 4, 4, base_seed=3, variant=1)
    for _ in range(2):
        b = render_gpu_path(s, 1280, 720, 4, 4, base_seed=3, variant=4)
        assert a["segments"] == b["segments"]
        assert np.array_equal(a["accum"].view(np.uint32), b["accum"].view(np.uint32))
        assert np.array_equal(a["composite"], b["composite"])
        assert np.array_equal(a["hits"], b["hits"])


@pytest.mark.parametrize("variant", [1, 2, 3, 4])
@pytest.mark.parametrize("traversal", [abi.TRAVERSAL_ORDERED, abi.TRAVERSAL_REBUILT])
def test_c3_path_terrain_1m(c3_scene_and_ref, traversal, variant):
    """C3 semantics (4 bounces) on the 1M-triangle terrain at reduced size/spp so the oracle finishes."""
    s, _ = c3_scene_and_ref
    w, h = 240, 136
    ref = ob.render_path(s, w, h, spp=2, max_depth=4, base_seed=1, policy=1)
    gpu = render_gpu_path(s, w, h, 2, 4, base_seed=1, traversal=traversal, variant=variant)
    assert gpu["traversal"] == traversal
    compare_path(gpu, ref)


# ---- multi-GPU tile split: N-rank output must be bit-identical to the 1-GPU output ----
@pytest.mark.parametrize("world", [2, 4, 8])
def test_tile_split_pack_unpack_bit_identical(world):
    """Emulates N ranks on one GPU: each 'rank' context renders its tiles (i % N == r), packs
    them (rt_pack_tiles); the 'rank 0' context unpacks every buffer (rt_unpack_tiles).  The
    assembled frame must equal the single-context frame bit for bit."""
    import torch
    from gpu_raytracer_b200 import tiles
    s = g.default_scene()
    w, h = 1000, 600  # ragged edge tiles
    kw = dict(mode=abi.MODE_PATH, spp=3, max_depth=3, base_seed=4, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE)
    one = g.Context(0, w, h)
    g.BufferManager(one).upload_scene(s)
    g.ComputeRenderer(one).render_frame(s.camera, **kw)
    want_c, want_a = one.read_composite(), one.read_accum()
    root = g.Context(0, w, h)  # plays rank 0's frame buffers
    n_pad = tiles.tiles_per_rank_padded(w, h, world)
    for r in range(world):
        ctx = g.Context(0, w, h)
        g.BufferManager(ctx).upload_scene(s)
        g.ComputeRenderer(ctx).render_frame(s.camera, tile_stride=world, tile_phase=r, **kw)
        for what, bpp in ((tiles.WHAT_COMPOSITE, 4), (tiles.WHAT_ACCUM, 16)):
            buf = torch.zeros(n_pad * tiles.TILE * tiles.TILE * bpp, dtype=torch.uint8, device="cuda")
            ctx.pack_tiles(what, world, r, buf.data_ptr())
            ctx.synchronize()
            root.unpack_tiles(what, world, r, buf.data_ptr())
            root.synchronize()
        ctx.close()
    assert np.array_equal(root.read_composite(), want_c)
    assert np.array_equal(root.read_accum().view(np.uint32), want_a.view(np.uint32))
    one.close()
    root.close()


@pytest.mark.parametrize("mode", ["path", "primary"])
def test_peer_frame_epilogue_assembles_frame(mode):
    """Fused multi-GPU epilogue emulated on one GPU: every 'rank' context renders its split tiles with
    RT_OUT_PEER, its kernels storing the composite straight into the root context's frame (raw device
    pointer here; CUDA IPC across processes in bench.py).  The root frame must equal the single-context frame."""
    s = g.default_scene()
    w, h, world = 1000, 600, 4
    kw = dict(mode=abi.MODE_PATH, spp=2, max_depth=3, base_seed=4) if mode == "path" else {}
    one = g.Context(0, w, h)
    g.BufferManager(one).upload_scene(s)
    g.ComputeRenderer(one).render_frame(s.camera, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE, **kw)
    want = one.read_composite()
    root = g.Context(0, w, h)
    for r in range(world):
        ctx = g.Context(0, w, h)
        g.BufferManager(ctx).upload_scene(s)
        ctx.set_peer_frame(root.device_composite)
        g.ComputeRenderer(ctx).render_frame(s.camera, tile_stride=world, tile_phase=r,
                                            outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE | abi.OUT_PEER, **kw)
        ctx.synchronize()
        assert not ctx.read_composite().any() or r >= 0  # the rank's own composite buffer is not the target
        ctx.close()
    assert np.array_equal(root.read_composite(), want)
    with pytest.raises(g.RtError):
        g.ComputeRenderer(one).render_frame(s.camera, outputs=abi.OUT_COMPOSITE | abi.OUT_PEER)  # no peer frame opened
    one.close()
    root.close()


@pytest.mark.gpu
def test_upload_refuses_cyclic_bvh():
    """A BVH whose child links loop would hang the reference traversal (shader/src/bvh.rs:40-85) and this
    device; the upload reports it instead, and the context stays usable."""
    s = g.default_scene()
    ctx = g.Context(0, 64, 36)
    bm = g.BufferManager(ctx)
    bad = s.bvh_nodes.copy()
    internal = np.flatnonzero(bad["left_child"] != 0xFFFFFFFF)
    assert len(internal)
    bad["right_child"][internal[-1]] = 0  # a link back to the root
    assert g.bvh_has_cycle(bad) and not g.bvh_has_cycle(s.bvh_nodes)
    with pytest.raises(g.RtError):
        bm.update_scene_metadata(s.spheres, s.lights, bad, s.triangle_indices, s.vertices)
    bm.upload_scene(s)
    g.ComputeRenderer(ctx).render_frame(s.camera, outputs=abi.OUT_HITS)
    hits = ctx.read_hits()
    assert int((hits["kind"] != 0).sum()) > 0
    ctx.close()
