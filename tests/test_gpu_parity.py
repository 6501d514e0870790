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
    # run_compute -> present (main_fs, shader/src/lib.rs:368-391): the composite follows the dispatches
    assert np.array_equal(ctx.read_composite(), ref["composite"])
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


@pytest.mark.parametrize("variant", [1, 4])
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


def test_refused_upload_keeps_the_previous_scene():
    """A refused upload that is LARGER than the current capacity must not disturb the scene in place
    (validation runs before any buffer is touched)."""
    small = g.default_scene()
    ctx = g.Context(0, 160, 90)
    bm, cr = g.BufferManager(ctx), g.ComputeRenderer(ctx)
    bm.upload_scene(small)
    cr.render_frame(small.camera, outputs=abi.OUT_HITS | abi.OUT_ACCUM)
    want_h, want_a = ctx.read_hits(), ctx.read_accum()
    big = g.terrain_scene(60, 50)  # far beyond the default metadata capacity -> needs_resize
    bad = big.bvh_nodes.copy()
    internal = np.flatnonzero(bad["left_child"] != 0xFFFFFFFF)
    bad["right_child"][internal[-1]] = 0
    with pytest.raises(g.RtError):
        bm.update_scene_metadata(big.spheres, big.lights, bad, big.triangle_indices, big.vertices)
    for trav in (abi.TRAVERSAL_LITERAL, abi.TRAVERSAL_AUTO):
        cr.render_frame(small.camera, outputs=abi.OUT_HITS | abi.OUT_ACCUM, traversal=trav)
        assert np.array_equal(ctx.read_hits(), want_h)
        assert np.array_equal(ctx.read_accum().view(np.uint32), want_a.view(np.uint32))
    ctx.close()


# ---- reference paths no test reached before round 2 ----------------------------
def gpu_dispatch_rects(ctx, scene, offsets, rects, w, h, tpb, wavefront=None, channels=(0, 1, 2), pc_offsets=None):
    """rt_dispatch with caller-built push constants, one per rectangle and channel (the literal drop-in call)."""
    from gpu_raytracer_b200.host import push_constants_new
    cr = g.ComputeRenderer(ctx)
    for (x, y, rw, rh) in rects:
        for ch in channels:
            pc = push_constants_new([w, h], scene.camera, len(scene.triangles), len(scene.materials), [x, y], [rw, rh], [1, 1],
                                    tpb, offsets if pc_offsets is None else pc_offsets, ch, wavefront=wavefront)
            cr.dispatch(pc)


def compare_masked(ctx, ref, channels=(0, 1, 2)):
    m = ref["mask"]
    got = ctx.read_hits()
    for f in ("kind", "prim", "t_bits", "material_id"):
        assert np.array_equal(got[f][m], ref["hits"][f][m]), f
    acc = ctx.read_accum()
    for c in channels:
        assert np.array_equal(ctx.read_channel(c)[m], ref["channels"][c][m]), c
        assert np.array_equal(acc[..., c].view(np.uint32)[m], ref["radiance"][..., c].view(np.uint32)[m]), c


@pytest.mark.parametrize("tpb,pc_tpb", [(1000, 1000), (700, 700), (1000, 2000)])
def test_multi_buffer_triangle_fetch(tpb, pc_tpb):
    """shader/src/triangle_access.rs:18-61 at buffer indices 1 and 2, and past them: 2 520 triangles over bound
    buffers of 1000 (three buffers, all valid) or 700 (indices >= 2100 fall into buffer 3 -> invalid: holes in the
    terrain), and push constants whose divisor differs from the bound buffers' length (the shader trusts it).
    Every traversal policy must agree with the oracle on the same bound buffers."""
    s = g.terrain_scene(36, 35, strategy=abi.BVH_CHUNKED)
    assert len(s.triangles) == 2520
    w, h = 320, 180
    bound = ob.BoundScene(s, triangles_per_buffer=tpb)
    rects = [(0, 0, w, h)]
    ref = ob.render_rects(bound, w, h, rects, triangles_per_buffer=pc_tpb)
    kinds = ref["hits"]["kind"]
    assert (kinds == 2).mean() > (0.3 if tpb == pc_tpb else 0.02)  # (1000, 2000): only indices < 1000 resolve to their own triangle
    prim = ref["hits"]["prim"][kinds == 2]
    if tpb == pc_tpb == 1000:
        assert (prim >= 2000).any() and ((prim >= 1000) & (prim < 2000)).any() and (prim < 1000).any()
    if tpb == 700:
        assert not (prim >= 2100).any() and (prim >= 1400).any()  # buffer 3 is never bound
    ctx = g.Context(0, w, h)
    bm = g.BufferManager(ctx, triangles_per_buffer=tpb)
    assert bm.triangles_per_buffer == tpb
    off = bm.upload_scene(s)
    gpu_dispatch_rects(ctx, s, off, rects, w, h, pc_tpb)
    # a divisor that differs from the uploaded split cannot use the re-laid-out scene: literal policy
    assert ctx.last_traversal == (abi.TRAVERSAL_LITERAL if tpb != pc_tpb else abi.TRAVERSAL_REBUILT)
    compare_masked(ctx, ref)
    if tpb == pc_tpb:
        for trav in (abi.TRAVERSAL_LITERAL, abi.TRAVERSAL_ORDERED, abi.TRAVERSAL_REBUILT):
            g.ComputeRenderer(ctx).render_frame(s.camera, traversal=trav, outputs=abi.OUT_HITS | abi.OUT_ACCUM)
            assert ctx.last_traversal == trav
            check_hits(ctx.read_hits(), ref["hits"])
            assert np.array_equal(ctx.read_accum()[..., :3].view(np.uint32), ref["radiance"].view(np.uint32))
        # the progressive loop passes the context's own divisor (src/compute.rs:239)
        cr = g.ComputeRenderer(ctx)
        pending = cr.run_compute(s.camera, True)
        while pending > 0:
            pending = cr.run_compute(s.camera, False)
        check_hits(ctx.read_hits(), ref["hits"])
    with pytest.raises(g.RtError):
        g.BufferManager(ctx, triangles_per_buffer=64)  # fixed once triangles are uploaded (src/buffers.rs:73)
    ctx.close()


def test_missing_triangle_bindings_alias_buffer_zero():
    """src/renderer.rs:680-693: bindings 3 and 4 fall back to triangle buffer 0 when only one or two buffers exist,
    so a triangle_indices entry in their range tests the aliased triangle under the out-of-range index."""
    s = g.terrain_scene(30, 25, strategy=abi.BVH_CHUNKED)  # 1500 triangles -> two buffers of 1000
    n = len(s.triangles)
    idx = s.triangle_indices.copy()
    sel = np.flatnonzero(idx < 500)[::3]
    idx[sel] = idx[sel] + 2000  # buffer index 2 (not bound: aliases buffer 0), same local triangle
    s.triangle_indices = idx
    w, h = 256, 144
    bound = ob.BoundScene(s, triangles_per_buffer=1000)
    ref = ob.render_rects(bound, w, h, [(0, 0, w, h)])
    prim = ref["hits"]["prim"][ref["hits"]["kind"] == 2]
    assert (prim >= 2000).any() and n == 1500
    ctx = g.Context(0, w, h)
    g.BufferManager(ctx, triangles_per_buffer=1000).upload_scene(s)
    for trav in (abi.TRAVERSAL_LITERAL, abi.TRAVERSAL_ORDERED, abi.TRAVERSAL_REBUILT):
        g.ComputeRenderer(ctx).render_frame(s.camera, traversal=trav, outputs=abi.OUT_HITS | abi.OUT_ACCUM)
        check_hits(ctx.read_hits(), ref["hits"])
        assert np.array_equal(ctx.read_accum()[..., :3].view(np.uint32), ref["radiance"].view(np.uint32))
    ctx.close()


def test_brute_force_without_bvh():
    """shader/src/lib.rs:272-296: bvh_nodes_count == 0 -> every triangle in index order, seeded with the sphere t."""
    import test_oracle
    s = test_oracle.random_soup(300, 7)
    d = g.default_scene()
    s.spheres = d.spheres[:3]
    s.bvh_nodes = np.zeros(0, abi.BvhNode)
    s.triangle_indices = np.zeros(0, np.uint32)
    ref = ob.render(s, 160, 90)
    assert set(np.unique(ref["hits"]["kind"])) >= {1, 2}
    gpu = render_gpu_fused(s, 160, 90)
    assert gpu["traversal"] == abi.TRAVERSAL_LITERAL
    compare(gpu, ref)
    wf = ob.render_path(s, 96, 54, spp=2, max_depth=3, base_seed=4)
    compare_path(render_gpu_path(s, 96, 54, 2, 3, base_seed=4), wf)


def test_dispatch_with_caller_built_push_constants(default_ref):
    """The literal drop-in call (src/compute.rs:212-251): flags and offsets come from the caller's push constants."""
    from gpu_raytracer_b200.host import push_constants_new
    s, _ = default_ref
    w, h = 320, 180
    rects = [(0, 0, 128, 128), (128, 0, 128, 128), (256, 0, 64, 128), (0, 128, 128, 52), (128, 128, 128, 52), (256, 128, 64, 52)]
    ctx = g.Context(0, w, h)
    off = g.BufferManager(ctx).upload_scene(s)
    tpb = abi.MAX_TRIANGLES_PER_BUFFER
    # wavefront mode per dispatch, depth 0 of 4 (lib.rs:92-149)
    ref = ob.render_rects(s, w, h, rects, wavefront=(0, 4, 77))
    gpu_dispatch_rects(ctx, s, off, rects, w, h, tpb, wavefront=(0, 4, 77))
    compare_masked(ctx, ref)
    assert np.array_equal(ctx.read_composite(), ref["composite"])
    # current_bounce_depth > max_bounce_depth: the loop body never runs -> black (lib.rs:115-118)
    gpu_dispatch_rects(ctx, s, off, rects, w, h, tpb, wavefront=(5, 4, 77))
    dead = ob.render_rects(s, w, h, rects, wavefront=(5, 4, 77))
    assert not dead["radiance"].any()
    for c in range(3):
        assert np.array_equal(ctx.read_channel(c), dead["channels"][c])
    assert not ctx.read_accum()[..., :3].any()
    # a dispatch whose current depth is not 0 still traces the primary ray only (lib.rs:120-146)
    ref2 = ob.render_rects(s, w, h, rects, wavefront=(2, 4, 77))
    gpu_dispatch_rects(ctx, s, off, rects, w, h, tpb, wavefront=(2, 4, 77))
    compare_masked(ctx, ref2)
    # offsets in the push constants that differ from the uploaded ones: three spheres, no lights
    o2 = off.copy()
    o2["spheres_count"], o2["lights_count"] = 3, 0
    ref3 = ob.render_rects(s, w, h, rects, offsets=o2)
    gpu_dispatch_rects(ctx, s, off, rects, w, h, tpb, pc_offsets=o2)
    assert ctx.last_traversal == abi.TRAVERSAL_LITERAL
    compare_masked(ctx, ref3)
    assert (ref3["hits"]["prim"][ref3["hits"]["kind"] == 1] < 3).all()
    # channel > 2: `.expect("Invalid color channel")` (src/compute.rs:228-229) -> status, not a panic
    pc = push_constants_new([w, h], s.camera, len(s.triangles), len(s.materials), [0, 0], [128, 128], [1, 1], tpb, off, 3)
    with pytest.raises(g.RtError) as e:
        g.ComputeRenderer(ctx).dispatch(pc)
    assert e.value.code == -4
    ctx.close()


@pytest.mark.parametrize("traversal", [abi.TRAVERSAL_LITERAL, abi.TRAVERSAL_ORDERED, abi.TRAVERSAL_AUTO])
def test_uploaded_small_leaf_tree(traversal):
    """the host builder's small-leaf tree (<= 4 triangles per leaf, reference node format)."""
    s = g.terrain_scene(120, 90, strategy=abi.BVH_SMALL_LEAF)
    assert 1 < s.bvh_nodes["triangle_count"].max() <= 4
    ref = ob.render(s, 320, 180)
    gpu = render_gpu_fused(s, 320, 180, traversal=traversal)
    compare(gpu, ref)
    refp = ob.render_path(s, 160, 90, spp=2, max_depth=3, base_seed=5)
    compare_path(render_gpu_path(s, 160, 90, 2, 3, base_seed=5, traversal=traversal), refp)


def test_material_id_between_count_and_capacity_is_a_zero_material():
    """materials.len() in the shader is the bound buffer's length = capacity (64), zero-filled past the uploaded
    count (src/buffers.rs:100-106, 339-352; src/renderer.rs:698): id 10 of 4 shades as a zero material (black,
    but a hit), id 64 is magenta (lib.rs:307-309)."""
    s = g.default_scene()
    s.spheres["material_id"][0] = 10
    s.spheres["material_id"][1] = 64
    ref = ob.render(s, 320, 180)
    hit10 = (ref["hits"]["kind"] == 1) & (ref["hits"]["prim"] == 0)
    hit64 = (ref["hits"]["kind"] == 1) & (ref["hits"]["prim"] == 1)
    assert hit10.any() and not ref["radiance"][hit10].any()
    assert hit64.any() and (ref["radiance"][hit64] == np.array([1, 0, 1], np.float32)).all()
    compare(render_gpu_fused(s, 320, 180), ref)
    refp = ob.render_path(s, 160, 90, spp=2, max_depth=3, base_seed=5)
    compare_path(render_gpu_path(s, 160, 90, 2, 3, base_seed=5), refp)
    # growth: 70 materials -> capacity 140 (src/buffers.rs:343)
    s2 = g.default_scene()
    s2.materials = np.concatenate([s2.materials] * 18)[:70]
    s2.spheres["material_id"][0] = 139
    s2.spheres["material_id"][1] = 140
    ref2 = ob.render(s2, 160, 90)
    compare(render_gpu_fused(s2, 160, 90), ref2)


def test_rebuilt_policy_far_and_offset_cameras_sweep():
    """Adversarial sweep for the REBUILT policy's conservative boxes: cameras 1e1 ... 1e6 scene sizes away and a
    tiny scene far from the origin; hit records must equal the LITERAL policy's (rays beyond RT_REBUILT_FAR scene
    magnitudes take the literal policy inside the kernel)."""
    s = g.terrain_scene(120, 90, strategy=abi.BVH_CHUNKED)
    ctx = g.Context(0, 256, 144)
    g.BufferManager(ctx).upload_scene(s)
    cr = g.ComputeRenderer(ctx)
    ext = float(np.abs(s.vertices["position"]).max())
    for dist in (10.0, 15.0, 17.0, 100.0, 1e3, 1e4, 1e6):
        cam = s.camera.copy()
        cam["position"] = -cam["direction"] * np.float32(dist * ext)
        cam["fov"] = np.float32(max(2.0 / dist, 1e-3) * 40.0)
        out = []
        for trav in (abi.TRAVERSAL_LITERAL, abi.TRAVERSAL_REBUILT):
            cr.render_frame(cam, traversal=trav, outputs=abi.OUT_HITS | abi.OUT_ACCUM)
            assert ctx.last_traversal == trav
            out.append((ctx.read_hits(), ctx.read_accum()))
        check_hits(out[1][0], out[0][0])
        assert np.array_equal(out[1][1].view(np.uint32), out[0][1].view(np.uint32))
        cr.render_frame(cam, mode=abi.MODE_PATH, spp=1, max_depth=2, traversal=abi.TRAVERSAL_REBUILT, outputs=abi.OUT_HITS | abi.OUT_ACCUM)
        a = ctx.read_accum()
        cr.render_frame(cam, mode=abi.MODE_PATH, spp=1, max_depth=2, traversal=abi.TRAVERSAL_LITERAL, outputs=abi.OUT_HITS | abi.OUT_ACCUM)
        assert np.array_equal(a.view(np.uint32), ctx.read_accum().view(np.uint32)), dist
    ctx.close()
    # a scene with a non-finite coordinate keeps the uploaded-tree policies
    s2 = g.terrain_scene(20, 20, strategy=abi.BVH_CHUNKED)
    s2.vertices["position"][5, 1] = np.inf
    s2 = s2.rebuild_bvh(abi.BVH_CHUNKED)
    ctx = g.Context(0, 64, 36)
    g.BufferManager(ctx).upload_scene(s2)
    g.ComputeRenderer(ctx).render_frame(s2.camera, outputs=abi.OUT_HITS)
    assert ctx.last_traversal != abi.TRAVERSAL_REBUILT
    ctx.close()


# ---- BASELINE configs at their configured sizes ------------------------------------------------------------------
def spread_rows(h, n):
    return sorted(set(int(round((i + 0.5) * h / n)) for i in range(n)))


def test_c3_configured_size_rows_vs_literal_oracle():
    """C3 as configured (1920x1080, 16 spp, 4 bounces, 1 000 000 triangles): rows spread over the frame from the
    LITERAL extension oracle against the wave engine's frame, accumulation bits and resolved pixels."""
    s = g.terrain_scene(1000, 500)
    w, h = 1920, 1080
    gpu = render_gpu_path(s, w, h, 16, 4, base_seed=1)
    assert gpu["traversal"] == abi.TRAVERSAL_REBUILT
    bound = ob.BoundScene(s)
    for y in spread_rows(h, 3):
        ref = ob.render_path(bound, w, h, spp=16, max_depth=4, base_seed=1, rows=(y, y + 1))
        assert np.array_equal(gpu["accum"][y].view(np.uint32), ref["accum"][y].view(np.uint32)), y
        assert np.array_equal(gpu["composite"][y], ref["resolved"][y])
        for f in ("kind", "prim", "t_bits", "material_id"):
            assert np.array_equal(gpu["hits"][f][y], ref["hits"][f][y])


def test_c4_configured_size_rows_vs_oracle():
    """C4 as configured: Cornell box, 1024x1024, 64 spp, depth 8."""
    s = g.cornell_scene()
    w = h = 1024
    gpu = render_gpu_path(s, w, h, 64, 8, base_seed=1)
    bound = ob.BoundScene(s)
    for y in spread_rows(h, 4):
        ref = ob.render_path(bound, w, h, spp=64, max_depth=8, base_seed=1, rows=(y, y + 1))
        assert np.array_equal(gpu["accum"][y].view(np.uint32), ref["accum"][y].view(np.uint32)), y
        assert np.array_equal(gpu["composite"][y], ref["resolved"][y])


def test_c5_10m_triangles_4k():
    """C5: 10 000 000 triangles (two bound triangle buffers: indices >= 8 388 608 live in buffer 1) at 3840x2160.
    Whole frame: REBUILT (AUTO) against the GPU's LITERAL policy; 16 rows spread over the frame plus patches
    around pixels whose primary hit is a buffer-1 triangle against the literal oracle; then C5's path mode
    (4 bounces) on a row against the extension oracle."""
    s = g.terrain_scene(2500, 2000)
    assert len(s.triangles) == 10_000_000 and s.bvh_nodes["triangle_count"].max() == 1000
    w, h = 3840, 2160
    ctx = g.Context(0, w, h)
    g.BufferManager(ctx).upload_scene(s)
    cr = g.ComputeRenderer(ctx)
    cr.render_frame(s.camera, outputs=abi.OUT_HITS | abi.OUT_ACCUM | abi.OUT_CHANNELS | abi.OUT_COMPOSITE)
    assert ctx.last_traversal == abi.TRAVERSAL_REBUILT
    hits, acc = ctx.read_hits(), ctx.read_accum()
    chans = [ctx.read_channel(c) for c in range(3)]
    cr.render_frame(s.camera, traversal=abi.TRAVERSAL_LITERAL, outputs=abi.OUT_HITS | abi.OUT_ACCUM)
    check_hits(ctx.read_hits(), hits)
    assert np.array_equal(ctx.read_accum().view(np.uint32), acc.view(np.uint32))
    hi = (hits["kind"] == 2) & (hits["prim"] >= abi.MAX_TRIANGLES_PER_BUFFER)
    assert hi.any() and ((hits["kind"] == 2) & (hits["prim"] < abi.MAX_TRIANGLES_PER_BUFFER)).any()
    # the literal oracle costs ~10 s per 4K row and channel here (1000-triangle leaves, no pruning): 16 row segments
    # of 480 pixels spread over the frame in x and y plus patches around buffer-1 hits, red channel; two segments
    # with all three channel passes
    rects = [((i % 8) * 480, y, 480, 1) for i, y in enumerate(spread_rows(h, 16))]
    ys, xs = np.nonzero(hi)
    for k in range(0, len(ys), max(1, len(ys) // 8)):
        rects.append((max(0, min(w - 64, int(xs[k]) - 32)), int(ys[k]), 64, 1))
    bound = ob.BoundScene(s)
    ref = ob.render_rects(bound, w, h, rects, channels=(0,))
    m = ref["mask"]
    assert (ref["hits"]["prim"][m & (ref["hits"]["kind"] == 2)] >= abi.MAX_TRIANGLES_PER_BUFFER).any()
    for f in ("kind", "prim", "t_bits", "material_id"):
        assert np.array_equal(hits[f][m], ref["hits"][f][m]), f
    assert np.array_equal(chans[0][m], ref["channels"][0][m])
    assert np.array_equal(acc[..., 0].view(np.uint32)[m], ref["radiance"][..., 0].view(np.uint32)[m])
    ref3 = ob.render_rects(bound, w, h, [rects[5], rects[-1]])
    m = ref3["mask"]
    for c in range(3):
        assert np.array_equal(chans[c][m], ref3["channels"][c][m])
        assert np.array_equal(acc[..., c].view(np.uint32)[m], ref3["radiance"][..., c].view(np.uint32)[m])
    # C5 path mode: 4 bounces, 1 spp
    cr.render_frame(s.camera, mode=abi.MODE_PATH, spp=1, max_depth=4, base_seed=1, outputs=abi.OUT_HITS | abi.OUT_ACCUM | abi.OUT_COMPOSITE)
    assert ctx.last_traversal == abi.TRAVERSAL_REBUILT
    pacc, pcomp = ctx.read_accum(), ctx.read_composite()
    for y in spread_rows(h, 1):
        refp = ob.render_path(bound, w, h, spp=1, max_depth=4, base_seed=1, rows=(y, y + 1))
        assert np.array_equal(pacc[y].view(np.uint32), refp["accum"][y].view(np.uint32)), y
        assert np.array_equal(pcomp[y], refp["resolved"][y])
    ctx.close()


def test_wavefront_counters_per_bounce_depth(default_ref):
    """WavefrontCounters (shared/src/lib.rs:187-194) of a wave-engine frame against the oracle's per-depth counts."""
    s, _ = default_ref
    w, h = 320, 180
    ref = ob.render_path(s, w, h, spp=3, max_depth=9, base_seed=21)
    ctx = g.Context(0, w, h)
    g.BufferManager(ctx).upload_scene(s)
    g.ComputeRenderer(ctx).render_frame(s.camera, mode=abi.MODE_PATH, spp=3, max_depth=9, base_seed=21, variant=4)
    wc = ctx.last_wavefront_counters()
    assert list(wc["rays_per_bounce"]) == [int(x) for x in ref["depth_rays"]]
    assert wc["rays_per_bounce"][0] == w * h * 3 and wc["total_rays_generated"] == ref["counters"]["rays"] == ctx.last_segments
    assert wc["max_bounce_depth"] == 9 and wc["frame_seed"] == 21
    assert wc["active_bounce_depths"] == sum(1 << d for d in range(8) if ref["depth_rays"][d])
    g.ComputeRenderer(ctx).render_frame(s.camera, mode=abi.MODE_PATH, spp=1, max_depth=2, variant=1)  # megakernel
    with pytest.raises(g.RtError):
        ctx.last_wavefront_counters()
    ctx.close()


def test_explicit_tile_lists_assemble_the_frame():
    """Multi-GPU load balancing: contexts render explicit, unequal tile lists (rt_set_tile_list); rank 0 assembles
    them with rt_pack_tile_list / rt_unpack_tile_list and, separately, through the peer-store epilogue.  Both frames
    must equal the single-context frame bit for bit."""
    import torch
    from gpu_raytracer_b200 import tiles
    s = g.default_scene()
    w, h, world = 1000, 600, 3
    kw = dict(mode=abi.MODE_PATH, spp=2, max_depth=3, base_seed=4)
    one = g.Context(0, w, h)
    g.BufferManager(one).upload_scene(s)
    g.ComputeRenderer(one).render_frame(s.camera, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE, **kw)
    want_c, want_a = one.read_composite(), one.read_accum()
    lists = tiles.rebalance(tiles.initial_lists(w, h, world), [3.0, 1.0, 2.0])
    assert len(set(len(l) for l in lists)) > 1 and sorted(t for l in lists for t in l) == list(range(16 * 10))
    root, peer_root = g.Context(0, w, h), g.Context(0, w, h)
    for r in range(world):
        ctx = g.Context(0, w, h)
        g.BufferManager(ctx).upload_scene(s)
        ctx.set_tile_list(lists[r])
        g.ComputeRenderer(ctx).render_frame(s.camera, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE, **kw)
        for what, bpp in ((tiles.WHAT_COMPOSITE, 4), (tiles.WHAT_ACCUM, 16)):
            buf = torch.zeros(len(lists[r]) * tiles.TILE * tiles.TILE * bpp, dtype=torch.uint8, device="cuda")
            ctx.pack_tiles(what, 1, 0, buf.data_ptr())  # the context's own list
            ctx.synchronize()
            root.unpack_tile_list(what, lists[r], buf.data_ptr())
            root.synchronize()
        ctx.set_peer_frame(peer_root.device_composite)
        g.ComputeRenderer(ctx).render_frame(s.camera, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE | abi.OUT_PEER, **kw)
        ctx.synchronize()
        # primary frames follow the list too
        g.ComputeRenderer(ctx).render_frame(s.camera, outputs=abi.OUT_HITS)
        hits = ctx.read_hits()
        ctx.close()
    assert np.array_equal(root.read_composite(), want_c)
    assert np.array_equal(root.read_accum().view(np.uint32), want_a.view(np.uint32))
    assert np.array_equal(peer_root.read_composite(), want_c)
    with pytest.raises(g.RtError):
        one.set_tile_list([16 * 10])  # not a tile of this frame
    one.set_tile_list([3, 5])
    one.resize(320, 200)  # tile numbers belong to the old size: cleared
    g.ComputeRenderer(one).render_frame(s.camera, outputs=abi.OUT_HITS)
    assert (one.read_hits()["kind"] != 0).any()
    for c in (one, root, peer_root):
        c.close()


@pytest.mark.parametrize("scene_name", ["default", "terrain"])
def test_primary_frames_through_the_wave_traversal(monkeypatch, scene_name):
    """RTB200_PRIMARY_WAVE=1: legacy and wavefront-depth-0 frames through wv_extend<GEN> + wv_shade_primary must equal
    the fused kernel's (and the oracle's) bit for bit, ragged tiles and a tile split included."""
    s = g.default_scene() if scene_name == "default" else g.terrain_scene(120, 90, strategy=abi.BVH_CHUNKED)
    w, h = 1000, 600
    ref = ob.render(s, w, h)
    refw = ob.render(s, w, h, wavefront=(0, 4, 9))
    monkeypatch.setenv("RTB200_PRIMARY_WAVE", "1")
    compare(render_gpu_fused(s, w, h), ref)
    compare(render_gpu_fused(s, w, h, mode=abi.MODE_WAVEFRONT0), refw)
    ctx = g.Context(0, w, h)
    g.BufferManager(ctx).upload_scene(s)
    for ph in range(3):
        g.ComputeRenderer(ctx).render_frame(s.camera, tile_stride=3, tile_phase=ph, outputs=abi.OUT_HITS | abi.OUT_ACCUM | abi.OUT_COMPOSITE)
    check_hits(ctx.read_hits(), ref["hits"])
    assert np.array_equal(ctx.read_composite(), ref["composite"])
    ctx.close()


@pytest.mark.parametrize("literal_walk", [False, True])
def test_literal_fallback_rays_take_the_fallback_kernel(monkeypatch, literal_walk):
    """Rays the internal tree cannot serve (axis-aligned directions: zero components; a camera 40 scene sizes away)
    are traced by wv_fallback, one warp per ray (pruned ORDERED search, or the literal walk with
    RTB200_FALLBACK_LITERAL=1); whole frames of them must still equal the oracle."""
    if literal_walk:
        monkeypatch.setenv("RTB200_FALLBACK_LITERAL", "1")
    s = g.terrain_scene(60, 45, strategy=abi.BVH_CHUNKED)
    w, h = 97, 55  # odd: the centre column / row have cx == 0 / cy == 0 exactly
    cam = s.camera.copy()
    cam["position"], cam["direction"], cam["up"] = [0.3, 30.0, 0.2], [0, -1, 0], [0, 0, -1]  # centre row / column: d.x == 0 or d.z == 0
    far = s.camera.copy()
    far["position"] = -far["direction"] * np.float32(40.0 * float(np.abs(s.vertices["position"]).max()))
    far["fov"] = np.float32(2.5)
    for c in (cam, far):
        ref = ob.render_path(s, w, h, spp=2, max_depth=3, base_seed=6, camera=c)
        ctx = g.Context(0, w, h)
        g.BufferManager(ctx).upload_scene(s)
        cr = g.ComputeRenderer(ctx)
        ctx.count_next_launch()
        cr.render_frame(c, mode=abi.MODE_PATH, spp=2, max_depth=3, base_seed=6, traversal=abi.TRAVERSAL_REBUILT, variant=4,
                        outputs=abi.OUT_ACCUM | abi.OUT_HITS | abi.OUT_COMPOSITE)
        n_fb = ctx.read_counters()["literal_fallbacks"]
        assert n_fb > (w * h if c is far else 10)
        assert ctx.last_segments == ref["counters"]["rays"]
        assert np.array_equal(ctx.read_accum().view(np.uint32), ref["accum"].view(np.uint32))
        check_hits(ctx.read_hits(), ref["hits"])
        ctx.close()


def test_filtered_research_rays_equal_the_oracle():
    """Rays whose internal-tree winner fails its certification are searched again by wv_fallback with the filter, one
    warp per ray, the lanes sharing the node tests (wv_filtered_search_warp).  Scenes that produce such rays (a field
    of tori over a flat ground; a terrain seen from 10 / 15 scene sizes away) must still equal the oracle bit for bit,
    instrumented and plain launches alike, and the megakernel (literal walk for those rays)."""
    field = g.field_scene(n_tori=120, strategy=abi.BVH_CHUNKED)
    terrain = g.terrain_scene(120, 90, strategy=abi.BVH_CHUNKED)
    ext = float(np.abs(terrain.vertices["position"]).max())
    cases = [(field, field.camera, 384, 216, 8, 4)]
    for dist in (10.0, 15.0):
        cam = terrain.camera.copy()
        cam["position"] = -cam["direction"] * np.float32(dist * ext)
        cam["fov"] = np.float32(2.0 / dist * 40.0)
        cases.append((terrain, cam, 256, 144, 4, 3))
    total_fb = 0
    for s, cam, w, h, spp, depth in cases:
        ref = ob.render_path(s, w, h, spp=spp, max_depth=depth, base_seed=11, camera=cam)
        ctx = g.Context(0, w, h)
        g.BufferManager(ctx).upload_scene(s)
        cr = g.ComputeRenderer(ctx)
        kw = dict(mode=abi.MODE_PATH, spp=spp, max_depth=depth, base_seed=11, traversal=abi.TRAVERSAL_REBUILT,
                  outputs=abi.OUT_ACCUM | abi.OUT_HITS | abi.OUT_COMPOSITE)
        ctx.count_next_launch()
        cr.render_frame(cam, variant=4, **kw)
        n_fb = ctx.read_counters()["literal_fallbacks"]
        print("filtered re-search: %s %dx%d: %d of %d segments" % (s.name, w, h, n_fb, ctx.last_segments))
        total_fb += n_fb
        for variant in (None, 4, 1):  # the instrumented frame above; a plain wave-engine frame; the megakernel
            if variant is not None:
                cr.render_frame(cam, variant=variant, **kw)
            assert ctx.last_segments == ref["counters"]["rays"]
            assert np.array_equal(ctx.read_accum().view(np.uint32), ref["accum"].view(np.uint32)), (s.name, variant)
            check_hits(ctx.read_hits(), ref["hits"])
        ctx.close()
    assert total_fb > 0, "no certification failure in any case: the filtered search was not exercised"


def test_cuda_array_views_and_stream_ordered_copy():
    """Context.composite_cuda_array / accum_cuda_array: zero-copy views of rt_device_composite / rt_device_accum for
    CUDA-array consumers.  A copy enqueued on the context's stream right after a frame (no rt_read_* wait) must
    deliver that frame: the asynchronous read-back of bench.py's pipelined e2e mode."""
    import torch
    s = g.default_scene()
    w, h = 333, 222
    ctx = g.Context(0, w, h)
    g.BufferManager(ctx).upload_scene(s)
    comp = torch.as_tensor(ctx.composite_cuda_array(), device="cuda:0")
    acc = torch.as_tensor(ctx.accum_cuda_array(), device="cuda:0")
    assert comp.shape == (h, w, 4) and comp.dtype == torch.uint8 and comp.data_ptr() == ctx.device_composite
    assert acc.shape == (h, w, 4) and acc.dtype == torch.float32 and acc.data_ptr() == ctx.device_accum
    stream = torch.cuda.ExternalStream(ctx.stream, device=0)
    host_c = torch.empty((h, w, 4), dtype=torch.uint8).pin_memory()
    host_a = torch.empty((h, w, 4), dtype=torch.float32).pin_memory()
    for seed in (3, 4):
        g.ComputeRenderer(ctx).render_frame(s.camera, mode=abi.MODE_PATH, spp=4, max_depth=3, base_seed=seed,
                                            outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE)
        with torch.cuda.stream(stream):
            host_c.copy_(comp, non_blocking=True)
            host_a.copy_(acc, non_blocking=True)
            done = torch.cuda.Event()
            done.record(stream)
        done.synchronize()
        assert np.array_equal(host_c.numpy(), ctx.read_composite())
        assert np.array_equal(host_a.numpy().view(np.uint32), ctx.read_accum().view(np.uint32))
    # torch's pinned-memory allocator records an event on the streams a block was used on when it is freed: free the
    # host tensors (and the views) while the context's stream still exists
    del host_c, host_a, comp, acc, done, stream
    ctx.close()


def test_two_contexts_double_buffered_from_two_threads():
    """The e2e mode of bench.py: two contexts on one GPU, an uploader thread preparing the next scene (rt_update_* +
    rt_prepare_scene) in one context while the main thread renders in the other.  Frames must equal the serial ones."""
    from concurrent.futures import ThreadPoolExecutor
    w, h = 320, 180
    scenes = [g.default_scene(), g.terrain_scene(60, 45, strategy=abi.BVH_CHUNKED), g.cornell_scene(),
              g.terrain_scene(40, 30, strategy=abi.BVH_SMALL_LEAF), g.default_scene()]
    kw = dict(mode=abi.MODE_PATH, spp=2, max_depth=3, base_seed=5, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE)
    want = []
    one = g.Context(0, w, h)
    for s in scenes:
        bm = g.BufferManager(one)
        bm.mark_all_dirty()
        bm.upload_scene(s)
        g.ComputeRenderer(one).render_frame(s.camera, **kw)
        want.append((one.read_accum(), one.read_composite(), one.last_traversal))
    one.close()
    lanes = [g.Context(0, w, h), g.Context(0, w, h)]
    bms = [g.BufferManager(c) for c in lanes]

    def upload(k, s):
        bms[k].mark_all_dirty()
        bms[k].upload_scene(s, prepare=True)

    pool = ThreadPoolExecutor(max_workers=1)
    pending = pool.submit(upload, 0, scenes[0])
    for i, s in enumerate(scenes):
        pending.result()
        if i + 1 < len(scenes):
            pending = pool.submit(upload, (i + 1) % 2, scenes[i + 1])  # runs beside this frame, other context
        c = lanes[i % 2]
        g.ComputeRenderer(c).render_frame(s.camera, **kw)
        acc, comp = c.read_accum(), c.read_composite()
        assert c.last_traversal == want[i][2]
        assert np.array_equal(acc.view(np.uint32), want[i][0].view(np.uint32)), i
        assert np.array_equal(comp, want[i][1]), i
    pool.shutdown()
    for c in lanes:
        c.close()
