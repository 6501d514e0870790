"""CPU tests of the oracle (oracle/rt_oracle.cpp): analytic known-answer tests for
the shader arithmetic the reference leaves untested, the committed
golden fixtures, brute-force-vs-BVH agreement, and the equivalence of the kernel's
ORDERED traversal policy with the reference's literal policy."""
import math
import os
import sys

import numpy as np
import pytest

import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi, host
import oracle_binding as ob

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden  # noqa: E402

f32 = np.float32


def t_of(rec):
    return float(np.uint32(rec["t_bits"]).view(np.float32))


def test_center_ray_hits_sphere0_at_5_5():
    # default camera (0,0,5)->(0,0,-1); pixel (1,1) of a 3x3 image is the exact centre ray;
    # sphere 0 at (0,0,-1) r=0.5 => t = 5.5 exactly; normal (0,0,1)
    r = ob.render(g.default_scene(), 3, 3, threads=1)
    rec = r["hits"][1, 1]
    assert rec["kind"] == 1 and rec["prim"] == 0 and rec["material_id"] == 0 and t_of(rec) == 5.5
    # shading by hand (lighting.rs / material.rs), in f32 with the reference's op order
    p = np.array([0, 0, -0.5], f32)
    n = np.array([0, 0, 1], f32)
    albedo = np.array([0.8, 0.3, 0.3], f32)
    to_l = np.array([5, 7, 4], f32) - p
    d2 = f32(f32(to_l[0] * to_l[0]) + f32(to_l[1] * to_l[1])) + f32(to_l[2] * to_l[2])
    dist = np.sqrt(f32(d2))
    ldir = to_l * f32(f32(1.0) / dist)
    att = f32(np.float16(f32(1.0) / f32(f32(1.0) + f32(f32(dist * dist) * f32(0.01)))))
    ndl = f32(f32(n[0] * ldir[0]) + f32(n[1] * ldir[1])) + f32(n[2] * ldir[2])
    inten = f32(f32(max(ndl, f32(0)) * f32(1.0)) * att)
    want = f32(albedo * f32(0.1)) + f32(f32(f32(albedo / f32(math.pi)) * inten) * f32(1.0)) * f32(1.0)
    want = want * f32(1.0)  # light colour (1,1,1) and validity mask
    got = r["radiance"][1, 1]
    assert np.array_equal(got, want.astype(f32)), (got, want)
    assert list(r["composite"][1, 1]) == [int(np.rint(want[0] * 255)), int(np.rint(want[1] * 255)), int(np.rint(want[2] * 255)), 255]


def one_triangle_scene(v0, v1, v2, cam_pos, cam_dir, with_bvh=True):
    s = g.default_scene()
    s.spheres = np.zeros(0, abi.Sphere)
    s.vertices = np.zeros(3, abi.Vertex)
    s.vertices["position"] = [v0, v1, v2]
    s.triangles = np.zeros(1, abi.Triangle)
    s.triangles[0] = (0, 1, 2, 0)
    s.camera["position"], s.camera["direction"] = cam_pos, cam_dir
    s.rebuild_bvh()
    if not with_bvh:
        s.bvh_nodes = np.zeros(0, abi.BvhNode)
        s.triangle_indices = np.zeros(0, np.uint32)
    return s


@pytest.mark.parametrize("with_bvh", [True, False])
def test_axis_aligned_ray_vs_unit_triangle(with_bvh):
    # ray from (0.25,0.25,3) along -z hits the triangle (0,0,0),(1,0,0),(0,1,0) at t=3, two-sided
    s = one_triangle_scene([0, 0, 0], [1, 0, 0], [0, 1, 0], [0.25, 0.25, 3], [0, 0, -1], with_bvh)
    r = ob.render(s, 1, 1, threads=1)
    rec = r["hits"][0, 0]
    assert rec["kind"] == 2 and rec["prim"] == 0 and t_of(rec) == 3.0
    s2 = one_triangle_scene([0, 0, 0], [0, 1, 0], [1, 0, 0], [0.25, 0.25, 3], [0, 0, -1], with_bvh)  # flipped winding
    assert t_of(ob.render(s2, 1, 1, threads=1)["hits"][0, 0]) == 3.0
    # outside the triangle (u+v>1) and behind the origin -> miss with t = f32::MAX
    s3 = one_triangle_scene([0, 0, 0], [1, 0, 0], [0, 1, 0], [0.75, 0.75, 3], [0, 0, -1], with_bvh)
    rec = ob.render(s3, 1, 1, threads=1)["hits"][0, 0]
    assert rec["kind"] == 0 and rec["prim"] == 0xFFFFFFFF and rec["t_bits"] == 0x7F7FFFFF and rec["material_id"] == 0xFFFFFFFF
    s4 = one_triangle_scene([0, 0, 0], [1, 0, 0], [0, 1, 0], [0.25, 0.25, -3], [0, 0, -1], with_bvh)
    assert ob.render(s4, 1, 1, threads=1)["hits"][0, 0]["kind"] == 0


def test_parallel_ray_rejected_by_epsilon():
    # ray in the triangle's plane: |a| < 1e-5 -> miss (intersection.rs:109-111)
    s = one_triangle_scene([0, 0, 0], [1, 0, 0], [0, 1, 0], [-2, 0.25, 0], [1, 0, 0])
    assert ob.render(s, 1, 1, threads=1)["hits"][0, 0]["kind"] == 0


def test_miss_colour_legacy_vs_wavefront_sky():
    s = g.default_scene()
    s.camera["direction"] = [0, 0, 1]  # look away from everything
    leg = ob.render(s, 4, 4, threads=1)
    assert not leg["radiance"].any() and np.all(leg["composite"] == [0, 0, 0, 255])
    wf = ob.render(s, 4, 4, wavefront=(0, 4, 1), threads=1)
    assert np.array_equal(wf["radiance"][0, 0], np.array([0.1, 0.2, 0.3], f32))  # wavefront.rs:146-151
    assert list(wf["composite"][0, 0]) == [26, 51, 76, 255]
    # current depth above max depth: nothing runs (lib.rs:115-118) -> black
    dead = ob.render(s, 4, 4, wavefront=(5, 4, 1), threads=1)
    assert not dead["radiance"].any()


def test_invalid_material_is_magenta():
    s = g.default_scene()
    s.spheres["material_id"][0] = 99
    r = ob.render(s, 3, 3, threads=1)
    assert np.array_equal(r["radiance"][1, 1], np.array([1, 0, 1], f32))  # lib.rs:307-309


def test_glass_dispersion_term():
    # sphere 2 (glass, ior 1.5, T=0.9 in f16): out_c = L_c*(1-T) + ({.2,.2,.3}_c * ((ior+off_c-1)/(ior-1))) * T
    s = g.default_scene()
    s.camera["position"] = [1, 0, 5]
    r = ob.render(s, 3, 3, threads=1)
    rec = r["hits"][1, 1]
    assert rec["kind"] == 1 and rec["prim"] == 2 and rec["material_id"] == 2
    T = f32(np.float16(0.9))
    ior = f32(1.5)
    base = np.array([0.2, 0.2, 0.3], f32)
    offs = np.array([-0.018, 0.0, 0.035], f32)
    # lighting for this pixel from an opaque twin material
    s2 = g.default_scene()
    s2.camera["position"] = [1, 0, 5]
    s2.materials[2] = host.material_new([0.2, 0.3, 0.8], 0.0, 0.0, [0, 0, 0], 1.5, 0.0)
    L = ob.render(s2, 3, 3, threads=1)["radiance"][1, 1]
    for c in range(3):
        disp = f32(f32(f32(ior + offs[c]) - f32(1)) / f32(ior - f32(1)))
        want = f32(f32(L[c] * f32(f32(1) - T)) + f32(f32(base[c] * disp) * T))
        assert r["radiance"][1, 1, c] == want


def test_slab_zero_direction_and_origin_on_plane():
    # direction (0,0,-1): the x,y slabs use inv = +inf.  With the origin exactly on the
    # box face x=0, t1.x = 0*inf = NaN is dropped by f32::min/max, leaving tmin.x = tmax.x = +inf,
    # so `tmin_max <= tmax_min` fails: the reference MISSES this box (intersection.rs:151-164).
    s = one_triangle_scene([0, 0, 0], [1, 0, 0], [0, 1, 0], [0.0, 0.5, 3], [0, 0, -1])
    for policy in (0, 1):
        assert ob.render(s, 1, 1, threads=1, policy=policy)["hits"][0, 0]["kind"] == 0
    # without a BVH the same ray hits: u == 0 lies on the edge and `u < 0.0` does not reject it
    s = one_triangle_scene([0, 0, 0], [1, 0, 0], [0, 1, 0], [0.0, 0.5, 3], [0, 0, -1], with_bvh=False)
    rec = ob.render(s, 1, 1, threads=1)["hits"][0, 0]
    assert rec["kind"] == 2 and t_of(rec) == 3.0
    # strictly inside the slab the zero components give -inf/+inf and the box is hit
    s = one_triangle_scene([0, 0, 0], [1, 0, 0], [0, 1, 0], [0.001, 0.5, 3], [0, 0, -1])
    for policy in (0, 1):
        rec = ob.render(s, 1, 1, threads=1, policy=policy)["hits"][0, 0]
        assert rec["kind"] == 2 and t_of(rec) == 3.0


@pytest.mark.parametrize("name", sorted(make_golden.CASES))
def test_golden_fixtures(name):
    want = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    got = make_golden.run(name)
    for k in want.files:
        assert np.array_equal(got[k], want[k]), (name, k)


def random_soup(n, seed, with_spheres=True):
    rng = np.random.default_rng(seed)
    s = g.default_scene()
    c = rng.uniform(-3, 3, (n, 3)).astype(f32)
    c[:, 2] -= 4
    d = rng.uniform(-0.6, 0.6, (n, 3, 3)).astype(f32)
    s.vertices = np.zeros(3 * n, abi.Vertex)
    s.vertices["position"] = (c[:, None, :] + d).reshape(-1, 3)
    s.triangles = np.zeros(n, abi.Triangle)
    s.triangles["v0_index"], s.triangles["v1_index"], s.triangles["v2_index"] = np.arange(n) * 3, np.arange(n) * 3 + 1, np.arange(n) * 3 + 2
    s.triangles["material_id"] = rng.integers(0, 4, n)
    if not with_spheres:
        s.spheres = np.zeros(0, abi.Sphere)
    return s


@pytest.mark.parametrize("strategy", [abi.BVH_STANDARD, abi.BVH_SMALL_LEAF, abi.BVH_CHUNKED])
def test_bruteforce_vs_bvh_and_ordered_policy_agree(strategy):
    s = random_soup(700, 11).rebuild_bvh(strategy)
    w, h = 160, 120
    lit = ob.render(s, w, h)
    ordered = ob.render(s, w, h, policy=1)
    brute = random_soup(700, 11)
    brute.bvh_nodes, brute.triangle_indices = np.zeros(0, abi.BvhNode), np.zeros(0, np.uint32)
    bf = ob.render(brute, w, h)
    assert (lit["hits"]["kind"] == 2).mean() > 0.3
    # the t of the closest hit is topology independent; ids can differ from brute force only on exact t ties
    assert np.array_equal(lit["hits"]["t_bits"], bf["hits"]["t_bits"])
    assert np.array_equal(lit["hits"]["kind"], bf["hits"]["kind"])
    for f in ("kind", "prim", "t_bits", "material_id"):
        assert np.array_equal(lit["hits"][f], ordered["hits"][f]), f
    assert np.array_equal(lit["radiance"].view(np.uint32), ordered["radiance"].view(np.uint32))
    assert ordered["counters"]["tri_tests"] <= lit["counters"]["tri_tests"]


def test_equal_t_tie_goes_to_first_in_dfs_order():
    # two coincident triangles (same vertices, different material): both report the same t;
    # the reference keeps the first one visited (strict `<`), i.e. the smaller triangle_indices position
    for order in ((0, 1), (1, 0)):
        s = one_triangle_scene([0, 0, 0], [1, 0, 0], [0, 1, 0], [0.25, 0.25, 3], [0, 0, -1])
        s.triangles = np.zeros(2, abi.Triangle)
        s.triangles[0] = (0, 1, 2, 0)
        s.triangles[1] = (0, 1, 2, 1)
        s.rebuild_bvh()
        # force the leaf order
        leaves = np.nonzero(s.bvh_nodes["left_child"] == 0xFFFFFFFF)[0]
        assert len(leaves) == 2
        s.triangle_indices[:] = order
        for policy in (0, 1):
            rec = ob.render(s, 1, 1, threads=1, policy=policy)["hits"][0, 0]
            assert rec["prim"] == order[0] and rec["material_id"] == order[0], (order, policy)


def test_ordered_policy_equals_literal_on_terrains():
    for sc, (w, h) in ((g.terrain_scene(60, 40, strategy=abi.BVH_STANDARD), (320, 180)),
                       (g.terrain_scene(120, 90, strategy=abi.BVH_CHUNKED), (240, 136)),
                       (g.terrain_scene(90, 70, strategy=abi.BVH_SMALL_LEAF), (320, 180)), (g.cornell_scene(), (200, 200))):
        a, b = ob.render(sc, w, h, channels=(1,)), ob.render(sc, w, h, channels=(1,), policy=1)
        for f in ("kind", "prim", "t_bits", "material_id"):
            assert np.array_equal(a["hits"][f], b["hits"][f]), (sc.name, f)


# ---- EXTENSION (SPEC.md): path tracing on the CPU oracle -----------------------
def test_path_spp1_depth0_equals_wavefront_depth0():
    s = g.default_scene()
    wf = ob.render(s, 96, 54, wavefront=(0, 4, 9))
    pt = ob.render_path(s, 96, 54, spp=1, max_depth=0, base_seed=9)
    assert np.array_equal(pt["accum"][..., :3].view(np.uint32), wf["radiance"].view(np.uint32))
    assert np.all(pt["accum"][..., 3] == 1.0)
    assert np.array_equal(pt["resolved"], wf["composite"])
    for f in ("kind", "prim", "t_bits", "material_id"):
        assert np.array_equal(pt["hits"][f], wf["hits"][f])
    assert pt["counters"]["rays"] == 96 * 54


def test_path_progressive_accumulation_is_resumable_and_thread_independent():
    s = g.cornell_scene()
    full = ob.render_path(s, 48, 48, spp=6, max_depth=4, base_seed=3)
    part = ob.render_path(s, 48, 48, spp=2, max_depth=4, base_seed=3, threads=1)
    part = ob.render_path(s, 48, 48, spp=4, max_depth=4, base_seed=3, first_sample=2, accum=part["accum"], threads=3)
    assert np.array_equal(full["accum"].view(np.uint32), part["accum"].view(np.uint32))
    assert np.all(full["accum"][..., 3] == 6.0)
    assert full["counters"]["rays"] > 48 * 48 * 6 * 2  # bounces happen
    assert np.isfinite(full["accum"]).all() and full["accum"][..., :3].min() >= 0
    # literal and ordered traversal policies agree on incoherent bounce rays too
    o = ob.render_path(s, 48, 48, spp=6, max_depth=4, base_seed=3, policy=1)
    assert np.array_equal(full["accum"].view(np.uint32), o["accum"].view(np.uint32))


def test_path_depth_increases_radiance_and_segments_bounded():
    s = g.cornell_scene()
    d0 = ob.render_path(s, 32, 32, spp=4, max_depth=0, base_seed=1)
    d8 = ob.render_path(s, 32, 32, spp=4, max_depth=8, base_seed=1)
    assert d8["accum"][..., :3].sum() > d0["accum"][..., :3].sum()
    assert d0["counters"]["rays"] == 32 * 32 * 4
    assert 32 * 32 * 4 < d8["counters"]["rays"] <= 32 * 32 * 4 * 9


def test_material_id_below_bound_capacity_is_a_zero_material():
    # materials.len() is the bound buffer's length (capacity 64, zero-filled past the 4 uploaded materials;
    # src/buffers.rs:100-106, src/renderer.rs:698, lib.rs:307): id 10 shades as a zero material -> black hit
    s = g.default_scene()
    s.spheres["material_id"][0] = 10
    r = ob.render(s, 3, 3, threads=1)
    assert r["hits"][1, 1]["kind"] == 1 and r["hits"][1, 1]["material_id"] == 10
    assert not r["radiance"][1, 1].any()
    s.spheres["material_id"][0] = 64
    assert np.array_equal(ob.render(s, 3, 3, threads=1)["radiance"][1, 1], np.array([1, 0, 1], f32))
    assert ob.materials_capacity(64) == 64 and ob.materials_capacity(65) == 130


def test_bound_triangle_buffers_alias_and_tail():
    # one triangle, buffers of 4 entries: index 0 valid; index 1..3 = zero-filled tail (degenerate, no hit);
    # index 4 and 8 (buffers 1, 2: not created, bound to buffer 0 again, src/renderer.rs:680-693) alias triangle 0;
    # index 12 (buffer 3) is invalid
    for idx, want in ((0, True), (2, False), (4, True), (8, True), (12, False)):
        s = one_triangle_scene([0, 0, 0], [1, 0, 0], [0, 1, 0], [0.25, 0.25, 3], [0, 0, -1])
        s.triangle_indices = np.array([idx], np.uint32)
        b = ob.BoundScene(s, triangles_per_buffer=4)
        rec = ob.render(b, 1, 1, threads=1)["hits"][0, 0]
        assert (rec["kind"] == 2) == want, idx
        if want:
            assert rec["prim"] == idx and t_of(rec) == 3.0


def test_oracle_threads_share_short_bands():
    # 2 rows must still give every thread work (64-pixel spans): same pixels whatever the thread count
    s = g.default_scene()
    a = ob.render_path(s, 320, 180, spp=2, max_depth=3, rows=(90, 92), threads=1)
    b = ob.render_path(s, 320, 180, spp=2, max_depth=3, rows=(90, 92), threads=7)
    assert np.array_equal(a["accum"].view(np.uint32), b["accum"].view(np.uint32))
    assert a["counters"] == b["counters"]
