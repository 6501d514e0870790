"""The oracle (oracle/rt_oracle.cpp) against a second, independent numpy restatement of the reference shader
(tests/py_reference.py): hit records and fp32 radiance of whole small frames must agree bit for bit, for both
ray-generation modes, the BVH and the brute-force path, every light type, glass / metal / diffuse / invalid
materials, an axis-aligned camera (zero direction components: the NaN behaviour of the slab test) and triangle
indices that resolve through the bound-buffer rule."""
import numpy as np
import pytest

import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi, host
import oracle_binding as ob
import py_reference as pr
from test_oracle import random_soup


def agree(bound, w, h, camera=None, wavefront=None, tpb=None):
    kw = {} if tpb is None else dict(triangles_per_buffer=tpb)
    if tpb is None:
        got = ob.render(bound, w, h, camera=camera, wavefront=None if wavefront is None else (wavefront[0], wavefront[1], 7))
    else:
        got = ob.render_rects(bound, w, h, [(0, 0, w, h)], camera=camera, **kw)
    want = pr.render(bound, w, h, camera=camera, wavefront=wavefront, tpb=tpb)
    hits = got["hits"]
    if wavefront is not None and wavefront[0] > wavefront[1]:  # nothing is traced: no hit records, a black frame
        assert not got["radiance"].any() and not want["radiance"].any() and not hits["kind"].any()
        return want
    assert np.array_equal(hits["kind"], want["kind"])
    assert np.array_equal(hits["prim"], want["prim"])
    hit = want["kind"] != 0
    assert np.array_equal(hits["t_bits"][hit], want["t"].view(np.uint32)[hit])
    assert np.array_equal(hits["material_id"][hit], want["material_id"][hit])
    assert np.array_equal(got["radiance"].view(np.uint32), want["radiance"].view(np.uint32))
    return want


def test_default_scene_both_modes():
    b = ob.BoundScene(g.default_scene())
    r = agree(b, 160, 90)
    assert (r["kind"] == 1).any() and (r["kind"] == 2).any() and (r["kind"] == 0).any()
    agree(b, 160, 90, wavefront=(0, 4))
    agree(b, 33, 17, wavefront=(5, 4))  # current bounce beyond the maximum: no iteration runs, black


def test_terrain_chunked_and_small_leaf_trees():
    for strat in (abi.BVH_CHUNKED, abi.BVH_SMALL_LEAF):
        b = ob.BoundScene(g.terrain_scene(24, 18, strategy=strat))
        r = agree(b, 96, 54)
        assert (r["kind"] == 2).sum() > 500
        agree(b, 96, 54, wavefront=(0, 4))


def test_soup_with_spheres_every_light_type_and_material_class():
    s = random_soup(150, 3)
    s.lights = np.concatenate([np.atleast_1d(s.lights),
                               np.atleast_1d(host.light_directional([0.3, -1.0, -0.2], [1.0, 0.9, 0.8], 0.7)),
                               np.atleast_1d(host.light_spot([1.0, 4.0, 2.0], [-0.2, -1.0, -0.5], [0.6, 0.8, 1.0], 2.5, 50.0, 0.3, 0.6)),
                               np.atleast_1d(host.light_directional([0.0, 0.0, 0.0], [1.0, 1.0, 1.0], 1.0))])  # zero direction: NaN scrubbed by max
    assert sorted(int(t) for t in s.lights["light_type"]) == [0, 0, 1, 2]
    s.triangles["material_id"][::7] = 9      # inside the bound capacity: a zero material
    s.triangles["material_id"][::11] = 200   # beyond it: magenta
    b = ob.BoundScene(s.rebuild_bvh(abi.BVH_STANDARD))
    r = agree(b, 120, 80)
    assert (r["kind"] == 1).any() and (r["kind"] == 2).any()
    agree(b, 120, 80, wavefront=(0, 4))


def test_brute_force_without_bvh():
    s = random_soup(60, 5)
    s.bvh_nodes = np.zeros(0, abi.BvhNode)
    s.triangle_indices = np.zeros(0, np.uint32)
    agree(ob.BoundScene(s), 80, 60)


def test_axis_aligned_camera_zero_direction_components():
    s = g.terrain_scene(20, 16, strategy=abi.BVH_CHUNKED)
    cam = s.camera.copy()
    cam["position"], cam["direction"], cam["up"] = [0.3, 30.0, 0.2], [0, -1, 0], [0, 0, -1]
    r = agree(ob.BoundScene(s), 41, 23, camera=cam)  # odd sizes: the centre column / row have cx == 0 / cy == 0 exactly
    assert (r["kind"] == 2).any()


def test_bound_buffer_rule_in_both_restatements():
    s = random_soup(90, 8, with_spheres=False).rebuild_bvh(abi.BVH_STANDARD)
    b = ob.BoundScene(s, triangles_per_buffer=40)   # three bound buffers of 40: indices 80-89 live in buffer 2
    agree(b, 64, 48, tpb=40)
    agree(b, 64, 48, tpb=25)                        # push constants disagree with the bound length: buffer > 2 is invalid


# ---- EXTENSION semantics: the oracle's path tracer against a restatement written from the SPEC.md text ----------------
@pytest.mark.parametrize("case", ["default", "cornell", "terrain"])
def test_path_tracer_follows_the_spec_text(case):
    """orc_render_path (the checker of every RT_MODE_PATH result on the GPU) against SPEC.md steps 1-4 restated in
    numpy: accumulated fp32 radiance of every pixel and the segment count must agree bit for bit, through jittered
    samples, Russian roulette (depth >= 3), the rejection-sampled bounce direction, glass, metal and diffuse continuations."""
    s = {"default": g.default_scene, "cornell": g.cornell_scene,
         "terrain": lambda: g.terrain_scene(16, 12, strategy=abi.BVH_CHUNKED)}[case]()
    w, h, spp, depth = (48, 27, 3, 5) if case != "cornell" else (32, 32, 2, 6)
    b = ob.BoundScene(s)
    got = ob.render_path(b, w, h, spp=spp, max_depth=depth, base_seed=21)
    want = pr.render_path(b, w, h, spp, depth, 21)
    assert got["counters"]["rays"] == want["segments"]
    assert np.array_equal(got["accum"].view(np.uint32), want["accum"].view(np.uint32))


def test_random_scenes_and_cameras_sweep():
    """Seeded sweep: random triangle soups (with and without spheres) under the three host builders, cameras anywhere
    around them (normalised and un-normalised forward vectors, an axis-aligned one, fov 20-100 degrees), ragged frame
    sizes, both ray-generation modes, and a path-traced frame now and then."""
    rng = np.random.default_rng(123)
    for it in range(12):
        s = random_soup(int(rng.integers(5, 200)), int(rng.integers(1, 10 ** 6)), with_spheres=bool(rng.integers(0, 2)))
        s = s.rebuild_bvh([abi.BVH_STANDARD, abi.BVH_SMALL_LEAF, abi.BVH_CHUNKED][it % 3])
        cam = s.camera.copy()
        cam["position"] = rng.uniform(-6, 6, 3).astype(np.float32)
        fwd = np.array([0, 0, -4], np.float32) + rng.uniform(-1, 1, 3).astype(np.float32) - cam["position"]
        if it % 5 == 0:
            fwd = np.array([0, 0, -1], np.float32)
        cam["direction"] = fwd / np.linalg.norm(fwd) if it % 2 else fwd  # the shader uses `direction` as given
        cam["fov"] = np.float32(rng.uniform(20, 100))
        b = ob.BoundScene(s)
        w, h = int(rng.integers(8, 70)), int(rng.integers(8, 50))
        agree(b, w, h, camera=cam)
        agree(b, w, h, camera=cam, wavefront=(0, 4))
        if it % 4 == 0:
            got = ob.render_path(b, w, h, spp=2, max_depth=4, base_seed=it, camera=cam)
            want = pr.render_path(b, w, h, 2, 4, it, camera=cam)
            assert got["counters"]["rays"] == want["segments"]
            assert np.array_equal(got["accum"].view(np.uint32), want["accum"].view(np.uint32))


@pytest.mark.parametrize("name", ["default_160x90", "default_wavefront_96x54", "terrain_40x30_chunkedbvh_128x72",
                                  "terrain_24x20_standardbvh_130x70", "cornell_96x96"])
def test_committed_golden_vectors_are_reproduced_by_the_second_restatement(name):
    """tests/golden/*.npz were written by the oracle (make_golden.py); the independent numpy restatement must produce
    the same committed hit records and radiance bits, so the fixtures do not rest on one transcription alone."""
    import os
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.join(here, "golden"))
    import make_golden
    mk, w, h, wf = make_golden.CASES[name]
    want = np.load(os.path.join(here, "golden", name + ".npz"))
    got = pr.render(ob.BoundScene(mk()), w, h, wavefront=None if wf is None else (wf[0], wf[1]))
    hits = want["hits"]
    assert np.array_equal(hits["kind"], got["kind"]) and np.array_equal(hits["prim"], got["prim"])
    hit = got["kind"] != 0
    assert np.array_equal(hits["t_bits"][hit], got["t"].view(np.uint32)[hit])
    assert np.array_equal(hits["material_id"][hit], got["material_id"][hit])
    assert np.array_equal(want["radiance_bits"], got["radiance"].view(np.uint32))
