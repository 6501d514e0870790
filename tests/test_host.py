"""Host-side encoders and BVH builders (CPU): f16 packing, constructors, default
scene, and the reference's BVH unit tests restated (src/bvh.rs:383-523)."""
import numpy as np

import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi, host, native
import oracle_binding as ob


def test_f16_conversions_match_ieee_rne_and_oracle():
    lib = native.load()
    # every binary16 value -> f32 exactly (numpy float16 is IEEE binary16)
    bits = np.arange(65536, dtype=np.uint16)
    want = bits.view(np.float16).astype(np.float32)
    got = np.array([lib.rt_f16_to_f32(int(b)) for b in bits], np.float32)
    orc = np.array([ob.lib().orc_f16_to_f32(int(b)) for b in bits], np.float32)
    nan = np.isnan(want)
    assert np.array_equal(got.view(np.uint32)[~nan], want.view(np.uint32)[~nan])
    assert np.array_equal(orc.view(np.uint32)[~nan], want.view(np.uint32)[~nan])
    assert np.isnan(got[nan]).all() and np.isnan(orc[nan]).all()
    # f32 -> f16 RNE: random bit patterns + the rounding boundaries
    rng = np.random.default_rng(7)
    xs = rng.integers(0, 2 ** 32, 40000, dtype=np.uint64).astype(np.uint32).view(np.float32)
    edge = np.array([0.0, -0.0, 1.0, 65504.0, 65519.99, 65520.0, 65536.0, 1e-8, 2.0 ** -25, np.nextafter(np.float32(2.0 ** -25), np.float32(1)),
                     2.0 ** -24, 2.0 ** -14, 0.1, 1.5, 0.9, 0.5, 1.0 / 3, np.inf, -np.inf, 6.1e-5, 6.0e-8, 1.0009765625, 1.00048828125,
                     1.00146484375], np.float32)
    xs = np.concatenate([xs, edge, rng.uniform(-2, 2, 20000).astype(np.float32), rng.uniform(0, 1e-4, 5000).astype(np.float32)])
    with np.errstate(over="ignore"):
        want16 = xs.astype(np.float16).view(np.uint16)
    for x, w in zip(xs, want16):
        a, b = lib.rt_f32_to_f16(float(x)), ob.lib().orc_f32_to_f16(float(x))
        if np.isnan(x):
            assert (a & 0x7C00) == 0x7C00 and (a & 0x3FF) and (b & 0x7C00) == 0x7C00 and (b & 0x3FF)
        else:
            assert a == w and b == w, (x, hex(a), hex(b), hex(w))


def test_material_constructors():
    m = host.material_glass([0.2, 0.3, 0.8], 1.5, 0.9)  # shared/src/lib.rs:326-328
    assert m["metallic_roughness_f16"] == 0
    assert m["ior_transmission_f16"] == (0x3E00 | (int(np.float16(0.9).view(np.uint16)) << 16))
    assert m["specular_factor"] == 1.0 and np.isposinf(m["attenuation_distance"]) and m["glossiness_factor"] == 1.0
    assert list(m["texture_indices"]) == [0xFFFFFFFF] * 8 and list(m["diffuse_factor"]) == list(m["albedo"])
    d = host.material_diffuse([0.8, 0.3, 0.3])  # metallic 0, roughness 1, ior 1.5, transmission 0
    assert d["metallic_roughness_f16"] == (0x3C00 << 16) and d["ior_transmission_f16"] == 0x3E00
    mt = host.material_metallic([0.8, 0.8, 0.2], 0.1)
    assert mt["metallic_roughness_f16"] == (0x3C00 | (int(np.float16(0.1).view(np.uint16)) << 16))
    assert mt["glossiness_factor"] == np.float32(1.0) - np.float32(0.1)


def test_light_constructors():
    p = host.light_point([5, 7, 4], [1, 1, 1], 1.0)  # shared/src/lib.rs:524-549
    assert p["light_type"] == 1 and list(p["direction"]) == [0, 0, 0] and p["range_packed"] == 0x7C00
    assert p["cone_angles_packed"] == 0
    d = host.light_directional([0, -1, 0], [1, 1, 1], 2.0)
    assert d["light_type"] == 0 and list(d["position"]) == [0, 0, 0] and d["range_packed"] == 0x7C00
    s = host.light_spot([1, 2, 3], [0, -1, 0], [1, 0, 0], 3.0, 10.0, 0.5, 1.0)
    assert s["light_type"] == 2 and s["range_packed"] == int(np.float16(10).view(np.uint16))
    assert s["cone_angles_packed"] == (0x3800 | (0x3C00 << 16))


def test_default_scene_matches_reference_constants():
    s = g.default_scene()  # shared/src/lib.rs:1242-1286
    assert len(s.spheres) == 6 and len(s.triangles) == 2 and len(s.vertices) == 6 and len(s.materials) == 4
    assert list(s.spheres["material_id"]) == [0, 1, 2, 2, 1, 3] and np.all(s.spheres["radius"] == 0.5)
    assert list(s.spheres["center"][5]) == [-1, 2, -5]
    assert [tuple(t) for t in s.triangles.tolist()] == [(0, 1, 2, 0), (3, 4, 5, 1)]
    assert list(s.vertices["position"][3]) == [1.5, 0.5, -3.0]
    assert list(s.lights["position"][0]) == [5, 7, 4] and s.lights["light_type"][0] == 1
    cam = s.camera
    assert list(cam["position"]) == [0, 0, 5] and list(cam["direction"]) == [0, 0, -1] and cam["fov"] == 45.0
    assert list(s.materials["emission"][3]) == [0.5, 0.5, 1.0]


def _tri_scene(n, rng):
    v = np.zeros(3 * n, abi.Vertex)
    v["position"] = rng.uniform(-5, 5, (3 * n, 3)).astype(np.float32)
    t = np.zeros(n, abi.Triangle)
    t["v0_index"], t["v1_index"], t["v2_index"] = np.arange(n) * 3, np.arange(n) * 3 + 1, np.arange(n) * 3 + 2
    return t, v


def test_bvh_empty_single_multi():
    # src/bvh.rs:425-434 empty -> one empty leaf
    nodes, idx = g.build_bvh(np.zeros(0, abi.Triangle), np.zeros(0, abi.Vertex))
    assert len(nodes) == 1 and len(idx) == 0 and nodes[0]["left_child"] == 0xFFFFFFFF and nodes[0]["triangle_count"] == 0
    assert np.all(np.isposinf(nodes[0]["bounds"]["min"])) and np.all(np.isneginf(nodes[0]["bounds"]["max"]))
    rng = np.random.default_rng(1)
    # :437-452 single triangle -> one leaf
    t, v = _tri_scene(1, rng)
    nodes, idx = g.build_bvh(t, v)
    assert len(nodes) == 1 and list(idx) == [0] and nodes[0]["triangle_count"] == 1 and nodes[0]["right_child"] == 0xFFFFFFFF
    # :455-483 three triangles -> indices are a permutation of {0,1,2}
    t, v = _tri_scene(3, rng)
    nodes, idx = g.build_bvh(t, v)
    assert sorted(idx) == [0, 1, 2] and len(nodes) == 5
    # :486-508 root bounds contain the scene
    pts = v["position"]
    assert np.all(nodes[0]["bounds"]["min"] <= pts.min(0)) and np.all(nodes[0]["bounds"]["max"] >= pts.max(0))


def check_tree(nodes, idx, tris, verts, max_leaf):
    """Structural invariants of the reference format: pre-order not required, but
    every triangle exactly once, leaf boxes tight, parents contain children."""
    seen = np.zeros(len(tris), int)
    stack, depth_max, prev_end = [(0, 0)], 0, 0
    P = verts["position"]
    while stack:
        n, d = stack.pop()
        depth_max = max(depth_max, d)
        nd = nodes[n]
        if nd["left_child"] == 0xFFFFFFFF:
            s, c = int(nd["triangle_start"]), int(nd["triangle_count"])
            assert 1 <= c <= max_leaf and s == prev_end  # DFS leaf order == index order
            prev_end = s + c
            ids = idx[s:s + c]
            seen[ids] += 1
            pts = np.concatenate([P[tris["v0_index"][ids]], P[tris["v1_index"][ids]], P[tris["v2_index"][ids]]])
            assert np.array_equal(nd["bounds"]["min"], pts.min(0)) and np.array_equal(nd["bounds"]["max"], pts.max(0))
        else:
            for ch in (nd["right_child"], nd["left_child"]):
                if ch != 0xFFFFFFFF:
                    assert np.all(nodes[ch]["bounds"]["min"] >= nd["bounds"]["min"])
                    assert np.all(nodes[ch]["bounds"]["max"] <= nd["bounds"]["max"])
                    stack.append((int(ch), d + 1))
    assert np.all(seen == 1)
    return depth_max


def test_bvh_standard_and_small_leaf_invariants():
    rng = np.random.default_rng(3)
    t, v = _tri_scene(2000, rng)
    nodes, idx = g.build_bvh(t, v, abi.BVH_STANDARD)
    assert len(nodes) == 2 * 2000 - 1  # one triangle per leaf (src/bvh.rs:366-369)
    assert check_tree(nodes, idx, t, v, 1) < 60
    # pre-order: left child is the next node (convert_node_recursive, src/bvh.rs:298-375)
    internal = nodes["left_child"] != 0xFFFFFFFF
    assert np.all(nodes["left_child"][internal] == np.nonzero(internal)[0] + 1)
    nodes, idx = g.build_bvh(t, v, abi.BVH_SMALL_LEAF)
    assert check_tree(nodes, idx, t, v, 4) < 60
    # degenerate input: all triangles identical -> median fallback must terminate
    t2 = np.repeat(t[:1], 300)
    nodes, idx = g.build_bvh(t2, v, abi.BVH_STANDARD)
    assert len(nodes) == 599 and check_tree(nodes, idx, t2, v, 1) < 60


def chunked_reference_restatement(tris, verts):
    """Independent pure-Python restatement of build_chunked +
    build_simple_top_level_bvh (src/bvh.rs:154-247) to pin the C++ builder."""
    n = len(tris)
    per_leaf = max(n // 10000, 32)
    P = verts["position"]

    def leaf_box(a, b):
        ids = np.arange(a, b)
        pts = np.concatenate([P[tris["v0_index"][ids]], P[tris["v1_index"][ids]], P[tris["v2_index"][ids]]])
        return pts.min(0), pts.max(0)
    level, indices = [], []
    for base in range(0, n, per_leaf):
        cnt = min(per_leaf, n - base)
        mn, mx = leaf_box(base, base + cnt)
        level.append(dict(mn=mn, mx=mx, l=0xFFFFFFFF, r=0xFFFFFFFF, s=len(indices), c=cnt))
        indices += list(range(base, base + cnt))
    if len(level) <= 1:
        return level, indices
    out = []
    while len(level) > 1:
        nxt = []
        for i in range(0, len(level), 2):
            if i + 1 < len(level):
                a, b = level[i], level[i + 1]
                li = len(out)
                out += [a, b]
                nxt.append(dict(mn=np.minimum(a["mn"], b["mn"]), mx=np.maximum(a["mx"], b["mx"]), l=li, r=li + 1, s=0, c=0))
            else:
                a = level[i]
                li = len(out)
                out.append(a)
                nxt.append(dict(mn=a["mn"], mx=a["mx"], l=li, r=0xFFFFFFFF, s=0, c=0))
        level = nxt
    out.append(level[0])
    out.reverse()
    off = len(out) - 1
    res = []
    for nd in out:
        nd = dict(nd)
        if not (nd["l"] == 0xFFFFFFFF and nd["r"] == 0xFFFFFFFF):
            if nd["l"] != 0xFFFFFFFF:
                nd["l"] = off - nd["l"]
            if nd["r"] != 0xFFFFFFFF:
                nd["r"] = off - nd["r"]
        res.append(nd)
    return res, indices


def test_bvh_chunked_matches_python_restatement():
    for nx, nz in ((40, 33), (64, 32), (37, 5)):
        sc = g.terrain_scene(nx, nz, strategy=None)
        nodes, idx = g.build_bvh(sc.triangles, sc.vertices, abi.BVH_CHUNKED)
        want, widx = chunked_reference_restatement(sc.triangles, sc.vertices)
        assert list(idx) == widx and len(nodes) == len(want)
        for nd, w in zip(nodes, want):
            assert np.array_equal(nd["bounds"]["min"], w["mn"]) and np.array_equal(nd["bounds"]["max"], w["mx"])
            assert (nd["left_child"], nd["right_child"], nd["triangle_start"], nd["triangle_count"]) == (w["l"], w["r"], w["s"], w["c"])


def test_bvh_auto_threshold_and_headline_counts():
    sc = g.terrain_scene(250, 200)  # 100 000 triangles -> standard (src/bvh.rs:117)
    assert len(sc.triangles) == 100000 and len(sc.bvh_nodes) == 199999
    sc = g.terrain_scene(251, 200)  # 100 400 -> chunked, leaf size max(n/10000, 32) = 32
    assert sc.bvh_nodes["triangle_count"].max() == 32
    sc = g.terrain_scene(1000, 500)
    assert len(sc.bvh_nodes) == 20005 and sc.bvh_nodes["triangle_count"].max() == 100
    assert (sc.bvh_nodes["left_child"] == 0xFFFFFFFF).sum() == 10000


def test_pack_scene_metadata_offsets():
    s = g.default_scene()
    meta, off = g.pack_scene_metadata(s)  # src/buffers.rs:257-268
    assert off["spheres_offset"] == 0 and off["lights_offset"] == 6 * 5 and off["bvh_nodes_offset"] == 30 + 13
    assert off["triangle_indices_offset"] == 43 + 3 * 12 and off["vertices_offset"] == 79 + 2
    assert meta.size == 81 + 18 and off["vertices_count"] == 6
    assert meta[30 + 3] == 1 and meta[79] == s.triangle_indices[0]


def test_bvh_cycle_detection():
    """rt_bvh_has_cycle: the upload refuses trees on which shader/src/bvh.rs:40-85 would never terminate."""
    NONE = 0xFFFFFFFF
    for strat in (abi.BVH_STANDARD, abi.BVH_CHUNKED, abi.BVH_SMALL_LEAF):
        sc = g.terrain_scene(40, 33, strategy=strat)
        assert not g.bvh_has_cycle(sc.bvh_nodes)
    assert not g.bvh_has_cycle(g.default_scene().bvh_nodes)
    assert not g.bvh_has_cycle(np.zeros(0, abi.BvhNode))

    def node(l, r):
        n = np.zeros(1, abi.BvhNode)
        n["left_child"], n["right_child"] = l, r
        return n
    leaf = node(NONE, NONE)
    # self loop, two-node loop, loop through the right child only
    assert g.bvh_has_cycle(np.concatenate([node(0, NONE)]))
    assert g.bvh_has_cycle(np.concatenate([node(1, 2), node(0, NONE), leaf]))
    assert g.bvh_has_cycle(np.concatenate([node(1, 2), leaf, node(1, 0)]))
    # a DAG (node 2 reachable twice) and links that the traversal ignores (out of range, leaf's right child) are fine
    assert not g.bvh_has_cycle(np.concatenate([node(1, 2), node(2, NONE), leaf]))
    assert not g.bvh_has_cycle(np.concatenate([node(1, 7), node(NONE, 0)]))
    # an unreachable cycle does not matter
    assert not g.bvh_has_cycle(np.concatenate([node(1, NONE), leaf, node(3, NONE), node(2, NONE)]))
    # a 100 000-node chain must not overflow any stack
    n = 100000
    chain = np.zeros(n, abi.BvhNode)
    chain["left_child"] = np.arange(1, n + 1); chain["right_child"] = NONE
    chain["left_child"][-1] = NONE
    assert not g.bvh_has_cycle(chain)
    chain["left_child"][-1] = 5
    assert g.bvh_has_cycle(chain)


def test_bench_reference_arm_contract_on_cpu():
    """`bench.py --impl reference` (the driver's reference arm) needs no GPU: it times the oracle port on the host
    threads and prints ONE JSON line with the arm's keys; under torchrun only rank 0 prints.  It must not map
    librtb200.so (scene generation comes from librtb200_host.so)."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    child = ("import runpy, sys\n"
             "sys.argv = ['bench.py', '--impl', 'reference', '--workload', 'c4', '--steps', '1', '--warmup', '0']\n"
             "code = 0\n"
             "try:\n"
             "    runpy.run_path('bench.py', run_name='__main__')\n"
             "except SystemExit as e:\n"
             "    code = e.code or 0\n"
             "maps = open('/proc/self/maps').read()\n"
             "sys.stderr.write('MAPS_OK\\n' if 'librtb200.so' not in maps and 'librtb200_host.so' in maps else 'MAPS_BAD\\n')\n"
             "sys.exit(code)\n")
    out = subprocess.run([sys.executable, "-c", child], capture_output=True, text=True, env=env, cwd=root, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "MAPS_OK" in out.stderr, "the reference arm mapped the CUDA library (or not the host-only one)"
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "Mrays/s" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]
    # rank 1 of a 2-rank launch exits 0 without work or output
    env2 = dict(env, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1", MASTER_ADDR="127.0.0.1", MASTER_PORT="29599")
    out2 = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "2", "--workload", "c4",
                           "--steps", "1", "--warmup", "0"], capture_output=True, text=True, env=env2, cwd=root, timeout=600)
    assert out2.returncode == 0 and not [l for l in out2.stdout.splitlines() if l.startswith("{")]


def test_bench_product_arm_fails_loudly_without_a_gpu():
    """The b200 arm of bench.py has no CPU fallback: without a CUDA device it exits non-zero and prints no metric line."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--workload", "c4", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, env=env, cwd=root, timeout=600)
    assert out.returncode != 0
    assert not [l for l in out.stdout.splitlines() if l.startswith("{")]
