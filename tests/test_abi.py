"""ABI tests (CPU): struct layouts, library exports, and the reference's 10
`shared` unit tests restated (shared/src/lib.rs:1328-1455) against the C header
and host helpers."""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi, native, host

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_compiles_as_c_and_cpp_with_static_asserts():
    src = '#include "rt_shared.h"\n#include "rtb200.h"\n#include "rtb200_host.h"\nint main(void){return 0;}\n'
    for comp, ext in (("gcc", ".c"), ("g++", ".cpp")):
        with tempfile.TemporaryDirectory() as d:
            p = os.path.join(d, "t" + ext)
            open(p, "w").write(src)
            subprocess.check_call([comp, "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", p, "-o",
                                   os.path.join(d, "t.o")])


def test_numpy_dtypes_match_reference_sizes():
    # sizes from the Rust layouts
    for name, (dt, size) in abi.SIZES.items():
        assert dt.itemsize == size, name
    assert abi.Material.fields["metallic_roughness_f16"][1] == 12
    assert abi.Material.fields["emission"][1] == 16
    assert abi.Material.fields["ior_transmission_f16"][1] == 28
    assert abi.Material.fields["material_type"][1] == 84
    assert abi.Light.fields["direction"][1] == 32
    assert abi.BvhNode.fields["left_child"][1] == 32
    assert abi.WavefrontRay.fields["active"][1] == 72
    pc = abi.PushConstants.fields
    assert [pc[k][1] for k in ("camera", "triangle_count", "tile_offset", "tile_size_packed", "total_tiles",
                               "triangles_per_buffer", "metadata_offsets", "packed_flags", "frame_seed")] == \
        [8, 48, 56, 64, 68, 76, 80, 120, 124]


def test_library_loads_and_exports_every_declared_symbol():
    lib = native.load()  # no GPU needed to load
    for sym in native.API_SYMBOLS + native.HOST_SYMBOLS:
        assert hasattr(lib, sym), sym
    # every rt_* prototype in the headers is in the symbol lists
    import re
    declared = set()
    for hname in ("rtb200.h", "rtb200_host.h"):
        text = open(os.path.join(ROOT, "include", hname)).read()
        declared |= set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text))
    declared -= {"rt_pack_tile_size", "rt_pack_flags", "rt_tile_count", "rt_tiles_per_frame"}  # static inline
    assert declared == set(native.API_SYMBOLS + native.HOST_SYMBOLS), declared ^ set(native.API_SYMBOLS + native.HOST_SYMBOLS)


def test_no_gpu_means_loud_failure_not_fallback():
    import conftest
    if conftest.has_gpu():
        return
    try:
        g.Context(0)
        assert False, "Context() must fail without a CUDA device"
    except g.RtError as e:
        assert "no CPU fallback" in str(e)


def test_push_constants_round_trip():
    # shared/src/lib.rs:1434-1455 (test_push_constants_helpers) restated
    cam = host.camera_default()
    off = np.zeros((), abi.SceneMetadataOffsets)
    pc = host.push_constants_new([800, 600], cam, 10, 3, [128, 256], [128, 88], [7, 5], 8388608, off, 2)
    assert pc["tile_size_packed"] == (128 | (88 << 16))
    assert pc["packed_flags"] & 0xFF == 2 and (pc["packed_flags"] >> 8) & 0xFF == 0
    assert (pc["packed_flags"] >> 16) & 0xFF == 4 and (pc["packed_flags"] >> 24) == 0  # PushConstants::new :1088
    assert pc["frame_seed"] == 0 and tuple(pc["total_tiles"]) == (7, 5)
    pw = host.push_constants_new([800, 600], cam, 10, 3, [0, 0], [70000, 3], [1, 1], 5, off, 1, wavefront=(2, 7, 99))
    assert pw["tile_size_packed"] == (65535 | (3 << 16))  # pack_tile_size clamps to u16 (:1135-1139)
    assert pw["packed_flags"] == (1 | (2 << 8) | (7 << 16) | (1 << 24)) and pw["frame_seed"] == 99


def test_tile_helper():
    # TileHelper, shared/src/lib.rs:1187-1203, via the header's inline functions compiled here
    src = r'''
#include <stdio.h>
#include "rt_shared.h"
int main(void){ unsigned tx,ty; rt_tile_count(1280,720,128,&tx,&ty); printf("%u %u ",tx,ty);
 rt_tile_count(1920,1080,128,&tx,&ty); printf("%u %u ",tx,ty); rt_tile_count(3840,2160,128,&tx,&ty); printf("%u %u ",tx,ty);
 unsigned v[]={0,1,16,17,60,64,65,135,256,257,510,1024,1025,100000}; for(int i=0;i<14;i++) printf("%u ",rt_tiles_per_frame(v[i]));
 printf("%u %u", rt_pack_flags(2,3,4,1), rt_pack_tile_size(128,56)); return 0; }'''
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "t.c")
        open(p, "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), p, "-o", os.path.join(d, "t")])
        out = subprocess.check_output([os.path.join(d, "t")]).decode().split()
    assert out[:6] == ["10", "6", "15", "9", "30", "17"]
    # 0->max(0,1)=1, 1->1, 16->16, 17->2, 60->7, 64->8, 65->2, 135->4, 256->8, 257->4, 510->7, 1024->16, 1025->1
    assert out[6:20] == ["1", "1", "16", "2", "7", "8", "2", "4", "8", "4", "7", "16", "1", "1"]
    assert out[20:] == [str(2 | (3 << 8) | (4 << 16) | (1 << 24)), str(128 | (56 << 16))]


def _aabb(mn, mx):
    a = np.zeros((), abi.Aabb)
    a["min"], a["max"] = mn, mx
    return a


def test_aabb_helpers():
    # shared/src/lib.rs:1368-1407 restated: triangle AABB, union, center, surface area
    lib = native.load()
    v = np.zeros(3, abi.Vertex)
    v["position"] = [[0, 0, 0], [1, 2, -1], [-1, 0.5, 3]]
    t = np.zeros((), abi.Triangle)
    t["v0_index"], t["v1_index"], t["v2_index"] = 0, 1, 2
    out = np.zeros((), abi.Aabb)
    lib.rt_triangle_aabb(t.ctypes.data, v.ctypes.data, out.ctypes.data)
    assert list(out["min"]) == [-1, 0, -1] and list(out["max"]) == [1, 2, 3]
    a, b = _aabb([0, 0, 0], [1, 1, 1]), _aabb([-1, 0.5, 0.5], [0.5, 2, 0.75])
    u = np.zeros((), abi.Aabb)
    lib.rt_aabb_union(a.ctypes.data, b.ctypes.data, u.ctypes.data)
    assert list(u["min"]) == [-1, 0, 0] and list(u["max"]) == [1, 2, 1]
    c = np.zeros(3, np.float32)
    lib.rt_aabb_center(_aabb([0, 0, 0], [2, 4, 6]).ctypes.data, c.ctypes.data)
    assert list(c) == [1, 2, 3]
    assert lib.rt_aabb_surface_area(_aabb([0, 0, 0], [1, 2, 3]).ctypes.data) == 22.0
    e = np.zeros((), abi.Aabb)
    lib.rt_aabb_empty(e.ctypes.data)
    assert np.all(np.isposinf(e["min"])) and np.all(np.isneginf(e["max"]))
    lib.rt_aabb_union(e.ctypes.data, a.ctypes.data, u.ctypes.data)  # union with empty is identity
    assert list(u["min"]) == [0, 0, 0] and list(u["max"]) == [1, 1, 1]


def test_committed_ncu_summary_belongs_to_the_committed_sources():
    """The roofline's traffic / hit-rate figures in bench.py come from profiles/ncu_summary.json; every workload
    recorded there must have been captured from the native sources in this tree (bench.source_hash), so a change
    to csrc/ or include/ without a fresh capture fails here instead of shipping stale evidence."""
    import importlib.util
    import json
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_for_hash", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    with open(os.path.join(root, "profiles", "ncu_summary.json")) as f:
        summary = json.load(f)
    assert summary, "profiles/ncu_summary.json is empty"
    want = bench.source_hash()
    stale = {w: rec.get("src_hash") for w, rec in summary.items() if rec.get("src_hash") != want}
    assert not stale, "captured from another build than %s: %s" % (want, stale)
    assert "c3" in summary
