"""Thin numpy wrappers over the host half of librtb200.so (include/rtb200_host.h):
the `shared` crate's encoders, the BVH builders and the procedural scenes."""
import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import abi
from .native import load_host as load  # host-only library: no CUDA code is mapped by scene preparation


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f3(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.float32).reshape(3))


def f32_to_f16_bits(v):
    return int(load().rt_f32_to_f16(float(v)))


def f16_bits_to_f32(b):
    return float(load().rt_f16_to_f32(int(b)))


def camera_default():
    cam = np.zeros((), dtype=abi.Camera)
    load().rt_camera_default(_p(cam))
    return cam


def material_new(albedo, metallic, roughness, emission, ior, transmission):
    m = np.zeros((), dtype=abi.Material)
    a, e = _f3(albedo), _f3(emission)
    load().rt_material_new(_p(m), _p(a), metallic, roughness, _p(e), ior, transmission)
    return m


def material_diffuse(albedo):
    return material_new(albedo, 0.0, 1.0, [0, 0, 0], 1.5, 0.0)


def material_metallic(albedo, roughness):
    return material_new(albedo, 1.0, roughness, [0, 0, 0], 1.5, 0.0)


def material_glass(albedo, ior, transmission):
    return material_new(albedo, 0.0, 0.0, [0, 0, 0], ior, transmission)


def material_emissive(albedo, emission):
    return material_new(albedo, 0.0, 1.0, emission, 1.5, 0.0)


def light_directional(direction, color, intensity):
    l = np.zeros((), dtype=abi.Light)
    d, c = _f3(direction), _f3(color)
    load().rt_light_directional(_p(l), _p(d), _p(c), intensity)
    return l


def light_point(position, color, intensity, rng=float("inf")):
    l = np.zeros((), dtype=abi.Light)
    p, c = _f3(position), _f3(color)
    load().rt_light_point(_p(l), _p(p), _p(c), intensity, rng)
    return l


def light_spot(position, direction, color, intensity, rng, inner, outer):
    l = np.zeros((), dtype=abi.Light)
    p, d, c = _f3(position), _f3(direction), _f3(color)
    load().rt_light_spot(_p(l), _p(p), _p(d), _p(c), intensity, rng, inner, outer)
    return l


def push_constants_new(resolution, camera, triangle_count, material_count, tile_offset, tile_size, total_tiles,
                       triangles_per_buffer, offsets, color_channel, wavefront=None):
    """PushConstants::new (shared/src/lib.rs:1076-1102); wavefront=(cur, max, seed) selects new_wavefront."""
    pc = np.zeros((), dtype=abi.PushConstants)
    res = np.asarray(resolution, dtype=np.float32)
    to, ts, tt = (np.asarray(x, dtype=np.uint32) for x in (tile_offset, tile_size, total_tiles))
    cam = np.ascontiguousarray(camera)
    off = np.ascontiguousarray(offsets)
    if wavefront is None:
        load().rt_push_constants_new(_p(pc), _p(res), _p(cam), triangle_count, material_count, _p(to), _p(ts), _p(tt),
                                     triangles_per_buffer, _p(off), color_channel)
    else:
        cur, mx, seed = wavefront
        load().rt_push_constants_new_wavefront(_p(pc), _p(res), _p(cam), triangle_count, material_count, _p(to),
                                               _p(ts), _p(tt), triangles_per_buffer, _p(off), color_channel, cur, mx,
                                               seed)
    return pc


@dataclass
class Scene:
    """The arrays SceneState owns (src/scene.rs:6-17)."""
    camera: np.ndarray
    spheres: np.ndarray
    triangles: np.ndarray
    vertices: np.ndarray
    materials: np.ndarray
    lights: np.ndarray
    bvh_nodes: np.ndarray = None
    triangle_indices: np.ndarray = None
    textures: np.ndarray = field(default_factory=lambda: np.zeros(0, dtype=abi.TextureInfo))
    texture_data: np.ndarray = field(default_factory=lambda: np.zeros(0, dtype=np.uint8))
    name: str = ""

    def rebuild_bvh(self, strategy=abi.BVH_AUTO):
        """SceneState::rebuild_bvh (src/scene.rs:122-127)."""
        self.bvh_nodes, self.triangle_indices = build_bvh(self.triangles, self.vertices, strategy)
        return self


def build_bvh(triangles, vertices, strategy=abi.BVH_AUTO):
    """BvhBuilder::build (src/bvh.rs:104-122) -> (nodes, triangle_indices)."""
    lib = load()
    triangles = np.ascontiguousarray(triangles, dtype=abi.Triangle)
    vertices = np.ascontiguousarray(vertices, dtype=abi.Vertex)
    n = len(triangles)
    nodes = np.zeros(int(lib.rt_bvh_max_nodes(n)), dtype=abi.BvhNode)
    idx = np.zeros(max(n, 1), dtype=np.uint32)
    nn, ni = C.c_uint32(0), C.c_uint32(0)
    rc = lib.rt_bvh_build(_p(triangles), n, _p(vertices), len(vertices), strategy, _p(nodes), C.byref(nn),
                          _p(idx), C.byref(ni))
    if rc != 0:
        raise ValueError(f"rt_bvh_build failed ({rc})")
    return nodes[:nn.value].copy(), idx[:ni.value].copy()


def bvh_has_cycle(nodes):
    """True if the child links reachable from node 0 form a cycle (such an upload is refused: the reference
    traversal, shader/src/bvh.rs:40-85, would not terminate on it)."""
    nodes = np.ascontiguousarray(nodes, dtype=abi.BvhNode)
    return bool(load().rt_bvh_has_cycle(_p(nodes), len(nodes)))


def default_scene(strategy=abi.BVH_AUTO):
    """SceneBuilder::build_default_scene + SceneState::new (shared/src/lib.rs:1242-1286, src/scene.rs:20-40)."""
    s = np.zeros(6, dtype=abi.Sphere)
    t = np.zeros(2, dtype=abi.Triangle)
    v = np.zeros(6, dtype=abi.Vertex)
    m = np.zeros(4, dtype=abi.Material)
    l = np.zeros(1, dtype=abi.Light)
    load().rt_default_scene(_p(s), _p(t), _p(v), _p(m), _p(l))
    return Scene(camera_default(), s, t, v, m, l, name="default").rebuild_bvh(strategy)


def terrain_scene(nx, nz, seed=1, strategy=abi.BVH_AUTO):
    """Terrain(nx, nz, seed): 2*nx*nz triangles."""
    v = np.zeros((nx + 1) * (nz + 1), dtype=abi.Vertex)
    t = np.zeros(2 * nx * nz, dtype=abi.Triangle)
    m = np.zeros(3, dtype=abi.Material)
    l = np.zeros(2, dtype=abi.Light)
    cam = np.zeros((), dtype=abi.Camera)
    load().rt_scene_terrain(nx, nz, seed, _p(v), _p(t), _p(m), _p(l), _p(cam))
    sc = Scene(cam, np.zeros(0, dtype=abi.Sphere), t, v, m, l, name=f"terrain{nx}x{nz}")
    return sc.rebuild_bvh(strategy) if strategy is not None else sc


def field_scene(n_tori=1996, seg=25, ring=10, gnx=40, gnz=25, seed=1, strategy=abi.BVH_AUTO):
    """Field: tori at mixed scales scattered through a volume over a ground grid (not a height field).  The default
    is exactly 1 000 000 triangles."""
    v = np.zeros(n_tori * seg * ring + (gnx + 1) * (gnz + 1), dtype=abi.Vertex)
    t = np.zeros(2 * n_tori * seg * ring + 2 * gnx * gnz, dtype=abi.Triangle)
    m = np.zeros(3, dtype=abi.Material)
    l = np.zeros(2, dtype=abi.Light)
    cam = np.zeros((), dtype=abi.Camera)
    load().rt_scene_field(n_tori, seg, ring, gnx, gnz, seed, _p(v), _p(t), _p(m), _p(l), _p(cam))
    sc = Scene(cam, np.zeros(0, dtype=abi.Sphere), t, v, m, l, name=f"field{n_tori}")
    return sc.rebuild_bvh(strategy) if strategy is not None else sc


def cornell_scene(strategy=abi.BVH_AUTO):
    v = np.zeros(72, dtype=abi.Vertex)
    t = np.zeros(36, dtype=abi.Triangle)
    m = np.zeros(4, dtype=abi.Material)
    l = np.zeros(1, dtype=abi.Light)
    cam = np.zeros((), dtype=abi.Camera)
    nv = C.c_uint32(0)
    load().rt_scene_cornell(_p(v), C.byref(nv), _p(t), _p(m), _p(l), _p(cam))
    return Scene(cam, np.zeros(0, dtype=abi.Sphere), t, v[:nv.value].copy(), m, l, name="cornell").rebuild_bvh(strategy)


class GltfError(ValueError):
    """GltfError (src/gltf_loader.rs:14-20)."""


def _copy_array(ptr, n, dtype):
    if not n:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_uint8 * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=n).copy()


def load_gltf(source, scene_index=None, strategy=abi.BVH_AUTO, base_dir=None):
    """SceneState::load_from_gltf (src/scene.rs:43-69): GltfLoader::load_from_path / load_from_glb +
    extract_scene (src/gltf_loader.rs:53-125), first camera or Camera::new, then BvhBuilder::build.
    `source` is a path (.gltf / .glb) or bytes."""
    lib = load()
    h = C.c_void_p()
    err = C.create_string_buffer(512)
    si = -1 if scene_index is None else int(scene_index)
    if isinstance(source, (bytes, bytearray, memoryview)):
        data = np.frombuffer(bytes(source), dtype=np.uint8)
        rc = lib.rt_gltf_load_memory(_p(data), len(data), (base_dir or ".").encode(), si, C.byref(h), err, len(err))
    else:
        rc = lib.rt_gltf_load_file(str(source).encode(), si, C.byref(h), err, len(err))
    if rc != 0:
        raise GltfError(err.value.decode(errors="replace"))
    try:
        cnt = np.zeros(10, dtype=np.uint64)
        lib.rt_gltf_counts(h, _p(cnt))
        nt, nv, nm, ns, nl, nc, ntex, ntb, _nscenes, nwarn = (int(x) for x in cnt)
        cams = _copy_array(lib.rt_gltf_cameras(h), nc, abi.Camera)
        sc = Scene(cams[0].copy() if nc else camera_default(),
                   _copy_array(lib.rt_gltf_spheres(h), ns, abi.Sphere), _copy_array(lib.rt_gltf_triangles(h), nt, abi.Triangle),
                   _copy_array(lib.rt_gltf_vertices(h), nv, abi.Vertex), _copy_array(lib.rt_gltf_materials(h), nm, abi.Material),
                   _copy_array(lib.rt_gltf_lights(h), nl, abi.Light),
                   textures=_copy_array(lib.rt_gltf_textures(h), ntex, abi.TextureInfo),
                   texture_data=_copy_array(lib.rt_gltf_texture_data(h), ntb, np.uint8), name="gltf")
        sc.cameras = cams
        sc.warnings = [lib.rt_gltf_warning(h, i).decode() for i in range(nwarn)]
    finally:
        lib.rt_gltf_free(h)
    return sc.rebuild_bvh(strategy) if strategy is not None else sc


def write_image(path, rgba8):
    """Output stage: .ppm or .png from an (H, W, 4) uint8 image (composite / channel read-back)."""
    a = np.ascontiguousarray(rgba8, dtype=np.uint8)
    h, w = a.shape[:2]
    fn = load().rt_write_png_rgba8 if str(path).lower().endswith(".png") else load().rt_write_ppm_rgba8
    if fn(str(path).encode(), _p(a), w, h) != 0:
        raise OSError(f"cannot write {path}")


def write_pfm(path, accum):
    """rgb / sample count of the (H, W, 4) float32 accumulation buffer as a PFM file."""
    a = np.ascontiguousarray(accum, dtype=np.float32)
    h, w = a.shape[:2]
    if load().rt_write_pfm_accum(str(path).encode(), _p(a), w, h) != 0:
        raise OSError(f"cannot write {path}")


def pack_scene_metadata(scene):
    """The combined scene-metadata buffer and offsets, exactly as
    BufferManager::update_scene_metadata lays them out (src/buffers.rs:213-268).
    Host-side helper for callers that need the packed view (e.g. the test oracle)."""
    parts = [scene.spheres, scene.lights, scene.bvh_nodes, scene.triangle_indices, scene.vertices]
    blobs = [np.ascontiguousarray(p).view(np.uint8).reshape(-1) for p in parts]
    meta = np.concatenate(blobs).view(np.uint32).copy() if sum(b.size for b in blobs) else np.zeros(0, np.uint32)
    off = np.zeros((), dtype=abi.SceneMetadataOffsets)
    o = 0
    for name, p, b in zip(("spheres", "lights", "bvh_nodes", "triangle_indices", "vertices"), parts, blobs):
        off[name + "_offset"] = o // 4
        off[name + "_count"] = len(p)
        o += b.size
    return meta, off
