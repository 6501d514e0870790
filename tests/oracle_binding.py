"""TEST INFRASTRUCTURE: ctypes binding of oracle/librt_oracle.so (the CPU
restatement of the reference shader).  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs import this."""
import ctypes as C
import os

import numpy as np

from gpu_raytracer_b200 import abi
from gpu_raytracer_b200.host import pack_scene_metadata, push_constants_new

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = None


class OrcBindings(C.Structure):
    _fields_ = [("scene_metadata", C.c_void_p), ("scene_metadata_len", C.c_uint64),
                ("triangles", C.c_void_p * 3), ("triangles_len", C.c_uint64 * 3),
                ("materials", C.c_void_p), ("materials_len", C.c_uint64)]


class OrcCounters(C.Structure):
    _fields_ = [("aabb_tests", C.c_uint64), ("tri_tests", C.c_uint64), ("sphere_tests", C.c_uint64),
                ("rays", C.c_uint64)]


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(ROOT, "oracle", "librt_oracle.so")
        _lib = C.CDLL(path)
        _lib.orc_dispatch.restype = C.c_int
        _lib.orc_dispatch.argtypes = [C.POINTER(OrcBindings), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                      C.POINTER(OrcCounters)]
        _lib.orc_dispatch_policy.restype = C.c_int
        _lib.orc_dispatch_policy.argtypes = _lib.orc_dispatch.argtypes + [C.c_int]
        _lib.orc_render_path.restype = C.c_int
        _lib.orc_render_path.argtypes = [C.POINTER(OrcBindings), C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                         C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int,
                                         C.POINTER(OrcCounters), C.c_int]
        _lib.orc_last_depth_rays.restype = None
        _lib.orc_last_depth_rays.argtypes = [C.c_void_p]
        _lib.orc_resolve.restype = None
        _lib.orc_resolve.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        _lib.orc_composite.restype = None
        _lib.orc_composite.argtypes = [C.c_void_p] * 4 + [C.c_uint64]
        _lib.orc_hardware_threads.restype = C.c_int
        _lib.orc_f32_to_f16.restype = C.c_uint16
        _lib.orc_f32_to_f16.argtypes = [C.c_float]
        _lib.orc_f16_to_f32.restype = C.c_float
        _lib.orc_f16_to_f32.argtypes = [C.c_uint16]
    return _lib


def hardware_threads():
    return int(lib().orc_hardware_threads())


def materials_capacity(n, capacity=64):
    """BufferManager::update_materials growth rule (src/buffers.rs:339-352) for a fresh manager."""
    return capacity if n <= capacity else max(n * 2, 64)


class BoundScene:
    """The reference's bound buffers for one scene (bindings 1-5), as a fresh BufferManager + bind group
    would hold them:
      * triangles: max(1, ceil(n / tpb)) buffers of exactly `tpb` zero-initialised entries each
        (src/buffers.rs:93-99, 274-325; wgpu zero-fills new buffers); binding k = buffer k if it exists,
        else buffer 0 again (src/renderer.rs:680-693 `.unwrap_or(&buffers.triangle_buffers[0])`);
      * materials: one buffer of `materials_capacity` entries (64, doubled on growth), zero-filled past
        the uploaded count — `materials.len()` in the shader is that bound length (shader/src/lib.rs:307)."""

    def __init__(self, scene, triangles_per_buffer=abi.MAX_TRIANGLES_PER_BUFFER):
        self.scene = scene
        self.meta, self.offsets = pack_scene_metadata(scene)
        self.tpb = triangles_per_buffer
        tris = np.ascontiguousarray(scene.triangles, abi.Triangle)
        n_buf = max(1, (len(tris) + self.tpb - 1) // self.tpb)
        self.tri_buffers = []
        for i in range(n_buf):
            buf = np.zeros(self.tpb, abi.Triangle)  # calloc: untouched pages stay virtual
            part = tris[i * self.tpb:(i + 1) * self.tpb]
            buf[:len(part)] = part
            self.tri_buffers.append(buf)
        mats = np.ascontiguousarray(scene.materials, abi.Material)
        self.materials = np.zeros(materials_capacity(len(mats)), abi.Material)
        self.materials[:len(mats)] = mats
        b = OrcBindings()
        b.scene_metadata = self.meta.ctypes.data
        b.scene_metadata_len = self.meta.size
        for i in range(3):
            buf = self.tri_buffers[i] if i < n_buf else self.tri_buffers[0]
            b.triangles[i] = buf.ctypes.data
            b.triangles_len[i] = len(buf)
        b.materials = self.materials.ctypes.data
        b.materials_len = len(self.materials)
        self.bindings = b


def dispatch(bound, pc, image, radiance=None, hits=None, threads=None, counters=None, policy=0):
    """One reference dispatch (src/compute.rs:245-250) on the CPU."""
    threads = threads or hardware_threads()
    pc = np.ascontiguousarray(pc, abi.PushConstants)
    rc = lib().orc_dispatch_policy(C.byref(bound.bindings), pc.ctypes.data,
                                   image.ctypes.data if image is not None else None,
                                   radiance.ctypes.data if radiance is not None else None,
                                   hits.ctypes.data if hits is not None else None, threads,
                                   C.byref(counters) if counters is not None else None, policy)
    assert rc == 0
    return rc


def tile_rects(width, height, tile=abi.TILE_SIZE):
    """calculate_tile_dimensions over all tiles (src/compute.rs:194-209)."""
    tx, ty = (width + tile - 1) // tile, (height + tile - 1) // tile
    for i in range(tx * ty):
        ox, oy = (i % tx) * tile, (i // tx) * tile
        yield ox, oy, min(tile, width - ox), min(tile, height - oy), (tx, ty)


def render(scene, width, height, wavefront=None, threads=None, camera=None, channels=(0, 1, 2), rows=None, policy=0):
    """The reference's whole progressive render on the CPU: every tile x 3 colour
    channels (src/compute.rs:157-191) then the main_fs composite.  `rows=(y0,y1)`
    restricts to the tiles intersecting those rows (bounded cpu_baseline samples).
    Returns dict(channels[3] HxWx4 u8, composite, radiance HxWx3 f32, hits, counters)."""
    bound = scene if isinstance(scene, BoundScene) else BoundScene(scene)
    sc = bound.scene
    cam = sc.camera if camera is None else camera
    chans = [np.zeros((height, width, 4), np.uint8) for _ in range(3)]
    radiance = np.zeros((height, width, 3), np.float32)
    hits = np.zeros((height, width), abi.HitRecord)
    counters = OrcCounters()
    for ox, oy, tw, th, tt in tile_rects(width, height):
        if rows is not None and (oy + th <= rows[0] or oy >= rows[1]):
            continue
        for ch in channels:
            pc = push_constants_new([width, height], cam, len(sc.triangles), len(sc.materials), [ox, oy], [tw, th], tt,
                                    bound.tpb, bound.offsets, ch, wavefront=wavefront)
            dispatch(bound, pc, chans[ch], radiance, hits, threads, counters, policy)
    comp = np.zeros((height, width, 4), np.uint8)
    lib().orc_composite(chans[0].ctypes.data, chans[1].ctypes.data, chans[2].ctypes.data, comp.ctypes.data,
                        width * height)
    return dict(channels=chans, composite=comp, radiance=radiance, hits=hits,
                counters=dict(aabb_tests=counters.aabb_tests, tri_tests=counters.tri_tests,
                              sphere_tests=counters.sphere_tests, rays=counters.rays))


def render_rects(scene, width, height, rects, camera=None, channels=(0, 1, 2), threads=None, policy=0, wavefront=None,
                 triangles_per_buffer=None, offsets=None):
    """Reference dispatches over arbitrary pixel rectangles [(x, y, w, h), ...] of the frame (each one
    `process_color_channel` call per channel with that tile offset / size, src/compute.rs:212-251): lets a test
    check a sparse set of rows or patches of a frame that is too large for the oracle as a whole.
    `triangles_per_buffer` / `offsets` override what goes into the push constants (the shader trusts them)."""
    bound = scene if isinstance(scene, BoundScene) else BoundScene(scene)
    sc = bound.scene
    cam = sc.camera if camera is None else camera
    chans = [np.zeros((height, width, 4), np.uint8) for _ in range(3)]
    radiance = np.zeros((height, width, 3), np.float32)
    hits = np.zeros((height, width), abi.HitRecord)
    mask = np.zeros((height, width), bool)
    counters = OrcCounters()
    for (x, y, w, h) in rects:
        mask[y:y + h, x:x + w] = True
        for ch in channels:
            pc = push_constants_new([width, height], cam, len(sc.triangles), len(sc.materials), [x, y], [w, h], [1, 1],
                                    bound.tpb if triangles_per_buffer is None else triangles_per_buffer,
                                    bound.offsets if offsets is None else offsets, ch, wavefront=wavefront)
            dispatch(bound, pc, chans[ch], radiance, hits, threads, counters, policy)
    comp = np.zeros((height, width, 4), np.uint8)
    lib().orc_composite(chans[0].ctypes.data, chans[1].ctypes.data, chans[2].ctypes.data, comp.ctypes.data,
                        width * height)
    return dict(channels=chans, composite=comp, radiance=radiance, hits=hits, mask=mask,
                counters=dict(aabb_tests=counters.aabb_tests, tri_tests=counters.tri_tests,
                              sphere_tests=counters.sphere_tests, rays=counters.rays))


def render_path(scene, width, height, spp=1, max_depth=0, base_seed=1, first_sample=0, accum=None, rows=None,
                threads=None, camera=None, policy=0, want_hits=True):
    """EXTENSION (SPEC.md): path-traced accumulation on the CPU.  Returns dict(accum HxWx4 f32,
    resolved HxWx4 u8, hits, counters)."""
    bound = scene if isinstance(scene, BoundScene) else BoundScene(scene)
    sc = bound.scene
    cam = sc.camera if camera is None else camera
    threads = threads or hardware_threads()
    reset = accum is None
    if accum is None:
        accum = np.zeros((height, width, 4), np.float32)
    hits = np.zeros((height, width), abi.HitRecord) if want_hits else None
    y0, y1 = rows if rows is not None else (0, height)
    pc = push_constants_new([width, height], cam, len(sc.triangles), len(sc.materials), [0, 0], [width, height], [1, 1],
                            bound.tpb, bound.offsets, 0, wavefront=(0, max_depth, base_seed))
    counters = OrcCounters()
    rc = lib().orc_render_path(C.byref(bound.bindings), pc.ctypes.data, spp, first_sample, max_depth, base_seed,
                               int(reset), y0, y1, accum.ctypes.data, hits.ctypes.data if want_hits else None, threads,
                               C.byref(counters), policy)
    assert rc == 0
    resolved = np.zeros((height, width, 4), np.uint8)
    with np.errstate(all="ignore"):
        lib().orc_resolve(accum.ctypes.data, resolved.ctypes.data, width * height)
    depth_rays = np.zeros(8, np.uint64)
    lib().orc_last_depth_rays(depth_rays.ctypes.data)
    return dict(accum=accum, resolved=resolved, hits=hits, depth_rays=depth_rays,
                counters=dict(aabb_tests=counters.aabb_tests, tri_tests=counters.tri_tests,
                              sphere_tests=counters.sphere_tests, rays=counters.rays))
