"""Python mirror of the reference's dispatch surface over the C ABI:
`BufferManager` (src/buffers.rs) and `ComputeRenderer` (src/compute.rs).
Every method is one call into librtb200.so; nothing is computed in Python."""
import ctypes as C

import numpy as np

from . import abi
from .native import load


class RtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"librtb200 error {code}: {msg}")
        self.code = code


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


class _CudaArray:
    """Carrier of `__cuda_array_interface__` (version 3) for a device buffer a context owns."""

    def __init__(self, ptr, shape, typestr, owner):
        self._owner = owner  # keeps the context alive as long as the view
        self.__cuda_array_interface__ = {"shape": tuple(int(x) for x in shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 3, "strides": None}


class Context:
    """RenderState.device/queue + output textures (src/renderer.rs:93-246, 452-475)."""

    def __init__(self, device=0, width=None, height=None):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.rt_create(device, C.byref(h))
        if rc != 0:
            raise RtError(rc, self.lib.rt_last_error(None).decode())
        self.handle = h
        self.width = self.height = 0
        if width:
            self.resize(width, height)

    def _ck(self, rc):
        if rc < 0:
            raise RtError(rc, self.lib.rt_last_error(self.handle).decode())
        return rc

    def close(self):
        if getattr(self, "handle", None):
            self.lib.rt_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def resize(self, width, height):
        self._ck(self.lib.rt_resize(self.handle, width, height))
        self.width, self.height = width, height

    def synchronize(self):
        self._ck(self.lib.rt_synchronize(self.handle))

    # ---- read-back
    def read_channel(self, ch):
        out = np.empty((self.height, self.width, 4), np.uint8)
        self._ck(self.lib.rt_read_channel_rgba8(self.handle, ch, _p(out)))
        return out

    def read_composite(self, out=None):
        if out is None:
            out = np.empty((self.height, self.width, 4), np.uint8)
        self._ck(self.lib.rt_read_composite_rgba8(self.handle, _p(out)))
        return out

    def read_accum(self):
        out = np.empty((self.height, self.width, 4), np.float32)
        self._ck(self.lib.rt_read_accum_f32(self.handle, _p(out)))
        return out

    def write_accum(self, a):
        a = np.ascontiguousarray(a, np.float32)
        assert a.shape == (self.height, self.width, 4)
        self._ck(self.lib.rt_write_accum_f32(self.handle, _p(a)))

    def read_hits(self):
        out = np.empty((self.height, self.width), abi.HitRecord)
        self._ck(self.lib.rt_read_hit_records(self.handle, _p(out)))
        return out

    @property
    def stream(self):
        return int(self.lib.rt_stream(self.handle) or 0)

    def pack_tiles(self, what, stride, phase, dst_device_ptr):
        self._ck(self.lib.rt_pack_tiles(self.handle, what, stride, phase, C.c_void_p(dst_device_ptr)))

    def unpack_tiles(self, what, stride, phase, src_device_ptr):
        self._ck(self.lib.rt_unpack_tiles(self.handle, what, stride, phase, C.c_void_p(src_device_ptr)))

    def set_tile_list(self, tiles):
        """Explicit tile assignment of this context (multi-GPU load balancing); empty / None clears it."""
        t = np.ascontiguousarray(tiles if tiles is not None else [], np.uint32)
        self._ck(self.lib.rt_set_tile_list(self.handle, _p(t), len(t)))

    def pack_tile_list(self, what, tiles, dst_device_ptr):
        t = np.ascontiguousarray(tiles, np.uint32)
        self._ck(self.lib.rt_pack_tile_list(self.handle, what, _p(t), len(t), C.c_void_p(dst_device_ptr)))

    def pack_tile_list_static(self, what, tiles, dst_device_ptr):
        """rt_pack_tile_list_static: no allocation (static device scratch, <= abi.STATIC_TILE_LIST_MAX tiles).  The list
        is kept alive on this object until the next call: the copy is stream-ordered."""
        self._static_list = np.ascontiguousarray(tiles, np.uint32)
        self._ck(self.lib.rt_pack_tile_list_static(self.handle, what, _p(self._static_list), len(self._static_list),
                                                   C.c_void_p(dst_device_ptr)))

    def unpack_tile_list(self, what, tiles, src_device_ptr):
        t = np.ascontiguousarray(tiles, np.uint32)
        self._ck(self.lib.rt_unpack_tile_list(self.handle, what, _p(t), len(t), C.c_void_p(src_device_ptr)))

    def count_next_launch(self):
        self._ck(self.lib.rt_count_next_launch(self.handle))

    def read_counters(self):
        out = np.zeros(9, np.uint64)
        self._ck(self.lib.rt_read_counters(self.handle, _p(out)))
        return dict(box_tests=int(out[0]), tri_tests=int(out[1]), segments=int(out[2]), shaded=int(out[3]),
                    literal_fallbacks=int(out[4]), s_phases=int(out[5]), s_ready_lanes=int(out[6]),
                    t_iterations=int(out[7]), t_active_lanes=int(out[8]))

    # ---- fused multi-GPU epilogue (peer stores into rank 0's composite)
    def export_frame_handle(self):
        h = (C.c_uint8 * 64)()
        self._ck(self.lib.rt_export_frame_handle(self.handle, h))
        return bytes(h)

    def open_peer_frame(self, handle64):
        buf = (C.c_uint8 * 64).from_buffer_copy(handle64)
        self._ck(self.lib.rt_open_peer_frame(self.handle, buf))

    def set_peer_frame(self, device_ptr):
        self._ck(self.lib.rt_set_peer_frame(self.handle, C.c_void_p(device_ptr)))

    def close_peer_frame(self):
        self._ck(self.lib.rt_close_peer_frame(self.handle))

    @property
    def device_composite(self):
        return int(self.lib.rt_device_composite(self.handle))

    @property
    def device_accum(self):
        return int(self.lib.rt_device_accum(self.handle))

    def composite_cuda_array(self):
        """Zero-copy view of the composite image (rt_device_composite) for CUDA-array consumers: an object carrying
        `__cuda_array_interface__` (shape (H, W, 4) uint8), e.g. `torch.as_tensor(ctx.composite_cuda_array(),
        device="cuda")`.  A host that reads frames back without stalling enqueues its own copy of this view on
        `ctx.stream` (stream-ordered after the frame) instead of the waiting rt_read_composite_rgba8.  The view is
        valid until the next rt_resize / rt_destroy."""
        return _CudaArray(self.device_composite, (self.height, self.width, 4), "|u1", self)

    def accum_cuda_array(self):
        """The accumulation buffer (rt_device_accum) the same way: shape (H, W, 4) float32 (rgb sums, sample count)."""
        return _CudaArray(self.device_accum, (self.height, self.width, 4), "<f4", self)

    def probe_gather_bandwidth(self, footprint_bytes):
        """GB/s of random independent 32-byte sector loads over a footprint (traversal roofline denominators)."""
        out = C.c_float(0.0)
        self._ck(self.lib.rt_probe_gather_bandwidth(self.handle, int(footprint_bytes), C.byref(out)))
        return float(out.value)

    def last_stage_ms(self):
        """Per-stage GPU ms of the last wave-engine frame: dict(generate, extend, shade, resolve, launches)."""
        out = np.zeros(8, np.float32)
        self._ck(self.lib.rt_last_stage_ms(self.handle, _p(out)))
        return dict(generate=float(out[0]), extend=float(out[1]), shade=float(out[2]), resolve=float(out[3]),
                    launches=[int(x) for x in out[4:8]])

    def last_launch_table(self):
        """[(stage name, ms)] of every launch interval of the last single-sub-frame wave-engine frame."""
        ms, st = np.zeros(256, np.float32), np.zeros(256, np.int32)
        n = self._ck(self.lib.rt_last_launch_table(self.handle, _p(ms), _p(st), 256))
        names = ("generate", "extend", "shade", "resolve", "shade-split", "fallback")
        return [(names[int(st[i])], float(ms[i])) for i in range(min(n, 256))]

    @property
    def last_kernel_ms(self):
        return float(self.lib.rt_last_kernel_ms(self.handle))

    @property
    def launch_count(self):
        return int(self.lib.rt_launch_count(self.handle))

    @property
    def last_segments(self):
        out = C.c_uint64(0)
        self._ck(self.lib.rt_last_segments(self.handle, C.byref(out)))
        return int(out.value)

    def last_wavefront_counters(self):
        """WavefrontCounters of the last wave-engine path frame (shared/src/lib.rs:187-194)."""
        out = np.zeros((), abi.WavefrontCounters)
        self._ck(self.lib.rt_last_wavefront_counters(self.handle, out.ctypes.data_as(C.c_void_p)))
        return out

    @property
    def last_build_ms(self):
        return float(self.lib.rt_last_build_ms(self.handle))

    @property
    def last_traversal(self):
        return int(self.lib.rt_last_traversal(self.handle))


class BufferManager:
    """src/buffers.rs — same method names, same return values (`resized` bools)."""

    def __init__(self, ctx, triangles_per_buffer=None):
        """`triangles_per_buffer`: BufferManager::max_triangles_per_buffer (src/buffers.rs:49-53) for a host
        built with another buffer-size limit; only before the first triangle upload."""
        self.ctx = ctx
        if triangles_per_buffer is not None:
            ctx._ck(ctx.lib.rt_set_triangles_per_buffer(ctx.handle, int(triangles_per_buffer)))
        self.triangles_per_buffer = int(ctx.lib.rt_triangles_per_buffer(ctx.handle))

    def update_scene_metadata(self, spheres, lights, bvh_nodes, triangle_indices, vertices):
        c = self.ctx
        arrs = [np.ascontiguousarray(spheres, abi.Sphere), np.ascontiguousarray(lights, abi.Light),
                np.ascontiguousarray(bvh_nodes, abi.BvhNode), np.ascontiguousarray(triangle_indices, np.uint32),
                np.ascontiguousarray(vertices, abi.Vertex)]
        off = np.zeros((), abi.SceneMetadataOffsets)
        rc = c._ck(c.lib.rt_update_scene_metadata(c.handle, _p(arrs[0]), len(arrs[0]), _p(arrs[1]), len(arrs[1]),
                                                  _p(arrs[2]), len(arrs[2]), _p(arrs[3]), len(arrs[3]), _p(arrs[4]),
                                                  len(arrs[4]), off.ctypes.data_as(C.c_void_p)))
        return bool(rc), off

    def update_triangles(self, triangles):
        t = np.ascontiguousarray(triangles, abi.Triangle)
        return bool(self.ctx._ck(self.ctx.lib.rt_update_triangles(self.ctx.handle, _p(t), len(t))))

    def update_materials(self, materials):
        m = np.ascontiguousarray(materials, abi.Material)
        return bool(self.ctx._ck(self.ctx.lib.rt_update_materials(self.ctx.handle, _p(m), len(m))))

    def update_textures(self, textures):
        t = np.ascontiguousarray(textures, abi.TextureInfo)
        return bool(self.ctx._ck(self.ctx.lib.rt_update_textures(self.ctx.handle, _p(t), len(t))))

    def update_texture_data(self, data):
        d = np.ascontiguousarray(data, np.uint8)
        return bool(self.ctx._ck(self.ctx.lib.rt_update_texture_data(self.ctx.handle, _p(d), d.size)))

    def mark_scene_metadata_dirty(self):
        self.ctx.lib.rt_mark_scene_metadata_dirty(self.ctx.handle)

    def mark_triangles_dirty(self):
        self.ctx.lib.rt_mark_triangles_dirty(self.ctx.handle)

    def mark_materials_dirty(self):
        self.ctx.lib.rt_mark_materials_dirty(self.ctx.handle)

    def mark_textures_dirty(self):
        self.ctx.lib.rt_mark_textures_dirty(self.ctx.handle)

    def mark_texture_data_dirty(self):
        self.ctx.lib.rt_mark_texture_data_dirty(self.ctx.handle)

    def mark_all_dirty(self):
        self.ctx.lib.rt_mark_all_dirty(self.ctx.handle)

    def needs_update(self):
        return bool(self.ctx.lib.rt_needs_update(self.ctx.handle))

    def prepare_scene(self):
        """rt_prepare_scene: run the device-side re-layout / LBVH build of the uploaded scene now."""
        self.ctx._ck(self.ctx.lib.rt_prepare_scene(self.ctx.handle))

    def upload_scene(self, scene, prepare=False):
        """ComputeRenderer::update_buffers_and_bind_groups (src/compute.rs:109-134)."""
        _, off = self.update_scene_metadata(scene.spheres, scene.lights, scene.bvh_nodes, scene.triangle_indices,
                                            scene.vertices)
        self.update_triangles(scene.triangles)
        self.update_materials(scene.materials)
        self.update_textures(scene.textures)
        self.update_texture_data(scene.texture_data)
        if prepare:
            self.prepare_scene()
        return off


class ComputeRenderer:
    """src/compute.rs — run_compute (progressive), one literal dispatch, and the fused frame."""

    def __init__(self, ctx):
        self.ctx = ctx

    def dispatch(self, push_constants):
        pc = np.ascontiguousarray(push_constants, abi.PushConstants)
        self.ctx._ck(self.ctx.lib.rt_dispatch(self.ctx.handle, pc.ctypes.data_as(C.c_void_p)))

    def run_compute(self, camera, needs_recompute=False):
        cam = np.ascontiguousarray(camera, abi.Camera)
        return self.ctx._ck(self.ctx.lib.rt_run_compute(self.ctx.handle, cam.ctypes.data_as(C.c_void_p),
                                                        int(needs_recompute)))

    def render_frame(self, camera, mode=abi.MODE_LEGACY, outputs=abi.OUT_ACCUM | abi.OUT_COMPOSITE, spp=1,
                     first_sample=0, max_depth=0, base_seed=1, reset_accum=True, tile_stride=0, tile_phase=0,
                     traversal=abi.TRAVERSAL_AUTO, variant=0):
        cam = np.ascontiguousarray(camera, abi.Camera)
        fp = np.zeros((), abi.FrameParams)
        fp["mode"], fp["outputs"], fp["spp"], fp["first_sample"] = mode, outputs, spp, first_sample
        fp["max_depth"], fp["base_seed"], fp["reset_accum"] = max_depth, base_seed, int(reset_accum)
        fp["tile_stride"], fp["tile_phase"], fp["traversal"] = tile_stride, tile_phase, traversal
        fp["_reserved"][0] = variant  # path mode engine: 0 auto, 1 megakernel, 4 streaming wavefront (wave)
        self.ctx._ck(self.ctx.lib.rt_render_frame(self.ctx.handle, cam.ctypes.data_as(C.c_void_p),
                                                  fp.ctypes.data_as(C.c_void_p)))
