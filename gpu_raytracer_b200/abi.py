"""numpy dtypes mirroring include/rt_shared.h, i.e. the reference `shared` crate's
#[repr(C)] structs (shared/src/lib.rs:38-227).  Sizes are asserted at import."""
import numpy as np

f32, u32 = np.float32, np.uint32

Camera = np.dtype([("position", f32, 3), ("direction", f32, 3), ("up", f32, 3), ("fov", f32)])
Material = np.dtype([
    ("albedo", f32, 3), ("metallic_roughness_f16", u32), ("emission", f32, 3), ("ior_transmission_f16", u32),
    ("specular_factor", f32), ("specular_color", f32, 3), ("attenuation_distance", f32),
    ("attenuation_color", f32, 3), ("thickness_factor", f32), ("diffuse_factor", f32, 3),
    ("glossiness_factor", f32), ("material_type", u32), ("texture_indices", u32, 8), ("_padding", f32, 2)])
Light = np.dtype([("position", f32, 3), ("light_type", u32), ("color", f32, 3), ("intensity", f32),
                  ("direction", f32, 3), ("range_packed", u32), ("cone_angles_packed", u32)])
TextureInfo = np.dtype([("width", u32), ("height", u32), ("format", u32), ("mip_levels", u32), ("offset", u32),
                        ("size", u32), ("_padding", u32, 2)])
Sphere = np.dtype([("center", f32, 3), ("radius", f32), ("material_id", u32)])
Vertex = np.dtype([("position", f32, 3)])
Triangle = np.dtype([("v0_index", u32), ("v1_index", u32), ("v2_index", u32), ("material_id", u32)])
Aabb = np.dtype([("min", f32, 3), ("_padding0", f32), ("max", f32, 3), ("_padding1", f32)])
BvhNode = np.dtype([("bounds", Aabb), ("left_child", u32), ("right_child", u32), ("triangle_start", u32),
                    ("triangle_count", u32)])
WavefrontRay = np.dtype([("origin", f32, 3), ("ray_type", u32), ("direction", f32, 3), ("bounce_depth", u32),
                         ("throughput", f32, 3), ("medium_ior", f32), ("pixel_coord", u32, 2), ("inv_pdf", f32),
                         ("t_min", f32), ("t_max", f32), ("wavelength_channel", u32), ("active", u32)])
WavefrontCounters = np.dtype([("total_rays_generated", u32), ("rays_per_bounce", u32, 8),
                              ("active_bounce_depths", u32), ("max_bounce_depth", u32), ("frame_seed", u32),
                              ("_padding", u32, 3)])
SceneMetadataOffsets = np.dtype([
    ("spheres_offset", u32), ("spheres_count", u32), ("lights_offset", u32), ("lights_count", u32),
    ("bvh_nodes_offset", u32), ("bvh_nodes_count", u32), ("triangle_indices_offset", u32),
    ("triangle_indices_count", u32), ("vertices_offset", u32), ("vertices_count", u32)])
PushConstants = np.dtype([
    ("resolution", f32, 2), ("camera", Camera), ("triangle_count", u32), ("material_count", u32),
    ("tile_offset", u32, 2), ("tile_size_packed", u32), ("total_tiles", u32, 2), ("triangles_per_buffer", u32),
    ("metadata_offsets", SceneMetadataOffsets), ("packed_flags", u32), ("frame_seed", u32)])
HitRecord = np.dtype([("kind", u32), ("prim", u32), ("t_bits", u32), ("material_id", u32)])
FrameParams = np.dtype([("mode", u32), ("outputs", u32), ("spp", u32), ("first_sample", u32), ("max_depth", u32),
                        ("base_seed", u32), ("reset_accum", u32), ("tile_stride", u32), ("tile_phase", u32),
                        ("traversal", u32), ("_reserved", u32, 2)])

SIZES = {"Camera": (Camera, 40), "Material": (Material, 128), "Light": (Light, 52), "TextureInfo": (TextureInfo, 32),
         "Sphere": (Sphere, 20), "Vertex": (Vertex, 12), "Triangle": (Triangle, 16), "Aabb": (Aabb, 32),
         "BvhNode": (BvhNode, 48), "WavefrontRay": (WavefrontRay, 76), "WavefrontCounters": (WavefrontCounters, 60),
         "SceneMetadataOffsets": (SceneMetadataOffsets, 40), "PushConstants": (PushConstants, 128),
         "HitRecord": (HitRecord, 16), "FrameParams": (FrameParams, 48)}
for _name, (_dt, _sz) in SIZES.items():
    assert _dt.itemsize == _sz, (_name, _dt.itemsize, _sz)

# RaytracerConfig, shared/src/lib.rs:18-35
TILE_SIZE = 128
SPLIT_TILE = 64  # include/rtb200.h RT_SPLIT_TILE: unit of the multi-GPU screen split
STATIC_TILE_LIST_MAX = 16384  # include/rtb200.h RT_STATIC_TILE_LIST_MAX: rt_pack_tile_list_static scratch entries
THREAD_GROUP_SIZE = (16, 16)
MIN_RAY_DISTANCE = np.float32(0.00001)
MAX_TRIANGLES_PER_BUFFER = 8388608
INVALID_INDEX = 0xFFFFFFFF

OUT_ACCUM, OUT_HITS, OUT_CHANNELS, OUT_COMPOSITE, OUT_PEER = 1, 2, 4, 8, 16
MODE_LEGACY, MODE_WAVEFRONT0, MODE_PATH = 0, 1, 2
TRAVERSAL_AUTO, TRAVERSAL_LITERAL, TRAVERSAL_ORDERED, TRAVERSAL_REBUILT = 0, 1, 2, 3
BVH_AUTO, BVH_STANDARD, BVH_CHUNKED, BVH_SMALL_LEAF = 0, 1, 2, 3
