"""ctypes loader for librtb200.so.  The library is the product: if it is missing
this module raises — there is no Python or CPU fallback for the render path."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTB200_LIB", os.path.join(_HERE, "librtb200.so"))  # override: tuning builds only

# every symbol include/rtb200.h and include/rtb200_host.h declare
API_SYMBOLS = [
    "rt_create", "rt_destroy", "rt_last_error", "rt_resize", "rt_synchronize", "rt_update_scene_metadata",
    "rt_update_triangles", "rt_update_materials", "rt_update_textures", "rt_update_texture_data",
    "rt_mark_scene_metadata_dirty", "rt_mark_triangles_dirty", "rt_mark_materials_dirty", "rt_mark_textures_dirty",
    "rt_mark_texture_data_dirty", "rt_mark_all_dirty", "rt_needs_update", "rt_triangles_per_buffer",
    "rt_set_triangles_per_buffer",
    "rt_get_metadata_offsets", "rt_prepare_scene", "rt_dispatch", "rt_run_compute", "rt_render_frame", "rt_read_channel_rgba8",
    "rt_read_composite_rgba8", "rt_read_accum_f32", "rt_write_accum_f32", "rt_read_hit_records", "rt_device_accum",
    "rt_device_composite", "rt_count_next_launch", "rt_read_counters", "rt_last_kernel_ms", "rt_launch_count",
    "rt_last_traversal", "rt_last_build_ms", "rt_last_segments", "rt_last_wavefront_counters", "rt_last_stage_ms", "rt_last_launch_table", "rt_probe_gather_bandwidth",
    "rt_export_frame_handle", "rt_open_peer_frame", "rt_set_peer_frame", "rt_close_peer_frame", "rt_pack_tiles",
    "rt_unpack_tiles", "rt_set_tile_list", "rt_pack_tile_list", "rt_unpack_tile_list", "rt_pack_tile_list_static",
    "rt_stream"]
HOST_SYMBOLS = [
    "rt_f32_to_f16", "rt_f16_to_f32", "rt_camera_default", "rt_material_new", "rt_material_diffuse",
    "rt_material_metallic", "rt_material_glass", "rt_material_emissive", "rt_light_directional", "rt_light_point",
    "rt_light_spot", "rt_push_constants_new", "rt_push_constants_new_wavefront", "rt_default_scene",
    "rt_triangle_aabb", "rt_aabb_empty", "rt_aabb_union", "rt_aabb_center", "rt_aabb_surface_area",
    "rt_bvh_max_nodes", "rt_bvh_build", "rt_bvh_has_cycle", "rt_scene_terrain", "rt_scene_cornell", "rt_scene_field",
    "rt_gltf_load_file", "rt_gltf_load_memory", "rt_gltf_free", "rt_gltf_counts", "rt_gltf_triangles", "rt_gltf_vertices",
    "rt_gltf_materials", "rt_gltf_spheres", "rt_gltf_lights", "rt_gltf_cameras", "rt_gltf_textures",
    "rt_gltf_texture_data", "rt_gltf_warning", "rt_write_ppm_rgba8", "rt_write_png_rgba8", "rt_write_pfm_accum"]

# The host-side encoders / builders / loaders (include/rtb200_host.h: no CUDA in them) are also built as a library of
# their own, so that a process that only prepares scenes (tests of the host logic, the CPU reference arm of bench.py)
# never maps the CUDA kernels.  The product library still contains them: one .so is all a host links.
HOST_LIB_PATH = os.path.join(_HERE, "librtb200_host.so")
_lib = None
_host_lib = None


def _signatures():
    vp, u32, u64, i32, f32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_float
    return {
        "rt_create": (i32, [i32, C.POINTER(vp)]), "rt_destroy": (None, [vp]), "rt_last_error": (C.c_char_p, [vp]),
        "rt_resize": (i32, [vp, u32, u32]), "rt_synchronize": (i32, [vp]),
        "rt_update_scene_metadata": (i32, [vp, vp, u32, vp, u32, vp, u32, vp, u32, vp, u32, vp]),
        "rt_update_triangles": (i32, [vp, vp, u64]), "rt_update_materials": (i32, [vp, vp, u64]),
        "rt_update_textures": (i32, [vp, vp, u64]), "rt_update_texture_data": (i32, [vp, vp, u64]),
        "rt_mark_scene_metadata_dirty": (None, [vp]), "rt_mark_triangles_dirty": (None, [vp]),
        "rt_mark_materials_dirty": (None, [vp]), "rt_mark_textures_dirty": (None, [vp]),
        "rt_mark_texture_data_dirty": (None, [vp]), "rt_mark_all_dirty": (None, [vp]),
        "rt_needs_update": (i32, [vp]), "rt_triangles_per_buffer": (u32, [vp]),
        "rt_set_triangles_per_buffer": (i32, [vp, u32]),
        "rt_get_metadata_offsets": (i32, [vp, vp]), "rt_prepare_scene": (i32, [vp]), "rt_dispatch": (i32, [vp, vp]),
        "rt_run_compute": (i32, [vp, vp, i32]), "rt_render_frame": (i32, [vp, vp, vp]),
        "rt_read_channel_rgba8": (i32, [vp, u32, vp]), "rt_read_composite_rgba8": (i32, [vp, vp]),
        "rt_read_accum_f32": (i32, [vp, vp]), "rt_write_accum_f32": (i32, [vp, vp]),
        "rt_read_hit_records": (i32, [vp, vp]), "rt_device_accum": (vp, [vp]), "rt_device_composite": (vp, [vp]),
        "rt_last_kernel_ms": (f32, [vp]), "rt_count_next_launch": (i32, [vp]), "rt_read_counters": (i32, [vp, vp]),
        "rt_launch_count": (u64, [vp]), "rt_last_traversal": (i32, [vp]), "rt_last_build_ms": (f32, [vp]),
        "rt_last_segments": (i32, [vp, vp]), "rt_last_wavefront_counters": (i32, [vp, vp]), "rt_last_stage_ms": (i32, [vp, vp]), "rt_last_launch_table": (i32, [vp, vp, vp, u32]),
        "rt_probe_gather_bandwidth": (i32, [vp, u64, vp]), "rt_export_frame_handle": (i32, [vp, vp]),
        "rt_open_peer_frame": (i32, [vp, vp]), "rt_set_peer_frame": (i32, [vp, vp]),
        "rt_close_peer_frame": (i32, [vp]), "rt_pack_tiles": (i32, [vp, u32, u32, u32, vp]),
        "rt_unpack_tiles": (i32, [vp, u32, u32, u32, vp]), "rt_stream": (vp, [vp]),
        "rt_set_tile_list": (i32, [vp, vp, u32]), "rt_pack_tile_list": (i32, [vp, u32, vp, u32, vp]),
        "rt_unpack_tile_list": (i32, [vp, u32, vp, u32, vp]), "rt_pack_tile_list_static": (i32, [vp, u32, vp, u32, vp]),
        "rt_f32_to_f16": (C.c_uint16, [f32]), "rt_f16_to_f32": (f32, [C.c_uint16]),
        "rt_camera_default": (None, [vp]),
        "rt_material_new": (None, [vp, vp, f32, f32, vp, f32, f32]), "rt_material_diffuse": (None, [vp, vp]),
        "rt_material_metallic": (None, [vp, vp, f32]), "rt_material_glass": (None, [vp, vp, f32, f32]),
        "rt_material_emissive": (None, [vp, vp, vp]), "rt_light_directional": (None, [vp, vp, vp, f32]),
        "rt_light_point": (None, [vp, vp, vp, f32, f32]),
        "rt_light_spot": (None, [vp, vp, vp, vp, f32, f32, f32, f32]),
        "rt_push_constants_new": (None, [vp, vp, vp, u32, u32, vp, vp, vp, u32, vp, u32]),
        "rt_push_constants_new_wavefront": (None, [vp, vp, vp, u32, u32, vp, vp, vp, u32, vp, u32, u32, u32, u32]),
        "rt_default_scene": (None, [vp, vp, vp, vp, vp]), "rt_triangle_aabb": (None, [vp, vp, vp]),
        "rt_aabb_empty": (None, [vp]), "rt_aabb_union": (None, [vp, vp, vp]), "rt_aabb_center": (None, [vp, vp]),
        "rt_aabb_surface_area": (f32, [vp]), "rt_bvh_max_nodes": (u64, [u64]),
        "rt_bvh_build": (i32, [vp, u64, vp, u64, i32, vp, vp, vp, vp]), "rt_bvh_has_cycle": (i32, [vp, u32]),
        "rt_scene_terrain": (None, [u32, u32, u32, vp, vp, vp, vp, vp]),
        "rt_scene_cornell": (None, [vp, vp, vp, vp, vp, vp]),
        "rt_scene_field": (None, [u32, u32, u32, u32, u32, u32, vp, vp, vp, vp, vp]),
        "rt_gltf_load_file": (i32, [C.c_char_p, i32, C.POINTER(vp), C.c_char_p, u64]),
        "rt_gltf_load_memory": (i32, [vp, u64, C.c_char_p, i32, C.POINTER(vp), C.c_char_p, u64]),
        "rt_gltf_free": (None, [vp]), "rt_gltf_counts": (None, [vp, vp]),
        "rt_gltf_triangles": (vp, [vp]), "rt_gltf_vertices": (vp, [vp]), "rt_gltf_materials": (vp, [vp]),
        "rt_gltf_spheres": (vp, [vp]), "rt_gltf_lights": (vp, [vp]), "rt_gltf_cameras": (vp, [vp]),
        "rt_gltf_textures": (vp, [vp]), "rt_gltf_texture_data": (vp, [vp]), "rt_gltf_warning": (C.c_char_p, [vp, u64]),
        "rt_write_ppm_rgba8": (i32, [C.c_char_p, vp, u32, u32]), "rt_write_png_rgba8": (i32, [C.c_char_p, vp, u32, u32]),
        "rt_write_pfm_accum": (i32, [C.c_char_p, vp, u32, u32]),
    }


def _declare(lib, names):
    sig = _signatures()
    for name in names:
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype, fn.argtypes = sig[name]
    return lib


def load():
    """Load librtb200.so (once) and declare its prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  gpu_raytracer_b200 has no fallback path.")
    _lib = _declare(C.CDLL(LIB_PATH), API_SYMBOLS + HOST_SYMBOLS)
    return _lib


def load_host():
    """The host-only library (scene encoders, BVH builder, glTF loader, image writers); the product library if the
    host-only one has not been built (it holds the same functions) or a tuning build is selected."""
    global _host_lib
    if _host_lib is None:
        if os.path.exists(HOST_LIB_PATH) and "RTB200_LIB" not in os.environ:
            _host_lib = _declare(C.CDLL(HOST_LIB_PATH), HOST_SYMBOLS)
        else:
            _host_lib = load()
    return _host_lib
