/* rt_shared.h — the reference `shared` crate's #[repr(C)] ABI as plain C structs.
 *
 * Every struct here mirrors, byte for byte, a `#[repr(C)]` Pod struct of
 * /root/reference/shared/src/lib.rs (cited per struct).  Sizes and offsets are
 * static-asserted against the values the Rust layout rules give (align 4
 * everywhere, no implicit padding).  Usable from C, C++ and CUDA.
 */
#ifndef RT_SHARED_H
#define RT_SHARED_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* RaytracerConfig, shared/src/lib.rs:18-35 */
#define RT_TILE_SIZE 128u
#define RT_THREAD_GROUP_X 16u
#define RT_THREAD_GROUP_Y 16u
#define RT_MIN_RAY_DISTANCE 0.00001f
#define RT_MAX_PUSH_CONSTANT_SIZE 128u
#define RT_INVALID_INDEX 0xFFFFFFFFu
/* src/buffers.rs:49-53: 128 MiB / sizeof(Triangle) */
#define RT_MAX_TRIANGLES_PER_BUFFER 8388608u

/* shared/src/lib.rs:38-45 */
typedef struct RtCamera {
    float position[3];
    float direction[3];
    float up[3];
    float fov; /* degrees */
} RtCamera;

/* shared/src/lib.rs:51-66 */
typedef struct RtMaterial {
    float albedo[3];
    uint32_t metallic_roughness_f16; /* metallic low16, roughness high16 */
    float emission[3];
    uint32_t ior_transmission_f16;   /* ior low16, transmission high16 */
    float specular_factor;
    float specular_color[3];
    float attenuation_distance;
    float attenuation_color[3];
    float thickness_factor;
    float diffuse_factor[3];
    float glossiness_factor;
    uint32_t material_type;
    uint32_t texture_indices[8];
    float _padding[2];
} RtMaterial;

/* shared/src/lib.rs:72-82 */
typedef struct RtLight {
    float position[3];
    uint32_t light_type; /* 0 directional, 1 point, 2 spot */
    float color[3];
    float intensity;
    float direction[3];
    uint32_t range_packed;
    uint32_t cone_angles_packed;
} RtLight;

/* shared/src/lib.rs:87-95 */
typedef struct RtTextureInfo {
    uint32_t width, height, format, mip_levels, offset, size;
    uint32_t _padding[2];
} RtTextureInfo;

/* shared/src/lib.rs:101-106 */
typedef struct RtSphere {
    float center[3];
    float radius;
    uint32_t material_id;
} RtSphere;

/* shared/src/lib.rs:112-115 */
typedef struct RtVertex {
    float position[3];
} RtVertex;

/* shared/src/lib.rs:121-127 */
typedef struct RtTriangle {
    uint32_t v0_index, v1_index, v2_index, material_id;
} RtTriangle;

/* shared/src/lib.rs:145-150 */
typedef struct RtAabb {
    float min[3];
    float _padding0;
    float max[3];
    float _padding1;
} RtAabb;

/* shared/src/lib.rs:155-161 */
typedef struct RtBvhNode {
    RtAabb bounds;
    uint32_t left_child;   /* 0xFFFFFFFF if leaf */
    uint32_t right_child;
    uint32_t triangle_start;
    uint32_t triangle_count;
} RtBvhNode;

/* shared/src/lib.rs:167-181 */
typedef struct RtWavefrontRay {
    float origin[3];
    uint32_t ray_type;
    float direction[3];
    uint32_t bounce_depth;
    float throughput[3];
    float medium_ior;
    uint32_t pixel_coord[2];
    float inv_pdf;
    float t_min;
    float t_max;
    uint32_t wavelength_channel;
    uint32_t active;
} RtWavefrontRay;

/* shared/src/lib.rs:187-194 */
typedef struct RtWavefrontCounters {
    uint32_t total_rays_generated;
    uint32_t rays_per_bounce[8];
    uint32_t active_bounce_depths;
    uint32_t max_bounce_depth;
    uint32_t frame_seed;
    uint32_t _padding[3];
} RtWavefrontCounters;

/* shared/src/lib.rs:199-210 — offsets are in u32 units */
typedef struct RtSceneMetadataOffsets {
    uint32_t spheres_offset, spheres_count;
    uint32_t lights_offset, lights_count;
    uint32_t bvh_nodes_offset, bvh_nodes_count;
    uint32_t triangle_indices_offset, triangle_indices_count;
    uint32_t vertices_offset, vertices_count;
} RtSceneMetadataOffsets;

/* shared/src/lib.rs:215-227 */
typedef struct RtPushConstants {
    float resolution[2];
    RtCamera camera;
    uint32_t triangle_count;
    uint32_t material_count;
    uint32_t tile_offset[2];
    uint32_t tile_size_packed; /* width low16, height high16 */
    uint32_t total_tiles[2];
    uint32_t triangles_per_buffer;
    RtSceneMetadataOffsets metadata_offsets;
    uint32_t packed_flags; /* channel 0-7, cur depth 8-15, max depth 16-23, mode 24-31 */
    uint32_t frame_seed;
} RtPushConstants;

#ifdef __cplusplus
#define RT_SA(c, m) static_assert(c, m)
#else
#define RT_SA(c, m) _Static_assert(c, m)
#endif
RT_SA(sizeof(RtCamera) == 40, "Camera");
RT_SA(sizeof(RtMaterial) == 128, "Material");
RT_SA(offsetof(RtMaterial, metallic_roughness_f16) == 12, "Material.mr");
RT_SA(offsetof(RtMaterial, emission) == 16, "Material.emission");
RT_SA(offsetof(RtMaterial, ior_transmission_f16) == 28, "Material.ior_tr");
RT_SA(offsetof(RtMaterial, specular_factor) == 32, "Material.spec");
RT_SA(offsetof(RtMaterial, material_type) == 84, "Material.type");
RT_SA(offsetof(RtMaterial, texture_indices) == 88, "Material.tex");
RT_SA(sizeof(RtLight) == 52, "Light");
RT_SA(offsetof(RtLight, direction) == 32, "Light.direction");
RT_SA(offsetof(RtLight, range_packed) == 44, "Light.range");
RT_SA(offsetof(RtLight, cone_angles_packed) == 48, "Light.cones");
RT_SA(sizeof(RtTextureInfo) == 32, "TextureInfo");
RT_SA(sizeof(RtSphere) == 20, "Sphere");
RT_SA(sizeof(RtVertex) == 12, "Vertex");
RT_SA(sizeof(RtTriangle) == 16, "Triangle");
RT_SA(sizeof(RtAabb) == 32, "Aabb");
RT_SA(sizeof(RtBvhNode) == 48, "BvhNode");
RT_SA(offsetof(RtBvhNode, left_child) == 32, "BvhNode.left");
RT_SA(sizeof(RtWavefrontRay) == 76, "WavefrontRay");
RT_SA(offsetof(RtWavefrontRay, active) == 72, "WavefrontRay.active");
RT_SA(sizeof(RtWavefrontCounters) == 60, "WavefrontCounters");
RT_SA(sizeof(RtSceneMetadataOffsets) == 40, "SceneMetadataOffsets");
RT_SA(sizeof(RtPushConstants) == 128, "PushConstants");
RT_SA(offsetof(RtPushConstants, camera) == 8, "PC.camera");
RT_SA(offsetof(RtPushConstants, triangle_count) == 48, "PC.triangle_count");
RT_SA(offsetof(RtPushConstants, tile_offset) == 56, "PC.tile_offset");
RT_SA(offsetof(RtPushConstants, tile_size_packed) == 64, "PC.tile_size");
RT_SA(offsetof(RtPushConstants, total_tiles) == 68, "PC.total_tiles");
RT_SA(offsetof(RtPushConstants, triangles_per_buffer) == 76, "PC.tpb");
RT_SA(offsetof(RtPushConstants, metadata_offsets) == 80, "PC.offsets");
RT_SA(offsetof(RtPushConstants, packed_flags) == 120, "PC.flags");
RT_SA(offsetof(RtPushConstants, frame_seed) == 124, "PC.seed");

/* PushConstants::pack_tile_size, shared/src/lib.rs:1135-1139 */
static inline uint32_t rt_pack_tile_size(uint32_t w, uint32_t h) {
    uint32_t ww = w < 65535u ? w : 65535u, hh = h < 65535u ? h : 65535u;
    return (ww & 0xFFFFu) | ((hh & 0xFFFFu) << 16);
}
/* PushConstants::pack_flags, shared/src/lib.rs:1152-1159 */
static inline uint32_t rt_pack_flags(uint32_t channel, uint32_t cur_depth,
                                     uint32_t max_depth, uint32_t mode) {
    return (channel & 0xFFu) | ((cur_depth & 0xFFu) << 8) |
           ((max_depth & 0xFFu) << 16) | ((mode & 0xFFu) << 24);
}
static inline uint32_t rt_pc_color_channel(const RtPushConstants* p) { return p->packed_flags & 0xFFu; }
static inline uint32_t rt_pc_current_bounce(const RtPushConstants* p) { return (p->packed_flags >> 8) & 0xFFu; }
static inline uint32_t rt_pc_max_bounce(const RtPushConstants* p) { return (p->packed_flags >> 16) & 0xFFu; }
static inline uint32_t rt_pc_wavefront_mode(const RtPushConstants* p) { return (p->packed_flags >> 24) & 0xFFu; }

/* TileHelper, shared/src/lib.rs:1187-1203 */
static inline void rt_tile_count(uint32_t w, uint32_t h, uint32_t tile, uint32_t* tx, uint32_t* ty) {
    *tx = (w + tile - 1) / tile;
    *ty = (h + tile - 1) / tile;
}
static inline uint32_t rt_tiles_per_frame(uint32_t total) {
    uint32_t r;
    if (total <= 16u) r = total;
    else if (total <= 64u) r = total / 8u;
    else if (total <= 256u) r = total / 32u;
    else if (total <= 1024u) r = total / 64u;
    else r = 1u;
    return r < 1u ? 1u : r;
}

#ifdef __cplusplus
}
#endif
#endif /* RT_SHARED_H */
