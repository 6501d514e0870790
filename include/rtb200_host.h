/* rtb200_host.h — host-side encoders, BVH builders and procedural scenes of
 * librtb200.so (pure C++, no CUDA needed to call them).  These mirror the host
 * code either side of the hot path: the `shared` crate's constructors/packers
 * (shared/src/lib.rs) and the BVH builder (src/bvh.rs). */
#ifndef RTB200_HOST_H
#define RTB200_HOST_H
#include <stdint.h>
#include "rt_shared.h"
#ifdef __cplusplus
extern "C" {
#endif

/* half 1.8.3 `f16::from_f32(..).to_bits()` / `to_f32` (IEEE RNE), shared/src/lib.rs:250-252 */
uint16_t rt_f32_to_f16(float v);
float rt_f16_to_f32(uint16_t bits);

/* Camera::new, shared/src/lib.rs:231-238 */
void rt_camera_default(RtCamera* out);
/* Material::new, shared/src/lib.rs:255-291 (+ diffuse/metallic/glass/emissive :316-332) */
void rt_material_new(RtMaterial* out, const float albedo[3], float metallic, float roughness,
                     const float emission[3], float ior, float transmission);
void rt_material_diffuse(RtMaterial* out, const float albedo[3]);
void rt_material_metallic(RtMaterial* out, const float albedo[3], float roughness);
void rt_material_glass(RtMaterial* out, const float albedo[3], float ior, float transmission);
void rt_material_emissive(RtMaterial* out, const float albedo[3], const float emission[3]);
/* Light::{directional,point,spot}, shared/src/lib.rs:497-586 */
void rt_light_directional(RtLight* out, const float direction[3], const float color[3], float intensity);
void rt_light_point(RtLight* out, const float position[3], const float color[3], float intensity, float range);
void rt_light_spot(RtLight* out, const float position[3], const float direction[3], const float color[3],
                   float intensity, float range, float inner_cone, float outer_cone);
/* PushConstants::new / new_wavefront, shared/src/lib.rs:1076-1131 */
void rt_push_constants_new(RtPushConstants* out, const float resolution[2], const RtCamera* camera,
                           uint32_t triangle_count, uint32_t material_count, const uint32_t tile_offset[2],
                           const uint32_t tile_size[2], const uint32_t total_tiles[2],
                           uint32_t triangles_per_buffer, const RtSceneMetadataOffsets* offsets,
                           uint32_t color_channel);
void rt_push_constants_new_wavefront(RtPushConstants* out, const float resolution[2], const RtCamera* camera,
                                     uint32_t triangle_count, uint32_t material_count,
                                     const uint32_t tile_offset[2], const uint32_t tile_size[2],
                                     const uint32_t total_tiles[2], uint32_t triangles_per_buffer,
                                     const RtSceneMetadataOffsets* offsets, uint32_t color_channel,
                                     uint32_t current_bounce_depth, uint32_t max_bounce_depth, uint32_t frame_seed);
/* SceneBuilder::build_default_scene, shared/src/lib.rs:1242-1286: 6 spheres,
 * 2 triangles, 6 vertices, 4 materials, 1 light.  Arrays must have that room. */
void rt_default_scene(RtSphere spheres[6], RtTriangle triangles[2], RtVertex vertices[6], RtMaterial materials[4],
                      RtLight lights[1]);
/* Triangle::bounding_box (shared/src/lib.rs), Aabb::{empty,union,center,surface_area} :752-801 */
void rt_triangle_aabb(const RtTriangle* t, const RtVertex* vertices, RtAabb* out);
void rt_aabb_empty(RtAabb* out);
void rt_aabb_union(const RtAabb* a, const RtAabb* b, RtAabb* out);
void rt_aabb_center(const RtAabb* a, float out[3]);
float rt_aabb_surface_area(const RtAabb* a);

/* ---- BVH builders emitting the reference node format (src/bvh.rs) */
enum {
    RT_BVH_AUTO = 0,     /* BvhBuilder::build, src/bvh.rs:104-122: >100 000 triangles -> chunked, else standard */
    RT_BVH_STANDARD = 1, /* build_standard format (:125-151, :298-375): pre-order, one triangle per leaf */
    RT_BVH_CHUNKED = 2,  /* build_chunked (:154-247), restated exactly */
    RT_BVH_SMALL_LEAF = 3/* same node format, binned SAH with <=4 triangles per leaf, any size */
};
/* upper bound on nodes the builders write for n triangles */
uint64_t rt_bvh_max_nodes(uint64_t n_triangles);
/* Returns 0 on success.  nodes must hold rt_bvh_max_nodes(n) entries, indices n entries. */
int rt_bvh_build(const RtTriangle* triangles, uint64_t n_triangles, const RtVertex* vertices, uint64_t n_vertices,
                 int strategy, RtBvhNode* nodes, uint32_t* n_nodes, uint32_t* indices, uint32_t* n_indices);

/* 1 if the child links reachable from node 0 form a cycle (the reference traversal, shader/src/bvh.rs:40-85,
 * would never terminate on such a tree; rt_update_scene_metadata refuses it), else 0. */
int rt_bvh_has_cycle(const RtBvhNode* nodes, uint32_t n_nodes);

/* ---- procedural scenes (bench / test inputs) */
/* Terrain(nx, nz, seed): (nx+1)(nz+1) vertices, 2*nx*nz triangles row-major, materials striped by row
 * (diffuse / metallic / glass).  vertices: (nx+1)*(nz+1); triangles: 2*nx*nz; materials: 3; lights: 2. */
void rt_scene_terrain(uint32_t nx, uint32_t nz, uint32_t seed, RtVertex* vertices, RtTriangle* triangles,
                      RtMaterial materials[3], RtLight lights[2], RtCamera* camera);
/* Cornell box: 5 walls + 2 boxes + ceiling light quad = 36 triangles, 4 materials, 1 light.
 * vertices: room for 72; returns the vertex count through n_vertices. */
/* Field: a non-height-field test scene (tree-quality study): n_tori closed tori of seg x ring quads at mixed scales
 * and orientations scattered through a volume, above a flat ground grid.  Vertices n_tori*seg*ring + (gnx+1)(gnz+1),
 * triangles 2*n_tori*seg*ring + 2*gnx*gnz.  No reference analogue (test/bench input). */
void rt_scene_field(uint32_t n_tori, uint32_t seg, uint32_t ring, uint32_t gnx, uint32_t gnz, uint32_t seed,
                    RtVertex* vertices, RtTriangle* triangles, RtMaterial materials[3], RtLight lights[2], RtCamera* camera);
void rt_scene_cornell(RtVertex* vertices, uint32_t* n_vertices, RtTriangle triangles[36], RtMaterial materials[4],
                      RtLight lights[1], RtCamera* camera);

/* ---- glTF 2.0 / GLB ingestion (src/gltf_loader.rs:53-590, src/scene.rs:43-69): JSON + external
 * .bin / base64 / GLB buffers, node hierarchy (matrix or TRS), TRIANGLES / STRIP / FAN with u8 / u16 /
 * u32 or no indices, bit-exact position de-duplication per primitive, cameras, KHR_lights_punctual,
 * pbrMetallicRoughness + KHR_materials_{pbrSpecularGlossiness,transmission,ior,specular,volume}.
 * Images are not decoded (dimensions from the PNG/JPEG header, RGBA8 block zero-filled): the shader
 * never reads textures.  scene_index < 0 = the default scene.  Returns 0, or -1 with a message in err. */
typedef struct RtGltfScene RtGltfScene;
typedef struct RtGltfCounts {
    uint64_t triangles, vertices, materials, spheres, lights, cameras, textures, texture_bytes, scenes, warnings;
} RtGltfCounts;
int rt_gltf_load_file(const char* path, int scene_index, RtGltfScene** out, char* err, uint64_t err_len);
int rt_gltf_load_memory(const uint8_t* data, uint64_t n, const char* base_dir, int scene_index, RtGltfScene** out,
                        char* err, uint64_t err_len);
void rt_gltf_free(RtGltfScene* scene);
void rt_gltf_counts(const RtGltfScene* scene, RtGltfCounts* out);
const RtTriangle* rt_gltf_triangles(const RtGltfScene* scene);
const RtVertex* rt_gltf_vertices(const RtGltfScene* scene);
const RtMaterial* rt_gltf_materials(const RtGltfScene* scene);
const RtSphere* rt_gltf_spheres(const RtGltfScene* scene);   /* always empty, like LoadedScene.spheres */
const RtLight* rt_gltf_lights(const RtGltfScene* scene);
const RtCamera* rt_gltf_cameras(const RtGltfScene* scene);
const RtTextureInfo* rt_gltf_textures(const RtGltfScene* scene);
const uint8_t* rt_gltf_texture_data(const RtGltfScene* scene);
const char* rt_gltf_warning(const RtGltfScene* scene, uint64_t i);

/* ---- output stage (the reference presents to a window only, shader/src/lib.rs:368-391): image files
 * from the buffers rt_read_composite_rgba8 / rt_read_channel_rgba8 / rt_read_accum_f32 return. */
int rt_write_ppm_rgba8(const char* path, const uint8_t* rgba, uint32_t width, uint32_t height);
int rt_write_png_rgba8(const char* path, const uint8_t* rgba, uint32_t width, uint32_t height);
int rt_write_pfm_accum(const char* path, const float* accum, uint32_t width, uint32_t height);

#ifdef __cplusplus
}
#endif
#endif
