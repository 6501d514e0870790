// rt_host.cpp — host-side half of librtb200.so: the `shared` crate's encoders
// (shared/src/lib.rs), the procedural bench/test scenes and the
// AABB helpers.  BVH builders live in rt_bvh.cpp.  No CUDA in this file.
#include <math.h>
#include <string.h>
#include <vector>

#include "../../include/rtb200_host.h"

extern "C" {

// ---- f16 (half 1.8.3 semantics: IEEE binary16, round to nearest even) --------
uint16_t rt_f32_to_f16(float v) {
    uint32_t x;
    memcpy(&x, &v, 4);
    uint32_t sign = (x & 0x80000000u) >> 16;
    uint32_t ax = x & 0x7FFFFFFFu;
    if (ax >= 0x7F800000u) // inf or nan (nan keeps a quiet payload bit)
        return (uint16_t)(sign | 0x7C00u | (ax > 0x7F800000u ? (0x0200u | ((ax >> 13) & 0x3FFu)) : 0u));
    if (ax >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u); // rounds to >= 65520 -> inf
    if (ax < 0x33000001u) return (uint16_t)sign;               // <= 2^-25 -> zero (ties to even)
    if (ax < 0x38800000u) {                                    // subnormal half
        uint32_t mant = (ax & 0x7FFFFFu) | 0x800000u;
        int shift = 126 - (int)(ax >> 23); // 14..24
        uint32_t q = mant >> shift, rem = mant & ((1u << shift) - 1u), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (q & 1u))) q++;
        return (uint16_t)(sign | q);
    }
    uint32_t q = (ax - 0x38000000u) >> 13, rem = ax & 0x1FFFu;
    if (rem > 0x1000u || (rem == 0x1000u && (q & 1u))) q++;
    return (uint16_t)(sign | q);
}

float rt_f16_to_f32(uint16_t bits) {
    uint32_t sign = ((uint32_t)bits & 0x8000u) << 16, exp = (bits >> 10) & 0x1Fu, man = bits & 0x3FFu, out;
    if (exp == 0x1Fu) out = sign | 0x7F800000u | (man << 13);
    else if (exp) out = sign | ((exp + 112u) << 23) | (man << 13);
    else if (!man) out = sign;
    else { // subnormal: value = man * 2^-24, exact in f32
        float f = (float)man * 5.9604644775390625e-08f;
        memcpy(&out, &f, 4);
        out |= sign;
    }
    float r;
    memcpy(&r, &out, 4);
    return r;
}

void rt_camera_default(RtCamera* c) {
    const float p[3] = {0.0f, 0.0f, 5.0f}, d[3] = {0.0f, 0.0f, -1.0f}, u[3] = {0.0f, 1.0f, 0.0f};
    memcpy(c->position, p, 12); memcpy(c->direction, d, 12); memcpy(c->up, u, 12);
    c->fov = 45.0f;
}

void rt_material_new(RtMaterial* m, const float albedo[3], float metallic, float roughness, const float emission[3],
                     float ior, float transmission) {
    memset(m, 0, sizeof *m);
    memcpy(m->albedo, albedo, 12);
    m->metallic_roughness_f16 = (uint32_t)rt_f32_to_f16(metallic) | ((uint32_t)rt_f32_to_f16(roughness) << 16);
    memcpy(m->emission, emission, 12);
    m->ior_transmission_f16 = (uint32_t)rt_f32_to_f16(ior) | ((uint32_t)rt_f32_to_f16(transmission) << 16);
    m->specular_factor = 1.0f;
    m->specular_color[0] = m->specular_color[1] = m->specular_color[2] = 1.0f;
    m->attenuation_distance = INFINITY;
    m->attenuation_color[0] = m->attenuation_color[1] = m->attenuation_color[2] = 1.0f;
    m->thickness_factor = 0.0f;
    memcpy(m->diffuse_factor, albedo, 12);
    m->glossiness_factor = 1.0f - roughness;
    m->material_type = 0;
    for (int i = 0; i < 8; i++) m->texture_indices[i] = 0xFFFFFFFFu;
}
static const float kZero3[3] = {0.0f, 0.0f, 0.0f};
void rt_material_diffuse(RtMaterial* m, const float a[3]) { rt_material_new(m, a, 0.0f, 1.0f, kZero3, 1.5f, 0.0f); }
void rt_material_metallic(RtMaterial* m, const float a[3], float r) { rt_material_new(m, a, 1.0f, r, kZero3, 1.5f, 0.0f); }
void rt_material_glass(RtMaterial* m, const float a[3], float ior, float tr) { rt_material_new(m, a, 0.0f, 0.0f, kZero3, ior, tr); }
void rt_material_emissive(RtMaterial* m, const float a[3], const float e[3]) { rt_material_new(m, a, 0.0f, 1.0f, e, 1.5f, 0.0f); }

static void light_fill(RtLight* l, const float pos[3], uint32_t type, const float color[3], float intensity,
                       const float dir[3], float range, float inner, float outer) {
    memcpy(l->position, pos, 12);
    l->light_type = type;
    memcpy(l->color, color, 12);
    l->intensity = intensity;
    memcpy(l->direction, dir, 12);
    l->range_packed = (uint32_t)rt_f32_to_f16(range);
    l->cone_angles_packed = (uint32_t)rt_f32_to_f16(inner) | ((uint32_t)rt_f32_to_f16(outer) << 16);
}
void rt_light_directional(RtLight* l, const float d[3], const float c[3], float i) { light_fill(l, kZero3, 0, c, i, d, INFINITY, 0.0f, 0.0f); }
void rt_light_point(RtLight* l, const float p[3], const float c[3], float i, float range) { light_fill(l, p, 1, c, i, kZero3, range, 0.0f, 0.0f); }
void rt_light_spot(RtLight* l, const float p[3], const float d[3], const float c[3], float i, float range, float inner,
                   float outer) { light_fill(l, p, 2, c, i, d, range, inner, outer); }

static void pc_fill(RtPushConstants* pc, const float res[2], const RtCamera* cam, uint32_t ntri, uint32_t nmat,
                    const uint32_t toff[2], const uint32_t tsize[2], const uint32_t ttiles[2], uint32_t tpb,
                    const RtSceneMetadataOffsets* off, uint32_t flags, uint32_t seed) {
    pc->resolution[0] = res[0]; pc->resolution[1] = res[1];
    pc->camera = *cam;
    pc->triangle_count = ntri; pc->material_count = nmat;
    pc->tile_offset[0] = toff[0]; pc->tile_offset[1] = toff[1];
    pc->tile_size_packed = rt_pack_tile_size(tsize[0], tsize[1]);
    pc->total_tiles[0] = ttiles[0]; pc->total_tiles[1] = ttiles[1];
    pc->triangles_per_buffer = tpb;
    pc->metadata_offsets = *off;
    pc->packed_flags = flags;
    pc->frame_seed = seed;
}
void rt_push_constants_new(RtPushConstants* pc, const float res[2], const RtCamera* cam, uint32_t ntri, uint32_t nmat,
                           const uint32_t toff[2], const uint32_t tsize[2], const uint32_t ttiles[2], uint32_t tpb,
                           const RtSceneMetadataOffsets* off, uint32_t channel) {
    pc_fill(pc, res, cam, ntri, nmat, toff, tsize, ttiles, tpb, off, rt_pack_flags(channel, 0, 4, 0), 0);
}
void rt_push_constants_new_wavefront(RtPushConstants* pc, const float res[2], const RtCamera* cam, uint32_t ntri,
                                     uint32_t nmat, const uint32_t toff[2], const uint32_t tsize[2],
                                     const uint32_t ttiles[2], uint32_t tpb, const RtSceneMetadataOffsets* off,
                                     uint32_t channel, uint32_t cur, uint32_t maxd, uint32_t seed) {
    pc_fill(pc, res, cam, ntri, nmat, toff, tsize, ttiles, tpb, off, rt_pack_flags(channel, cur, maxd, 1), seed);
}

// TriangleLegacy::to_indexed (shared/src/lib.rs:713-748): bit-exact position dedup
// (Vertex derives PartialEq over f32s), first occurrence keeps its index.
static uint32_t find_or_add(std::vector<RtVertex>& vs, const float p[3]) {
    for (size_t i = 0; i < vs.size(); i++)
        if (vs[i].position[0] == p[0] && vs[i].position[1] == p[1] && vs[i].position[2] == p[2]) return (uint32_t)i;
    RtVertex v;
    memcpy(v.position, p, 12);
    vs.push_back(v);
    return (uint32_t)vs.size() - 1;
}

void rt_default_scene(RtSphere spheres[6], RtTriangle triangles[2], RtVertex vertices[6], RtMaterial materials[4],
                      RtLight lights[1]) {
    const float red[3] = {0.8f, 0.3f, 0.3f}, yellow[3] = {0.8f, 0.8f, 0.2f}, blue[3] = {0.2f, 0.3f, 0.8f},
                white[3] = {1.0f, 1.0f, 1.0f}, glow[3] = {0.5f, 0.5f, 1.0f};
    rt_material_diffuse(&materials[0], red);
    rt_material_metallic(&materials[1], yellow, 0.1f);
    rt_material_glass(&materials[2], blue, 1.5f, 0.9f);
    rt_material_emissive(&materials[3], white, glow);
    const float sc[6][3] = {{0, 0, -1}, {-1, 0, -1}, {1, 0, -1}, {2, 0, -3}, {-2, 0, -4}, {-1, 2, -5}};
    const uint32_t sm[6] = {0, 1, 2, 2, 1, 3};
    for (int i = 0; i < 6; i++) { memcpy(spheres[i].center, sc[i], 12); spheres[i].radius = 0.5f; spheres[i].material_id = sm[i]; }
    const float tv[2][3][3] = {{{0.0f, 1.0f, -2.0f}, {-0.5f, 0.0f, -2.0f}, {0.5f, 0.0f, -2.0f}},
                               {{1.5f, 0.5f, -3.0f}, {1.0f, -0.5f, -3.0f}, {2.0f, -0.5f, -3.0f}}};
    const uint32_t tm[2] = {0, 1};
    std::vector<RtVertex> vs;
    for (int t = 0; t < 2; t++) {
        triangles[t].v0_index = find_or_add(vs, tv[t][0]);
        triangles[t].v1_index = find_or_add(vs, tv[t][1]);
        triangles[t].v2_index = find_or_add(vs, tv[t][2]);
        triangles[t].material_id = tm[t];
    }
    for (size_t i = 0; i < vs.size() && i < 6; i++) vertices[i] = vs[i];
    const float lp[3] = {5.0f, 7.0f, 4.0f};
    rt_light_point(&lights[0], lp, white, 1.0f, INFINITY);
}

// ---- AABB helpers (shared/src/lib.rs:752-801; f32::min/max == fminf/fmaxf) ----
void rt_aabb_empty(RtAabb* a) {
    a->min[0] = a->min[1] = a->min[2] = INFINITY;
    a->max[0] = a->max[1] = a->max[2] = -INFINITY;
    a->_padding0 = a->_padding1 = 0.0f;
}
void rt_aabb_union(const RtAabb* a, const RtAabb* b, RtAabb* out) {
    RtAabb r;
    for (int i = 0; i < 3; i++) { r.min[i] = fminf(a->min[i], b->min[i]); r.max[i] = fmaxf(a->max[i], b->max[i]); }
    r._padding0 = r._padding1 = 0.0f;
    *out = r;
}
void rt_aabb_center(const RtAabb* a, float out[3]) {
    for (int i = 0; i < 3; i++) out[i] = (a->min[i] + a->max[i]) * 0.5f;
}
float rt_aabb_surface_area(const RtAabb* a) {
    float dx = a->max[0] - a->min[0], dy = a->max[1] - a->min[1], dz = a->max[2] - a->min[2];
    return 2.0f * (dx * dy + dy * dz + dz * dx);
}
void rt_triangle_aabb(const RtTriangle* t, const RtVertex* v, RtAabb* out) {
    const float *a = v[t->v0_index].position, *b = v[t->v1_index].position, *c = v[t->v2_index].position;
    for (int i = 0; i < 3; i++) {
        out->min[i] = fminf(fminf(a[i], b[i]), c[i]);
        out->max[i] = fmaxf(fmaxf(a[i], b[i]), c[i]);
    }
    out->_padding0 = out->_padding1 = 0.0f;
}

} // extern "C"
