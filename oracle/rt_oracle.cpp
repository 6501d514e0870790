/* rt_oracle.cpp — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference's per-pixel compute shader `main_cs`
 * (every .rs file under /root/reference/shader/src), consuming the same bound buffers and the
 * same 128-byte push constants the reference dispatch supplies
 * (src/compute.rs:212-251, src/buffers.rs:157-271).  Each function cites the
 * reference lines it follows.  It is the parity checker for the CUDA path and
 * the timed `cpu_baseline`; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.  The product path never
 * does.
 *
 * PARITY UNPINNED: the reference's shader crate has zero tests and no golden
 * vectors, and cannot be compiled here (no rustc; needs rust-gpu nightly +
 * crates.io).  The restatement is pinned only by the analytic known-answer
 * tests in tests/ and by the reference's 17 structural unit tests restated
 * there.
 *
 * Build: g++ -O2 -ffp-contract=off -fno-fast-math -shared -fPIC (oracle/Makefile)
 */
#include <atomic>
#include <thread>
#include <vector>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#include "../include/rt_shared.h"
#include "rt_oracle_math.h"
#include "rt_oracle.h"

namespace {

/* The bound resources of one dispatch (shader/src/lib.rs:26-37). */
struct Bind {
    const uint32_t* meta;   /* binding 1: scene_metadata */
    uint64_t meta_len;      /* in u32 */
    const RtTriangle* tri[3]; /* bindings 2-4 */
    uint64_t tri_len[3];
    const RtMaterial* materials; /* binding 5 */
    uint64_t materials_len;
    const RtPushConstants* pc;
};

struct Counters {
    uint64_t aabb_tests = 0, tri_tests = 0, sphere_tests = 0, rays = 0;
    uint64_t depth_rays[8] = {0, 0, 0, 0, 0, 0, 0, 0}; /* path mode: segments per bounce depth (>= 7 share the last) */
};
static uint64_t g_last_depth_rays[8]; /* of the last orc_render_path call (orc_last_depth_rays) */

struct Ray {
    V3 origin, direction;
};

/* shader/src/intersection.rs:10-38 plus the oracle's extra primitive id. */
struct Hit {
    bool hit;
    float t;
    V3 point, normal;
    uint32_t material_id;
    uint32_t kind; /* ORC_MISS / ORC_SPHERE / ORC_TRIANGLE */
    uint32_t prim; /* sphere index or triangle index */
};
static inline Hit miss() {
    return Hit{false, FLT_MAX, v3(0, 0, 0), v3(0, 0, 0), 0xFFFFFFFFu, ORC_MISS, 0xFFFFFFFFu};
}

/* Robust-buffer-access style read: the reference would read out of bounds
 * (undefined); the oracle returns 0 so it never faults.  Valid uploads never
 * get here. */
static inline uint32_t meta_at(const Bind& b, uint64_t i) { return i < b.meta_len ? b.meta[i] : 0u; }
static inline float meta_f(const Bind& b, uint64_t i) { return u2f(meta_at(b, i)); }

/* shader/src/scene_access.rs:33-57 */
static inline V3 sphere_center(const Bind& b, uint32_t i) {
    uint64_t o = (uint64_t)b.pc->metadata_offsets.spheres_offset + (uint64_t)i * 5;
    return v3(meta_f(b, o), meta_f(b, o + 1), meta_f(b, o + 2));
}
static inline float sphere_radius(const Bind& b, uint32_t i) {
    return meta_f(b, (uint64_t)b.pc->metadata_offsets.spheres_offset + (uint64_t)i * 5 + 3);
}
static inline uint32_t sphere_material(const Bind& b, uint32_t i) {
    return meta_at(b, (uint64_t)b.pc->metadata_offsets.spheres_offset + (uint64_t)i * 5 + 4);
}
/* shader/src/scene_access.rs:110-159 (node stride 12 u32) */
static inline uint64_t node_off(const Bind& b, uint32_t n) {
    return (uint64_t)b.pc->metadata_offsets.bvh_nodes_offset + (uint64_t)n * 12;
}
/* shader/src/scene_access.rs:169-177 (vertex stride 3 u32) */
static inline V3 vertex_pos(const Bind& b, uint32_t v) {
    uint64_t o = (uint64_t)b.pc->metadata_offsets.vertices_offset + (uint64_t)v * 3;
    return v3(meta_f(b, o), meta_f(b, o + 1), meta_f(b, o + 2));
}

/* shader/src/ray.rs:14-19 */
static inline Ray ray_new(V3 o, V3 d) { return Ray{o, normalize(d)}; }
static inline V3 ray_at(const Ray& r, float t) { return r.origin + r.direction * t; }

/* shader/src/ray.rs:22-53 — legacy camera ray; normalises twice. */
static Ray ray_from_screen(uint32_t px, uint32_t py, const RtPushConstants& pc) {
    uint32_t width = (uint32_t)pc.resolution[0];
    uint32_t height = (uint32_t)pc.resolution[1];
    float u = ((float)px + 0.5f) / (float)width;
    float v = ((float)py + 0.5f) / (float)height;
    float aspect = (float)width / (float)height;
    float fov_scale = tanf(pc.camera.fov * 0.5f * 3.14159265358979323846f / 180.0f);
    float cx = (u * 2.0f - 1.0f) * aspect * fov_scale;
    float cy = (1.0f - v * 2.0f) * fov_scale;
    V3 forward = v3a(pc.camera.direction), up = v3a(pc.camera.up);
    V3 right = cross(forward, up);
    V3 true_up = cross(right, forward);
    V3 dir = forward + right * cx + true_up * cy;
    V3 dirn = normalize(dir);
    return ray_new(v3a(pc.camera.position), dirn);
}

/* shader/src/wavefront.rs:75-112 — wavefront camera ray; divides by the f32
 * resolution directly and normalises once. */
static Ray ray_wavefront_camera(uint32_t px, uint32_t py, const RtPushConstants& pc) {
    float u = ((float)px + 0.5f) / pc.resolution[0];
    float v = ((float)py + 0.5f) / pc.resolution[1];
    float aspect = pc.resolution[0] / pc.resolution[1];
    float fov_scale = tanf(pc.camera.fov * 0.5f * 3.14159265358979323846f / 180.0f);
    float cx = (u * 2.0f - 1.0f) * aspect * fov_scale;
    float cy = (1.0f - v * 2.0f) * fov_scale;
    V3 forward = v3a(pc.camera.direction), up = v3a(pc.camera.up);
    V3 right = cross(forward, up);
    V3 true_up = cross(right, forward);
    V3 dir = forward + right * cx + true_up * cy;
    return Ray{v3a(pc.camera.position), normalize(dir)};
}

/* shader/src/intersection.rs:52-87 */
static Hit test_sphere(const Ray& ray, const Bind& b, uint32_t i, float max_t) {
    V3 c = sphere_center(b, i);
    float r = sphere_radius(b, i);
    V3 oc = ray.origin - c;
    float a = dot(ray.direction, ray.direction);
    float bq = 2.0f * dot(oc, ray.direction);
    float cq = dot(oc, oc) - r * r;
    float disc = bq * bq - 4.0f * a * cq;
    if (disc < 0.0f) return miss();
    float sq = sqrtf(disc);
    float t1 = (-bq - sq) / (2.0f * a);
    float t2 = (-bq + sq) / (2.0f * a);
    float t = t1 > RT_MIN_RAY_DISTANCE ? t1 : t2;
    if (t > RT_MIN_RAY_DISTANCE && t < max_t) {
        V3 p = ray_at(ray, t);
        V3 n = normalize(p - c);
        return Hit{true, t, p, n, sphere_material(b, i), ORC_SPHERE, i};
    }
    return miss();
}

/* shader/src/intersection.rs:91-138 — Möller–Trumbore, two-sided. */
static Hit test_triangle(const Ray& ray, V3 v0, V3 v1, V3 v2, uint32_t material_id, float max_t,
                         uint32_t prim) {
    V3 e1 = v1 - v0, e2 = v2 - v0;
    V3 h = cross(ray.direction, e2);
    float a = dot(e1, h);
    if (fabsf(a) < RT_MIN_RAY_DISTANCE) return miss();
    float f = 1.0f / a;
    V3 s = ray.origin - v0;
    float u = f * dot(s, h);
    if (u < 0.0f || u > 1.0f) return miss();
    V3 q = cross(s, e1);
    float v = f * dot(ray.direction, q);
    if (v < 0.0f || u + v > 1.0f) return miss();
    float t = f * dot(e2, q);
    if (t > RT_MIN_RAY_DISTANCE && t < max_t) {
        V3 p = ray_at(ray, t);
        V3 n = normalize(cross(e1, e2));
        return Hit{true, t, p, n, material_id, ORC_TRIANGLE, prim};
    }
    return miss();
}

/* shader/src/intersection.rs:151-164 — slab test; no t_max pruning. */
static inline bool ray_aabb(V3 o, V3 d, V3 bmin, V3 bmax) {
    V3 inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    V3 t1 = (bmin - o) * inv;
    V3 t2 = (bmax - o) * inv;
    V3 tmin = vmin(t1, t2), tmax = vmax(t1, t2);
    float tmin_max = fmaxf(fmaxf(tmin.x, tmin.y), tmin.z);
    float tmax_min = fminf(fminf(tmax.x, tmax.y), tmax.z);
    return tmax_min >= 0.0f && tmin_max <= tmax_min;
}

/* shader/src/triangle_access.rs:18-61 */
static inline bool fetch_triangle(const Bind& b, uint32_t index, V3* v0, V3* v1, V3* v2, uint32_t* mat) {
    uint32_t tpb = b.pc->triangles_per_buffer;
    if (tpb == 0) return false; /* reference would trap on /0; uploads never send 0 */
    uint32_t bi = index / tpb;
    uint64_t li = index % tpb;
    if (bi > 2 || li >= b.tri_len[bi]) return false;
    RtTriangle t = b.tri[bi][li];
    *v0 = vertex_pos(b, t.v0_index);
    *v1 = vertex_pos(b, t.v1_index);
    *v2 = vertex_pos(b, t.v2_index);
    *mat = t.material_id;
    return true;
}

/* shader/src/bvh.rs:91-133 */
static Hit test_leaf(const Ray& ray, const Bind& b, uint32_t node, float max_t, Counters& c) {
    uint64_t no = node_off(b, node);
    uint32_t start = meta_at(b, no + 10), count = meta_at(b, no + 11);
    Hit result = miss();
    float closest = max_t;
    for (uint32_t i = 0; i < count; i++) {
        if (start + i >= b.pc->metadata_offsets.triangle_indices_count) break;
        uint32_t ti = meta_at(b, (uint64_t)b.pc->metadata_offsets.triangle_indices_offset + start + i);
        V3 v0, v1, v2;
        uint32_t mat;
        if (fetch_triangle(b, ti, &v0, &v1, &v2, &mat)) {
            c.tri_tests++;
            Hit h = test_triangle(ray, v0, v1, v2, mat, closest, ti);
            if (h.hit) { closest = h.t; result = h; }
        }
    }
    return result;
}

/* shader/src/bvh.rs:18-88 — literal stack DFS incl. the `stack_ptr < 63` quirk. */
static Hit traverse(const Ray& ray, const Bind& b, float max_t, Counters& c) {
    uint32_t node_count = b.pc->metadata_offsets.bvh_nodes_count;
    if (node_count == 0) return miss();
    Hit result = miss();
    float closest = max_t;
    uint32_t stack[64];
    for (int i = 0; i < 64; i++) stack[i] = 0xFFFFFFFFu;
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0) {
        uint32_t n = stack[--sp];
        if (n == 0xFFFFFFFFu || n >= node_count) continue;
        uint64_t no = node_off(b, n);
        V3 bmin = v3(meta_f(b, no), meta_f(b, no + 1), meta_f(b, no + 2));
        V3 bmax = v3(meta_f(b, no + 4), meta_f(b, no + 5), meta_f(b, no + 6));
        c.aabb_tests++;
        if (!ray_aabb(ray.origin, ray.direction, bmin, bmax)) continue;
        uint32_t left = meta_at(b, no + 8), right = meta_at(b, no + 9);
        if (left == 0xFFFFFFFFu) {
            Hit h = test_leaf(ray, b, n, closest, c);
            if (h.hit) { closest = h.t; result = h; }
        } else {
            if (right != 0xFFFFFFFFu && sp < 63) stack[sp++] = right;
            if (left != 0xFFFFFFFFu && sp < 63) stack[sp++] = left;
        }
    }
    return result;
}

/* NOT in the reference: the CUDA kernel's declared ORDERED policy (near-first,
 * conservatively t-pruned, equal-t ties to the smaller triangle_indices position),
 * restated on the reference node layout.  Used (a) to count the box / triangle
 * tests of that policy for the roofline's bytes-per-ray and (b) to check on the
 * CPU that the policy returns exactly the literal policy's hits.  Only valid on
 * BVHs the product's upload validation accepts (tree, depth < 60, leaves in
 * increasing triangle_start order). */
static const float PRUNE_SCALE = 1.001f, PRUNE_ABS = 1.0e-4f;
static inline bool slab_entry(V3 o, V3 inv, V3 bmin, V3 bmax, float* entry) {
    V3 t1 = (bmin - o) * inv, t2 = (bmax - o) * inv;
    V3 tmin = vmin(t1, t2), tmax = vmax(t1, t2);
    float tmin_max = fmaxf(fmaxf(tmin.x, tmin.y), tmin.z);
    float tmax_min = fminf(fminf(tmax.x, tmax.y), tmax.z);
    *entry = tmin_max;
    return tmax_min >= 0.0f && tmin_max <= tmax_min;
}
static Hit traverse_ordered(const Ray& ray, const Bind& b, float max_t, Counters& c) {
    uint32_t node_count = b.pc->metadata_offsets.bvh_nodes_count;
    uint32_t n_idx = b.pc->metadata_offsets.triangle_indices_count;
    Hit result = miss();
    if (node_count == 0) return result;
    float closest = max_t;
    uint32_t best_pos = 0;
    V3 inv = v3(1.0f / ray.direction.x, 1.0f / ray.direction.y, 1.0f / ray.direction.z);
    auto box_of = [&](uint32_t n, V3* mn, V3* mx) {
        uint64_t no = node_off(b, n);
        *mn = v3(meta_f(b, no), meta_f(b, no + 1), meta_f(b, no + 2));
        *mx = v3(meta_f(b, no + 4), meta_f(b, no + 5), meta_f(b, no + 6));
    };
    V3 mn, mx;
    float e;
    box_of(0, &mn, &mx);
    c.aabb_tests++;
    if (!slab_entry(ray.origin, inv, mn, mx, &e)) return result;
    uint32_t stack[64];
    int sp = 0;
    uint32_t cur = 0;
    for (;;) {
        uint64_t no = node_off(b, cur);
        uint32_t left = meta_at(b, no + 8), right = meta_at(b, no + 9);
        if (left == 0xFFFFFFFFu) {
            uint32_t start = meta_at(b, no + 10), count = meta_at(b, no + 11);
            for (uint32_t i = 0; i < count && start + i < n_idx; i++) {
                uint32_t pos = start + i;
                uint32_t ti = meta_at(b, (uint64_t)b.pc->metadata_offsets.triangle_indices_offset + pos);
                V3 v0, v1, v2;
                uint32_t mat;
                if (!fetch_triangle(b, ti, &v0, &v1, &v2, &mat)) continue;
                c.tri_tests++;
                Hit h = test_triangle(ray, v0, v1, v2, mat, FLT_MAX, ti); /* any t > eps */
                if (h.hit && h.t < max_t &&
                    (h.t < closest || (h.t == closest && result.hit && pos < best_pos))) {
                    closest = h.t; result = h; best_pos = pos;
                }
            }
        } else {
            bool lv = left < node_count, rv = right != 0xFFFFFFFFu && right < node_count;
            float el = 0, er = 0;
            bool hl = false, hr = false;
            float limit = closest * PRUNE_SCALE + PRUNE_ABS;
            if (lv) { box_of(left, &mn, &mx); c.aabb_tests++; hl = slab_entry(ray.origin, inv, mn, mx, &el) && el <= limit; }
            if (rv) { box_of(right, &mn, &mx); c.aabb_tests++; hr = slab_entry(ray.origin, inv, mn, mx, &er) && er <= limit; }
            if (hl && hr) {
                bool right_first = er < el;
                stack[sp++] = right_first ? left : right;
                cur = right_first ? right : left;
                continue;
            } else if (hl) { cur = left; continue; }
            else if (hr) { cur = right; continue; }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    return result;
}

/* shader/src/lib.rs:272-296 */
static Hit brute_force(const Ray& ray, const Bind& b, float max_t, Counters& c) {
    Hit result = miss();
    float closest = max_t;
    for (uint32_t i = 0; i < b.pc->triangle_count; i++) {
        V3 v0, v1, v2;
        uint32_t mat;
        if (fetch_triangle(b, i, &v0, &v1, &v2, &mat)) {
            c.tri_tests++;
            Hit h = test_triangle(ray, v0, v1, v2, mat, closest, i);
            if (h.hit) { closest = h.t; result = h; }
        }
    }
    return result;
}

/* shader/src/lib.rs:174-249 (twin: wavefront.rs:214-289) */
static thread_local int g_policy = 0; /* 0 literal (reference), 1 ordered (kernel's policy) */
static Hit find_closest(const Ray& ray, const Bind& b, Counters& c) {
    float closest = FLT_MAX - 2.0f;
    /* lib.rs:252-269 */
    Hit sphere = miss();
    {
        float cs = closest;
        for (uint32_t i = 0; i < b.pc->metadata_offsets.spheres_count; i++) {
            c.sphere_tests++;
            Hit h = test_sphere(ray, b, i, cs);
            if (h.hit) { cs = h.t; sphere = h; }
        }
    }
    if (sphere.hit) closest = sphere.t;
    Hit tri = b.pc->metadata_offsets.bvh_nodes_count > 0 ? (g_policy == 1 ? traverse_ordered(ray, b, closest, c) : traverse(ray, b, closest, c))
                                                         : brute_force(ray, b, closest, c);
    /* branchless select, lib.rs:216-248 */
    uint32_t sphere_closer = sphere.t < tri.t ? 1u : 0u;
    uint32_t both = (tri.hit && sphere.hit) ? 1u : 0u;
    uint32_t sphere_only = (sphere.hit && !tri.hit) ? 1u : 0u;
    uint32_t tri_only = (tri.hit && !sphere.hit) ? 1u : 0u;
    uint32_t use_sphere = both * sphere_closer + sphere_only;
    uint32_t use_tri = both * (1u - sphere_closer) + tri_only;
    if (use_sphere) return sphere;
    if (use_tri) return tri;
    return miss();
}

/* shader/src/material.rs:6-84 */
struct Mat {
    V3 albedo, emission;
    float metallic, roughness, ior, transmission;
};
static inline Mat mat_eval(const RtMaterial& m) {
    Mat r;
    r.albedo = v3a(m.albedo);
    r.emission = v3a(m.emission);
    r.metallic = f16_bits_to_f32(m.metallic_roughness_f16 & 0xFFFFu);
    r.roughness = f16_bits_to_f32(m.metallic_roughness_f16 >> 16);
    r.ior = f16_bits_to_f32(m.ior_transmission_f16 & 0xFFFFu);
    r.transmission = f16_bits_to_f32(m.ior_transmission_f16 >> 16);
    return r;
}
/* material.rs:42-58 */
static inline float ior_for_channel(const Mat& m, uint32_t ch) {
    static const float table[4] = {-0.018f, 0.0f, 0.035f, 0.0f};
    return m.ior + table[ch < 3 ? ch : 3];
}
/* material.rs:76-83 */
static inline V3 evaluate_brdf(const Mat& m, float I) {
    V3 diffuse = m.albedo / 3.14159265358979323846f;
    float is_m = (m.metallic > 0.5f) ? 1.0f : 0.0f;
    V3 metallic_c = m.albedo * I * 0.5f;
    V3 dielectric_c = diffuse * I;
    return metallic_c * is_m + dielectric_c * (1.0f - is_m);
}

/* shader/src/lighting.rs:50-139 (+ scene_access.rs:60-107, light stride 13 u32) */
static V3 light_contribution(const Bind& b, const Hit& h, const Mat& m, uint32_t li) {
    float index_valid = (li < b.pc->metadata_offsets.lights_count) ? 1.0f : 0.0f;
    uint64_t o = (uint64_t)b.pc->metadata_offsets.lights_offset + (uint64_t)li * 13;
    V3 lpos = v3(meta_f(b, o), meta_f(b, o + 1), meta_f(b, o + 2));
    uint32_t ltype = meta_at(b, o + 3);
    V3 lcol = v3(meta_f(b, o + 4), meta_f(b, o + 5), meta_f(b, o + 6));
    float lint = meta_f(b, o + 7);
    V3 ldir = v3(meta_f(b, o + 8), meta_f(b, o + 9), meta_f(b, o + 10));
    /* lighting.rs:97-110 */
    V3 dir_light_dir = -normalize(ldir);
    float dir_intensity = fmaxf(dot(h.normal, dir_light_dir), 0.0f) * lint;
    /* lighting.rs:113-139 */
    V3 to_light = lpos - h.point;
    float distance = length(to_light);
    V3 pdir = normalize(to_light);
    float att32 = 1.0f / (1.0f + distance * distance * 0.01f);
    float att = f16_bits_to_f32(f32_to_f16_bits(att32));
    float point_intensity = fmaxf(dot(h.normal, pdir), 0.0f) * lint * att;
    float spot_factor = fmaxf(dot(-normalize(ldir), pdir), 0.0f);
    float spot_intensity = point_intensity * spot_factor;
    /* lighting.rs:79-93 */
    float is_dir = (ltype == 0) ? 1.0f : 0.0f, is_point = (ltype == 1) ? 1.0f : 0.0f,
          is_spot = (ltype == 2) ? 1.0f : 0.0f;
    float I = dir_intensity * is_dir + point_intensity * is_point + spot_intensity * is_spot;
    V3 contrib = evaluate_brdf(m, I);
    float valid = ((I > 0.0f) ? 1.0f : 0.0f) * index_valid;
    return contrib * lcol * valid;
}

/* shader/src/lighting.rs:20-47 */
static V3 calculate_lighting(const Bind& b, const Hit& h, const Mat& m) {
    V3 total = v3(0, 0, 0);
    total = total + m.albedo * 0.1f;
    for (uint32_t i = 0; i < b.pc->metadata_offsets.lights_count; i++)
        total = total + light_contribution(b, h, m, i);
    return total + m.emission;
}

/* shader/src/lib.rs:300-338; wavefront twin wavefront.rs:168-211 multiplies by
 * `throughput` when `tp` is non-null. */
static V3 calculate_shading(const Bind& b, const Hit& h, uint32_t channel, const V3* tp) {
    if ((uint64_t)h.material_id >= b.materials_len) return v3(1.0f, 0.0f, 1.0f);
    Mat m = mat_eval(b.materials[h.material_id]);
    V3 lighting = calculate_lighting(b, h, m);
    float T = fminf(fmaxf(m.transmission, 0.0f), 1.0f);
    V3 out;
    if (T > 0.0f) {
        float wl = ior_for_channel(m, channel);
        float disp = (wl - 1.0f) / (m.ior - 1.0f);
        V3 transmitted = v3(0.2f, 0.2f, 0.3f) * disp;
        out = lighting * (1.0f - T) + transmitted * T;
    } else {
        out = lighting;
    }
    return tp ? out * (*tp) : out;
}

/* shader/src/lib.rs:342-349 */
static inline V3 filter_channel(V3 c, uint32_t ch) {
    switch (ch) {
        case 0: return v3(c.x, 0, 0);
        case 1: return v3(0, c.y, 0);
        case 2: return v3(0, 0, c.z);
        default: return c;
    }
}

/* Rgba8Unorm image store (lib.rs:85-88; src/renderer.rs:258,464): clamp to
 * [0,1], scale by 255, round to nearest; NaN -> 0. */
static inline uint8_t unorm8(float f) {
    if (!(f > 0.0f)) return 0; /* also NaN */
    if (f >= 1.0f) return 255;
    return (uint8_t)rintf(f * 255.0f);
}

/* One invocation of main_cs past the bounds check (lib.rs:44-80): returns the
 * un-filtered `final_color` and the closest-hit record. */
static V3 shade_pixel(const Bind& b, uint32_t px, uint32_t py, Hit* out_hit, Counters& c) {
    const RtPushConstants& pc = *b.pc;
    uint32_t channel = rt_pc_color_channel(&pc);
    if (rt_pc_wavefront_mode(&pc) != 0) {
        /* lib.rs:92-149: only the iteration bounce == current_bounce_depth runs. */
        Ray ray = ray_wavefront_camera(px, py, pc);
        V3 acc = v3(0, 0, 0);
        *out_hit = miss();
        if (rt_pc_current_bounce(&pc) <= rt_pc_max_bounce(&pc)) {
            c.rays++;
            Hit h = find_closest(ray, b, c);
            *out_hit = h;
            V3 tp = v3(1.0f, 1.0f, 1.0f);
            V3 contrib = h.hit ? calculate_shading(b, h, channel, &tp) : v3(0.1f, 0.2f, 0.3f) * tp;
            acc = acc + contrib;
        }
        return acc;
    }
    Ray ray = ray_from_screen(px, py, pc);
    c.rays++;
    Hit h = find_closest(ray, b, c);
    *out_hit = h;
    return h.hit ? calculate_shading(b, h, channel, nullptr) : v3(0, 0, 0);
}

/* ---- EXTENSION (not in the reference; SPEC.md) ------------------------------ */
/* shader/src/wavefront.rs:46-72 */
struct SimpleRng {
    uint32_t seed;
    uint32_t next_u32() { seed = seed * 1664525u + 1013904223u; return seed; }
    float next_f32() { return (float)(next_u32() >> 8) / 16777216.0f; }
    float next_f32_signed() { return next_f32() * 2.0f - 1.0f; }
};

struct PathParams {
    uint32_t spp, first_sample, max_depth, base_seed;
};

/* SPEC.md steps 1-3 for one sample; returns L, counts segments; optionally the primary hit. */
static V3 path_sample(const Bind& b, uint32_t px, uint32_t py, uint32_t s, const PathParams& pp, Counters& c,
                      Hit* primary) {
    const RtPushConstants& pc = *b.pc;
    uint32_t W = (uint32_t)pc.resolution[0];
    SimpleRng rng{pp.base_seed + s + px + py * W};
    float jx = 0.5f, jy = 0.5f;
    if (s != 0) { jx = rng.next_f32(); jy = rng.next_f32(); }
    Ray ray;
    { /* generate_camera_ray with the sub-pixel offset, wavefront.rs:75-112 */
        float u = ((float)px + jx) / pc.resolution[0];
        float v = ((float)py + jy) / pc.resolution[1];
        float aspect = pc.resolution[0] / pc.resolution[1];
        float fov_scale = tanf(pc.camera.fov * 0.5f * 3.14159265358979323846f / 180.0f);
        float cx = (u * 2.0f - 1.0f) * aspect * fov_scale;
        float cy = (1.0f - v * 2.0f) * fov_scale;
        V3 forward = v3a(pc.camera.direction), up = v3a(pc.camera.up);
        V3 right = cross(forward, up);
        V3 true_up = cross(right, forward);
        ray = Ray{v3a(pc.camera.position), normalize(forward + right * cx + true_up * cy)};
    }
    V3 tp = v3(1.0f, 1.0f, 1.0f), L = v3(0, 0, 0);
    uint32_t depth = 0;
    for (;;) {
        c.rays++;
        c.depth_rays[depth < 7 ? depth : 7]++;
        Hit h = find_closest(ray, b, c);
        if (depth == 0 && primary) *primary = h;
        if (!h.hit) { L = L + v3(0.1f, 0.2f, 0.3f) * tp; break; }
        if ((uint64_t)h.material_id >= b.materials_len) { L = L + v3(1.0f, 0.0f, 1.0f) * tp; break; }
        V3 C = v3(calculate_shading(b, h, 0, nullptr).x, calculate_shading(b, h, 1, nullptr).y,
                  calculate_shading(b, h, 2, nullptr).z);
        L = L + C * tp;
        if (depth == pp.max_depth) break;
        if (depth >= 3) {
            float p = fminf(fmaxf(fmaxf(fmaxf(tp.x, tp.y), tp.z), 0.05f), 1.0f);
            float r = rng.next_f32();
            if (r > p) break;
            tp = v3(tp.x / p, tp.y / p, tp.z / p);
        }
        float r0 = rng.next_f32();
        V3 u = v3(0, 0, 0);
        for (int k = 0; k < 8; k++) {
            float a = rng.next_f32_signed(), bb = rng.next_f32_signed(), cc = rng.next_f32_signed();
            float q = a * a + bb * bb + cc * cc;
            if (q > 1e-6f && q <= 1.0f) { u = v3(a, bb, cc) * (1.0f / sqrtf(q)); break; }
        }
        Mat m = mat_eval(b.materials[h.material_id]);
        V3 d = ray.direction, n = h.normal;
        float cosi = dot(n, d);
        V3 nf = cosi < 0.0f ? n : -n;
        V3 w = nf + u;
        if (dot(w, w) < 1e-8f) w = nf;
        w = normalize(w);
        float T = fminf(fmaxf(m.transmission, 0.0f), 1.0f);
        V3 dir;
        float side = 1.0f;
        if (T > 0.0f && r0 < T) {
            float eta = cosi < 0.0f ? 1.0f / m.ior : m.ior;
            float c1 = -dot(d, nf);
            float k = 1.0f - eta * eta * (1.0f - c1 * c1);
            if (k < 0.0f) dir = d - nf * (2.0f * dot(d, nf));
            else { dir = d * eta + nf * (eta * c1 - sqrtf(k)); side = -1.0f; }
            dir = normalize(dir);
        } else if (m.metallic > 0.5f) {
            V3 mr = d - nf * (2.0f * dot(d, nf));
            dir = mr * (1.0f - m.roughness) + w * m.roughness;
            if (dot(dir, dir) < 1e-8f || dot(dir, nf) <= 0.0f) dir = w;
            dir = normalize(dir);
        } else {
            dir = w;
        }
        tp = tp * m.albedo;
        ray = Ray{h.point + nf * (side * 1e-3f), dir};
        depth++;
    }
    return L;
}

/* Work items are 64-pixel spans of a row (not whole rows): a band of 2 rows still
 * feeds 60 items to the pool at 1080p, so a short cpu_baseline sample uses every
 * host thread it is given. */
static const uint32_t SPAN = 64;
template <class F>
static void parallel_spans(uint32_t rows, uint32_t width, int threads, F fn) {
    const uint32_t per_row = (width + SPAN - 1) / SPAN;
    const uint64_t items = (uint64_t)rows * per_row;
    auto run = [&](uint64_t it, int t) {
        uint32_t r = (uint32_t)(it / per_row), x0 = (uint32_t)(it % per_row) * SPAN;
        fn(r, x0, x0 + SPAN < width ? x0 + SPAN : width, t);
    };
    if (threads <= 1 || items <= 1) {
        for (uint64_t it = 0; it < items; it++) run(it, 0);
        return;
    }
    std::atomic<uint64_t> next(0);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++)
        pool.emplace_back([&, t]() {
            for (;;) {
                uint64_t it = next.fetch_add(1);
                if (it >= items) break;
                run(it, t);
            }
        });
    for (auto& th : pool) th.join();
}

static Bind make_bind(const OrcBindings* ob, const RtPushConstants* pc) {
    Bind b;
    b.meta = ob->scene_metadata;
    b.meta_len = ob->scene_metadata_len;
    for (int i = 0; i < 3; i++) { b.tri[i] = ob->triangles[i]; b.tri_len[i] = ob->triangles_len[i]; }
    b.materials = ob->materials;
    b.materials_len = ob->materials_len;
    b.pc = pc;
    return b;
}

} /* namespace */

extern "C" {

int orc_hardware_threads(void) {
    unsigned n = std::thread::hardware_concurrency();
    return n ? (int)n : 1;
}

/* One reference dispatch: `dispatch_workgroups(ceil(w/16), ceil(h/16), 1)` of
 * main_cs with these push constants (src/compute.rs:245-250).  Every in-bounds
 * invocation (lib.rs:152-163) writes one rgba8 texel of `image` (W*H*4 bytes,
 * the channel's storage texture).  Optional outputs: `radiance` (W*H*3 fp32,
 * the un-filtered final_color, only component `channel` is written for
 * channel<3), `hits` (W*H records). */
int orc_dispatch(const OrcBindings* ob, const RtPushConstants* pc, uint8_t* image, float* radiance,
                 OrcHit* hits, int threads, OrcCounters* counters) {
    return orc_dispatch_policy(ob, pc, image, radiance, hits, threads, counters, 0);
}

/* policy 0: the reference's literal traversal; 1: the kernel's ORDERED policy. */
int orc_dispatch_policy(const OrcBindings* ob, const RtPushConstants* pc, uint8_t* image, float* radiance,
                        OrcHit* hits, int threads, OrcCounters* counters, int policy) {
    if (!ob || !pc) return -1;
    Bind b = make_bind(ob, pc);
    uint32_t width = (uint32_t)pc->resolution[0], height = (uint32_t)pc->resolution[1];
    uint32_t tw = pc->tile_size_packed & 0xFFFFu, th = (pc->tile_size_packed >> 16) & 0xFFFFu;
    /* invocation grid actually launched (multiples of the 16x16 group) */
    uint32_t gw = ((tw + 15u) / 16u) * 16u, gh = ((th + 15u) / 16u) * 16u;
    uint32_t channel = rt_pc_color_channel(pc);
    if (threads < 1) threads = 1;
    std::vector<Counters> cs((size_t)threads);
    parallel_spans(gh, gw, threads, [&](uint32_t iy, uint32_t ix0, uint32_t ix1, int tid) {
        Counters& c = cs[(size_t)tid];
        g_policy = policy;
        for (uint32_t ix = ix0; ix < ix1; ix++) {
            uint32_t px = pc->tile_offset[0] + ix, py = pc->tile_offset[1] + iy;
            if (!(ix < tw && iy < th && px < width && py < height)) continue; /* lib.rs:152-163 */
            Hit h;
            V3 col = shade_pixel(b, px, py, &h, c);
            V3 f = filter_channel(col, channel);
            size_t p = (size_t)py * width + px;
            if (image) {
                image[p * 4 + 0] = unorm8(f.x);
                image[p * 4 + 1] = unorm8(f.y);
                image[p * 4 + 2] = unorm8(f.z);
                image[p * 4 + 3] = 255;
            }
            if (radiance) {
                if (channel == 0 || channel > 2) radiance[p * 3 + 0] = col.x;
                if (channel == 1 || channel > 2) radiance[p * 3 + 1] = col.y;
                if (channel == 2 || channel > 2) radiance[p * 3 + 2] = col.z;
            }
            if (hits) {
                hits[p].kind = h.kind;
                hits[p].prim = h.prim;
                hits[p].t_bits = f2u(h.t);
                hits[p].material_id = h.material_id;
            }
        }
    });
    if (counters) {
        for (auto& c : cs) {
            counters->aabb_tests += c.aabb_tests;
            counters->tri_tests += c.tri_tests;
            counters->sphere_tests += c.sphere_tests;
            counters->rays += c.rays;
        }
    }
    return 0;
}

/* EXTENSION (SPEC.md): path tracing with spp / bounces / accumulation over rows
 * [y0, y1) of the frame.  `pc` supplies resolution, camera, counts and offsets (tile
 * fields and flags are ignored).  accum: W*H float4 (rgb sum, sample count), updated in
 * sample order; reset != 0 starts from zero.  hits: primary hit of sample 0 (optional). */
int orc_render_path(const OrcBindings* ob, const RtPushConstants* pc, uint32_t spp, uint32_t first_sample,
                    uint32_t max_depth, uint32_t base_seed, int reset, uint32_t y0, uint32_t y1, float* accum,
                    OrcHit* hits, int threads, OrcCounters* counters, int policy) {
    if (!ob || !pc || !accum) return -1;
    Bind b = make_bind(ob, pc);
    uint32_t width = (uint32_t)pc->resolution[0], height = (uint32_t)pc->resolution[1];
    if (y1 > height) y1 = height;
    if (y0 >= y1) return 0;
    if (threads < 1) threads = 1;
    PathParams pp{spp, first_sample, max_depth, base_seed};
    std::vector<Counters> cs((size_t)threads);
    parallel_spans(y1 - y0, width, threads, [&](uint32_t iy, uint32_t px0, uint32_t px1, int tid) {
        Counters& c = cs[(size_t)tid];
        g_policy = policy;
        uint32_t py = y0 + iy;
        for (uint32_t px = px0; px < px1; px++) {
            size_t p = (size_t)py * width + px;
            float* a = accum + p * 4;
            if (reset) a[0] = a[1] = a[2] = a[3] = 0.0f;
            for (uint32_t s = first_sample; s < first_sample + spp; s++) {
                Hit h;
                V3 L = path_sample(b, px, py, s, pp, c, (s == 0 && hits) ? &h : nullptr);
                a[0] = a[0] + L.x; a[1] = a[1] + L.y; a[2] = a[2] + L.z; a[3] = a[3] + 1.0f;
                if (s == 0 && hits) {
                    hits[p].kind = h.kind; hits[p].prim = h.prim; hits[p].t_bits = f2u(h.t); hits[p].material_id = h.material_id;
                }
            }
        }
    });
    if (counters)
        for (auto& c : cs) {
            counters->aabb_tests += c.aabb_tests; counters->tri_tests += c.tri_tests;
            counters->sphere_tests += c.sphere_tests; counters->rays += c.rays;
        }
    for (int d = 0; d < 8; d++) { g_last_depth_rays[d] = 0; for (auto& c : cs) g_last_depth_rays[d] += c.depth_rays[d]; }
    return 0;
}

/* Ray segments per bounce depth of the last orc_render_path call: what WavefrontCounters.rays_per_bounce
 * (shared/src/lib.rs:187-194) would hold if the reference's wavefront loop were live. */
void orc_last_depth_rays(uint64_t out[8]) { for (int d = 0; d < 8; d++) out[d] = g_last_depth_rays[d]; }

/* resolve: rgb / w -> Rgba8Unorm (SPEC.md step 4) */
void orc_resolve(const float* accum, uint8_t* out, uint64_t pixels) {
    for (uint64_t i = 0; i < pixels; i++) {
        float w = accum[i * 4 + 3];
        out[i * 4 + 0] = unorm8(accum[i * 4 + 0] / w);
        out[i * 4 + 1] = unorm8(accum[i * 4 + 1] / w);
        out[i * 4 + 2] = unorm8(accum[i * 4 + 2] / w);
        out[i * 4 + 3] = 255;
    }
}

/* main_fs (lib.rs:368-391) at 1:1 sampling: out = (R.x, G.y, B.z, 1). */
void orc_composite(const uint8_t* r, const uint8_t* g, const uint8_t* b, uint8_t* out, uint64_t pixels) {
    for (uint64_t i = 0; i < pixels; i++) {
        out[i * 4 + 0] = r[i * 4 + 0];
        out[i * 4 + 1] = g[i * 4 + 1];
        out[i * 4 + 2] = b[i * 4 + 2];
        out[i * 4 + 3] = 255;
    }
}

uint16_t orc_f32_to_f16(float f) { return f32_to_f16_bits(f); }
float orc_f16_to_f32(uint16_t h) { return f16_bits_to_f32(h); }

} /* extern "C" */
