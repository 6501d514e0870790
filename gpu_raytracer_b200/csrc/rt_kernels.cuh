// rt_kernels.cuh — device code of the per-pixel ray tracing path (sm_100a).
//
// Replaces the reference compute shader `main_cs` (shader/src/lib.rs:25-89) and
// everything it calls.  Two closest-hit policies produce identical results:
//   * LITERAL: the reference's own traversal — unordered, un-pruned stack DFS
//     over the 48-byte BvhNode array read from the scene-metadata buffer
//     (shader/src/bvh.rs:18-133), including the `stack_ptr < 63` quirk.
//   * ORDERED: near-first, conservatively t-pruned traversal over a 64-byte
//     two-child node layout and 48-byte pre-gathered triangles built at upload
//     (rt_api.cu), fetched with 128-bit loads.  Equal-t ties are resolved as the
//     reference's visit order would ("first in DFS order wins" == smaller
//     position in triangle_indices), so primitive ids stay bit-exact.
// No tensor cores: this is pointer chasing plus scalar fp32, not a contraction.
#pragma once
#include <cuda_fp16.h>
#include <float.h>
#include "../../include/rt_shared.h"
#include "../../include/rtb200.h"
#include "rt_math.cuh"

#define RT_LEAF_FLAG 0x80000000u
#define RT_CHILD_INTERNAL 0xFFFFFFFFu
#define RT_CHILD_ABSENT 0xFFFFFFFEu
// Conservative pruning margin of the ORDERED policy: a node is culled only if its
// slab entry distance exceeds closest_t * RT_PRUNE_SCALE + RT_PRUNE_ABS.  The
// margin absorbs the rounding difference between the slab t of a box face and the
// Möller–Trumbore t of a triangle lying in that face (flat terrain triangles have
// zero-thickness boxes), so a culled node can never hold the reference's winner.
#define RT_PRUNE_SCALE 1.001f
#define RT_PRUNE_ABS 1.0e-4f

struct KParams {
    // image + camera (push constants: shared/src/lib.rs:215-227)
    uint32_t width, height;   // resolution as u32 (`resolution[i] as u32`)
    float res_x, res_y;       // resolution as f32
    float cam_pos[3], cam_dir[3], cam_up[3];
    float fov_scale;          // tan(fov*0.5*PI/180), evaluated on the host (glibc tanf)
    // work: pixel rectangle of one dispatch, or tiles of a fused frame
    uint32_t rect_x0, rect_y0, rect_w, rect_h;
    uint32_t tiles_x, n_tiles_total, tile_stride, tile_phase; // fused frame: RT_SPLIT_TILE^2 tile split
    const uint32_t* tile_list; // explicit assignment (rt_set_tile_list): entry j = a tile of the frame; n_tiles_total = its
                               // length and stride / phase walk the list.  NULL: entry j is tile j (tile i -> i % stride)
    uint32_t use_tiles;       // 1: fused frame over tiles; 0: single rect
    uint32_t mode;            // RT_MODE_*
    uint32_t outputs;         // RT_OUT_*
    uint32_t channel;         // single dispatch: colour channel (0..2, >2 = unfiltered); fused: ignored
    uint32_t cur_depth, max_depth; // wavefront flags
    uint32_t reset_accum;
    // scene in the reference layout
    const uint32_t* meta;
    uint64_t meta_len;
    RtSceneMetadataOffsets off;
    const RtTriangle* tris;
    uint64_t n_tris;
    uint32_t triangles_per_buffer;     // value in the push constants
    uint32_t ctx_triangles_per_buffer; // length of every bound triangle buffer (BufferManager.triangles_per_buffer)
    uint32_t n_tri_buffers;            // bound buffers that exist; binding k >= this aliases buffer 0 (src/renderer.rs:680-693)
    uint32_t triangle_count;           // push_constants.triangle_count (brute-force loop bound)
    const RtMaterial* mats;
    uint32_t n_mats;
    // re-laid-out scene (ORDERED policy)
    const float4* wnodes;  // 4 x float4 per internal node
    const float4* wtris;   // 3 x float4 per triangle position
    uint32_t root_ref;     // internal index, or RT_LEAF_FLAG|start, or RT_CHILD_ABSENT
    float root_min[3], root_max[3];
    // internally rebuilt small-leaf LBVH (REBUILT policy, rt_lbvh.cuh)
    const float4* bnodes;          // 4 x float4 per radix-tree node
    const float4* btris;           // 3 x float4 per sorted slot; last word = triangle_indices position
    const uint32_t* leaf_of_pos;   // uploaded leaf node of every triangle_indices position
    uint32_t broot_ref;
    float broot_min[3], broot_max[3];
    float rebuilt_far;             // REBUILT: ray origins with a coordinate beyond this take the literal policy
    const float4* cert_box;        // [2 * uploaded node]: (min, max) of the uploaded leaves, aligned copies of the metadata words
    const uint4* bqnodes;          // the same tree as 2 x uint4 per node: 16-bit boxes on the grid below (rt_lbvh.cuh)
    float bq_origin[3], bq_scale[3];
    // outputs
    float4* accum;
    RtHitRecord* hits;
    uchar4* chan[3];
    uchar4* composite;
    // instrumented launches only: [0] box tests, [1] triangle tests, [2] ray segments, [3] shaded hits,
    // [4] rays that re-ran the literal policy (REBUILT certification failed)
    unsigned long long* counters;
    // RT_MODE_PATH (EXTENSION, SPEC.md)
    uint32_t spp, first_sample, path_max_depth, base_seed;
    uint32_t fallback_ordered;       // wv_fallback: 1 = pruned ORDERED search over wnodes / wtris, 0 = the literal walk
    uint32_t legacy_camera;          // wave engine primary frames: 1 = Ray::from_screen_coordinates (legacy), 0 = wavefront camera ray
    uint32_t n_work;                 // pixels in this launch's work list (4096 per assigned split tile)
    unsigned int* work_counter;      // persistent-thread work fetch
    unsigned long long* seg_counter; // ray segments traced (the Mrays/s numerator)
};

struct Ray {
    V3 o, d;
};

// One 64-byte BVH node = two 256-bit read-only loads (sm_100 LDG.E.256; SASS: LDG.E.ENL2.256.CONSTANT).
// ncu on the traversal kernels showed the L1 data stage (wavefronts per load instruction x distinct
// line) as the busiest unit, so a node costs two load instructions per lane instead of four.
__device__ __forceinline__ void ldg_node(const float4* __restrict__ np, float4& q0, float4& q1, float4& q2, float4& q3) {
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(q0.x), "=f"(q0.y), "=f"(q0.z), "=f"(q0.w), "=f"(q1.x), "=f"(q1.y), "=f"(q1.z), "=f"(q1.w) : "l"(np));
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(q2.x), "=f"(q2.y), "=f"(q2.z), "=f"(q2.w), "=f"(q3.x), "=f"(q3.y), "=f"(q3.z), "=f"(q3.w) : "l"(np + 2));
}

// One 32-byte quantised node = one 256-bit read-only load.
__device__ __forceinline__ void ldg_qnode(const uint4* __restrict__ np, uint32_t (&w)[8]) {
    asm("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(np));
}

struct Hit {
    float t;
    uint32_t kind;  // 0 miss 1 sphere 2 triangle
    uint32_t prim;  // sphere index / triangle index
    uint32_t material_id;
    V3 point, normal;
    uint32_t n_box, n_tri, n_fallback; // instrumentation
};

__device__ __forceinline__ uint32_t meta_at(const KParams& p, uint64_t i) {
    return i < p.meta_len ? __ldg(p.meta + i) : 0u;
}
__device__ __forceinline__ float meta_f(const KParams& p, uint64_t i) { return __uint_as_float(meta_at(p, i)); }

// Same node load with an L2 evict_last policy: the wave engine streams gigabytes of path state through L2
// between traversal launches (ncu: L2 hit rate of the deep, sparse launches fell to 24 % and each node step
// became an HBM round trip); BVH lines marked evict_last stay resident while the streaming lines, loaded and
// stored with evict-first (.cs) hints, go first.
__device__ __forceinline__ unsigned long long l2_keep_policy() {
    unsigned long long pol;
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void ldg_node_keep(const float4* __restrict__ np, unsigned long long pol, float4& q0, float4& q1, float4& q2, float4& q3) {
    asm("ld.global.nc.L2::cache_hint.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8], %9;"
        : "=f"(q0.x), "=f"(q0.y), "=f"(q0.z), "=f"(q0.w), "=f"(q1.x), "=f"(q1.y), "=f"(q1.z), "=f"(q1.w) : "l"(np), "l"(pol));
    asm("ld.global.nc.L2::cache_hint.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8], %9;"
        : "=f"(q2.x), "=f"(q2.y), "=f"(q2.z), "=f"(q2.w), "=f"(q3.x), "=f"(q3.y), "=f"(q3.z), "=f"(q3.w) : "l"(np + 2), "l"(pol));
}
__device__ __forceinline__ float4 ldg_keep(const float4* __restrict__ p, unsigned long long pol) {
    float4 r;
    asm("ld.global.nc.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p), "l"(pol));
    return r;
}

// ---------------------------------------------------------------- camera rays
// shader/src/ray.rs:22-53 (legacy: double normalise, u32 resolution) and
// shader/src/wavefront.rs:75-112 (wavefront: single normalise, f32 resolution).
__device__ __forceinline__ Ray camera_ray(const KParams& p, uint32_t px, uint32_t py, float jx, float jy,
                                          bool wavefront) {
    float w = wavefront ? p.res_x : (float)p.width;
    float h = wavefront ? p.res_y : (float)p.height;
    float u = ((float)px + jx) / w;
    float v = ((float)py + jy) / h;
    float aspect = w / h;
    float cx = (u * 2.0f - 1.0f) * aspect * p.fov_scale;
    float cy = (1.0f - v * 2.0f) * p.fov_scale;
    V3 forward = mka(p.cam_dir), up = mka(p.cam_up);
    V3 right = cross3(forward, up);
    V3 true_up = cross3(right, forward);
    V3 dir = forward + right * cx + true_up * cy;
    V3 dn = normalize3(dir);
    if (!wavefront) dn = normalize3(dn); // Ray::new normalises again (ray.rs:14-19)
    Ray r;
    r.o = mka(p.cam_pos);
    r.d = dn;
    return r;
}

// ---------------------------------------------------------------- primitives
// shader/src/intersection.rs:52-87
__device__ __forceinline__ bool sphere_t(const Ray& ray, V3 c, float r, float max_t, float* t_out) {
    V3 oc = ray.o - c;
    float a = dot3(ray.d, ray.d);
    float b = 2.0f * dot3(oc, ray.d);
    float cq = dot3(oc, oc) - r * r;
    float disc = b * b - 4.0f * a * cq;
    if (disc < 0.0f) return false;
    float sq = sqrtf(disc);
    float t1 = (-b - sq) / (2.0f * a);
    float t2 = (-b + sq) / (2.0f * a);
    float t = t1 > RT_MIN_RAY_DISTANCE ? t1 : t2;
    if (t > RT_MIN_RAY_DISTANCE && t < max_t) { *t_out = t; return true; }
    return false;
}

// shader/src/intersection.rs:91-138 with e1/e2 given.  Returns t (accepted iff
// RT_MIN_RAY_DISTANCE < t; the caller applies the `< max_t` / tie rule) or -1.
__device__ __forceinline__ float triangle_t(const Ray& ray, V3 v0, V3 e1, V3 e2) {
    V3 h = cross3(ray.d, e2);
    float a = dot3(e1, h);
    if (fabsf(a) < RT_MIN_RAY_DISTANCE) return -1.0f;
    float f = 1.0f / a;
    V3 s = ray.o - v0;
    float u = f * dot3(s, h);
    if (u < 0.0f || u > 1.0f) return -1.0f;
    V3 q = cross3(s, e1);
    float v = f * dot3(ray.d, q);
    if (v < 0.0f || u + v > 1.0f) return -1.0f;
    float t = f * dot3(e2, q);
    return (t > RT_MIN_RAY_DISTANCE) ? t : -1.0f;
}

// shader/src/intersection.rs:151-164; returns tmin_max through *entry.
__device__ __forceinline__ bool slab(V3 o, V3 inv, V3 bmin, V3 bmax, float* entry) {
    V3 t1 = (bmin - o) * inv;
    V3 t2 = (bmax - o) * inv;
    float tminx = fminf(t1.x, t2.x), tminy = fminf(t1.y, t2.y), tminz = fminf(t1.z, t2.z);
    float tmaxx = fmaxf(t1.x, t2.x), tmaxy = fmaxf(t1.y, t2.y), tmaxz = fmaxf(t1.z, t2.z);
    float tmin_max = fmaxf(fmaxf(tminx, tminy), tminz);
    float tmax_min = fminf(fminf(tmaxx, tmaxy), tmaxz);
    *entry = tmin_max;
    return tmax_min >= 0.0f && tmin_max <= tmax_min;
}

// Conservative-box variant for the internally rebuilt tree only (never for boxes whose test
// must equal the reference's): one FMA per plane.  Differs from slab() by rounding; the
// rebuilt boxes are inflated by 2^-16 relative, far more than that difference.
__device__ __forceinline__ bool slab_fma(V3 inv, V3 oinv, V3 bmin, V3 bmax, float* entry) {
    float t1x = __fmaf_rn(bmin.x, inv.x, -oinv.x), t1y = __fmaf_rn(bmin.y, inv.y, -oinv.y), t1z = __fmaf_rn(bmin.z, inv.z, -oinv.z);
    float t2x = __fmaf_rn(bmax.x, inv.x, -oinv.x), t2y = __fmaf_rn(bmax.y, inv.y, -oinv.y), t2z = __fmaf_rn(bmax.z, inv.z, -oinv.z);
    float tmin_max = fmaxf(fmaxf(fminf(t1x, t2x), fminf(t1y, t2y)), fminf(t1z, t2z));
    float tmax_min = fminf(fminf(fmaxf(t1x, t2x), fmaxf(t1y, t2y)), fmaxf(t1z, t2z));
    *entry = tmin_max;
    return tmax_min >= 0.0f && tmin_max <= tmax_min;
}

// ------------------------------------------------- LITERAL closest-hit policy
// shader/src/triangle_access.rs:18-61 over the buffers the reference binds: max(1, ceil(n / B)) buffers of
// exactly B = BufferManager.triangles_per_buffer zero-initialised entries each (src/buffers.rs:93-99, 274-325),
// binding k = buffer k if it exists, else buffer 0 again (src/renderer.rs:680-693).  The buffer index / local
// index arithmetic uses the push-constant value `tpb_pc`.  Returns false for "invalid" (buffer > 2 or local
// index past the bound length); otherwise *flat is the position in the flat uploaded array (it may lie past the
// uploaded count: the zero-filled tail of a buffer).
__host__ __device__ __forceinline__ bool resolve_bound_triangle(uint32_t index, uint32_t tpb_pc, uint32_t tpb_ctx,
                                                                uint32_t n_buffers, uint64_t* flat) {
    if (tpb_pc == 0u) return false; // the reference would trap on the division; uploads never send 0
    uint32_t bi = index / tpb_pc;
    uint64_t li = index % tpb_pc;
    if (bi > 2u || li >= tpb_ctx) return false;
    if (bi >= n_buffers) bi = 0u;
    *flat = (uint64_t)bi * tpb_ctx + li;
    return true;
}
__device__ __forceinline__ bool fetch_triangle_ref(const KParams& p, uint32_t index, V3* v0, V3* v1, V3* v2,
                                                   uint32_t* mat) {
    uint64_t gi;
    if (!resolve_bound_triangle(index, p.triangles_per_buffer, p.ctx_triangles_per_buffer, p.n_tri_buffers, &gi)) return false;
    uint4 t = make_uint4(0u, 0u, 0u, 0u); // zero-initialised tail of a bound buffer
    if (gi < p.n_tris) t = __ldg(reinterpret_cast<const uint4*>(p.tris) + gi);
    uint64_t vo = p.off.vertices_offset;
    *v0 = mk(meta_f(p, vo + (uint64_t)t.x * 3), meta_f(p, vo + (uint64_t)t.x * 3 + 1), meta_f(p, vo + (uint64_t)t.x * 3 + 2));
    *v1 = mk(meta_f(p, vo + (uint64_t)t.y * 3), meta_f(p, vo + (uint64_t)t.y * 3 + 1), meta_f(p, vo + (uint64_t)t.y * 3 + 2));
    *v2 = mk(meta_f(p, vo + (uint64_t)t.z * 3), meta_f(p, vo + (uint64_t)t.z * 3 + 1), meta_f(p, vo + (uint64_t)t.z * 3 + 2));
    *mat = t.w;
    return true;
}

struct TriBest {
    float t;       // closest accepted t so far (seeded with max_t)
    uint32_t prim; // triangle index of the winner
    uint32_t mat;
    V3 v0, e1, e2; // winner's geometry (for point/normal)
    bool hit;
    uint32_t n_box, n_tri, n_fallback; // instrumentation (dead code unless COUNT)
};

__device__ __forceinline__ void literal_test(const KParams& p, const Ray& ray, uint32_t tri_index, TriBest& b) {
    V3 v0, v1, v2;
    uint32_t mat;
    if (!fetch_triangle_ref(p, tri_index, &v0, &v1, &v2, &mat)) return;
    b.n_tri++;
    V3 e1 = v1 - v0, e2 = v2 - v0;
    float t = triangle_t(ray, v0, e1, e2);
    if (t > 0.0f && t < b.t) {
        b.t = t; b.prim = tri_index; b.mat = mat; b.v0 = v0; b.e1 = e1; b.e2 = e2; b.hit = true;
    }
}

// shader/src/bvh.rs:18-133 (stack quirk included) / lib.rs:272-296 (brute force)
__device__ void closest_triangle_literal(const KParams& p, const Ray& ray, float max_t, TriBest& b) {
    b.t = max_t; b.hit = false; b.prim = 0xFFFFFFFFu; b.mat = 0xFFFFFFFFu; b.n_box = 0; b.n_tri = 0;
    uint32_t node_count = p.off.bvh_nodes_count;
    if (node_count == 0u) {
        for (uint32_t i = 0; i < p.triangle_count; i++) literal_test(p, ray, i, b);
        return;
    }
    V3 inv = mk(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
    uint32_t stack[64];
    int sp = 0;
    stack[sp++] = 0u;
    while (sp > 0) {
        uint32_t n = stack[--sp];
        if (n == 0xFFFFFFFFu || n >= node_count) continue;
        uint64_t no = (uint64_t)p.off.bvh_nodes_offset + (uint64_t)n * 12;
        V3 bmin = mk(meta_f(p, no), meta_f(p, no + 1), meta_f(p, no + 2));
        V3 bmax = mk(meta_f(p, no + 4), meta_f(p, no + 5), meta_f(p, no + 6));
        float entry;
        b.n_box++;
        if (!slab(ray.o, inv, bmin, bmax, &entry)) continue;
        uint32_t left = meta_at(p, no + 8), right = meta_at(p, no + 9);
        if (left == 0xFFFFFFFFu) {
            uint32_t start = meta_at(p, no + 10), count = meta_at(p, no + 11);
            for (uint32_t i = 0; i < count; i++) {
                if (start + i >= p.off.triangle_indices_count) break;
                uint32_t ti = meta_at(p, (uint64_t)p.off.triangle_indices_offset + start + i);
                literal_test(p, ray, ti, b);
            }
        } else {
            if (right != 0xFFFFFFFFu && sp < 63) stack[sp++] = right;
            if (left != 0xFFFFFFFFu && sp < 63) stack[sp++] = left;
        }
    }
}

// The same search by a whole warp for ONE ray (every lane passes the same ray): the walk over the nodes is
// uniform, the triangles of a leaf (or of the brute-force loop) are dealt to the lanes, and the result is the
// minimum of (t, position in the reference's visit order) over all accepted triangles, which is exactly what the
// sequential strict `t < closest` scan returns (the first triangle in visit order that attains the minimum).
// Used by wv_fallback for the rare rays of the wave engine whose REBUILT winner fails certification (or that the
// internal tree cannot serve): alone on one lane such a ray takes ~0.25 ms on 100-triangle leaves.
// Must be called by all 32 lanes of a converged warp; every lane returns the same result.
__device__ __noinline__ void closest_triangle_literal_warp(const KParams& p, const Ray& ray, float max_t, TriBest& b) {
    const unsigned full = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    TriBest m;
    m.t = max_t; m.hit = false; m.prim = 0xFFFFFFFFu; m.mat = 0xFFFFFFFFu; m.n_box = 0; m.n_tri = 0; m.n_fallback = 0;
    m.v0 = mk(0, 0, 0); m.e1 = mk(0, 0, 0); m.e2 = mk(0, 0, 0);
    uint32_t my_seq = 0xFFFFFFFFu, seq = 0u;
    auto test = [&](uint32_t tri_index, uint32_t s) {
        V3 v0, v1, v2;
        uint32_t mat;
        if (!fetch_triangle_ref(p, tri_index, &v0, &v1, &v2, &mat)) return;
        m.n_tri++;
        V3 e1 = v1 - v0, e2 = v2 - v0;
        float t = triangle_t(ray, v0, e1, e2);
        if (t > 0.0f && t < m.t) { // a lane sees its triangles in visit order: strict less keeps the first
            m.t = t; m.prim = tri_index; m.mat = mat; m.v0 = v0; m.e1 = e1; m.e2 = e2; m.hit = true; my_seq = s;
        }
    };
    const uint32_t node_count = p.off.bvh_nodes_count;
    if (node_count == 0u) {
        for (uint32_t i = lane; i < p.triangle_count; i += 32u) test(i, i);
    } else {
        V3 inv = mk(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
        uint32_t stack[64];
        int sp = 0;
        stack[sp++] = 0u;
        while (sp > 0) {
            uint32_t n = stack[--sp];
            if (n == 0xFFFFFFFFu || n >= node_count) continue;
            uint64_t no = (uint64_t)p.off.bvh_nodes_offset + (uint64_t)n * 12;
            V3 bmin = mk(meta_f(p, no), meta_f(p, no + 1), meta_f(p, no + 2));
            V3 bmax = mk(meta_f(p, no + 4), meta_f(p, no + 5), meta_f(p, no + 6));
            float entry;
            m.n_box++;
            if (!slab(ray.o, inv, bmin, bmax, &entry)) continue;
            uint32_t left = meta_at(p, no + 8), right = meta_at(p, no + 9);
            if (left == 0xFFFFFFFFu) {
                uint32_t start = meta_at(p, no + 10), count = meta_at(p, no + 11);
                for (uint32_t i = lane; i < count; i += 32u) {
                    if (start + i >= p.off.triangle_indices_count) break;
                    test(meta_at(p, (uint64_t)p.off.triangle_indices_offset + start + i), seq + i);
                }
                seq += count;
            } else {
                if (right != 0xFFFFFFFFu && sp < 63) stack[sp++] = right;
                if (left != 0xFFFFFFFFu && sp < 63) stack[sp++] = left;
            }
        }
    }
    float bt = m.t;
    uint32_t bs = my_seq, nt = m.n_tri;
    for (int o = 16; o; o >>= 1) {
        float ot = __shfl_xor_sync(full, bt, o);
        uint32_t os = __shfl_xor_sync(full, bs, o);
        nt += __shfl_xor_sync(full, nt, o);
        if (ot < bt || (ot == bt && os < bs)) { bt = ot; bs = os; }
    }
    const unsigned wm = __ballot_sync(full, m.hit && my_seq == bs);
    b.t = max_t; b.hit = false; b.prim = 0xFFFFFFFFu; b.mat = 0xFFFFFFFFu; b.n_box = m.n_box; b.n_tri = nt;
    if (wm) {
        const int src = __ffs(wm) - 1;
        b.hit = true;
        b.t = __shfl_sync(full, m.t, src); b.prim = __shfl_sync(full, m.prim, src); b.mat = __shfl_sync(full, m.mat, src);
        b.v0 = mk(__shfl_sync(full, m.v0.x, src), __shfl_sync(full, m.v0.y, src), __shfl_sync(full, m.v0.z, src));
        b.e1 = mk(__shfl_sync(full, m.e1.x, src), __shfl_sync(full, m.e1.y, src), __shfl_sync(full, m.e1.z, src));
        b.e2 = mk(__shfl_sync(full, m.e2.x, src), __shfl_sync(full, m.e2.y, src), __shfl_sync(full, m.e2.z, src));
    }
}

// ------------------------------------------------- ORDERED closest-hit policy
// Wide node i = 4 x float4:
//   q0 = (L.min.x, L.min.y, L.min.z, L.max.x)   q1 = (L.max.y, L.max.z, R.min.x, R.min.y)
//   q2 = (R.min.z, R.max.x, R.max.y, R.max.z)   q3 = (Lref, Rref, -, -) as u32
// ref: internal node index, or RT_LEAF_FLAG|first triangle position, or RT_CHILD_ABSENT.
// Triangle position j = 3 x float4: (v0.xyz, e1.x) (e1.yz, e2.xy) (e2.z, material, tri_index, remaining)
// where `remaining` counts the triangles left in the leaf including this one.
__device__ __forceinline__ void ordered_leaf(const KParams& p, const Ray& ray, uint32_t start, TriBest& b,
                                             uint32_t& best_pos) {
    uint32_t pos = start;
    for (;;) {
        const float4* tp = p.wtris + (uint64_t)pos * 3;
        float4 a = __ldg(tp), c = __ldg(tp + 1), d = __ldg(tp + 2);
        V3 v0 = mk(a.x, a.y, a.z), e1 = mk(a.w, c.x, c.y), e2 = mk(c.z, c.w, d.x);
        float t = triangle_t(ray, v0, e1, e2);
        b.n_tri++;
        // reference: strict `t < closest` in visit order; with re-ordered visits the
        // equal-t winner is the smaller position in triangle_indices.
        if (t > 0.0f && (t < b.t || (t == b.t && b.hit && pos < best_pos))) {
            b.t = t; b.hit = true; best_pos = pos;
            b.mat = __float_as_uint(d.y); b.prim = __float_as_uint(d.z);
            b.v0 = v0; b.e1 = e1; b.e2 = e2;
        }
        uint32_t remaining = __float_as_uint(d.w);
        if (remaining <= 1u) break;
        pos++;
    }
}

__device__ void closest_triangle_ordered(const KParams& p, const Ray& ray, float max_t, TriBest& b) {
    b.t = max_t; b.hit = false; b.prim = 0xFFFFFFFFu; b.mat = 0xFFFFFFFFu; b.n_box = 1; b.n_tri = 0;
    uint32_t best_pos = 0u;
    uint32_t cur = p.root_ref;
    if (cur == RT_CHILD_ABSENT) return;
    V3 inv = mk(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
    {
        float entry;
        if (!slab(ray.o, inv, mka(p.root_min), mka(p.root_max), &entry)) return;
    }
    uint32_t stack[64];
    int sp = 0;
    // "while-while" (Aila & Laine): an inner loop over internal nodes that a lane leaves when
    // it reaches a leaf, then the leaf's triangles.  Lanes re-converge after the node loop, so
    // the Möller–Trumbore code runs with every lane that found a leaf instead of the one or two
    // lanes that happen to be at a leaf in a given step (ncu: 2 of 32 lanes before this change).
    const uint32_t DONE = 0xFFFFFFFFu; // has RT_LEAF_FLAG set: ends the node loop
    for (;;) {
        while (!(cur & RT_LEAF_FLAG)) {
            const float4* np = p.wnodes + (uint64_t)cur * 4;
            float4 q0, q1, q2, q3;
            ldg_node(np, q0, q1, q2, q3);
            uint32_t lref = __float_as_uint(q3.x), rref = __float_as_uint(q3.y);
            float el, er;
            b.n_box += 2;
            bool hl = slab(ray.o, inv, mk(q0.x, q0.y, q0.z), mk(q0.w, q1.x, q1.y), &el);
            bool hr = slab(ray.o, inv, mk(q1.z, q1.w, q2.x), mk(q2.y, q2.z, q2.w), &er);
            float limit = b.t * RT_PRUNE_SCALE + RT_PRUNE_ABS;
            hl = hl && (lref != RT_CHILD_ABSENT) && (el <= limit);
            hr = hr && (rref != RT_CHILD_ABSENT) && (er <= limit);
            if (hl && hr) {
                // near child first; equal entry -> left first (the reference's order)
                bool right_first = er < el;
                stack[sp++] = right_first ? lref : rref;
                cur = right_first ? rref : lref;
            } else if (hl) cur = lref;
            else if (hr) cur = rref;
            else cur = sp ? stack[--sp] : DONE;
        }
        if (cur == DONE) break;
        ordered_leaf(p, ray, cur & ~RT_LEAF_FLAG, b, best_pos);
        cur = sp ? stack[--sp] : DONE;
    }
}

// The ORDERED search by a whole warp for ONE ray (every lane passes the same ray; all 32 lanes of a converged warp
// must call it; every lane returns the same result): the near-first, t-pruned walk is uniform, the triangles of a
// leaf are dealt to the lanes and folded with a warp-wide minimum of (t, position).  The ORDERED policy returns the
// reference's (= the LITERAL policy's) hit for every ray on a validated upload, so wv_fallback uses this for the
// rays the internal tree cannot serve: pruned, it visits a few dozen nodes where the un-pruned literal walk of a
// grazing ray visits a thousand (measured: 0.25-0.4 ms for ONE such ray, on one lane or on a warp alike).
__device__ __forceinline__ void closest_triangle_ordered_warp(const KParams& p, const Ray& ray, float max_t, TriBest& b) {
    const unsigned full = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    b.t = max_t; b.hit = false; b.prim = 0xFFFFFFFFu; b.mat = 0xFFFFFFFFu; b.n_box = 1; b.n_tri = 0; b.n_fallback = 0;
    b.v0 = mk(0, 0, 0); b.e1 = mk(0, 0, 0); b.e2 = mk(0, 0, 0);
    uint32_t best_pos = 0u, n_tri = 0u;
    uint32_t cur = p.root_ref;
    if (cur == RT_CHILD_ABSENT) return;
    V3 inv = mk(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
    {
        float entry;
        if (!slab(ray.o, inv, mka(p.root_min), mka(p.root_max), &entry)) return;
    }
    uint32_t stack[64];
    int sp = 0;
    const uint32_t DONE = 0xFFFFFFFFu;
    for (;;) {
        while (!(cur & RT_LEAF_FLAG)) { // uniform: the same node step as closest_triangle_ordered
            const float4* np = p.wnodes + (uint64_t)cur * 4;
            float4 q0, q1, q2, q3;
            ldg_node(np, q0, q1, q2, q3);
            uint32_t lref = __float_as_uint(q3.x), rref = __float_as_uint(q3.y);
            float el, er;
            b.n_box += 2;
            bool hl = slab(ray.o, inv, mk(q0.x, q0.y, q0.z), mk(q0.w, q1.x, q1.y), &el);
            bool hr = slab(ray.o, inv, mk(q1.z, q1.w, q2.x), mk(q2.y, q2.z, q2.w), &er);
            float limit = b.t * RT_PRUNE_SCALE + RT_PRUNE_ABS;
            hl = hl && (lref != RT_CHILD_ABSENT) && (el <= limit);
            hr = hr && (rref != RT_CHILD_ABSENT) && (er <= limit);
            if (hl && hr) {
                bool right_first = er < el;
                stack[sp++] = right_first ? lref : rref;
                cur = right_first ? rref : lref;
            } else if (hl) cur = lref;
            else if (hr) cur = rref;
            else cur = sp ? stack[--sp] : DONE;
        }
        if (cur == DONE) break;
        const uint32_t start = cur & ~RT_LEAF_FLAG;
        const uint32_t cnt = __float_as_uint(__ldg(p.wtris + (uint64_t)start * 3 + 2).w); // `remaining` of the first record
        float mt = 0.0f;
        uint32_t mpos = 0xFFFFFFFFu; // this lane's best of the leaf: valid iff mpos != ~0
        for (uint32_t i = lane; i < cnt; i += 32u) {
            const uint32_t pos = start + i;
            const float4* tp = p.wtris + (uint64_t)pos * 3;
            float4 a = __ldg(tp), c = __ldg(tp + 1), d = __ldg(tp + 2);
            float t = triangle_t(ray, mk(a.x, a.y, a.z), mk(a.w, c.x, c.y), mk(c.z, c.w, d.x));
            n_tri++;
            if (t > 0.0f && (mpos == 0xFFFFFFFFu || t < mt)) { mt = t; mpos = pos; } // increasing pos per lane: strict less keeps the smaller
        }
        for (int o = 16; o; o >>= 1) {
            float ot = __shfl_xor_sync(full, mt, o);
            uint32_t op = __shfl_xor_sync(full, mpos, o);
            if (op != 0xFFFFFFFFu && (mpos == 0xFFFFFFFFu || ot < mt || (ot == mt && op < mpos))) { mt = ot; mpos = op; }
        }
        // the same acceptance rule as ordered_leaf: strictly closer, or equally close at a smaller position
        if (mpos != 0xFFFFFFFFu && (mt < b.t || (mt == b.t && b.hit && mpos < best_pos))) { b.t = mt; b.hit = true; best_pos = mpos; }
        cur = sp ? stack[--sp] : DONE;
    }
    for (int o = 16; o; o >>= 1) n_tri += __shfl_xor_sync(full, n_tri, o);
    b.n_tri = n_tri;
    if (b.hit) {
        const float4* tp = p.wtris + (uint64_t)best_pos * 3;
        float4 a = __ldg(tp), c = __ldg(tp + 1), d = __ldg(tp + 2);
        b.v0 = mk(a.x, a.y, a.z); b.e1 = mk(a.w, c.x, c.y); b.e2 = mk(c.z, c.w, d.x);
        b.mat = __float_as_uint(d.y); b.prim = __float_as_uint(d.z);
    }
}

// ------------------------------------------------- REBUILT closest-hit policy
// Traverses the internal LBVH (conservatively inflated boxes, leaves of <= 4
// triangles), then certifies the winner W against the reference traversal:
//   * reference result = first-in-DFS-order minimum-t hit among the positions whose
//     uploaded leaf and all its ancestors pass the reference slab test (no pruning in
//     the reference, so that set does not depend on the search order);
//   * W is the minimum over ALL positions the Möller–Trumbore test accepts (ties to
//     the smaller position == earlier in DFS order), so if W's own chain passes, W is
//     the reference result;
//   * upload validated that every ancestor box contains the leaf box component-wise;
//     (b - o) * inv is monotone in b under IEEE rounding, so for rays without zero
//     direction components "leaf box passes" implies "every ancestor passes".
// Hence one exact slab test on W's uploaded leaf box certifies W.  Rays with a zero direction
// component (or a far origin) take the literal policy.  A ray whose winner fails the test is
// searched again with FILTER = true: a triangle is then only accepted if its own uploaded leaf
// box passes the exact slab test, i.e. (same monotonicity argument) if the reference traversal
// tests it; the minimum over exactly those triangles IS the reference result, found at the
// internal tree's speed instead of the un-pruned literal walk's (0.25-0.4 ms for one grazing ray
// on the terrain's 100-triangle leaves).
__device__ __forceinline__ bool rebuilt_leaf_certified(const KParams& p, const Ray& ray, V3 inv, uint32_t pos) {
    const uint32_t leaf = __ldg(p.leaf_of_pos + pos);
    const float4 blo = __ldg(p.cert_box + (uint64_t)leaf * 2), bhi = __ldg(p.cert_box + (uint64_t)leaf * 2 + 1);
    float entry;
    return slab(ray.o, inv, mk(blo.x, blo.y, blo.z), mk(bhi.x, bhi.y, bhi.z), &entry);
}
template <bool FILTER>
__device__ __forceinline__ void closest_triangle_rebuilt_t(const KParams& p, const Ray& ray, float max_t, TriBest& b) {
    // zero (or so tiny that o/d overflows) direction components: certification argument and the
    // FMA slab do not apply -> literal policy
    // origins far outside the scene: the rounding of the exact tests outgrows the boxes' pad (rt_lbvh.cuh)
    if (!FILTER && (!(fabsf(ray.d.x) >= 1e-30f && fabsf(ray.d.y) >= 1e-30f && fabsf(ray.d.z) >= 1e-30f) ||
        !(fmaxf(fmaxf(fabsf(ray.o.x), fabsf(ray.o.y)), fabsf(ray.o.z)) <= p.rebuilt_far))) { closest_triangle_literal(p, ray, max_t, b); b.n_fallback = 1; return; }
    b.t = max_t; b.hit = false; b.prim = 0xFFFFFFFFu; b.mat = 0xFFFFFFFFu; b.n_box = 1; b.n_tri = 0;
    uint32_t best_pos = 0u;
    uint32_t cur = p.broot_ref;
    if (cur == RT_CHILD_ABSENT) return;
    V3 inv = mk(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
    {
        float entry;
        if (!slab(ray.o, inv, mka(p.broot_min), mka(p.broot_max), &entry)) return;
    }
    uint32_t stack[64];
    int sp = 0;
    // One loop, leaf or node per step: measured faster than the while-while form for this
    // policy (incoherent bounce rays on the LBVH: 1520 vs 1207 Mrays/s on C3), the opposite of
    // the ORDERED policy on 1-triangle leaves.  The internal tree's boxes only have to be
    // conservative (not bit-equal to anything), so the slab test uses one FMA per plane:
    // t = b * inv - o * inv.
    const V3 oinv = mk(ray.o.x * inv.x, ray.o.y * inv.y, ray.o.z * inv.z);
    for (;;) {
        if (cur & RT_LEAF_FLAG) {
            uint32_t slot = cur & 0x0FFFFFFFu, cnt = ((cur >> 28) & 7u) + 1u;
            for (uint32_t i = 0; i < cnt; i++, slot++) {
                const float4* tp = p.btris + (uint64_t)slot * 3;
                float4 a = __ldg(tp), c = __ldg(tp + 1), d = __ldg(tp + 2);
                V3 v0 = mk(a.x, a.y, a.z), e1 = mk(a.w, c.x, c.y), e2 = mk(c.z, c.w, d.x);
                float t = triangle_t(ray, v0, e1, e2);
                b.n_tri++;
                uint32_t pos = __float_as_uint(d.w);
                if (t > 0.0f && (t < b.t || (t == b.t && b.hit && pos < best_pos)) &&
                    (!FILTER || rebuilt_leaf_certified(p, ray, inv, pos))) {
                    b.t = t; b.hit = true; best_pos = pos;
                    b.mat = __float_as_uint(d.y); b.prim = __float_as_uint(d.z);
                    b.v0 = v0; b.e1 = e1; b.e2 = e2;
                }
            }
        } else {
            const float4* np = p.bnodes + (uint64_t)cur * 4;
            float4 q0, q1, q2, q3;
            ldg_node(np, q0, q1, q2, q3);
            uint32_t lref = __float_as_uint(q3.x), rref = __float_as_uint(q3.y);
            float el, er;
            b.n_box += 2;
            bool hl = slab_fma(inv, oinv, mk(q0.x, q0.y, q0.z), mk(q0.w, q1.x, q1.y), &el);
            bool hr = slab_fma(inv, oinv, mk(q1.z, q1.w, q2.x), mk(q2.y, q2.z, q2.w), &er);
            float limit = b.t * RT_PRUNE_SCALE + RT_PRUNE_ABS;
            hl = hl && (el <= limit);
            hr = hr && (er <= limit);
            if (hl && hr) {
                bool right_first = er < el;
                stack[sp++] = right_first ? lref : rref;
                cur = right_first ? rref : lref;
                continue;
            } else if (hl) { cur = lref; continue; }
            else if (hr) { cur = rref; continue; }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    if (!FILTER && b.hit) {
        b.n_box++;
        if (!rebuilt_leaf_certified(p, ray, inv, best_pos)) b.n_fallback = 2; // the caller searches again with the filter
    }
}
// (ptxas 12.9 crashes when the filtered search or closest_triangle_ordered_warp are emitted as out-of-line functions:
// both are inlined into wv_fallback, their only caller; the per-lane megakernels keep the literal walk for the
// rare failures so that their code does not grow.)
__device__ void closest_triangle_rebuilt(const KParams& p, const Ray& ray, float max_t, TriBest& b) {
    closest_triangle_rebuilt_t<false>(p, ray, max_t, b);
    if (b.n_fallback == 2u) {
        uint32_t nb = b.n_box, nt = b.n_tri;
        closest_triangle_literal(p, ray, max_t, b);
        b.n_box += nb; b.n_tri += nt; b.n_fallback = 1;
    }
}

// ---------------------------------------------------------------- closest hit
// shader/src/lib.rs:174-269: spheres first (brute force), then triangles seeded
// with the sphere t; the triangle wins only if strictly closer.
// POLICY: 0 literal, 1 ordered (uploaded tree), 2 rebuilt (internal LBVH)
template <int POLICY>
__device__ __forceinline__ Hit find_closest(const KParams& p, const Ray& ray) {
    Hit h;
    h.t = FLT_MAX; h.kind = 0u; h.prim = 0xFFFFFFFFu; h.material_id = 0xFFFFFFFFu;
    h.point = mk(0, 0, 0); h.normal = mk(0, 0, 0);
    float closest = FLT_MAX - 2.0f;
    uint32_t sphere_idx = 0xFFFFFFFFu;
    for (uint32_t i = 0; i < p.off.spheres_count; i++) {
        uint64_t o = (uint64_t)p.off.spheres_offset + (uint64_t)i * 5;
        float t;
        if (sphere_t(ray, mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2)), meta_f(p, o + 3), closest, &t)) {
            closest = t; sphere_idx = i;
        }
    }
    TriBest b;
    b.n_fallback = 0;
    if (POLICY == 2) closest_triangle_rebuilt(p, ray, closest, b);
    else if (POLICY == 1) closest_triangle_ordered(p, ray, closest, b);
    else closest_triangle_literal(p, ray, closest, b);
    h.n_box = b.n_box; h.n_tri = b.n_tri; h.n_fallback = b.n_fallback;
    if (b.hit) { // b.t < sphere t by construction
        h.t = b.t; h.kind = 2u; h.prim = b.prim; h.material_id = b.mat;
        h.point = ray.o + ray.d * b.t;
        h.normal = normalize3(cross3(b.e1, b.e2)); // geometric, un-flipped (intersection.rs:132)
    } else if (sphere_idx != 0xFFFFFFFFu) {
        uint64_t o = (uint64_t)p.off.spheres_offset + (uint64_t)sphere_idx * 5;
        V3 c = mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2));
        h.t = closest; h.kind = 1u; h.prim = sphere_idx; h.material_id = meta_at(p, o + 4);
        h.point = ray.o + ray.d * closest;
        h.normal = normalize3(h.point - c);
    }
    return h;
}

// -------------------------------------------------------------------- shading
struct MatEval {
    V3 albedo, emission;
    float metallic, roughness, ior, transmission;
};
// shader/src/material.rs:6-63; only the first 32 bytes of the 128-byte Material are read.
__device__ __forceinline__ MatEval load_material(const KParams& p, uint32_t id) {
    const float4* mp = reinterpret_cast<const float4*>(p.mats + id);
    float4 a = __ldg(mp), b = __ldg(mp + 1);
    MatEval m;
    m.albedo = mk(a.x, a.y, a.z);
    m.emission = mk(b.x, b.y, b.z);
    uint32_t mr = __float_as_uint(a.w), it = __float_as_uint(b.w);
    m.metallic = f16lo(mr); m.roughness = f16hi(mr);
    m.ior = f16lo(it); m.transmission = f16hi(it);
    return m;
}
// shader/src/material.rs:76-83
__device__ __forceinline__ V3 evaluate_brdf(const MatEval& m, float I) {
    V3 diffuse = m.albedo / 3.14159265358979323846f;
    float is_m = (m.metallic > 0.5f) ? 1.0f : 0.0f;
    V3 metallic_c = m.albedo * I * 0.5f;
    V3 dielectric_c = diffuse * I;
    return metallic_c * is_m + dielectric_c * (1.0f - is_m);
}
// shader/src/lighting.rs:20-139.  fmaxf(x, 0) scrubs the NaN that
// normalize((0,0,0)) produces for point lights.
__device__ __forceinline__ V3 calculate_lighting(const KParams& p, const Hit& h, const MatEval& m) {
    V3 total = mk(0.0f, 0.0f, 0.0f);
    total = total + m.albedo * 0.1f;
    for (uint32_t li = 0; li < p.off.lights_count; li++) {
        uint64_t o = (uint64_t)p.off.lights_offset + (uint64_t)li * 13;
        V3 lpos = mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2));
        uint32_t ltype = meta_at(p, o + 3);
        V3 lcol = mk(meta_f(p, o + 4), meta_f(p, o + 5), meta_f(p, o + 6));
        float lint = meta_f(p, o + 7);
        V3 ldir = mk(meta_f(p, o + 8), meta_f(p, o + 9), meta_f(p, o + 10));
        V3 nld = -normalize3(ldir);
        float dir_intensity = fmaxf(dot3(h.normal, nld), 0.0f) * lint;
        V3 to_light = lpos - h.point;
        float distance = length3(to_light);
        V3 pdir = normalize3(to_light);
        float att = round_through_f16(1.0f / (1.0f + distance * distance * 0.01f));
        float point_intensity = fmaxf(dot3(h.normal, pdir), 0.0f) * lint * att;
        float spot_factor = fmaxf(dot3(nld, pdir), 0.0f);
        float spot_intensity = point_intensity * spot_factor;
        float is_dir = (ltype == 0u) ? 1.0f : 0.0f, is_point = (ltype == 1u) ? 1.0f : 0.0f,
              is_spot = (ltype == 2u) ? 1.0f : 0.0f;
        float I = dir_intensity * is_dir + point_intensity * is_point + spot_intensity * is_spot;
        V3 contrib = evaluate_brdf(m, I);
        float valid = (I > 0.0f) ? 1.0f : 0.0f; // index_valid is 1 inside the loop bound
        total = total + contrib * lcol * valid;
    }
    return total + m.emission;
}
// shader/src/lib.rs:300-338 for all three colour channels at once: channel c only
// enters through ior_for_channel(c) and the final component select, and every
// Vec3 op is component-wise, so out.c equals pass c's value bit for bit.
// `channel` < 3 evaluates that single pass; >= 3 the fused triple.
__device__ __forceinline__ V3 shade_hit_m(const KParams& p, const Hit& h, const MatEval& m, uint32_t channel);
__device__ __forceinline__ V3 shade_hit(const KParams& p, const Hit& h, uint32_t channel) {
    if (h.material_id >= p.n_mats) return mk(1.0f, 0.0f, 1.0f);
    MatEval m = load_material(p, h.material_id);
    return shade_hit_m(p, h, m, channel);
}
__device__ __forceinline__ V3 shade_hit_m(const KParams& p, const Hit& h, const MatEval& m, uint32_t channel) {
    V3 lighting = calculate_lighting(p, h, m);
    float T = fminf(fmaxf(m.transmission, 0.0f), 1.0f);
    if (!(T > 0.0f)) return lighting;
    const float offs[4] = {-0.018f, 0.0f, 0.035f, 0.0f};
    float denom = m.ior - 1.0f;
    V3 a = lighting * (1.0f - T);
    if (channel < 3u || channel == 0xFFu) {
        if (channel == 0xFFu) {
            float d0 = ((m.ior + offs[0]) - 1.0f) / denom, d1 = ((m.ior + offs[1]) - 1.0f) / denom,
                  d2 = ((m.ior + offs[2]) - 1.0f) / denom;
            return mk(a.x + (0.2f * d0) * T, a.y + (0.2f * d1) * T, a.z + (0.3f * d2) * T);
        }
        float d = ((m.ior + offs[channel]) - 1.0f) / denom;
        return a + mk(0.2f, 0.2f, 0.3f) * d * T;
    }
    float d = ((m.ior + offs[3]) - 1.0f) / denom;
    return a + mk(0.2f, 0.2f, 0.3f) * d * T;
}

// ----------------------------------------------------------------- the kernel
// Work decomposition: a 256-thread block covers a 16x16 pixel block; each warp an
// 8x4 sub-block (coherent primary rays).  In a fused frame the block list
// enumerates the RT_SPLIT_TILE^2 split tiles (numbered like the reference's 128x128 tiles,
// src/compute.rs:194-209) assigned to this context (tile i with i % tile_stride ==
// tile_phase), 16 blocks per tile.
__device__ __forceinline__ bool pixel_of_thread(const KParams& p, uint32_t* px, uint32_t* py) {
    uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    uint32_t lx = (warp & 1u) * 8u + (lane & 7u), ly = (warp >> 1) * 4u + (lane >> 3);
    uint32_t x0, y0, w, h;
    uint32_t blk = blockIdx.x;
    if (p.use_tiles) {
        uint32_t tile_local = blk >> 4, sub = blk & 15u; // 16 blocks of 16x16 pixels per 64x64 split tile
        uint32_t tile = p.tile_phase + tile_local * p.tile_stride;
        if (tile >= p.n_tiles_total) return false;
        if (p.tile_list) tile = __ldg(p.tile_list + tile);
        uint32_t tx = tile % p.tiles_x, ty = tile / p.tiles_x;
        x0 = tx * RT_SPLIT_TILE; y0 = ty * RT_SPLIT_TILE; w = RT_SPLIT_TILE; h = RT_SPLIT_TILE;
        lx += (sub & 3u) * 16u; ly += (sub >> 2) * 16u;
    } else {
        uint32_t bw = (p.rect_w + 15u) >> 4;
        x0 = p.rect_x0; y0 = p.rect_y0; w = p.rect_w; h = p.rect_h;
        lx += (blk % bw) * 16u; ly += (blk / bw) * 16u;
    }
    *px = x0 + lx; *py = y0 + ly;
    // is_pixel_in_bounds, shader/src/lib.rs:152-163
    return lx < w && ly < h && *px < p.width && *py < p.height;
}

// Output stage of one pixel: filter_color_by_channel + Rgba8Unorm store of one reference dispatch
// (lib.rs:82-88, 342-349), or the fused frame's three channel images + main_fs composite (lib.rs:368-391) +
// accumulation; hit record.  Shared by rt_trace_shade and the wave engine's primary-frame shading kernel.
__device__ __forceinline__ void store_pixel(const KParams& p, uint32_t px, uint32_t py, V3 color, const Hit& h, bool fused,
                                            uint32_t channel) {
    const size_t pix = (size_t)py * p.width + px;
    if (!fused) {
        // filter_color_by_channel (lib.rs:342-349) + Rgba8Unorm store
        uchar4 o = make_uchar4(0, 0, 0, 255);
        if (channel == 0u) o.x = (unsigned char)unorm8(color.x);
        else if (channel == 1u) o.y = (unsigned char)unorm8(color.y);
        else o.z = (unsigned char)unorm8(color.z);
        p.chan[channel][pix] = o;
        {   // main_fs (lib.rs:368-391) samples the three channel textures: component c of the composite is
            // whatever the last channel-c dispatch wrote, so the dispatch itself keeps it up to date
            unsigned char* cp = reinterpret_cast<unsigned char*>(p.composite + pix);
            cp[channel] = channel == 0u ? o.x : (channel == 1u ? o.y : o.z);
            cp[3] = 255;
        }
        if (p.outputs & RT_OUT_ACCUM) {
            float4 a = p.accum[pix];
            if (channel == 0u) a.x = color.x; else if (channel == 1u) a.y = color.y; else a.z = color.z;
            a.w = 1.0f;
            p.accum[pix] = a;
        }
    } else {
        uint32_t r8 = unorm8(color.x), g8 = unorm8(color.y), b8 = unorm8(color.z);
        if (p.outputs & RT_OUT_CHANNELS) {
            p.chan[0][pix] = make_uchar4((unsigned char)r8, 0, 0, 255);
            p.chan[1][pix] = make_uchar4(0, (unsigned char)g8, 0, 255);
            p.chan[2][pix] = make_uchar4(0, 0, (unsigned char)b8, 255);
        }
        if (p.outputs & RT_OUT_COMPOSITE) // main_fs, lib.rs:368-391
            p.composite[pix] = make_uchar4((unsigned char)r8, (unsigned char)g8, (unsigned char)b8, 255);
        if (p.outputs & RT_OUT_ACCUM) {
            float4 a = make_float4(color.x, color.y, color.z, 1.0f);
            if (!p.reset_accum) {
                float4 old = p.accum[pix];
                a.x = old.x + color.x; a.y = old.y + color.y; a.z = old.z + color.z; a.w = old.w + 1.0f;
            }
            p.accum[pix] = a;
        }
    }
    if (p.outputs & RT_OUT_HITS) {
        uint4 rec = make_uint4(h.kind, h.prim, __float_as_uint(h.t), h.material_id);
        reinterpret_cast<uint4*>(p.hits)[pix] = rec;
    }
}

template <int POLICY, bool COUNT>
__global__ void __launch_bounds__(256) rt_trace_shade(const __grid_constant__ KParams p) {
    uint32_t px, py;
    if (!pixel_of_thread(p, &px, &py)) return;
    uint32_t n_seg = 0;
    const bool wavefront = p.mode == RT_MODE_WAVEFRONT0;
    const bool fused = p.use_tiles != 0u;
    const uint32_t channel = fused ? 0xFFu : p.channel;
    V3 color = mk(0.0f, 0.0f, 0.0f);
    Hit h;
    h.t = FLT_MAX; h.kind = 0u; h.prim = 0xFFFFFFFFu; h.material_id = 0xFFFFFFFFu; h.n_box = 0; h.n_tri = 0; h.n_fallback = 0;
    if (!wavefront) {
        Ray ray = camera_ray(p, px, py, 0.5f, 0.5f, false);
        n_seg = 1;
        h = find_closest<POLICY>(p, ray);
        if (h.kind != 0u) color = shade_hit(p, h, channel); // miss: (0,0,0), lib.rs:77
    } else if (p.cur_depth <= p.max_depth) {
        // run_wavefront_raytracing, lib.rs:92-149: throughput (1,1,1); sky on miss
        Ray ray = camera_ray(p, px, py, 0.5f, 0.5f, true);
        n_seg = 1;
        h = find_closest<POLICY>(p, ray);
        V3 tp = mk(1.0f, 1.0f, 1.0f);
        V3 c = (h.kind != 0u) ? shade_hit(p, h, channel) * tp : mk(0.1f, 0.2f, 0.3f) * tp;
        color = color + c;
    }
    store_pixel(p, px, py, color, h, fused, channel);
    if (p.outputs & RT_OUT_PEER) __threadfence_system(); // the composite store went to another GPU's frame
    if (COUNT) {
        atomicAdd(p.counters + 0, (unsigned long long)h.n_box);
        atomicAdd(p.counters + 1, (unsigned long long)h.n_tri);
        atomicAdd(p.counters + 2, (unsigned long long)n_seg);
        atomicAdd(p.counters + 3, (unsigned long long)(h.kind != 0u ? 1u : 0u));
        if (h.n_fallback) atomicAdd(p.counters + 4, 1ull);
    }
}

// ------------------------------------------------ EXTENSION: path tracing kernel
// SPEC.md.  Persistent threads: the grid is sized to fill the machine once; every lane
// owns one pixel at a time, runs that pixel's samples in order (so the fp32
// accumulation order is fixed and the result is bit-reproducible) and, when the pixel
// is finished, fetches the next one from a global counter with one warp-aggregated
// atomicAdd (ballot + popc).  Lanes at different bounce depths share the same
// traversal loop, so a finished path is replaced immediately instead of idling until
// the slowest path of the warp ends (path regeneration).
struct SimpleRng { // shader/src/wavefront.rs:46-72
    uint32_t seed;
    __device__ __forceinline__ uint32_t next_u32() { seed = seed * 1664525u + 1013904223u; return seed; }
    __device__ __forceinline__ float next_f32() { return (float)(next_u32() >> 8) / 16777216.0f; }
    __device__ __forceinline__ float next_f32_signed() { return next_f32() * 2.0f - 1.0f; }
};

// SPEC.md steps 3.2-3.8 for one path segment whose closest hit is `h` — the ONE definition of the extension's
// bounce semantics on the device (rt_path_trace and wv_shade both call it; the oracle's twin is path_sample in
// oracle/rt_oracle.cpp).  Adds the segment's radiance to L; returns true when the path ends here, else ray /
// throughput / depth describe the continuation segment.
__device__ __forceinline__ bool path_bounce(const KParams& p, const Hit& h, Ray& ray, V3& tp, V3& L, SimpleRng& rng,
                                            uint32_t& depth) {
    if (h.kind == 0u) { L = L + mk(0.1f, 0.2f, 0.3f) * tp; return true; }              // 3.2 miss: sky
    if (h.material_id >= p.n_mats) { L = L + mk(1.0f, 0.0f, 1.0f) * tp; return true; } // 3.3 invalid material
    MatEval m = load_material(p, h.material_id);
    L = L + shade_hit_m(p, h, m, 0xFFu) * tp;                                          // 3.3
    if (depth == p.path_max_depth) return true;                                        // 3.4
    if (depth >= 3u) {                                                                 // 3.5 Russian roulette
        float pr = fminf(fmaxf(fmaxf(fmaxf(tp.x, tp.y), tp.z), 0.05f), 1.0f);
        float r = rng.next_f32();
        if (r > pr) return true;
        tp = mk(tp.x / pr, tp.y / pr, tp.z / pr);
    }
    float r0 = rng.next_f32();                                                         // 3.6
    V3 u = mk(0.0f, 0.0f, 0.0f);
    for (int k = 0; k < 8; k++) {
        float a = rng.next_f32_signed(), bb = rng.next_f32_signed(), cc = rng.next_f32_signed();
        float q = a * a + bb * bb + cc * cc;
        if (q > 1e-6f && q <= 1.0f) { u = mk(a, bb, cc) * (1.0f / sqrtf(q)); break; }
    }
    V3 d = ray.d, n = h.normal;                                                        // 3.7
    float cosi = dot3(n, d);
    V3 nf = cosi < 0.0f ? n : -n;
    V3 w = nf + u;
    if (dot3(w, w) < 1e-8f) w = nf;
    w = normalize3(w);
    float T = fminf(fmaxf(m.transmission, 0.0f), 1.0f);                                // 3.8
    V3 dir;
    float side = 1.0f;
    if (T > 0.0f && r0 < T) {
        float eta = cosi < 0.0f ? 1.0f / m.ior : m.ior;
        float c1 = -dot3(d, nf);
        float k = 1.0f - eta * eta * (1.0f - c1 * c1);
        if (k < 0.0f) dir = d - nf * (2.0f * dot3(d, nf));
        else { dir = d * eta + nf * (eta * c1 - sqrtf(k)); side = -1.0f; }
        dir = normalize3(dir);
    } else if (m.metallic > 0.5f) {
        V3 mr = d - nf * (2.0f * dot3(d, nf));
        dir = mr * (1.0f - m.roughness) + w * m.roughness;
        if (dot3(dir, dir) < 1e-8f || dot3(dir, nf) <= 0.0f) dir = w;
        dir = normalize3(dir);
    } else {
        dir = w;
    }
    tp = tp * m.albedo;
    ray.o = h.point + nf * (side * 1e-3f);
    ray.d = dir;
    depth++;
    return false;
}

__device__ __forceinline__ bool work_to_pixel(const KParams& p, uint32_t idx, uint32_t* px, uint32_t* py) {
    uint32_t tile_local = idx >> 12, within = idx & 4095u; // 4096 work items per 64x64 split tile
    uint32_t tile = p.tile_phase + tile_local * p.tile_stride;
    if (p.tile_list) tile = __ldg(p.tile_list + tile);
    uint32_t tx = tile % p.tiles_x, ty = tile / p.tiles_x;
    uint32_t sub = within >> 5, l = within & 31u; // 8x4 pixel groups, 8 across x 16 down per tile
    *px = tx * RT_SPLIT_TILE + (sub & 7u) * 8u + (l & 7u);
    *py = ty * RT_SPLIT_TILE + (sub >> 3) * 4u + (l >> 3);
    return *px < p.width && *py < p.height;
}

template <int POLICY, bool COUNT>
__global__ void __launch_bounds__(256) rt_path_trace(const __grid_constant__ KParams p) {
    const uint32_t lane = threadIdx.x & 31u;
    unsigned long long c_box = 0, c_tri = 0, c_shaded = 0, c_fb = 0;
    const unsigned full = 0xFFFFFFFFu;
    bool have = false, exhausted = false, new_sample = false;
    uint32_t px = 0, py = 0, s = 0, depth = 0, n_seg = 0;
    const uint32_t s_end = p.first_sample + p.spp;
    SimpleRng rng; rng.seed = 0;
    Ray ray; ray.o = mk(0, 0, 0); ray.d = mk(0, 0, 1);
    V3 tp = mk(1, 1, 1), L = mk(0, 0, 0), sum = mk(0, 0, 0);
    float wsum = 0.0f;
    for (;;) {
        unsigned idle = __ballot_sync(full, !have && !exhausted);
        if (idle) {
            uint32_t base = 0;
            int leader = __ffs(idle) - 1;
            if ((int)lane == leader) base = atomicAdd(p.work_counter, (unsigned int)__popc(idle));
            base = __shfl_sync(full, base, leader);
            if (!have && !exhausted) {
                uint32_t idx = base + __popc(idle & ((1u << lane) - 1u));
                if (idx >= p.n_work) exhausted = true;
                else if (work_to_pixel(p, idx, &px, &py)) {
                    have = true; new_sample = true; s = p.first_sample;
                    sum = mk(0, 0, 0); wsum = 0.0f;
                    if (!p.reset_accum) {
                        float4 old = p.accum[(size_t)py * p.width + px];
                        sum = mk(old.x, old.y, old.z); wsum = old.w;
                    }
                }
            }
        }
        if (__all_sync(full, !have)) {
            if (__all_sync(full, exhausted)) break;
            continue;
        }
        if (!have) continue;
        if (new_sample) { // SPEC.md steps 1-2
            rng.seed = p.base_seed + s + px + py * p.width;
            float jx = 0.5f, jy = 0.5f;
            if (s != 0u) { jx = rng.next_f32(); jy = rng.next_f32(); }
            ray = camera_ray(p, px, py, jx, jy, true);
            tp = mk(1.0f, 1.0f, 1.0f); L = mk(0.0f, 0.0f, 0.0f); depth = 0; new_sample = false;
        }
        Hit h = find_closest<POLICY>(p, ray); // step 3.1
        n_seg++;
        if (COUNT) { c_box += h.n_box; c_tri += h.n_tri; c_shaded += (h.kind != 0u); c_fb += h.n_fallback; }
        if (depth == 0u && s == 0u && (p.outputs & RT_OUT_HITS))
            reinterpret_cast<uint4*>(p.hits)[(size_t)py * p.width + px] =
                make_uint4(h.kind, h.prim, __float_as_uint(h.t), h.material_id);
        const bool end = path_bounce(p, h, ray, tp, L, rng, depth); // steps 3.2-3.8
        if (end) { // step 4
            sum = sum + L; wsum = wsum + 1.0f; s++;
            if (s == s_end) {
                const size_t pix = (size_t)py * p.width + px;
                if (p.outputs & RT_OUT_ACCUM) p.accum[pix] = make_float4(sum.x, sum.y, sum.z, wsum);
                if (p.outputs & RT_OUT_COMPOSITE)
                    p.composite[pix] = make_uchar4((unsigned char)unorm8(sum.x / wsum), (unsigned char)unorm8(sum.y / wsum),
                                                   (unsigned char)unorm8(sum.z / wsum), 255);
                have = false;
            } else new_sample = true;
        }
    }
    if (p.outputs & RT_OUT_PEER) __threadfence_system(); // composite stores went to another GPU's frame
    // one atomic per warp for the segment count
    for (int o = 16; o; o >>= 1) n_seg += __shfl_xor_sync(full, n_seg, o);
    if (lane == 0u && n_seg) atomicAdd(p.seg_counter, (unsigned long long)n_seg);
    if (COUNT) {
        for (int o = 16; o; o >>= 1) {
            c_box += __shfl_xor_sync(full, c_box, o); c_tri += __shfl_xor_sync(full, c_tri, o);
            c_shaded += __shfl_xor_sync(full, c_shaded, o); c_fb += __shfl_xor_sync(full, c_fb, o);
        }
        if (lane == 0u) {
            atomicAdd(p.counters + 0, c_box); atomicAdd(p.counters + 1, c_tri);
            atomicAdd(p.counters + 2, (unsigned long long)n_seg); atomicAdd(p.counters + 3, c_shaded);
            atomicAdd(p.counters + 4, c_fb);
        }
    }
}
