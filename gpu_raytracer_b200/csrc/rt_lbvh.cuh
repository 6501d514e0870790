// rt_lbvh.cuh — on-device rebuild of a small-leaf BVH over the uploaded triangle
// positions (REBUILT traversal policy).
//
// Why: the reference's builder for > 100 000 triangles (src/bvh.rs:154-247) emits
// leaves of 100-1000 consecutive input triangles; traversing that tree costs
// hundreds of triangle tests per ray.  A drop-in caller still uploads that tree,
// so at upload this file builds an internal LBVH (Morton order + Karras radix
// tree, leaves of <= RT_LBVH_LEAF triangles) over the same triangle positions.
// The kernel traverses the internal tree and then proves the winner is one the
// reference traversal would also have tested (see closest_triangle_rebuilt in
// rt_kernels.cuh); otherwise that ray re-runs the literal policy.
#pragma once
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef RT_LBVH_LEAF
#define RT_LBVH_LEAF 2u
#endif
#ifndef RT_LBVH_FLAT
#define RT_LBVH_FLAT 0.125f // an axis whose centroid extent is below this fraction of the largest one is left out of the Morton code (0: never)
#endif
// REBUILT leaf ref: bit31 | (count-1) << 28 | first sorted slot (28 bits)
#define RT_LBVH_MAX_PRIMS (1u << 28)
// rays starting further than this many scene magnitudes from the coordinate origin take the literal policy
#define RT_REBUILT_FAR 16.0f

namespace lbvh {

__device__ __forceinline__ unsigned int f2ord(float f) { // order-preserving float -> uint
    unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned int u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}
__device__ __forceinline__ unsigned long long expand21(unsigned long long v) { // 21 bits -> every third bit
    v &= 0x1FFFFFull;
    v = (v | v << 32) & 0x1F00000000FFFFull;
    v = (v | v << 16) & 0x1F0000FF0000FFull;
    v = (v | v << 8) & 0x100F00F00F00F00Full;
    v = (v | v << 4) & 0x10C30C30C30C30C3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}

// bounds[0..2] = ordered min of centroids, bounds[3..5] = ordered max
__global__ void centroid_bounds(const float4* __restrict__ pbox, uint32_t n, unsigned int* __restrict__ bounds) {
    // grid-stride over the boxes, warp shuffle + shared-memory fold per block, then 6 atomics per BLOCK: with one
    // warp's worth of atomics per 32 triangles the six addresses serialised the whole kernel (125 us for 1 M
    // triangles in the ncu launch list, a fifth of the upload-time build)
    __shared__ unsigned int s_mn[3][8], s_mx[3][8];
    unsigned int mn[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, mx[3] = {0u, 0u, 0u};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float4 lo = pbox[(size_t)i * 2], hi = pbox[(size_t)i * 2 + 1];
        const float c[3] = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z)};
        for (int k = 0; k < 3; k++) { unsigned int o = f2ord(c[k]); mn[k] = min(mn[k], o); mx[k] = max(mx[k], o); }
    }
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o; o >>= 1) {
            mn[k] = min(mn[k], __shfl_xor_sync(0xFFFFFFFFu, mn[k], o));
            mx[k] = max(mx[k], __shfl_xor_sync(0xFFFFFFFFu, mx[k], o));
        }
        if (lane == 0u) { s_mn[k][warp] = mn[k]; s_mx[k][warp] = mx[k]; }
    }
    __syncthreads();
    if (threadIdx.x < 3u) {
        const int k = (int)threadIdx.x;
        unsigned int a = 0xFFFFFFFFu, b = 0u;
        for (uint32_t w = 0; w < (blockDim.x >> 5); w++) { a = min(a, s_mn[k][w]); b = max(b, s_mx[k][w]); }
        atomicMin(bounds + k, a);
        atomicMax(bounds + 3 + k, b);
    }
}

__global__ void morton_keys(const float4* __restrict__ pbox, uint32_t n, const unsigned int* __restrict__ bounds,
                            unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 lo = pbox[(size_t)i * 2], hi = pbox[(size_t)i * 2 + 1];
    float c[3] = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z)};
    unsigned long long q[3];
    // One scale for all three axes (the largest centroid extent): the Morton cells stay cubes, so a
    // flat scene (terrain: y extent << x, z extents) is split in x/z only until the cells are as
    // small as its height range.  Per-axis normalisation would spend every third split on height
    // slices that overlap completely in x/z (measured on C3: 92.6 -> see DESIGN.md box tests/ray).
    float ext = 0.0f, ek[3];
    for (int k = 0; k < 3; k++) { ek[k] = ord2f(bounds[3 + k]) - ord2f(bounds[k]); ext = fmaxf(ext, ek[k]); }
    uint32_t axes = 0u; // axes that take part in the code
    for (int k = 0; k < 3; k++) {
        float mn = ord2f(bounds[k]);
        float u = ext > 0.0f ? (c[k] - mn) / ext : 0.0f;
        u = fminf(fmaxf(u, 0.0f), 1.0f);
        q[k] = (unsigned long long)min(2097151.0f, u * 2097152.0f);
        if (ek[k] >= ext * RT_LBVH_FLAT) axes |= 1u << k;
    }
    if (axes == 7u) keys[i] = expand21(q[0]) << 2 | expand21(q[1]) << 1 | expand21(q[2]);
    else {
        // A sheet-like scene (terrain: the height range is a small fraction of the ground extents) is a 2-D
        // manifold: inside a Morton cell of side s the surface spans much less than s in the thin axis, so a
        // split across that axis leaves two children with the same footprint (SAH: worse than halving the
        // footprint as long as the cell is wider than its content is tall).  Such axes are left out of the
        // code altogether; the remaining ones interleave as before.  Oracle-counted on the terrain scenes:
        // 7 % fewer box tests per ray at the same triangle tests (scripts/tree_study.py).
        unsigned long long key = 0ull;
        for (int b = 20; b >= 0; b--)
            for (int k = 0; k < 3; k++)
                if (axes >> k & 1u) key = key << 1 | (q[k] >> b & 1ull);
        keys[i] = key;
    }
    vals[i] = i;
}

// Karras 2012 radix tree over sorted keys; equal keys are split by index.
__device__ __forceinline__ int delta(const unsigned long long* __restrict__ k, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    unsigned long long a = k[i], b = k[j];
    if (a == b) return 64 + __clz((unsigned int)i ^ (unsigned int)j);
    return __clzll((long long)(a ^ b));
}
// child encoding inside the builder: bit31 set = Karras leaf (sorted slot), else internal index
__global__ void build_radix_tree(const unsigned long long* __restrict__ keys, uint32_t n, uint2* __restrict__ children,
                                 uint2* __restrict__ range, uint32_t* __restrict__ parent_internal,
                                 uint32_t* __restrict__ parent_leaf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int ni = (int)n - 1;
    if (i >= ni) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    uint32_t left = (lo == gamma) ? (0x80000000u | (uint32_t)gamma) : (uint32_t)gamma;
    uint32_t right = (hi == gamma + 1) ? (0x80000000u | (uint32_t)(gamma + 1)) : (uint32_t)(gamma + 1);
    children[i] = make_uint2(left, right);
    range[i] = make_uint2((uint32_t)lo, (uint32_t)hi);
    if (left & 0x80000000u) parent_leaf[gamma] = (uint32_t)i; else parent_internal[gamma] = (uint32_t)i;
    if (right & 0x80000000u) parent_leaf[gamma + 1] = (uint32_t)i; else parent_internal[gamma + 1] = (uint32_t)i;
    if (i == 0) parent_internal[0] = 0xFFFFFFFFu;
}

// bottom-up fit: the second thread to reach a node computes its box.
// ROTATE: before it moves on, the same thread tries the four tree rotations at that node (Kensler 2008): swap one
// child K with a grandchild G under the other child O when that shrinks O's box area (O = {K, sibling of G}).  The
// whole subtree below the node is finished at that point and no thread above has started, so the rewrite is race-free.
// Subtrees that will be collapsed into leaves (<= RT_LBVH_LEAF triangles: contiguous sorted slots) only move as a
// whole, and a rotation that would leave O with <= RT_LBVH_LEAF triangles (a leaf of non-contiguous slots) is not
// taken.  After a rotation range[O] only carries O's triangle COUNT (y - x + 1); its x is meaningless, and it is
// only ever read for nodes of <= RT_LBVH_LEAF triangles, which are never rewritten.  Measured on a scene that is not a
// height field (1 M triangles of tori, scripts/tree_study.py / bench --workload field): 5 % fewer box tests per ray;
// nothing to gain on the terrains (x/z Morton code).
__device__ __forceinline__ float box_area(float4 lo, float4 hi) {
    float dx = fmaxf(hi.x - lo.x, 0.0f), dy = fmaxf(hi.y - lo.y, 0.0f), dz = fmaxf(hi.z - lo.z, 0.0f);
    return dx * dy + dy * dz + dz * dx;
}
__global__ void fit_boxes(const float4* __restrict__ pbox, const uint32_t* __restrict__ sorted_pos, uint32_t n,
                          uint2* children, uint2* range, uint32_t* parent_internal, uint32_t* parent_leaf,
                          float4* __restrict__ nbox, unsigned int* __restrict__ visits, int rotate) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n || n < 2) return;
    auto load_box = [&](uint32_t c, float4& lo, float4& hi) {
        const float4* s = (c & 0x80000000u) ? pbox + (size_t)sorted_pos[c & 0x7FFFFFFFu] * 2 : nbox + (size_t)c * 2;
        lo = __ldcg(s); hi = __ldcg(s + 1); // plain L2 loads: nbox entries were written by other threads in this launch
    };
    auto count_of = [&](uint32_t c) -> uint32_t {
        if (c & 0x80000000u) return 1u;
        uint2 r = __ldcg(range + c);
        return r.y - r.x + 1u;
    };
    uint32_t node = parent_leaf[k];
    while (node != 0xFFFFFFFFu) {
        __threadfence();
        if (atomicAdd(visits + node, 1u) == 0u) return; // first arrival waits for the sibling
        uint2 ch = __ldcg(children + node);
        float4 alo, ahi, blo, bhi;
        load_box(ch.x, alo, ahi);
        load_box(ch.y, blo, bhi);
        __stcg(nbox + (size_t)node * 2, make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.0f));
        __stcg(nbox + (size_t)node * 2 + 1, make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.0f));
        if (rotate) {
            float best = 0.0f;
            int best_side = -1, best_gs = 0;
            float4 best_lo = alo, best_hi = ahi;
            for (int side = 0; side < 2; side++) {
                const uint32_t keep = side == 0 ? ch.x : ch.y, other = side == 0 ? ch.y : ch.x;
                if ((other & 0x80000000u) || count_of(other) <= RT_LBVH_LEAF) continue; // O is (or becomes) a leaf
                const float4 klo = side == 0 ? alo : blo, khi = side == 0 ? ahi : bhi;
                const float cur = box_area(side == 0 ? blo : alo, side == 0 ? bhi : ahi);
                const uint2 och = __ldcg(children + other);
                const uint32_t kc = count_of(keep);
                for (int gs = 0; gs < 2; gs++) {
                    const uint32_t stay = gs == 0 ? och.y : och.x;
                    if (kc + count_of(stay) <= RT_LBVH_LEAF) continue; // O would collapse into a non-contiguous leaf
                    float4 slo, shi;
                    load_box(stay, slo, shi);
                    const float4 nlo = make_float4(fminf(klo.x, slo.x), fminf(klo.y, slo.y), fminf(klo.z, slo.z), 0.0f);
                    const float4 nhi = make_float4(fmaxf(khi.x, shi.x), fmaxf(khi.y, shi.y), fmaxf(khi.z, shi.z), 0.0f);
                    const float d = box_area(nlo, nhi) - cur;
                    if (d < best) { best = d; best_side = side; best_gs = gs; best_lo = nlo; best_hi = nhi; }
                }
            }
            if (best_side >= 0) {
                const uint32_t keep = best_side == 0 ? ch.x : ch.y, other = best_side == 0 ? ch.y : ch.x;
                uint2 och = __ldcg(children + other);
                const uint32_t gc = best_gs == 0 ? och.x : och.y, stay = best_gs == 0 ? och.y : och.x;
                if (best_gs == 0) och.x = keep; else och.y = keep;
                if (best_side == 0) ch.x = gc; else ch.y = gc;
                __stcg(children + other, och);
                __stcg(children + node, ch);
                __stcg(range + other, make_uint2(0u, count_of(keep) + count_of(stay) - 1u)); // count only (see above)
                __stcg(nbox + (size_t)other * 2, best_lo);
                __stcg(nbox + (size_t)other * 2 + 1, best_hi);
                if (keep & 0x80000000u) parent_leaf[keep & 0x7FFFFFFFu] = other; else parent_internal[keep] = other;
                if (gc & 0x80000000u) parent_leaf[gc & 0x7FFFFFFFu] = node; else parent_internal[gc] = node;
            }
        }
        node = parent_internal[node];
    }
}

// depth of the radix tree (stack bound of the traversal): every leaf walks to the root
__global__ void tree_depth(const uint32_t* __restrict__ parent_internal, const uint32_t* __restrict__ parent_leaf, uint32_t n,
                           unsigned int* __restrict__ out_max) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int d = 0;
    if (k < n && n >= 2) {
        uint32_t node = parent_leaf[k];
        while (node != 0xFFFFFFFFu) { d++; node = parent_internal[node]; }
    }
    for (int o = 16; o; o >>= 1) d = max(d, __shfl_xor_sync(0xFFFFFFFFu, d, o));
    if ((threadIdx.x & 31u) == 0 && d) atomicMax(out_max, d);
}

// Conservative boxes.  A Möller–Trumbore hit implies that the true ray passes within a few ulps of the triangle,
// where the ulp is that of the LARGEST coordinate the test handles: the scene's coordinates and the ray origin
// (s = o - v0, the cross and dot products over it), not the box's own.  Every box is therefore padded by the same
// scene-wide amount, 2^-16 of the scene magnitude M = max |coordinate| of the root box, and rays whose origin lies
// further than RT_REBUILT_FAR * M from the coordinate origin do not use this tree at all (they take the literal
// policy; see closest_triangle_rebuilt / wv_extend): inside that range the rounding of the exact test (a few
// 2^-24 * 17 M) and of the FMA slab test (2^-23 * 16 M) stay below 2^-18 M, a quarter of the pad.
__host__ __device__ __forceinline__ float scene_pad(float4 lo, float4 hi) {
    float m = fmaxf(fmaxf(fmaxf(fabsf(lo.x), fabsf(hi.x)), fmaxf(fabsf(lo.y), fabsf(hi.y))), fmaxf(fabsf(lo.z), fabsf(hi.z)));
    return m * 1.52587890625e-05f + 1e-30f;
}
__device__ __forceinline__ void inflate(float4& lo, float4& hi, float pad) {
    lo.x -= pad; lo.y -= pad; lo.z -= pad; hi.x += pad; hi.y += pad; hi.z += pad;
}

// 1 for every radix-tree node that becomes a traversal node (covers more than RT_LBVH_LEAF triangles); an
// exclusive scan of these flags gives the compact index: the emitted nodes are stored densely, in radix-tree
// (= Morton) order, instead of leaving the collapsed subtrees' slots as holes between them (with 2-triangle
// leaves about half of the slots), which halves the node footprint in L2 / L1.
__global__ void mark_emitted(const uint2* __restrict__ range, uint32_t n, uint32_t* __restrict__ flags) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t f = 0u;
    if (n >= 2 && i < n - 1) { uint2 r = range[i]; f = (r.y - r.x + 1u > RT_LBVH_LEAF) ? 1u : 0u; }
    flags[i] = f;
}

__device__ __forceinline__ uint32_t emit_ref(uint32_t child, const uint2* __restrict__ range, const uint32_t* __restrict__ compact) {
    if (child & 0x80000000u) return 0x80000000u | (child & 0x7FFFFFFFu); // single triangle: count-1 = 0
    uint2 r = range[child];
    uint32_t cnt = r.y - r.x + 1u;
    if (cnt <= RT_LBVH_LEAF) return 0x80000000u | ((cnt - 1u) << 28) | r.x; // collapsed subtree
    return compact[child];
}

// two-child 64-byte nodes in the layout rt_kernels.cuh traverses
__global__ void emit_wide(const float4* __restrict__ pbox, const uint32_t* __restrict__ sorted_pos, uint32_t n,
                          const uint2* __restrict__ children, const uint2* __restrict__ range,
                          const float4* __restrict__ nbox, const uint32_t* __restrict__ compact,
                          const float* __restrict__ grid, float4* __restrict__ wnodes) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < 2 || i >= n - 1) return;
    uint2 r = range[i];
    if (r.y - r.x + 1u <= RT_LBVH_LEAF) return; // inside / is a collapsed leaf: never visited
    uint2 ch = children[i];
    const float pad = grid[6];
    float4 lo[2], hi[2];
    uint32_t c[2] = {ch.x, ch.y};
    for (int k = 0; k < 2; k++) {
        const float4* s = (c[k] & 0x80000000u) ? pbox + (size_t)sorted_pos[c[k] & 0x7FFFFFFFu] * 2 : nbox + (size_t)c[k] * 2;
        lo[k] = s[0]; hi[k] = s[1];
        inflate(lo[k], hi[k], pad);
    }
    float4* q = wnodes + (size_t)compact[i] * 4;
    q[0] = make_float4(lo[0].x, lo[0].y, lo[0].z, hi[0].x);
    q[1] = make_float4(hi[0].y, hi[0].z, lo[1].x, lo[1].y);
    q[2] = make_float4(lo[1].z, hi[1].x, hi[1].y, hi[1].z);
    q[3] = make_float4(__uint_as_float(emit_ref(ch.x, range, compact)), __uint_as_float(emit_ref(ch.y, range, compact)), 0.0f, 0.0f);
}

// ---- 32-byte nodes with 16-bit boxes (the wave engine's extend kernel) ------------------------------------
// ncu on wv_extend: the SM's L1 data stage is the busiest unit and every lane of a warp fetches a different
// node, so a node should be ONE 32-byte sector instead of two.  Both child boxes are stored on a fixed grid of
// 65000 cells per axis over the (inflated) root box: word a of a child = min_a | max_a << 16.  The boxes only
// have to be conservative (the winner is certified against the uploaded tree afterwards): each plane is the
// fp32 plane after lbvh::inflate, moved outwards to the next grid line and one more cell.  The first cell
// covers the directed rounding of the division, the second the rounding of the ray-side constant
// B2 = (O - o) / d - 2^23 S / d (half a cell, see wv_extend), the world-space pad everything the 64-byte
// layout's pad covers.
// grid[0..2] = origin O, grid[3..5] = cell size S, grid[6] = the scene-wide pad (scene_pad of the root box)
__global__ void quant_grid(const float4* __restrict__ root_box, float* __restrict__ grid) {
    float4 lo = root_box[0], hi = root_box[1];
    const float pad = scene_pad(lo, hi);
    grid[6] = pad;
    inflate(lo, hi, pad);
    const float l[3] = {lo.x, lo.y, lo.z}, h[3] = {hi.x, hi.y, hi.z};
    for (int k = 0; k < 3; k++) {
        float s = fmaxf((h[k] - l[k]) / 65000.0f, 1e-30f);
        grid[3 + k] = s;
        grid[k] = l[k] - 8.0f * s;
    }
}
__device__ __forceinline__ uint32_t quant_pair(float lo, float hi, float O, float S) {
    float ql = floorf((lo - O) / S) - 1.0f, qh = ceilf((hi - O) / S) + 1.0f;
    ql = fminf(fmaxf(ql, 0.0f), 65535.0f);
    qh = fminf(fmaxf(qh, 0.0f), 65535.0f);
    return (uint32_t)ql | ((uint32_t)qh << 16);
}
__global__ void emit_quant(const float4* __restrict__ pbox, const uint32_t* __restrict__ sorted_pos, uint32_t n,
                           const uint2* __restrict__ children, const uint2* __restrict__ range,
                           const float4* __restrict__ nbox, const uint32_t* __restrict__ compact,
                           const float* __restrict__ grid, uint4* __restrict__ qnodes) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < 2 || i >= n - 1) return;
    uint2 r = range[i];
    if (r.y - r.x + 1u <= RT_LBVH_LEAF) return;
    uint2 ch = children[i];
    const float pad = grid[6];
    uint32_t c[2] = {ch.x, ch.y}, w[6];
    for (int k = 0; k < 2; k++) {
        const float4* s = (c[k] & 0x80000000u) ? pbox + (size_t)sorted_pos[c[k] & 0x7FFFFFFFu] * 2 : nbox + (size_t)c[k] * 2;
        float4 lo = s[0], hi = s[1];
        inflate(lo, hi, pad);
        w[k * 3] = quant_pair(lo.x, hi.x, grid[0], grid[3]);
        w[k * 3 + 1] = quant_pair(lo.y, hi.y, grid[1], grid[4]);
        w[k * 3 + 2] = quant_pair(lo.z, hi.z, grid[2], grid[5]);
    }
    uint4* q = qnodes + (size_t)compact[i] * 2;
    q[0] = make_uint4(w[0], w[1], w[2], w[3]);
    q[1] = make_uint4(w[4], w[5], emit_ref(ch.x, range, compact), emit_ref(ch.y, range, compact));
}

// triangle records in sorted order; the last word carries the triangle_indices position
__global__ void gather_sorted(const float4* __restrict__ wtris, const uint32_t* __restrict__ sorted_pos, uint32_t n,
                              const uint32_t* __restrict__ leaf_of_pos, float4* __restrict__ out) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint32_t pos = sorted_pos[k];
    float4 a = wtris[(size_t)pos * 3], b = wtris[(size_t)pos * 3 + 1], c = wtris[(size_t)pos * 3 + 2];
    c.w = __uint_as_float(pos);
    if (leaf_of_pos[pos] == 0xFFFFFFFFu) { // not in any reachable leaf: the reference never tests it
        a.w = 0.0f; b = make_float4(0.0f, 0.0f, 0.0f, 0.0f); c.x = 0.0f;
    }
    out[(size_t)k * 3] = a; out[(size_t)k * 3 + 1] = b; out[(size_t)k * 3 + 2] = c;
}

} // namespace lbvh
