// rt_wave.cuh — EXTENSION (SPEC.md) path tracing, streaming wavefront engine ("wave").
//
// Same per-path arithmetic and bit-identical results as rt_path_trace (rt_kernels.cuh);
// the execution shape is chosen from what ncu showed on the megakernel (profiles/r1h_*):
// 10 of 32 lanes active per issued instruction (triangle tests 3.6 lanes, stack pops 5.7),
// and more L1 data-stage wavefronts spent on the local-memory traversal stack than on the
// BVH itself.  Here
//   * all samples of a wave are in flight at once as compact path records (64 B state +
//     8 B hit), ping-ponged by depth, so there is one extend + one shade launch per depth
//     for the whole frame instead of per sample, and the extend kernel's idle lanes are
//     refilled by loading 40 bytes instead of running the shading code;
//   * wv_extend keeps box tests and Möller–Trumbore tests in separate, densely populated
//     instruction streams: a lane that reaches a leaf postpones it and (speculatively) keeps
//     descending until it holds a second leaf; the warp leaves the node loop when too few
//     lanes can still descend, then all lanes test their postponed triangles together;
//   * the traversal stack is a per-thread column of shared memory (entry k of the 32 lanes
//     of a warp are 32 consecutive words: conflict-free whatever the lanes' depths are);
//   * the REBUILT certification (one exact slab test on the winner's uploaded leaf box)
//     moves to wv_shade, which runs convergent over compacted paths;
//   * sample radiances go to Lbuf[sample][pixel] and wv_resolve adds them in sample order,
//     so fp32 accumulation stays bit-identical to the sequential definition (SPEC.md step 4).
#pragma once
#include <assert.h>
#include "rt_kernels.cuh"

// -DRT_CHECKED builds (make tune NAME=checked FLAGS=-DRT_CHECKED; scripts/sanitize.sh): device-side bounds checks on
// every index this engine computes into its queues, stacks and per-warp lists.  compute-sanitizer is closed on this
// pool, so the parity subset is run against this build instead; a violated check traps the kernel.
#ifdef RT_CHECKED
#define WV_CHECK(cond) assert(cond)
#else
#define WV_CHECK(cond) ((void)0)
#endif

struct WvCtrl {
    unsigned int n[2];          // live paths at the current / next depth (ping-pong)
    unsigned int head;          // wv_extend work cursor
    unsigned int _pad;
    unsigned long long segments;
    unsigned long long c[4];    // instrumentation: box tests, triangle tests, shaded hits, literal fallbacks
    unsigned int misses[2];     // rays of the current / next depth that hit nothing (picks wv_shade's form)
    unsigned int _pad2[2];
    unsigned int n_fb;          // paths of the current depth handed to wv_fallback (the reference's own traversal)
    unsigned int _pad3[3];
    unsigned int rays_at_depth[8]; // ray segments traced per bounce depth (depths >= 7 share the last entry): the
                                   // reference's WavefrontCounters.rays_per_bounce (shared/src/lib.rs:187-194)
};

struct WvBuffers {
    float4* state[2];  // 4 x float4 per path: (o, rng) (d, depth) (throughput, work idx) (L, local sample)
    uint2* hit[2];     // per path: (t bits, id); in: sphere seed, out: closest hit
    float4* lbuf;      // [local sample][work idx]: finished sample radiance
    uint32_t* fb_list; // queue positions of the current depth that wv_shade left to wv_fallback
    WvCtrl* ctrl;
    uint32_t s0, spp_wave; // samples [s0, s0 + spp_wave) are in this wave
    uint32_t last_wave;    // resolve writes the composite
    uint32_t first_wave;   // resolve honours reset_accum
    uint32_t valid_pixels; // work items of this sub-frame that lie inside the image (fused generation: segment count of depth 0)
    uint32_t depth;        // bounce depth of this launch (bookkeeping only: rays_at_depth)
    uint32_t capacity;     // paths this sub-frame's queues / lists hold (RT_CHECKED bounds checks)
    uint32_t split_quarters; // wv_shade takes its split form when misses * 4 > rays * split_quarters (>= 5: never)
};

// hit id encoding
#define WV_ID_MISS 0xFFFFFFFFu
#define WV_KIND_TRI 0u      // id = slot in btris / wtris
#define WV_KIND_SPHERE 1u   // id = 1<<30 | sphere index
#define WV_KIND_LITERAL 2u  // id = 2<<30 | (sphere index or 0x3FFFFFFF): wv_shade runs the literal policy
#define WV_NO_SPHERE 0x3FFFFFFFu

#ifndef WV_STK
#define WV_STK 32          // shared-memory stack entries per thread (32 KB per 256-thread block).  Trees deeper than
#endif                     // WV_STK - 2 levels run the DEEP instantiation, whose further entries live in local memory
#ifndef WV_REFILL
#define WV_REFILL 10       // refill when this many lanes have finished their ray
#endif
#ifndef WV_NODE_MIN
#define WV_NODE_MIN 20     // leave the node loop when fewer lanes than this can still descend
#endif
#ifndef WV_TRI_AGAIN
#define WV_TRI_AGAIN 16    // repeat the triangle phase while this many lanes hold a second leaf
#endif
#ifndef WV_SPEC
#define WV_SPEC 1          // speculative descent past the first postponed leaf
#endif
#ifndef WV_L2_KEEP
#define WV_L2_KEEP 0       // 1: BVH loads carry an L2 evict_last policy (measured: no effect on C3 / C4, kept for HBM-resident scenes)
#endif
#ifndef WV_CS_STORES
#define WV_CS_STORES 0     // 1: path-state stores with the evict-first (.cs) hint (measured: C3 +1 %, C4 -7 %)
#endif
#if WV_CS_STORES
#define WV_ST(ptr, val) __stcs((ptr), (val))
#else
#define WV_ST(ptr, val) (*(ptr) = (val))
#endif
#ifndef WV_UNROLL
#define WV_UNROLL 2        // node steps per loop-exit vote (measured on C3: 1 -> 3864, 2 -> 4004, 3 -> 3886, 4 -> 3934 Mrays/s)
#endif
#ifndef WV_QUANT
#define WV_QUANT 1         // 1: REBUILT policy traverses the 32-byte nodes with 16-bit boxes (one sector per node visit)
#endif
#ifndef WV_FUSE_GEN
#define WV_FUSE_GEN 1      // 1: no wv_generate launch; the depth-0 extend / shade compute the primary ray from the queue position
#endif
#ifndef WV_SHADE_BLOCKS
#define WV_SHADE_BLOCKS 3  // resident blocks per SM the shade kernel is compiled for (4: 64 registers, some spills; measured +-0 alone)
#endif
#ifndef WV_BLOCKS
#define WV_BLOCKS 4        // resident 256-thread blocks per SM the extend kernel is compiled for
#endif

// spheres first (shader/src/lib.rs:183-190, 252-269): returns the seed t and the sphere index
__device__ __forceinline__ void wv_sphere_seed(const KParams& p, const Ray& ray, float* closest, uint32_t* sphere_idx) {
    float c = FLT_MAX - 2.0f;
    uint32_t si = WV_NO_SPHERE;
    for (uint32_t i = 0; i < p.off.spheres_count; i++) {
        uint64_t o = (uint64_t)p.off.spheres_offset + (uint64_t)i * 5;
        float t;
        if (sphere_t(ray, mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2)), meta_f(p, o + 3), c, &t)) { c = t; si = i; }
    }
    *closest = c; *sphere_idx = si;
}
__device__ __forceinline__ uint2 wv_seed_record(const KParams& p, const Ray& ray) {
    if (p.off.spheres_count == 0u) return make_uint2(__float_as_uint(FLT_MAX - 2.0f), WV_ID_MISS);
    float c; uint32_t si;
    wv_sphere_seed(p, ray, &c, &si);
    return make_uint2(__float_as_uint(c), si == WV_NO_SPHERE ? WV_ID_MISS : ((WV_KIND_SPHERE << 30) | si));
}

// Position q of a wave's primary-ray list -> work item, local sample, camera ray and the rng state after the
// jitter draws (SPEC.md steps 1-2).  Warp-sized runs of q are one 8x4 pixel group for one sample, consecutive
// runs the samples of that group.  Used by wv_generate, or (WV_FUSE_GEN) by the depth-0 extend and shade
// launches themselves, which then never materialise the 64-byte primary path records.
__device__ __forceinline__ bool wv_primary_index(const KParams& p, const WvBuffers& w, uint32_t q, uint32_t* idx, uint32_t* sl,
                                                 uint32_t* px, uint32_t* py) {
    const uint32_t wi = q >> 5, group = wi / w.spp_wave;
    *sl = wi - group * w.spp_wave;
    *idx = group * 32u + (q & 31u);
    return work_to_pixel(p, *idx, px, py);
}
__device__ __forceinline__ bool wv_primary(const KParams& p, const WvBuffers& w, uint32_t q, uint32_t* idx, uint32_t* sl,
                                           Ray* ray, uint32_t* rng_seed) {
    uint32_t px = 0, py = 0;
    if (!wv_primary_index(p, w, q, idx, sl, &px, &py)) return false;
    const uint32_t s = w.s0 + *sl;
    SimpleRng rng; rng.seed = p.base_seed + s + px + py * p.width;
    float jx = 0.5f, jy = 0.5f;
    if (s != 0u) { jx = rng.next_f32(); jy = rng.next_f32(); }
    *ray = camera_ray(p, px, py, jx, jy, p.legacy_camera == 0u);
    *rng_seed = rng.seed;
    return true;
}

// ---- wave start: SPEC.md steps 1-2 for every (pixel, sample) of this wave ---------------
// Warp = one 8x4 pixel group for one sample; consecutive warps = the samples of that group.
__global__ void __launch_bounds__(256) wv_generate(const __grid_constant__ KParams p, WvBuffers w) {
    const unsigned full = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t total_warps = (uint64_t)(p.n_work >> 5) * w.spp_wave;
    for (uint64_t wi = (uint64_t)blockIdx.x * 8u + (threadIdx.x >> 5); wi < total_warps; wi += (uint64_t)gridDim.x * 8u) {
        uint32_t group = (uint32_t)(wi / w.spp_wave), sl = (uint32_t)(wi % w.spp_wave);
        uint32_t idx = group * 32u + lane, px = 0, py = 0;
        bool ok = work_to_pixel(p, idx, &px, &py);
        unsigned m = __ballot_sync(full, ok);
        if (!m) continue;
        uint32_t base = 0;
        int leader = __ffs(m) - 1;
        if ((int)lane == leader) base = atomicAdd(&w.ctrl->n[0], (unsigned int)__popc(m));
        base = __shfl_sync(full, base, leader);
        if (!ok) continue;
        uint32_t slot = base + __popc(m & ((1u << lane) - 1u));
        uint32_t s = w.s0 + sl;
        SimpleRng rng; rng.seed = p.base_seed + s + px + py * p.width;
        float jx = 0.5f, jy = 0.5f;
        if (s != 0u) { jx = rng.next_f32(); jy = rng.next_f32(); }
        Ray ray = camera_ray(p, px, py, jx, jy, true);
        float4* st = w.state[0] + (size_t)slot * 4;
        WV_ST(st, make_float4(ray.o.x, ray.o.y, ray.o.z, __uint_as_float(rng.seed)));
        WV_ST(st + 1, make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(0u)));
        WV_ST(st + 2, make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(idx)));
        WV_ST(st + 3, make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(sl)));
        WV_ST(w.hit[0] + slot, wv_seed_record(p, ray));
    }
}

// ---- extend: closest triangle of every ray of the current depth ---------------------------
// GEN: depth 0 of a wave with fused generation; queue position q is primary ray q (wv_primary)
// DEEP: the tree may be deeper than the shared-memory stack (otherwise every stack access is one predicated
// shared-memory instruction and the node step has no branches)
template <int POLICY, bool COUNT, bool GEN, bool DEEP> // POLICY 1 (ORDERED, uploaded tree) or 2 (REBUILT, internal LBVH)
__global__ void __launch_bounds__(256, WV_BLOCKS) wv_extend(const __grid_constant__ KParams p, WvBuffers w, int qi) {
    __shared__ uint32_t s_stack[WV_STK][256];
    uint32_t lstack[DEEP ? 64 - WV_STK : 1];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const unsigned full = 0xFFFFFFFFu;
    constexpr uint32_t EMPTY = 0xFFFFFFFFu; // has RT_LEAF_FLAG set: never an internal node
    const float4* __restrict__ nodes = POLICY == 2 ? p.bnodes : p.wnodes;
    const float4* __restrict__ tris = POLICY == 2 ? p.btris : p.wtris;
    constexpr bool QUANT = POLICY == 2 && WV_QUANT;
    // QUANT: plane q of axis a lies at t = q * A_a + B_a with A = S / d, B = (O - o) / d.  The 16-bit q is turned
    // into the float 2^23 + q by one byte permute (0x4B00'qqqq), so t = fma(2^23 + q, A, B - 2^23 A); the same
    // permute picks the near or the far plane of the axis by the sign of d (selectors sn / sn ^ 0x22), which
    // replaces the min / max of the two plane distances.  ix.. hold A, kx.. hold B - 2^23 A.
    uint32_t snx = 0x7410u, sny = 0x7410u, snz = 0x7410u;
    const uint32_t root_ref = POLICY == 2 ? p.broot_ref : p.root_ref;
    const uint32_t n = GEN ? p.n_work * w.spp_wave : w.ctrl->n[qi];
    const float4* __restrict__ state = w.state[qi];
    uint2* __restrict__ hit = w.hit[qi];
    if (blockIdx.x == 0 && tid == 0) { w.ctrl->n[qi ^ 1] = 0; w.ctrl->misses[qi ^ 1] = 0; w.ctrl->n_fb = 0; } // wv_shade of this depth appends there
    uint32_t n_miss = 0;
#if WV_L2_KEEP
    const unsigned long long keep = l2_keep_policy();
#endif

    uint32_t slot = EMPTY;      // queue position of the ray this lane traces
    bool exhausted = false, literal = false;
    float ox = 0, oy = 0, oz = 0, dx = 0, dy = 0, dz = 1, ix = 0, iy = 0, iz = 0, kx = 0, ky = 0, kz = 0;
    float best_t = FLT_MAX, limit = FLT_MAX;
    uint32_t best_slot = EMPTY, best_pos = 0u, seed_id = WV_ID_MISS;
    uint32_t cur = EMPTY, pend = EMPTY;
    int sp = 1;
    unsigned long long c_box = 0, c_tri = 0;
    // warp-uniform work-fetch state
    uint32_t chunk_next = 0, chunk_end = 0;
    bool queue_done = n == 0u;
    const uint32_t chunk = n >= gridDim.x * 8u * 128u * 4u ? 128u : 32u;

    // entry 0 of every column holds EMPTY and the stack starts at entry 1: popping the last entry yields EMPTY
    // ("traversal finished") without an emptiness test; a finished lane is never popped again.
    // !DEEP keeps the stack top as a shared-memory pointer (no index arithmetic), DEEP as an index.
    s_stack[0][tid] = EMPTY;
    uint32_t* sptr = &s_stack[1][tid];
    auto push = [&](uint32_t v) {
        if (!DEEP) { *sptr = v; sptr += 256; return; }
        if (sp < WV_STK) s_stack[sp][tid] = v; else lstack[sp - WV_STK] = v;
        sp++;
    };
    auto pop = [&]() -> uint32_t {
        if (!DEEP) { sptr -= 256; WV_CHECK(sptr >= &s_stack[0][tid]); return *sptr; }
        sp--;
        return sp < WV_STK ? s_stack[sp][tid] : lstack[sp - WV_STK];
    };
    auto flush = [&]() { // store the finished ray's record
        if (slot != EMPTY) {
            if (literal) WV_ST(hit + slot, make_uint2(__float_as_uint(best_t), (WV_KIND_LITERAL << 30) | (seed_id == WV_ID_MISS ? WV_NO_SPHERE : (seed_id & WV_NO_SPHERE))));
            else if (best_slot != EMPTY) WV_ST(hit + slot, make_uint2(__float_as_uint(best_t), best_slot));
            else if (GEN) WV_ST(hit + slot, make_uint2(__float_as_uint(best_t), seed_id)); // nobody seeded the record
            if (!literal && best_slot == EMPTY && seed_id == WV_ID_MISS) n_miss++;
            slot = EMPTY;
        }
    };

    for (;;) {
        // ---- refill finished lanes from the queue
        const bool done = cur == EMPTY && pend == EMPTY;
        const unsigned idle = __ballot_sync(full, done && !exhausted);
        const unsigned act = __ballot_sync(full, !done);
        if (idle && (__popc(idle) >= WV_REFILL || act == 0u)) {
            // the warp reserves `chunk` consecutive queue positions at a time: one atomic per chunk
            // instead of one per refill, and consecutive positions are neighbouring rays
            const uint32_t need = (uint32_t)__popc(idle);
            uint32_t rest_base = 0, rest_n = 0;
            if (chunk_end - chunk_next < need && !queue_done) {
                uint32_t nb = 0;
                if (lane == 0u) nb = atomicAdd(&w.ctrl->head, (unsigned int)chunk);
                nb = __shfl_sync(full, nb, 0);
                rest_base = chunk_next; rest_n = chunk_end - chunk_next; // left-overs of the old chunk go first
                if (nb >= n) { chunk_next = chunk_end = 0; queue_done = true; }
                else {
                    chunk_next = nb; chunk_end = nb + chunk;
                    if (chunk_end >= n) { chunk_end = n; queue_done = true; }
                }
            }
            const uint32_t r = (uint32_t)__popc(idle & ((1u << lane) - 1u));
            uint32_t q = EMPTY;
            if (r < rest_n) q = rest_base + r;
            else if (chunk_next + (r - rest_n) < chunk_end) q = chunk_next + (r - rest_n);
            {
                uint32_t take = need > rest_n ? need - rest_n : 0u;
                chunk_next = chunk_next + take < chunk_end ? chunk_next + take : chunk_end;
            }
            if (done && !exhausted) {
                flush();
                if (q == EMPTY) exhausted = queue_done;
                else {
                    bool valid = true;
                    if (GEN) {
                        uint32_t g_idx, g_sl, g_seed;
                        Ray pr;
                        valid = wv_primary(p, w, q, &g_idx, &g_sl, &pr, &g_seed);
                        if (valid) {
                            uint2 seed = wv_seed_record(p, pr);
                            ox = pr.o.x; oy = pr.o.y; oz = pr.o.z; dx = pr.d.x; dy = pr.d.y; dz = pr.d.z;
                            best_t = __uint_as_float(seed.x); seed_id = seed.y;
                        }
                    } else {
                        float4 a = __ldcs(state + (size_t)q * 4), b = __ldcs(state + (size_t)q * 4 + 1);
                        uint2 seed = __ldcs(hit + q);
                        ox = a.x; oy = a.y; oz = a.z; dx = b.x; dy = b.y; dz = b.z;
                        best_t = __uint_as_float(seed.x); seed_id = seed.y;
                    }
                    if (valid) {
                        WV_CHECK(q < n);
                        slot = q;
                        limit = best_t * RT_PRUNE_SCALE + RT_PRUNE_ABS;
                        best_slot = EMPTY; best_pos = 0u; sp = 1; sptr = &s_stack[1][tid]; pend = EMPTY; literal = false;
                        cur = root_ref == RT_CHILD_ABSENT ? EMPTY : root_ref;
                        if (POLICY == 2) {
                            // zero / tiny direction components: neither the certification argument nor the
                            // FMA slab applies (closest_triangle_rebuilt) -> wv_shade runs the literal policy
                            // origins far outside the scene: the exact tests' rounding outgrows the boxes' pad (rt_lbvh.cuh)
                            if (!(fabsf(dx) >= 1e-30f && fabsf(dy) >= 1e-30f && fabsf(dz) >= 1e-30f) ||
                                !(fmaxf(fmaxf(fabsf(ox), fabsf(oy)), fabsf(oz)) <= p.rebuilt_far)) { literal = true; cur = EMPTY; }
                            // boxes of the internal tree are conservative (inflated 2^-16 relative), so
                            // the slab test may use t = b * inv - o * inv with one FMA per plane
                            ix = 1.0f / dx; iy = 1.0f / dy; iz = 1.0f / dz;
                            kx = ox * ix; ky = oy * iy; kz = oz * iz;
                            float e;
                            if (COUNT) c_box++;
                            if (cur != EMPTY && !slab_fma(mk(ix, iy, iz), mk(kx, ky, kz), mka(p.broot_min), mka(p.broot_max), &e)) cur = EMPTY;
                            if (QUANT) {
                                // tiny components: 2^23 A must stay finite
                                if (!(fabsf(dx) >= 1e-18f && fabsf(dy) >= 1e-18f && fabsf(dz) >= 1e-18f)) { literal = true; cur = EMPTY; }
                                const float bx = (p.bq_origin[0] - ox) * ix, by = (p.bq_origin[1] - oy) * iy, bz = (p.bq_origin[2] - oz) * iz;
                                ix = p.bq_scale[0] * ix; iy = p.bq_scale[1] * iy; iz = p.bq_scale[2] * iz;
                                kx = __fmaf_rn(-8388608.0f, ix, bx); ky = __fmaf_rn(-8388608.0f, iy, by); kz = __fmaf_rn(-8388608.0f, iz, bz);
                                snx = dx < 0.0f ? 0x7432u : 0x7410u; sny = dy < 0.0f ? 0x7432u : 0x7410u; snz = dz < 0.0f ? 0x7432u : 0x7410u;
                            }
                        } else {
                            ix = 1.0f / dx; iy = 1.0f / dy; iz = 1.0f / dz; // exact: intersection.rs:152
                            float e;
                            if (COUNT) c_box++;
                            if (cur != EMPTY && !slab(mk(ox, oy, oz), mk(ix, iy, iz), mka(p.root_min), mka(p.root_max), &e)) cur = EMPTY;
                        }
                        if (cur != EMPTY && (cur & RT_LEAF_FLAG)) { pend = cur; cur = EMPTY; } // the root is a leaf
                    } // else: a work item outside the image (edge tiles); the lane stays idle and refills next round
                }
            }
        }
        const unsigned live = __ballot_sync(full, !(cur == EMPTY && pend == EMPTY));
        if (live == 0u) {
            if (__all_sync(full, exhausted)) break;
            continue;
        }
        // ---- node loop: every lane holding an internal node descends; a lane that reaches a leaf
        // postpones it and keeps going until it holds a second one
        const int nl = __popc(live);
        const int node_min = nl - (nl >> 2) < WV_NODE_MIN ? nl - (nl >> 2) : WV_NODE_MIN;
        for (;;) {
#pragma unroll
            for (int u = 0; u < WV_UNROLL; u++)
            if (!(cur & RT_LEAF_FLAG)) {
                float4 q0, q1, q2, q3;
                uint32_t lref, rref;
                float el, er;
                bool hl, hr;
                if (COUNT) c_box += 2;
                if (QUANT) {
                    uint32_t qw[8];
                    ldg_qnode(reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(p.bqnodes) + (uint64_t)cur * 32u), qw);
                    lref = qw[6]; rref = qw[7];
#define WV_QF(word, sel) __uint_as_float(__byte_perm((word), 0x4B000000u, (sel)))
                    el = fmaxf(fmaxf(__fmaf_rn(WV_QF(qw[0], snx), ix, kx), __fmaf_rn(WV_QF(qw[1], sny), iy, ky)), __fmaf_rn(WV_QF(qw[2], snz), iz, kz));
                    er = fmaxf(fmaxf(__fmaf_rn(WV_QF(qw[3], snx), ix, kx), __fmaf_rn(WV_QF(qw[4], sny), iy, ky)), __fmaf_rn(WV_QF(qw[5], snz), iz, kz));
                    const float xl = fminf(fminf(__fmaf_rn(WV_QF(qw[0], snx ^ 0x22u), ix, kx), __fmaf_rn(WV_QF(qw[1], sny ^ 0x22u), iy, ky)), __fmaf_rn(WV_QF(qw[2], snz ^ 0x22u), iz, kz));
                    const float xr = fminf(fminf(__fmaf_rn(WV_QF(qw[3], snx ^ 0x22u), ix, kx), __fmaf_rn(WV_QF(qw[4], sny ^ 0x22u), iy, ky)), __fmaf_rn(WV_QF(qw[5], snz ^ 0x22u), iz, kz));
#undef WV_QF
                    hl = xl >= 0.0f && el <= xl;
                    hr = xr >= 0.0f && er <= xr;
                } else {
                const float4* np = nodes + (uint64_t)cur * 4;
#if WV_L2_KEEP
                ldg_node_keep(np, keep, q0, q1, q2, q3);
#else
                ldg_node(np, q0, q1, q2, q3);
#endif
                lref = __float_as_uint(q3.x); rref = __float_as_uint(q3.y);
                if (POLICY == 2) {
                    hl = slab_fma(mk(ix, iy, iz), mk(kx, ky, kz), mk(q0.x, q0.y, q0.z), mk(q0.w, q1.x, q1.y), &el);
                    hr = slab_fma(mk(ix, iy, iz), mk(kx, ky, kz), mk(q1.z, q1.w, q2.x), mk(q2.y, q2.z, q2.w), &er);
                } else {
                    hl = slab(mk(ox, oy, oz), mk(ix, iy, iz), mk(q0.x, q0.y, q0.z), mk(q0.w, q1.x, q1.y), &el) && lref != RT_CHILD_ABSENT;
                    hr = slab(mk(ox, oy, oz), mk(ix, iy, iz), mk(q1.z, q1.w, q2.x), mk(q2.y, q2.z, q2.w), &er) && rref != RT_CHILD_ABSENT;
                }
                }
                hl = hl && (el <= limit);
                hr = hr && (er <= limit);
                if (!DEEP) { // branch-free step: the stack ops are one predicated shared-memory access each
                    const bool both = hl && hr, any = hl || hr;
                    const bool rf = hr && (!hl || er < el); // the right child is the nearer (or the only) one
                    const uint32_t nearc = rf ? rref : lref, farc = rf ? lref : rref;
                    WV_CHECK(!both || sptr <= &s_stack[WV_STK - 1][tid]);
                    if (both) { *sptr = farc; sptr += 256; }
                    if (!any) sptr -= 256;
                    WV_CHECK(sptr >= &s_stack[0][tid] && sptr <= &s_stack[WV_STK - 1][tid] + 256);
                    const uint32_t top = *sptr; // read only when nothing was hit (predicated)
                    cur = any ? nearc : top;
                } else if (hl && hr) {
                    bool right_first = er < el;
                    push(right_first ? lref : rref);
                    cur = right_first ? rref : lref;
                } else if (hl) cur = lref;
                else if (hr) cur = rref;
                else cur = pop();
                if (WV_SPEC && (cur & RT_LEAF_FLAG) && cur != EMPTY && pend == EMPTY) { pend = cur; cur = pop(); }
            }
            if (__popc(__ballot_sync(full, !(cur & RT_LEAF_FLAG))) < node_min) break;
        }
        // ---- triangle phase: all lanes test their postponed leaf together
        for (;;) {
            uint32_t leaf = pend;
            if (!WV_SPEC && leaf == EMPTY && (cur & RT_LEAF_FLAG) && cur != EMPTY) { leaf = cur; cur = pop(); }
            if (leaf != EMPTY) {
                uint32_t ts = POLICY == 2 ? (leaf & 0x0FFFFFFFu) : (leaf & ~RT_LEAF_FLAG);
                uint32_t rem = POLICY == 2 ? ((leaf >> 28) & 7u) + 1u : 0u;
                Ray ray; ray.o = mk(ox, oy, oz); ray.d = mk(dx, dy, dz);
                for (;;) {
                    const float4* tpn = tris + (uint64_t)ts * 3;
#if WV_L2_KEEP
                    float4 a = ldg_keep(tpn, keep), c = ldg_keep(tpn + 1, keep), d = ldg_keep(tpn + 2, keep);
#else
                    float4 a = __ldg(tpn), c = __ldg(tpn + 1), d = __ldg(tpn + 2);
#endif
                    float t = triangle_t(ray, mk(a.x, a.y, a.z), mk(a.w, c.x, c.y), mk(c.z, c.w, d.x));
                    uint32_t pos = POLICY == 2 ? __float_as_uint(d.w) : ts;
                    if (COUNT) c_tri++;
                    // strict `t < closest` in visit order; with re-ordered visits the equal-t winner is
                    // the smaller triangle_indices position (== first in the reference's DFS order)
                    if (t > 0.0f && (t < best_t || (t == best_t && best_slot != EMPTY && pos < best_pos))) {
                        best_t = t; best_slot = ts; best_pos = pos;
                        limit = best_t * RT_PRUNE_SCALE + RT_PRUNE_ABS;
                    }
                    if (POLICY == 2) { if (--rem == 0u) break; }
                    else if (__float_as_uint(d.w) <= 1u) break;
                    ts++;
                }
                pend = EMPTY;
                if (WV_SPEC && (cur & RT_LEAF_FLAG) && cur != EMPTY) { pend = cur; cur = pop(); }
            }
            if (!WV_SPEC) break;
            if (__popc(__ballot_sync(full, pend != EMPTY)) < WV_TRI_AGAIN) break;
        }
    }
    flush();
    for (int o = 16; o; o >>= 1) n_miss += __shfl_xor_sync(full, n_miss, o);
    if (lane == 0u && n_miss) atomicAdd(&w.ctrl->misses[qi], n_miss);
    if (COUNT) {
        for (int o = 16; o; o >>= 1) { c_box += __shfl_xor_sync(full, c_box, o); c_tri += __shfl_xor_sync(full, c_tri, o); }
        if (lane == 0u) { atomicAdd(&w.ctrl->c[0], c_box); atomicAdd(&w.ctrl->c[1], c_tri); }
    }
}


// Second half of shading one path of the current depth, given its closest hit: primary-hit record, SPEC.md steps
// 3.2-3.8 (path_bounce) and 4; survivors become the compact path list of the next depth (warp-aggregated append,
// order within the warp kept).  Called by all 32 lanes of a converged warp; `have` marks the lanes holding a path.
template <bool COUNT>
__device__ __forceinline__ void wv_finish_path(const KParams& p, const WvBuffers& w, int qi, bool have, const Hit& h, Ray& ray,
                                               V3& tp, V3& L, SimpleRng& rng, uint32_t& depth, uint32_t idx, uint32_t sl) {
    const unsigned full = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    bool cont = false;
    if (have) {
        if (COUNT && h.kind != 0u) atomicAdd(&w.ctrl->c[2], 1ull);
        if (depth == 0u && w.s0 + sl == 0u && (p.outputs & RT_OUT_HITS)) {
            uint32_t px, py;
            work_to_pixel(p, idx, &px, &py);
            reinterpret_cast<uint4*>(p.hits)[(size_t)py * p.width + px] = make_uint4(h.kind, h.prim, __float_as_uint(h.t), h.material_id);
        }
        const bool end = path_bounce(p, h, ray, tp, L, rng, depth); // SPEC.md steps 3.2-3.8 (rt_kernels.cuh)
        cont = !end;
        if (end) { WV_CHECK(sl < w.spp_wave && idx < p.n_work); WV_ST(w.lbuf + (size_t)sl * p.n_work + idx, make_float4(L.x, L.y, L.z, 1.0f)); } // SPEC.md step 4, summed by wv_resolve
    }
    unsigned m = __ballot_sync(full, cont);
    if (m) {
        uint32_t base = 0;
        int leader = __ffs(m) - 1;
        if ((int)lane == leader) base = atomicAdd(&w.ctrl->n[qi ^ 1], (unsigned int)__popc(m));
        base = __shfl_sync(full, base, leader);
        if (cont) {
            uint32_t ns = base + __popc(m & ((1u << lane) - 1u));
            WV_CHECK(ns < w.capacity);
            float4* st = w.state[qi ^ 1] + (size_t)ns * 4;
            WV_ST(st, make_float4(ray.o.x, ray.o.y, ray.o.z, __uint_as_float(rng.seed)));
            WV_ST(st + 1, make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(depth)));
            WV_ST(st + 2, make_float4(tp.x, tp.y, tp.z, __uint_as_float(idx)));
            WV_ST(st + 3, make_float4(L.x, L.y, L.z, __uint_as_float(sl)));
            WV_ST(w.hit[qi ^ 1] + ns, wv_seed_record(p, ray));
        }
    }
}

// ---- fallback: the paths of the current depth that wv_shade could not serve from the re-laid-out scene (REBUILT
// winner failed its certification; zero / tiny direction component; origin far outside the scene) get the
// reference's result here, one warp per ray (closest_triangle_ordered_warp / _literal_warp), then the same second half of
// the shading.  Launched after the shade launches of every depth; normally the list is empty (C3: 5 of 70 M
// segments).  Kept out of wv_shade because (a) one such ray on one lane takes ~0.25 ms on 100-triangle leaves and
// holds a small launch (a 1/8 frame share) up for that long, and (b) the literal traversal's registers and
// 256-byte stack slowed every shade launch by ~13 % when it was inlined there.
template <bool COUNT, bool GEN>
__global__ void __launch_bounds__(256) wv_fallback(const __grid_constant__ KParams p, WvBuffers w, int qi) {
    const unsigned full = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t n_fb = w.ctrl->n_fb;
    const uint32_t warps = gridDim.x * 8u;
    for (uint32_t j = blockIdx.x * 8u + (threadIdx.x >> 5); j < n_fb; j += warps) {
        const uint32_t entry = w.fb_list[j], i = entry & 0x7FFFFFFFu;
        const bool filtered = (entry & 0x80000000u) != 0u; // uniform over the warp
        // every lane holds the same path (uniform loads)
        Ray ray; ray.o = mk(0, 0, 0); ray.d = mk(0, 0, 1);
        V3 tp = mk(1.0f, 1.0f, 1.0f), L = mk(0, 0, 0);
        SimpleRng rng; rng.seed = 0;
        uint32_t depth = 0, idx = 0, sl = 0;
        if (GEN) wv_primary(p, w, i, &idx, &sl, &ray, &rng.seed);
        else {
            const float4* st = w.state[qi] + (size_t)i * 4;
            float4 s0 = __ldcs(st), s1 = __ldcs(st + 1), s2 = __ldcs(st + 2), s3 = __ldcs(st + 3);
            ray.o = mk(s0.x, s0.y, s0.z); ray.d = mk(s1.x, s1.y, s1.z);
            rng.seed = __float_as_uint(s0.w);
            depth = __float_as_uint(s1.w); idx = __float_as_uint(s2.w); sl = __float_as_uint(s3.w);
            tp = mk(s2.x, s2.y, s2.z); L = mk(s3.x, s3.y, s3.z);
        }
        float sphere_seed; uint32_t sphere_idx;
        wv_sphere_seed(p, ray, &sphere_seed, &sphere_idx); // lib.rs:183-190: spheres first, then triangles seeded with it
        TriBest b;
        // the ORDERED policy over the uploaded tree returns the reference's hit for every ray (pinned to the
        // oracle and to LITERAL by the parity suite) and prunes; without a usable re-layout, the literal walk
        // a winner that failed certification: search the internal tree again, accepting only triangles whose own
        // uploaded leaf box passes the exact slab test (closest_triangle_rebuilt_t<true>: that minimum is the reference
        // result); all lanes run it on the same ray
        if (filtered) closest_triangle_rebuilt_t<true>(p, ray, sphere_seed, b);
        else if (p.fallback_ordered) closest_triangle_ordered_warp(p, ray, sphere_seed, b);
        else closest_triangle_literal_warp(p, ray, sphere_seed, b);
        Hit h;
        h.t = FLT_MAX; h.kind = 0u; h.prim = 0xFFFFFFFFu; h.material_id = 0xFFFFFFFFu;
        h.point = mk(0, 0, 0); h.normal = mk(0, 0, 0);
        if (b.hit) {
            h.t = b.t; h.kind = 2u; h.prim = b.prim; h.material_id = b.mat;
            h.point = ray.o + ray.d * b.t;
            h.normal = normalize3(cross3(b.e1, b.e2));
        } else if (sphere_idx != WV_NO_SPHERE) {
            uint64_t o = (uint64_t)p.off.spheres_offset + (uint64_t)sphere_idx * 5;
            V3 c = mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2));
            h.t = sphere_seed; h.kind = 1u; h.prim = sphere_idx; h.material_id = meta_at(p, o + 4);
            h.point = ray.o + ray.d * sphere_seed;
            h.normal = normalize3(h.point - c);
        }
        if (COUNT && lane == 0u) { atomicAdd(&w.ctrl->c[3], 1ull); atomicAdd(&w.ctrl->c[0], (unsigned long long)b.n_box); atomicAdd(&w.ctrl->c[1], (unsigned long long)b.n_tri); }
        wv_finish_path<COUNT>(p, w, qi, lane == 0u, h, ray, tp, L, rng, depth, idx, sl);
    }
    (void)full;
}

// ---- shade: SPEC.md steps 3.2-3.8 and 4 for every path of the current depth ----------------
// Both forms (SPLIT = false / true) are launched for every depth; each looks at the miss count wv_extend left in
// ctrl->misses[qi] and only the matching one runs (the other returns at once): the split form when more than a
// quarter of the rays left the scene.  Results are identical either way; on a closed scene the queueing only
// costs (Cornell box: 5 % of the frame), on the open terrain it saves a sixth of the shading time.
template <int POLICY, bool COUNT, bool GEN, bool SPLIT>
__device__ __forceinline__ void wv_shade_body(const KParams& p, const WvBuffers& w, int qi, uint32_t (*s_list)[3][64]);

#ifndef WV_SHADE_ONE_LAUNCH
#define WV_SHADE_ONE_LAUNCH 1 // 1: one shade launch per depth holding both forms (the device picks); 0: one launch per form
#endif
// SPLIT here: 0 in-place form only, 1 split form only (two launches per depth, each returns at once if the miss
// ratio asks for the other), 2 both forms in one launch
template <int POLICY, bool COUNT, bool GEN, int SPLIT>
__global__ void __launch_bounds__(256, WV_SHADE_BLOCKS) wv_shade(const __grid_constant__ KParams p, WvBuffers w, int qi) {
    __shared__ uint32_t s_list[SPLIT == 0 ? 1 : 8][3][64];
    const uint32_t n = GEN ? p.n_work * w.spp_wave : w.ctrl->n[qi];
    if (SPLIT != 1 && blockIdx.x == 0 && threadIdx.x == 0) { // wv_extend of this depth has finished (first launch of the depth)
        w.ctrl->head = 0;
        const unsigned long long traced = GEN ? (unsigned long long)w.valid_pixels * w.spp_wave : (unsigned long long)n;
        w.ctrl->segments += traced;
        w.ctrl->rays_at_depth[w.depth < 7u ? w.depth : 7u] += (unsigned int)traced;
    }
    const bool want_split = (unsigned long long)w.ctrl->misses[qi] * 4ull > (unsigned long long)n * w.split_quarters;
    if (SPLIT == 2) {
        if (want_split) wv_shade_body<POLICY, COUNT, GEN, true>(p, w, qi, s_list);
        else wv_shade_body<POLICY, COUNT, GEN, false>(p, w, qi, s_list);
    } else if (want_split == (SPLIT == 1)) wv_shade_body<POLICY, COUNT, GEN, SPLIT == 1>(p, w, qi, s_list);
}

template <int POLICY, bool COUNT, bool GEN, bool SPLIT>
__device__ __forceinline__ void wv_shade_body(const KParams& p, const WvBuffers& w, int qi, uint32_t (*s_list)[3][64]) {
    const unsigned full = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t n = GEN ? p.n_work * w.spp_wave : w.ctrl->n[qi];
    const float4* __restrict__ tris = POLICY == 2 ? p.btris : p.wtris;
    // Paths whose ray left the scene (on open scenes most of depth >= 1) only add the sky term; if they shared
    // warps with the hits, the shading code would run with a handful of lanes (ncu: 7 of 32 at depth 1 of C3).
    // Every warp owns a small queue in shared memory: phase A of each 32-path chunk finishes the misses and queues
    // the others; phase B runs the full shading on 32 queued paths at a time, in queue order.  Warps never wait
    // for each other (a dense chunk goes straight through: 32 queued, 32 shaded).
    // A queue entry is (path, hit record); the hit records of the next chunk are loaded before the current one is
    // shaded, so the only global-memory wait on the dense path stays the path-state load of phase B.
    uint32_t (*const my_list)[64] = s_list[SPLIT ? (threadIdx.x >> 5) : 0];
    uint32_t queued_now = 0; // warp-uniform
    const uint32_t n_round = (n + 31u) & ~31u;
    const uint32_t warp_stride = gridDim.x * 256u;
    const uint32_t first = blockIdx.x * 256u + (threadIdx.x & ~31u);
    uint2 hr_next = make_uint2(0u, WV_ID_MISS);
    if (SPLIT && first + lane < n) hr_next = __ldcs(w.hit[qi] + first + lane);
    for (uint32_t base = first; base < n_round; base += warp_stride) {
        if (SPLIT) { // ---- phase A
            const uint32_t i = base + lane;
            const uint2 hr = hr_next;
            if (i + warp_stride < n && i + warp_stride > i) hr_next = __ldcs(w.hit[qi] + i + warp_stride);
            bool heavy = false;
            if (i < n) {
                uint32_t idx = 0, sl = 0, px = 0, py = 0, depth = 0;
                bool valid = true;
                if (GEN) valid = wv_primary_index(p, w, i, &idx, &sl, &px, &py);
                if (valid) {
                    if (hr.y != WV_ID_MISS) heavy = true;
                    else {
                        V3 tp = mk(1.0f, 1.0f, 1.0f), L = mk(0.0f, 0.0f, 0.0f);
                        if (!GEN) {
                            const float4* st = w.state[qi] + (size_t)i * 4;
                            const float4 s2 = __ldcs(st + 2), s3 = __ldcs(st + 3);
                            tp = mk(s2.x, s2.y, s2.z); idx = __float_as_uint(s2.w);
                            L = mk(s3.x, s3.y, s3.z); sl = __float_as_uint(s3.w);
                            if (p.outputs & RT_OUT_HITS) depth = __float_as_uint(__ldcs(st + 1).w);
                        }
                        if (depth == 0u && w.s0 + sl == 0u && (p.outputs & RT_OUT_HITS)) {
                            if (!GEN) work_to_pixel(p, idx, &px, &py);
                            reinterpret_cast<uint4*>(p.hits)[(size_t)py * p.width + px] = make_uint4(0u, 0xFFFFFFFFu, __float_as_uint(FLT_MAX), 0xFFFFFFFFu);
                        }
                        L = L + mk(0.1f, 0.2f, 0.3f) * tp;
                        { WV_CHECK(sl < w.spp_wave && idx < p.n_work); WV_ST(w.lbuf + (size_t)sl * p.n_work + idx, make_float4(L.x, L.y, L.z, 1.0f)); }
                    }
                }
            }
            const unsigned hm = __ballot_sync(full, heavy);
            if (heavy) {
                const uint32_t e = queued_now + (uint32_t)__popc(hm & ((1u << lane) - 1u));
                WV_CHECK(e < 64u);
                my_list[0][e] = i; my_list[1][e] = hr.x; my_list[2][e] = hr.y;
            }
            queued_now += (uint32_t)__popc(hm);
            __syncwarp();
        }
        // ---- phase B: 32 queued paths at a time; after the warp's last chunk until its queue is empty
        const bool last_chunk = base + warp_stride >= n_round;
        for (;;) {
        bool have;
        uint32_t i;
        uint2 hr = make_uint2(0u, WV_ID_MISS);
        if (SPLIT) {
            const uint32_t queued = queued_now;
            if (!(queued >= 32u || (last_chunk && queued > 0u))) break;
            const uint32_t take = queued < 32u ? queued : 32u;
            have = lane < take;
            i = have ? my_list[0][lane] : 0u;
            if (have) hr = make_uint2(my_list[1][lane], my_list[2][lane]);
            if (queued > 32u) { // the rest moves to the front
                const bool mv = lane + 32u < queued;
                const uint32_t m0 = mv ? my_list[0][lane + 32u] : 0u, m1 = mv ? my_list[1][lane + 32u] : 0u, m2 = mv ? my_list[2][lane + 32u] : 0u;
                __syncwarp();
                if (mv) { my_list[0][lane] = m0; my_list[1][lane] = m1; my_list[2][lane] = m2; }
            }
            queued_now = queued - take;
            __syncwarp(); // the next chunk's phase A appends behind the moved entries
        } else { // every path of the chunk, in place
            i = base + lane;
            have = i < n;
        }
        bool cont = false;
        Ray ray; ray.o = mk(0, 0, 0); ray.d = mk(0, 0, 1);
        V3 tp = mk(0, 0, 0), L = mk(0, 0, 0);
        SimpleRng rng; rng.seed = 0;
        uint32_t depth = 0, idx = 0, sl = 0;
        if (GEN) { // the same primary ray the depth-0 extend traced for position i
            if (have) have = wv_primary(p, w, i, &idx, &sl, &ray, &rng.seed);
            tp = mk(1.0f, 1.0f, 1.0f);
        }
        if (have) {
            if (!SPLIT) hr = __ldcs(w.hit[qi] + i);
            if (!GEN) {
                const float4* st = w.state[qi] + (size_t)i * 4;
                float4 s0 = __ldcs(st), s1 = __ldcs(st + 1), s2 = __ldcs(st + 2), s3 = __ldcs(st + 3);
                ray.o = mk(s0.x, s0.y, s0.z); ray.d = mk(s1.x, s1.y, s1.z);
                rng.seed = __float_as_uint(s0.w);
                depth = __float_as_uint(s1.w); idx = __float_as_uint(s2.w); sl = __float_as_uint(s3.w);
                tp = mk(s2.x, s2.y, s2.z); L = mk(s3.x, s3.y, s3.z);
            }
            // ---- rebuild the Hit from the record
            Hit h;
            h.t = FLT_MAX; h.kind = 0u; h.prim = 0xFFFFFFFFu; h.material_id = 0xFFFFFFFFu;
            h.point = mk(0, 0, 0); h.normal = mk(0, 0, 0);
            const uint32_t kind = hr.y >> 30;
            bool lit = hr.y != WV_ID_MISS && kind == WV_KIND_LITERAL;
            TriBest b;
            b.hit = false; b.n_box = 0; b.n_tri = 0; b.n_fallback = 0;
            if (hr.y != WV_ID_MISS && kind == WV_KIND_TRI) {
                const float4* tpn = tris + (uint64_t)hr.y * 3;
                float4 a = __ldg(tpn), c = __ldg(tpn + 1), d = __ldg(tpn + 2);
                b.hit = true; b.t = __uint_as_float(hr.x);
                b.v0 = mk(a.x, a.y, a.z); b.e1 = mk(a.w, c.x, c.y); b.e2 = mk(c.z, c.w, d.x);
                b.mat = __float_as_uint(d.y); b.prim = __float_as_uint(d.z);
                if (POLICY == 2) { // certification against the uploaded tree (see closest_triangle_rebuilt)
                    uint32_t leaf = __ldg(p.leaf_of_pos + __float_as_uint(d.w));
                    const float4 blo = __ldg(p.cert_box + (uint64_t)leaf * 2), bhi = __ldg(p.cert_box + (uint64_t)leaf * 2 + 1);
                    V3 inv = mk(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
                    float entry;
                    lit = !slab(ray.o, inv, mk(blo.x, blo.y, blo.z), mk(bhi.x, bhi.y, bhi.z), &entry);
                }
            }
            float sphere_seed = FLT_MAX - 2.0f;
            uint32_t sphere_idx = WV_NO_SPHERE;
            if (hr.y != WV_ID_MISS && kind == WV_KIND_SPHERE) { sphere_seed = __uint_as_float(hr.x); sphere_idx = hr.y & WV_NO_SPHERE; }
            if (lit) { // rare: the internal tree's answer does not stand for this ray -> wv_fallback (launched after this kernel)
                // bit 31: the ray went through the internal tree and its winner failed certification (filtered re-search);
                // clear: the tree could not be used at all (zero / tiny direction component, far origin)
                const uint32_t fe = atomicAdd(&w.ctrl->n_fb, 1u);
                WV_CHECK(fe < w.capacity);
                w.fb_list[fe] = i | ((hr.y != WV_ID_MISS && kind == WV_KIND_TRI) ? 0x80000000u : 0u);
            } else {
            if (b.hit) {
                h.t = b.t; h.kind = 2u; h.prim = b.prim; h.material_id = b.mat;
                h.point = ray.o + ray.d * b.t;
                h.normal = normalize3(cross3(b.e1, b.e2));
            } else if (sphere_idx != WV_NO_SPHERE) {
                uint64_t o = (uint64_t)p.off.spheres_offset + (uint64_t)sphere_idx * 5;
                V3 c = mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2));
                h.t = sphere_seed; h.kind = 1u; h.prim = sphere_idx; h.material_id = meta_at(p, o + 4);
                h.point = ray.o + ray.d * sphere_seed;
                h.normal = normalize3(h.point - c);
            }
            if (COUNT && h.kind != 0u) atomicAdd(&w.ctrl->c[2], 1ull);
            if (depth == 0u && w.s0 + sl == 0u && (p.outputs & RT_OUT_HITS)) {
                uint32_t px, py;
                work_to_pixel(p, idx, &px, &py);
                reinterpret_cast<uint4*>(p.hits)[(size_t)py * p.width + px] = make_uint4(h.kind, h.prim, __float_as_uint(h.t), h.material_id);
            }
            const bool end = path_bounce(p, h, ray, tp, L, rng, depth); // SPEC.md steps 3.2-3.8 (rt_kernels.cuh)
            cont = !end;
            if (end) { WV_CHECK(sl < w.spp_wave && idx < p.n_work); WV_ST(w.lbuf + (size_t)sl * p.n_work + idx, make_float4(L.x, L.y, L.z, 1.0f)); } // SPEC.md step 4, summed by wv_resolve
            } // !lit
        }
        // ---- survivors become the compact path list of the next depth (order within the warp kept; the same
        // append as wv_finish_path, which wv_fallback uses for the rays it serves)
        unsigned m = __ballot_sync(full, cont);
        if (m) {
            uint32_t base = 0;
            int leader = __ffs(m) - 1;
            if ((int)lane == leader) base = atomicAdd(&w.ctrl->n[qi ^ 1], (unsigned int)__popc(m));
            base = __shfl_sync(full, base, leader);
            if (cont) {
                uint32_t ns = base + __popc(m & ((1u << lane) - 1u));
                WV_CHECK(ns < w.capacity);
            WV_CHECK(ns < w.capacity);
                float4* st = w.state[qi ^ 1] + (size_t)ns * 4;
                WV_ST(st, make_float4(ray.o.x, ray.o.y, ray.o.z, __uint_as_float(rng.seed)));
                WV_ST(st + 1, make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(depth)));
                WV_ST(st + 2, make_float4(tp.x, tp.y, tp.z, __uint_as_float(idx)));
                WV_ST(st + 3, make_float4(L.x, L.y, L.z, __uint_as_float(sl)));
                WV_ST(w.hit[qi ^ 1] + ns, wv_seed_record(p, ray));
            }
        }
        if (!SPLIT) break;
        } // phase B
    }
}

// ---- resolve: add the wave's sample radiances to the accumulation buffer in sample order ----
__global__ void __launch_bounds__(256) wv_resolve(const __grid_constant__ KParams p, WvBuffers w) {
    if (blockIdx.x == 0 && threadIdx.x == 0) { w.ctrl->misses[0] = 0; w.ctrl->misses[1] = 0; } // next wave starts clean
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < p.n_work; idx += gridDim.x * blockDim.x) {
        uint32_t px, py;
        if (!work_to_pixel(p, idx, &px, &py)) continue;
        const size_t pix = (size_t)py * p.width + px;
        float4 a = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (!(w.first_wave && p.reset_accum)) a = p.accum[pix];
        for (uint32_t sl = 0; sl < w.spp_wave; sl++) {
            float4 l = __ldcs(w.lbuf + (size_t)sl * p.n_work + idx);
            a.x = a.x + l.x; a.y = a.y + l.y; a.z = a.z + l.z; a.w = a.w + 1.0f;
        }
        if (!w.last_wave || (p.outputs & RT_OUT_ACCUM)) p.accum[pix] = a;
        if (w.last_wave && (p.outputs & RT_OUT_COMPOSITE))
            p.composite[pix] = make_uchar4((unsigned char)unorm8(a.x / a.w), (unsigned char)unorm8(a.y / a.w),
                                           (unsigned char)unorm8(a.z / a.w), 255);
    }
    if (p.outputs & RT_OUT_PEER) __threadfence_system(); // composite stores went to another GPU's frame
}

// ---- primary frames with the reference's semantics (RT_MODE_LEGACY / RT_MODE_WAVEFRONT0) through the wave
// engine's traversal: wv_extend<GEN> finds the closest triangle of every primary ray (shared-memory stack
// column, one-sector nodes, speculative leaf postponement), then this kernel, convergent with one thread per
// primary ray in queue order, certifies the winner (REBUILT), shades it exactly like main_cs (lib.rs:59-89 /
// 92-149 at depth 0) and runs the output stage.  Results are bit-identical to rt_trace_shade's.
template <int POLICY, bool COUNT>
__global__ void __launch_bounds__(256) wv_shade_primary(const __grid_constant__ KParams p, WvBuffers w) {
    const uint32_t n = p.n_work;
    const float4* __restrict__ tris = POLICY == 2 ? p.btris : p.wtris;
    if (blockIdx.x == 0 && threadIdx.x == 0) { w.ctrl->head = 0; w.ctrl->segments += (unsigned long long)w.valid_pixels; }
    const bool wavefront = p.mode == RT_MODE_WAVEFRONT0;
    unsigned long long c_shaded = 0, c_fb = 0, c_box = 0, c_tri = 0;
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x) {
        uint32_t idx, sl, seed, px, py;
        Ray ray;
        if (!wv_primary(p, w, q, &idx, &sl, &ray, &seed)) continue;
        work_to_pixel(p, idx, &px, &py);
        const uint2 hr = __ldcs(w.hit[0] + q);
        Hit h;
        h.t = FLT_MAX; h.kind = 0u; h.prim = 0xFFFFFFFFu; h.material_id = 0xFFFFFFFFu;
        h.point = mk(0, 0, 0); h.normal = mk(0, 0, 0); h.n_box = 0; h.n_tri = 0; h.n_fallback = 0;
        V3 color = mk(0.0f, 0.0f, 0.0f);
        if (!wavefront || p.cur_depth <= p.max_depth) {
            const uint32_t kind = hr.y >> 30;
            bool lit = hr.y != WV_ID_MISS && kind == WV_KIND_LITERAL;
            TriBest b;
            b.hit = false; b.n_box = 0; b.n_tri = 0; b.n_fallback = 0;
            if (hr.y != WV_ID_MISS && kind == WV_KIND_TRI) {
                const float4* tpn = tris + (uint64_t)hr.y * 3;
                float4 a = __ldg(tpn), c = __ldg(tpn + 1), d = __ldg(tpn + 2);
                b.hit = true; b.t = __uint_as_float(hr.x);
                b.v0 = mk(a.x, a.y, a.z); b.e1 = mk(a.w, c.x, c.y); b.e2 = mk(c.z, c.w, d.x);
                b.mat = __float_as_uint(d.y); b.prim = __float_as_uint(d.z);
                if (POLICY == 2) { // certification against the uploaded tree (see closest_triangle_rebuilt)
                    uint32_t leaf = __ldg(p.leaf_of_pos + __float_as_uint(d.w));
                    const float4 blo = __ldg(p.cert_box + (uint64_t)leaf * 2), bhi = __ldg(p.cert_box + (uint64_t)leaf * 2 + 1);
                    V3 inv = mk(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
                    float entry;
                    lit = !slab(ray.o, inv, mk(blo.x, blo.y, blo.z), mk(bhi.x, bhi.y, bhi.z), &entry);
                }
            }
            float sphere_seed = FLT_MAX - 2.0f;
            uint32_t sphere_idx = WV_NO_SPHERE;
            if (hr.y != WV_ID_MISS && kind == WV_KIND_SPHERE) { sphere_seed = __uint_as_float(hr.x); sphere_idx = hr.y & WV_NO_SPHERE; }
            if (lit) { // rare: the reference's own traversal for this ray
                wv_sphere_seed(p, ray, &sphere_seed, &sphere_idx);
                closest_triangle_literal(p, ray, sphere_seed, b);
                if (COUNT) { c_fb++; c_box += b.n_box; c_tri += b.n_tri; }
            }
            if (b.hit) {
                h.t = b.t; h.kind = 2u; h.prim = b.prim; h.material_id = b.mat;
                h.point = ray.o + ray.d * b.t;
                h.normal = normalize3(cross3(b.e1, b.e2));
            } else if (sphere_idx != WV_NO_SPHERE) {
                uint64_t o = (uint64_t)p.off.spheres_offset + (uint64_t)sphere_idx * 5;
                V3 c = mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2));
                h.t = sphere_seed; h.kind = 1u; h.prim = sphere_idx; h.material_id = meta_at(p, o + 4);
                h.point = ray.o + ray.d * sphere_seed;
                h.normal = normalize3(h.point - c);
            }
            if (COUNT && h.kind != 0u) c_shaded++;
            if (!wavefront) {
                if (h.kind != 0u) color = shade_hit(p, h, 0xFFu); // miss: (0,0,0), lib.rs:77
            } else { // run_wavefront_raytracing, lib.rs:92-149: throughput (1,1,1); sky on miss
                V3 tp = mk(1.0f, 1.0f, 1.0f);
                V3 c = (h.kind != 0u) ? shade_hit(p, h, 0xFFu) * tp : mk(0.1f, 0.2f, 0.3f) * tp;
                color = color + c;
            }
        }
        store_pixel(p, px, py, color, h, true, 0xFFu);
    }
    if (p.outputs & RT_OUT_PEER) __threadfence_system(); // the composite store went to another GPU's frame
    if (COUNT) {
        if (c_shaded) atomicAdd(&w.ctrl->c[2], c_shaded);
        if (c_fb) { atomicAdd(&w.ctrl->c[3], c_fb); atomicAdd(&w.ctrl->c[0], c_box); atomicAdd(&w.ctrl->c[1], c_tri); }
    }
}
