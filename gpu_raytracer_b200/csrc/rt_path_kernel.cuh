// rt_path_kernel.cuh — EXTENSION (SPEC.md) path tracing, second-generation kernel.
//
// Same semantics and bit-exact results as rt_path_trace (rt_kernels.cuh); different
// execution shape.  ncu on rt_path_trace showed 7.8 of 32 lanes active per issued
// instruction: lanes that finish their ray early (sky misses, short traversals)
// wait for the slowest traversal of the warp, once per segment.  Here every lane is a
// small state machine and the warp alternates between two phases:
//   T  traversal steps (one node pop = two box tests, or <= 4 triangle tests) for all
//      lanes that are traversing; the phase ends as soon as RT_REFILL lanes have
//      finished their ray (or nobody is traversing);
//   S  for the finished lanes only, together: certify / shade the hit, sample the
//      continuation or start the next sample / fetch the next pixel (one
//      warp-aggregated atomicAdd), and arm the next ray.  Unfinished lanes keep their
//      traversal state (node, stack, best hit) in registers / shared memory.
// The traversal stack is a per-thread column of a shared-memory array (conflict-free:
// entry k of the 32 lanes of a warp are 32 consecutive words) spilling to local
// memory beyond RT_SSTACK entries.  Per-pixel sample order stays fixed (each lane owns
// one pixel at a time), so fp32 accumulation is bit-reproducible.
#pragma once
#include "rt_kernels.cuh"

#define RT_SSTACK 16   // shared-memory stack entries per thread (16 KB per 256-thread block)
#ifndef RT_REFILL
#define RT_REFILL 12   // leave the traversal phase when this many lanes have finished their ray
#endif
#ifndef RT_TRI_BATCH
#define RT_TRI_BATCH 16 // leave the node loop when this many lanes have triangles pending
#endif

enum { LS_NEED_WORK = 0, LS_TRAVERSING = 1, LS_READY = 2, LS_EXHAUSTED = 3 };

template <int POLICY, bool COUNT> // POLICY 1 (ORDERED) or 2 (REBUILT)
__global__ void __launch_bounds__(256, 3) rt_path_trace_v2(const __grid_constant__ KParams p) {
    __shared__ uint32_t s_stack[RT_SSTACK][256];
    uint32_t lstack[64 - RT_SSTACK];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const unsigned full = 0xFFFFFFFFu;
    const float4* __restrict__ nodes = POLICY == 2 ? p.bnodes : p.wnodes;
    const float4* __restrict__ tris = POLICY == 2 ? p.btris : p.wtris;
    const uint32_t root_ref = POLICY == 2 ? p.broot_ref : p.root_ref;

    // ---- per-lane state
    int state = LS_NEED_WORK;
    uint32_t px = 0, py = 0, s = 0, depth = 0, n_seg = 0;
    const uint32_t s_end = p.first_sample + p.spp;
    SimpleRng rng; rng.seed = 0;
    Ray ray; ray.o = mk(0, 0, 0); ray.d = mk(0, 0, 1);
    V3 inv = mk(0, 0, 0);
    V3 tp = mk(1, 1, 1), L = mk(0, 0, 0), sum = mk(0, 0, 0);
    float wsum = 0.0f;
    // traversal state
    uint32_t cur = 0; int sp = 0;
    uint32_t tri_slot = 0, tri_rem = 0;
    float best_t = FLT_MAX, sphere_t_seed = FLT_MAX;
    uint32_t best_slot = 0xFFFFFFFFu, best_pos = 0u, sphere_idx = 0xFFFFFFFFu;
    bool use_literal = false;
    unsigned long long c_box = 0, c_tri = 0, c_shaded = 0, c_fb = 0;
    unsigned long long c_sphase = 0, c_sready = 0, c_titer = 0, c_tactive = 0; // warp-level, lane 0 only

    auto push = [&](uint32_t v) { if (sp < RT_SSTACK) s_stack[sp][tid] = v; else lstack[sp - RT_SSTACK] = v; sp++; };
    auto pop = [&]() -> uint32_t { sp--; return sp < RT_SSTACK ? s_stack[sp][tid] : lstack[sp - RT_SSTACK]; };

    for (;;) {
        // ================================ phase S ================================
        bool arm = false; // this lane has a fresh ray to arm for traversal
        if (COUNT) { unsigned r = __ballot_sync(full, state == LS_READY); c_sphase++; c_sready += __popc(r); }
        if (state == LS_READY) {
            // ---- resolve the closest hit of the finished segment
            Hit h;
            h.t = FLT_MAX; h.kind = 0u; h.prim = 0xFFFFFFFFu; h.material_id = 0xFFFFFFFFu;
            h.point = mk(0, 0, 0); h.normal = mk(0, 0, 0);
            TriBest b;
            b.hit = false; b.t = sphere_t_seed; b.n_box = 0; b.n_tri = 0; b.n_fallback = 0;
            bool lit = use_literal;
            if (!lit && best_slot != 0xFFFFFFFFu) {
                const float4* tpn = tris + (uint64_t)best_slot * 3;
                float4 a = __ldg(tpn), c = __ldg(tpn + 1), d = __ldg(tpn + 2);
                b.hit = true; b.t = best_t;
                b.v0 = mk(a.x, a.y, a.z); b.e1 = mk(a.w, c.x, c.y); b.e2 = mk(c.z, c.w, d.x);
                b.mat = __float_as_uint(d.y); b.prim = __float_as_uint(d.z);
                if (POLICY == 2) { // certification against the uploaded tree (see closest_triangle_rebuilt)
                    uint32_t leaf = __ldg(p.leaf_of_pos + best_pos);
                    uint64_t no = (uint64_t)p.off.bvh_nodes_offset + (uint64_t)leaf * 12;
                    float entry;
                    if (COUNT) c_box++;
                    lit = !slab(ray.o, inv, mk(meta_f(p, no), meta_f(p, no + 1), meta_f(p, no + 2)),
                                mk(meta_f(p, no + 4), meta_f(p, no + 5), meta_f(p, no + 6)), &entry);
                }
            }
            if (lit) {
                closest_triangle_literal(p, ray, sphere_t_seed, b);
                if (COUNT) { c_fb++; c_box += b.n_box; c_tri += b.n_tri; }
            }
            if (b.hit) {
                h.t = b.t; h.kind = 2u; h.prim = b.prim; h.material_id = b.mat;
                h.point = ray.o + ray.d * b.t;
                h.normal = normalize3(cross3(b.e1, b.e2));
            } else if (sphere_idx != 0xFFFFFFFFu) {
                uint64_t o = (uint64_t)p.off.spheres_offset + (uint64_t)sphere_idx * 5;
                V3 c = mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2));
                h.t = sphere_t_seed; h.kind = 1u; h.prim = sphere_idx; h.material_id = meta_at(p, o + 4);
                h.point = ray.o + ray.d * sphere_t_seed;
                h.normal = normalize3(h.point - c);
            }
            n_seg++;
            if (COUNT) c_shaded += (h.kind != 0u);
            if (depth == 0u && s == 0u && (p.outputs & RT_OUT_HITS))
                reinterpret_cast<uint4*>(p.hits)[(size_t)py * p.width + px] =
                    make_uint4(h.kind, h.prim, __float_as_uint(h.t), h.material_id);
            // ---- SPEC.md step 3.2 - 3.8
            bool end = false;
            if (h.kind == 0u) { L = L + mk(0.1f, 0.2f, 0.3f) * tp; end = true; }
            else if (h.material_id >= p.n_mats) { L = L + mk(1.0f, 0.0f, 1.0f) * tp; end = true; }
            else {
                MatEval m = load_material(p, h.material_id);
                L = L + shade_hit_m(p, h, m, 0xFFu) * tp;
                if (depth == p.path_max_depth) end = true;
                if (!end && depth >= 3u) {
                    float pr = fminf(fmaxf(fmaxf(fmaxf(tp.x, tp.y), tp.z), 0.05f), 1.0f);
                    float r = rng.next_f32();
                    if (r > pr) end = true;
                    else tp = mk(tp.x / pr, tp.y / pr, tp.z / pr);
                }
                if (!end) {
                    float r0 = rng.next_f32();
                    V3 u = mk(0.0f, 0.0f, 0.0f);
                    for (int k = 0; k < 8; k++) {
                        float a = rng.next_f32_signed(), bb = rng.next_f32_signed(), cc = rng.next_f32_signed();
                        float q = a * a + bb * bb + cc * cc;
                        if (q > 1e-6f && q <= 1.0f) { u = mk(a, bb, cc) * (1.0f / sqrtf(q)); break; }
                    }
                    V3 d = ray.d, n = h.normal;
                    float cosi = dot3(n, d);
                    V3 nf = cosi < 0.0f ? n : -n;
                    V3 w = nf + u;
                    if (dot3(w, w) < 1e-8f) w = nf;
                    w = normalize3(w);
                    float T = fminf(fmaxf(m.transmission, 0.0f), 1.0f);
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
                    arm = true;
                }
            }
            if (end) { // step 4
                sum = sum + L; wsum = wsum + 1.0f; s++;
                if (s == s_end) {
                    const size_t pix = (size_t)py * p.width + px;
                    if (p.outputs & RT_OUT_ACCUM) p.accum[pix] = make_float4(sum.x, sum.y, sum.z, wsum);
                    if (p.outputs & RT_OUT_COMPOSITE)
                        p.composite[pix] = make_uchar4((unsigned char)unorm8(sum.x / wsum), (unsigned char)unorm8(sum.y / wsum),
                                                       (unsigned char)unorm8(sum.z / wsum), 255);
                    state = LS_NEED_WORK;
                } else state = LS_NEED_WORK + 100; // same pixel, next sample (temporary marker)
            }
        }
        // ---- work fetch (one warp-aggregated atomic) for lanes without a pixel
        {
            unsigned idle = __ballot_sync(full, state == LS_NEED_WORK);
            while (idle) {
                uint32_t base = 0;
                int leader = __ffs(idle) - 1;
                if ((int)lane == leader) base = atomicAdd(p.work_counter, (unsigned int)__popc(idle));
                base = __shfl_sync(full, base, leader);
                if (state == LS_NEED_WORK) {
                    uint32_t idx = base + __popc(idle & ((1u << lane) - 1u));
                    if (idx >= p.n_work) state = LS_EXHAUSTED;
                    else if (work_to_pixel(p, idx, &px, &py)) {
                        state = LS_NEED_WORK + 100; s = p.first_sample;
                        sum = mk(0, 0, 0); wsum = 0.0f;
                        if (!p.reset_accum) {
                            float4 old = p.accum[(size_t)py * p.width + px];
                            sum = mk(old.x, old.y, old.z); wsum = old.w;
                        }
                    }
                }
                idle = __ballot_sync(full, state == LS_NEED_WORK); // pixels outside the image: fetch again
            }
        }
        if (state == LS_NEED_WORK + 100) { // SPEC.md steps 1-2: new sample
            rng.seed = p.base_seed + s + px + py * p.width;
            float jx = 0.5f, jy = 0.5f;
            if (s != 0u) { jx = rng.next_f32(); jy = rng.next_f32(); }
            ray = camera_ray(p, px, py, jx, jy, true);
            tp = mk(1.0f, 1.0f, 1.0f); L = mk(0.0f, 0.0f, 0.0f); depth = 0;
            arm = true;
        }
        if (arm) { // ---- arm the traversal of the fresh ray (find_closest prologue, lib.rs:174-190)
            state = LS_TRAVERSING;
            float closest = FLT_MAX - 2.0f;
            sphere_idx = 0xFFFFFFFFu;
            for (uint32_t i = 0; i < p.off.spheres_count; i++) {
                uint64_t o = (uint64_t)p.off.spheres_offset + (uint64_t)i * 5;
                float t;
                if (sphere_t(ray, mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2)), meta_f(p, o + 3), closest, &t)) {
                    closest = t; sphere_idx = i;
                }
            }
            sphere_t_seed = closest; best_t = closest; best_slot = 0xFFFFFFFFu; best_pos = 0u;
            sp = 0; tri_rem = 0; cur = root_ref;
            use_literal = false;
            inv = mk(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
            if (POLICY == 2 && (ray.d.x == 0.0f || ray.d.y == 0.0f || ray.d.z == 0.0f)) { use_literal = true; state = LS_READY; }
            else if (cur == RT_CHILD_ABSENT) state = LS_READY;
            else {
                float entry;
                if (COUNT) c_box++;
                bool ok = POLICY == 2 ? slab(ray.o, inv, mka(p.broot_min), mka(p.broot_max), &entry)
                                      : slab(ray.o, inv, mka(p.root_min), mka(p.root_max), &entry);
                if (!ok) state = LS_READY;
            }
        }
        if (__all_sync(full, state == LS_EXHAUSTED)) break;

        // ================================ phase T ================================
        // "while-while": a node loop in which lanes that reach a leaf park (tri_rem > 0) until
        // RT_TRI_BATCH lanes have triangles pending, then a triangle loop that tests one triangle
        // per lane per iteration.  Keeps box tests and Möller–Trumbore tests in separate,
        // densely populated instruction streams instead of one divergent if/else per step.
        for (;;) {
            for (;;) { // ---- node loop
                if (state == LS_TRAVERSING && tri_rem == 0u) {
                    if (cur == 0xFFFFFFFFu) { // need the next node
                        if (sp == 0) state = LS_READY;
                        else cur = pop();
                    }
                    if (state == LS_TRAVERSING) {
                        if (cur & RT_LEAF_FLAG) { // enter a leaf; its triangles wait for the triangle loop
                            if (POLICY == 2) { tri_slot = cur & 0x0FFFFFFFu; tri_rem = ((cur >> 28) & 7u) + 1u; }
                            else { tri_slot = cur & ~RT_LEAF_FLAG; tri_rem = __float_as_uint(__ldg(tris + (uint64_t)tri_slot * 3 + 2).w); }
                            cur = 0xFFFFFFFFu;
                        } else { // internal node: two box tests, near child first
                            const float4* np = nodes + (uint64_t)cur * 4;
                            float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2), q3 = __ldg(np + 3);
                            uint32_t lref = __float_as_uint(q3.x), rref = __float_as_uint(q3.y);
                            float el, er;
                            if (COUNT) c_box += 2;
                            bool hl = slab(ray.o, inv, mk(q0.x, q0.y, q0.z), mk(q0.w, q1.x, q1.y), &el);
                            bool hr = slab(ray.o, inv, mk(q1.z, q1.w, q2.x), mk(q2.y, q2.z, q2.w), &er);
                            float limit = best_t * RT_PRUNE_SCALE + RT_PRUNE_ABS;
                            hl = hl && (lref != RT_CHILD_ABSENT) && (el <= limit);
                            hr = hr && (rref != RT_CHILD_ABSENT) && (er <= limit);
                            if (hl && hr) {
                                bool right_first = er < el;
                                push(right_first ? lref : rref);
                                cur = right_first ? rref : lref;
                            } else if (hl) cur = lref;
                            else if (hr) cur = rref;
                            else cur = 0xFFFFFFFFu;
                        }
                    }
                }
                unsigned node_l = __ballot_sync(full, state == LS_TRAVERSING && tri_rem == 0u);
                unsigned tri_l = __ballot_sync(full, state == LS_TRAVERSING && tri_rem != 0u);
                unsigned ready_l = __ballot_sync(full, state == LS_READY);
                if (COUNT) { c_titer++; c_tactive += __popc(node_l); }
                if (node_l == 0u || __popc(tri_l) >= RT_TRI_BATCH || __popc(ready_l) >= RT_REFILL) break;
            }
            for (;;) { // ---- triangle loop: one triangle per lane per iteration
                bool has = state == LS_TRAVERSING && tri_rem != 0u;
                if (!__any_sync(full, has)) break;
                if (has) {
                    const float4* tpn = tris + (uint64_t)tri_slot * 3;
                    float4 a = __ldg(tpn), c = __ldg(tpn + 1), d = __ldg(tpn + 2);
                    float t = triangle_t(ray, mk(a.x, a.y, a.z), mk(a.w, c.x, c.y), mk(c.z, c.w, d.x));
                    uint32_t pos = POLICY == 2 ? __float_as_uint(d.w) : tri_slot;
                    if (t > 0.0f && (t < best_t || (t == best_t && best_slot != 0xFFFFFFFFu && pos < best_pos))) {
                        best_t = t; best_slot = tri_slot; best_pos = pos;
                    }
                    if (COUNT) c_tri++;
                    tri_slot++; tri_rem--;
                }
            }
            unsigned ready = __ballot_sync(full, state == LS_READY);
            unsigned trav = __ballot_sync(full, state == LS_TRAVERSING);
            if (trav == 0u || __popc(ready) >= RT_REFILL) break;
        }
    }
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
            atomicAdd(p.counters + 8, c_sphase); atomicAdd(p.counters + 9, c_sready);
            atomicAdd(p.counters + 10, c_titer); atomicAdd(p.counters + 11, c_tactive);
        }
    }
}
