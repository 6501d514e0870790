// rt_wavefront.cuh — EXTENSION (SPEC.md) path tracing as a wavefront pipeline.
//
// The reference declares this design and stubs it: 76-byte `WavefrontRay`s
// (shared/src/lib.rs:163-181), `WavefrontCounters` (:183-194), a host loop over bounce
// depths (`run_wavefront_compute`, src/compute.rs:413-479, never called) and
// `generate_continuation_rays` returning 0 (shader/src/wavefront.rs:340-355).  Here it is
// real, with the same per-path arithmetic as SPEC.md / rt_path_trace (bit-identical):
//
//   for sample s:  wf_generate                      camera rays -> queue 0
//     for depth d: wf_extend  (persistent threads)  closest hit of every queued ray
//                  wf_shade   (one thread per ray)  shade, accumulate, sample the
//                                                   continuation, append survivors to the
//                                                   other queue (warp-aggregated atomics)
//
// Why: in a megakernel a finished lane can only be refilled by running the long,
// IEEE-exact shading code with few lanes active (ncu: 8 of 32 lanes per instruction).
// Here a finished lane of wf_extend is refilled by loading 32 bytes from the queue, so
// the traversal loops stay populated, and wf_shade runs convergent over compacted rays.
// Everything stays on the device: queue lengths live in WfCtrl, no host round trip
// inside a frame.  Per-pixel accumulation order is fixed (sample s is finished before
// s+1 starts), so results are bit-reproducible.
#pragma once
#include "rt_kernels.cuh"

struct WfCtrl {
    unsigned int n[2];     // queue lengths (ping-pong)
    unsigned int head;     // wf_extend work-fetch cursor
    unsigned int _pad;
    unsigned long long segments;
    unsigned long long c[5]; // instrumentation: box tests, triangle tests, shaded, fallbacks, refills
};

struct WfBuffers {
    float4* state;         // 4 x float4 per path: (o, rng) (d, depth) (tp, pixel) (L, -)
    uint4* hit;            // per path: t bits, kind (0 miss, 1 sphere, 2 slot in wtris/btris, 3 literal: triangle index), id, -
    uint32_t* queue[2];
    WfCtrl* ctrl;
};

#ifndef WF_REFILL
#define WF_REFILL 8        // refill when this many lanes of the warp are idle
#endif
#ifndef WF_TRI_BATCH
#define WF_TRI_BATCH 8     // leave the node loop when this many lanes have triangles pending
#endif

// ---- sample start: SPEC.md steps 1-2 for every pixel of this context's tiles ----------
__global__ void __launch_bounds__(256) wf_generate(const __grid_constant__ KParams p, WfBuffers w, uint32_t s) {
    uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t px = 0, py = 0;
    bool ok = idx < p.n_work && work_to_pixel(p, idx, &px, &py);
    const unsigned full = 0xFFFFFFFFu;
    unsigned m = __ballot_sync(full, ok);
    if (!m) return;
    uint32_t lane = threadIdx.x & 31u, base = 0;
    int leader = __ffs(m) - 1;
    if ((int)lane == leader) base = atomicAdd(&w.ctrl->n[0], (unsigned int)__popc(m));
    base = __shfl_sync(full, base, leader);
    if (!ok) return;
    uint32_t slot = base + __popc(m & ((1u << lane) - 1u)); // path slot == queue position (compact)
    SimpleRng rng; rng.seed = p.base_seed + s + px + py * p.width;
    float jx = 0.5f, jy = 0.5f;
    if (s != 0u) { jx = rng.next_f32(); jy = rng.next_f32(); }
    Ray ray = camera_ray(p, px, py, jx, jy, true);
    float4* st = w.state + (size_t)slot * 4;
    st[0] = make_float4(ray.o.x, ray.o.y, ray.o.z, __uint_as_float(rng.seed));
    st[1] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(0u));
    st[2] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(py * p.width + px));
    st[3] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    w.queue[0][slot] = slot;
}

// ---- extend: closest hit of every ray in queue[qi] -------------------------------------
template <int POLICY, bool COUNT> // POLICY 1 (ORDERED) or 2 (REBUILT)
__global__ void __launch_bounds__(256, 3) wf_extend(const __grid_constant__ KParams p, WfBuffers w, int qi) {
    __shared__ uint32_t s_stack[16][256];
    uint32_t lstack[48];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const unsigned full = 0xFFFFFFFFu;
    const float4* __restrict__ nodes = POLICY == 2 ? p.bnodes : p.wnodes;
    const float4* __restrict__ tris = POLICY == 2 ? p.btris : p.wtris;
    const uint32_t root_ref = POLICY == 2 ? p.broot_ref : p.root_ref;
    const uint32_t n = w.ctrl->n[qi];
    const uint32_t* __restrict__ queue = w.queue[qi];
    if (blockIdx.x == 0 && tid == 0) w.ctrl->n[qi ^ 1] = 0; // wf_shade of this depth appends there

    bool active = false, exhausted = false;
    uint32_t path = 0;
    Ray ray; ray.o = mk(0, 0, 0); ray.d = mk(0, 0, 1);
    V3 inv = mk(0, 0, 0);
    uint32_t cur = 0xFFFFFFFFu; int sp = 0;
    uint32_t tri_slot = 0, tri_rem = 0;
    float best_t = FLT_MAX, sphere_seed = FLT_MAX;
    uint32_t best_slot = 0xFFFFFFFFu, best_pos = 0u, sphere_idx = 0xFFFFFFFFu;
    bool use_literal = false;
    unsigned long long c_box = 0, c_tri = 0, c_fb = 0, c_refill = 0;
    auto push = [&](uint32_t v) { if (sp < 16) s_stack[sp][tid] = v; else lstack[sp - 16] = v; sp++; };
    auto pop = [&]() -> uint32_t { sp--; return sp < 16 ? s_stack[sp][tid] : lstack[sp - 16]; };

    for (;;) {
        // ---- refill idle lanes from the queue (cheap: 32 bytes per ray)
        unsigned idle = __ballot_sync(full, !active && !exhausted);
        unsigned act = __ballot_sync(full, active);
        if (idle && (__popc(idle) >= WF_REFILL || act == 0u)) {
            uint32_t base = 0;
            int leader = __ffs(idle) - 1;
            if ((int)lane == leader) base = atomicAdd(&w.ctrl->head, (unsigned int)__popc(idle));
            base = __shfl_sync(full, base, leader);
            if (COUNT && lane == 0u) c_refill++;
            if (!active && !exhausted) {
                uint32_t qpos = base + __popc(idle & ((1u << lane) - 1u));
                if (qpos >= n) exhausted = true;
                else {
                    path = queue[qpos];
                    float4 a = w.state[(size_t)path * 4], b = w.state[(size_t)path * 4 + 1];
                    ray.o = mk(a.x, a.y, a.z); ray.d = mk(b.x, b.y, b.z);
                    active = true;
                    // find_closest prologue (shader/src/lib.rs:174-190): spheres first, triangles seeded
                    float closest = FLT_MAX - 2.0f;
                    sphere_idx = 0xFFFFFFFFu;
                    for (uint32_t i = 0; i < p.off.spheres_count; i++) {
                        uint64_t o = (uint64_t)p.off.spheres_offset + (uint64_t)i * 5;
                        float t;
                        if (sphere_t(ray, mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2)), meta_f(p, o + 3), closest, &t)) {
                            closest = t; sphere_idx = i;
                        }
                    }
                    sphere_seed = closest; best_t = closest; best_slot = 0xFFFFFFFFu; best_pos = 0u;
                    sp = 0; tri_rem = 0; cur = root_ref; use_literal = false;
                    inv = mk(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
                    if (POLICY == 2 && (ray.d.x == 0.0f || ray.d.y == 0.0f || ray.d.z == 0.0f)) { use_literal = true; cur = 0xFFFFFFFFu; }
                    else if (cur == RT_CHILD_ABSENT) cur = 0xFFFFFFFFu;
                    else {
                        float entry;
                        if (COUNT) c_box++;
                        bool ok = POLICY == 2 ? slab(ray.o, inv, mka(p.broot_min), mka(p.broot_max), &entry)
                                              : slab(ray.o, inv, mka(p.root_min), mka(p.root_max), &entry);
                        if (!ok) cur = 0xFFFFFFFFu;
                    }
                }
            }
        }
        if (__all_sync(full, !active)) {
            if (__all_sync(full, exhausted)) break;
            continue;
        }
        // ---- node loop: lanes that reach a leaf park until WF_TRI_BATCH lanes have triangles pending
        for (;;) {
            if (active && tri_rem == 0u) {
                bool finished = false;
                if (cur == 0xFFFFFFFFu) {
                    if (sp == 0) finished = true;
                    else cur = pop();
                }
                if (finished) { // ---- resolve + store the hit record, free the lane
                    uint32_t kind = 0u, id = 0xFFFFFFFFu;
                    float t = FLT_MAX;
                    bool lit = use_literal;
                    if (!lit && best_slot != 0xFFFFFFFFu) {
                        kind = 2u; id = best_slot; t = best_t;
                        if (POLICY == 2) { // certify against the uploaded tree (closest_triangle_rebuilt)
                            uint32_t leaf = __ldg(p.leaf_of_pos + best_pos);
                            uint64_t no = (uint64_t)p.off.bvh_nodes_offset + (uint64_t)leaf * 12;
                            float entry;
                            if (COUNT) c_box++;
                            lit = !slab(ray.o, inv, mk(meta_f(p, no), meta_f(p, no + 1), meta_f(p, no + 2)),
                                        mk(meta_f(p, no + 4), meta_f(p, no + 5), meta_f(p, no + 6)), &entry);
                        }
                    }
                    if (lit) {
                        TriBest b;
                        closest_triangle_literal(p, ray, sphere_seed, b);
                        if (COUNT) { c_fb++; c_box += b.n_box; c_tri += b.n_tri; }
                        if (b.hit) { kind = 3u; id = b.prim; t = b.t; } else { kind = 0u; id = 0xFFFFFFFFu; t = FLT_MAX; }
                    }
                    if (kind == 0u && sphere_idx != 0xFFFFFFFFu) { kind = 1u; id = sphere_idx; t = sphere_seed; }
                    w.hit[path] = make_uint4(__float_as_uint(t), kind, id, 0u);
                    active = false;
                } else if (cur & RT_LEAF_FLAG) {
                    if (POLICY == 2) { tri_slot = cur & 0x0FFFFFFFu; tri_rem = ((cur >> 28) & 7u) + 1u; }
                    else { tri_slot = cur & ~RT_LEAF_FLAG; tri_rem = __float_as_uint(__ldg(tris + (uint64_t)tri_slot * 3 + 2).w); }
                    cur = 0xFFFFFFFFu;
                } else {
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
            unsigned node_l = __ballot_sync(full, active && tri_rem == 0u);
            unsigned tri_l = __ballot_sync(full, active && tri_rem != 0u);
            unsigned idle_l = __ballot_sync(full, !active && !exhausted);
            if (node_l == 0u || __popc(tri_l) >= WF_TRI_BATCH || __popc(idle_l) >= WF_REFILL) break;
        }
        // ---- triangle loop: one Möller–Trumbore test per lane per iteration
        for (;;) {
            bool has = active && tri_rem != 0u;
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
    }
    if (COUNT) {
        for (int o = 16; o; o >>= 1) {
            c_box += __shfl_xor_sync(full, c_box, o); c_tri += __shfl_xor_sync(full, c_tri, o);
            c_fb += __shfl_xor_sync(full, c_fb, o); c_refill += __shfl_xor_sync(full, c_refill, o);
        }
        if (lane == 0u) {
            atomicAdd(&w.ctrl->c[0], c_box); atomicAdd(&w.ctrl->c[1], c_tri);
            atomicAdd(&w.ctrl->c[3], c_fb); atomicAdd(&w.ctrl->c[4], c_refill);
        }
    }
}

// ---- shade: SPEC.md steps 3.2-3.8 and 4 for every ray of queue[qi]; survivors -> queue[qi^1]
template <bool COUNT>
__global__ void __launch_bounds__(256) wf_shade(const __grid_constant__ KParams p, WfBuffers w, int qi, uint32_t s) {
    const unsigned full = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t n = w.ctrl->n[qi];
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { w.ctrl->head = 0; w.ctrl->segments += n; } // wf_extend of this depth has finished
    if ((i & ~31u) >= n) return; // whole warp beyond the queue
    bool valid = i < n;
    bool cont = false;
    uint32_t path = 0;
    if (valid) {
        path = w.queue[qi][i];
        float4* st = w.state + (size_t)path * 4;
        float4 s0 = st[0], s1 = st[1], s2 = st[2], s3 = st[3];
        uint4 hr = w.hit[path];
        Ray ray; ray.o = mk(s0.x, s0.y, s0.z); ray.d = mk(s1.x, s1.y, s1.z);
        SimpleRng rng; rng.seed = __float_as_uint(s0.w);
        uint32_t depth = __float_as_uint(s1.w), pix = __float_as_uint(s2.w);
        V3 tp = mk(s2.x, s2.y, s2.z), L = mk(s3.x, s3.y, s3.z);
        // ---- rebuild the Hit from the record
        Hit h;
        h.t = __uint_as_float(hr.x); h.kind = 0u; h.prim = 0xFFFFFFFFu; h.material_id = 0xFFFFFFFFu;
        h.point = mk(0, 0, 0); h.normal = mk(0, 0, 0);
        if (hr.y == 2u) { // triangle slot in the re-laid-out array
            const float4* tpn = (p.wf_rebuilt ? p.btris : p.wtris) + (uint64_t)hr.z * 3;
            float4 a = __ldg(tpn), c = __ldg(tpn + 1), d = __ldg(tpn + 2);
            h.kind = 2u; h.material_id = __float_as_uint(d.y); h.prim = __float_as_uint(d.z);
            h.point = ray.o + ray.d * h.t;
            h.normal = normalize3(cross3(mk(a.w, c.x, c.y), mk(c.z, c.w, d.x)));
        } else if (hr.y == 3u) { // literal-policy winner: triangle index in the reference layout
            V3 v0, v1, v2; uint32_t mat = 0;
            fetch_triangle_ref(p, hr.z, &v0, &v1, &v2, &mat);
            h.kind = 2u; h.material_id = mat; h.prim = hr.z;
            h.point = ray.o + ray.d * h.t;
            h.normal = normalize3(cross3(v1 - v0, v2 - v0));
        } else if (hr.y == 1u) {
            uint64_t o = (uint64_t)p.off.spheres_offset + (uint64_t)hr.z * 5;
            V3 c = mk(meta_f(p, o), meta_f(p, o + 1), meta_f(p, o + 2));
            h.kind = 1u; h.prim = hr.z; h.material_id = meta_at(p, o + 4);
            h.point = ray.o + ray.d * h.t;
            h.normal = normalize3(h.point - c);
        }
        if (COUNT && h.kind != 0u) atomicAdd(&w.ctrl->c[2], 1ull);
        if (depth == 0u && s == 0u && (p.outputs & RT_OUT_HITS))
            reinterpret_cast<uint4*>(p.hits)[pix] = make_uint4(h.kind, h.prim, __float_as_uint(h.t), h.material_id);
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
                V3 d = ray.d, nn = h.normal;
                float cosi = dot3(nn, d);
                V3 nf = cosi < 0.0f ? nn : -nn;
                V3 wv = nf + u;
                if (dot3(wv, wv) < 1e-8f) wv = nf;
                wv = normalize3(wv);
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
                    dir = mr * (1.0f - m.roughness) + wv * m.roughness;
                    if (dot3(dir, dir) < 1e-8f || dot3(dir, nf) <= 0.0f) dir = wv;
                    dir = normalize3(dir);
                } else {
                    dir = wv;
                }
                tp = tp * m.albedo;
                V3 o = h.point + nf * (side * 1e-3f);
                st[0] = make_float4(o.x, o.y, o.z, __uint_as_float(rng.seed));
                st[1] = make_float4(dir.x, dir.y, dir.z, __uint_as_float(depth + 1u));
                st[2] = make_float4(tp.x, tp.y, tp.z, __uint_as_float(pix));
                st[3] = make_float4(L.x, L.y, L.z, 0.0f);
                cont = true;
            }
        }
        if (end) { // SPEC.md step 4: accumulate in sample order (sample s is the only writer of pix now)
            float4 a = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (!(s == p.first_sample && p.reset_accum)) a = p.accum[pix];
            a.x = a.x + L.x; a.y = a.y + L.y; a.z = a.z + L.z; a.w = a.w + 1.0f;
            p.accum[pix] = a;
            if (s + 1u == p.first_sample + p.spp && (p.outputs & RT_OUT_COMPOSITE))
                p.composite[pix] = make_uchar4((unsigned char)unorm8(a.x / a.w), (unsigned char)unorm8(a.y / a.w),
                                               (unsigned char)unorm8(a.z / a.w), 255);
        }
    }
    // ---- append survivors to the other queue: one atomic per warp, order within the warp kept
    unsigned m = __ballot_sync(full, cont);
    if (m) {
        uint32_t base = 0;
        int leader = __ffs(m) - 1;
        if ((int)lane == leader) base = atomicAdd(&w.ctrl->n[qi ^ 1], (unsigned int)__popc(m));
        base = __shfl_sync(full, base, leader);
        if (cont) w.queue[qi ^ 1][base + __popc(m & ((1u << lane) - 1u))] = path;
    }
}
