// rt_bvh.cpp — BVH builders that emit the reference's node format
// (`BvhNode` 48 B + `triangle_indices`, shared/src/lib.rs:155-161), i.e. the
// output contract of src/bvh.rs that the hot path consumes.
//
//  * RT_BVH_CHUNKED restates BvhBuilder::build_chunked + build_simple_top_level_bvh
//    (src/bvh.rs:154-247) exactly: it is fully deterministic and uses no
//    third-party code, so node-for-node equality with the reference is expected.
//  * RT_BVH_STANDARD emits the *format* of build_standard (src/bvh.rs:125-151,
//    298-375): pre-order nodes, left child = parent + 1, exactly one triangle per
//    leaf, leaf AABB = the triangle's box, indices in left-first leaf order.  The
//    reference delegates the topology to the `bvh` crate 0.11.0 (not vendored, no
//    test pins it); here a binned SAH chooses the splits.  Closest-hit results do
//    not depend on the topology.
//  * RT_BVH_SMALL_LEAF: same format with <=4 triangles per leaf, for any size.
#include <math.h>
#include <string.h>
#include <algorithm>
#include <vector>

#include "../../include/rtb200_host.h"

namespace {

struct Box {
    float mn[3], mx[3];
    void reset() { mn[0] = mn[1] = mn[2] = INFINITY; mx[0] = mx[1] = mx[2] = -INFINITY; }
    void grow(const Box& o) { for (int i = 0; i < 3; i++) { mn[i] = fminf(mn[i], o.mn[i]); mx[i] = fmaxf(mx[i], o.mx[i]); } }
    void grow_pt(const float* p) { for (int i = 0; i < 3; i++) { mn[i] = fminf(mn[i], p[i]); mx[i] = fmaxf(mx[i], p[i]); } }
    float area() const {
        float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
        if (!(dx >= 0.0f)) return 0.0f;
        return 2.0f * (dx * dy + dy * dz + dz * dx);
    }
};

RtAabb to_aabb(const Box& b) {
    RtAabb a;
    memcpy(a.min, b.mn, 12); memcpy(a.max, b.mx, 12);
    a._padding0 = a._padding1 = 0.0f;
    return a;
}
RtBvhNode make_leaf(const RtAabb& b, uint32_t start, uint32_t count) { // BvhNode::leaf, shared/src/lib.rs:806-814
    RtBvhNode n; n.bounds = b; n.left_child = n.right_child = RT_INVALID_INDEX; n.triangle_start = start; n.triangle_count = count; return n;
}
RtBvhNode make_internal(const RtAabb& b, uint32_t l, uint32_t r) { // BvhNode::internal, :817-825
    RtBvhNode n; n.bounds = b; n.left_child = l; n.right_child = r; n.triangle_start = 0; n.triangle_count = 0; return n;
}
bool node_is_leaf(const RtBvhNode& n) { return n.left_child == RT_INVALID_INDEX && n.right_child == RT_INVALID_INDEX; } // :828-830

// ---- build_chunked, src/bvh.rs:154-247 ---------------------------------------
void build_chunked(const RtTriangle* tris, uint64_t n, const RtVertex* verts, std::vector<RtBvhNode>& nodes,
                   std::vector<uint32_t>& indices) {
    uint64_t per_leaf = std::max<uint64_t>(n / 10000, 32); // :156
    std::vector<RtBvhNode> level;
    for (uint64_t base = 0; base < n; base += per_leaf) { // triangles.chunks(per_leaf)
        uint64_t cnt = std::min<uint64_t>(per_leaf, n - base);
        RtAabb box; rt_aabb_empty(&box);
        for (uint64_t i = 0; i < cnt; i++) { RtAabb tb; rt_triangle_aabb(&tris[base + i], verts, &tb); rt_aabb_union(&box, &tb, &box); }
        uint32_t start = (uint32_t)indices.size();
        for (uint64_t i = 0; i < cnt; i++) indices.push_back((uint32_t)(base + i));
        level.push_back(make_leaf(box, start, (uint32_t)cnt));
    }
    if (level.size() <= 1) { nodes = level; return; } // :183-185
    // build_simple_top_level_bvh, :192-247
    std::vector<RtBvhNode> out;
    while (level.size() > 1) {
        std::vector<RtBvhNode> next;
        for (size_t i = 0; i < level.size(); i += 2) {
            if (i + 1 < level.size()) {
                RtAabb u; rt_aabb_union(&level[i].bounds, &level[i + 1].bounds, &u);
                uint32_t l = (uint32_t)out.size();
                out.push_back(level[i]); out.push_back(level[i + 1]);
                next.push_back(make_internal(u, l, l + 1));
            } else { // odd node out: single-child parent (:215-219)
                uint32_t idx = (uint32_t)out.size();
                out.push_back(level[i]);
                next.push_back(make_internal(level[i].bounds, idx, RT_INVALID_INDEX));
            }
        }
        level.swap(next);
    }
    if (!level.empty()) out.push_back(level[0]);
    std::reverse(out.begin(), out.end()); // :229
    uint32_t offset = (uint32_t)out.size() - 1;
    for (auto& nd : out)
        if (!node_is_leaf(nd)) {
            if (nd.left_child != RT_INVALID_INDEX) nd.left_child = offset - nd.left_child;
            if (nd.right_child != RT_INVALID_INDEX) nd.right_child = offset - nd.right_child;
        }
    nodes.swap(out);
}

// ---- binned SAH, pre-order output --------------------------------------------
struct Sah {
    const RtTriangle* tris; const RtVertex* verts;
    std::vector<Box> tbox; std::vector<float> cen; // 3 per triangle
    std::vector<uint32_t> perm;
    std::vector<RtBvhNode>* nodes; std::vector<uint32_t>* indices;
    uint32_t max_leaf;
    static constexpr int BINS = 16;

    uint32_t build(uint32_t lo, uint32_t hi, int depth) {
        Box bounds, cb; bounds.reset(); cb.reset();
        for (uint32_t i = lo; i < hi; i++) { bounds.grow(tbox[perm[i]]); cb.grow_pt(&cen[(size_t)perm[i] * 3]); }
        uint32_t me = (uint32_t)nodes->size();
        uint32_t n = hi - lo;
        if (n <= max_leaf) {
            uint32_t start = (uint32_t)indices->size();
            for (uint32_t i = lo; i < hi; i++) indices->push_back(perm[i]);
            nodes->push_back(make_leaf(to_aabb(bounds), start, n));
            return me;
        }
        nodes->push_back(make_internal(to_aabb(bounds), 0, 0));
        uint32_t mid = lo + n / 2;
        bool done = false;
        if (depth < 32) { // deeper than that: plain median keeps the tree under the 60-level traversal limit
            float best = INFINITY; int best_axis = -1, best_bin = -1;
            for (int ax = 0; ax < 3; ax++) {
                float ext = cb.mx[ax] - cb.mn[ax];
                if (!(ext > 0.0f)) continue;
                Box bb[BINS]; uint32_t bc[BINS];
                for (int b = 0; b < BINS; b++) { bb[b].reset(); bc[b] = 0; }
                float scale = (float)BINS / ext;
                for (uint32_t i = lo; i < hi; i++) {
                    int b = (int)((cen[(size_t)perm[i] * 3 + ax] - cb.mn[ax]) * scale);
                    b = b < 0 ? 0 : (b >= BINS ? BINS - 1 : b);
                    bb[b].grow(tbox[perm[i]]); bc[b]++;
                }
                float la[BINS - 1]; uint32_t lc[BINS - 1];
                Box acc; acc.reset(); uint32_t cnt = 0;
                for (int b = 0; b < BINS - 1; b++) { acc.grow(bb[b]); cnt += bc[b]; la[b] = acc.area(); lc[b] = cnt; }
                acc.reset(); cnt = 0;
                for (int b = BINS - 1; b > 0; b--) {
                    acc.grow(bb[b]); cnt += bc[b];
                    if (lc[b - 1] == 0 || cnt == 0) continue;
                    float cost = la[b - 1] * (float)lc[b - 1] + acc.area() * (float)cnt;
                    if (cost < best) { best = cost; best_axis = ax; best_bin = b; }
                }
            }
            if (best_axis >= 0) {
                float ext = cb.mx[best_axis] - cb.mn[best_axis], scale = (float)BINS / ext, mn = cb.mn[best_axis];
                int ax = best_axis, bin = best_bin;
                auto it = std::partition(perm.begin() + lo, perm.begin() + hi, [&](uint32_t t) {
                    int b = (int)((cen[(size_t)t * 3 + ax] - mn) * scale);
                    b = b < 0 ? 0 : (b >= BINS ? BINS - 1 : b);
                    return b < bin;
                });
                uint32_t m = (uint32_t)(it - perm.begin());
                if (m > lo && m < hi) { mid = m; done = true; }
            }
        }
        if (!done) { // object median along the widest centroid axis
            int ax = 0;
            float e0 = cb.mx[0] - cb.mn[0], e1 = cb.mx[1] - cb.mn[1], e2 = cb.mx[2] - cb.mn[2];
            if (e1 > e0) { ax = 1; e0 = e1; }
            if (e2 > e0) ax = 2;
            std::nth_element(perm.begin() + lo, perm.begin() + mid, perm.begin() + hi, [&](uint32_t a, uint32_t b) {
                float ca = cen[(size_t)a * 3 + ax], cbv = cen[(size_t)b * 3 + ax];
                return ca < cbv || (ca == cbv && a < b);
            });
        }
        uint32_t l = build(lo, mid, depth + 1);
        uint32_t r = build(mid, hi, depth + 1);
        (*nodes)[me].left_child = l; (*nodes)[me].right_child = r;
        return me;
    }
};

void build_sah(const RtTriangle* tris, uint64_t n, const RtVertex* verts, uint32_t max_leaf, std::vector<RtBvhNode>& nodes,
               std::vector<uint32_t>& indices) {
    Sah s;
    s.tris = tris; s.verts = verts; s.max_leaf = max_leaf; s.nodes = &nodes; s.indices = &indices;
    s.tbox.resize(n); s.cen.resize(n * 3); s.perm.resize(n);
    for (uint64_t i = 0; i < n; i++) {
        RtAabb a; rt_triangle_aabb(&tris[i], verts, &a);
        memcpy(s.tbox[i].mn, a.min, 12); memcpy(s.tbox[i].mx, a.max, 12);
        for (int k = 0; k < 3; k++) s.cen[i * 3 + k] = 0.5f * (a.min[k] + a.max[k]);
        s.perm[i] = (uint32_t)i;
    }
    nodes.reserve(2 * n); indices.reserve(n);
    s.build(0, (uint32_t)n, 0);
}

} // namespace

extern "C" {

uint64_t rt_bvh_max_nodes(uint64_t n) { return n ? 2 * n + 64 : 1; }

int rt_bvh_build(const RtTriangle* tris, uint64_t n, const RtVertex* verts, uint64_t n_verts, int strategy, RtBvhNode* nodes_out,
                 uint32_t* n_nodes, uint32_t* indices_out, uint32_t* n_indices) {
    if (!nodes_out || !n_nodes || !n_indices || (n && (!tris || !verts || !indices_out))) return -1;
    if (n > 0x7FFFFFFFull) return -1;
    for (uint64_t i = 0; i < n; i++)
        if (tris[i].v0_index >= n_verts || tris[i].v1_index >= n_verts || tris[i].v2_index >= n_verts) return -2;
    std::vector<RtBvhNode> nodes;
    std::vector<uint32_t> indices;
    if (n == 0) { // src/bvh.rs:105-114: one empty leaf
        RtAabb e; rt_aabb_empty(&e);
        nodes.push_back(make_leaf(e, 0, 0));
    } else {
        if (strategy == RT_BVH_AUTO) strategy = n > 100000 ? RT_BVH_CHUNKED : RT_BVH_STANDARD; // :117-121
        if (strategy == RT_BVH_CHUNKED) build_chunked(tris, n, verts, nodes, indices);
        else if (strategy == RT_BVH_STANDARD) build_sah(tris, n, verts, 1, nodes, indices);
        else if (strategy == RT_BVH_SMALL_LEAF) build_sah(tris, n, verts, 4, nodes, indices);
        else return -1;
    }
    memcpy(nodes_out, nodes.data(), nodes.size() * sizeof(RtBvhNode));
    if (!indices.empty()) memcpy(indices_out, indices.data(), indices.size() * 4);
    *n_nodes = (uint32_t)nodes.size();
    *n_indices = (uint32_t)indices.size();
    return 0;
}

} // extern "C"

// A BVH whose child links form a cycle reachable from the root makes the reference traversal
// (shader/src/bvh.rs:40-85: pop, test, push both children) run forever as long as the boxes on the
// cycle are hit; on this device that is a hung context, so rt_update_scene_metadata refuses such an upload.  A DAG (a node
// reachable twice without a cycle) terminates and is served by the LITERAL policy like in the reference.
extern "C" int rt_bvh_has_cycle(const RtBvhNode* nodes, uint32_t n_nodes) {
    if (n_nodes == 0 || !nodes) return 0;
    std::vector<uint8_t> state(n_nodes, 0); // 0 new, 1 on the DFS path, 2 finished
    struct Frame { uint32_t node; uint8_t next; };
    std::vector<Frame> path;
    path.push_back({0u, 0});
    state[0] = 1;
    while (!path.empty()) {
        Frame& f = path.back();
        const RtBvhNode& nd = nodes[f.node];
        uint32_t child = RT_INVALID_INDEX;
        if (nd.left_child != RT_INVALID_INDEX) { // internal node (bvh.rs:60)
            while (f.next < 2 && child == RT_INVALID_INDEX) {
                const uint32_t c = f.next == 0 ? nd.left_child : nd.right_child;
                f.next++;
                if (c != RT_INVALID_INDEX && c < n_nodes) child = c;
            }
        }
        if (child == RT_INVALID_INDEX) { state[f.node] = 2; path.pop_back(); continue; }
        if (state[child] == 1) return 1;
        if (state[child] == 0) { state[child] = 1; path.push_back({child, 0}); }
    }
    return 0;
}

