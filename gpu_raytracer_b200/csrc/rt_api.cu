// rt_api.cu — the C ABI of librtb200.so (include/rtb200.h): context, buffer
// manager, upload-time validation + device re-layout, dispatch, read-back.
//
// Mirrors the reference's dispatch surface: BufferManager (src/buffers.rs) and
// ComputeRenderer (src/compute.rs) with ProgressiveState (src/renderer.rs:821-855).
// There is no CPU fallback anywhere in this file: every render entry point
// launches rt_trace_shade / rt_path_trace on the context's CUDA stream.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

#include "../../include/rtb200_host.h"
#include "rt_kernels.cuh"
#include "rt_lbvh.cuh"
#include "rt_wave.cuh"

namespace {

thread_local std::string g_create_error;

template <class T>
struct DevBuf {
    T* ptr = nullptr;
    size_t capacity = 0; // elements
    ~DevBuf() { if (ptr) cudaFree(ptr); }
    // returns 1 if reallocated, 0 if not, <0 cuda error
    // The new buffer is allocated before the old one is released: a failed growth leaves the
    // previous contents (and capacity) in place.  Contents are NOT carried over.
    int reserve(size_t n) {
        if (n <= capacity && ptr) return 0;
        size_t cap = n ? n : 1;
        T* np = nullptr;
        cudaError_t e = cudaMalloc((void**)&np, cap * sizeof(T));
        if (e != cudaSuccess) { cudaGetLastError(); return -1; }
        if (ptr) cudaFree(ptr);
        ptr = np;
        capacity = cap;
        return 1;
    }
};

struct WideBvh { // host-side result of validation + node re-layout
    bool ordered_ok = false;
    std::string why_not;
    std::vector<uint32_t> index;    // uploaded node -> compact internal-node index (0xFFFFFFFF: leaf / unreachable)
    uint32_t n_internal = 0;
    std::vector<uint2> leaves;      // (start, effective count) of every reachable non-empty leaf
    std::vector<uint32_t> leaf_nodes; // uploaded node index of each entry of `leaves`
    bool nested_ok = false;         // every child box inside its parent's box, no NaN (REBUILT policy needs it)
    uint32_t max_leaf = 0;          // largest effective leaf
    uint32_t root_ref = RT_CHILD_ABSENT;
    float root_min[3] = {0, 0, 0}, root_max[3] = {0, 0, 0};
    uint32_t max_depth = 0;
};

} // namespace

struct RtContext {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    std::string err;
    uint64_t launches = 0;
    int last_traversal = 0;

    // BufferManager state (src/buffers.rs:13-45)
    DevBuf<uint32_t> meta, meta_back; // live scene / staging buffer of the next upload (swapped on success)
    DevBuf<uint32_t> wide_index;      // uploaded node -> compact internal index (relayout_wide_nodes)
    size_t meta_len = 0; // u32
    size_t scene_metadata_capacity_bytes = 0;
    RtSceneMetadataOffsets offsets{};
    bool have_meta = false;
    DevBuf<RtTriangle> tris;
    size_t n_tris = 0;
    size_t tri_buffers = 1; // reference starts with one bound buffer (src/buffers.rs:92-99)
    uint32_t tpb = RT_MAX_TRIANGLES_PER_BUFFER; // BufferManager.triangles_per_buffer (src/buffers.rs:49-53, 73)
    DevBuf<RtMaterial> mats;
    size_t n_mats = 0, mats_capacity = 64;
    DevBuf<RtTextureInfo> textures;
    size_t n_textures = 0, textures_capacity = 64;
    DevBuf<uint32_t> texdata;
    size_t texdata_len = 0, texdata_capacity = 256 * 1024;
    bool meta_dirty = true, tris_dirty = true, mats_dirty = true, tex_dirty = true, texdata_dirty = true;
    size_t last_spheres = 0, last_lights = 0, last_nodes = 0, last_indices = 0, last_vertices = 0,
           last_tris = 0, last_mats = 0, last_textures = 0, last_texdata = 0;

    // re-laid-out scene for the ORDERED policy
    WideBvh wide;
    DevBuf<float4> wnodes;
    DevBuf<float4> wtris;
    DevBuf<uint2> wleaves;
    DevBuf<uint32_t> dflag;
    DevBuf<unsigned long long> dcounters;
    bool count_next = false; // next launch is instrumented
    bool relayout_dirty = true;
    bool relayout_ok = false;
    // internally rebuilt LBVH (REBUILT policy)
    DevBuf<uint32_t> wleaf_nodes, leaf_of_pos, lb_vals, lb_vals_sorted, lb_parent_i, lb_parent_l, lb_visits, lb_bounds;
    DevBuf<unsigned long long> lb_keys, lb_keys_sorted;
    DevBuf<unsigned char> lb_temp;
    DevBuf<uint2> lb_children, lb_range;
    DevBuf<float4> pbox, lb_nbox, bnodes, btris;
    DevBuf<uint4> bqnodes;   // 32-byte quantised copy of bnodes (wave engine)
    DevBuf<float4> cert_box; // [2 * uploaded node]: boxes of the uploaded leaves (REBUILT certification)
    DevBuf<float> lb_grid;   // its grid: origin[3], cell[3]
    float bq_grid[6] = {0, 0, 0, 1, 1, 1};
    uint32_t lbvh_depth = 0; // levels of the internal tree (picks the wave engine's stack form)
    bool rebuilt_ok = false;
    uint32_t broot_ref = RT_CHILD_ABSENT;
    float broot_min[3] = {0, 0, 0}, broot_max[3] = {0, 0, 0};
    float4 rb_box[2] = {};     // read-back of the internal tree's root box / depth (build_rebuilt -> finish_rebuilt)
    unsigned int rb_depth = 0;
    float rebuilt_far = 0.0f; // rays whose origin has a coordinate beyond this take the literal policy (rt_lbvh.cuh)
    float last_build_ms = -1.0f;
    bool last_was_wavefront = false;
    uint32_t last_wave_max_depth = 0, last_wave_seed = 0;
    // streaming wavefront engine (rt_wave.cuh)
    DevBuf<float4> wv_state[2], wv_lbuf;
    DevBuf<uint32_t> wv_fb;  // per path: queue positions handed to wv_fallback
    DevBuf<uint2> wv_hit[2];
    DevBuf<WvCtrl> wv_ctrl;
    std::vector<cudaStream_t> wv_streams; // extra streams of the sub-frames (launch_wave)
    std::vector<cudaEvent_t> wv_join;
    cudaEvent_t wv_fork = nullptr;
    std::vector<cudaEvent_t> wv_events;  // one before the first and one after every launch of the last wave frame
    std::vector<int> wv_event_stage;     // stage of the launch that ends at event i+1: 0 generate, 1 extend, 2 shade, 3 resolve

    // outputs (src/renderer.rs:452-475) + progressive state (src/renderer.rs:821-855)
    uint32_t width = 0, height = 0;
    DevBuf<float4> accum;
    DevBuf<RtHitRecord> hits;
    DevBuf<uchar4> chan[3];
    DevBuf<uchar4> composite;
    uint32_t tiles_x = 0, tiles_y = 0, tiles_per_frame = 0, current_tile = 0;
    uchar4* peer_composite = nullptr; // rank 0's composite through peer memory (fused multi-GPU epilogue)
    bool peer_is_ipc = false;
    uint32_t stiles_x = 0, stiles_y = 0; // RT_SPLIT_TILE^2 split tiles of the fused frame / multi-GPU split
    DevBuf<uint32_t> tile_list, copy_list; // rt_set_tile_list assignment; scratch list of rt_(un)pack_tile_list
    std::vector<uint32_t> tile_list_host;
    uint32_t tile_list_n = 0;
    bool needs_recompute = true, is_progressive = false;
};

namespace {

int fail(RtContext* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf; else g_create_error = buf;
    return code;
}
#define CK(c, call)                                                                         \
    do {                                                                                    \
        cudaError_t e_ = (call);                                                            \
        if (e_ != cudaSuccess)                                                              \
            return fail(c, RT_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_));    \
    } while (0)

// Walks the uploaded BVH exactly as the reference's DFS would if every box were
// hit (shader/src/bvh.rs:40-85) and emits the two-child node layout.  The ORDERED
// policy is enabled only if that walk is a tree (no node reached twice), stays
// shallower than the reference's 63-entry stack limit (so the `stack_ptr < 63`
// quirk can never trigger), and visits leaves in non-overlapping, increasing
// triangle_start order (so "first in DFS order" == "smaller position").
void build_wide(const RtBvhNode* nodes, uint32_t n_nodes, uint32_t n_indices, WideBvh& w) {
    w = WideBvh();
    if (n_nodes == 0) { w.why_not = "no BVH nodes (brute-force path)"; return; }
    auto eff_count = [&](const RtBvhNode& nd) -> uint32_t {
        if (nd.triangle_start >= n_indices) return 0u; // bvh.rs:109-111 break
        uint64_t room = (uint64_t)n_indices - nd.triangle_start;
        return (uint32_t)(nd.triangle_count < room ? nd.triangle_count : room);
    };
    auto is_leaf = [&](const RtBvhNode& nd) { return nd.left_child == RT_INVALID_INDEX; }; // bvh.rs:60
    auto valid_child = [&](uint32_t c) { return c != RT_INVALID_INDEX && c < n_nodes; };   // bvh.rs:45
    std::vector<uint32_t>& wide_index = w.index;
    wide_index.assign(n_nodes, 0xFFFFFFFFu);
    std::vector<uint8_t> seen(n_nodes, 0);
    struct Item { uint32_t node, depth; };
    std::vector<Item> stack;
    stack.push_back({0u, 0u});
    uint64_t prev_end = 0;
    uint32_t n_internal = 0;
    bool nested = true;
    auto box_ok = [](const RtAabb& a) {
        for (int k = 0; k < 3; k++) if (a.min[k] != a.min[k] || a.max[k] != a.max[k]) return false;
        return true;
    };
    auto inside = [](const RtAabb& c, const RtAabb& p) {
        for (int k = 0; k < 3; k++) if (!(c.min[k] >= p.min[k] && c.max[k] <= p.max[k])) return false;
        return true;
    };
    // pass 1: DFS in reference order; number internal nodes; check leaves
    while (!stack.empty()) {
        Item it = stack.back();
        stack.pop_back();
        const RtBvhNode& nd = nodes[it.node];
        if (seen[it.node]) { w.why_not = "BVH is not a tree (node reachable twice)"; return; }
        seen[it.node] = 1;
        if (it.depth > w.max_depth) w.max_depth = it.depth;
        if (it.depth >= 60) { w.why_not = "BVH deeper than 60 levels"; return; }
        nested = nested && box_ok(nd.bounds);
        if (is_leaf(nd)) {
            uint32_t cnt = eff_count(nd);
            if (cnt) {
                if (nd.triangle_start < prev_end) { w.why_not = "leaf ranges not increasing in DFS order"; return; }
                prev_end = (uint64_t)nd.triangle_start + cnt;
                if (nd.triangle_start & RT_LEAF_FLAG) { w.why_not = "triangle_start >= 2^31"; return; }
                w.leaves.push_back(make_uint2(nd.triangle_start, cnt));
                w.leaf_nodes.push_back(it.node);
                if (cnt > w.max_leaf) w.max_leaf = cnt;
            }
        } else {
            wide_index[it.node] = n_internal++;
            if (valid_child(nd.right_child)) { stack.push_back({nd.right_child, it.depth + 1}); nested = nested && inside(nodes[nd.right_child].bounds, nd.bounds); }
            if (valid_child(nd.left_child)) { stack.push_back({nd.left_child, it.depth + 1}); nested = nested && inside(nodes[nd.left_child].bounds, nd.bounds); }
        }
    }
    w.n_internal = n_internal;
    const RtBvhNode& root = nodes[0];
    memcpy(w.root_min, root.bounds.min, 12);
    memcpy(w.root_max, root.bounds.max, 12);
    if (is_leaf(root)) w.root_ref = eff_count(root) ? (RT_LEAF_FLAG | root.triangle_start) : RT_CHILD_ABSENT;
    else w.root_ref = wide_index[0];
    w.ordered_ok = true;
    w.nested_ok = nested;
}

// ---- device re-layout kernels -------------------------------------------------
// One thread per uploaded node: internal nodes (compact index from the host's validation walk) become one
// 64-byte two-child node each: (L box, R box, child refs), boxes copied bit for bit from the uploaded children.
__global__ void relayout_wide_nodes(const uint32_t* __restrict__ meta, RtSceneMetadataOffsets off,
                                    const uint32_t* __restrict__ wide_index, float4* __restrict__ wnodes) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n_nodes = off.bvh_nodes_count, n_indices = off.triangle_indices_count;
    if (i >= n_nodes) return;
    const uint32_t wi = wide_index[i];
    if (wi == 0xFFFFFFFFu) return;
    const uint32_t* nb = meta + off.bvh_nodes_offset;
    const float inf = __int_as_float(0x7f800000);
    float4 q[4];
    float bx[2][6];
    uint32_t ref[2];
    for (int k = 0; k < 2; k++) {
        const uint32_t c = nb[(uint64_t)i * 12 + 8 + k];
        for (int a = 0; a < 3; a++) { bx[k][a] = inf; bx[k][3 + a] = -inf; }
        ref[k] = RT_CHILD_ABSENT;
        if (c != RT_INVALID_INDEX && c < n_nodes) { // bvh.rs:45
            const uint32_t* cn = nb + (uint64_t)c * 12;
            for (int a = 0; a < 3; a++) { bx[k][a] = __uint_as_float(cn[a]); bx[k][3 + a] = __uint_as_float(cn[4 + a]); }
            if (cn[8] == RT_INVALID_INDEX) { // leaf (bvh.rs:60): effective count per bvh.rs:109-111
                const uint32_t start = cn[10], count = cn[11];
                const uint32_t eff = start >= n_indices ? 0u : (count < n_indices - start ? count : n_indices - start);
                ref[k] = eff ? (RT_LEAF_FLAG | start) : RT_CHILD_ABSENT;
            } else ref[k] = wide_index[c];
        }
    }
    q[0] = make_float4(bx[0][0], bx[0][1], bx[0][2], bx[0][3]);
    q[1] = make_float4(bx[0][4], bx[0][5], bx[1][0], bx[1][1]);
    q[2] = make_float4(bx[1][2], bx[1][3], bx[1][4], bx[1][5]);
    q[3] = make_float4(__uint_as_float(ref[0]), __uint_as_float(ref[1]), 0.0f, 0.0f);
    for (int k = 0; k < 4; k++) wnodes[(uint64_t)wi * 4 + k] = q[k];
}


// One thread per triangle_indices position: gather the indexed triangle into
// (v0, e1, e2, material, index).  e1 = v1 - v0 and e2 = v2 - v0 are the same two
// IEEE subtractions intersection.rs:102-103 performs per test, hoisted to upload.
__global__ void relayout_triangles(const uint32_t* __restrict__ meta, RtSceneMetadataOffsets off,
                                   const RtTriangle* __restrict__ tris, uint64_t n_tris, uint32_t tpb, uint32_t n_buffers,
                                   float4* __restrict__ out, float4* __restrict__ pbox, uint32_t* __restrict__ flag) {
    uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= off.triangle_indices_count) return;
    uint32_t ti = meta[(uint64_t)off.triangle_indices_offset + pos];
    float4 a = make_float4(0, 0, 0, 0), b = a, c = a;
    float4 blo = a, bhi = a;
    c.z = __uint_as_float(ti);
    // triangle_access.rs:26-48 over the bound buffers (resolve_bound_triangle): an index that resolves to the
    // zero-filled tail of a buffer is a degenerate triangle the Möller–Trumbore test always rejects (a == 0),
    // so it is stored like an invalid one (all zero: never hit)
    uint64_t gi;
    bool valid = resolve_bound_triangle(ti, tpb, tpb, n_buffers, &gi) && gi < n_tris;
    if (valid) {
        RtTriangle t = tris[gi];
        if (t.v0_index >= off.vertices_count || t.v1_index >= off.vertices_count || t.v2_index >= off.vertices_count) {
            atomicOr(flag, 1u); // the reference would read out of bounds
        } else {
            const uint32_t* vb = meta + off.vertices_offset;
            float v0x = __uint_as_float(vb[(uint64_t)t.v0_index * 3]), v0y = __uint_as_float(vb[(uint64_t)t.v0_index * 3 + 1]),
                  v0z = __uint_as_float(vb[(uint64_t)t.v0_index * 3 + 2]);
            float v1x = __uint_as_float(vb[(uint64_t)t.v1_index * 3]), v1y = __uint_as_float(vb[(uint64_t)t.v1_index * 3 + 1]),
                  v1z = __uint_as_float(vb[(uint64_t)t.v1_index * 3 + 2]);
            float v2x = __uint_as_float(vb[(uint64_t)t.v2_index * 3]), v2y = __uint_as_float(vb[(uint64_t)t.v2_index * 3 + 1]),
                  v2z = __uint_as_float(vb[(uint64_t)t.v2_index * 3 + 2]);
            a = make_float4(v0x, v0y, v0z, v1x - v0x);
            b = make_float4(v1y - v0y, v1z - v0z, v2x - v0x, v2y - v0y);
            c.x = v2z - v0z;
            c.y = __uint_as_float(t.material_id);
            blo = make_float4(fminf(fminf(v0x, v1x), v2x), fminf(fminf(v0y, v1y), v2y), fminf(fminf(v0z, v1z), v2z), 0.0f);
            bhi = make_float4(fmaxf(fmaxf(v0x, v1x), v2x), fmaxf(fmaxf(v0y, v1y), v2y), fmaxf(fmaxf(v0z, v1z), v2z), 0.0f);
        }
    }
    c.w = __uint_as_float(1u); // `remaining`, fixed up per leaf below
    out[(uint64_t)pos * 3] = a;
    out[(uint64_t)pos * 3 + 1] = b;
    out[(uint64_t)pos * 3 + 2] = c;
    if (pbox) { pbox[(uint64_t)pos * 2] = blo; pbox[(uint64_t)pos * 2 + 1] = bhi; }
}
// One warp per leaf: remaining[start + i] = count - i.
// cert_box (REBUILT policy only): the leaf's uploaded box as two aligned float4 at [2 * node], the same
// words the certification would otherwise gather one by one from the metadata buffer.
__global__ void relayout_leaf_counts(const uint2* __restrict__ leaves, const uint32_t* __restrict__ leaf_nodes,
                                     uint32_t n_leaves, float4* __restrict__ out, uint32_t* __restrict__ leaf_of_pos,
                                     const uint32_t* __restrict__ meta, uint64_t meta_len, uint32_t nodes_offset,
                                     float4* __restrict__ cert_box) {
    uint32_t leaf = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (leaf >= n_leaves) return;
    uint2 l = leaves[leaf];
    uint32_t node = leaf_nodes[leaf];
    if (cert_box && lane < 2u) {
        uint64_t o = (uint64_t)nodes_offset + (uint64_t)node * 12 + lane * 4u;
        float v[3];
        for (int k = 0; k < 3; k++) v[k] = o + k < meta_len ? __uint_as_float(meta[o + k]) : 0.0f;
        cert_box[(uint64_t)node * 2 + lane] = make_float4(v[0], v[1], v[2], 0.0f);
    }
    for (uint32_t i = lane; i < l.y; i += 32u) {
        reinterpret_cast<uint32_t*>(out + (uint64_t)(l.x + i) * 3 + 2)[3] = l.y - i;
        leaf_of_pos[l.x + i] = node;
    }
}

// ---- multi-GPU tile gather helpers ---------------------------------------------
// pack: this context's tiles (i % stride == phase) -> tile-major compact buffer
// [local tile][128][128]; unpack: the inverse, into the full frame.  T = uchar4 / float4.
template <class T, bool PACK>
__global__ void tile_copy(T* __restrict__ frame, T* __restrict__ compact, uint32_t width, uint32_t height, uint32_t tiles_x,
                          uint32_t n_tiles_total, uint32_t stride, uint32_t phase, const uint32_t* __restrict__ tile_list) {
    uint32_t tile_local = blockIdx.y;
    uint32_t tile = phase + tile_local * stride;
    if (tile >= n_tiles_total) return;
    if (tile_list) tile = tile_list[tile];
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; // 0..4095
    uint32_t x = (tile % tiles_x) * RT_SPLIT_TILE + (i & 63u), y = (tile / tiles_x) * RT_SPLIT_TILE + (i >> 6);
    size_t ci = (size_t)tile_local * 4096u + i;
    if (x < width && y < height) {
        if (PACK) compact[ci] = frame[(size_t)y * width + x];
        else frame[(size_t)y * width + x] = compact[ci];
    } else if (PACK) {
        T z; memset(&z, 0, sizeof z); compact[ci] = z;
    }
}

} // namespace

// =============================================================== C ABI ======
extern "C" {

int rt_create(int device, RtContext** out) {
    if (!out) return fail(nullptr, RT_ERR_INVALID_ARGUMENT, "rt_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, RT_ERR_CUDA, "rt_create: no CUDA device (%s); this library has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= count) return fail(nullptr, RT_ERR_INVALID_ARGUMENT, "rt_create: device %d of %d", device, count);
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(nullptr, RT_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    RtContext* c = new RtContext();
    c->device = device;
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreate(&c->ev0)) != cudaSuccess || (e = cudaEventCreate(&c->ev1)) != cudaSuccess) {
        fail(nullptr, RT_ERR_CUDA, "rt_create: %s", cudaGetErrorString(e));
        delete c;
        return RT_ERR_CUDA;
    }
    // BufferManager::new default capacities, src/buffers.rs:66-84
    c->scene_metadata_capacity_bytes = 64 * sizeof(RtSphere) + 32 * sizeof(RtLight) + 256 * sizeof(RtBvhNode) +
                                       1024 * sizeof(uint32_t) + 1024 * sizeof(RtVertex);
    if (c->meta.reserve(c->scene_metadata_capacity_bytes / 4) < 0 || c->mats.reserve(c->mats_capacity) < 0 ||
        c->textures.reserve(c->textures_capacity) < 0 || c->texdata.reserve(c->texdata_capacity) < 0 ||
        c->tris.reserve(64) < 0 || c->dflag.reserve(1) < 0 || c->dcounters.reserve(32) < 0) {
        fail(nullptr, RT_ERR_CUDA, "rt_create: cudaMalloc failed");
        delete c;
        return RT_ERR_CUDA;
    }
    // wgpu zero-initialises new buffers; the shader's materials.len() is the bound capacity, so ids in
    // [uploaded count, capacity) index zero materials (src/buffers.rs:100-106, src/renderer.rs:698)
    if (cudaMemsetAsync(c->mats.ptr, 0, c->mats_capacity * sizeof(RtMaterial), c->stream) != cudaSuccess) {
        fail(nullptr, RT_ERR_CUDA, "rt_create: cudaMemset failed");
        delete c;
        return RT_ERR_CUDA;
    }
    *out = c;
    return RT_OK;
}

void rt_destroy(RtContext* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (cudaStream_t st : c->wv_streams) cudaStreamSynchronize(st);
    rt_close_peer_frame(c); // before the stream it synchronises goes away
    for (cudaStream_t st : c->wv_streams) cudaStreamDestroy(st);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    for (cudaEvent_t e : c->wv_events) cudaEventDestroy(e);
    for (cudaEvent_t e : c->wv_join) cudaEventDestroy(e);
    if (c->wv_fork) cudaEventDestroy(c->wv_fork);
    delete c;
}

const char* rt_last_error(const RtContext* c) { return c ? c->err.c_str() : g_create_error.c_str(); }

int rt_resize(RtContext* c, uint32_t w, uint32_t h) {
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (w == 0 || h == 0 || w > 65535u || h > 65535u) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_resize: %ux%u", w, h);
    CK(c, cudaSetDevice(c->device));
    size_t n = (size_t)w * h;
    if (c->accum.reserve(n) < 0 || c->hits.reserve(n) < 0 || c->composite.reserve(n) < 0 || c->chan[0].reserve(n) < 0 ||
        c->chan[1].reserve(n) < 0 || c->chan[2].reserve(n) < 0)
        return fail(c, RT_ERR_CUDA, "rt_resize: cudaMalloc failed for %ux%u", w, h);
    c->width = w; c->height = h;
    CK(c, cudaMemsetAsync(c->accum.ptr, 0, n * sizeof(float4), c->stream));
    CK(c, cudaMemsetAsync(c->hits.ptr, 0, n * sizeof(RtHitRecord), c->stream));
    CK(c, cudaMemsetAsync(c->composite.ptr, 0, n * 4, c->stream));
    for (int i = 0; i < 3; i++) CK(c, cudaMemsetAsync(c->chan[i].ptr, 0, n * 4, c->stream));
    // ProgressiveState::resize, src/renderer.rs:838-848
    rt_tile_count(w, h, RT_TILE_SIZE, &c->tiles_x, &c->tiles_y);
    rt_tile_count(w, h, RT_SPLIT_TILE, &c->stiles_x, &c->stiles_y);
    c->tile_list_n = 0; c->tile_list_host.clear(); // tile numbers belong to the old frame size
    c->tiles_per_frame = rt_tiles_per_frame(c->tiles_x * c->tiles_y);
    c->needs_recompute = true; c->current_tile = 0; c->is_progressive = false;
    return RT_OK;
}

int rt_synchronize(RtContext* c) {
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_update_scene_metadata(RtContext* c, const RtSphere* spheres, uint32_t ns, const RtLight* lights, uint32_t nl,
                             const RtBvhNode* nodes, uint32_t nn, const uint32_t* tri_idx, uint32_t ni,
                             const RtVertex* verts, uint32_t nv, RtSceneMetadataOffsets* out_offsets) {
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if ((ns && !spheres) || (nl && !lights) || (nn && !nodes) || (ni && !tri_idx) || (nv && !verts))
        return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_update_scene_metadata: NULL array with non-zero count");
    CK(c, cudaSetDevice(c->device));
    // src/buffers.rs:167-184
    size_t ss = (size_t)ns * sizeof(RtSphere), ls = (size_t)nl * sizeof(RtLight), bs = (size_t)nn * sizeof(RtBvhNode),
           is = (size_t)ni * 4, vs = (size_t)nv * sizeof(RtVertex);
    size_t total = ss + ls + bs + is + vs;
    if (total / 4 > 0xFFFFFFFFull) return fail(c, RT_ERR_INVALID_ARGUMENT, "scene metadata exceeds u32 offsets");
    bool needs_resize = total > c->scene_metadata_capacity_bytes;
    bool any_changed = ns != c->last_spheres || nl != c->last_lights || nn != c->last_nodes || ni != c->last_indices ||
                       nv != c->last_vertices;
    const bool upload = c->meta_dirty || needs_resize || any_changed; // src/buffers.rs:212
    size_t new_capacity = c->scene_metadata_capacity_bytes;
    if (needs_resize) { // src/buffers.rs:186-204: double, with the default floors
        size_t sc = (size_t)ns * 2 > 64 ? (size_t)ns * 2 : 64, lc = (size_t)nl * 2 > 32 ? (size_t)nl * 2 : 32,
               bc = (size_t)nn * 2 > 256 ? (size_t)nn * 2 : 256, ic = (size_t)ni * 2 > 1024 ? (size_t)ni * 2 : 1024,
               vc = (size_t)nv * 2 > 1024 ? (size_t)nv * 2 : 1024;
        new_capacity = sc * sizeof(RtSphere) + lc * sizeof(RtLight) + bc * sizeof(RtBvhNode) + ic * 4 + vc * sizeof(RtVertex);
    }
    if (upload) {
        // The new scene is staged into the BACK buffer while the host validates it; the live buffer, its offsets
        // and its re-layout are only replaced (pointer swap) once nothing can refuse the upload any more, so a
        // refused upload leaves the previously uploaded scene fully in place.  The DMA from the caller's (pinned)
        // arrays runs concurrently with the validation walk below.
        if (c->meta_back.reserve(new_capacity / 4) < 0)
            return fail(c, RT_ERR_CUDA, "scene metadata cudaMalloc(%zu) failed", new_capacity);
        char* d = reinterpret_cast<char*>(c->meta_back.ptr);
        size_t o = 0;
        cudaError_t ce = cudaSuccess;
        auto put = [&](const void* src, size_t bytes) {
            if (bytes && ce == cudaSuccess) ce = cudaMemcpyAsync(d + o, src, bytes, cudaMemcpyHostToDevice, c->stream);
            o += bytes;
        };
        put(spheres, ss); put(lights, ls); put(nodes, bs); put(tri_idx, is); put(verts, vs);
        // validation + compact numbering of the internal nodes (host walk, reference DFS order)
        WideBvh wide;
        build_wide(nodes, nn, ni, wide);
        // a cycle among the child links is only possible when that walk did not certify a tree
        const bool cyclic = !wide.ordered_ok && rt_bvh_has_cycle(nodes, nn);
        if (!cyclic && ce == cudaSuccess && wide.ordered_ok) {
            if ((wide.n_internal * 4 > c->wnodes.capacity || wide.leaves.size() > c->wleaves.capacity ||
                 wide.leaf_nodes.size() > c->wleaf_nodes.capacity) && cudaStreamSynchronize(c->stream) != cudaSuccess)
                ce = cudaErrorUnknown; // kernels may still read the auxiliary buffers a growth replaces
            if (ce == cudaSuccess && (c->wide_index.reserve(nn) < 0 || c->wnodes.reserve((size_t)wide.n_internal * 4) < 0 ||
                                      c->wleaves.reserve(wide.leaves.size()) < 0 || c->wleaf_nodes.reserve(wide.leaf_nodes.size()) < 0))
                ce = cudaErrorMemoryAllocation;
            if (ce == cudaSuccess && nn) ce = cudaMemcpyAsync(c->wide_index.ptr, wide.index.data(), (size_t)nn * 4, cudaMemcpyHostToDevice, c->stream);
            if (ce == cudaSuccess && !wide.leaves.empty()) {
                ce = cudaMemcpyAsync(c->wleaves.ptr, wide.leaves.data(), wide.leaves.size() * sizeof(uint2), cudaMemcpyHostToDevice, c->stream);
                if (ce == cudaSuccess)
                    ce = cudaMemcpyAsync(c->wleaf_nodes.ptr, wide.leaf_nodes.data(), wide.leaf_nodes.size() * 4, cudaMemcpyHostToDevice, c->stream);
            }
        }
        // Borrowed host memory (src/buffers.rs:236 `queue.write_buffer` copy semantics): the DMA reads the
        // caller's arrays (and `wide`) until the stream reaches this point -> wait before returning.
        cudaError_t se = cudaStreamSynchronize(c->stream);
        if (cyclic)
            return fail(c, RT_ERR_INVALID_ARGUMENT, "BVH child links form a cycle: the reference traversal would not terminate");
        if (ce != cudaSuccess || se != cudaSuccess) {
            // the live metadata buffer is untouched, but the auxiliary re-layout buffers may be half-written:
            // the previous scene falls back to the literal policy until its next upload
            c->wide.ordered_ok = false; c->wide.why_not = "a later upload failed half-way";
            c->relayout_dirty = true;
            return fail(c, RT_ERR_CUDA, "scene metadata upload failed: %s", cudaGetErrorString(ce != cudaSuccess ? ce : se));
        }
        std::swap(c->meta.ptr, c->meta_back.ptr);
        std::swap(c->meta.capacity, c->meta_back.capacity);
        c->scene_metadata_capacity_bytes = new_capacity;
        c->wide = std::move(wide);
        c->relayout_dirty = true; c->relayout_ok = false; c->rebuilt_ok = false;
        c->meta_len = total / 4;
        c->meta_dirty = false;
        c->last_spheres = ns; c->last_lights = nl; c->last_nodes = nn; c->last_indices = ni; c->last_vertices = nv;
    }
    // src/buffers.rs:257-268
    RtSceneMetadataOffsets off;
    off.spheres_offset = 0; off.spheres_count = ns;
    off.lights_offset = (uint32_t)(ss / 4); off.lights_count = nl;
    off.bvh_nodes_offset = (uint32_t)((ss + ls) / 4); off.bvh_nodes_count = nn;
    off.triangle_indices_offset = (uint32_t)((ss + ls + bs) / 4); off.triangle_indices_count = ni;
    off.vertices_offset = (uint32_t)((ss + ls + bs + is) / 4); off.vertices_count = nv;
    c->offsets = off;
    c->have_meta = true;
    if (out_offsets) *out_offsets = off;
    return needs_resize ? 1 : 0;
}

int rt_update_triangles(RtContext* c, const RtTriangle* tris, uint64_t n) {
    if (!c || (n && !tris)) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_update_triangles: NULL") : RT_ERR_INVALID_ARGUMENT;
    CK(c, cudaSetDevice(c->device));
    // src/buffers.rs:274-303: one more bound buffer per triangles_per_buffer triangles (never released)
    size_t needed = (size_t)((n + c->tpb - 1) / c->tpb);
    if (needed < 1) needed = 1;
    const bool needs_resize = c->tri_buffers < needed;
    bool changed = n != c->last_tris;
    if (c->tris_dirty || needs_resize || changed) {
        if ((size_t)n > c->tris.capacity) CK(c, cudaStreamSynchronize(c->stream)); // kernels may still read the old array
        if (c->tris.reserve((size_t)n) < 0) return fail(c, RT_ERR_CUDA, "triangle cudaMalloc(%llu) failed", (unsigned long long)n);
        if (needs_resize) c->tri_buffers = needed;
        if (n) CK(c, cudaMemcpyAsync(c->tris.ptr, tris, (size_t)n * sizeof(RtTriangle), cudaMemcpyHostToDevice, c->stream));
        CK(c, cudaStreamSynchronize(c->stream));
        c->n_tris = (size_t)n;
        c->tris_dirty = false;
        c->last_tris = (size_t)n;
        c->relayout_dirty = true;
    }
    return needs_resize ? 1 : 0;
}

int rt_update_materials(RtContext* c, const RtMaterial* mats, uint64_t n) {
    if (!c || (n && !mats)) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_update_materials: NULL") : RT_ERR_INVALID_ARGUMENT;
    CK(c, cudaSetDevice(c->device));
    bool needs_resize = n > c->mats_capacity; // src/buffers.rs:339-352
    if (needs_resize) {
        const size_t cap = (size_t)n * 2 > 64 ? (size_t)n * 2 : 64;
        CK(c, cudaStreamSynchronize(c->stream)); // kernels may still read the old buffer
        if (c->mats.reserve(cap) < 0) return fail(c, RT_ERR_CUDA, "materials cudaMalloc failed");
        c->mats_capacity = cap;
        // a new wgpu buffer is zero-initialised; entries [n, capacity) stay zero materials the shader can index
        CK(c, cudaMemsetAsync(c->mats.ptr, 0, cap * sizeof(RtMaterial), c->stream));
    }
    if (c->mats_dirty || needs_resize || n != c->last_mats) {
        if (n) CK(c, cudaMemcpyAsync(c->mats.ptr, mats, (size_t)n * sizeof(RtMaterial), cudaMemcpyHostToDevice, c->stream));
        CK(c, cudaStreamSynchronize(c->stream));
        c->n_mats = (size_t)n; c->mats_dirty = false; c->last_mats = (size_t)n;
    }
    return needs_resize ? 1 : 0;
}

int rt_update_textures(RtContext* c, const RtTextureInfo* tex, uint64_t n) {
    if (!c || (n && !tex)) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_update_textures: NULL") : RT_ERR_INVALID_ARGUMENT;
    CK(c, cudaSetDevice(c->device));
    bool needs_resize = n > c->textures_capacity; // src/buffers.rs:381-394
    if (needs_resize) {
        c->textures_capacity = (size_t)n * 2 > 64 ? (size_t)n * 2 : 64;
        if (c->textures.reserve(c->textures_capacity) < 0) return fail(c, RT_ERR_CUDA, "textures cudaMalloc failed");
    }
    if (c->tex_dirty || needs_resize || n != c->last_textures) {
        if (n) CK(c, cudaMemcpyAsync(c->textures.ptr, tex, (size_t)n * sizeof(RtTextureInfo), cudaMemcpyHostToDevice, c->stream));
        CK(c, cudaStreamSynchronize(c->stream));
        c->n_textures = (size_t)n; c->tex_dirty = false; c->last_textures = (size_t)n;
    }
    return needs_resize ? 1 : 0;
}

int rt_update_texture_data(RtContext* c, const uint8_t* bytes, uint64_t nb) {
    if (!c || (nb && !bytes)) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_update_texture_data: NULL") : RT_ERR_INVALID_ARGUMENT;
    CK(c, cudaSetDevice(c->device));
    // src/buffers.rs:422-432: 4 bytes little-endian per u32, last word zero-padded.
    size_t words = (size_t)((nb + 3) / 4);
    bool needs_resize = words > c->texdata_capacity;
    if (needs_resize) {
        c->texdata_capacity = words * 2 > 256 * 1024 ? words * 2 : 256 * 1024;
        if (c->texdata.reserve(c->texdata_capacity) < 0) return fail(c, RT_ERR_CUDA, "texture data cudaMalloc failed");
    }
    if (c->texdata_dirty || needs_resize || nb != c->last_texdata) {
        if (words) {
            CK(c, cudaMemsetAsync(c->texdata.ptr + (words - 1), 0, 4, c->stream));
            CK(c, cudaMemcpyAsync(c->texdata.ptr, bytes, (size_t)nb, cudaMemcpyHostToDevice, c->stream)); // x86 is LE
        }
        CK(c, cudaStreamSynchronize(c->stream));
        c->texdata_len = words; c->texdata_dirty = false; c->last_texdata = (size_t)nb;
    }
    return needs_resize ? 1 : 0;
}

void rt_mark_scene_metadata_dirty(RtContext* c) { if (c) c->meta_dirty = true; }
void rt_mark_triangles_dirty(RtContext* c) { if (c) c->tris_dirty = true; }
void rt_mark_materials_dirty(RtContext* c) { if (c) c->mats_dirty = true; }
void rt_mark_textures_dirty(RtContext* c) { if (c) c->tex_dirty = true; }
void rt_mark_texture_data_dirty(RtContext* c) { if (c) c->texdata_dirty = true; }
void rt_mark_all_dirty(RtContext* c) {
    if (c) c->meta_dirty = c->tris_dirty = c->mats_dirty = c->tex_dirty = c->texdata_dirty = true;
}
int rt_needs_update(const RtContext* c) {
    return c && (c->meta_dirty || c->tris_dirty || c->mats_dirty || c->tex_dirty || c->texdata_dirty) ? 1 : 0;
}
uint32_t rt_triangles_per_buffer(const RtContext* c) { return c ? c->tpb : RT_MAX_TRIANGLES_PER_BUFFER; }
int rt_set_triangles_per_buffer(RtContext* c, uint32_t n) {
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (n == 0) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_set_triangles_per_buffer: 0");
    // the reference fixes the value when the BufferManager is constructed (src/buffers.rs:73): only before the first upload
    if (c->last_tris != 0 || c->n_tris != 0 || c->tri_buffers != 1)
        return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_set_triangles_per_buffer: triangles were already uploaded");
    c->tpb = n;
    c->tris_dirty = true; c->relayout_dirty = true;
    return RT_OK;
}
int rt_get_metadata_offsets(const RtContext* c, RtSceneMetadataOffsets* out) {
    if (!c || !out || !c->have_meta) return RT_ERR_NOT_READY;
    *out = c->offsets;
    return RT_OK;
}

} // extern "C"

namespace {

// Builds the internal small-leaf LBVH over the n triangle positions (rt_lbvh.cuh).
int build_rebuilt(RtContext* c, uint32_t n) {
    c->rebuilt_ok = false;
    if (n == 0 || n >= RT_LBVH_MAX_PRIMS) return RT_OK;
    size_t temp_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                    (uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 63, c->stream);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, c->stream);
    if (scan_bytes > temp_bytes) temp_bytes = scan_bytes;
    if (c->lb_keys.reserve(n) < 0 || c->lb_keys_sorted.reserve(n) < 0 || c->lb_vals.reserve(n) < 0 ||
        c->lb_vals_sorted.reserve(n) < 0 || c->lb_temp.reserve(temp_bytes) < 0 || c->lb_children.reserve(n) < 0 ||
        c->lb_range.reserve(n) < 0 || c->lb_parent_i.reserve(n) < 0 || c->lb_parent_l.reserve(n) < 0 ||
        c->lb_visits.reserve(n) < 0 || c->lb_bounds.reserve(8) < 0 || c->lb_nbox.reserve((size_t)n * 2) < 0 ||
        c->bnodes.reserve((size_t)n * 4) < 0 || c->btris.reserve((size_t)n * 3) < 0 || c->bqnodes.reserve((size_t)n * 2) < 0 ||
        c->lb_grid.reserve(8) < 0)
        return fail(c, RT_ERR_CUDA, "LBVH cudaMalloc failed for %u triangles", n);
    const uint32_t g = (n + 255) / 256;
    // bounds[0..2] = 0xFFFFFFFF (ordered min), [3..7] = 0 (ordered max; [6] = tree depth)
    CK(c, cudaMemsetAsync(c->lb_bounds.ptr, 0xFF, 12, c->stream));
    CK(c, cudaMemsetAsync(c->lb_bounds.ptr + 3, 0, 20, c->stream));
    lbvh::centroid_bounds<<<g < 1184u ? g : 1184u, 256, 0, c->stream>>>(c->pbox.ptr, n, c->lb_bounds.ptr);
    lbvh::morton_keys<<<g, 256, 0, c->stream>>>(c->pbox.ptr, n, c->lb_bounds.ptr, c->lb_keys.ptr, c->lb_vals.ptr);
    c->launches += 2;
    if (cub::DeviceRadixSort::SortPairs(c->lb_temp.ptr, temp_bytes, c->lb_keys.ptr, c->lb_keys_sorted.ptr, c->lb_vals.ptr,
                                        c->lb_vals_sorted.ptr, (int)n, 0, 63, c->stream) != cudaSuccess)
        return fail(c, RT_ERR_CUDA, "LBVH radix sort failed");
    if (n >= 2) {
        CK(c, cudaMemsetAsync(c->lb_visits.ptr, 0, (size_t)n * 4, c->stream));
        lbvh::build_radix_tree<<<g, 256, 0, c->stream>>>(c->lb_keys_sorted.ptr, n, c->lb_children.ptr, c->lb_range.ptr,
                                                         c->lb_parent_i.ptr, c->lb_parent_l.ptr);
        // tree rotations while fitting (rt_lbvh.cuh); RTB200_LBVH_ROTATE=0 switches them off (tuning / A-B measurements)
        const char* rot_env = getenv("RTB200_LBVH_ROTATE");
        lbvh::fit_boxes<<<g, 256, 0, c->stream>>>(c->pbox.ptr, c->lb_vals_sorted.ptr, n, c->lb_children.ptr, c->lb_range.ptr,
                                                  c->lb_parent_i.ptr, c->lb_parent_l.ptr, c->lb_nbox.ptr, c->lb_visits.ptr,
                                                  (rot_env && atoi(rot_env) == 0) ? 0 : 1);
        // compact index of every emitted node: flags in lb_visits (free after the fit), scan into lb_vals (free after the sort)
        lbvh::mark_emitted<<<g, 256, 0, c->stream>>>(c->lb_range.ptr, n, c->lb_visits.ptr);
        if (cub::DeviceScan::ExclusiveSum(c->lb_temp.ptr, temp_bytes, c->lb_visits.ptr, c->lb_vals.ptr, (int)n, c->stream) != cudaSuccess)
            return fail(c, RT_ERR_CUDA, "LBVH node compaction scan failed");
        lbvh::quant_grid<<<1, 1, 0, c->stream>>>(c->lb_nbox.ptr, c->lb_grid.ptr);
        lbvh::emit_wide<<<g, 256, 0, c->stream>>>(c->pbox.ptr, c->lb_vals_sorted.ptr, n, c->lb_children.ptr, c->lb_range.ptr,
                                                  c->lb_nbox.ptr, c->lb_vals.ptr, c->lb_grid.ptr, c->bnodes.ptr);
        lbvh::emit_quant<<<g, 256, 0, c->stream>>>(c->pbox.ptr, c->lb_vals_sorted.ptr, n, c->lb_children.ptr, c->lb_range.ptr,
                                                   c->lb_nbox.ptr, c->lb_vals.ptr, c->lb_grid.ptr, c->bqnodes.ptr);
        lbvh::tree_depth<<<g, 256, 0, c->stream>>>(c->lb_parent_i.ptr, c->lb_parent_l.ptr, n, c->lb_bounds.ptr + 6);
        c->launches += 8;
    }
    lbvh::gather_sorted<<<g, 256, 0, c->stream>>>(c->wtris.ptr, c->lb_vals_sorted.ptr, n, c->leaf_of_pos.ptr, c->btris.ptr);
    c->launches++;
    // root box, grid and depth are read back by finish_rebuilt once the caller has waited for the stream
    CK(c, cudaMemcpyAsync(&c->rb_depth, c->lb_bounds.ptr + 6, 4, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaMemcpyAsync(c->rb_box, n >= 2 ? c->lb_nbox.ptr : c->pbox.ptr, sizeof c->rb_box, cudaMemcpyDeviceToHost, c->stream));
    if (n >= 2) CK(c, cudaMemcpyAsync(c->bq_grid, c->lb_grid.ptr, sizeof c->bq_grid, cudaMemcpyDeviceToHost, c->stream));
    return RT_OK;
}

void finish_rebuilt(RtContext* c, uint32_t n) {
    if (n == 0 || n >= RT_LBVH_MAX_PRIMS) return;
    const float4* rb = c->rb_box;
    const unsigned int tree_depth = c->rb_depth;
    const float v[6] = {rb[0].x, rb[0].y, rb[0].z, rb[1].x, rb[1].y, rb[1].z};
    const float pad = lbvh::scene_pad(rb[0], rb[1]); // the same scene-wide inflation every box of the tree got
    for (int k = 0; k < 3; k++) { c->broot_min[k] = v[k] - pad; c->broot_max[k] = v[3 + k] + pad; }
    bool finite = true;
    for (float x : v) finite = finite && isfinite(x);
    c->rebuilt_far = RT_REBUILT_FAR * (pad * 65536.0f); // RT_REBUILT_FAR scene magnitudes
    // root: the whole tree collapses into one leaf when n <= RT_LBVH_LEAF
    c->broot_ref = n <= RT_LBVH_LEAF ? (RT_LEAF_FLAG | ((n - 1u) << 28)) : 0u;
    if (getenv("RTB200_DEBUG")) fprintf(stderr, "[rtb200] internal LBVH: %u triangles, depth %u\n", n, tree_depth);
    c->lbvh_depth = tree_depth;
    // the traversal stacks hold 64 entries (many coincident centroids can deepen the tree); a scene with
    // non-finite coordinates has no usable grid or pad: such uploads keep the ORDERED / LITERAL policies
    c->rebuilt_ok = tree_depth < 60 && finite;
}

// Bring the ORDERED / REBUILT policy buffers up to date (all on the GPU).
int prepare_relayout(RtContext* c) {
    if (!c->relayout_dirty) return RT_OK;
    c->relayout_ok = false;
    c->rebuilt_ok = false;
    if (c->have_meta && c->wide.ordered_ok) {
        uint32_t ni = c->offsets.triangle_indices_count;
        // REBUILT is worth it (and certifiable) only for big-leaf, properly nested trees
        bool want_rebuilt = c->wide.nested_ok && c->wide.max_leaf > 2 * RT_LBVH_LEAF && ni < RT_LBVH_MAX_PRIMS;
        // experiment switch (off unless set): rebuild small-leaf uploads of at least this many triangles too
        // (one-sector nodes + 2-triangle leaves against the uploaded tree's fp32 nodes)
        if (const char* e = getenv("RTB200_REBUILD_MIN_TRIANGLES")) {
            const long long m = atoll(e);
            if (m > 0 && (long long)ni >= m && c->wide.nested_ok && ni < RT_LBVH_MAX_PRIMS) want_rebuilt = true;
        }
        cudaEvent_t e0 = c->ev0, e1 = c->ev1;
        CK(c, cudaEventRecord(e0, c->stream));
        if (c->wtris.reserve((size_t)ni * 3 + 3) < 0 || c->leaf_of_pos.reserve((size_t)ni + 1) < 0 ||
            (want_rebuilt && (c->pbox.reserve((size_t)ni * 2 + 2) < 0 ||
                              c->cert_box.reserve((size_t)c->offsets.bvh_nodes_count * 2 + 2) < 0)))
            return fail(c, RT_ERR_CUDA, "relayout cudaMalloc failed");
        CK(c, cudaMemsetAsync(c->dflag.ptr, 0, 4, c->stream));
        CK(c, cudaMemsetAsync(c->leaf_of_pos.ptr, 0xFF, ((size_t)ni + 1) * 4, c->stream));
        if (c->wide.n_internal) {
            relayout_wide_nodes<<<(c->offsets.bvh_nodes_count + 255) / 256, 256, 0, c->stream>>>(c->meta.ptr, c->offsets, c->wide_index.ptr,
                                                                                                 c->wnodes.ptr);
            c->launches++;
        }
        if (ni) {
            relayout_triangles<<<(ni + 255) / 256, 256, 0, c->stream>>>(c->meta.ptr, c->offsets, c->tris.ptr, c->n_tris,
                                                                        c->tpb, (uint32_t)c->tri_buffers, c->wtris.ptr,
                                                                        want_rebuilt ? c->pbox.ptr : nullptr, c->dflag.ptr);
            c->launches++;
            uint32_t nl = (uint32_t)c->wide.leaves.size();
            if (nl) {
                relayout_leaf_counts<<<(nl * 32 + 255) / 256, 256, 0, c->stream>>>(c->wleaves.ptr, c->wleaf_nodes.ptr, nl,
                                                                                   c->wtris.ptr, c->leaf_of_pos.ptr, c->meta.ptr,
                                                                                   c->meta_len, c->offsets.bvh_nodes_offset,
                                                                                   want_rebuilt ? c->cert_box.ptr : nullptr);
                c->launches++;
            }
        }
        // the internal tree is enqueued before the validation flag is read back: one host wait for both
        if (want_rebuilt) {
            int r = build_rebuilt(c, ni);
            if (r != RT_OK) return r;
        }
        uint32_t flag = 0;
        CK(c, cudaMemcpyAsync(&flag, c->dflag.ptr, 4, cudaMemcpyDeviceToHost, c->stream));
        CK(c, cudaEventRecord(e1, c->stream));
        CK(c, cudaEventSynchronize(e1));
        CK(c, cudaGetLastError());
        cudaEventElapsedTime(&c->last_build_ms, e0, e1);
        if (flag & 1u) { c->rebuilt_ok = false; return fail(c, RT_ERR_INVALID_SCENE, "a triangle references a vertex index >= vertices_count"); }
        c->relayout_ok = true;
        if (want_rebuilt) finish_rebuilt(c, ni);
    }
    c->relayout_dirty = false;
    return RT_OK;
}

void fill_scene(RtContext* c, KParams& k) {
    k.meta = c->meta.ptr; k.meta_len = c->meta_len; k.off = c->offsets;
    k.tris = c->tris.ptr; k.n_tris = c->n_tris;
    k.ctx_triangles_per_buffer = c->tpb; k.n_tri_buffers = (uint32_t)c->tri_buffers;
    // materials.len() in the shader is the bound buffer's length = capacity (lib.rs:307, src/renderer.rs:698)
    k.mats = c->mats.ptr; k.n_mats = (uint32_t)c->mats_capacity;
    k.wnodes = c->wnodes.ptr; k.wtris = c->wtris.ptr;
    k.root_ref = c->wide.root_ref;
    memcpy(k.root_min, c->wide.root_min, 12); memcpy(k.root_max, c->wide.root_max, 12);
    k.bnodes = c->bnodes.ptr; k.btris = c->btris.ptr; k.leaf_of_pos = c->leaf_of_pos.ptr;
    k.broot_ref = c->broot_ref;
    memcpy(k.broot_min, c->broot_min, 12); memcpy(k.broot_max, c->broot_max, 12);
    k.cert_box = c->cert_box.ptr; k.rebuilt_far = c->rebuilt_far;
    k.fallback_ordered = (c->relayout_ok && !getenv("RTB200_FALLBACK_LITERAL")) ? 1u : 0u;
    k.bqnodes = c->bqnodes.ptr; memcpy(k.bq_origin, c->bq_grid, 12); memcpy(k.bq_scale, c->bq_grid + 3, 12);
    k.accum = c->accum.ptr; k.hits = c->hits.ptr;
    for (int i = 0; i < 3; i++) k.chan[i] = c->chan[i].ptr;
    k.composite = c->composite.ptr;
    k.counters = c->count_next ? c->dcounters.ptr : nullptr;
}

void fill_camera(KParams& k, const RtCamera& cam) {
    memcpy(k.cam_pos, cam.position, 12); memcpy(k.cam_dir, cam.direction, 12); memcpy(k.cam_up, cam.up, 12);
    // ray.rs:34 / wavefront.rs:89 — evaluated once on the host with glibc tanf (the
    // same function the oracle calls); CUDA's tanf is not bit-identical to it.
    k.fov_scale = tanf(cam.fov * 0.5f * 3.14159265358979323846f / 180.0f);
}

template <int POLICY>
void launch_policy(RtContext* c, const KParams& k, uint32_t blocks) {
    if (k.counters) rt_trace_shade<POLICY, true><<<blocks, 256, 0, c->stream>>>(k);
    else rt_trace_shade<POLICY, false><<<blocks, 256, 0, c->stream>>>(k);
}

int launch(RtContext* c, const KParams& k, uint32_t blocks, uint32_t traversal) {
    if (blocks == 0) return RT_OK;
    // the re-laid-out buffers bake in the uploaded offsets and the context's buffer
    // split; a dispatch whose push constants differ from them takes the literal path
    bool same_scene = k.off.bvh_nodes_count > 0 && k.triangles_per_buffer == c->tpb &&
                      memcmp(&k.off, &c->offsets, sizeof(RtSceneMetadataOffsets)) == 0;
    bool can_ordered = c->relayout_ok && same_scene, can_rebuilt = c->rebuilt_ok && same_scene;
    int policy;
    if (traversal == RT_TRAVERSAL_LITERAL) policy = 0;
    else if (traversal == RT_TRAVERSAL_ORDERED) policy = can_ordered ? 1 : -1;
    else if (traversal == RT_TRAVERSAL_REBUILT) policy = can_rebuilt ? 2 : -1;
    else policy = can_rebuilt ? 2 : (can_ordered ? 1 : 0);
    if (policy < 0)
        return fail(c, RT_ERR_INVALID_SCENE, "requested traversal unavailable: %s",
                    !same_scene ? "push constants differ from the uploaded scene"
                                : (c->wide.why_not.empty() ? "BVH not eligible (small leaves, not nested, or too large)" : c->wide.why_not.c_str()));
    CK(c, cudaEventRecord(c->ev0, c->stream));
    if (policy == 2) launch_policy<2>(c, k, blocks);
    else if (policy == 1) launch_policy<1>(c, k, blocks);
    else launch_policy<0>(c, k, blocks);
    CK(c, cudaGetLastError());
    CK(c, cudaEventRecord(c->ev1, c->stream));
    c->timed = true;
    c->launches++;
    c->last_traversal = policy == 2 ? RT_TRAVERSAL_REBUILT : (policy == 1 ? RT_TRAVERSAL_ORDERED : RT_TRAVERSAL_LITERAL);
    c->count_next = false;
    return RT_OK;
}

int pick_policy(RtContext* c, const KParams& k, uint32_t traversal) {
    bool same_scene = k.off.bvh_nodes_count > 0 && k.triangles_per_buffer == c->tpb &&
                      memcmp(&k.off, &c->offsets, sizeof(RtSceneMetadataOffsets)) == 0;
    bool can_ordered = c->relayout_ok && same_scene, can_rebuilt = c->rebuilt_ok && same_scene;
    if (traversal == RT_TRAVERSAL_LITERAL) return 0;
    if (traversal == RT_TRAVERSAL_ORDERED) return can_ordered ? 1 : -1;
    if (traversal == RT_TRAVERSAL_REBUILT) return can_rebuilt ? 2 : -1;
    return can_rebuilt ? 2 : (can_ordered ? 1 : 0);
}

// EXTENSION path tracing: persistent grid, one resident wave of 256-thread blocks
int launch_path(RtContext* c, const KParams& k, uint32_t traversal) {
    if (k.n_work == 0) return RT_OK;
    int policy = pick_policy(c, k, traversal);
    if (policy < 0) return fail(c, RT_ERR_INVALID_SCENE, "requested traversal unavailable for this scene");
    const void* fn = policy == 2 ? (const void*)rt_path_trace<2, false> : policy == 1 ? (const void*)rt_path_trace<1, false> : (const void*)rt_path_trace<0, false>;
    int per_sm = 0, sms = 0;
    CK(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 256, 0));
    CK(c, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    uint32_t blocks = (uint32_t)(per_sm > 0 ? per_sm : 1) * (uint32_t)sms;
    uint32_t need = (k.n_work + 255u) / 256u;
    if (blocks > need) blocks = need;
    CK(c, cudaEventRecord(c->ev0, c->stream));
    if (k.counters) {
        if (policy == 2) rt_path_trace<2, true><<<blocks, 256, 0, c->stream>>>(k);
        else if (policy == 1) rt_path_trace<1, true><<<blocks, 256, 0, c->stream>>>(k);
        else rt_path_trace<0, true><<<blocks, 256, 0, c->stream>>>(k);
    } else {
        if (policy == 2) rt_path_trace<2, false><<<blocks, 256, 0, c->stream>>>(k);
        else if (policy == 1) rt_path_trace<1, false><<<blocks, 256, 0, c->stream>>>(k);
        else rt_path_trace<0, false><<<blocks, 256, 0, c->stream>>>(k);
    }
    CK(c, cudaGetLastError());
    CK(c, cudaEventRecord(c->ev1, c->stream));
    c->timed = true;
    c->launches++;
    c->last_traversal = policy == 2 ? RT_TRAVERSAL_REBUILT : (policy == 1 ? RT_TRAVERSAL_ORDERED : RT_TRAVERSAL_LITERAL);
    c->count_next = false;
    return RT_OK;
}

// EXTENSION path tracing, streaming wavefront engine (rt_wave.cuh): all samples of a wave in
// flight at once; per wave one generate, per depth one extend + one shade, one resolve.
// Queue lengths stay on the device (no host sync inside the frame).
__global__ void wv_publish(const WvCtrl* ctrl, int n_sub, unsigned long long* dcounters, int count) {
    unsigned long long seg = 0, cc[4] = {0, 0, 0, 0};
    for (int i = 0; i < n_sub; i++) { seg += ctrl[i].segments; for (int j = 0; j < 4; j++) cc[j] += ctrl[i].c[j]; }
    dcounters[6] = seg; // rt_last_segments
    for (int d = 0; d < 8; d++) { unsigned long long r = 0; for (int i = 0; i < n_sub; i++) r += ctrl[i].rays_at_depth[d]; dcounters[16 + d] = r; }
    if (count) { dcounters[0] = cc[0]; dcounters[1] = cc[1]; dcounters[2] = seg; dcounters[3] = cc[2]; dcounters[4] = cc[3]; }
}

typedef void (*WvKernel)(const KParams, WvBuffers, int);
template <int P, bool D>
static WvKernel wv_extend_pick(bool count, bool gen) {
    if (count) return gen ? wv_extend<P, true, true, D> : wv_extend<P, true, false, D>;
    return gen ? wv_extend<P, false, true, D> : wv_extend<P, false, false, D>;
}
static WvKernel wv_extend_kernel(int policy, bool count, bool gen, bool deep) {
    if (policy == 2) return deep ? wv_extend_pick<2, true>(count, gen) : wv_extend_pick<2, false>(count, gen);
    return deep ? wv_extend_pick<1, true>(count, gen) : wv_extend_pick<1, false>(count, gen);
}
template <int P, int S>
static WvKernel wv_shade_pick(bool count, bool gen) {
    if (count) return gen ? wv_shade<P, true, true, S> : wv_shade<P, true, false, S>;
    return gen ? wv_shade<P, false, true, S> : wv_shade<P, false, false, S>;
}
static WvKernel wv_shade_kernel(int policy, bool count, bool gen, int split) { // split: 0 / 1 one form, 2 both (rt_wave.cuh)
    if (policy == 2) return split == 2 ? wv_shade_pick<2, 2>(count, gen) : split ? wv_shade_pick<2, 1>(count, gen) : wv_shade_pick<2, 0>(count, gen);
    return split == 2 ? wv_shade_pick<1, 2>(count, gen) : split ? wv_shade_pick<1, 1>(count, gen) : wv_shade_pick<1, 0>(count, gen);
}

int launch_wave(RtContext* c, KParams k, uint32_t traversal) {
    if (k.n_work == 0) return RT_OK;
    int policy = pick_policy(c, k, traversal);
    if (policy < 1) return -100; // literal policy: the caller falls back to the megakernel
    size_t cap = 40u << 20; // paths in flight per wave (64 B state x2 + 8 B hit x2 + 16 B radiance each)
    if (const char* e = getenv("RTB200_WAVE_PATHS")) { long long v = atoll(e); if (v > 0) cap = (size_t)v; }
    uint32_t spp_wave = (uint32_t)(cap / k.n_work);
    if (spp_wave < 1u) spp_wave = 1u;
    if (spp_wave > k.spp) spp_wave = k.spp;
    // Sub-frames: the context's tiles are dealt round-robin to n_sub independent wavefronts on their own
    // streams.  Every per-depth extend launch ends in a tail (a few long rays keep a handful of warps busy
    // for ~0.1-0.2 ms); with two sub-frames the other one's kernels fill the SMs meanwhile.  Worth it when
    // Measured on C3: +3 % at 35 M paths on one GPU, more when the context's share of the frame is small
    // (multi-GPU tile split); 3 or 4 sub-frames are no better than 2.
    const uint32_t my_tiles = k.n_work >> 12;
    uint32_t n_sub = my_tiles >= 2u ? 2u : 1u;
    if (const char* e = getenv("RTB200_WAVE_SPLIT")) { int v = atoi(e); if (v >= 1 && v <= 4) n_sub = (uint32_t)v; }
    if (n_sub > my_tiles) n_sub = my_tiles;
    const size_t P = (size_t)k.n_work * spp_wave;
    if (c->wv_state[0].reserve(P * 4) < 0 || c->wv_state[1].reserve(P * 4) < 0 || c->wv_hit[0].reserve(P) < 0 ||
        c->wv_hit[1].reserve(P) < 0 || c->wv_lbuf.reserve(P) < 0 || c->wv_fb.reserve(P) < 0 || c->wv_ctrl.reserve(4) < 0)
        return fail(c, RT_ERR_CUDA, "wave engine cudaMalloc failed for %zu paths", P);
    while (c->wv_streams.size() + 1 < n_sub) {
        cudaStream_t st; cudaEvent_t ev;
        CK(c, cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        CK(c, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        c->wv_streams.push_back(st); c->wv_join.push_back(ev);
    }
    if (!c->wv_fork) CK(c, cudaEventCreateWithFlags(&c->wv_fork, cudaEventDisableTiming));
    const bool count = k.counters != nullptr;
    int per_sm = 0, sms = 0;
    // the shared-memory stack holds one deferred sibling per level plus the sentinel entry
    const uint32_t tree_levels = policy == 2 ? c->lbvh_depth : c->wide.max_depth;
    bool deep = tree_levels + 2u > WV_STK;
    if (const char* e = getenv("RTB200_WAVE_DEEP")) deep = deep || atoi(e) != 0; // tests: force the overflow-capable form
    const void* fn = (const void*)wv_extend_kernel(policy, false, false, deep);
    CK(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 256, 0));
    CK(c, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    CK(c, cudaEventRecord(c->ev0, c->stream));
    CK(c, cudaMemsetAsync(c->wv_ctrl.ptr, 0, 4 * sizeof(WvCtrl), c->stream));
    if (n_sub > 1) CK(c, cudaEventRecord(c->wv_fork, c->stream));
    const uint32_t s_end = k.first_sample + k.spp;
    // which wv_shade form a depth takes (rt_wave.cuh): split when misses * 4 > rays * quarters
    bool one_shade_launch = WV_SHADE_ONE_LAUNCH != 0;
    if (const char* e = getenv("RTB200_SHADE_LAUNCHES")) one_shade_launch = atoi(e) == 1;
    uint32_t split_gen = 1u, split_deep = 1u;
    if (const char* e = getenv("RTB200_SPLIT_GEN")) { int v = atoi(e); if (v >= 0 && v <= 5) split_gen = (uint32_t)v; }
    if (const char* e = getenv("RTB200_SPLIT_DEEP")) { int v = atoi(e); if (v >= 0 && v <= 5) split_deep = (uint32_t)v; }
    // per-stage timing (single sub-frame only): an event between consecutive launches (rt_last_stage_ms)
    size_t n_ev = 0;
    c->wv_event_stage.clear();
    const bool stage_timing = n_sub == 1;
    auto mark = [&](int stage) -> int {
        if (!stage_timing) return 0;
        if (n_ev >= c->wv_events.size()) {
            cudaEvent_t e;
            if (cudaEventCreate(&e) != cudaSuccess) return -1;
            c->wv_events.push_back(e);
        }
        if (cudaEventRecord(c->wv_events[n_ev++], c->stream) != cudaSuccess) return -1;
        if (stage >= 0) c->wv_event_stage.push_back(stage);
        return 0;
    };
    if (mark(-1)) return fail(c, RT_ERR_CUDA, "wave engine: event record failed");
    size_t path_base = 0;
    for (uint32_t sub = 0; sub < n_sub; sub++) {
        cudaStream_t st = sub == 0 ? c->stream : c->wv_streams[sub - 1];
        if (sub) CK(c, cudaStreamWaitEvent(st, c->wv_fork, 0));
        KParams ks = k;
        ks.tile_stride = k.tile_stride * n_sub;
        ks.tile_phase = k.tile_phase + sub * k.tile_stride;
        const uint32_t tiles_s = (my_tiles - sub + n_sub - 1) / n_sub; // local tiles sub, sub + n_sub, ...
        ks.n_work = tiles_s * 4096u;
        const size_t Ps = (size_t)ks.n_work * spp_wave;
        WvBuffers w;
        w.valid_pixels = 0; // pixels of this sub-frame's tiles that lie inside the image
        for (uint32_t t = 0; t < tiles_s; t++) {
            uint32_t tile = ks.tile_phase + t * ks.tile_stride;
            if (ks.tile_list) tile = c->tile_list_host[tile];
            const uint32_t tx = tile % ks.tiles_x, ty = tile / ks.tiles_x;
            const uint32_t x0 = tx * RT_SPLIT_TILE, y0 = ty * RT_SPLIT_TILE;
            const uint32_t tw = x0 >= ks.width ? 0u : (ks.width - x0 < RT_SPLIT_TILE ? ks.width - x0 : RT_SPLIT_TILE);
            const uint32_t th = y0 >= ks.height ? 0u : (ks.height - y0 < RT_SPLIT_TILE ? ks.height - y0 : RT_SPLIT_TILE);
            w.valid_pixels += tw * th;
        }
        w.state[0] = c->wv_state[0].ptr + path_base * 4; w.state[1] = c->wv_state[1].ptr + path_base * 4;
        w.hit[0] = c->wv_hit[0].ptr + path_base; w.hit[1] = c->wv_hit[1].ptr + path_base;
        w.lbuf = c->wv_lbuf.ptr + path_base; w.ctrl = c->wv_ctrl.ptr + sub; w.fb_list = c->wv_fb.ptr + path_base;
        w.capacity = (uint32_t)Ps;
        path_base += Ps;
        uint32_t ext_per_sm = (uint32_t)(per_sm > 0 ? per_sm : 1);
        if (const char* e = getenv("RTB200_EXT_BLOCKS_PER_SM")) { int v = atoi(e); if (v >= 1 && (uint32_t)v < ext_per_sm) ext_per_sm = (uint32_t)v; } // tuning
        uint32_t ext_blocks = ext_per_sm * (uint32_t)sms;
        uint32_t full_blocks = (uint32_t)((Ps + 255) / 256);
        if (ext_blocks > full_blocks) ext_blocks = full_blocks;
        uint32_t stream_blocks = (uint32_t)sms * 8u;
        if (const char* e = getenv("RTB200_SHADE_BLOCKS_PER_SM")) { int v = atoi(e); if (v >= 1 && v <= 16) stream_blocks = (uint32_t)sms * (uint32_t)v; } // tuning
        if (stream_blocks > full_blocks) stream_blocks = full_blocks;
        const uint32_t fb_blocks = stream_blocks < (uint32_t)sms * 2u ? stream_blocks : (uint32_t)sms * 2u;
        for (uint32_t s0 = k.first_sample; s0 < s_end; s0 += spp_wave) {
            w.s0 = s0; w.spp_wave = s_end - s0 < spp_wave ? s_end - s0 : spp_wave;
            w.first_wave = s0 == k.first_sample; w.last_wave = s0 + w.spp_wave == s_end;
            if (!w.first_wave) CK(c, cudaMemsetAsync(w.ctrl, 0, 16, st)); // n[0], n[1], head
            if (!w.first_wave) mark(-1);
#if !WV_FUSE_GEN
            wv_generate<<<stream_blocks, 256, 0, st>>>(ks, w);
            c->launches++;
#endif
            mark(0);
            for (uint32_t d = 0; d <= k.path_max_depth; d++) {
                const int qi = (int)(d & 1u);
                const bool gen = WV_FUSE_GEN && d == 0u; // depth 0 computes its primary rays from the queue position
                w.depth = d;
                w.split_quarters = gen ? split_gen : split_deep;
                wv_extend_kernel(policy, count, gen, deep)<<<ext_blocks, 256, 0, st>>>(ks, w, qi);
                mark(1);
                // both forms are launched; the one that does not match the depth's miss ratio returns at once
                if (one_shade_launch) {
                    wv_shade_kernel(policy, count, gen, 2)<<<stream_blocks, 256, 0, st>>>(ks, w, qi);
                    mark(2);
                } else {
                    wv_shade_kernel(policy, count, gen, 0)<<<stream_blocks, 256, 0, st>>>(ks, w, qi);
                    mark(2);
                    wv_shade_kernel(policy, count, gen, 1)<<<stream_blocks, 256, 0, st>>>(ks, w, qi);
                    mark(4);
                    c->launches++;
                }
                // the rays those left to the reference's own traversal (normally none: the kernel reads an empty list)
                if (count) { if (gen) wv_fallback<true, true><<<fb_blocks, 256, 0, st>>>(ks, w, qi); else wv_fallback<true, false><<<fb_blocks, 256, 0, st>>>(ks, w, qi); }
                else { if (gen) wv_fallback<false, true><<<fb_blocks, 256, 0, st>>>(ks, w, qi); else wv_fallback<false, false><<<fb_blocks, 256, 0, st>>>(ks, w, qi); }
                mark(5);
                c->launches += 3;
            }
            wv_resolve<<<stream_blocks, 256, 0, st>>>(ks, w);
            mark(3);
            c->launches++;
        }
        if (sub) { CK(c, cudaEventRecord(c->wv_join[sub - 1], st)); CK(c, cudaStreamWaitEvent(c->stream, c->wv_join[sub - 1], 0)); }
    }
    CK(c, cudaGetLastError());
    // publish the segment count (and instrumentation) where rt_last_segments / rt_read_counters look
    wv_publish<<<1, 1, 0, c->stream>>>(c->wv_ctrl.ptr, (int)n_sub, c->dcounters.ptr, count ? 1 : 0);
    c->launches++;
    CK(c, cudaEventRecord(c->ev1, c->stream));
    c->timed = true;
    c->last_traversal = policy == 2 ? RT_TRAVERSAL_REBUILT : RT_TRAVERSAL_ORDERED;
    c->count_next = false;
    c->last_was_wavefront = true;
    c->last_wave_max_depth = k.path_max_depth; c->last_wave_seed = k.base_seed;
    return RT_OK;
}

// Primary frames (RT_MODE_LEGACY / RT_MODE_WAVEFRONT0) of a fused rt_render_frame through the wave engine's
// traversal: one wv_extend<GEN> launch + one wv_shade_primary launch (rt_wave.cuh).  Returns -100 when no
// re-laid-out policy is available (the caller then launches rt_trace_shade with the literal policy).
int launch_primary_wave(RtContext* c, KParams k, uint32_t traversal, uint32_t my_tiles) {
    if (my_tiles == 0) return RT_OK;
    int policy = pick_policy(c, k, traversal);
    if (policy < 1) return -100;
    // Measured on B200 (profiles/r2b_probe_primary.log): for ONE sample per pixel the fused kernel rt_trace_shade
    // is faster than this two-launch form (C3 scene 0.388 vs 0.420 ms, C2 0.285 vs 0.383 ms: 2 M rays leave each of
    // the 151 552 persistent threads ~14 rays, and the hit records make a round trip through memory), so the
    // fused kernel stays the default and this path is opt-in (tests run both).
    if (!getenv("RTB200_PRIMARY_WAVE")) return -100;
    k.n_work = my_tiles * 4096u;
    k.spp = 1; k.first_sample = 0; k.base_seed = 0; k.path_max_depth = 0;
    k.legacy_camera = k.mode == RT_MODE_WAVEFRONT0 ? 0u : 1u;
    const size_t P = k.n_work;
    if (c->wv_hit[0].reserve(P) < 0 || c->wv_ctrl.reserve(4) < 0)
        return fail(c, RT_ERR_CUDA, "wave engine cudaMalloc failed for %zu primary rays", P);
    const bool count = k.counters != nullptr;
    const uint32_t tree_levels = policy == 2 ? c->lbvh_depth : c->wide.max_depth;
    bool deep = tree_levels + 2u > WV_STK;
    if (const char* e = getenv("RTB200_WAVE_DEEP")) deep = deep || atoi(e) != 0;
    int per_sm = 0, sms = 0;
    const void* fn = (const void*)wv_extend_kernel(policy, false, true, deep);
    CK(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 256, 0));
    CK(c, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    WvBuffers w;
    memset(&w, 0, sizeof w);
    w.hit[0] = c->wv_hit[0].ptr; w.hit[1] = c->wv_hit[0].ptr; // depth 0 only: nothing is appended
    w.ctrl = c->wv_ctrl.ptr;
    w.s0 = 0; w.spp_wave = 1; w.first_wave = 1; w.last_wave = 1; w.capacity = (uint32_t)P;
    for (uint32_t t = 0; t < my_tiles; t++) {
        uint32_t tile = k.tile_phase + t * k.tile_stride;
        if (k.tile_list) tile = c->tile_list_host[tile];
        const uint32_t tx = tile % k.tiles_x, ty = tile / k.tiles_x;
        const uint32_t x0 = tx * RT_SPLIT_TILE, y0 = ty * RT_SPLIT_TILE;
        const uint32_t tw = x0 >= k.width ? 0u : (k.width - x0 < RT_SPLIT_TILE ? k.width - x0 : RT_SPLIT_TILE);
        const uint32_t th = y0 >= k.height ? 0u : (k.height - y0 < RT_SPLIT_TILE ? k.height - y0 : RT_SPLIT_TILE);
        w.valid_pixels += tw * th;
    }
    uint32_t full_blocks = (uint32_t)((P + 255) / 256);
    uint32_t ext_blocks = (uint32_t)(per_sm > 0 ? per_sm : 1) * (uint32_t)sms;
    if (ext_blocks > full_blocks) ext_blocks = full_blocks;
    uint32_t shade_blocks = (uint32_t)sms * 8u;
    if (shade_blocks > full_blocks) shade_blocks = full_blocks;
    CK(c, cudaEventRecord(c->ev0, c->stream));
    CK(c, cudaMemsetAsync(c->wv_ctrl.ptr, 0, sizeof(WvCtrl), c->stream));
    wv_extend_kernel(policy, count, true, deep)<<<ext_blocks, 256, 0, c->stream>>>(k, w, 0);
    if (policy == 2) { if (count) wv_shade_primary<2, true><<<shade_blocks, 256, 0, c->stream>>>(k, w); else wv_shade_primary<2, false><<<shade_blocks, 256, 0, c->stream>>>(k, w); }
    else { if (count) wv_shade_primary<1, true><<<shade_blocks, 256, 0, c->stream>>>(k, w); else wv_shade_primary<1, false><<<shade_blocks, 256, 0, c->stream>>>(k, w); }
    CK(c, cudaGetLastError());
    c->launches += 2;
    if (count) { wv_publish<<<1, 1, 0, c->stream>>>(c->wv_ctrl.ptr, 1, c->dcounters.ptr, 1); c->launches++; }
    CK(c, cudaEventRecord(c->ev1, c->stream));
    c->timed = true;
    c->last_traversal = policy == 2 ? RT_TRAVERSAL_REBUILT : RT_TRAVERSAL_ORDERED;
    c->count_next = false;
    c->last_was_wavefront = false;
    return RT_OK;
}

// Microbenchmark behind the traversal roofline: every lane loads one random, independent
// 32-byte sector per iteration (ld.global.nc.v8.f32, the node-fetch instruction) from a
// buffer of `sectors` sectors.  A small footprint measures the SM's L1 rate for divergent
// sector gathers, a footprint between L1 and L2 size the L2 rate, a larger one HBM.
__global__ void __launch_bounds__(256) probe_gather(const float4* __restrict__ buf, uint32_t sector_mask, uint32_t iters,
                                                    float* __restrict__ out) {
    uint32_t x = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    float acc = 0.0f;
    for (uint32_t i = 0; i < iters; i += 4) {
        float4 a[4], b[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            x = x * 1664525u + 1013904223u;
            const float4* sp = buf + (size_t)((x >> 7) & sector_mask) * 2;
            asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=f"(a[u].x), "=f"(a[u].y), "=f"(a[u].z), "=f"(a[u].w), "=f"(b[u].x), "=f"(b[u].y), "=f"(b[u].z), "=f"(b[u].w)
                         : "l"(sp));
        }
#pragma unroll
        for (int u = 0; u < 4; u++) acc += a[u].x + b[u].w;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

int check_ready(RtContext* c) {
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (c->width == 0) return fail(c, RT_ERR_NOT_READY, "rt_resize has not been called");
    if (!c->have_meta) return fail(c, RT_ERR_NOT_READY, "rt_update_scene_metadata has not been called");
    return RT_OK;
}

} // namespace

extern "C" {

// Bring the device-side re-layout (ORDERED / REBUILT policy buffers, internal LBVH) of the uploaded scene up to date
// now instead of inside the next dispatch / frame: a host that double-buffers scenes over two contexts calls this from
// its upload thread so that the next frame's launch finds everything ready.
int rt_prepare_scene(RtContext* c) {
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (!c->have_meta) return fail(c, RT_ERR_NOT_READY, "rt_update_scene_metadata has not been called");
    CK(c, cudaSetDevice(c->device));
    return prepare_relayout(c);
}

int rt_dispatch(RtContext* c, const RtPushConstants* pc) {
    int r = check_ready(c);
    if (r) return r;
    if (!pc) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_dispatch: pc is NULL");
    uint32_t channel = rt_pc_color_channel(pc);
    // src/compute.rs:227-229: get_compute_bind_group(channel).expect("Invalid color channel")
    if (channel > 2u) return fail(c, RT_ERR_INVALID_CHANNEL, "Invalid color channel %u", channel);
    if ((uint32_t)pc->resolution[0] != c->width || (uint32_t)pc->resolution[1] != c->height)
        return fail(c, RT_ERR_INVALID_ARGUMENT, "push-constant resolution %gx%g != output size %ux%u", pc->resolution[0],
                    pc->resolution[1], c->width, c->height);
    CK(c, cudaSetDevice(c->device));
    if ((r = prepare_relayout(c)) != RT_OK) return r;
    KParams k;
    memset(&k, 0, sizeof k);
    fill_scene(c, k);
    k.off = pc->metadata_offsets; // the shader trusts the push constants (scene_access.rs)
    // offsets pointing outside the uploaded buffer read as 0 (meta_at bounds check)
    k.width = c->width; k.height = c->height; k.res_x = pc->resolution[0]; k.res_y = pc->resolution[1];
    fill_camera(k, pc->camera);
    k.rect_x0 = pc->tile_offset[0]; k.rect_y0 = pc->tile_offset[1];
    k.rect_w = pc->tile_size_packed & 0xFFFFu; k.rect_h = (pc->tile_size_packed >> 16) & 0xFFFFu;
    k.use_tiles = 0;
    k.mode = rt_pc_wavefront_mode(pc) != 0 ? RT_MODE_WAVEFRONT0 : RT_MODE_LEGACY;
    k.cur_depth = rt_pc_current_bounce(pc); k.max_depth = rt_pc_max_bounce(pc);
    k.outputs = RT_OUT_CHANNELS | RT_OUT_HITS | RT_OUT_ACCUM;
    k.channel = channel;
    k.triangles_per_buffer = pc->triangles_per_buffer;
    k.triangle_count = pc->triangle_count;
    uint32_t bw = (k.rect_w + 15u) / 16u, bh = (k.rect_h + 15u) / 16u; // src/compute.rs:248-250
    return launch(c, k, bw * bh, RT_TRAVERSAL_AUTO);
}

int rt_render_frame(RtContext* c, const RtCamera* cam, const RtFrameParams* fp) {
    int r = check_ready(c);
    if (r) return r;
    if (!cam || !fp) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_render_frame: NULL argument");
    if (fp->mode > RT_MODE_PATH) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_render_frame: mode %u unknown", fp->mode);
    if (fp->mode == RT_MODE_PATH && (fp->spp == 0 || fp->max_depth > 255))
        return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_render_frame: path mode needs spp >= 1 and max_depth <= 255");
    if (fp->mode == RT_MODE_PATH && fp->_reserved[0] != 0 && fp->_reserved[0] != 1 && fp->_reserved[0] != 4)
        return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_render_frame: path engine %u unknown (0 auto, 1 megakernel, 4 streaming wavefront)", fp->_reserved[0]);
    CK(c, cudaSetDevice(c->device));
    if ((r = prepare_relayout(c)) != RT_OK) return r;
    KParams k;
    memset(&k, 0, sizeof k);
    fill_scene(c, k);
    k.width = c->width; k.height = c->height; k.res_x = (float)c->width; k.res_y = (float)c->height;
    fill_camera(k, *cam);
    k.use_tiles = 1;
    k.tiles_x = c->stiles_x; k.n_tiles_total = c->stiles_x * c->stiles_y;
    if (c->tile_list_n) { k.tile_list = c->tile_list.ptr; k.n_tiles_total = c->tile_list_n; } // stride / phase walk the list
    k.tile_stride = fp->tile_stride ? fp->tile_stride : 1u;
    k.tile_phase = fp->tile_stride ? fp->tile_phase : 0u;
    if (k.tile_phase >= k.tile_stride) return fail(c, RT_ERR_INVALID_ARGUMENT, "tile_phase >= tile_stride");
    k.mode = fp->mode; k.outputs = fp->outputs; k.channel = 0xFFu;
    if (fp->outputs & RT_OUT_PEER) { // fused multi-GPU epilogue: composite stores go to rank 0's frame
        if (!c->peer_composite) return fail(c, RT_ERR_NOT_READY, "RT_OUT_PEER without rt_open_peer_frame / rt_set_peer_frame");
        k.composite = c->peer_composite;
    }
    k.cur_depth = 0; k.max_depth = 4; // PushConstants::new, shared/src/lib.rs:1088
    k.reset_accum = fp->reset_accum;
    k.triangles_per_buffer = c->tpb;
    k.triangle_count = (uint32_t)c->n_tris;
    uint32_t my_tiles = k.n_tiles_total > k.tile_phase ? (k.n_tiles_total - k.tile_phase + k.tile_stride - 1) / k.tile_stride : 0;
    if (fp->mode == RT_MODE_PATH) {
        k.spp = fp->spp; k.first_sample = fp->first_sample; k.path_max_depth = fp->max_depth; k.base_seed = fp->base_seed;
        k.n_work = my_tiles * 4096u;
        k.seg_counter = c->dcounters.ptr + 6;
        k.work_counter = reinterpret_cast<unsigned int*>(c->dcounters.ptr + 7);
        CK(c, cudaMemsetAsync(c->dcounters.ptr + 6, 0, 2 * sizeof(unsigned long long), c->stream));
        c->last_was_wavefront = false;
        c->wv_event_stage.clear();
        // _reserved[0] selects the path-tracing engine: 0 auto = the streaming wavefront engine (rt_wave.cuh)
        // where a re-laid-out traversal policy is available, else the megakernel; 1 megakernel with path
        // regeneration (rt_path_trace: the LITERAL policy's engine and the cross-check of the wave engine);
        // 4 streaming wavefront, megakernel if unavailable
        if (fp->_reserved[0] == 4 || fp->_reserved[0] == 0) {
            int wr = launch_wave(c, k, fp->traversal);
            if (wr != -100) return wr;
        }
        return launch_path(c, k, fp->traversal);
    }
    {   // reference-semantics primary frame: the wave engine's traversal where a re-laid-out policy exists
        int wr = launch_primary_wave(c, k, fp->traversal, my_tiles);
        if (wr != -100) return wr;
    }
    return launch(c, k, my_tiles * 16u, fp->traversal);
}

int rt_run_compute(RtContext* c, const RtCamera* cam, int needs_recompute) {
    int r = check_ready(c);
    if (r) return r;
    if (!cam) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_run_compute: camera is NULL");
    if (needs_recompute) { c->needs_recompute = true; c->current_tile = 0; c->is_progressive = false; } // trigger_recompute
    // handle_progressive_rendering_setup, src/compute.rs:53-82
    if (c->needs_recompute && !c->is_progressive) { c->is_progressive = true; c->current_tile = 0; c->needs_recompute = false; }
    if (!c->is_progressive) return 0;
    uint32_t total = c->tiles_x * c->tiles_y;
    if (c->current_tile >= total) { c->is_progressive = false; return 0; } // check_rendering_completion :85-100
    uint32_t remaining = total - c->current_tile;
    uint32_t this_frame = c->tiles_per_frame < remaining ? c->tiles_per_frame : remaining; // :103-106
    for (uint32_t i = 0; i < this_frame; i++) { // execute_compute_pass :157-162
        uint32_t tile = c->current_tile + i;
        uint32_t tx = tile % c->tiles_x, ty = tile / c->tiles_x; // calculate_tile_dimensions :194-209
        uint32_t ox = tx * RT_TILE_SIZE, oy = ty * RT_TILE_SIZE;
        uint32_t tw = RT_TILE_SIZE < c->width - ox ? RT_TILE_SIZE : c->width - ox;
        uint32_t th = RT_TILE_SIZE < c->height - oy ? RT_TILE_SIZE : c->height - oy;
        for (uint32_t ch = 0; ch < 3; ch++) { // process_tile :184-190
            RtPushConstants pc; // PushConstants::new, src/compute.rs:232-243
            memset(&pc, 0, sizeof pc);
            pc.resolution[0] = (float)c->width; pc.resolution[1] = (float)c->height;
            pc.camera = *cam;
            pc.triangle_count = (uint32_t)c->n_tris; pc.material_count = (uint32_t)c->n_mats;
            pc.tile_offset[0] = ox; pc.tile_offset[1] = oy;
            pc.tile_size_packed = rt_pack_tile_size(tw, th);
            pc.total_tiles[0] = c->tiles_x; pc.total_tiles[1] = c->tiles_y;
            pc.triangles_per_buffer = c->tpb; // buffer_manager.triangles_per_buffer, src/compute.rs:239
            pc.metadata_offsets = c->offsets;
            pc.packed_flags = rt_pack_flags(ch, 0, 4, 0);
            pc.frame_seed = 0;
            if ((r = rt_dispatch(c, &pc)) != RT_OK) return r;
        }
    }
    c->current_tile += this_frame;
    return (int)(total - c->current_tile);
}

static int read_back(RtContext* c, void* dst, const void* src, size_t bytes) {
    if (!c || !dst) return RT_ERR_INVALID_ARGUMENT;
    if (c->width == 0) return fail(c, RT_ERR_NOT_READY, "rt_resize has not been called");
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}
int rt_read_channel_rgba8(RtContext* c, uint32_t channel, void* dst) {
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (channel > 2u) return fail(c, RT_ERR_INVALID_CHANNEL, "Invalid color channel %u", channel);
    return read_back(c, dst, c->chan[channel].ptr, (size_t)c->width * c->height * 4);
}
int rt_read_composite_rgba8(RtContext* c, void* dst) { return c ? read_back(c, dst, c->composite.ptr, (size_t)c->width * c->height * 4) : RT_ERR_INVALID_ARGUMENT; }
int rt_read_accum_f32(RtContext* c, void* dst) { return c ? read_back(c, dst, c->accum.ptr, (size_t)c->width * c->height * 16) : RT_ERR_INVALID_ARGUMENT; }
int rt_read_hit_records(RtContext* c, void* dst) { return c ? read_back(c, dst, c->hits.ptr, (size_t)c->width * c->height * 16) : RT_ERR_INVALID_ARGUMENT; }
int rt_write_accum_f32(RtContext* c, const void* src) {
    if (!c || !src) return RT_ERR_INVALID_ARGUMENT;
    if (c->width == 0) return fail(c, RT_ERR_NOT_READY, "rt_resize has not been called");
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemcpyAsync(c->accum.ptr, src, (size_t)c->width * c->height * 16, cudaMemcpyHostToDevice, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}
void* rt_device_accum(RtContext* c) { return c ? c->accum.ptr : nullptr; }
void* rt_device_composite(RtContext* c) { return c ? c->composite.ptr : nullptr; }
float rt_last_kernel_ms(RtContext* c) {
    if (!c || !c->timed) return -1.0f;
    cudaSetDevice(c->device);
    if (cudaEventSynchronize(c->ev1) != cudaSuccess) return -1.0f;
    float ms = -1.0f;
    if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) != cudaSuccess) return -1.0f;
    return ms;
}
// Instrumentation: the next render/dispatch launch counts box tests, triangle tests,
// ray segments and shaded hits (atomics; not for timing).  Read with rt_read_counters.
int rt_count_next_launch(RtContext* c) {
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemsetAsync(c->dcounters.ptr, 0, 6 * sizeof(unsigned long long), c->stream));
    CK(c, cudaMemsetAsync(c->dcounters.ptr + 8, 0, 8 * sizeof(unsigned long long), c->stream));
    c->count_next = true;
    return RT_OK;
}
int rt_read_counters(RtContext* c, uint64_t out[9]) {
    if (!c || !out) return RT_ERR_INVALID_ARGUMENT;
    CK(c, cudaSetDevice(c->device));
    unsigned long long h[12];
    CK(c, cudaMemcpyAsync(h, c->dcounters.ptr, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 5; i++) out[i] = h[i];
    for (int i = 0; i < 4; i++) out[5 + i] = h[8 + i]; // path kernel: S phases, ready lanes, T iterations, traversing lanes
    return RT_OK;
}
uint64_t rt_launch_count(const RtContext* c) { return c ? c->launches : 0; }
int rt_last_traversal(const RtContext* c) { return c ? c->last_traversal : 0; }
void* rt_stream(RtContext* c) { return c ? (void*)c->stream : nullptr; }

// what: 0 = composite rgba8 (4 B/px), 1 = accumulation float4 (16 B/px)
// list != NULL: a host array of n tiles (staged into the context's scratch list buffer); else stride / phase over
// the context's own assignment (its tile list if one is set, otherwise tile i -> i % stride)
static int tile_copy_launch(RtContext* c, uint32_t what, uint32_t stride, uint32_t phase, const uint32_t* list, uint32_t n,
                            void* compact, bool pack) {
    if (!c || !compact) return RT_ERR_INVALID_ARGUMENT;
    if (c->width == 0) return fail(c, RT_ERR_NOT_READY, "rt_resize has not been called");
    if (stride == 0 || phase >= stride || what > 1) return fail(c, RT_ERR_INVALID_ARGUMENT, "tile copy: bad stride/phase/what");
    CK(c, cudaSetDevice(c->device));
    const uint32_t frame_tiles = c->stiles_x * c->stiles_y;
    const uint32_t* dlist = nullptr;
    uint32_t total = frame_tiles;
    if (list) {
        for (uint32_t i = 0; i < n; i++) if (list[i] >= frame_tiles) return fail(c, RT_ERR_INVALID_ARGUMENT, "tile %u of %u", list[i], frame_tiles);
        if (n > c->copy_list.capacity) CK(c, cudaStreamSynchronize(c->stream));
        if (c->copy_list.reserve(n ? n : 1) < 0) return fail(c, RT_ERR_CUDA, "tile list cudaMalloc failed");
        if (n) CK(c, cudaMemcpyAsync(c->copy_list.ptr, list, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
        CK(c, cudaStreamSynchronize(c->stream)); // `list` is borrowed for the call
        dlist = c->copy_list.ptr; total = n;
    } else if (c->tile_list_n) { dlist = c->tile_list.ptr; total = c->tile_list_n; }
    uint32_t mine = total > phase ? (total - phase + stride - 1) / stride : 0;
    if (!mine) return RT_OK;
    dim3 grid(4096 / 256, mine);
    if (what == 0) {
        if (pack) tile_copy<uchar4, true><<<grid, 256, 0, c->stream>>>(c->composite.ptr, (uchar4*)compact, c->width, c->height, c->stiles_x, total, stride, phase, dlist);
        else tile_copy<uchar4, false><<<grid, 256, 0, c->stream>>>(c->composite.ptr, (uchar4*)compact, c->width, c->height, c->stiles_x, total, stride, phase, dlist);
    } else {
        if (pack) tile_copy<float4, true><<<grid, 256, 0, c->stream>>>(c->accum.ptr, (float4*)compact, c->width, c->height, c->stiles_x, total, stride, phase, dlist);
        else tile_copy<float4, false><<<grid, 256, 0, c->stream>>>(c->accum.ptr, (float4*)compact, c->width, c->height, c->stiles_x, total, stride, phase, dlist);
    }
    CK(c, cudaGetLastError());
    c->launches++;
    return RT_OK;
}
int rt_pack_tiles(RtContext* c, uint32_t what, uint32_t stride, uint32_t phase, void* dst_device) {
    return tile_copy_launch(c, what, stride, phase, nullptr, 0, dst_device, true);
}
int rt_unpack_tiles(RtContext* c, uint32_t what, uint32_t stride, uint32_t phase, const void* src_device) {
    return tile_copy_launch(c, what, stride, phase, nullptr, 0, const_cast<void*>(src_device), false);
}
int rt_pack_tile_list(RtContext* c, uint32_t what, const uint32_t* tiles, uint32_t n, void* dst_device) {
    if (n && !tiles) return RT_ERR_INVALID_ARGUMENT;
    static const uint32_t none = 0;
    return tile_copy_launch(c, what, 1, 0, tiles ? tiles : &none, n, dst_device, true);
}
// ---- rt_pack_tile_list without any allocation -----------------------------------------------------------------
// rt_pack_tile_list stages the caller's tile list in a device scratch buffer that it (re)allocates with cudaMalloc
// when the list grows.  Callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc) use this variant: the list is staged in ONE file-scope device array of RT_STATIC_TILE_LIST_MAX
// entries (64 KiB; every GPU that loads the library holds it until exit).  In host code the array's name is a host-side
// stand-in, not a device address: the address comes from cudaGetSymbolAddress, taken once per device with that device
// current.  All calls on a GPU share the array (it keeps the last call's list; every call overwrites the n entries it
// uses, nothing has to start at zero), so calls that overlap on the GPU (other streams / contexts of the same device)
// race.  No cudaMalloc / cudaFree / allocating library call, and no host wait.
static __device__ __align__(256) uint32_t g_static_tile_list[RT_STATIC_TILE_LIST_MAX];
static uint32_t* g_static_tile_list_addr[64]; // per device ordinal; filled on the first call for that device

int rt_pack_tile_list_static(RtContext* c, uint32_t what, const uint32_t* tiles, uint32_t n, void* dst_device) {
    if (!c || !dst_device || (n && !tiles)) return c ? fail(c, RT_ERR_INVALID_ARGUMENT, "rt_pack_tile_list_static: NULL argument") : RT_ERR_INVALID_ARGUMENT;
    if (c->width == 0) return fail(c, RT_ERR_NOT_READY, "rt_resize has not been called");
    if (what > 1) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_pack_tile_list_static: what = %u (0 composite, 1 accumulation)", what);
    if (n > RT_STATIC_TILE_LIST_MAX)
        return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_pack_tile_list_static: %u tiles exceed the static scratch of %u entries "
                                                "(use rt_pack_tile_list, which allocates)", n, (unsigned)RT_STATIC_TILE_LIST_MAX);
    if (c->device < 0 || c->device >= 64) return fail(c, RT_ERR_INVALID_ARGUMENT, "rt_pack_tile_list_static: device ordinal %d", c->device);
    const uint32_t frame_tiles = c->stiles_x * c->stiles_y;
    for (uint32_t i = 0; i < n; i++) if (tiles[i] >= frame_tiles) return fail(c, RT_ERR_INVALID_ARGUMENT, "tile %u of %u", tiles[i], frame_tiles);
    CK(c, cudaSetDevice(c->device));
    uint32_t* list = g_static_tile_list_addr[c->device];
    if (!list) { // set-up, once per device, with that device current
        CK(c, cudaGetSymbolAddress((void**)&list, g_static_tile_list));
        if (!list) return fail(c, RT_ERR_CUDA, "cudaGetSymbolAddress returned NULL");
        g_static_tile_list_addr[c->device] = list;
    }
    if (n == 0) return RT_OK;
    // stream-ordered like every other call: `tiles` is read when the stream reaches this copy (pageable memory is
    // staged before the call returns, pinned memory must stay valid until then)
    CK(c, cudaMemcpyAsync(list, tiles, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    dim3 grid(4096 / 256, n);
    if (what == 0) tile_copy<uchar4, true><<<grid, 256, 0, c->stream>>>(c->composite.ptr, (uchar4*)dst_device, c->width, c->height, c->stiles_x, n, 1u, 0u, list);
    else tile_copy<float4, true><<<grid, 256, 0, c->stream>>>(c->accum.ptr, (float4*)dst_device, c->width, c->height, c->stiles_x, n, 1u, 0u, list);
    CK(c, cudaGetLastError());
    c->launches++;
    return RT_OK;
}

int rt_unpack_tile_list(RtContext* c, uint32_t what, const uint32_t* tiles, uint32_t n, const void* src_device) {
    if (n && !tiles) return RT_ERR_INVALID_ARGUMENT;
    static const uint32_t none = 0;
    return tile_copy_launch(c, what, 1, 0, tiles ? tiles : &none, n, const_cast<void*>(src_device), false);
}
// Explicit tile assignment of this context for fused frames (multi-GPU load balancing): `tiles` = the split tiles
// (numbered row-major, RT_SPLIT_TILE^2 pixels) it renders, in the order it walks them.  n = 0 returns to the
// arithmetic assignment tile i -> i % tile_stride.
int rt_set_tile_list(RtContext* c, const uint32_t* tiles, uint32_t n) {
    if (!c || (n && !tiles)) return RT_ERR_INVALID_ARGUMENT;
    if (c->width == 0) return fail(c, RT_ERR_NOT_READY, "rt_resize has not been called");
    CK(c, cudaSetDevice(c->device));
    const uint32_t frame_tiles = c->stiles_x * c->stiles_y;
    for (uint32_t i = 0; i < n; i++) if (tiles[i] >= frame_tiles) return fail(c, RT_ERR_INVALID_ARGUMENT, "tile %u of %u", tiles[i], frame_tiles);
    CK(c, cudaStreamSynchronize(c->stream)); // frames in flight still read the old list
    if (n) {
        if (c->tile_list.reserve(n) < 0) return fail(c, RT_ERR_CUDA, "tile list cudaMalloc failed");
        CK(c, cudaMemcpyAsync(c->tile_list.ptr, tiles, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
        CK(c, cudaStreamSynchronize(c->stream));
    }
    c->tile_list_host.assign(tiles, tiles + n);
    c->tile_list_n = n;
    return RT_OK;
}

int rt_export_frame_handle(RtContext* c, void* out_handle64) {
    if (!c || !out_handle64) return RT_ERR_INVALID_ARGUMENT;
    if (c->width == 0) return fail(c, RT_ERR_NOT_READY, "rt_resize has not been called");
    CK(c, cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle is 64 bytes");
    CK(c, cudaIpcGetMemHandle(&h, c->composite.ptr));
    memcpy(out_handle64, &h, 64);
    return RT_OK;
}
int rt_close_peer_frame(RtContext* c) {
    if (!c) return RT_ERR_INVALID_ARGUMENT;
    if (c->peer_composite && c->peer_is_ipc) { cudaSetDevice(c->device); cudaStreamSynchronize(c->stream); cudaIpcCloseMemHandle(c->peer_composite); }
    c->peer_composite = nullptr; c->peer_is_ipc = false;
    return RT_OK;
}
int rt_open_peer_frame(RtContext* c, const void* handle64) {
    if (!c || !handle64) return RT_ERR_INVALID_ARGUMENT;
    CK(c, cudaSetDevice(c->device));
    rt_close_peer_frame(c);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void* p = nullptr;
    CK(c, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->peer_composite = (uchar4*)p; c->peer_is_ipc = true;
    return RT_OK;
}
int rt_set_peer_frame(RtContext* c, void* device_composite) {
    if (!c || !device_composite) return RT_ERR_INVALID_ARGUMENT;
    rt_close_peer_frame(c);
    c->peer_composite = (uchar4*)device_composite; c->peer_is_ipc = false;
    return RT_OK;
}

// ray segments traced by the last RT_MODE_PATH launch (waits for it)
int rt_last_segments(RtContext* c, uint64_t* out) {
    if (!c || !out) return RT_ERR_INVALID_ARGUMENT;
    CK(c, cudaSetDevice(c->device));
    unsigned long long h = 0;
    CK(c, cudaMemcpyAsync(&h, c->dcounters.ptr + 6, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    *out = h;
    return RT_OK;
}
float rt_last_build_ms(const RtContext* c) { return c ? c->last_build_ms : -1.0f; }

// WavefrontCounters of the last streaming-wavefront frame (shared/src/lib.rs:187-194, 1006-1043): what the
// reference's host loop (src/compute.rs:413-479) only simulates with a 0.7^depth decay is counted here.
int rt_last_wavefront_counters(RtContext* c, RtWavefrontCounters* out) {
    if (!c || !out) return RT_ERR_INVALID_ARGUMENT;
    if (!c->last_was_wavefront) return fail(c, RT_ERR_NOT_READY, "the last launch was not a wave-engine frame");
    CK(c, cudaSetDevice(c->device));
    unsigned long long h[8];
    CK(c, cudaMemcpyAsync(h, c->dcounters.ptr + 16, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    memset(out, 0, sizeof *out);
    unsigned long long total = 0;
    for (int d = 0; d < 8; d++) {
        out->rays_per_bounce[d] = h[d] > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)h[d]; // u32 like the reference; saturating
        if (h[d]) out->active_bounce_depths |= 1u << d;                                  // WavefrontCounters::add_rays
        total += h[d];
    }
    out->total_rays_generated = total > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)total;
    out->max_bounce_depth = c->last_wave_max_depth;
    out->frame_seed = c->last_wave_seed;
    return RT_OK;
}

// GPU time per stage of the last streaming-wavefront frame (rt_wave.cuh), summed over its
// launches: out[0] generate, [1] extend, [2] shade, [3] resolve (ms); out[4..7] launch counts.
int rt_last_stage_ms(RtContext* c, float out[8]) {
    if (!c || !out) return RT_ERR_INVALID_ARGUMENT;
    for (int i = 0; i < 8; i++) out[i] = 0.0f;
    if (!c->last_was_wavefront || c->wv_event_stage.empty()) return fail(c, RT_ERR_NOT_READY, "the last launch was not a wave-engine frame");
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaStreamSynchronize(c->stream));
    size_t e = 0;
    for (size_t i = 0; i < c->wv_event_stage.size(); i++) {
        int st = c->wv_event_stage[i];
        if (st == 0 && i != 0) e++; // a new wave starts with its own opening event
        float ms = 0.0f;
        CK(c, cudaEventElapsedTime(&ms, c->wv_events[e], c->wv_events[e + 1]));
        const int bucket = st >= 4 ? 2 : st; // shade = its in-place form (2), its split form (4) and the fallback launch (5)
        out[bucket] += ms; if (st < 4) out[4 + st] += 1.0f;
        e++;
    }
    return RT_OK;
}

// The same events one by one: GPU ms and stage (0 generate, 1 extend, 2 shade, 3 resolve) of every launch interval
// of the last single-sub-frame wave-engine frame, in launch order.  Returns the number of intervals (<= cap written).
int rt_last_launch_table(RtContext* c, float* ms, int32_t* stage, uint32_t cap) {
    if (!c || !ms || !stage) return RT_ERR_INVALID_ARGUMENT;
    if (!c->last_was_wavefront || c->wv_event_stage.empty()) return fail(c, RT_ERR_NOT_READY, "the last launch was not a single-sub-frame wave-engine frame");
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaStreamSynchronize(c->stream));
    size_t e = 0;
    uint32_t n = 0;
    for (size_t i = 0; i < c->wv_event_stage.size(); i++) {
        int st = c->wv_event_stage[i];
        if (st == 0 && i != 0) e++;
        float t = 0.0f;
        CK(c, cudaEventElapsedTime(&t, c->wv_events[e], c->wv_events[e + 1]));
        if (n < cap) { ms[n] = t; stage[n] = st; }
        n++; e++;
    }
    return (int)n;
}

// Measured gather bandwidth (GB/s) for random independent 32-byte sector loads over a
// footprint of `footprint_bytes` (rounded down to a power of two, >= 4 KiB): the roofline
// denominators of the traversal kernels (L1-resident, L2-resident, HBM-resident node sets).
int rt_probe_gather_bandwidth(RtContext* c, uint64_t footprint_bytes, float* out_gbs) {
    if (!c || !out_gbs) return RT_ERR_INVALID_ARGUMENT;
    CK(c, cudaSetDevice(c->device));
    uint64_t fp = 4096;
    while (fp * 2 <= footprint_bytes && fp < (1ull << 34)) fp *= 2;
    int sms = 0;
    CK(c, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    const uint32_t blocks = (uint32_t)sms * 8u, iters = 4096u;
    float4* buf = nullptr; float* out = nullptr;
    CK(c, cudaMalloc((void**)&buf, fp));
    if (cudaMalloc((void**)&out, (size_t)blocks * 256 * 4) != cudaSuccess) { cudaFree(buf); return fail(c, RT_ERR_CUDA, "probe cudaMalloc failed"); }
    cudaMemsetAsync(buf, 0, fp, c->stream);
    const uint32_t mask = (uint32_t)(fp / 32 - 1);
    probe_gather<<<blocks, 256, 0, c->stream>>>(buf, mask, 256u, out); // warm-up
    cudaEventRecord(c->ev0, c->stream);
    probe_gather<<<blocks, 256, 0, c->stream>>>(buf, mask, iters, out);
    cudaEventRecord(c->ev1, c->stream);
    cudaError_t e = cudaEventSynchronize(c->ev1);
    float ms = 0.0f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    cudaFree(buf); cudaFree(out);
    c->launches += 2;
    c->timed = false;
    if (e != cudaSuccess || !(ms > 0.0f)) return fail(c, RT_ERR_CUDA, "probe_gather failed: %s", cudaGetErrorString(e));
    *out_gbs = (float)((double)blocks * 256.0 * iters * 32.0 / (ms * 1e-3) / 1e9);
    return RT_OK;
}

} // extern "C"
