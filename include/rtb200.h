/* rtb200.h — C ABI of librtb200.so: the B200 (sm_100a) implementation of the
 * reference's per-pixel ray tracing path, behind the reference's own dispatch
 * surface.  Plain pointers and sizes only; every entry point cites the
 * reference interface it replaces (paths relative to the reference tree).
 *
 * Conventions: functions return 0 (RT_OK) / a positive documented value on
 * success and a negative RtStatus on failure; the text is in rt_last_error().
 * Host memory passed in is borrowed for the duration of the call only
 * (same as wgpu `queue.write_buffer`, src/buffers.rs:236).  All GPU work is
 * asynchronous on the context's stream (same as `queue.submit`,
 * src/compute.rs:165); rt_synchronize / rt_read_* wait for it.
 * A context is bound to one GPU and is not thread-safe; use one per GPU.
 * There is no CPU fallback: without a CUDA device rt_create fails.
 */
#ifndef RTB200_H
#define RTB200_H
#include <stdint.h>
#include "rt_shared.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef enum RtStatus {
    RT_OK = 0,
    RT_ERR_INVALID_ARGUMENT = -1,
    RT_ERR_CUDA = -2,
    RT_ERR_INVALID_SCENE = -3,   /* upload-time validation failed */
    RT_ERR_INVALID_CHANNEL = -4, /* src/compute.rs:228-229 `.expect("Invalid color channel")` */
    RT_ERR_NOT_READY = -5,
    RT_ERR_NCCL = -6
} RtStatus;

typedef struct RtContext RtContext;

/* Closest-hit record per pixel (the reference's IntersectionResult,
 * shader/src/intersection.rs:10-38, plus a primitive id). */
typedef struct RtHitRecord {
    uint32_t kind;        /* 0 miss, 1 sphere, 2 triangle */
    uint32_t prim;        /* sphere index / triangle index; 0xFFFFFFFF on miss */
    uint32_t t_bits;      /* IEEE-754 bits of t; f32::MAX on miss */
    uint32_t material_id; /* 0xFFFFFFFF on miss */
} RtHitRecord;

enum {
    RT_OUT_ACCUM = 1u,     /* float4 accumulation buffer (rgb sum, sample count) */
    RT_OUT_HITS = 2u,      /* RtHitRecord per pixel (sample 0 / primary ray) */
    RT_OUT_CHANNELS = 4u,  /* the three Rgba8Unorm channel images (src/renderer.rs:452-475) */
    RT_OUT_COMPOSITE = 8u, /* main_fs composite (R.x,G.y,B.z,1), shader/src/lib.rs:368-391 */
    RT_OUT_PEER = 16u      /* with RT_OUT_COMPOSITE: store the composite into the peer frame opened with
                            * rt_open_peer_frame / rt_set_peer_frame (multi-GPU fused epilogue) instead of this
                            * context's own composite buffer */
};
enum {
    RT_MODE_LEGACY = 0,    /* main_cs legacy branch, shader/src/lib.rs:59-79 (what src/main.rs runs) */
    RT_MODE_WAVEFRONT0 = 1,/* main_cs wavefront branch at depth 0, lib.rs:92-149 */
    RT_MODE_PATH = 2       /* EXTENSION (reference stubs it): bounces + spp, SPEC.md */
};
/* LITERAL: the reference's own DFS.  ORDERED: near-first + t-pruned over the uploaded
 * tree.  REBUILT: an internal small-leaf LBVH built on the GPU at upload, each winner
 * certified against the uploaded tree.  All three return identical hit records.
 * AUTO picks REBUILT for big-leaf trees, else ORDERED, else LITERAL. */
enum { RT_TRAVERSAL_AUTO = 0, RT_TRAVERSAL_LITERAL = 1, RT_TRAVERSAL_ORDERED = 2, RT_TRAVERSAL_REBUILT = 3 };

/* Unit of the multi-GPU screen split (RtFrameParams.tile_stride / tile_phase, rt_pack_tiles): the frame is
 * cut into RT_SPLIT_TILE x RT_SPLIT_TILE pixel tiles numbered row-major, like the reference's own
 * 128 x 128 progressive tiles (src/compute.rs:194-209) but finer, so that 8 ranks get equal shares of a
 * 1080p frame (510 tiles instead of 135). */
#define RT_SPLIT_TILE 64u

/* Whole-frame parameters of the fused path (no reference analogue: replaces the
 * per-tile x per-channel loop of src/compute.rs:157-191 by one launch). */
typedef struct RtFrameParams {
    uint32_t mode;         /* RT_MODE_* */
    uint32_t outputs;      /* RT_OUT_* mask */
    uint32_t spp;          /* samples added by this call (>=1) */
    uint32_t first_sample; /* index of the first sample (progressive resume) */
    uint32_t max_depth;    /* RT_MODE_PATH: max bounce depth */
    uint32_t base_seed;    /* frame_seed of sample s is base_seed + s */
    uint32_t reset_accum;  /* 1: overwrite accumulation; 0: add to it */
    uint32_t tile_stride;  /* multi-GPU tile split: this context renders the RT_SPLIT_TILE^2 */
    uint32_t tile_phase;   /*   tiles i with i % tile_stride == tile_phase (0,0 = all) */
    uint32_t traversal;    /* RT_TRAVERSAL_* */
    uint32_t _reserved[2]; /* [0]: RT_MODE_PATH engine (0 auto = streaming wavefront where a re-laid-out policy is
                            * available, else megakernel; tuning/tests: 1 megakernel, 4 streaming wavefront); [1]: 0 */
} RtFrameParams;

/* Environment switches read by rt_render_frame (tuning and tests only; results never change):
 *   RTB200_WAVE_SPLIT=1..4  sub-frames of the streaming wavefront engine (default 2; 1 enables rt_last_stage_ms)
 *   RTB200_WAVE_PATHS=N     paths in flight per wave (default 40 M)
 *   RTB200_WAVE_DEEP=1      force the traversal-stack form used for trees deeper than 30 levels
 *   RTB200_PRIMARY_WAVE=1   primary frames through wv_extend + wv_shade_primary instead of the fused rt_trace_shade
 *   RTB200_EXT_BLOCKS_PER_SM, RTB200_SHADE_BLOCKS_PER_SM, RTB200_SPLIT_GEN, RTB200_SPLIT_DEEP   launch-shape tuning
 *   RTB200_DEBUG=1          print the internal tree's size and depth at upload */

/* ---- context: stands in for RenderState.device/queue + the three output
 * textures (src/renderer.rs:93-246, 452-475) and BufferManager::new
 * (src/buffers.rs:66-154). */
int rt_create(int device, RtContext** out);
void rt_destroy(RtContext* ctx);
const char* rt_last_error(const RtContext* ctx); /* ctx may be NULL: last create error */
/* RenderState::resize + ProgressiveState::resize (src/renderer.rs:477, 838) */
int rt_resize(RtContext* ctx, uint32_t width, uint32_t height);
int rt_synchronize(RtContext* ctx);

/* ---- BufferManager (src/buffers.rs).  Return 1 if the device buffer was
 * (re)allocated, 0 if not, <0 on error — the reference returns `bool resized`. */
/* update_scene_metadata, src/buffers.rs:157-271; offsets per :257-268 */
int rt_update_scene_metadata(RtContext* ctx, const RtSphere* spheres, uint32_t n_spheres,
                             const RtLight* lights, uint32_t n_lights, const RtBvhNode* nodes,
                             uint32_t n_nodes, const uint32_t* triangle_indices, uint32_t n_indices,
                             const RtVertex* vertices, uint32_t n_vertices,
                             RtSceneMetadataOffsets* out_offsets);
/* update_triangles, src/buffers.rs:274-336 (split over <=3 bindings there; one flat array here) */
int rt_update_triangles(RtContext* ctx, const RtTriangle* triangles, uint64_t n);
/* update_materials, src/buffers.rs:339-378 */
int rt_update_materials(RtContext* ctx, const RtMaterial* materials, uint64_t n);
/* update_textures, src/buffers.rs:381-419 (bound but never read by the shader) */
int rt_update_textures(RtContext* ctx, const RtTextureInfo* textures, uint64_t n);
/* update_texture_data, src/buffers.rs:422-470 (bytes packed little-endian into u32) */
int rt_update_texture_data(RtContext* ctx, const uint8_t* bytes, uint64_t n_bytes);
/* mark_*_dirty / needs_update, src/buffers.rs:473-501 */
void rt_mark_scene_metadata_dirty(RtContext* ctx);
void rt_mark_triangles_dirty(RtContext* ctx);
void rt_mark_materials_dirty(RtContext* ctx);
void rt_mark_textures_dirty(RtContext* ctx);
void rt_mark_texture_data_dirty(RtContext* ctx);
void rt_mark_all_dirty(RtContext* ctx);
int rt_needs_update(const RtContext* ctx);
/* BufferManager.triangles_per_buffer (src/buffers.rs:49-53, 103): length of every bound triangle buffer and the
 * divisor run_compute puts into the push constants (src/compute.rs:239).  The reference derives it from a
 * 128 MB buffer limit when the manager is constructed (8 388 608); rt_set_triangles_per_buffer stands for
 * constructing the manager with another limit and is only accepted before the first triangle upload.  The
 * flat device array keeps the reference's binding rule as index arithmetic: buffer k of max(1, ceil(n / B))
 * holds triangles [k B, (k+1) B), zero-filled past n; a binding whose buffer does not exist aliases buffer 0
 * (src/renderer.rs:680-693); buffer index > 2 is invalid (shader/src/triangle_access.rs:26-48). */
uint32_t rt_triangles_per_buffer(const RtContext* ctx);
int rt_set_triangles_per_buffer(RtContext* ctx, uint32_t triangles_per_buffer);
/* offsets of the last metadata upload (what run_compute passes on, src/compute.rs:109-134) */
int rt_get_metadata_offsets(const RtContext* ctx, RtSceneMetadataOffsets* out);

/* ComputeRenderer::update_buffers_and_bind_groups ends with the bind-group rebuild (src/compute.rs:109-134,
 * src/renderer.rs:496-520); its counterpart here is the device-side re-layout of the uploaded scene (two-child nodes,
 * pre-gathered triangles, internal LBVH), which otherwise runs lazily inside the next dispatch / frame.  Calling it
 * right after the rt_update_* calls lets a host that double-buffers scenes over two contexts do all per-scene work on
 * its upload thread.  Idempotent; waits for the build. */
int rt_prepare_scene(RtContext* ctx);

/* ---- dispatch (src/compute.rs) */
/* One reference dispatch: process_color_channel, src/compute.rs:212-251 —
 * main_cs over ceil(w/16) x ceil(h/16) groups with these push constants; writes
 * the channel's Rgba8Unorm image. */
int rt_dispatch(RtContext* ctx, const RtPushConstants* pc);
/* ComputeRenderer::run_compute, src/compute.rs:12-50: progressive step — renders
 * the next tiles_per_frame tiles x 3 channels (TileHelper, shared/src/lib.rs:
 * 1187-1203; ProgressiveState, src/renderer.rs:821-855).  needs_recompute!=0
 * restarts at tile 0.  Returns the number of tiles still pending after the call. */
int rt_run_compute(RtContext* ctx, const RtCamera* camera, int needs_recompute);
/* Fused whole frame: all tiles, three channels, composite and accumulation in
 * one launch. */
int rt_render_frame(RtContext* ctx, const RtCamera* camera, const RtFrameParams* params);

/* ---- outputs (the reference never reads back; these are the headless
 * replacement of the surface present, src/renderer.rs:778-818) */
int rt_read_channel_rgba8(RtContext* ctx, uint32_t channel, void* dst); /* W*H*4 bytes */
int rt_read_composite_rgba8(RtContext* ctx, void* dst);                /* W*H*4 bytes */
int rt_read_accum_f32(RtContext* ctx, void* dst);                      /* W*H*16 bytes */
int rt_write_accum_f32(RtContext* ctx, const void* src);               /* resume */
int rt_read_hit_records(RtContext* ctx, void* dst);                    /* W*H*16 bytes */
/* raw device pointers for zero-copy consumers (torch / NCCL) */
void* rt_device_accum(RtContext* ctx);
void* rt_device_composite(RtContext* ctx);
/* GPU time of the last rt_render_frame / rt_dispatch launch in ms (CUDA events
 * on the context stream); <0 if none. */
float rt_last_kernel_ms(RtContext* ctx);
/* Instrumentation (not for timing): make the next launch count [0] box tests,
 * [1] triangle tests, [2] ray segments, [3] shaded hits, [4] rays whose REBUILT winner
 * failed certification and re-ran the literal policy, [5..8] path kernel: shading phases,
 * lanes served by them, traversal iterations, lanes active in them (warp-level); read them back. */
int rt_count_next_launch(RtContext* ctx);
int rt_read_counters(RtContext* ctx, uint64_t out[9]);
/* number of kernels this library has launched on ctx since creation */
uint64_t rt_launch_count(const RtContext* ctx);
/* which traversal the last launch used (RT_TRAVERSAL_LITERAL / _ORDERED / _REBUILT) */
int rt_last_traversal(const RtContext* ctx);
/* ---- multi-GPU tile split (no reference analogue: the reference is single-GPU;
 * it iterates independent 128x128 tiles the same way, src/compute.rs:157-162, 194-209).
 * Each context renders the tiles i % stride == phase (RtFrameParams); pack copies
 * them into a tile-major compact device buffer [local tile][RT_SPLIT_TILE][RT_SPLIT_TILE] for a
 * collective, unpack scatters such a buffer back into the full frame (rank 0).
 * what: 0 = composite rgba8, 1 = accumulation float4.  Asynchronous on rt_stream(). */
int rt_pack_tiles(RtContext* ctx, uint32_t what, uint32_t stride, uint32_t phase, void* dst_device);
int rt_unpack_tiles(RtContext* ctx, uint32_t what, uint32_t stride, uint32_t phase, const void* src_device);
/* Explicit tile assignment (multi-GPU load balancing on the previous frame's per-rank times): the split tiles this
 * context renders in fused frames, in the order it walks them; RtFrameParams.tile_stride / tile_phase then index
 * positions of this list (0, 0 = all of it) and rt_pack_tiles / rt_unpack_tiles follow it.  n = 0 clears it.
 * rt_pack_tile_list / rt_unpack_tile_list copy the tiles of ANY list (another rank's, on rank 0) between the frame
 * and a tile-major compact device buffer [k][RT_SPLIT_TILE][RT_SPLIT_TILE].  rt_resize clears the list. */
int rt_set_tile_list(RtContext* ctx, const uint32_t* tiles, uint32_t n);
int rt_pack_tile_list(RtContext* ctx, uint32_t what, const uint32_t* tiles, uint32_t n, void* dst_device);
int rt_unpack_tile_list(RtContext* ctx, uint32_t what, const uint32_t* tiles, uint32_t n, const void* src_device);
/* rt_pack_tile_list that allocates nothing.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same result as rt_pack_tile_list (tile k of `tiles` -> dst_device[k][RT_SPLIT_TILE][RT_SPLIT_TILE]), but
 * the list is staged in one file-scope static device array of RT_STATIC_TILE_LIST_MAX entries (64 KiB per GPU that
 * loads the library, held until exit) instead of a scratch buffer grown with cudaMalloc; n beyond that limit is refused
 * with RT_ERR_INVALID_ARGUMENT.  The call makes no cudaMalloc / cudaFree / allocating library call and does not wait
 * on the host: `tiles` must stay valid until the context's stream has passed the call when it lies in pinned memory.
 * Every call on a GPU shares that array and leaves its list in it (each call overwrites the entries it uses; nothing
 * needs clearing): calls that overlap on the GPU, from other contexts or streams of the same device, race. */
#define RT_STATIC_TILE_LIST_MAX 16384u
int rt_pack_tile_list_static(RtContext* ctx, uint32_t what, const uint32_t* tiles, uint32_t n, void* dst_device);
/* Fused multi-GPU epilogue (no reference analogue): instead of pack -> NCCL gather -> unpack, the kernels of
 * ranks != 0 store their finished composite pixels straight into rank 0's frame through peer memory (P2P
 * stores over NVLink; the frame is W*H*4 bytes in total).  Rank 0 exports its composite buffer as a 64-byte
 * CUDA IPC handle (valid until rt_resize / rt_destroy); every other rank opens it once and then renders
 * with RT_OUT_PEER.  A stream-ordered barrier (e.g. a 4-byte NCCL all-reduce on rt_stream()) after the frame
 * orders rank 0's read-back after everybody's stores.  rt_set_peer_frame takes a raw device pointer instead
 * (same-process contexts). */
int rt_export_frame_handle(RtContext* ctx, void* out_handle64);
int rt_open_peer_frame(RtContext* ctx, const void* handle64);
int rt_set_peer_frame(RtContext* ctx, void* device_composite);
int rt_close_peer_frame(RtContext* ctx);
/* the context's cudaStream_t, so callers can order their own work (NCCL) after it */
void* rt_stream(RtContext* ctx);
/* EXTENSION: ray segments traced by the last RT_MODE_PATH launch (SPEC.md) */
int rt_last_segments(RtContext* ctx, uint64_t* out);
/* EXTENSION: the reference's WavefrontCounters (shared/src/lib.rs:187-194; host loop src/compute.rs:413-479, which
 * only simulates the per-depth ray counts) for the last RT_MODE_PATH frame of the streaming wavefront engine:
 * rays_per_bounce[d] = ray segments traced at bounce depth d (depths >= 7 share entry 7), total, active-depth
 * bitmask, max depth and base seed.  Counts saturate at 2^32 - 1.  RT_ERR_NOT_READY after another engine. */
int rt_last_wavefront_counters(RtContext* ctx, RtWavefrontCounters* out);
/* GPU time of the last upload-time re-layout + LBVH build, ms (<0 if none) */
float rt_last_build_ms(const RtContext* ctx);
/* EXTENSION, streaming wavefront engine: GPU time per stage of the last RT_MODE_PATH frame,
 * summed over its launches (CUDA events between consecutive launches on the context
 * stream): out[0] generate, [1] extend (closest hit), [2] shade, [3] resolve, in ms;
 * out[4..7] the number of launches of each stage.  RT_ERR_NOT_READY if the last frame used
 * another engine. */
int rt_last_stage_ms(RtContext* ctx, float out[8]);
/* The same intervals one by one, in launch order: ms[i] and stage[i] (0 generate, 1 extend, 2 shade in place,
 * 3 resolve, 4 shade split form, 5 fallback launch; of 2 and 4 one returns at once, 5 normally finds an empty list);
 * returns their number (at most `cap` are written). */
int rt_last_launch_table(RtContext* ctx, float* ms, int32_t* stage, uint32_t cap);
/* Microbenchmark for the traversal roofline (no reference analogue): GB/s of random,
 * independent 32-byte sector loads (the node-fetch instruction) over a buffer of
 * footprint_bytes (rounded down to a power of two): small = L1 gather rate, between L1 and
 * L2 size = L2 gather rate, larger = HBM gather rate. */
int rt_probe_gather_bandwidth(RtContext* ctx, uint64_t footprint_bytes, float* out_gbs);

#ifdef __cplusplus
}
#endif
#endif
