/* rt_oracle.h — TEST INFRASTRUCTURE ONLY: C interface of the CPU oracle
 * (see rt_oracle.cpp).  Loaded through ctypes by tests/, smoke() and the
 * cpu_baseline leg of bench.py; never by the product path. */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H
#include <stdint.h>
#include "../include/rt_shared.h"
#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_MISS = 0, ORC_SPHERE = 1, ORC_TRIANGLE = 2 };

/* The storage-buffer bindings 1-5 of main_cs (shader/src/lib.rs:28-33). */
typedef struct OrcBindings {
    const uint32_t* scene_metadata;
    uint64_t scene_metadata_len; /* u32 units */
    const RtTriangle* triangles[3];
    uint64_t triangles_len[3];
    const RtMaterial* materials;
    uint64_t materials_len;
} OrcBindings;

/* Closest-hit record the oracle adds (the reference returns no primitive id). */
typedef struct OrcHit {
    uint32_t kind;        /* ORC_* */
    uint32_t prim;        /* sphere index / triangle index; 0xFFFFFFFF on miss */
    uint32_t t_bits;      /* IEEE bits of t (f32::MAX on miss) */
    uint32_t material_id; /* 0xFFFFFFFF on miss */
} OrcHit;

typedef struct OrcCounters {
    uint64_t aabb_tests, tri_tests, sphere_tests, rays;
} OrcCounters;

int orc_hardware_threads(void);
int orc_dispatch(const OrcBindings* ob, const RtPushConstants* pc, uint8_t* image, float* radiance,
                 OrcHit* hits, int threads, OrcCounters* counters);
int orc_dispatch_policy(const OrcBindings* ob, const RtPushConstants* pc, uint8_t* image, float* radiance,
                        OrcHit* hits, int threads, OrcCounters* counters, int policy);
int orc_render_path(const OrcBindings* ob, const RtPushConstants* pc, uint32_t spp, uint32_t first_sample,
                    uint32_t max_depth, uint32_t base_seed, int reset, uint32_t y0, uint32_t y1, float* accum,
                    OrcHit* hits, int threads, OrcCounters* counters, int policy);
void orc_last_depth_rays(uint64_t out[8]);
void orc_resolve(const float* accum, uint8_t* out, uint64_t pixels);
void orc_composite(const uint8_t* r, const uint8_t* g, const uint8_t* b, uint8_t* out, uint64_t pixels);
uint16_t orc_f32_to_f16(float f);
float orc_f16_to_f32(uint16_t h);

#ifdef __cplusplus
}
#endif
#endif
