// rt_math.cuh — scalar vector arithmetic with the exact operation order of the
// reference's glam (scalar-math) calls.  This translation unit is compiled with
// -fmad=false and without --use_fast_math so that every a*b+c stays two IEEE
// roundings (Rust never contracts), and `/`, sqrtf are IEEE-correct.
// fminf/fmaxf have Rust f32::min/max semantics (the non-NaN operand wins), which
// the lighting code relies on (shader/src/lighting.rs:104,128,131).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 mka(const float* p) { return mk(p[0], p[1], p[2]); }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross3(V3 a, V3 b) {
    return mk(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
__device__ __forceinline__ float length3(V3 a) { return sqrtf(dot3(a, a)); }
__device__ __forceinline__ V3 normalize3(V3 a) { return a * (1.0f / length3(a)); }

__device__ __forceinline__ float f16lo(uint32_t packed) {
    return __half2float(__ushort_as_half((unsigned short)(packed & 0xFFFFu)));
}
__device__ __forceinline__ float f16hi(uint32_t packed) {
    return __half2float(__ushort_as_half((unsigned short)(packed >> 16)));
}
// f32 -> f16 -> f32 round trip, round-to-nearest-even (shader/src/lighting.rs:125-127)
__device__ __forceinline__ float round_through_f16(float f) { return __half2float(__float2half_rn(f)); }

// Rgba8Unorm store conversion: clamp [0,1], *255, round to nearest even; NaN -> 0
__device__ __forceinline__ uint32_t unorm8(float f) {
    if (!(f > 0.0f)) return 0u;
    if (f >= 1.0f) return 255u;
    return (uint32_t)rintf(f * 255.0f);
}
