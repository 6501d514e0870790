/* rt_oracle_math.h — TEST INFRASTRUCTURE ONLY (CPU oracle; never shipped or
 * imported by the product path).
 *
 * Scalar restatement of the vector arithmetic the reference shader gets from
 * glam (scalar-math) through `spirv_std::glam` (shader/src/lib.rs:13) and of
 * the f16 conversions it gets from `spirv_std::float` (shader/src/lighting.rs:1,
 * shader/src/material.rs:1).  glam / spirv-std are third-party crates that are
 * NOT under /root/reference (no lockfile; versions unpinned: `spirv-std 0.9`,
 * `glam 0.29` per shader/Cargo.toml).  Their published scalar semantics are
 * restated here:
 *   dot       = x*x' + y*y' + z*z'            (left to right)
 *   cross     = (y z' - z y', z x' - x z', x y' - y x')
 *   length    = sqrt(dot(v,v))
 *   normalize = v * (1.0 / length(v))         (multiply by reciprocal)
 *   min/max   = Rust f32::min/max (NaN-dropping) == C fminf/fmaxf
 * PARITY UNPINNED: the reference holds no test or golden vector for any of this
 * arithmetic.
 *
 * Build with -ffp-contract=off -fno-fast-math: Rust never contracts a*b+c.
 */
#ifndef RT_ORACLE_MATH_H
#define RT_ORACLE_MATH_H
#include <math.h>
#include <stdint.h>
#include <string.h>

struct V3 {
    float x, y, z;
};
static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 v3a(const float* p) { return V3{p[0], p[1], p[2]}; }
static inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
static inline V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
static inline V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
static inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) {
    return V3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
static inline float length(V3 a) { return sqrtf(dot(a, a)); }
static inline V3 normalize(V3 a) { return a * (1.0f / length(a)); }
static inline V3 vmin(V3 a, V3 b) { return V3{fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)}; }
static inline V3 vmax(V3 a, V3 b) { return V3{fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)}; }

static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

/* IEEE binary16 <-> binary32, round-to-nearest-even (GLSL PackHalf2x16 /
 * half 1.8.3 `f16::from_f32`).  Written out by hand so the oracle does not
 * depend on compiler _Float16 support. */
static inline uint16_t f32_to_f16_bits(float f) {
    uint32_t x = f2u(f);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t exp = (x >> 23) & 0xFFu;
    uint32_t man = x & 0x7FFFFFu;
    if (exp == 0xFFu) { /* inf / nan */
        if (man == 0) return (uint16_t)(sign | 0x7C00u);
        uint32_t m = man >> 13;
        return (uint16_t)(sign | 0x7C00u | m | (m == 0 ? 1u : 0u) | 0x200u);
    }
    int e = (int)exp - 127 + 15;
    if (e >= 0x1F) return (uint16_t)(sign | 0x7C00u); /* overflow -> inf */
    if (e <= 0) {
        if (e < -10) return (uint16_t)sign; /* underflow -> signed zero */
        man |= 0x800000u;
        uint32_t shift = (uint32_t)(14 - e);
        uint32_t half_man = man >> shift;
        uint32_t rem = man & ((1u << shift) - 1u);
        uint32_t halfway = 1u << (shift - 1);
        if (rem > halfway || (rem == halfway && (half_man & 1u))) half_man++;
        return (uint16_t)(sign | half_man);
    }
    uint32_t half = (uint32_t)(e << 10) | (man >> 13);
    uint32_t rem = man & 0x1FFFu;
    if (rem > 0x1000u || (rem == 0x1000u && (half & 1u))) half++;
    return (uint16_t)(sign | half);
}
static inline float f16_bits_to_f32(uint32_t h) {
    uint32_t sign = (h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1Fu;
    uint32_t man = h & 0x3FFu;
    if (exp == 0) {
        if (man == 0) return u2f(sign);
        /* subnormal: normalise */
        int e = -1;
        do { e++; man <<= 1; } while ((man & 0x400u) == 0);
        man &= 0x3FFu;
        return u2f(sign | ((uint32_t)(127 - 15 - e) << 23) | (man << 13));
    }
    if (exp == 0x1Fu) return u2f(sign | 0x7F800000u | (man << 13));
    return u2f(sign | ((exp - 15 + 127) << 23) | (man << 13));
}
#endif
