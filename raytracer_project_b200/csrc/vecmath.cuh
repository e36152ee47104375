// vecmath.cuh — fp32 vector helpers and the counter-based Philox4x32-10 generator for the device path.
// The reference computes in double with one shared mt19937 (vec3.hpp:7-213, common.hpp:29-34); the
// device path is fp32 with selective fp64 (see sphere test in trace.cuh) and a stateless RNG keyed by
// (pixel, sample, bounce) so that any partition of the frame across GPUs draws the same numbers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define ZB_DEV __device__ __forceinline__

ZB_DEV float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
ZB_DEV float3 f3(float4 v) { return make_float3(v.x, v.y, v.z); }
ZB_DEV float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
ZB_DEV float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
ZB_DEV float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
ZB_DEV float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
ZB_DEV float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
ZB_DEV float3 operator*(float3 a, float s) { return f3(s * a.x, s * a.y, s * a.z); }
ZB_DEV float3& operator+=(float3& a, float3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
ZB_DEV float3& operator*=(float3& a, float3 b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; return a; }
ZB_DEV float3& operator*=(float3& a, float s) { a.x *= s; a.y *= s; a.z *= s; return a; }
ZB_DEV float dot(float3 a, float3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
ZB_DEV float3 cross(float3 a, float3 b) {
    return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
ZB_DEV float3 fma3(float s, float3 a, float3 b) { return f3(fmaf(s, a.x, b.x), fmaf(s, a.y, b.y), fmaf(s, a.z, b.z)); }
ZB_DEV float length(float3 a) { return sqrtf(dot(a, a)); }
// unit_vector (vec3.hpp:165-171): zero vector when the length is below 1e-8
ZB_DEV float3 unit_vector(float3 v) {
    float l2 = dot(v, v);
    return l2 < 1e-16f ? f3(0.f, 0.f, 0.f) : rsqrtf(l2) * v;
}
ZB_DEV float max3(float a, float b, float c) { return fmaxf(a, fmaxf(b, c)); }
ZB_DEV float min3(float a, float b, float c) { return fminf(a, fminf(b, c)); }

// sin and cos of 2 pi x for x in [0, 1): two MUFU evaluations on the angle moved to [-pi, pi), where their absolute error is
// ~2^-21 (5 instructions instead of the ~28 of sincospif; the results only steer sampled directions)
ZB_DEV void sincos_turn(float x, float* s, float* c) {
    const float a = 6.283185307179586f * (x - 0.5f);
    *s = -__sinf(a);
    *c = -__cosf(a);
}

// MUFU square root (~1 ulp), for values that only steer sampled directions
ZB_DEV float sqrt_fast(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// row-major 3x4 affine held in three float4 rows
struct Affine {
    float4 r0, r1, r2;
};
ZB_DEV float3 xform_point(const Affine& m, float3 p) {
    return f3(fmaf(m.r0.x, p.x, fmaf(m.r0.y, p.y, fmaf(m.r0.z, p.z, m.r0.w))),
              fmaf(m.r1.x, p.x, fmaf(m.r1.y, p.y, fmaf(m.r1.z, p.z, m.r1.w))),
              fmaf(m.r2.x, p.x, fmaf(m.r2.y, p.y, fmaf(m.r2.z, p.z, m.r2.w))));
}
ZB_DEV float3 xform_vector(const Affine& m, float3 v) {
    return f3(fmaf(m.r0.x, v.x, fmaf(m.r0.y, v.y, m.r0.z * v.z)), fmaf(m.r1.x, v.x, fmaf(m.r1.y, v.y, m.r1.z * v.z)),
              fmaf(m.r2.x, v.x, fmaf(m.r2.y, v.y, m.r2.z * v.z)));
}
// transpose of the linear part applied to v  (normal transform with m = world->object)
ZB_DEV float3 xform_vector_t(const Affine& m, float3 v) {
    return f3(fmaf(m.r0.x, v.x, fmaf(m.r1.x, v.y, m.r2.x * v.z)), fmaf(m.r0.y, v.x, fmaf(m.r1.y, v.y, m.r2.y * v.z)),
              fmaf(m.r0.z, v.x, fmaf(m.r1.z, v.y, m.r2.z * v.z)));
}

// ---------------------------------------------------------------- Philox4x32-10 (Salmon et al. 2011)
#ifndef ZB_PHILOX_ROUNDS
#define ZB_PHILOX_ROUNDS 10
#endif
// The key schedule (key + r * Weyl constants) is expanded on the host into 2 * rounds words that live in the kernel
// parameters, so each round is two wide multiplies and two three-input XORs with a constant-bank operand.
#define ZB_PHILOX_WORDS (2 * ZB_PHILOX_ROUNDS)
struct Philox {
    const uint32_t* rk;  // round keys: rk[2 r] = key0 + r * 0x9E3779B9, rk[2 r + 1] = key1 + r * 0xBB67AE85
    ZB_DEV uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
#pragma unroll
        for (int r = 0; r < ZB_PHILOX_ROUNDS; r++) {
            uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            c0 = hi1 ^ c1 ^ rk[2 * r];
            c1 = lo1;
            c2 = hi0 ^ c3 ^ rk[2 * r + 1];
            c3 = lo0;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};
inline void philox_expand_key(uint32_t key0, uint32_t key1, uint32_t* rk) {  // host side
    for (int r = 0; r < ZB_PHILOX_ROUNDS; r++) {
        rk[2 * r] = key0 + uint32_t(r) * 0x9E3779B9u;
        rk[2 * r + 1] = key1 + uint32_t(r) * 0xBB67AE85u;
    }
}
// uniform in [0,1) with 24 bits, the analogue of random_double() (common.hpp:29-34)
ZB_DEV float u01(uint32_t x) { return float(x >> 8) * (1.0f / 16777216.0f); }
// uniform in (0,1): safe argument for log()
ZB_DEV float u01_open(uint32_t x) { return (float(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }
