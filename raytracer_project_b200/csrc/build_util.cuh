// build_util.cuh — helpers shared by the device BVH builders (lbvh.cu, sah_build.cu)
#pragma once
#include <cstdint>

#include "vecmath.cuh"

namespace {

// order-preserving float <-> uint map so that atomicMin / atomicMax work on floats of either sign
ZB_DEV uint32_t f2ord(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
ZB_DEV float ord2f(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

struct TriBox { float3 lo, hi; };
// triangle::bounding_box (triangle.hpp:84-101): axes thinner than 1e-4 are padded by 1e-4 on both sides
ZB_DEV TriBox tri_box(const float4* t) {
    float3 a = f3(__ldg(t)), b = f3(__ldg(t + 1)), c = f3(__ldg(t + 2));
    TriBox r;
    r.lo = f3(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)));
    r.hi = f3(fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)));
    if (r.hi.x - r.lo.x < 1e-4f) { r.lo.x -= 1e-4f; r.hi.x += 1e-4f; }
    if (r.hi.y - r.lo.y < 1e-4f) { r.lo.y -= 1e-4f; r.hi.y += 1e-4f; }
    if (r.hi.z - r.lo.z < 1e-4f) { r.lo.z -= 1e-4f; r.hi.z += 1e-4f; }
    return r;
}

}  // namespace
