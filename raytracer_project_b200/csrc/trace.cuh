// trace.cuh — K2: closest-hit query over the flat two-level BVH (replaces the recursive virtual
// world.hit() of bvh.hpp:46-118 / hittable_list.hpp:28-41 and every primitive's hit()).
//
// One thread = one ray.  The TLAS leaf holds 32-byte items (inline spheres, or references to cubes,
// loose triangles, instances and media); an instance over a mesh switches the same loop into that
// mesh's BLAS with the ray moved to object space (directions are never renormalised, so t is
// invariant — translate.hpp:16-19, scale.hpp:21-24) and a sentinel on the stack brings it back.
// Node = 64 B holding both child boxes: four 16-byte non-coherent loads, then two slab tests; the
// nearer child is visited first and the other pushed on a short per-thread stack.
#pragma once
#include "../../include/zenith_b200.h"
#include "scene_dev.cuh"

#define ZB_STACK 64
// Stack budget of one ray: the sentinel, at most one deferred child per TLAS level, three entries for a dive into a
// mesh (rest of the leaf, exit marker, BLAS root) and one deferred child per BLAS level:
//   1 + ZB_MAX_TLAS_DEPTH + 3 + ZB_MAX_BLAS_DEPTH <= ZB_STACK   (include/zenith_b200.h; checked by zb_scene_upload)
static_assert(1 + ZB_MAX_TLAS_DEPTH + 3 + ZB_MAX_BLAS_DEPTH <= ZB_STACK, "traversal stack too small for the depth caps");
#ifndef ZB_BOX_CH
#define ZB_BOX_CH 1
#endif
// Slab tests use MUFU.RCP reciprocals (~1 ulp) of the ray direction; the far distance is inflated by
// 10 ulp so that the box test stays conservative (Ize 2013).  Measured -2 % frame time vs IEEE divides.
#define ZB_BOX_INFLATE 1.0000012f
#ifndef ZB_MEDIUM_ONE_SOLVE
#define ZB_MEDIUM_ONE_SOLVE 1
#endif
#ifndef ZB_BRANCHLESS_POP
#define ZB_BRANCHLESS_POP 0  // measured: Weekend +7 % time, mesh -1.3 % (profiles/r01_feature_sets.txt)
#endif
#ifndef ZB_PREFETCH_FAR
#define ZB_PREFETCH_FAR 0  // mesh feature sets: prefetch the deferred child's node / first triangle when it is pushed.
                           // Measured slower: C3 67.4 vs 66.0 ms, C5 at 16 spp 190.1 vs 184.0 ms (profiles/r02_traversal_experiments.txt)
#endif
#ifndef ZB_PREFETCH_POP
#define ZB_PREFETCH_POP 0  // BLAS leaves: prefetch the node the stack hands out next, before the triangle tests.
                           // Measured slower: C3 66.6 vs 66.1 ms, C5 at 16 spp 186.1 vs 184.2 ms (profiles/r02_traversal_experiments.txt)
#endif
#define ZB_EXIT_INSTANCE ((int)0x80000000)

struct Hit {
    float t;
    uint32_t item;  // index into items[] (TLAS leaf order); 0xFFFFFFFF = miss
    uint32_t prim;  // absolute triangle index for mesh / triangle hits
};

struct Counters {
    uint32_t rays, nodes, prims, insts;
};

// Per-ray source of the free-flight draw of constant_medium::hit (constant_medium.hpp:64): one draw
// per (ray, medium), so the result does not depend on traversal order.
struct MediumRng {
    Philox ph;
    uint32_t pixel, sample, bounce;
    ZB_DEV float draw(uint32_t medium) const {
        uint4 r = ph(pixel, sample, bounce, 0x10000u + (medium >> 2));
        uint32_t v = (medium & 3) == 0 ? r.x : (medium & 3) == 1 ? r.y : (medium & 3) == 2 ? r.z : r.w;
        return u01_open(v);
    }
};

// ---------------------------------------------------------------------------------- primitives

// sphere::hit (sphere.hpp:18-39): roots of a t^2 - 2 h t + c = 0 accepted only strictly inside
// (tmin, tmax), nearer root first.  fp32 with two safeguards for the r = 1000 ground sphere and for
// rays that start 1e-4 above a surface: the discriminant uses the perpendicular-distance form
// a (r^2 - |oc - (h/a) d|^2), and c = |oc|^2 - r^2 is evaluated in fp64 (B200 runs fp64 at half the
// fp32 rate) and the roots come from q = h + sgn(h) sqrt(disc): t = q/a and t = c/q.
ZB_DEV bool hit_sphere(float4 s, float3 o, float3 d, float tmin, float tmax, float& t_out) {
    float3 oc = f3(s.x - o.x, s.y - o.y, s.z - o.z);
    float a = dot(d, d);
    float h = dot(d, oc);
    float inv_a = __frcp_rn(a);
    float3 l = fma3(-h * inv_a, d, oc);
    float disc = fmaf(s.w, s.w, -dot(l, l));
    if (disc < 0.f) return false;
    double ocx = double(s.x) - double(o.x), ocy = double(s.y) - double(o.y), ocz = double(s.z) - double(o.z);
    float c = float(fma(ocx, ocx, fma(ocy, ocy, fma(ocz, ocz, -double(s.w) * double(s.w)))));
    float sq = sqrtf(a * disc);
    float q = h + copysignf(sq, h);
    float ta = q * inv_a;
    float tb = q != 0.f ? c / q : ta;
    float t0 = fminf(ta, tb), t1 = fmaxf(ta, tb);
    if (t0 > tmin && t0 < tmax) { t_out = t0; return true; }
    if (t1 > tmin && t1 < tmax) { t_out = t1; return true; }
    return false;
}

// cube::hit (cube.hpp:44-70): slabs +-half about the origin of the current frame, t = entry distance
// or tmin when the origin is inside; accepted whenever tmax >= tmin.
ZB_DEV bool hit_cube(float3 half, float3 o, float3 d, float tmin, float tmax, float& t_out) {
    float lo = tmin, hi = tmax;
    {
        float inv = 1.0f / d.x;
        float t0 = (-half.x - o.x) * inv, t1 = (half.x - o.x) * inv;
        if (inv < 0.f) { float s = t0; t0 = t1; t1 = s; }
        lo = fmaxf(t0, lo); hi = fminf(t1, hi);
        if (hi < lo) return false;
    }
    {
        float inv = 1.0f / d.y;
        float t0 = (-half.y - o.y) * inv, t1 = (half.y - o.y) * inv;
        if (inv < 0.f) { float s = t0; t0 = t1; t1 = s; }
        lo = fmaxf(t0, lo); hi = fminf(t1, hi);
        if (hi < lo) return false;
    }
    {
        float inv = 1.0f / d.z;
        float t0 = (-half.z - o.z) * inv, t1 = (half.z - o.z) * inv;
        if (inv < 0.f) { float s = t0; t0 = t1; t1 = s; }
        lo = fmaxf(t0, lo); hi = fminf(t1, hi);
        if (hi < lo) return false;
    }
    t_out = lo;
    return true;
}

// triangle::hit (triangle.hpp:17-66) restated as Moeller-Trumbore with the reference's acceptance
// rules: degenerate (|N| < 1e-8) and parallel (|N.d|/|N| < 1e-8) triangles rejected, t accepted
// INCLUSIVELY in [tmin, tmax], edges inclusive (a point on a shared edge hits both triangles).
ZB_DEV bool hit_tri(float3 v0, float3 v1, float3 v2, float3 o, float3 d, float tmin, float tmax, float& t_out) {
    float3 e1 = v1 - v0, e2 = v2 - v0;
    float3 n = cross(e1, e2);
    float n2 = dot(n, n);
    if (n2 < 1e-16f) return false;
    float nd = dot(n, d);
    if (nd * nd < 1e-16f * n2) return false;
    float inv = 1.0f / nd;
    float3 tv = v0 - o;
    float t = dot(n, tv) * inv;
    if (!(t >= tmin && t <= tmax)) return false;
    float3 q = cross(tv, d);
    float u = dot(q, e2) * inv;   // weight of v1
    float v = -dot(q, e1) * inv;  // weight of v2
    if (u < 0.f || v < 0.f || u + v > 1.0f) return false;
    t_out = t;
    return true;
}

struct InstanceHead {
    Affine w2o;
    uint32_t geom_kind, geom_index, flags, node_root, wide_root;
};
ZB_DEV InstanceHead load_instance_head(const SceneDev& sc, uint32_t idx) {
    const float4* p = sc.instances + 8 * size_t(idx);
    InstanceHead h;
    h.w2o.r0 = __ldg(p + 0);
    h.w2o.r1 = __ldg(p + 1);
    h.w2o.r2 = __ldg(p + 2);
    float4 m = __ldg(p + 6);
    float4 m2 = __ldg(p + 7);
    h.geom_kind = __float_as_uint(m.x);
    h.geom_index = __float_as_uint(m.y);
    h.flags = __float_as_uint(m.z);
    h.node_root = __float_as_uint(m.w);
    h.wide_root = __float_as_uint(m2.x);
    return h;
}

// boundary->hit() of a constant_medium for sphere / cube boundaries (optionally instanced)
template <uint32_t F = ZB_F_ALL>
ZB_DEV bool hit_simple_geom(const SceneDev& sc, uint32_t geom_kind, uint32_t geom_index, float3 o, float3 d,
                            float tmin, float tmax, float& t) {
    if (geom_kind == ZB_GEOM_SPHERE) return hit_sphere(__ldg(sc.spheres_obj + geom_index), o, d, tmin, tmax, t);
    if ((F & ZB_F_CUBE) && geom_kind == ZB_GEOM_CUBE) return hit_cube(f3(__ldg(sc.cubes + 2 * size_t(geom_index))), o, d, tmin, tmax, t);
    if ((F & ZB_F_TRI) && geom_kind == ZB_GEOM_TRIANGLE) {
        const float4* tp = sc.tris + 3 * size_t(geom_index);
        return hit_tri(f3(__ldg(tp)), f3(__ldg(tp + 1)), f3(__ldg(tp + 2)), o, d, tmin, tmax, t);
    }
    return false;
}

template <uint32_t F = ZB_F_ALL>
ZB_DEV bool hit_medium_boundary(const SceneDev& sc, float4 m0, float4 sph, float3 o, float3 d, float tmin, float tmax,
                                float& t) {
    uint32_t kind = __float_as_uint(m0.x), index = __float_as_uint(m0.y);
    if (kind == ZB_ITEM_SPHERE) return hit_sphere(sph, o, d, tmin, tmax, t);
    if ((F & ZB_F_CUBE) && kind == ZB_ITEM_CUBE) return hit_cube(f3(__ldg(sc.cubes + 2 * size_t(index))), o, d, tmin, tmax, t);
    if (!(F & ZB_F_INST)) return false;
    InstanceHead ih = load_instance_head(sc, index);
    return hit_simple_geom<F>(sc, ih.geom_kind, ih.geom_index, xform_point(ih.w2o, o), xform_vector(ih.w2o, d), tmin,
                           tmax, t);
}

// both roots of hit_sphere's quadratic in one evaluation (same arithmetic, so lo / hi are the values hit_sphere returns)
ZB_DEV bool sphere_roots(float4 s, float3 o, float3 d, float& lo, float& hi) {
    float3 oc = f3(s.x - o.x, s.y - o.y, s.z - o.z);
    float a = dot(d, d);
    float h = dot(d, oc);
    float inv_a = __frcp_rn(a);
    float3 l = fma3(-h * inv_a, d, oc);
    float disc = fmaf(s.w, s.w, -dot(l, l));
    if (disc < 0.f) return false;
    double ocx = double(s.x) - double(o.x), ocy = double(s.y) - double(o.y), ocz = double(s.z) - double(o.z);
    float c = float(fma(ocx, ocx, fma(ocy, ocy, fma(ocz, ocz, -double(s.w) * double(s.w)))));
    float sq = sqrtf(a * disc);
    float q = h + copysignf(sq, h);
    float ta = q * inv_a;
    float tb = q != 0.f ? c / q : ta;
    lo = fminf(ta, tb);
    hi = fmaxf(ta, tb);
    return true;
}

// constant_medium::hit (constant_medium.hpp:39-77)
template <uint32_t F = ZB_F_ALL>
ZB_DEV bool hit_medium(const SceneDev& sc, uint32_t medium, float3 o, float3 d, float tmin, float tmax,
                       const MediumRng& rng, float& t_out) {
    float4 m0 = __ldg(sc.media + 2 * size_t(medium));
    float4 sph = __ldg(sc.media + 2 * size_t(medium) + 1);
    float t1, t2;
    if (ZB_MEDIUM_ONE_SOLVE && __float_as_uint(m0.x) == ZB_ITEM_SPHERE) {
        // inline sphere boundary: boundary->hit(r, universe) is the nearer root, boundary->hit(r, [t1 + 1e-4, inf)) the
        // farther one if it lies beyond t1 + 1e-4 — one solve of the quadratic instead of two
        if (!sphere_roots(sph, o, d, t1, t2)) return false;
        if (!(t1 > -INFINITY && t1 < INFINITY)) {  // hit_sphere's strict interval test on the nearer root, then the farther
            if (!(t2 > -INFINITY && t2 < INFINITY)) return false;
            t1 = t2;
        }
        if (!(t2 > t1 + 0.0001f && t2 < INFINITY)) return false;
    } else {
        if (!hit_medium_boundary<F>(sc, m0, sph, o, d, -INFINITY, INFINITY, t1)) return false;
        if (!hit_medium_boundary<F>(sc, m0, sph, o, d, t1 + 0.0001f, INFINITY, t2)) return false;
    }
    if (t1 < tmin) t1 = tmin;
    if (t2 > tmax) t2 = tmax;
    if (t1 >= t2) return false;
    if (t1 < 0.f) t1 = 0.f;
    float len = length(d);
    float inside = (t2 - t1) * len;
    float dist = m0.z * __logf(rng.draw(medium));  // neg_inv_density * log(xi)
    if (dist > inside) return false;
    t_out = t1 + dist / len;
    return true;
}

// ---------------------------------------------------------------------------------- traversal

struct RayBox {
    float3 inv, ood;
    ZB_DEV void set(float3 o, float3 d) {
        float dx = fabsf(d.x) > 1e-20f ? d.x : copysignf(1e-20f, d.x);
        float dy = fabsf(d.y) > 1e-20f ? d.y : copysignf(1e-20f, d.y);
        float dz = fabsf(d.z) > 1e-20f ? d.z : copysignf(1e-20f, d.z);
        inv = f3(__fdividef(1.0f, dx), __fdividef(1.0f, dy), __fdividef(1.0f, dz));
        ood = f3(o.x * inv.x, o.y * inv.y, o.z * inv.z);
    }
};

// Stack codes: >= 0 inner node, < 0 leaf (~code), and two reserved leaf codes that no real leaf can
// take (they would need first >= 0x0FFFFFFF): ZB_EXIT_INSTANCE pops the ray back to world space,
// ZB_TRAVERSAL_DONE is the sentinel at the bottom of the stack.
#define ZB_TRAVERSAL_DONE ((int)0x80000001)
// Overflow detection without a branch in the node loop: the last stack slot holds a canary that only an overflowing
// push (or the dive guard) overwrites; trace() looks at it once per ray and counts into SceneDev::overflow.
#define ZB_STACK_CANARY ((int)0x80000002)
#define ZB_STACK_POISON ((int)0x80000003)

// primitive tests of TLAS leaf `code` (items first .. first + n).  Returns true when the ray has to dive
// into a mesh BLAS: the rest of the leaf, the exit marker and the BLAS root are then on the stack and
// (ro, rd) is the ray in the object space of instance item `blas_item`.
// F without any ZB_F_GEOM bit: every item is an inline sphere, only the geometry half of the item is read.
template <bool STATS, uint32_t F = ZB_F_ALL>
ZB_DEV bool tlas_leaf(const SceneDev& sc, uint32_t code, float3 o, float3 d, float tmin, const MediumRng& rng, Hit& hit,
                      int* stack, int& sp, float3& ro, float3& rd, uint32_t& blas_item, Counters& cnt) {
    const uint32_t first = code >> 3, end = first + (code & 7u) + 1u;
    for (uint32_t i = first; i < end; i++) {
        float4 g = __ldg(sc.items + 2 * size_t(i));
        if (!(F & ZB_F_GEOM)) {
            if (STATS) cnt.prims++;
            float ts;
            if (hit_sphere(g, o, d, tmin, hit.t, ts)) { hit.t = ts; hit.item = i; }
            continue;
        }
        float4 mf = __ldg(sc.items + 2 * size_t(i) + 1);
        uint32_t kind = __float_as_uint(mf.x) & 0xFFu, aux = __float_as_uint(mf.w);
        if (STATS) cnt.prims++;
        float t;
        if (kind == ZB_ITEM_SPHERE) {
            if (hit_sphere(g, o, d, tmin, hit.t, t)) { hit.t = t; hit.item = i; }
        } else if ((F & ZB_F_INST) && kind == ZB_ITEM_INSTANCE) {
            InstanceHead ih = load_instance_head(sc, aux);
            if (STATS) cnt.insts++;
            float3 lo = xform_point(ih.w2o, o), ld = xform_vector(ih.w2o, d);
            if ((F & ZB_F_MESH) && ih.geom_kind == ZB_GEOM_MESH) {
                // defer the rest of this leaf, then dive into the BLAS in object space
                if (sp + 3 > ZB_STACK) {  // cannot happen for a scene zb_scene_upload accepted; counted, never silent
                    stack[ZB_STACK - 1] = ZB_STACK_POISON;
                    return false;
                }
                if (i + 1 < end) stack[sp++] = int(~(((i + 1) << 3) | (end - i - 2)));
                stack[sp++] = ZB_EXIT_INSTANCE;
                stack[sp++] = int(ih.node_root);
                blas_item = i;
                ro = lo;
                rd = ld;
                return true;
            }
            if (hit_simple_geom<F>(sc, ih.geom_kind, ih.geom_index, lo, ld, tmin, hit.t, t)) {
                hit.t = t;
                hit.item = i;
                hit.prim = ih.geom_index;
            }
        } else if ((F & ZB_F_CUBE) && kind == ZB_ITEM_CUBE) {
            if (hit_cube(f3(__ldg(sc.cubes + 2 * size_t(aux))), o, d, tmin, hit.t, t)) { hit.t = t; hit.item = i; }
        } else if ((F & ZB_F_TRI) && kind == ZB_ITEM_TRIANGLE) {
            const float4* tp = sc.tris + 3 * size_t(aux);
            if (hit_tri(f3(__ldg(tp)), f3(__ldg(tp + 1)), f3(__ldg(tp + 2)), o, d, tmin, hit.t, t)) {
                hit.t = t;
                hit.item = i;
                hit.prim = aux;
            }
        } else if ((F & ZB_F_MEDIA) && kind == ZB_ITEM_MEDIUM) {
            if (hit_medium<F>(sc, aux, o, d, tmin, hit.t, rng, t)) { hit.t = t; hit.item = i; }
        }
    }
    return false;
}

// Traversal state of one ray.  Only the stack (owned by the caller, local memory) is addressed dynamically; every
// other member is a scalar the compiler keeps in registers as long as the object itself never has its address taken
// (all users are force-inlined).  node_step() = one visit of an inner node (both child boxes, near child first, far
// child pushed); leaf_step() = the primitive tests of the leaf / marker in `cur`, then the next entry is popped.
// The four-way child selection is written as selects plus one predicated push / pop: no divergent branch.
struct RayWalk {
    int cur, sp;
    bool in_blas;
    uint32_t blas_item;
    float3 ro, rd;  // ray in the current frame (world, or the object space of the instance being traversed)
    RayBox rb;
    Hit hit;

    ZB_DEV void begin(const SceneDev& sc, float3 o, float3 d, float tmax, int* stack) {
        hit.t = tmax;
        hit.item = 0xFFFFFFFFu;
        hit.prim = 0;
        stack[0] = ZB_TRAVERSAL_DONE;
        stack[ZB_STACK - 1] = ZB_STACK_CANARY;
        sp = 1;
        cur = sc.n_nodes ? 0 : ZB_TRAVERSAL_DONE;
        in_blas = false;
        blas_item = 0;
        ro = o;
        rd = d;
        rb.set(o, d);
    }

    template <bool STATS, bool PREFETCH = false>
    ZB_DEV void node_step(const SceneDev& sc, float tmin, int* stack, Counters& cnt) {
        const float4* np = sc.nodes + 4 * size_t(cur);
        const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2);
        const float2 nc = __ldg(reinterpret_cast<const float2*>(np + 3));
        if (STATS) cnt.nodes++;
#if ZB_BOX_CH
        // nodes re-encoded at upload as centre / half extent (zbk_convert_nodes): t = (c - o) / d -+ h / |d| needs
        // no per-axis min / max, which moves 12 of the 20 FMNMX of a visit from the half-rate ALU pipe (the
        // busiest pipe of this kernel, ncu: 66 %) to 6 more FFMA on the FMA pipe
        const float mlx = fmaf(n0.x, rb.inv.x, -rb.ood.x), mly = fmaf(n0.z, rb.inv.y, -rb.ood.y), mlz = fmaf(n2.x, rb.inv.z, -rb.ood.z);
        const float mrx = fmaf(n1.x, rb.inv.x, -rb.ood.x), mry = fmaf(n1.z, rb.inv.y, -rb.ood.y), mrz = fmaf(n2.z, rb.inv.z, -rb.ood.z);
        const float ax = fabsf(rb.inv.x), ay = fabsf(rb.inv.y), az = fabsf(rb.inv.z);
        float ln = fmaxf(max3(fmaf(-n0.y, ax, mlx), fmaf(-n0.w, ay, mly), fmaf(-n2.y, az, mlz)), tmin);
        float lf = fminf(min3(fmaf(n0.y, ax, mlx), fmaf(n0.w, ay, mly), fmaf(n2.y, az, mlz)), hit.t) * ZB_BOX_INFLATE;
        float rn = fmaxf(max3(fmaf(-n1.y, ax, mrx), fmaf(-n1.w, ay, mry), fmaf(-n2.w, az, mrz)), tmin);
        float rf = fminf(min3(fmaf(n1.y, ax, mrx), fmaf(n1.w, ay, mry), fmaf(n2.w, az, mrz)), hit.t) * ZB_BOX_INFLATE;
#else
        float lx0 = fmaf(n0.x, rb.inv.x, -rb.ood.x), lx1 = fmaf(n0.y, rb.inv.x, -rb.ood.x);
        float ly0 = fmaf(n0.z, rb.inv.y, -rb.ood.y), ly1 = fmaf(n0.w, rb.inv.y, -rb.ood.y);
        float lz0 = fmaf(n2.x, rb.inv.z, -rb.ood.z), lz1 = fmaf(n2.y, rb.inv.z, -rb.ood.z);
        float rx0 = fmaf(n1.x, rb.inv.x, -rb.ood.x), rx1 = fmaf(n1.y, rb.inv.x, -rb.ood.x);
        float ry0 = fmaf(n1.z, rb.inv.y, -rb.ood.y), ry1 = fmaf(n1.w, rb.inv.y, -rb.ood.y);
        float rz0 = fmaf(n2.z, rb.inv.z, -rb.ood.z), rz1 = fmaf(n2.w, rb.inv.z, -rb.ood.z);
        float ln = fmaxf(max3(fminf(lx0, lx1), fminf(ly0, ly1), fminf(lz0, lz1)), tmin);
        float lf = fminf(min3(fmaxf(lx0, lx1), fmaxf(ly0, ly1), fmaxf(lz0, lz1)), hit.t) * ZB_BOX_INFLATE;
        float rn = fmaxf(max3(fminf(rx0, rx1), fminf(ry0, ry1), fminf(rz0, rz1)), tmin);
        float rf = fminf(min3(fmaxf(rx0, rx1), fmaxf(ry0, ry1), fmaxf(rz0, rz1)), hit.t) * ZB_BOX_INFLATE;
#endif
        const bool hl = ln <= lf, hr = rn <= rf;
        const int cl = __float_as_int(nc.x), cr = __float_as_int(nc.y);
        const bool right_first = hr && (!hl || rn < ln);
        int next = right_first ? cr : cl;
#if ZB_BRANCHLESS_POP
        // push and pop without a divergent branch (and its convergence barrier): a predicated store of the far child, an
        // unconditional load of the slot below the top (in bounds: the sentinel keeps sp >= 1), selects for sp and cur
        const bool any = hl || hr, both = hl && hr;
        if (both) stack[sp] = right_first ? cl : cr;
        const int top = stack[sp - 1];
        sp = both ? min(sp + 1, ZB_STACK - 1) : any ? sp : sp - 1;
        cur = any ? next : top;
#else
        if (hl && hr) {
            // branch-free push (a cold overflow branch here cost 44 % of the headline frame, 38.2 vs 26.5 ms: it breaks the
            // predicated form of this loop).  A push at the last slot overwrites the canary begin() put there; trace() checks it.
            stack[sp] = right_first ? cl : cr;
            sp = min(sp + 1, ZB_STACK - 1);
            if (PREFETCH) {
                // mesh scenes are bound by the latency of dependent node fetches (ncu: long scoreboard 5 stall cycles per issue,
                // L1 hit 60 %): the deferred child is fetched again only when the stack pops it — ask for its line now
                const int far = right_first ? cl : cr;
                const void* p = far >= 0 ? static_cast<const void*>(sc.nodes + 4 * size_t(far))
                                         : static_cast<const void*>(sc.tris + 3 * size_t((~uint32_t(far)) >> 3));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
            }
        }
        if (!(hl || hr)) next = stack[--sp];
        cur = next;
#endif
    }

    // (o, d) is the world-space ray given to begin()
    template <bool STATS, uint32_t F = ZB_F_ALL>
    ZB_DEV void leaf_step(const SceneDev& sc, float3 o, float3 d, float tmin, const MediumRng& rng, int* stack, Counters& cnt) {
        if (!(F & ZB_F_MESH)) {
            tlas_leaf<STATS, F>(sc, ~uint32_t(cur), o, d, tmin, rng, hit, stack, sp, ro, rd, blas_item, cnt);
        } else if (cur == ZB_EXIT_INSTANCE) {
            in_blas = false;
            ro = o;
            rd = d;
            rb.set(o, d);
        } else if (in_blas) {
            // triangles of a mesh BLAS leaf, in object space
            const uint32_t code = ~uint32_t(cur);
            const uint32_t first = code >> 3, end = first + (code & 7u) + 1u;
#if ZB_PREFETCH_POP
            {   // the entry the stack will hand out after this leaf is certain to be fetched: ask for its line before the
                // triangle tests instead of after them
                const int nxt = stack[sp - 1];
                if (nxt >= 0) asm volatile("prefetch.global.L1 [%0];" ::"l"(sc.nodes + 4 * size_t(nxt)));
            }
#endif
            for (uint32_t i = first; i < end; i++) {
                const float4* tp = sc.tris + 3 * size_t(i);
                float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
                if (STATS) cnt.prims++;
                float t;
                if (hit_tri(f3(a), f3(b), f3(c), ro, rd, tmin, hit.t, t)) {
                    hit.t = t;
                    hit.item = blas_item;
                    hit.prim = i;
                }
            }
        } else if (tlas_leaf<STATS, F>(sc, ~uint32_t(cur), o, d, tmin, rng, hit, stack, sp, ro, rd, blas_item, cnt)) {
            in_blas = true;
            rb.set(ro, rd);
        }
        cur = stack[--sp];
    }
};

// ---------------------------------------------------------------------------------- 8-wide compressed tree
// Traversal of the zb_wide_node tree (host/wide_bvh.hpp; Ylitie, Karras, Laine 2017): one visit = five 16-byte loads and
// eight slab tests on 8-bit boxes.  A quantised coordinate q is turned into the float 1 + q 2^-15 by ONE byte permute
// (q into mantissa bits 8..15 of 0x3F800000) and the slab distance t = (p + q 2^e - o) / d into ONE FFMA, f * A + C with
// A = 2^(e+15) / d and C = (p - o) / d - A, both per node and axis.  The stack holds GROUPS, not nodes: (child base, hit
// bits of the inner children in bits 24..31 | imask) or (primitive base, hit bits of up to 24 primitives); inner children
// are taken highest bit first, which — their bit being 24 + (slot XOR ray octant) — is front to back.  Stack need per ray:
// one group per level plus three for a dive into a mesh (zb_flat_scene::wide_stack_need <= ZB_WSTACK - 1, checked at upload).
#define ZB_WSTACK 32
#define ZB_WSTACK_CANARY 0xFFFFFFFEu
#ifndef ZB_WIDE_ANY
#ifdef ZB_EXPERIMENTS
#define ZB_WIDE_ANY 1  // experiments build: kernels of every feature set carry the wide traversal (A/B on TLAS-only scenes)
#else
#define ZB_WIDE_ANY 0
#endif
#endif

ZB_DEV uint32_t ray_octant(float3 d) { return (d.x >= 0.f ? 4u : 0u) | (d.y >= 0.f ? 2u : 0u) | (d.z >= 0.f ? 1u : 0u); }

template <bool STATS, uint32_t F>
ZB_DEV bool trace_wide(const SceneDev& sc, float3 o, float3 d, float tmin, float tmax, const MediumRng& rng, Hit& hit, Counters& cnt) {
    uint2 stack[ZB_WSTACK];
    int sp = 0;
    stack[ZB_WSTACK - 1] = make_uint2(ZB_WSTACK_CANARY, ZB_WSTACK_CANARY);
    cnt.rays++;
    hit.t = tmax;
    hit.item = 0xFFFFFFFFu;
    hit.prim = 0;
    float3 ro = o, rd = d;
    RayBox rb;
    rb.set(o, d);
    uint32_t oct = ray_octant(d);
    bool in_blas = false;
    uint32_t blas_item = 0;
    uint2 G = make_uint2(0u, 0x80000000u), T = make_uint2(0u, 0u);
    for (;;) {
        if (G.y & 0xFF000000u) {
            // ---- take the nearest pending inner child of the group and visit it
            const uint32_t bit = 31u - uint32_t(__clz(G.y));
            G.y &= ~(1u << bit);
            const uint32_t slot = (bit - 24u) ^ oct;
            const uint32_t ni = G.x + uint32_t(__popc(G.y & 0xFFu & ((1u << slot) - 1u)));
            if (G.y & 0xFF000000u) {
                stack[sp] = G;
                sp = min(sp + 1, ZB_WSTACK - 1);
            }
            const float4* np = sc.wide + 5 * size_t(ni);
            const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
            if (STATS) cnt.nodes++;
            const uint32_t ew = __float_as_uint(n0.w);
            const float ax = __uint_as_float(((ew & 0xFFu) + 15u) << 23) * rb.inv.x;
            const float ay = __uint_as_float((((ew >> 8) & 0xFFu) + 15u) << 23) * rb.inv.y;
            const float az = __uint_as_float((((ew >> 16) & 0xFFu) + 15u) << 23) * rb.inv.z;
            const float cx = fmaf(n0.x, rb.inv.x, -rb.ood.x) - ax, cy = fmaf(n0.y, rb.inv.y, -rb.ood.y) - ay,
                        cz = fmaf(n0.z, rb.inv.z, -rb.ood.z) - az;
            const bool xn = rb.inv.x < 0.f, yn = rb.inv.y < 0.f, zn = rb.inv.z < 0.f;
            // words of the near / far planes per axis: lo planes are near for a positive direction
            const uint32_t lox[2] = {__float_as_uint(n2.x), __float_as_uint(n2.y)}, loy[2] = {__float_as_uint(n2.z), __float_as_uint(n2.w)},
                           loz[2] = {__float_as_uint(n3.x), __float_as_uint(n3.y)}, hix[2] = {__float_as_uint(n3.z), __float_as_uint(n3.w)},
                           hiy[2] = {__float_as_uint(n4.x), __float_as_uint(n4.y)}, hiz[2] = {__float_as_uint(n4.z), __float_as_uint(n4.w)};
            const uint32_t metaw[2] = {__float_as_uint(n1.z), __float_as_uint(n1.w)};
            uint32_t mask = 0;
#pragma unroll
            for (int w = 0; w < 2; w++) {
                const uint32_t nx = xn ? hix[w] : lox[w], fx = xn ? lox[w] : hix[w];
                const uint32_t ny = yn ? hiy[w] : loy[w], fy = yn ? loy[w] : hiy[w];
                const uint32_t nz = zn ? hiz[w] : loz[w], fz = zn ? loz[w] : hiz[w];
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    const uint32_t sel = 0x7604u | (uint32_t(b) << 4);
                    const float tnx = fmaf(__uint_as_float(__byte_perm(nx, 0x3F800000u, sel)), ax, cx);
                    const float tny = fmaf(__uint_as_float(__byte_perm(ny, 0x3F800000u, sel)), ay, cy);
                    const float tnz = fmaf(__uint_as_float(__byte_perm(nz, 0x3F800000u, sel)), az, cz);
                    const float tfx = fmaf(__uint_as_float(__byte_perm(fx, 0x3F800000u, sel)), ax, cx);
                    const float tfy = fmaf(__uint_as_float(__byte_perm(fy, 0x3F800000u, sel)), ay, cy);
                    const float tfz = fmaf(__uint_as_float(__byte_perm(fz, 0x3F800000u, sel)), az, cz);
                    const float tn = fmaxf(max3(tnx, tny, tnz), tmin);
                    const float tf = fminf(min3(tfx, tfy, tfz), hit.t) * ZB_BOX_INFLATE;
                    const uint32_t m = (metaw[w] >> (8 * b)) & 0xFFu;
                    uint32_t idx = m & 31u;
                    idx ^= idx >= 24u ? oct : 0u;
                    const uint32_t bits = ((1u << (m >> 5)) - 1u) << idx;
                    if (tn <= tf) mask |= bits;
                }
            }
            G = make_uint2(__float_as_uint(n1.x), (mask & 0xFF000000u) | (ew >> 24));
            T = make_uint2(__float_as_uint(n1.y), mask & 0x00FFFFFFu);
        } else {
            T = G;  // a primitive group popped from the stack
            G = make_uint2(0u, 0u);
        }
        // ---- the primitives of the group
        while (T.y) {
            const uint32_t j = uint32_t(__ffs(int(T.y))) - 1u;
            T.y &= T.y - 1u;
            const uint32_t i = T.x + j;
            if (STATS) cnt.prims++;
            float t;
            if ((F & ZB_F_MESH) && in_blas) {
                const float4* tp = sc.tris + 3 * size_t(i);
                const float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
                if (hit_tri(f3(a), f3(b), f3(c), ro, rd, tmin, hit.t, t)) {
                    hit.t = t;
                    hit.item = blas_item;
                    hit.prim = i;
                }
                continue;
            }
            const float4 g = __ldg(sc.items + 2 * size_t(i));
            if (!(F & ZB_F_GEOM)) {
                if (hit_sphere(g, o, d, tmin, hit.t, t)) { hit.t = t; hit.item = i; }
                continue;
            }
            const float4 mf = __ldg(sc.items + 2 * size_t(i) + 1);
            const uint32_t kind = __float_as_uint(mf.x) & 0xFFu, aux = __float_as_uint(mf.w);
            if (kind == ZB_ITEM_SPHERE) {
                if (hit_sphere(g, o, d, tmin, hit.t, t)) { hit.t = t; hit.item = i; }
            } else if ((F & ZB_F_INST) && kind == ZB_ITEM_INSTANCE) {
                const InstanceHead ih = load_instance_head(sc, aux);
                if (STATS) cnt.insts++;
                const float3 lo = xform_point(ih.w2o, o), ld = xform_vector(ih.w2o, d);
                if ((F & ZB_F_MESH) && ih.geom_kind == ZB_GEOM_MESH) {
                    // dive into the mesh: what is left of this level, the way back, then the BLAS root in object space
                    if (G.y & 0xFF000000u) { stack[sp] = G; sp = min(sp + 1, ZB_WSTACK - 1); }
                    if (T.y) { stack[sp] = T; sp = min(sp + 1, ZB_WSTACK - 1); }
                    stack[sp] = make_uint2(0xFFFFFFFFu, 0u);
                    sp = min(sp + 1, ZB_WSTACK - 1);
                    in_blas = true;
                    blas_item = i;
                    ro = lo;
                    rd = ld;
                    rb.set(ro, rd);
                    oct = ray_octant(rd);
                    G = make_uint2(ih.wide_root, 0x80000000u);
                    T = make_uint2(0u, 0u);
                    break;
                }
                if (hit_simple_geom<F>(sc, ih.geom_kind, ih.geom_index, lo, ld, tmin, hit.t, t)) {
                    hit.t = t;
                    hit.item = i;
                    hit.prim = ih.geom_index;
                }
            } else if ((F & ZB_F_CUBE) && kind == ZB_ITEM_CUBE) {
                if (hit_cube(f3(__ldg(sc.cubes + 2 * size_t(aux))), o, d, tmin, hit.t, t)) { hit.t = t; hit.item = i; }
            } else if ((F & ZB_F_TRI) && kind == ZB_ITEM_TRIANGLE) {
                const float4* tp = sc.tris + 3 * size_t(aux);
                if (hit_tri(f3(__ldg(tp)), f3(__ldg(tp + 1)), f3(__ldg(tp + 2)), o, d, tmin, hit.t, t)) {
                    hit.t = t;
                    hit.item = i;
                    hit.prim = aux;
                }
            } else if ((F & ZB_F_MEDIA) && kind == ZB_ITEM_MEDIUM) {
                if (hit_medium<F>(sc, aux, o, d, tmin, hit.t, rng, t)) { hit.t = t; hit.item = i; }
            }
        }
        // ---- next group
        if (!(G.y & 0xFF000000u)) {
            bool done = false;
            for (;;) {
                if (sp == 0) { done = true; break; }
                G = stack[--sp];
                if (G.y != 0u) break;
                // the way back out of a mesh: the world-space ray again
                in_blas = false;
                ro = o;
                rd = d;
                rb.set(o, d);
                oct = ray_octant(d);
            }
            if (done) break;
        }
    }
    if (stack[ZB_WSTACK - 1].x != ZB_WSTACK_CANARY) atomicAdd(sc.overflow, 1ull);
    return hit.item != 0xFFFFFFFFu;
}

// Closest hit of one ray, run to completion.  "while-while" traversal (Aila & Laine 2009): every lane first descends
// through inner nodes until it holds a leaf, then the warp runs the primitive tests together; this keeps more lanes
// on the same instruction than one loop that alternates between a node step and a leaf step.
template <bool STATS, uint32_t F = ZB_F_ALL>
ZB_DEV bool trace(const SceneDev& sc, float3 o, float3 d, float tmin, float tmax, const MediumRng& rng, Hit& hit,
                  Counters& cnt) {
    if constexpr ((F & ZB_F_MESH) != 0 || ZB_WIDE_ANY) {
        if (sc.wide) return trace_wide<STATS, F>(sc, o, d, tmin, tmax, rng, hit, cnt);  // warp-uniform: a launch parameter
    }
    int stack[ZB_STACK];
    cnt.rays++;
    RayWalk w;
    w.begin(sc, o, d, tmax, stack);
    while (w.cur != ZB_TRAVERSAL_DONE) {
        while (w.cur >= 0) w.node_step<STATS, (F & ZB_F_MESH) != 0 && ZB_PREFETCH_FAR != 0>(sc, tmin, stack, cnt);
        if (w.cur == ZB_TRAVERSAL_DONE) break;
        w.leaf_step<STATS, F>(sc, o, d, tmin, rng, stack, cnt);
    }
    hit = w.hit;
    if (stack[ZB_STACK - 1] != ZB_STACK_CANARY) atomicAdd(sc.overflow, 1ull);  // deferred subtree lost: the render fails
    return hit.item != 0xFFFFFFFFu;
}
