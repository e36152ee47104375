// wireframe.cu — BVH debug view: bvh_node::hit with debug_wire (bvh.hpp:46-110),
// aabb::is_on_edge (aabb.hpp:68-84) and the debug branch of ray_color (camera.hpp:937-954), on the product's own
// two-level tree.
//
// Reference semantics kept: only the primary ray counts; a node at the chosen level (level >= 0: depth == level,
// level -1: a node whose children are primitives — here a leaf range) is drawn as a frame where the ray's entry or
// exit point on its box lies within `thickness * (0.05 + 0.1 t_entry)` of two or three box planes; a frame hit ends
// the descent below that node and shrinks the ray interval (t = entry, or exit when only the exit point is on an
// edge); geometry below a level node takes the node's dim "volume" colour, geometry elsewhere its own emission; a
// mesh's tree restarts at depth 0 in object space (model::hit calls its bvh_node with the default depth,
// model.hpp:105-107) and an enclosing level node of the outer tree overrides what the inner tree found
// (bvh.hpp:96-101).  render_rows (camera.hpp:454-531) shows the regular background where the primary ray misses, the
// frame / volume colour where it hits one (those records carry a diffuse_light, which does not scatter), and for
// geometry outside every level node one scatter whose second segment runs ray_color's debug branch
// (camera.hpp:937-954).  Colours: (0.4, g, 1 - g) with g = min(0.15 depth, 1), x 4 for frames, x 0.1 for volumes.
// What differs by construction: the tree (binned SAH here, random-axis median split there — the reference's own
// picture changes from run to run, bvh.hpp:17), the visiting order (near child first here), and material_instance's
// replacement of rec.mat after the inner hit (material_instance.hpp:12-28), which hides a wrapped mesh's frames in
// the reference and is not reproduced.  Level 0 (the root box) does not depend on the tree: that case is pinned on
// the reference's output (tests/golden/wire_*.npz).
#include "kernels.h"
#include "shade.cuh"

namespace {

struct WireBest {
    float t;
    int kind;   // 0 nothing, 1 geometry, 2 frame
    int cover;  // geometry: depth of the enclosing level node or -1; frame: depth of the framed node
    int outer;  // frame: depth of an enclosing level node of an outer tree, or -1
    Hit hit;
};

ZB_DEV bool wire_on_edge(float3 p, float3 lo, float3 hi, float th) {  // aabb.hpp:68-84
    const int nx = (fabsf(p.x - lo.x) < th) || (fabsf(p.x - hi.x) < th);
    const int ny = (fabsf(p.y - lo.y) < th) || (fabsf(p.y - hi.y) < th);
    const int nz = (fabsf(p.z - lo.z) < th) || (fabsf(p.z - hi.z) < th);
    return nx + ny + nz >= 2;
}

// aabb::hit (aabb.hpp:44-66) on [tmin, tmax]: true when the clipped interval is non-empty
ZB_DEV bool wire_slab(float3 lo, float3 hi, float3 o, float3 d, float tmin, float tmax, float& tn, float& tf) {
    const float ix = 1.0f / d.x, iy = 1.0f / d.y, iz = 1.0f / d.z;
    float a = (lo.x - o.x) * ix, b = (hi.x - o.x) * ix;
    tn = fmaxf(tmin, fminf(a, b));
    tf = fminf(tmax, fmaxf(a, b));
    if (tf <= tn) return false;
    a = (lo.y - o.y) * iy; b = (hi.y - o.y) * iy;
    tn = fmaxf(tn, fminf(a, b));
    tf = fminf(tf, fmaxf(a, b));
    if (tf <= tn) return false;
    a = (lo.z - o.z) * iz; b = (hi.z - o.z) * iz;
    tn = fmaxf(tn, fminf(a, b));
    tf = fminf(tf, fmaxf(a, b));
    return tf > tn;
}

ZB_DEV void wire_child_box(const float4& n0, const float4& n1, const float4& n2, int side, float3& lo, float3& hi) {
#if ZB_BOX_CH  // nodes hold centre / half extent (zbk_convert_nodes)
    const float3 c = side ? f3(n1.x, n1.z, n2.z) : f3(n0.x, n0.z, n2.x), h = side ? f3(n1.y, n1.w, n2.w) : f3(n0.y, n0.w, n2.y);
    lo = c - h;
    hi = c + h;
#else
    lo = side ? f3(n1.x, n1.z, n2.z) : f3(n0.x, n0.z, n2.x);
    hi = side ? f3(n1.y, n1.w, n2.w) : f3(n0.y, n0.w, n2.y);
#endif
}

#define WIRE_EXIT 0x7FFFFFFF
#define WIRE_STACK 96

// One primary ray through the debug traversal.  Stack entries: a = node * 2 + side (a child box, stored in its
// parent), a = -2 - node (the root box of a tree = union of the root's child boxes) or WIRE_EXIT (back to world
// space); b = depth | (cover + 1) << 16; si = the instance item of a mesh's root entry.
ZB_DEV WireBest wire_trace(const SceneDev& sc, float3 o, float3 d, const MediumRng& rng, int level, float thickness) {
    WireBest best;
    best.t = INFINITY;
    best.kind = 0;
    best.cover = best.outer = -1;
    best.hit.t = INFINITY;
    best.hit.item = 0xFFFFFFFFu;
    best.hit.prim = 0;
    int sa[WIRE_STACK], sb[WIRE_STACK], sp = 0;
    uint32_t si[WIRE_STACK];  // tree-root entries of a mesh: the instance item (0xFFFFFFFF for the scene's own tree)
    float3 ro = o, rd = d;
    bool in_blas = false;
    uint32_t blas_item = 0;
    const float tmin = 0.001f;
    if (sc.n_nodes) { sa[0] = -2; sb[0] = 0; si[0] = 0xFFFFFFFFu; sp = 1; }
    while (sp > 0) {
        const int a = sa[--sp], b = sb[sp];
        const uint32_t root_item = si[sp];
        if (a == WIRE_EXIT) {
            in_blas = false;
            ro = o;
            rd = d;
            continue;
        }
        const int depth = b & 0xFFFF, cover_in = (b >> 16) - 1;
        float3 lo, hi;
        int code;
        if (a <= -2) {
            code = -2 - a;
            if (root_item != 0xFFFFFFFFu) {
                const float4 mf = __ldg(sc.items + 2 * size_t(root_item) + 1);
                const InstanceHead ih = load_instance_head(sc, __float_as_uint(mf.w));
                ro = xform_point(ih.w2o, o);
                rd = xform_vector(ih.w2o, d);
                in_blas = true;
                blas_item = root_item;
            }
            const float4* np = sc.nodes + 4 * size_t(code);
            const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2);
            float3 l0, h0, l1, h1;
            wire_child_box(n0, n1, n2, 0, l0, h0);
            wire_child_box(n0, n1, n2, 1, l1, h1);
            lo = f3(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z));
            hi = f3(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z));
        } else {
            const float4* np = sc.nodes + 4 * size_t(a >> 1);
            const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2);
            const float2 nc = __ldg(reinterpret_cast<const float2*>(np + 3));
            wire_child_box(n0, n1, n2, a & 1, lo, hi);
            code = __float_as_int((a & 1) ? nc.y : nc.x);
        }
        float tn, tf;
        if (!wire_slab(lo, hi, ro, rd, tmin, best.t, tn, tf)) continue;
        const bool level_here = level < 0 ? code < 0 : depth == level;
        if (level_here) {
            const float th = thickness * (0.05f + tn * 0.1f);
            const bool e_in = wire_on_edge(ro + (tn + 0.0001f) * rd, lo, hi, th);
            if (e_in || wire_on_edge(ro + (tf - 0.0001f) * rd, lo, hi, th)) {
                best.t = e_in ? tn : tf;
                best.kind = 2;
                best.cover = depth;
                best.outer = cover_in;
                continue;  // a frame hit ends the descent below this node (bvh.hpp:75-88)
            }
        }
        const int cover = cover_in >= 0 ? cover_in : (level_here ? depth : -1);
        const int meta = (depth + 1) | ((cover + 1) << 16);
        if (code >= 0) {
            // children: the nearer box is visited (popped) first
            const float4* np = sc.nodes + 4 * size_t(code);
            const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2);
            float3 l0, h0, l1, h1;
            wire_child_box(n0, n1, n2, 0, l0, h0);
            wire_child_box(n0, n1, n2, 1, l1, h1);
            float t0n, t0f, t1n, t1f;
            const bool k0 = wire_slab(l0, h0, ro, rd, tmin, best.t, t0n, t0f), k1 = wire_slab(l1, h1, ro, rd, tmin, best.t, t1n, t1f);
            const int first = (k1 && (!k0 || t1n < t0n)) ? 1 : 0;
            if (sp + 2 > WIRE_STACK) continue;
            sa[sp] = code * 2 + (first ^ 1); sb[sp] = meta; si[sp++] = 0;
            sa[sp] = code * 2 + first; sb[sp] = meta; si[sp++] = 0;
            continue;
        }
        // leaf range
        const uint32_t lc = ~uint32_t(code), first = lc >> 3, end = first + (lc & 7u) + 1u;
        if (in_blas) {
            for (uint32_t i = first; i < end; i++) {
                const float4* tp = sc.tris + 3 * size_t(i);
                float t;
                if (hit_tri(f3(__ldg(tp)), f3(__ldg(tp + 1)), f3(__ldg(tp + 2)), ro, rd, tmin, best.t, t)) {
                    best.t = t; best.kind = 1; best.cover = cover;
                    best.hit.t = t; best.hit.item = blas_item; best.hit.prim = i;
                }
            }
            continue;
        }
        for (uint32_t i = first; i < end; i++) {
            const float4 g = __ldg(sc.items + 2 * size_t(i)), mf = __ldg(sc.items + 2 * size_t(i) + 1);
            const uint32_t kind = __float_as_uint(mf.x) & 0xFFu, aux = __float_as_uint(mf.w);
            float t;
            bool got = false;
            uint32_t prim = 0;
            if (kind == ZB_ITEM_SPHERE) {
                got = hit_sphere(g, o, d, tmin, best.t, t);
            } else if (kind == ZB_ITEM_CUBE) {
                got = hit_cube(f3(__ldg(sc.cubes + 2 * size_t(aux))), o, d, tmin, best.t, t);
            } else if (kind == ZB_ITEM_TRIANGLE) {
                const float4* tp = sc.tris + 3 * size_t(aux);
                got = hit_tri(f3(__ldg(tp)), f3(__ldg(tp + 1)), f3(__ldg(tp + 2)), o, d, tmin, best.t, t);
                prim = aux;
            } else if (kind == ZB_ITEM_INSTANCE) {
                const InstanceHead ih = load_instance_head(sc, aux);
                if (ih.geom_kind == ZB_GEOM_MESH) {
                    // the mesh's own tree: depth 0 again, object space (the ray is transformed when the entry is popped)
                    if (sp + 2 <= WIRE_STACK) {
                        sa[sp] = WIRE_EXIT; sb[sp] = 0; si[sp++] = 0;
                        sa[sp] = -2 - int(ih.node_root); sb[sp] = (cover + 1) << 16; si[sp++] = i;
                    }
                } else {
                    const float3 lo2 = xform_point(ih.w2o, o), ld2 = xform_vector(ih.w2o, d);
                    got = hit_simple_geom(sc, ih.geom_kind, ih.geom_index, lo2, ld2, tmin, best.t, t);
                    prim = ih.geom_index;
                }
            } else {
                got = hit_medium(sc, aux, o, d, tmin, best.t, rng, t);
            }
            if (got) {
                best.t = t; best.kind = 1; best.cover = cover;
                best.hit.t = t; best.hit.item = i; best.hit.prim = prim;
            }
        }
    }
    return best;
}

// what rec.mat->emitted() returns for the record the debug traversal produced; `debug_light` = the record carries one
// of the traversal's own diffuse_light materials (frame or volume) and therefore never scatters
ZB_DEV float3 wire_emitted(const SceneDev& sc, const WireBest& best, float3 o, float3 d, bool& debug_light, Surface& sf) {
    const int vol = best.kind == 2 ? best.outer : best.cover;
    debug_light = true;
    if (vol >= 0) {
        const float g = fminf(float(vol) * 0.15f, 1.0f);
        return 0.1f * f3(0.4f, g, 1.0f - g);  // bvh.hpp:96-101
    }
    if (best.kind == 2) {
        const float g = fminf(float(best.cover) * 0.15f, 1.0f);
        return 4.0f * f3(0.4f, g, 1.0f - g);  // bvh.hpp:82-86
    }
    debug_light = false;
    make_surface(sc, best.hit, o, d, sf);
    return material_emitted(sc, sf);
}

// One camera sample of render_rows with global_settings::bvh_debug_mode set (camera.hpp:454-531): the primary hit goes
// through the debug traversal; a miss shows the regular background; a frame or volume record is a light and ends the
// path with its colour; other geometry scatters once and the second segment runs ray_color's debug branch
// (camera.hpp:937-954): black on a miss, the emission when its length exceeds 0.1, else (0.01, 0.01, 0.01).
ZB_DEV float3 wire_sample(const RenderParams& P, const Philox& ph, uint32_t pixel, uint32_t s, float3 o, float3 d, int level, float thickness) {
    const MediumRng mr{ph, pixel, s, 1u};
    const WireBest first = wire_trace(P.sc, o, d, mr, level, thickness);
    if (first.kind == 0) return background(P.sc, P.env, d);
    bool light;
    Surface sf;
    float3 L = wire_emitted(P.sc, first, o, d, light, sf);
    if (light || P.max_depth <= 1) return L;
    const uint4 rb = ph(pixel, s, 1u, 0u);
    float3 atten = f3(1, 1, 1), no, nd;
    if (!material_scatter(P.sc, sf, d, f3(u01(rb.x), u01(rb.y), u01(rb.z)), atten, no, nd)) return L;
    const MediumRng mr2{ph, pixel, s, 2u};
    const WireBest second = wire_trace(P.sc, no, nd, mr2, level, thickness);
    if (second.kind == 0) return L;
    const float3 e = wire_emitted(P.sc, second, no, nd, light, sf);
    return L + atten * (length(e) > 0.1f ? e : f3(0.01f, 0.01f, 0.01f));
}

__global__ void __launch_bounds__(128) k_wireframe(const __grid_constant__ RenderParams P, int level, float thickness) {
    const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 4 + (threadIdx.x >> 5);
    if (px >= P.cam.width || py >= P.cam.height) return;
    const uint32_t pixel = uint32_t(py) * uint32_t(P.cam.width) + uint32_t(px);
    const Philox ph{P.pk[0]};
    float3 sum = f3(0, 0, 0);
    for (int s = P.sample_begin; s < P.sample_end; s++) {
        // K1 raygen, the same draws as the beauty pass (camera.hpp:784-794, :817, :822)
        const uint4 r0 = ph(pixel, uint32_t(s), 0u, 0u);
        float3 d = P.cam.pix00_rel + (float(px) + u01(r0.x) - 0.5f) * P.cam.du + (float(py) + u01(r0.y) - 0.5f) * P.cam.dv;
        float3 o = P.cam.center;
        if (P.cam.defocus) {
            float rr = sqrt_fast(u01(r0.z)), sn, cs;
            sincos_turn(u01(r0.w), &sn, &cs);
            const float3 off = (rr * cs) * P.cam.disk_u + (rr * sn) * P.cam.disk_v;
            o += off;
            d = d - off;
        }
        sum += wire_sample(P, ph, pixel, uint32_t(s), o, d, level, thickness);
    }
    float4 acc = make_float4(sum.x, sum.y, sum.z, float(P.sample_end - P.sample_begin));
    if (P.accumulate) {
        const float4 old = P.accum[pixel];
        acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
    }
    P.accum[pixel] = acc;
}

}  // namespace

cudaError_t zbk_launch_wireframe(const RenderParams& P, int level, float thickness, cudaStream_t stream) {
    const dim3 grid((P.cam.width + 31) / 32, (P.cam.height + 3) / 4);
    k_wireframe<<<grid, 128, 0, stream>>>(P, level, thickness);
    return cudaGetLastError();
}
