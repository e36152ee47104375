// render_pool.cuh — k_render_pool: the megakernel with a per-warp RAY POOL in shared memory.
//
// Same work decomposition, RNG keys, estimator and fixed-point tile sums as k_render (kernels.cu); what
// changes is how the 32 lanes of a warp are kept busy.  k_render binds one path to one lane and runs
// "trace to completion, then shade" per lane, so in the node loop only 13.5 of 32 lanes are active (ncu,
// profiles/r01_k_render_v2_lines.txt): every lane waits for the longest traversal of its warp.  Here each
// warp keeps ZB_POOL_R (> 32) paths in shared memory and alternates two warp-wide phases:
//   SHADE  lane i shades slots i, i+32, ...: hit -> scatter / terminate, terminated slots are refilled with
//          the next (pixel, sample) of the tile queue and a camera ray.  All lanes do the same thing.
//   TRACE  the slots form a queue; a lane takes a slot, traverses, stores (t, item, prim) and takes the next.
//          Finished lanes are refilled in batches (when fewer than ZB_POOL_T lanes are still traversing), so
//          the cost of the divergent refill code is shared by several rays.
// No path state lives in registers across the two phases.
#pragma once
#include "shade.cuh"

#ifndef ZB_POOL_R
#define ZB_POOL_R 64  // path slots per warp
#endif
#ifndef ZB_POOL_T
#define ZB_POOL_T 22  // refill the finished lanes when fewer than this many lanes still traverse
#endif
#ifndef ZB_POOL_MIN_BLOCKS
#define ZB_POOL_MIN_BLOCKS 6
#endif

#define ZB_FIX_SCALE 4194304.0f
#define ZB_FIX_INV (1.0 / 4194304.0)

// 64-bit shared-memory add from two native 32-bit atomics (a 64-bit ATOMS add is a CAS spin loop):
// the low word wraps, and exactly the operation whose addition wrapped it carries one into the high
// word, so (hi, lo) ends as the exact sum whatever the order.  Radiance is never negative.
ZB_DEV void fix_add(unsigned long long* slot, float v) {
    const unsigned long long q = __float2ull_rn(fmaxf(v, 0.f) * ZB_FIX_SCALE);
    const uint32_t ql = uint32_t(q), qh = uint32_t(q >> 32);
    uint32_t* w = reinterpret_cast<uint32_t*>(slot);
    const uint32_t old = atomicAdd(w, ql);
    const uint32_t up = qh + (old + ql < old ? 1u : 0u);
    if (up) atomicAdd(w + 1, up);
}
ZB_DEV float fix_get(unsigned long long v) { return float(double((long long)v) * ZB_FIX_INV); }

// One warp's pool, structure-of-arrays so that lane i touching slot i (+32 k) is conflict free.
struct WarpPool {
    float o[3][ZB_POOL_R], d[3][ZB_POOL_R], beta[3][ZB_POOL_R];
    uint32_t meta[ZB_POOL_R];    // bit 31 alive | seg << 8 | pixel-in-tile (5 bits)
    uint32_t sample[ZB_POOL_R];  // global sample index
    float hit_t[ZB_POOL_R];
    uint32_t hit_item[ZB_POOL_R], hit_prim[ZB_POOL_R];
    unsigned long long sum[32][3];
    unsigned long long aux[32][7];  // albedo rgb, normal xyz, z
    int next_path;                  // tile queue: next (pixel, sample)
    int next_slot;                  // trace queue: next slot to trace
};
#define ZB_SLOT_ALIVE 0x80000000u

template <bool STATS>
__global__ void __launch_bounds__(ZB_BLOCK_THREADS, ZB_POOL_MIN_BLOCKS) k_render_pool(const __grid_constant__ RenderParams P) {
    constexpr int WARPS = ZB_BLOCK_THREADS / 32;
    constexpr int GROUPS = ZB_POOL_R / 32;
    __shared__ WarpPool s_pool[WARPS];

    const int tile = blockIdx.x * P.tile_ranks + P.tile_rank;
    const int tx = tile % P.tiles_x, ty = tile / P.tiles_x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int px0 = tx * ZB_TILE_W + (warp % (ZB_TILE_W / 8)) * 8, py0 = ty * ZB_TILE_H + (warp / (ZB_TILE_W / 8)) * 4;
    const bool aux_on = P.accum_aux != nullptr;
    WarpPool& W = s_pool[warp];

#pragma unroll
    for (int c = 0; c < 3; c++) W.sum[lane][c] = 0ull;
#pragma unroll
    for (int c = 0; c < 7; c++) W.aux[lane][c] = 0ull;
#pragma unroll
    for (int g = 0; g < GROUPS; g++) W.meta[g * 32 + lane] = 0u;
    if (lane == 0) { W.next_path = 0; W.next_slot = 0; }
    __syncwarp();

    const int n_s = P.sample_end - P.sample_begin;
    const int total = (ty < P.tiles_y && P.max_depth > 0) ? 32 * n_s : 0;
    const Philox ph{P.seed_lo, P.seed_hi};
    Counters cnt = {0, 0, 0, 0};
    uint32_t paths = 0;

    for (;;) {
        // ================================================================== SHADE + regenerate
        bool any_alive = false;
#pragma unroll 1
        for (int g = 0; g < GROUPS; g++) {
            const int slot = g * 32 + lane;
            uint32_t meta = W.meta[slot];
            bool alive = (meta & ZB_SLOT_ALIVE) != 0u;
            float3 o, d, beta;
            uint32_t s = 0;
            int p = int(meta & 31u), seg = int((meta >> 8) & 0xFFFFu);
            if (alive) {
                // ---- K3: one segment of ray_color (camera.hpp:928-1004) for the hit stored by TRACE
                o = f3(W.o[0][slot], W.o[1][slot], W.o[2][slot]);
                d = f3(W.d[0][slot], W.d[1][slot], W.d[2][slot]);
                beta = f3(W.beta[0][slot], W.beta[1][slot], W.beta[2][slot]);
                s = W.sample[slot];
                Hit hit;
                hit.t = W.hit_t[slot];
                hit.item = W.hit_item[slot];
                hit.prim = W.hit_prim[slot];
                const uint32_t pixel = uint32_t(py0 + (p >> 3)) * uint32_t(P.cam.width) + uint32_t(px0 + (p & 7));
                if (hit.item == 0xFFFFFFFFu) {
                    float3 c = beta * background(P.sc, P.env, d);
                    fix_add(&W.sum[p][0], c.x);
                    fix_add(&W.sum[p][1], c.y);
                    fix_add(&W.sum[p][2], c.z);
                    if (seg == 0 && aux_on && int(s) < P.aux_samples && P.use_normal) {  // camera.hpp:518-526
                        fix_add(&W.aux[p][3], 0.5f);
                        fix_add(&W.aux[p][4], 0.5f);
                        fix_add(&W.aux[p][5], 1.0f);
                    }
                    alive = false;
                } else {
                    Surface sf;
                    make_surface(P.sc, hit, o, d, sf);
                    if (seg == 0 && aux_on && int(s) < P.aux_samples) {  // K8 first-hit AOVs (camera.hpp:464-488)
                        if (P.use_albedo) {
                            float3 a = material_get_albedo(P.sc, sf);
                            fix_add(&W.aux[p][0], a.x);
                            fix_add(&W.aux[p][1], a.y);
                            fix_add(&W.aux[p][2], a.z);
                        }
                        if (P.use_normal) {
                            float3 n = unit_vector(sf.n);
                            fix_add(&W.aux[p][3], (dot(n, P.cam.u) + 1.f) * 0.5f);
                            fix_add(&W.aux[p][4], (dot(n, P.cam.v) + 1.f) * 0.5f);
                            fix_add(&W.aux[p][5], (dot(n, P.cam.w) + 1.f) * 0.5f);
                        }
                        if (P.use_z) fix_add(&W.aux[p][6], 1.0f - clamp01(hit.t * P.inv_z_max));
                    }
                    if (__float_as_uint(sf.m0.x) == ZB_MAT_LIGHT) {
                        float3 e = beta * material_emitted(P.sc, sf);
                        fix_add(&W.sum[p][0], e.x);
                        fix_add(&W.sum[p][1], e.y);
                        fix_add(&W.sum[p][2], e.z);
                    }
                    uint4 rb = ph(pixel, s, uint32_t(seg + 1), 0u);
                    float3 atten = f3(1, 1, 1), no, nd;
                    alive = material_scatter(P.sc, sf, d, f3(u01(rb.x), u01(rb.y), u01(rb.z)), atten, no, nd);
                    if (alive) {
                        beta *= atten;
                        o = no;
                        d = nd;
                        // ray_color's loop index is seg-1; its `i > 10` rules start at seg 12 (camera.hpp:968-983)
                        if (seg >= 12) {
                            float pr = fminf(fmaxf(max3(beta.x, beta.y, beta.z), 0.05f), 0.95f);
                            if (dot(beta, beta) < 1e-8f || u01(rb.w) > pr) alive = false;
                            beta *= 1.0f / pr;
                        }
                        if (++seg >= P.max_depth) alive = false;  // hard stop keeps the light gathered so far
                    }
                }
            }
            if (!alive) {
                // ---- regenerate: next (pixel, sample) of this warp's tile queue whose pixel is inside the image
                for (;;) {
                    int k = atomicAdd(&W.next_path, 1);
                    if (k >= total) break;
                    p = k / n_s;
                    int px = px0 + (p & 7), py = py0 + (p >> 3);
                    if (px >= P.cam.width || py >= P.cam.height) continue;
                    s = uint32_t(P.sample_begin + (k - p * n_s));
                    const uint32_t pixel = uint32_t(py) * uint32_t(P.cam.width) + uint32_t(px);
                    // K1 raygen: get_ray (camera.hpp:784-794), sample_square (:817), defocus disk (:822)
                    uint4 r0 = ph(pixel, s, 0u, 0u);
                    d = P.cam.pix00_rel + (float(px) + u01(r0.x) - 0.5f) * P.cam.du + (float(py) + u01(r0.y) - 0.5f) * P.cam.dv;
                    o = P.cam.center;
                    if (P.cam.defocus) {
                        float rr = sqrtf(u01(r0.z)), sn, cs;
                        sincospif(2.0f * u01(r0.w), &sn, &cs);
                        float3 off = (rr * cs) * P.cam.disk_u + (rr * sn) * P.cam.disk_v;
                        o += off;
                        d = d - off;
                    }
                    seg = 0;
                    beta = f3(1, 1, 1);
                    paths++;
                    alive = true;
                    break;
                }
            }
            if (alive) {
                W.o[0][slot] = o.x; W.o[1][slot] = o.y; W.o[2][slot] = o.z;
                W.d[0][slot] = d.x; W.d[1][slot] = d.y; W.d[2][slot] = d.z;
                W.beta[0][slot] = beta.x; W.beta[1][slot] = beta.y; W.beta[2][slot] = beta.z;
                W.sample[slot] = s;
                W.meta[slot] = ZB_SLOT_ALIVE | (uint32_t(seg) << 8) | uint32_t(p);
                any_alive = true;
            } else {
                W.meta[slot] = 0u;
            }
        }
        if (lane == 0) W.next_slot = 0;
        if (!__any_sync(0xFFFFFFFFu, any_alive)) break;
        __syncwarp();

        // ================================================================== TRACE (K2) over the slot queue
        {
            Traversal tv;
            tv.cur = ZB_TRAVERSAL_DONE;
            int slot = -1;
            float3 o = f3(0, 0, 0), d = f3(0, 0, 1);
            MediumRng mr{ph, 0u, 0u, 0u};
            for (;;) {
                const bool active = tv.cur != ZB_TRAVERSAL_DONE;
                const int n_act = __popc(__ballot_sync(0xFFFFFFFFu, active));
                const bool more = *reinterpret_cast<volatile int*>(&W.next_slot) < ZB_POOL_R;
                if (n_act == 0 && !more) break;
                if (more && n_act < ZB_POOL_T) {
                    // refill every finished lane: store its hit, take the next live slot of the queue
                    if (!active) {
                        if (slot >= 0) {
                            W.hit_t[slot] = tv.hit.t;
                            W.hit_item[slot] = tv.hit.item;
                            W.hit_prim[slot] = tv.hit.prim;
                            slot = -1;
                        }
                        for (;;) {
                            int k = atomicAdd(&W.next_slot, 1);
                            if (k >= ZB_POOL_R) break;
                            const uint32_t meta = W.meta[k];
                            if (!(meta & ZB_SLOT_ALIVE)) continue;
                            slot = k;
                            o = f3(W.o[0][k], W.o[1][k], W.o[2][k]);
                            d = f3(W.d[0][k], W.d[1][k], W.d[2][k]);
                            const int p = int(meta & 31u);
                            mr.pixel = uint32_t(py0 + (p >> 3)) * uint32_t(P.cam.width) + uint32_t(px0 + (p & 7));
                            mr.sample = W.sample[k];
                            mr.bounce = ((meta >> 8) & 0xFFFFu) + 1u;
                            tv.begin(P.sc, o, d, INFINITY);
                            cnt.rays++;
                            break;
                        }
                    }
                    continue;
                }
                if (active) tv.step<STATS>(P.sc, o, d, 0.001f, mr, cnt);
            }
            if (slot >= 0) {
                W.hit_t[slot] = tv.hit.t;
                W.hit_item[slot] = tv.hit.item;
                W.hit_prim[slot] = tv.hit.prim;
            }
        }
        __syncwarp();
    }
    __syncwarp();

    // ---- K5 accumulate: lane l writes pixel l of the sub-tile
    {
        const int px = px0 + (lane & 7), py = py0 + (lane >> 3);
        if (px < P.cam.width && py < P.cam.height && ty < P.tiles_y) {
            const size_t pix = size_t(py) * P.cam.width + px;
            float4 acc = make_float4(fix_get(W.sum[lane][0]), fix_get(W.sum[lane][1]), fix_get(W.sum[lane][2]), float(n_s));
            if (P.accumulate) {
                float4 old = P.accum[pix];
                acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
            }
            P.accum[pix] = acc;
            if (aux_on) {
                const size_t n = size_t(P.cam.width) * P.cam.height;
                float z = fix_get(W.aux[lane][6]);
                float4 a0 = make_float4(fix_get(W.aux[lane][0]), fix_get(W.aux[lane][1]), fix_get(W.aux[lane][2]), 0.f);
                float4 a1 = make_float4(fix_get(W.aux[lane][3]), fix_get(W.aux[lane][4]), fix_get(W.aux[lane][5]), 0.f);
                float4 a2 = make_float4(z, z, z, 0.f);
                if (P.accumulate) {
                    float4 b0 = P.accum_aux[pix], b1 = P.accum_aux[n + pix], b2 = P.accum_aux[2 * n + pix];
                    a0.x += b0.x; a0.y += b0.y; a0.z += b0.z;
                    a1.x += b1.x; a1.y += b1.y; a1.z += b1.z;
                    a2.x += b2.x; a2.y += b2.y; a2.z += b2.z;
                }
                P.accum_aux[pix] = a0;
                P.accum_aux[n + pix] = a1;
                P.accum_aux[2 * n + pix] = a2;
            }
        }
    }
    uint32_t v[5] = {cnt.rays, cnt.nodes, cnt.prims, cnt.insts, paths};
#pragma unroll
    for (int k = 0; k < 5; k++) {
        if (!STATS && (k == 1 || k == 2 || k == 3)) continue;
        uint32_t x = __reduce_add_sync(0xFFFFFFFFu, v[k]);
        if (lane == 0 && x) atomicAdd(P.counters + k, (unsigned long long)x);
    }
}
