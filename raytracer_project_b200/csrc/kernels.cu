// kernels.cu — sm_100a kernels of the path-tracing hot path and their launchers.
//   k_render      K6 megakernel: raygen (K1) + trace (K2) + shade (K3) per (pixel, sample), iterative
//                 ray_color (camera.hpp:928-1004) with the reference's termination rules, float4
//                 radiance sums in HBM (K5 accumulate), first-hit AOVs (K8, camera.hpp:464-488).
//   k_primary_ids K7: centre-of-pixel pinhole primary hit (camera.hpp:809-814 + :459).
//   k_resolve     K5: sums -> means, float4 or the reference's double RGB.
// Replaces render_rows (camera.hpp:430-554).  No tensor cores: nothing here is a dense contraction.
#include <algorithm>
#include <cstdlib>

#include "kernels.h"
#include "shade.cuh"

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

// ---------------------------------------------------------------------------------------------------------------
// Pieces shared by every render kernel (one copy each; all force-inlined).

// K1 raygen: get_ray (camera.hpp:784-794), sample_square (:817), defocus disk (:822) for (pixel, sample), from the Philox
// draw (pixel, s, bounce 0, stream 0) — so every kernel that regenerates a camera ray gets the same one.
ZB_DEV void camera_ray(const RenderParams& P, const Philox& ph, int px, int py, uint32_t pixel, uint32_t s, float3& o, float3& d) {
    const uint4 r0 = ph(pixel, s, 0u, 0u);
    d = P.cam.pix00_rel + (float(px) + u01(r0.x) - 0.5f) * P.cam.du + (float(py) + u01(r0.y) - 0.5f) * P.cam.dv;
    o = P.cam.center;
    if (P.cam.defocus) {
        float rr = sqrt_fast(u01(r0.z)), sn, cs;
        sincos_turn(u01(r0.w), &sn, &cs);
        const float3 off = (rr * cs) * P.cam.disk_u + (rr * sn) * P.cam.disk_v;
        o += off;
        d = d - off;
    }
}

// Path index g of a launch -> pixel and sample: tile, 8x4 sub-tile, pixel-major inside the sub-tile, sample (the order every
// kernel takes paths in).  Exact division of a 32-bit number by n: high word of x * ceil(2^64 / n) (error < 2^-32 of a
// quotient step); the two reciprocals cost one 64-bit division each per persistent thread instead of three 32-bit
// divisions per path.  Returns false for a pixel outside the image (ragged tile edge).
struct PathDecoder {
    uint32_t n_s, tiles_x;
    unsigned long long m_ns, m_tx;
    ZB_DEV void init(const RenderParams& P) {
        n_s = uint32_t(P.sample_end - P.sample_begin);
        tiles_x = uint32_t(P.tiles_x);
        m_ns = n_s > 1u ? ~0ull / n_s + 1ull : 0ull;
        m_tx = tiles_x > 1u ? ~0ull / tiles_x + 1ull : 0ull;
    }
    ZB_DEV bool decode(const RenderParams& P, unsigned long long g, unsigned long long total, int& px, int& py, uint32_t& s) const {
        // one division: q = g / n_s numbers the (sub-tile, pixel) pairs, 32 per sub-tile; in 32 bits whenever the
        // launch has fewer than 2^32 paths (a 64-bit division is a ~80-instruction routine)
        uint32_t q, sr;
        if (total <= 0xFFFFFFFFull) {
            const uint32_t g32 = uint32_t(g);
            q = n_s > 1u ? uint32_t(__umul64hi(g32, m_ns)) : g32;
            sr = g32 - q * n_s;
        } else {
            const unsigned long long q64 = g / n_s;
            q = uint32_t(q64);
            sr = uint32_t(g - q64 * n_s);
        }
        const uint32_t u = q >> 5, p = q & 31u;
        const int tile = int(u >> 2) * P.tile_ranks + P.tile_rank, w = int(u & 3u);
        const int ty = tiles_x > 1u ? int(__umul64hi(uint32_t(tile), m_tx)) : tile, tx = tile - ty * int(tiles_x);
        px = tx * ZB_TILE_W + (w % (ZB_TILE_W / 8)) * 8 + int(p & 7u);
        py = ty * ZB_TILE_H + (w / (ZB_TILE_W / 8)) * 4 + int(p >> 3);
        s = uint32_t(P.sample_begin) + sr;
        return px < P.cam.width && py < P.cam.height && ty < P.tiles_y;
    }
};

// State of one path between segments.
struct PathState {
    int seg;
    uint32_t pixel, s;
    float3 o, d, L, beta;
    float3 first_atten;  // LIGHT: attenuation of the primary scatter (camera.hpp:512, :515)
    int light_class;     // LIGHT: 1 reflection, 2 refraction, 0 neither
    ZB_DEV void start(uint32_t pixel_, uint32_t s_) {
        pixel = pixel_;
        s = s_;
        seg = 0;
        L = f3(0, 0, 0);
        beta = f3(1, 1, 1);
    }
};

// K3 shade: one segment of ray_color (camera.hpp:928-1004) after the closest-hit query of (st.o, st.d) returned `hit`;
// returns true when the path is finished (st.L then holds its radiance).  LIGHT = the reflection / refraction passes
// (camera.hpp:492-517): at segment 0 a second scatter of the primary hit decides the class of the sample; a primary miss
// or a sample of neither class ends at once.  `aov(sf, hit_something)` is called for the primary segment of beauty paths.
template <bool LIGHT, uint32_t F, class Aov>
ZB_DEV bool shade_segment(const RenderParams& P, const Philox& ph, const Hit& hit, PathState& st, Aov&& aov) {
    if (LIGHT && st.seg == 0) {
        st.light_class = 0;
        if (hit.item == 0xFFFFFFFFu) return true;
        Surface sf;
        make_surface<F>(P.sc, hit, st.o, st.d, sf);
        const uint4 rb = ph(st.pixel, st.s, 1u, 0u);
        float3 no, nd;
        st.first_atten = f3(1, 1, 1);
        if (!material_scatter<F>(P.sc, sf, st.d, f3(u01(rb.x), u01(rb.y), u01(rb.z)), st.first_atten, no, nd)) return true;
        const float3 ud = unit_vector(st.d), un = unit_vector(sf.n);
        const float3 mirror = fma3(-2.0f * dot(ud, un), un, ud);
        if (dot(unit_vector(nd), mirror) > 0.9f) st.light_class = 1;
        else if (dot(nd, sf.n) < 0.f) st.light_class = 2;
        if (!st.light_class) return true;  // the sample lands in neither buffer: skip its path
        st.o = no;
        st.d = nd;
        st.seg = 1;
        return P.max_depth <= 1;
    }
    if (hit.item == 0xFFFFFFFFu) {
        st.L += st.beta * ((F & ZB_F_ENV) ? background(P.sc, P.env, st.d) : P.env.background);
        if (!LIGHT && st.seg == 0) aov(static_cast<const Surface*>(nullptr), 0.f);
        return true;
    }
    Surface sf;
    make_surface<F>(P.sc, hit, st.o, st.d, sf);
    if (!LIGHT && st.seg == 0) aov(&sf, hit.t);
    st.L += st.beta * material_emitted<F>(P.sc, sf);
    const uint4 rb = ph(st.pixel, st.s, uint32_t(st.seg + 1), 0u);
    float3 atten = f3(1, 1, 1), no, nd;
    if (!material_scatter<F>(P.sc, sf, st.d, f3(u01(rb.x), u01(rb.y), u01(rb.z)), atten, no, nd)) return true;
    st.beta *= atten;
    st.o = no;
    st.d = nd;
    bool done = false;
    // ray_color's loop index is seg - 1; its `i > 10` rules start at seg 12 (camera.hpp:968-983)
    if (st.seg >= 12) {
        const float pr = fminf(fmaxf(max3(st.beta.x, st.beta.y, st.beta.z), 0.05f), 0.95f);
        if (dot(st.beta, st.beta) < 1e-8f || u01(rb.w) > pr) done = true;
        st.beta *= 1.0f / pr;
    }
    if (++st.seg >= P.max_depth) done = true;  // hard stop keeps the light gathered so far
    return done;
}

// radiance of a finished LIGHT path: luma clamp 0.2126 |c| <= 2 on the vector length (camera.hpp:500), then the primary attenuation
ZB_DEV float3 light_pass_value(const PathState& st) {
    float3 L = st.L;
    const float luma = 0.2126f * length(L);
    if (luma > 2.0f) L *= 2.0f / luma;
    return st.first_atten * L;
}

// first-hit AOV values (K8, camera.hpp:464-488, :518-526): albedo rgb, view-space normal xyz, 1 - clamp(t / z_max); sf == null: miss
template <uint32_t F, class Add>
ZB_DEV void aov_values(const RenderParams& P, const Surface* sf, float t, Add&& add) {
    if (!sf) {
        if (P.use_normal) { add(3, 0.5f); add(4, 0.5f); add(5, 1.0f); }
        return;
    }
    if (P.use_albedo) {
        const float3 a = material_get_albedo<F>(P.sc, *sf);
        add(0, a.x); add(1, a.y); add(2, a.z);
    }
    if (P.use_normal) {
        const float3 n = unit_vector(sf->n);
        add(3, (dot(n, P.cam.u) + 1.f) * 0.5f);
        add(4, (dot(n, P.cam.v) + 1.f) * 0.5f);
        add(5, (dot(n, P.cam.w) + 1.f) * 0.5f);
    }
    if (P.use_z) add(6, 1.0f - clamp01(t * P.inv_z_max));
}

// Work decomposition.  A block renders a 16x8 pixel tile (tiles are dealt to ranks round-robin for the
// tile-stripe partition); each of its four warps owns an 8x4 sub-tile and works through the
// 32 x n_samples paths of that sub-tile as a queue: a lane whose path has ended takes the next
// (pixel, sample) from the warp's counter, pixel-major, so at any time the 32 lanes trace samples of
// the same one or two pixels (coherent primary rays, similar bounce rays) and no lane idles while the
// queue is non-empty.  Lanes are never bound to a pixel, so per-pixel sums are kept in shared memory
// as 64-bit fixed point (2^-22 units): integer addition is associative, which makes the result
// independent of the order in which lanes finish — the frame is bit-reproducible and identical under
// any tile / sample partition of the work.
//
// LIGHT = true renders the reflection / refraction light passes of render_rows (camera.hpp:492-517) instead of
// the beauty pass: a second, independent scatter of the primary hit, the full ray_color estimator behind it
// (same termination rules), the luma clamp 0.2126 |c| <= 2, and the split by direction — "specular" when the
// scattered direction is within acos(0.9) of the mirror direction, "transmitted" when it crosses the surface.
// Reflection sums go to accum, refraction sums to plane 0 of accum_aux.
// F: scene features compiled in (ZB_F_*, include/zenith_b200.h); the light passes run the smallest covering set
template <bool STATS, bool LIGHT, uint32_t F = ZB_F_ALL>
__global__ void __launch_bounds__(ZB_BLOCK_THREADS, ZB_MIN_BLOCKS) k_render(const __grid_constant__ RenderParams P) {
    constexpr int WARPS = ZB_BLOCK_THREADS / 32;
    __shared__ unsigned long long s_sum[WARPS][32][3];
    __shared__ unsigned long long s_aux[WARPS][32][7];  // albedo rgb, normal xyz, z
    __shared__ int s_next[WARPS];

    const int tile = blockIdx.x * P.tile_ranks + P.tile_rank;
    const int tx = tile % P.tiles_x, ty = tile / P.tiles_x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int px0 = tx * ZB_TILE_W + (warp % (ZB_TILE_W / 8)) * 8, py0 = ty * ZB_TILE_H + (warp / (ZB_TILE_W / 8)) * 4;
    const bool aux_on = LIGHT || P.accum_aux != nullptr;

#pragma unroll
    for (int c = 0; c < 3; c++) s_sum[warp][lane][c] = 0ull;
    if (aux_on) {
#pragma unroll
        for (int c = 0; c < 7; c++) s_aux[warp][lane][c] = 0ull;
    }
    if (lane == 0) s_next[warp] = 0;
    __syncwarp();

    const int n_s = P.sample_end - P.sample_begin;
    const int total = (ty < P.tiles_y && P.max_depth > 0) ? 32 * n_s : 0;
    const Philox ph{P.pk[LIGHT ? 1 : 0]};  // the light passes draw their own stream
    Counters cnt = {0, 0, 0, 0};
    uint32_t paths = 0;

    // path state of this lane
    int p = 0;
    PathState st;
    st.start(0u, 0u);
    st.o = f3(0, 0, 0);
    st.d = f3(0, 0, 0);
    st.first_atten = f3(0, 0, 0);
    st.light_class = 0;

    // next path of this warp's queue whose pixel lies inside the image
    auto fetch = [&]() -> bool {
        for (;;) {
            int k = atomicAdd(&s_next[warp], 1);
            if (k >= total) return false;
            p = k / n_s;
            int px = px0 + (p & 7), py = py0 + (p >> 3);
            if (px >= P.cam.width || py >= P.cam.height) continue;
            st.start(uint32_t(py) * uint32_t(P.cam.width) + uint32_t(px), uint32_t(P.sample_begin + (k - p * n_s)));
            camera_ray(P, ph, px, py, st.pixel, st.s, st.o, st.d);
            paths++;
            return true;
        }
    };

    // Main loop: one iteration = one path segment of this lane (trace to completion, then shade).
    // Leaving the traversal early to refill finished lanes (Aila & Laine's ray-level persistence) was
    // measured and lost 7-40 % on this kernel: shading under a partial warp costs more than it saves.
    bool have = fetch();
    while (have) {
        Hit hit;
        {
            MediumRng mr{ph, st.pixel, st.s, uint32_t(st.seg + 1)};
            trace<STATS, F>(P.sc, st.o, st.d, 0.001f, INFINITY, mr, hit, cnt);
        }
        const bool done = shade_segment<LIGHT, F>(P, ph, hit, st, [&](const Surface* sf, float t) {
            if (aux_on && int(st.s) < P.aux_samples)
                aov_values<F>(P, sf, t, [&](int k, float v) { fix_add(&s_aux[warp][p][k], v); });
        });
        if (done) {
            if (LIGHT) {
                if (st.light_class) {
                    const float3 L = light_pass_value(st);
                    unsigned long long* dst = st.light_class == 1 ? s_sum[warp][p] : s_aux[warp][p];
                    fix_add(dst + 0, L.x);
                    fix_add(dst + 1, L.y);
                    fix_add(dst + 2, L.z);
                }
            } else {
                fix_add(&s_sum[warp][p][0], st.L.x);
                fix_add(&s_sum[warp][p][1], st.L.y);
                fix_add(&s_sum[warp][p][2], st.L.z);
            }
            have = fetch();
        }
    }
    __syncwarp();

    // ---- K5 accumulate: lane l writes pixel l of the sub-tile
    {
        const int px = px0 + (lane & 7), py = py0 + (lane >> 3);
        if (px < P.cam.width && py < P.cam.height && ty < P.tiles_y) {
            const size_t pix = size_t(py) * P.cam.width + px;
            float4 acc = make_float4(fix_get(s_sum[warp][lane][0]), fix_get(s_sum[warp][lane][1]),
                                     fix_get(s_sum[warp][lane][2]), float(n_s));
            if (P.accumulate) {
                float4 old = P.accum[pix];
                acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
            }
            P.accum[pix] = acc;
            if (LIGHT) {
                float4 a0 = make_float4(fix_get(s_aux[warp][lane][0]), fix_get(s_aux[warp][lane][1]), fix_get(s_aux[warp][lane][2]), float(n_s));
                if (P.accumulate) {
                    float4 b0 = P.accum_aux[pix];
                    a0.x += b0.x; a0.y += b0.y; a0.z += b0.z; a0.w += b0.w;
                }
                P.accum_aux[pix] = a0;
            } else if (aux_on) {
                const size_t n = size_t(P.cam.width) * P.cam.height;
                float z = fix_get(s_aux[warp][lane][6]);
                float4 a0 = make_float4(fix_get(s_aux[warp][lane][0]), fix_get(s_aux[warp][lane][1]), fix_get(s_aux[warp][lane][2]), 0.f);
                float4 a1 = make_float4(fix_get(s_aux[warp][lane][3]), fix_get(s_aux[warp][lane][4]), fix_get(s_aux[warp][lane][5]), 0.f);
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
    // one atomic per warp and counter
    uint32_t v[5] = {cnt.rays, cnt.nodes, cnt.prims, cnt.insts, paths};
#pragma unroll
    for (int k = 0; k < 5; k++) {
        if (!STATS && (k == 1 || k == 2 || k == 3)) continue;
        uint32_t x = __reduce_add_sync(0xFFFFFFFFu, v[k]);
        if (lane == 0 && x) atomicAdd(P.counters + k, (unsigned long long)x);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// k_render_pq — the beauty pass with a frame-wide path queue ("persistent queue" form of k_render<false,false>).
// k_render binds a warp to one 8x4 sub-tile: when the sub-tile's queue runs dry the lanes of the warp go idle one by
// one until the longest path ends, then the block ends and a new one starts.  That drain is paid once per sub-tile —
// under 1 % of the work at 64 samples per pixel, 13 % at the 10 spp of config C1 and 9 % on each GPU of an 8-GPU
// sample-slice frame (8 spp per GPU).  Here warps are persistent and take the frame's paths in batches of 64..256
// consecutive indices from one global counter (same order: tile, sub-tile, pixel-major, sample), so lanes only run
// dry at the end of the frame.  Lanes that need a path are served at the top of the segment loop, where the warp is
// converged, by ballot rank from the warp's batch (registers, no shared-memory atomics).  Per-pixel sums are 64-bit
// fixed-point RED.ADDs into a frame-sized buffer in L2 / HBM (same 2^-22 format as k_render's shared-memory sums, so
// the frame is bit-identical and order-independent); k_pq_resolve turns them into the float4 accumulation buffer.
#ifndef PQ_BATCH
#define PQ_BATCH 256u
#endif
// resident blocks per SM: 6 (80 registers) for scenes whose tree lives in L1 / L2, 7 (72 registers) when the tree is
// large enough for memory latency to dominate (measured: 1 M-triangle mesh +6 %, Weekend -1 %, quirks -6 %)
#define PQ_BIG_TREE_NODES 65536u

// MODE: PQ_BEAUTY, PQ_AUX (beauty + first-hit AOVs into 7 more sums per pixel: albedo rgb, normal xyz, z) or PQ_LIGHT
// (reflection sums into `fix`, refraction sums into `fix_aux`, 3 per pixel; see k_render<.., LIGHT>)
enum { PQ_BEAUTY = 0, PQ_AUX = 1, PQ_LIGHT = 2 };

ZB_DEV void gfix_add(unsigned long long* slot, float v) { atomicAdd(slot, __float2ull_rn(fmaxf(v, 0.f) * ZB_FIX_SCALE)); }

// F: scene features compiled in (ZB_F_*, include/zenith_b200.h); zbk_launch_render_pq picks the smallest instantiation that covers
// SceneDev::features and the environment of the render
template <int MIN_BLOCKS, int MODE, uint32_t F = ZB_F_ALL>
__global__ void __launch_bounds__(ZB_BLOCK_THREADS, MIN_BLOCKS) k_render_pq(const __grid_constant__ RenderParams P,
                                                                                  unsigned long long* __restrict__ fix,
                                                                                  unsigned long long* __restrict__ fix_aux,
                                                                                  unsigned long long* __restrict__ head,
                                                                                  unsigned long long total, unsigned batch) {
    constexpr bool AUX = MODE == PQ_AUX, LIGHT = MODE == PQ_LIGHT;
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const Philox ph{P.pk[LIGHT ? 1 : 0]};  // the light passes draw their own stream
    Counters cnt = {0, 0, 0, 0};
    uint32_t paths = 0;
    bool have = false, dry = false;
    PathState st;
    st.start(0u, 0u);
    st.o = f3(0, 0, 0);
    st.d = f3(0, 0, 1);
    st.first_atten = f3(0, 0, 0);
    st.light_class = 0;
    unsigned long long w_pos = 0, w_end = 0;  // the warp's batch of path indices (warp-uniform)
    PathDecoder dec;
    dec.init(P);
    for (;;) {
        // ---- lanes without a path take the next ones of the warp's batch; out-of-image pixels are skipped
        for (;;) {
            const unsigned need = __ballot_sync(0xFFFFFFFFu, !have && !dry);
            if (!need) break;
            const unsigned want = __popc(need);
            const unsigned long long avail = w_end - w_pos;
            unsigned long long base2 = 0;
            if (avail < want) {
                // fixed batches: shrinking them towards the end of the frame ("guided" scheduling) was measured and
                // cost more in the segment loop than the shorter tail returned (profiles/r01_pq.txt)
                if (lane == 0) base2 = atomicAdd(head, (unsigned long long)batch);
                base2 = __shfl_sync(0xFFFFFFFFu, base2, 0);
            }
            if (!have && !dry) {
                const unsigned r = __popc(need & lt);
                const unsigned long long g = r < avail ? w_pos + r : base2 + (r - avail);
                if (g >= total) {
                    dry = true;
                } else {
                    int px, py;
                    uint32_t s;
                    if (dec.decode(P, g, total, px, py, s)) {
                        st.start(uint32_t(py) * uint32_t(P.cam.width) + uint32_t(px), s);
                        camera_ray(P, ph, px, py, st.pixel, st.s, st.o, st.d);
                        paths++;
                        have = true;
                    }
                }
            }
            if (avail < want) {
                w_pos = base2 + (want - avail);
                w_end = base2 + batch;
            } else {
                w_pos += want;
            }
        }
#ifdef ZB_EXPERIMENTS
        // VERDICT r1 item 6 (experiment, tuning "pq_redeal"): coherence between bounces.  1 = every segment of the block starts
        // at a block-wide barrier (what any re-dealing inside a persistent block needs); 2 = the block's 128 rays are also
        // re-dealt to the lanes by the octant of their direction (counting sort through shared memory), so that a warp
        // traces rays of one or two octants.  Order-free fixed-point sums: the frame is bit-identical either way.
        if (MODE == PQ_BEAUTY && P.exp_flags) {
            if (!__syncthreads_or(have ? 1 : 0)) break;
            if (P.exp_flags & 2) {
                __shared__ int s_cnt[ZB_BLOCK_THREADS / 32][9];
                __shared__ float s_state[ZB_BLOCK_THREADS][16];
                const int warp = threadIdx.x >> 5;
                const uint32_t key = have ? ray_octant(st.d) : 8u;
                if (threadIdx.x < 36) s_cnt[threadIdx.x / 9][threadIdx.x % 9] = 0;
                __syncthreads();
                const unsigned peers = __match_any_sync(0xFFFFFFFFu, key);
                const int rank = __popc(peers & lt);
                if (rank == 0) s_cnt[warp][key] = __popc(peers);
                __syncthreads();
                int at = rank;
                for (uint32_t b = 0; b < key; b++)
                    for (int w2 = 0; w2 < ZB_BLOCK_THREADS / 32; w2++) at += s_cnt[w2][b];
                for (int w2 = 0; w2 < warp; w2++) at += s_cnt[w2][key];
                float* q = s_state[at];
                q[0] = __uint_as_float(st.pixel); q[1] = __uint_as_float(st.s); q[2] = __int_as_float(st.seg); q[3] = have ? 1.f : 0.f;
                q[4] = st.o.x; q[5] = st.o.y; q[6] = st.o.z; q[7] = st.d.x; q[8] = st.d.y; q[9] = st.d.z;
                q[10] = st.L.x; q[11] = st.L.y; q[12] = st.L.z; q[13] = st.beta.x; q[14] = st.beta.y; q[15] = st.beta.z;
                __syncthreads();
                const float* r = s_state[threadIdx.x];
                st.pixel = __float_as_uint(r[0]); st.s = __float_as_uint(r[1]); st.seg = __float_as_int(r[2]); have = r[3] != 0.f;
                st.o = f3(r[4], r[5], r[6]); st.d = f3(r[7], r[8], r[9]);
                st.L = f3(r[10], r[11], r[12]); st.beta = f3(r[13], r[14], r[15]);
            }
        } else
#endif
        if (!__any_sync(0xFFFFFFFFu, have)) break;
        if (have) {
            Hit hit;
            {
                MediumRng mr{ph, st.pixel, st.s, uint32_t(st.seg + 1)};
                trace<false, F>(P.sc, st.o, st.d, 0.001f, INFINITY, mr, hit, cnt);
            }
            const bool done = shade_segment<LIGHT, F>(P, ph, hit, st, [&](const Surface* sf, float t) {
                if (AUX && int(st.s) < P.aux_samples) {
                    unsigned long long* a = fix_aux + 7 * size_t(st.pixel);
                    aov_values<F>(P, sf, t, [&](int k, float v) { gfix_add(a + k, v); });
                }
            });
            if (done) {  // K5 accumulate
                if (LIGHT) {
                    if (st.light_class) {
                        const float3 L = light_pass_value(st);
                        unsigned long long* f = (st.light_class == 1 ? fix : fix_aux) + 3 * size_t(st.pixel);
                        gfix_add(f + 0, L.x);
                        gfix_add(f + 1, L.y);
                        gfix_add(f + 2, L.z);
                    }
                } else {
                    unsigned long long* f = fix + 3 * size_t(st.pixel);
                    gfix_add(f + 0, st.L.x);
                    gfix_add(f + 1, st.L.y);
                    gfix_add(f + 2, st.L.z);
                }
                have = false;
            }
        }
    }
    uint32_t v[2] = {cnt.rays, paths};
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const uint32_t x = __reduce_add_sync(0xFFFFFFFFu, v[k]);
        if (lane == 0 && x) atomicAdd(P.counters + (k ? 4 : 0), (unsigned long long)x);
    }
}

#ifdef ZB_EXPERIMENTS  // measured 1..22 % slower than k_render_pq at every threshold (profiles/r01_traversal_experiments.txt)
// k_render_pqr — k_render_pq's beauty pass with RESUMABLE traversal: the traversal state of a lane (RayWalk + its stack) lives
// across iterations of the warp's main loop, and the warp leaves the traversal loop as soon as fewer than `thresh` of its lanes
// still hold an unfinished ray while others wait with a finished one.  The finished lanes are shaded and get their next ray (or
// the next path of the batch) while the others keep their state, so the node loop runs with more lanes than when every
// segment ends with the slowest ray of the warp.  Same paths, same arithmetic per path, order-free fixed-point sums: the frame
// is bit-identical to k_render_pq's.
template <int MIN_BLOCKS, uint32_t F>
__global__ void __launch_bounds__(ZB_BLOCK_THREADS, MIN_BLOCKS) k_render_pqr(const __grid_constant__ RenderParams P,
                                                                                   unsigned long long* __restrict__ fix,
                                                                                   unsigned long long* __restrict__ head,
                                                                                   unsigned long long total, unsigned batch, unsigned thresh) {
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const Philox ph{P.pk[0]};
    Counters cnt = {0, 0, 0, 0};
    uint32_t paths = 0;
    bool have = false, dry = false;
    PathState st;
    st.start(0u, 0u);
    st.o = f3(0, 0, 0);
    st.d = f3(0, 0, 1);
    st.first_atten = f3(0, 0, 0);
    st.light_class = 0;
    unsigned long long w_pos = 0, w_end = 0;
    PathDecoder dec;
    dec.init(P);
    int stack[ZB_STACK];
    RayWalk w;
    w.begin(P.sc, st.o, st.d, INFINITY, stack);
    w.cur = ZB_TRAVERSAL_DONE;
    for (;;) {
        // ---- lanes without a path take the next ones of the warp's batch (as in k_render_pq)
        for (;;) {
            const unsigned need = __ballot_sync(0xFFFFFFFFu, !have && !dry);
            if (!need) break;
            const unsigned want = __popc(need);
            const unsigned long long avail = w_end - w_pos;
            unsigned long long base2 = 0;
            if (avail < want) {
                if (lane == 0) base2 = atomicAdd(head, (unsigned long long)batch);
                base2 = __shfl_sync(0xFFFFFFFFu, base2, 0);
            }
            if (!have && !dry) {
                const unsigned r = __popc(need & lt);
                const unsigned long long g = r < avail ? w_pos + r : base2 + (r - avail);
                if (g >= total) {
                    dry = true;
                } else {
                    int px, py;
                    uint32_t s;
                    if (dec.decode(P, g, total, px, py, s)) {
                        st.start(uint32_t(py) * uint32_t(P.cam.width) + uint32_t(px), s);
                        camera_ray(P, ph, px, py, st.pixel, st.s, st.o, st.d);
                        paths++;
                        have = true;
                        cnt.rays++;
                        w.begin(P.sc, st.o, st.d, INFINITY, stack);
                    }
                }
            }
            if (avail < want) {
                w_pos = base2 + (want - avail);
                w_end = base2 + batch;
            } else {
                w_pos += want;
            }
        }
        const unsigned holders = __ballot_sync(0xFFFFFFFFu, have);
        if (!holders) break;
        // ---- K2 traverse: while-while rounds until too few lanes still hold an unfinished ray
        for (;;) {
            const bool trav = have && w.cur != ZB_TRAVERSAL_DONE;
            const unsigned m = __ballot_sync(0xFFFFFFFFu, trav);
            if (!m || (unsigned(__popc(m)) < thresh && m != holders)) break;
            if (trav) {
                while (w.cur >= 0) w.node_step<false>(P.sc, 0.001f, stack, cnt);
                if (w.cur != ZB_TRAVERSAL_DONE) {
                    const MediumRng mr{ph, st.pixel, st.s, uint32_t(st.seg + 1)};
                    w.leaf_step<false, F>(P.sc, st.o, st.d, 0.001f, mr, stack, cnt);
                }
            }
        }
        // ---- K3 shade the lanes whose ray is finished, as in k_render_pq
        if (have && w.cur == ZB_TRAVERSAL_DONE) {
            const Hit hit = w.hit;
            const bool done = shade_segment<false, F>(P, ph, hit, st, [](const Surface*, float) {});
            if (done) {  // K5 accumulate
                unsigned long long* f = fix + 3 * size_t(st.pixel);
                gfix_add(f + 0, st.L.x);
                gfix_add(f + 1, st.L.y);
                gfix_add(f + 2, st.L.z);
                have = false;
            } else {
                cnt.rays++;
                w.begin(P.sc, st.o, st.d, INFINITY, stack);
            }
        }
    }
    uint32_t v[2] = {cnt.rays, paths};
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const uint32_t x = __reduce_add_sync(0xFFFFFFFFu, v[k]);
        if (lane == 0 && x) atomicAdd(P.counters + (k ? 4 : 0), (unsigned long long)x);
    }
}

#endif  // ZB_EXPERIMENTS

// First-hit AOVs (K8, camera.hpp:464-488, :518-526) as a pass of their own beside k_render_pq: one thread per pixel walks
// the pixel's first aux_samples samples, regenerates the camera ray of (pixel, sample) — the same Philox draws as the
// beauty pass, so the same primary hit — and sums albedo, view-space normal and z in 64-bit fixed point in registers
// (the same integers k_render adds up in shared memory, so the planes are bit-identical to its output).  Primary rays
// of an 8x4 sub-tile at the same sample index are coherent: 133 M of them cost ~3 ms on the headline frame.
template <uint32_t F>
__global__ void __launch_bounds__(ZB_BLOCK_THREADS, 6) k_aov_pass(const __grid_constant__ RenderParams P,
                                                                  unsigned long long* __restrict__ fix_aux) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * P.tile_ranks + P.tile_rank;
    const int tx = tile % P.tiles_x, ty = tile / P.tiles_x;
    const int px = tx * ZB_TILE_W + (warp % (ZB_TILE_W / 8)) * 8 + (lane & 7), py = ty * ZB_TILE_H + (warp / (ZB_TILE_W / 8)) * 4 + (lane >> 3);
    if (px >= P.cam.width || py >= P.cam.height || ty >= P.tiles_y) return;
    const uint32_t pixel = uint32_t(py) * uint32_t(P.cam.width) + uint32_t(px);
    const Philox ph{P.pk[0]};
    Counters cnt = {0, 0, 0, 0};
    unsigned long long acc[7] = {0, 0, 0, 0, 0, 0, 0};
    auto add = [&](int k, float v) { acc[k] += __float2ull_rn(fmaxf(v, 0.f) * ZB_FIX_SCALE); };
    const int s_end = min(P.sample_end, P.aux_samples);
    for (int s = P.sample_begin; s < s_end; s++) {
        float3 o, d;
        camera_ray(P, ph, px, py, pixel, uint32_t(s), o, d);
        Hit hit;
        const MediumRng mr{ph, pixel, uint32_t(s), 1u};
        trace<false, F>(P.sc, o, d, 0.001f, INFINITY, mr, hit, cnt);
        if (hit.item == 0xFFFFFFFFu) {
            aov_values<F>(P, nullptr, 0.f, add);
            continue;
        }
        Surface sf;
        make_surface<F>(P.sc, hit, o, d, sf);
        aov_values<F>(P, &sf, hit.t, add);
    }
    const size_t n = size_t(P.cam.width) * P.cam.height, i = pixel;
    if (fix_aux) {
        // fused-resolve / multi-GPU form (group.cu): the integer sums themselves, 7 words per pixel, added to what earlier
        // sample batches left there (one thread per pixel, launches are stream-ordered: no atomics needed)
#pragma unroll
        for (int k = 0; k < 7; k++) fix_aux[7 * i + k] += acc[k];
        return;
    }
    const float z = fix_get(acc[6]);
    float4 a0 = make_float4(fix_get(acc[0]), fix_get(acc[1]), fix_get(acc[2]), 0.f);
    float4 a1 = make_float4(fix_get(acc[3]), fix_get(acc[4]), fix_get(acc[5]), 0.f);
    float4 a2 = make_float4(z, z, z, 0.f);
    if (P.accumulate) {
        const float4 b0 = P.accum_aux[i], b1 = P.accum_aux[n + i], b2 = P.accum_aux[2 * n + i];
        a0.x += b0.x; a0.y += b0.y; a0.z += b0.z;
        a1.x += b1.x; a1.y += b1.y; a1.z += b1.z;
        a2.x += b2.x; a2.y += b2.y; a2.z += b2.z;
    }
    P.accum_aux[i] = a0;
    P.accum_aux[n + i] = a1;
    P.accum_aux[2 * n + i] = a2;
}

// fixed-point frame sums of k_render_pq, rows [row0, row1) -> the float4 accumulation buffer (xyz sums, w = samples) for
// this rank's tiles, or — when `mean_out` is given — straight to scaled means (xyz, w = 1) without touching P.accum
__global__ void __launch_bounds__(256) k_pq_resolve(const __grid_constant__ RenderParams P, const unsigned long long* __restrict__ fix,
                                                    const unsigned long long* __restrict__ fix_aux, int mode, int row0, int row1,
                                                    float scale, float4* __restrict__ mean_out) {
    const int px = blockIdx.x * 64 + (threadIdx.x & 63), py = row0 + blockIdx.y * 4 + (threadIdx.x >> 6);
    if (px >= P.cam.width || py >= row1) return;
    if (P.tile_ranks > 1) {  // tile stripes: the other ranks' pixels stay untouched
        const int tile = (py / ZB_TILE_H) * P.tiles_x + px / ZB_TILE_W;
        if (tile % P.tile_ranks != P.tile_rank) return;
    }
    const size_t i = size_t(py) * P.cam.width + px;
    float4 a = make_float4(fix_get(fix[3 * i]), fix_get(fix[3 * i + 1]), fix_get(fix[3 * i + 2]), float(P.sample_end - P.sample_begin));
    if (mean_out) {
        mean_out[i] = make_float4(a.x * scale, a.y * scale, a.z * scale, 1.0f);  // as k_resolve
        return;
    }
    if (P.accumulate) {
        const float4 old = P.accum[i];
        a.x += old.x; a.y += old.y; a.z += old.z; a.w += old.w;
    }
    P.accum[i] = a;
    if (!fix_aux) return;
    const size_t n = size_t(P.cam.width) * P.cam.height;
    if (mode == PQ_LIGHT) {  // refraction sums -> plane 0 of accum_aux, like k_render<.., LIGHT>
        float4 r = make_float4(fix_get(fix_aux[3 * i]), fix_get(fix_aux[3 * i + 1]), fix_get(fix_aux[3 * i + 2]), float(P.sample_end - P.sample_begin));
        if (P.accumulate) {
            const float4 old = P.accum_aux[i];
            r.x += old.x; r.y += old.y; r.z += old.z; r.w += old.w;
        }
        P.accum_aux[i] = r;
    } else {  // albedo, normal, z planes
        const unsigned long long* q = fix_aux + 7 * i;
        const float z = fix_get(q[6]);
        float4 a0 = make_float4(fix_get(q[0]), fix_get(q[1]), fix_get(q[2]), 0.f);
        float4 a1 = make_float4(fix_get(q[3]), fix_get(q[4]), fix_get(q[5]), 0.f);
        float4 a2 = make_float4(z, z, z, 0.f);
        if (P.accumulate) {
            const float4 b0 = P.accum_aux[i], b1 = P.accum_aux[n + i], b2 = P.accum_aux[2 * n + i];
            a0.x += b0.x; a0.y += b0.y; a0.z += b0.z;
            a1.x += b1.x; a1.y += b1.y; a1.z += b1.z;
            a2.x += b2.x; a2.y += b2.y; a2.z += b2.z;
        }
        P.accum_aux[i] = a0;
        P.accum_aux[n + i] = a1;
        P.accum_aux[2 * n + i] = a2;
    }
}

__global__ void __launch_bounds__(128) k_primary_ids(const __grid_constant__ RenderParams P, uint32_t* leaf_id,
                                                     uint32_t* prim_id, float* t_out) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.cam.width * P.cam.height) return;
    const int px = idx % P.cam.width, py = idx / P.cam.width;
    // get_center_ray (camera.hpp:809-814) in double, then narrowed once
    float3 d = f3(float(P.cam.pix00_rel_d[0] + px * P.cam.du_d[0] + py * P.cam.dv_d[0]),
                  float(P.cam.pix00_rel_d[1] + px * P.cam.du_d[1] + py * P.cam.dv_d[1]),
                  float(P.cam.pix00_rel_d[2] + px * P.cam.du_d[2] + py * P.cam.dv_d[2]));
    Counters cnt = {0, 0, 0, 0};
    MediumRng mr{Philox{P.pk[0]}, uint32_t(idx), 0u, 0u};
    Hit hit;
    uint32_t lid = 0xFFFFFFFFu, pid = 0;
    float t = INFINITY;
    if (trace<false>(P.sc, P.cam.center, d, 0.001f, INFINITY, mr, hit, cnt)) {
        Surface sf;
        make_surface(P.sc, hit, P.cam.center, d, sf);
        lid = sf.leaf_id;
        pid = sf.prim_id;
        t = hit.t;
    }
    if (leaf_id) leaf_id[idx] = lid;
    if (prim_id) prim_id[idx] = pid;
    if (t_out) t_out[idx] = t;
}

// Per-ray probe: closest hit + the reference's hit_record + albedo / emitted / background (zb_probe_rays).
__global__ void __launch_bounds__(128) k_probe(const __grid_constant__ RenderParams P, size_t n, const float* __restrict__ org,
                                               const float* __restrict__ dir, zb_probe* __restrict__ out) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float3 o = f3(org[3 * i], org[3 * i + 1], org[3 * i + 2]), d = f3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]);
    Counters cnt = {0, 0, 0, 0};
    MediumRng mr{Philox{P.pk[0]}, uint32_t(i), 0u, 0u};
    Hit hit;
    zb_probe r;
    memset(&r, 0, sizeof(r));
    if (trace<false>(P.sc, o, d, 0.001f, INFINITY, mr, hit, cnt)) {
        Surface sf;
        make_surface(P.sc, hit, o, d, sf, true);
        r.t = hit.t;
        r.leaf_id = sf.leaf_id;
        r.prim_id = sf.prim_id;
        r.front_face = sf.front_face ? 1u : 0u;
        r.p[0] = sf.p.x; r.p[1] = sf.p.y; r.p[2] = sf.p.z;
        r.normal[0] = sf.n.x; r.normal[1] = sf.n.y; r.normal[2] = sf.n.z;
        r.u = sf.u;
        r.v = sf.v;
        float3 a = material_get_albedo(P.sc, sf), e = material_emitted(P.sc, sf);
        r.albedo[0] = a.x; r.albedo[1] = a.y; r.albedo[2] = a.z;
        r.emitted[0] = e.x; r.emitted[1] = e.y; r.emitted[2] = e.z;
    } else {
        r.t = INFINITY;
        r.leaf_id = 0xFFFFFFFFu;
        float3 b = background(P.sc, P.env, d);
        r.albedo[0] = b.x; r.albedo[1] = b.y; r.albedo[2] = b.z;
    }
    out[i] = r;
}

// Known-answer probe (zb_kat_eval): single device functions of the shading code on explicit inputs, one thread per case.
__global__ void __launch_bounds__(128) k_kat(const __grid_constant__ RenderParams P, size_t n, const zb_kat_case* __restrict__ in,
                                             zb_kat_result* __restrict__ out) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const zb_kat_case c = in[i];
    const float* a = c.a;
    float3 r = f3(0, 0, 0);
    float w = 0.f;
    switch (c.op) {
        case ZB_KAT_REFRACT: {  // a: unit direction (3), normal facing it (3), eta ratio
            const float3 ud = f3(a[0], a[1], a[2]), wn = f3(a[3], a[4], a[5]);
            r = refract_dir(ud, wn, a[6], fminf(dot(-ud, wn), 1.0f));
        } break;
        case ZB_KAT_REFLECT: r = reflect_dir(f3(a[0], a[1], a[2]), f3(a[3], a[4], a[5])); break;
        case ZB_KAT_REFLECTANCE: r.x = schlick_reflectance(a[0], a[1]); break;
        case ZB_KAT_TEXTURE: r = texture_value(P.sc, c.index, a[0], a[1], f3(a[2], a[3], a[4])); break;
        case ZB_KAT_BUMP: {  // index: material; a: u, v, normal (3), tangent (3), bitangent (3), p (3)
            Surface sf;
            sf.u = a[0];
            sf.v = a[1];
            sf.n = f3(a[2], a[3], a[4]);
            sf.tangent = f3(a[5], a[6], a[7]);
            sf.bitangent = f3(a[8], a[9], a[10]);
            sf.p = f3(a[11], a[12], a[13]);
            r = bumped_normal(P.sc, load_mat0(P.sc, c.index), sf);
        } break;
        case ZB_KAT_SPHERE_UV: {  // a: unit outward normal; the uv of sphere_surface for a unit sphere at the origin, seen from outside
            Surface sf;
            const float3 on = f3(a[0], a[1], a[2]);
            sphere_surface(make_float4(0.f, 0.f, 0.f, 1.f), on, -on, sf, true);
            r.x = sf.u;
            r.y = sf.v;
        } break;
        case ZB_KAT_BACKGROUND: r = background(P.sc, P.env, f3(a[0], a[1], a[2])); break;
        default: w = -1.f; break;
    }
    out[i].r[0] = r.x;
    out[i].r[1] = r.y;
    out[i].r[2] = r.z;
    out[i].r[3] = w;
}

cudaError_t zbk_launch_kat(const RenderParams& P, size_t n, const zb_kat_case* in, zb_kat_result* out, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    k_kat<<<unsigned((n + 127) / 128), 128, 0, stream>>>(P, n, in, out);
    return cudaGetLastError();
}

cudaError_t zbk_launch_probe(const RenderParams& P, size_t n, const float* org, const float* dir, zb_probe* out,
                             cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    k_probe<<<unsigned((n + 127) / 128), 128, 0, stream>>>(P, n, org, dir, out);
    return cudaGetLastError();
}

// K5 resolve: mean = sum * scale.  Writes float4 and/or the reference's 3-double layout.
__global__ void __launch_bounds__(256) k_resolve(const float4* __restrict__ accum, size_t n, float scale,
                                                 float4* __restrict__ out_f4, double* __restrict__ out_f64) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = accum[i];
    float3 m = f3(a.x * scale, a.y * scale, a.z * scale);
    if (out_f4) out_f4[i] = make_float4(m.x, m.y, m.z, 1.0f);
    if (out_f64) {
        out_f64[3 * i + 0] = double(m.x);
        out_f64[3 * i + 1] = double(m.y);
        out_f64[3 * i + 2] = double(m.z);
    }
}

// Feature sets the render kernels are compiled for (ZB_F_*, include/zenith_b200.h), smallest first:
//   PLAIN   inline spheres, untextured materials, solid-colour environment (Weekend)
//   SPHTE   inline spheres with textures and / or a sky or HDR environment
//   BOXES   + loose triangles, cubes, instances of them, media (Cornell)
//   MESH0   instances, meshes, triangles (mesh configs), MESH: + textures
//   NOMESH  everything but meshes and triangles (exterior: HDR map, bump maps, instanced cubes / spheres, media)
//   NOMEDIA everything but participating media
// ZB_NO_LEAN=1 forces the general kernels (A/B runs, equality tests).
namespace {
constexpr uint32_t F_PLAIN = 0, F_BOXES = ZB_F_INST | ZB_F_CUBE | ZB_F_TRI | ZB_F_MEDIA, F_MESH0 = ZB_F_INST | ZB_F_MESH | ZB_F_TRI,
                   F_MESH = F_MESH0 | ZB_F_TEX, F_NOMESH = ZB_F_ALL & ~(ZB_F_MESH | ZB_F_TRI), F_SPHTE = ZB_F_TEX | ZB_F_ENV,
                   F_NOMEDIA = ZB_F_ALL & ~ZB_F_MEDIA;
uint32_t pick_feature_set(const RenderParams& P, const ZbTuning& tn) {
    if (tn.no_lean) return ZB_F_ALL;
    const uint32_t need = (P.sc.features & ZB_F_ALL) | (P.env.mode == ZB_ENV_SOLID_COLOR ? 0u : uint32_t(ZB_F_ENV));
    for (uint32_t f : {F_PLAIN, F_SPHTE, F_BOXES, F_MESH0, F_MESH, F_NOMESH, F_NOMEDIA})
        if ((need & ~f) == 0) return f;
    return ZB_F_ALL;
}
}  // namespace
#define ZB_FOR_FEATURE_SET(feat, CALL)                                      \
    switch (feat) {                                                         \
        case F_PLAIN: { constexpr uint32_t F = F_PLAIN; CALL; } break;      \
        case F_BOXES: { constexpr uint32_t F = F_BOXES; CALL; } break;      \
        case F_MESH0: { constexpr uint32_t F = F_MESH0; CALL; } break;      \
        case F_MESH: { constexpr uint32_t F = F_MESH; CALL; } break;        \
        case F_NOMESH: { constexpr uint32_t F = F_NOMESH; CALL; } break;    \
        case F_SPHTE: { constexpr uint32_t F = F_SPHTE; CALL; } break;      \
        case F_NOMEDIA: { constexpr uint32_t F = F_NOMEDIA; CALL; } break;  \
        default: { constexpr uint32_t F = ZB_F_ALL; CALL; } break;          \
    }

ZbTuning zbk_tuning_from_env() {
    ZbTuning t;
    auto num = [](const char* name, long long dflt) { const char* e = getenv(name); return e ? atoll(e) : dflt; };
    t.pq = int(num("ZB_PQ", t.pq));
    t.no_lean = num("ZB_NO_LEAN", 0) != 0;
    t.pq_blocks = int(num("ZB_PQ_BLOCKS", 0));
    t.pq_resume = int(num("ZB_PQ_RESUME", 0));
    t.wavefront = num("ZB_WAVEFRONT", 0) > 0;
    t.wf_slots = num("ZB_WF_SLOTS", 0);
    t.gather_p2p = int(num("ZB_GATHER_P2P", 1));
    t.gpu_builder = int(num("ZB_GPU_BUILDER", 1));
    t.build_timing = int(num("ZB_BUILD_TIMING", 0));
    t.group_timing = int(num("ZB_GROUP_TIMING", 0));
    t.build_check = int(num("ZB_BUILD_CHECK", 0));
    t.blas_isect_pct = int(num("ZB_BLAS_ISECT_PCT", 100));
    t.blas_leaf = int(num("ZB_DEVICE_BLAS_LEAF", 4));
    t.wide = int(num("ZB_WIDE", 1));
    t.pq_redeal = int(num("ZB_PQ_REDEAL", 0));
    return t;
}

cudaError_t zbk_launch_render(const RenderParams& P, int n_tiles_owned, bool stats, const ZbTuning& tn, cudaStream_t stream) {
    if (n_tiles_owned <= 0) return cudaSuccess;
    // A shared-memory ray-pool variant of this kernel was measured 20-40 % slower on the headline frame
    // (profiles/r01_pool_vs_queue.txt) and removed.
    if (P.light_pass) {
        const uint32_t feat = pick_feature_set(P, tn);
        ZB_FOR_FEATURE_SET(feat, (k_render<false, true, F><<<n_tiles_owned, ZB_BLOCK_THREADS, 0, stream>>>(P)));
    }
    else if (stats) k_render<true, false><<<n_tiles_owned, ZB_BLOCK_THREADS, 0, stream>>>(P);
    else k_render<false, false><<<n_tiles_owned, ZB_BLOCK_THREADS, 0, stream>>>(P);
    return cudaGetLastError();
}

static thread_local uint32_t g_last_pq_pick = 0;
uint32_t zbk_last_pq_pick() { return g_last_pq_pick; }

size_t zbk_pq_words(size_t pixels, int mode) { return pixels * (mode == PQ_AUX ? 10 : mode == PQ_LIGHT ? 6 : 3); }

cudaError_t zbk_pq_clear(unsigned long long* fix, size_t pixels, int mode, unsigned long long* heads, int n_heads, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(fix, 0, zbk_pq_words(pixels, mode) * sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    return cudaMemsetAsync(heads, 0, size_t(n_heads) * sizeof(unsigned long long), stream);
}

// beauty instantiation for feature set F at 6 / 7 / 8 resident blocks per SM (80 / 72 / 64 registers; 5, 9 and 10 were
// measured too and never won, profiles/r01_feature_sets.txt); the general kernel is built for 6 and 7 only
template <uint32_t F>
static void pq_launch_beauty(int per_sm, unsigned blocks, cudaStream_t stream, const RenderParams& P, unsigned long long* fix,
                             unsigned long long* head, unsigned long long total, unsigned batch) {
    if (per_sm <= 6) k_render_pq<6, PQ_BEAUTY, F><<<blocks, ZB_BLOCK_THREADS, 0, stream>>>(P, fix, nullptr, head, total, batch);
    else if (per_sm == 7 || F == ZB_F_ALL) k_render_pq<7, PQ_BEAUTY, F><<<blocks, ZB_BLOCK_THREADS, 0, stream>>>(P, fix, nullptr, head, total, batch);
    else if constexpr (F != ZB_F_ALL) k_render_pq<8, PQ_BEAUTY, F><<<blocks, ZB_BLOCK_THREADS, 0, stream>>>(P, fix, nullptr, head, total, batch);
}

// the queue kernel alone: the sums and `*head` must be zero (zbk_pq_clear); tiles P.tile_rank + i * P.tile_ranks,
// i < n_tiles.  The aux sums follow the 3 * pixels beauty sums in `fix`.
cudaError_t zbk_launch_render_pq(const RenderParams& P, int n_tiles_owned, int mode, unsigned long long* fix, unsigned long long* head,
                                 int sm_count, const ZbTuning& tn, cudaStream_t stream) {
    if (n_tiles_owned <= 0) return cudaSuccess;
    const size_t pixels = size_t(P.cam.width) * P.cam.height;
    unsigned long long* fix_aux = mode == PQ_BEAUTY ? nullptr : fix + 3 * pixels;
    const unsigned long long total = (unsigned long long)n_tiles_owned * 4ull * 32ull * (unsigned long long)(P.sample_end - P.sample_begin);
    const unsigned long long warps_needed = (total + 31) / 32;
    // beauty renders run the smallest compiled feature set that covers the scene and its environment (pick_feature_set);
    // ZB_PQ_BLOCKS=6..8 overrides the resident blocks per SM of the specialised instantiations
    const int env_blocks = tn.pq_blocks;
    const uint32_t feat = mode == PQ_BEAUTY ? pick_feature_set(P, tn) : uint32_t(ZB_F_ALL);
    const bool big_tree = P.sc.n_nodes >= PQ_BIG_TREE_NODES;
    // resident blocks per SM, measured per feature set at 1080p x 16 spp with 5 / 6 / 7 / 8 blocks (profiles/r01_feature_sets.txt):
    // Cornell 15.31 / 14.42 / 14.14 / 14.07 ms, exterior 18.74 / 18.17 / 17.42 / 17.83, mesh 17.95 / 16.86 / 16.67 / 16.04
    int per_sm = feat == ZB_F_ALL ? (big_tree ? 7 : 6) : (feat == F_NOMESH || feat == F_NOMEDIA) ? 7 : 8;
    if (feat != ZB_F_ALL && env_blocks >= 6 && env_blocks <= 8) per_sm = env_blocks;
    const unsigned blocks = unsigned(std::min<unsigned long long>((unsigned long long)sm_count * per_sm, (warps_needed + 3) / 4));
    // paths per grab: 256 when every warp will grab many times; short launches (few samples per pixel, small images,
    // one GPU's slice of a multi-GPU frame) end better balanced with smaller batches.  Measured, 1080p Weekend:
    // 64 spp 35.10 / 34.66 / 34.42 ms with 64 / 128 / 256, 8 spp 4.44 / 4.46 / 4.52, 1200x675 x 10 spp 2.23 / 2.25 / 2.36
    const unsigned long long per_warp = total / (4ull * blocks);
    const unsigned batch = per_warp >= 64ull * PQ_BATCH ? PQ_BATCH : per_warp >= 64ull * (PQ_BATCH / 2) ? PQ_BATCH / 2 : PQ_BATCH / 4;
    g_last_pq_pick = (uint32_t(feat == ZB_F_ALL ? std::min(std::max(per_sm, 6), 7) : per_sm) << 16) | (uint32_t(mode) << 8) | feat;
#define PQ_LAUNCH(B, M) k_render_pq<B, M><<<blocks, ZB_BLOCK_THREADS, 0, stream>>>(P, fix, fix_aux, head, total, batch)
#ifdef ZB_EXPERIMENTS
    const int resume = tn.pq_resume;  // experiment: resumable traversal (k_render_pqr)
    if (mode == PQ_BEAUTY && feat == F_PLAIN && resume > 0) {
        if (per_sm == 7) k_render_pqr<7, F_PLAIN><<<blocks, ZB_BLOCK_THREADS, 0, stream>>>(P, fix, head, total, batch, unsigned(resume));
        else if (per_sm == 6) k_render_pqr<6, F_PLAIN><<<blocks, ZB_BLOCK_THREADS, 0, stream>>>(P, fix, head, total, batch, unsigned(resume));
        else k_render_pqr<8, F_PLAIN><<<blocks, ZB_BLOCK_THREADS, 0, stream>>>(P, fix, head, total, batch, unsigned(resume));
    }
    else
#endif
    if (mode == PQ_AUX) { if (per_sm == 7) PQ_LAUNCH(7, PQ_AUX); else PQ_LAUNCH(6, PQ_AUX); }
    else if (mode == PQ_LIGHT) { if (per_sm == 7) PQ_LAUNCH(7, PQ_LIGHT); else PQ_LAUNCH(6, PQ_LIGHT); }
    else { ZB_FOR_FEATURE_SET(feat, (pq_launch_beauty<F>(per_sm, blocks, stream, P, fix, head, total, batch))); }
#undef PQ_LAUNCH
    return cudaGetLastError();
}

cudaError_t zbk_launch_aov_pass(const RenderParams& P, int n_tiles_owned, const ZbTuning& tn, cudaStream_t stream) {
    if (n_tiles_owned <= 0 || !P.accum_aux) return cudaSuccess;
    const uint32_t feat = pick_feature_set(P, tn);
    ZB_FOR_FEATURE_SET(feat, (k_aov_pass<F><<<n_tiles_owned, ZB_BLOCK_THREADS, 0, stream>>>(P, nullptr)));
    return cudaGetLastError();
}

cudaError_t zbk_launch_aov_pass_fixed(const RenderParams& P, int n_tiles_owned, unsigned long long* fix_aux, const ZbTuning& tn,
                                      cudaStream_t stream) {
    if (n_tiles_owned <= 0 || !fix_aux) return cudaSuccess;
    const uint32_t feat = pick_feature_set(P, tn);
    ZB_FOR_FEATURE_SET(feat, (k_aov_pass<F><<<n_tiles_owned, ZB_BLOCK_THREADS, 0, stream>>>(P, fix_aux)));
    return cudaGetLastError();
}

cudaError_t zbk_pq_resolve(const RenderParams& P, int mode, const unsigned long long* fix, int row0, int row1, float scale,
                           float4* mean_out, cudaStream_t stream) {
    if (row1 <= row0) return cudaSuccess;
    const size_t pixels = size_t(P.cam.width) * P.cam.height;
    const unsigned long long* fix_aux = mode == PQ_BEAUTY ? nullptr : fix + 3 * pixels;
    k_pq_resolve<<<dim3((P.cam.width + 63) / 64, (row1 - row0 + 3) / 4), 256, 0, stream>>>(P, fix, fix_aux, mode, row0, row1, scale, mean_out);
    return cudaGetLastError();
}

cudaError_t zbk_launch_primary_ids(const RenderParams& P, uint32_t* leaf_id, uint32_t* prim_id, float* t,
                                   cudaStream_t stream) {
    int n = P.cam.width * P.cam.height;
    k_primary_ids<<<(n + 127) / 128, 128, 0, stream>>>(P, leaf_id, prim_id, t);
    return cudaGetLastError();
}

// float4 means -> packed RGB floats (12 B / pixel): what crosses PCIe on the way to the reference's 3-double layout
__global__ void __launch_bounds__(256) k_pack_rgb(const float4* __restrict__ in, size_t n, float* __restrict__ out) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = in[i];
    out[3 * i + 0] = a.x;
    out[3 * i + 1] = a.y;
    out[3 * i + 2] = a.z;
}

cudaError_t zbk_launch_pack_rgb(const float4* in, size_t n, float* out, cudaStream_t stream) {
    if (!n) return cudaSuccess;
    k_pack_rgb<<<unsigned((n + 255) / 256), 256, 0, stream>>>(in, n, out);
    return cudaGetLastError();
}

cudaError_t zbk_launch_resolve(const float4* accum, size_t n, float scale, float4* out_f4, double* out_f64,
                               cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    k_resolve<<<unsigned((n + 255) / 256), 256, 0, stream>>>(accum, n, scale, out_f4, out_f64);
    return cudaGetLastError();
}

// Device-side re-encoding of the BVH nodes after upload / device build: (lo, hi) per axis -> (centre, half extent),
// rounded outwards so that the re-encoded box contains the original one.  An empty child (inverted box) becomes
// half extent -inf, which no ray can hit.  See ZB_BOX_CH in trace.cuh.
__global__ void __launch_bounds__(256) k_convert_nodes(float4* __restrict__ nodes, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 3 * n) return;
    const uint32_t node = i / 3, part = i - 3 * node;  // the three float4 that hold (lo, hi) pairs
    float4 v = nodes[4 * size_t(node) + part];
    auto enc = [](float lo, float hi, float& c, float& h) {
        if (!(lo <= hi)) { c = 0.f; h = -INFINITY; return; }
        c = 0.5f * lo + 0.5f * hi;
        h = fmaxf(__fsub_ru(hi, c), __fsub_ru(c, lo));
    };
    float4 o;
    enc(v.x, v.y, o.x, o.y);
    enc(v.z, v.w, o.z, o.w);
    nodes[4 * size_t(node) + part] = o;
}

bool zbk_nodes_center_half() { return ZB_BOX_CH != 0; }

cudaError_t zbk_convert_nodes(float4* nodes, uint32_t n, cudaStream_t stream) {
    if (!n || !ZB_BOX_CH) return cudaSuccess;
    k_convert_nodes<<<(3 * n + 255) / 256, 256, 0, stream>>>(nodes, n);
    return cudaGetLastError();
}
