// kernels.cu — sm_100a kernels of the path-tracing hot path and their launchers.
//   k_render      K6 megakernel: raygen (K1) + trace (K2) + shade (K3) per (pixel, sample), iterative
//                 ray_color (camera.hpp:928-1004) with the reference's termination rules, float4
//                 radiance sums in HBM (K5 accumulate), first-hit AOVs (K8, camera.hpp:464-488).
//   k_primary_ids K7: centre-of-pixel pinhole primary hit (camera.hpp:809-814 + :459).
//   k_resolve     K5: sums -> means, float4 or the reference's double RGB.
// Replaces render_rows (camera.hpp:430-554).  No tensor cores: nothing here is a dense contraction.
#include "kernels.h"
#include "shade.cuh"

// One thread owns one pixel and walks its samples in order, so the per-pixel sum is deterministic
// and needs no atomics.  A block renders a 16x8 pixel tile (each warp an 8x4 sub-tile, for coherent
// primary rays); tiles are dealt to ranks round-robin for the tile-stripe partition.
template <bool STATS>
__global__ void __launch_bounds__(ZB_BLOCK_THREADS) k_render(const __grid_constant__ RenderParams P) {
    const int tile = blockIdx.x * P.tile_ranks + P.tile_rank;
    const int tx = tile % P.tiles_x, ty = tile / P.tiles_x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int px = tx * ZB_TILE_W + (warp & 1) * 8 + (lane & 7);
    const int py = ty * ZB_TILE_H + (warp >> 1) * 4 + (lane >> 3);
    const bool active = px < P.cam.width && py < P.cam.height && ty < P.tiles_y;
    Counters cnt = {0, 0, 0, 0};
    uint32_t paths = 0;
    if (active) {
        const uint32_t pixel = uint32_t(py) * uint32_t(P.cam.width) + uint32_t(px);
        const Philox ph{P.seed_lo, P.seed_hi};
        float3 sum = f3(0, 0, 0), s_alb = f3(0, 0, 0), s_nrm = f3(0, 0, 0);
        float s_z = 0.f;
        const float3 base = P.cam.pix00_rel + float(px) * P.cam.du + float(py) * P.cam.dv;
        for (int s = P.sample_begin; s < P.sample_end; s++) {
            // ---- K1 raygen: get_ray (camera.hpp:784-794), sample_square (:817), defocus disk (:822)
            uint4 r0 = ph(pixel, uint32_t(s), 0u, 0u);
            float3 d = base + (u01(r0.x) - 0.5f) * P.cam.du + (u01(r0.y) - 0.5f) * P.cam.dv;
            float3 o = P.cam.center;
            if (P.cam.defocus) {
                float rr = sqrtf(u01(r0.z)), sn, cs;
                sincospif(2.0f * u01(r0.w), &sn, &cs);
                float3 off = (rr * cs) * P.cam.disk_u + (rr * sn) * P.cam.disk_v;
                o += off;
                d = d - off;
            }
            paths++;
            float3 L = f3(0, 0, 0), beta = f3(1, 1, 1);
            for (int seg = 0; seg < P.max_depth; seg++) {
                MediumRng mr{ph, pixel, uint32_t(s), uint32_t(seg + 1)};
                Hit hit;
                if (!trace<STATS>(P.sc, o, d, 0.001f, INFINITY, mr, hit, cnt)) {
                    L += beta * background(P.sc, P.env, d);
                    if (seg == 0 && s < P.aux_samples && P.use_normal) s_nrm += f3(0.5f, 0.5f, 1.0f);
                    break;
                }
                Surface sf;
                make_surface(P.sc, hit, o, d, sf);
                if (seg == 0 && s < P.aux_samples) {
                    if (P.use_albedo) s_alb += material_get_albedo(P.sc, sf);
                    if (P.use_normal) {
                        float3 n = unit_vector(sf.n);
                        s_nrm += f3((dot(n, P.cam.u) + 1.f) * 0.5f, (dot(n, P.cam.v) + 1.f) * 0.5f, (dot(n, P.cam.w) + 1.f) * 0.5f);
                    }
                    if (P.use_z) s_z += 1.0f - clamp01(hit.t * P.inv_z_max);
                }
                L += beta * material_emitted(P.sc, sf);
                uint4 rb = ph(pixel, uint32_t(s), uint32_t(seg + 1), 0u);
                float3 atten = f3(1, 1, 1), no, nd;
                if (!material_scatter(P.sc, sf, d, f3(u01(rb.x), u01(rb.y), u01(rb.z)), atten, no, nd)) break;
                beta *= atten;
                o = no;
                d = nd;
                // ray_color's loop index is seg-1; its `i > 10` rules start at seg 12 (camera.hpp:968-983)
                if (seg >= 12) {
                    if (dot(beta, beta) < 1e-8f) break;
                    float p = fminf(fmaxf(max3(beta.x, beta.y, beta.z), 0.05f), 0.95f);
                    if (u01(rb.w) > p) break;
                    beta *= 1.0f / p;
                }
            }
            sum += L;
        }
        float4 acc = make_float4(sum.x, sum.y, sum.z, float(P.sample_end - P.sample_begin));
        if (P.accumulate) {
            float4 old = P.accum[pixel];
            acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
        }
        P.accum[pixel] = acc;
        if (P.accum_aux) {
            const size_t n = size_t(P.cam.width) * P.cam.height;
            float4 a0 = make_float4(s_alb.x, s_alb.y, s_alb.z, 0.f), a1 = make_float4(s_nrm.x, s_nrm.y, s_nrm.z, 0.f),
                   a2 = make_float4(s_z, s_z, s_z, 0.f);
            if (P.accumulate) {
                float4 b0 = P.accum_aux[pixel], b1 = P.accum_aux[n + pixel], b2 = P.accum_aux[2 * n + pixel];
                a0.x += b0.x; a0.y += b0.y; a0.z += b0.z;
                a1.x += b1.x; a1.y += b1.y; a1.z += b1.z;
                a2.x += b2.x; a2.y += b2.y; a2.z += b2.z;
            }
            P.accum_aux[pixel] = a0;
            P.accum_aux[n + pixel] = a1;
            P.accum_aux[2 * n + pixel] = a2;
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
    MediumRng mr{Philox{P.seed_lo, P.seed_hi}, uint32_t(idx), 0u, 0u};
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

cudaError_t zbk_launch_render(const RenderParams& P, int n_tiles_owned, bool stats, cudaStream_t stream) {
    if (n_tiles_owned <= 0) return cudaSuccess;
    if (stats) k_render<true><<<n_tiles_owned, ZB_BLOCK_THREADS, 0, stream>>>(P);
    else k_render<false><<<n_tiles_owned, ZB_BLOCK_THREADS, 0, stream>>>(P);
    return cudaGetLastError();
}

cudaError_t zbk_launch_primary_ids(const RenderParams& P, uint32_t* leaf_id, uint32_t* prim_id, float* t,
                                   cudaStream_t stream) {
    int n = P.cam.width * P.cam.height;
    k_primary_ids<<<(n + 127) / 128, 128, 0, stream>>>(P, leaf_id, prim_id, t);
    return cudaGetLastError();
}

cudaError_t zbk_launch_resolve(const float4* accum, size_t n, float scale, float4* out_f4, double* out_f64,
                               cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    k_resolve<<<unsigned((n + 255) / 256), 256, 0, stream>>>(accum, n, scale, out_f4, out_f64);
    return cudaGetLastError();
}
