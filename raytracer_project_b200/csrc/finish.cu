// finish.cu — the finishing chain on the device-resident frame, hand-written for sm_100a.
// Replaces, on W*H float4 means left in HBM by k_resolve:
//   post_processor::analyze_framebuffer  color_processing.hpp:153-186   -> k_luma_stats
//   bloom_filter::generate_bloom_overlay bloom.hpp:18-68                -> k_bloom_h (bright pass fused) + k_bloom_v
//   camera::update_post_processing       camera.hpp:142-207             -> k_finish: exposure, + bloom overlay,
//   post_processor::apply_sharpening     color_processing.hpp:212-232      5-tap sharpening, process(), gamma,
//   post_processor::process              color_processing.hpp:76-150       8-bit pack (camera.hpp:773-775)
// Every kernel is a streaming pass over the frame with 16-byte accesses; neighbours come through shared memory
// (horizontal blur) or L1/L2 (vertical blur, sharpening taps).  HBM-bound unless HSV / ACES are switched on.
// Algorithmic bytes per pixel: k_luma_stats 16 (read); k_bloom_h 16 + 16; k_bloom_v 16 + 16;
// k_finish 16 (+16 with bloom) read + 16 (float4) and/or 4 (RGBA8) written.
// The reference computes in double (HSV in float); the device computes in fp32.
#include <algorithm>

#include "abi_internal.h"
#include "vecmath.cuh"

struct PostDev {
    float ev;           // 2^exposure (camera.hpp:160)
    float exposure;     // the plain multiplier process() applies again (color_processing.hpp:91)
    float saturation, contrast, hue_shift, vignette;
    float3 balance;
    int aces, dbg_r, dbg_g, dbg_b, dbg_lum, dbg_bvh;
    int bloom, bloom_radius;
    float bloom_threshold, bloom_intensity;
    int sharpen;
    float sharpen_amount;
};

ZB_DEV float luminance(float3 c) { return fmaf(0.2126f, c.x, fmaf(0.7152f, c.y, 0.0722f * c.z)); }  // vec3.hpp:106-108

// ---------------------------------------------------------------------------------- analyze_framebuffer
// out: [0] sum of log2 max(lum, 1e-4) as double bits, [1] max luminance as float bits, [2..257] histogram
__global__ void __launch_bounds__(256) k_luma_stats(const float4* __restrict__ img, size_t n, unsigned long long* __restrict__ out) {
    __shared__ int s_hist[256];
    __shared__ double s_sum[8];
    __shared__ float s_max[8];
    s_hist[threadIdx.x] = 0;
    __syncthreads();
    double sum = 0.0;
    float mx = 0.f;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
        float4 p = __ldg(img + i);
        float lum = luminance(f3(p));
        mx = fmaxf(mx, lum);
        float l2 = log2f(fmaxf(0.0001f, lum));  // full-precision log2: the histogram bin edges are sensitive to it
        sum += double(l2);
        float norm = (l2 - (-10.0f)) / 20.0f;
        int bin = min(max(int(norm * 255.0f), 0), 255);
        atomicAdd(&s_hist[bin], 1);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
        mx = fmaxf(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
    }
    if ((threadIdx.x & 31) == 0) { s_sum[threadIdx.x >> 5] = sum; s_max[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++) { sum += s_sum[w]; mx = fmaxf(mx, s_max[w]); }
        atomicAdd(reinterpret_cast<double*>(out), sum);
        atomicMax(reinterpret_cast<unsigned int*>(out + 1), __float_as_uint(mx));  // luminance >= 0: bit order = value order
    }
    int h = s_hist[threadIdx.x];
    if (h) atomicAdd(reinterpret_cast<int*>(out + 2) + threadIdx.x, h);
}

// ---------------------------------------------------------------------------------- bloom
// bright pass of bloom.hpp:27-38 on the exposed colour (src * 2^exposure), evaluated on the fly by the horizontal blur
ZB_DEV float3 bright_pass(float4 p, const PostDev& P) {
    float3 c = P.ev * f3(p);
    float lum = luminance(c);
    if (!(lum > P.bloom_threshold)) return f3(0, 0, 0);
    float factor = (lum - P.bloom_threshold) * P.bloom_intensity;
    return (factor / fmaxf(lum, 0.0001f)) * c;
}

// blur_pass (bloom.hpp:48-67): tent weights 1 - |o| / (r + 1), renormalised by the weights that fall inside the image.
// Horizontal pass: a block stages the bright pass of its 256 pixels + 2r halo in shared memory once (the bright pass
// is evaluated once per pixel, not once per tap), then every thread sums its 2r + 1 taps from there.
#define ZB_BLOOM_MAX_R 64
__global__ void __launch_bounds__(256) k_bloom_h(const float4* __restrict__ in, float4* __restrict__ out, int W, int H, PostDev P) {
    __shared__ float4 s_row[256 + 2 * ZB_BLOOM_MAX_R];
    const int r = P.bloom_radius, x0 = blockIdx.x * 256, y = blockIdx.y;
    const float4* row = in + size_t(y) * W;
    for (int i = threadIdx.x; i < 256 + 2 * r; i += 256) {
        const int sx = x0 - r + i;
        float3 b = f3(0, 0, 0);
        if (sx >= 0 && sx < W) b = bright_pass(__ldg(row + sx), P);
        s_row[i] = make_float4(b.x, b.y, b.z, 0.f);
    }
    __syncthreads();
    const int x = x0 + threadIdx.x;
    if (x >= W) return;
    float3 sum = f3(0, 0, 0);
    float total = 0.f;
    const float inv_r1 = 1.0f / float(r + 1);
    for (int o = -r; o <= r; o++) {
        if (x + o < 0 || x + o >= W) continue;
        const float w = 1.0f - float(abs(o)) * inv_r1;
        sum += w * f3(s_row[threadIdx.x + r + o]);
        total += w;
    }
    const float inv = total > 0.f ? 1.0f / total : 0.f;
    out[size_t(y) * W + x] = make_float4(sum.x * inv, sum.y * inv, sum.z * inv, 0.f);
}

// Vertical pass: one thread per column and ZB_BLOOM_ROWS consecutive rows; the rows a block touches are shared by
// its neighbours above and below through L2 (each input row is fetched from HBM once).
#define ZB_BLOOM_ROWS 8
__global__ void __launch_bounds__(256) k_bloom_v(const float4* __restrict__ in, float4* __restrict__ out, int W, int H, PostDev P) {
    const int x = blockIdx.x * 256 + threadIdx.x, y0 = blockIdx.y * ZB_BLOOM_ROWS;
    if (x >= W) return;
    const int r = P.bloom_radius;
    const float inv_r1 = 1.0f / float(r + 1);
    float3 sum[ZB_BLOOM_ROWS];
    float total[ZB_BLOOM_ROWS];
#pragma unroll
    for (int k = 0; k < ZB_BLOOM_ROWS; k++) { sum[k] = f3(0, 0, 0); total[k] = 0.f; }
    for (int sy = max(y0 - r, 0); sy <= min(y0 + ZB_BLOOM_ROWS - 1 + r, H - 1); sy++) {
        const float3 p = f3(__ldg(in + size_t(sy) * W + x));
#pragma unroll
        for (int k = 0; k < ZB_BLOOM_ROWS; k++) {
            const int o = abs(sy - (y0 + k));
            if (o <= r) {
                const float w = 1.0f - float(o) * inv_r1;
                sum[k] += w * p;
                total[k] += w;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < ZB_BLOOM_ROWS; k++) {
        if (y0 + k >= H) break;
        const float inv = total[k] > 0.f ? 1.0f / total[k] : 0.f;
        out[size_t(y0 + k) * W + x] = make_float4(sum[k].x * inv, sum[k].y * inv, sum[k].z * inv, 0.f);
    }
}

// ---------------------------------------------------------------------------------- process()
// common.hpp:67-72; x^(1/2.2) = 2^(log2(x) / 2.2) on the MUFU pipe (lg2.approx / ex2.approx, ~2 ulp on [0, 1]) instead
// of the 30-instruction powf: with three of these per pixel the libm version made k_finish issue-bound
ZB_DEV float gamma22(float v) { return v > 0.f ? exp2f(__log2f(v) * (1.0f / 2.2f)) : 0.f; }
ZB_DEV float3 clamp_gamma(float3 c) {
    return f3(gamma22(fminf(fmaxf(c.x, 0.f), 1.f)), gamma22(fminf(fmaxf(c.y, 0.f), 1.f)), gamma22(fminf(fmaxf(c.z, 0.f), 1.f)));
}
ZB_DEV float aces1(float v) {  // common.hpp:48-64
    if (isnan(v) || isinf(v)) return 0.f;
    float x = fmaxf(0.f, v);
    return __fdividef(x * (2.51f * x + 0.03f), x * (2.43f * x + 0.59f) + 0.14f);  // MUFU.RCP + FMUL (2 ulp); the IEEE divide is ~10 instructions
}
// color_processing.hpp:280-306
ZB_DEV float3 rgb_to_hsv(float3 c) {
    float mx = fmaxf(c.x, fmaxf(c.y, c.z)), mn = fminf(c.x, fminf(c.y, c.z));
    float d = mx - mn, h = 0.f;
    float s = mx < 1e-6f ? 0.f : __fdividef(d, mx);
    if (mx != mn) {
        const float inv_d = __fdividef(1.0f, d);
        if (mx == c.x) h = (c.y - c.z) * inv_d + (c.y < c.z ? 6.f : 0.f);
        else if (mx == c.y) h = (c.z - c.x) * inv_d + 2.f;
        else h = (c.x - c.y) * inv_d + 4.f;
        h *= (1.0f / 6.f);
    }
    return f3(h * 360.f, s, mx);
}
// color_processing.hpp:308-343
ZB_DEV float3 hsv_to_rgb(float3 hsv) {
    float h = hsv.x / 360.f, s = hsv.y, v = hsv.z;
    int i = int(h * 6.f);
    float f = h * 6.f - float(i);
    float p = v * (1.f - s), q = v * (1.f - f * s), t = v * (1.f - (1.f - f) * s);
    switch (i % 6) {
        case 0: return f3(v, t, p);
        case 1: return f3(q, v, p);
        case 2: return f3(p, v, t);
        case 3: return f3(p, q, v);
        case 4: return f3(t, p, v);
        case 5: return f3(v, p, q);
        default: return f3(0, 0, 0);
    }
}
// post_processor::process for a beauty pass (color_processing.hpp:76-150)
ZB_DEV float3 process(float3 c, float u, float v, const PostDev& P) {
    c = P.exposure * c;  // the plain exposure value once more, on top of 2^exposure (color_processing.hpp:91)
    c = c * P.balance;
    if (fabsf(P.contrast - 1.0f) > 0.001f) {  // pivot 0.18 (color_processing.hpp:236-244)
        c = f3(fmaxf(0.f, (c.x - 0.18f) * P.contrast + 0.18f), fmaxf(0.f, (c.y - 0.18f) * P.contrast + 0.18f),
               fmaxf(0.f, (c.z - 0.18f) * P.contrast + 0.18f));
    }
    if (P.vignette > 0.f) {
        float dist = sqrtf((u - 0.5f) * (u - 0.5f) + (v - 0.5f) * (v - 0.5f));
        c = fminf(fmaxf(1.0f - dist * P.vignette, 0.f), 1.f) * c;
    }
    if (fabsf(P.saturation - 1.0f) > 0.001f || fabsf(P.hue_shift) > 0.001f) {
        float luma = luminance(c);
        if (luma > 0.0001f) {
            float3 hsv = rgb_to_hsv(__fdividef(1.0f, luma) * c);
            // fmodf(h + shift, 360) as h - 360 trunc(h / 360): fmodf is a ~40-instruction loop, this is four
            hsv.x += P.hue_shift;
            hsv.x = fmaf(-360.f, truncf(hsv.x * (1.0f / 360.f)), hsv.x);
            if (hsv.x < 0.f) hsv.x += 360.f;
            hsv.y = fminf(fmaxf(hsv.y * P.saturation, 0.f), 1.f);
            c = luma * hsv_to_rgb(hsv);
        }
    }
    if (P.aces) c = f3(aces1(c.x), aces1(c.y), aces1(c.z));
    if (P.dbg_r | P.dbg_g | P.dbg_b | P.dbg_lum | P.dbg_bvh) {  // apply_debug_view (color_processing.hpp:246-278)
        if (P.dbg_lum) {
            float l = luminance(c);
            c = l >= 1.0f ? f3(1, 1, 1) : l > 0.95f ? f3(1, 0, 0) : l > 0.70f ? f3(1, 1, 0) : l > 0.40f ? f3(.5f, .5f, .5f)
              : l > 0.10f ? f3(0, .5f, 0) : l > 0.02f ? f3(0, 0, 1) : f3(0.1f, 0.f, 0.2f);
        } else if (!P.dbg_bvh) {
            c = f3(P.dbg_r ? c.x : 0.f, P.dbg_g ? c.y : 0.f, P.dbg_b ? c.z : 0.f);
        }
    }
    return clamp_gamma(c);
}

// update_post_processing (camera.hpp:142-207).  A thread finishes ZB_FIN_PX consecutive pixels of a row so that four
// 16-byte loads are in flight per thread and the RGBA8 result leaves as one 16-byte store; `overlay` is the blurred
// bloom plane or null.
#define ZB_FIN_PX 4
__global__ void __launch_bounds__(128) k_finish(const float4* __restrict__ src, const float4* __restrict__ overlay, int W, int H,
                                                int pass_kind, PostDev P, float4* __restrict__ out_f4, uchar4* __restrict__ out_u8) {
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * ZB_FIN_PX, y = blockIdx.y;
    if (x0 >= W) return;
    const size_t row = size_t(y) * W;
    const bool beauty = pass_kind == ZB_PASS_BEAUTY;
    auto tap = [&](size_t i) {
        float3 t = f3(__ldg(src + i));
        if (beauty) {
            t = P.ev * t;
            if (overlay) t += f3(__ldg(overlay + i));
        }
        return t;
    };
    float3 c[ZB_FIN_PX], res[ZB_FIN_PX];
#pragma unroll
    for (int k = 0; k < ZB_FIN_PX; k++) c[k] = x0 + k < W ? tap(row + x0 + k) : f3(0, 0, 0);
#pragma unroll
    for (int k = 0; k < ZB_FIN_PX; k++) res[k] = c[k];
    if (beauty && P.sharpen && y >= 1 && y < H - 1) {  // interior pixels only (color_processing.hpp:218-219)
        const float3 left = x0 >= 1 ? tap(row + x0 - 1) : f3(0, 0, 0);
        const float3 right = x0 + ZB_FIN_PX < W ? tap(row + x0 + ZB_FIN_PX) : f3(0, 0, 0);
#pragma unroll
        for (int k = 0; k < ZB_FIN_PX; k++) {
            const int x = x0 + k;
            if (x < 1 || x >= W - 1) continue;
            const float3 l = k > 0 ? c[k - 1] : left, r = k < ZB_FIN_PX - 1 ? c[k + 1] : right;
            const float3 sum = 5.0f * c[k] - tap(row + x - W) - tap(row + x + W) - l - r;
            res[k] = (1.0f - P.sharpen_amount) * c[k] + P.sharpen_amount * sum;
        }
    }
    const float inv_w = 1.0f / float(W - 1), v = float(y) / float(H - 1);
    uchar4 packed[ZB_FIN_PX];
#pragma unroll
    for (int k = 0; k < ZB_FIN_PX; k++) {
        float3 o;
        if (beauty) o = process(res[k], float(x0 + k) * inv_w, v, P);
        else if (pass_kind == ZB_PASS_LIGHT) o = process(P.ev * res[k], float(x0 + k) * inv_w, v, P);
        else o = clamp_gamma(res[k]);
        res[k] = o;
        packed[k] = make_uchar4((unsigned char)(255.999f * o.x), (unsigned char)(255.999f * o.y), (unsigned char)(255.999f * o.z), 255);
    }
    const bool full = x0 + ZB_FIN_PX <= W && (W % ZB_FIN_PX) == 0;  // aligned 16-byte store of four RGBA8 pixels
    if (out_f4) {
#pragma unroll
        for (int k = 0; k < ZB_FIN_PX; k++)
            if (x0 + k < W) out_f4[row + x0 + k] = make_float4(res[k].x, res[k].y, res[k].z, 1.0f);
    }
    if (out_u8) {
        if (full) {
            *reinterpret_cast<uint4*>(out_u8 + row + x0) = *reinterpret_cast<uint4*>(packed);
        } else {
#pragma unroll
            for (int k = 0; k < ZB_FIN_PX; k++)
                if (x0 + k < W) out_u8[row + x0 + k] = packed[k];
        }
    }
}

// ---------------------------------------------------------------------------------- fused beauty chain with bloom
// update_post_processing with bloom as ONE pass over the frame (radius <= ZB_FUSED_MAX_R; the reference's default is 4):
// a block owns a 64x32 tile and stages, in shared memory, (1) the bright pass of the tile + (r + 1) halo and the exposed
// colour of the tile + 1, (2) the horizontal tent blur of those rows, (3) the vertical blur of the tile + 1 — the bloom
// plane the 5-tap sharpening needs at its four neighbours — then (4) sharpens, runs process() and stores RGBA8 / float4.
// Same arithmetic, in the same order, as k_bloom_h + k_bloom_v + k_finish.  Algorithmic bytes: 16 in + 4 (RGBA8) or 16
// (float4) out per pixel instead of 100 (two float4 planes written and re-read, sharpening taps through L2); the halo
// (1.5x the tile at r = 4) is shared with the neighbouring tiles through L2.
#define ZB_FUSED_MAX_R 6
#define ZB_FT_W 64
#define ZB_FT_H 32
// Row pitches of the shared-memory planes are ODD: a warp whose lanes walk down a column (the horizontal blur's strips) then
// touches 32 different banks.
#define ZB_FT_CB (ZB_FT_W + 3)
#define ZB_FT_CA(r) ((ZB_FT_W + 2 * ((r) + 1)) | 1)

// One strip of S outputs of the tent blur (bloom.hpp:48-67) along a line of one channel, the 2R + S inputs held in
// registers: (2R + S) / S shared-memory loads per output instead of 2R + 1.  in[(k + o) * IS] is tap o of output k,
// out[k * OS] the output; g0 = image coordinate of output 0, G = image extent along the line.  CHECK: some tap or output may
// lie outside the image (border tiles) — weights of missing taps are left out of the normalisation, outputs outside are 0.
// The weights 1 - |o| / (R + 1) are added and applied in the reference's order o = -R .. R.
template <int R, int S, int IS, int OS, bool CHECK, bool ACCUM>
__device__ __forceinline__ void tent_strip(const float* __restrict__ in, float* __restrict__ out, int g0, int G, float inv_all) {
    constexpr float inv_r1 = 1.0f / float(R + 1);
    float win[S + 2 * R];
#pragma unroll
    for (int j = 0; j < S + 2 * R; j++) win[j] = in[(j - R) * IS];
#pragma unroll
    for (int k = 0; k < S; k++) {
        float sum = 0.f, inv = inv_all;
        if (!CHECK) {
#pragma unroll
            for (int o = -R; o <= R; o++) sum += (1.0f - float(o < 0 ? -o : o) * inv_r1) * win[k + o + R];
        } else {
            const int g = g0 + k;
            float total = 0.f;
            if (g >= 0 && g < G) {
#pragma unroll
                for (int o = -R; o <= R; o++) {
                    if (g + o < 0 || g + o >= G) continue;
                    const float w = 1.0f - float(o < 0 ? -o : o) * inv_r1;
                    sum += w * win[k + o + R];
                    total += w;
                }
            }
            inv = total > 0.f ? 1.0f / total : 0.f;
        }
        if (ACCUM) out[k * OS] += sum * inv;
        else out[k * OS] = sum * inv;
    }
}

template <int R>
__global__ void __launch_bounds__(256) k_finish_fused(const float4* __restrict__ src, int W, int H, PostDev P, float4* __restrict__ out_f4,
                                                      uchar4* __restrict__ out_u8) {
    extern __shared__ float smem[];
    // the radius is a template parameter: every index below divides by a compile-time constant
    constexpr int r = R, h = r + 1;
    constexpr int RA = ZB_FT_H + 2 * h, WA = ZB_FT_W + 2 * h, CA = ZB_FT_CA(R), WB = ZB_FT_W + 2, CB = ZB_FT_CB, RD = ZB_FT_H + 2;
    float* A = smem;                    // [3][RA][CA] bright pass (WA columns used)
    float* B = A + 3 * RA * CA;         // [3][RA][CB] horizontally blurred (WB columns used)
    float* D = B + 3 * RA * CB;         // [3][RD][CB] exposed colour ev * src, then t = ev * src + bloom
    const int x0 = blockIdx.x * ZB_FT_W, y0 = blockIdx.y * ZB_FT_H;
    constexpr float inv_r1 = 1.0f / float(r + 1);
    // ---- 1: bright pass over the tile + h halo, exposed colour over the tile + 1.  All of a thread's pixels are fetched
    // before the first is used (two blocks of 8 warps per SM cannot hide a DRAM round trip per pixel otherwise); pixels
    // outside the image are staged as black, which is what bright_pass and ev * src make of them
    constexpr int N1 = (RA * WA + 255) / 256;
    float4 px[N1];
#pragma unroll
    for (int k = 0; k < N1; k++) {
        const int i = threadIdx.x + k * 256;
        const int cy = i / WA, cx = i - cy * WA;
        const int gx = x0 - h + cx, gy = y0 - h + cy;
        px[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < RA * WA && gx >= 0 && gx < W && gy >= 0 && gy < H) px[k] = __ldg(src + size_t(gy) * W + gx);
    }
#pragma unroll
    for (int k = 0; k < N1; k++) {
        const int i = threadIdx.x + k * 256;
        if (i >= RA * WA) break;
        const int cy = i / WA, cx = i - cy * WA;
        const float3 b = bright_pass(px[k], P), e = P.ev * f3(px[k]);
        const int ia = cy * CA + cx;
        A[ia] = b.x;
        A[RA * CA + ia] = b.y;
        A[2 * RA * CA + ia] = b.z;
        const int dx = cx - r, dy = cy - r;  // position in the tile + 1 region
        if (dx >= 0 && dx < WB && dy >= 0 && dy < RD) {
            const int j = dy * CB + dx;
            D[j] = e.x;
            D[RD * CB + j] = e.y;
            D[2 * RD * CB + j] = e.z;
        }
    }
    __syncthreads();
    // A block whose staged region lies inside the image (all but the border tiles) skips the per-tap bounds checks; the sum
    // of the weights is then the same constant for every pixel (added in the same order as the checked loop adds them).
    const bool x_in = x0 - h >= 0 && x0 + ZB_FT_W + h <= W, y_in = y0 - h >= 0 && y0 + ZB_FT_H + h <= H;
    float w_all = 0.f;
#pragma unroll
    for (int o = -r; o <= r; o++) w_all += 1.0f - float(o < 0 ? -o : o) * inv_r1;
    const float inv_all = 1.0f / w_all;
    // ---- 2: horizontal blur of every staged row for the columns of the tile + 1: strips of SH outputs; consecutive lanes
    // take consecutive rows (odd pitch: conflict-free)
    constexpr int SH = 11, NSH = WB / SH;
    static_assert(NSH * SH == WB, "strip length divides the blurred width");
    for (int i = threadIdx.x; i < 3 * NSH * RA; i += 256) {
        const int cy = i % RA, t = i / RA, strip = t % NSH, c = t / NSH;
        const int x = strip * SH;
        const float* a = A + c * RA * CA + cy * CA + x + r;  // column of output x in A = x - 1 + h
        float* o = B + c * RA * CB + cy * CB + x;
        if (x_in) tent_strip<R, SH, 1, 1, false, false>(a, o, x0 - 1 + x, W, inv_all);
        else tent_strip<R, SH, 1, 1, true, false>(a, o, x0 - 1 + x, W, inv_all);
    }
    __syncthreads();
    // ---- 3: vertical blur for the rows of the tile + 1, added to the exposed colour in place (D becomes t): strips of SV
    // rows, consecutive lanes on consecutive columns
    constexpr int SV = 17, NSV = RD / SV;
    static_assert(NSV * SV == RD, "strip length divides the blurred height");
    for (int i = threadIdx.x; i < 3 * NSV * WB; i += 256) {
        const int x = i % WB, t = i / WB, strip = t % NSV, c = t / NSV;
        const int y = strip * SV;
        const float* b = B + c * RA * CB + (y + r) * CB + x;  // row of output y in B = y - 1 + h
        float* o = D + c * RD * CB + y * CB + x;
        if (y_in) tent_strip<R, SV, CB, CB, false, true>(b, o, y0 - 1 + y, H, inv_all);
        else tent_strip<R, SV, CB, CB, true, true>(b, o, y0 - 1 + y, H, inv_all);
    }
    __syncthreads();
    // ---- 4: 5-tap sharpening of interior pixels on t, process(), store.  One pixel per thread and step, consecutive
    // threads on consecutive pixels: conflict-free shared-memory reads, coalesced stores (a warp writes 128 B of RGBA8 or
    // 512 B of float4)
    auto tap = [&](int tx, int ty) {  // tile coordinates, -1 .. tile size
        const int j = (ty + 1) * CB + tx + 1;
        return f3(D[j], D[RD * CB + j], D[2 * RD * CB + j]);
    };
    const float inv_w = 1.0f / float(W - 1), inv_h = 1.0f / float(H - 1);
    for (int q = threadIdx.x; q < ZB_FT_W * ZB_FT_H; q += 256) {
        const int ty = q / ZB_FT_W, tx = q - ty * ZB_FT_W;
        const int y = y0 + ty, x = x0 + tx;
        if (y >= H || x >= W) continue;
        const float3 c = tap(tx, ty);
        float3 o = c;
        if (P.sharpen && y >= 1 && y < H - 1 && x >= 1 && x < W - 1) {  // color_processing.hpp:218-219
            const float3 sum = 5.0f * c - tap(tx, ty - 1) - tap(tx, ty + 1) - tap(tx - 1, ty) - tap(tx + 1, ty);
            o = (1.0f - P.sharpen_amount) * c + P.sharpen_amount * sum;
        }
        o = process(o, float(x) * inv_w, float(y) / float(H - 1), P);
        (void)inv_h;
        const size_t at = size_t(y) * W + x;
        if (out_f4) out_f4[at] = make_float4(o.x, o.y, o.z, 1.0f);
        if (out_u8) out_u8[at] = make_uchar4((unsigned char)(255.999f * o.x), (unsigned char)(255.999f * o.y), (unsigned char)(255.999f * o.z), 255);
    }
}

size_t fused_smem_bytes(int r) {
    const int h = r + 1, RA = ZB_FT_H + 2 * h, CA = ZB_FT_CA(r), CB = ZB_FT_CB, RD = ZB_FT_H + 2;
    return sizeof(float) * 3 * size_t(RA * CA + RA * CB + RD * CB);
}

// ---------------------------------------------------------------------------------- C ABI
namespace {

PostDev post_dev(const zb_post* p) {
    PostDev d{};
    d.ev = float(std::pow(2.0, double(p->exposure)));
    d.exposure = p->exposure;
    d.saturation = p->saturation;
    d.contrast = p->contrast;
    d.hue_shift = p->hue_shift;
    d.vignette = p->vignette_intensity;
    d.balance = make_float3(p->color_balance[0], p->color_balance[1], p->color_balance[2]);
    d.aces = p->use_aces_tone_mapping;
    d.dbg_r = p->debug_red; d.dbg_g = p->debug_green; d.dbg_b = p->debug_blue; d.dbg_lum = p->debug_luminance; d.dbg_bvh = p->debug_bvh;
    d.bloom = p->use_bloom;
    d.bloom_radius = std::max(0, std::min(p->bloom_radius, ZB_BLOOM_MAX_R));
    d.bloom_threshold = p->bloom_threshold;
    d.bloom_intensity = p->bloom_intensity;
    d.sharpen = p->use_sharpening && p->sharpen_amount > 0.0;
    d.sharpen_amount = float(p->sharpen_amount);
    return d;
}

int ensure_finish_scratch(zb_context* ctx, size_t pixels, bool planes) {
    if (!ctx->fin_stats) ZB_CUDA(cudaMalloc(&ctx->fin_stats, 130 * sizeof(unsigned long long)));
    if (planes && ctx->fin_pixels < pixels) {
        cudaFree(ctx->fin_a);
        cudaFree(ctx->fin_b);
        cudaFree(ctx->fin_in);
        ctx->fin_a = ctx->fin_b = ctx->fin_in = nullptr;
        ctx->fin_pixels = 0;
        ZB_CUDA(cudaMalloc(&ctx->fin_a, pixels * sizeof(float4)));
        ZB_CUDA(cudaMalloc(&ctx->fin_b, pixels * sizeof(float4)));
        ZB_CUDA(cudaMalloc(&ctx->fin_in, pixels * sizeof(float4)));
        ctx->fin_pixels = pixels;
    }
    return ZB_OK;
}

}  // namespace

extern "C" int zb_analyze_device(zb_context* ctx, const void* mean_f4, size_t n_pixels, zb_image_stats* out) {
    if (!ctx || !mean_f4 || !out) return fail(ZB_ERR_INVALID, "zb_analyze_device: NULL argument");
    ZB_CUDA(cudaSetDevice(ctx->device));
    int rc = ensure_finish_scratch(ctx, n_pixels, false);
    if (rc) return rc;
    ZB_CUDA(cudaMemsetAsync(ctx->fin_stats, 0, 130 * sizeof(unsigned long long), ctx->stream));
    if (n_pixels) {
        const unsigned blocks = unsigned(std::min<size_t>((n_pixels + 255) / 256, size_t(ctx->sm_count) * 8));
        k_luma_stats<<<blocks, 256, 0, ctx->stream>>>(static_cast<const float4*>(mean_f4), n_pixels, ctx->fin_stats);
        ZB_CUDA(cudaGetLastError());
        ctx->stats.launches++;
    }
    unsigned long long host[130];
    ZB_CUDA(cudaMemcpyAsync(host, ctx->fin_stats, sizeof(host), cudaMemcpyDeviceToHost, ctx->stream));
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));
    double sum;
    float mx;
    std::memcpy(&sum, &host[0], 8);
    std::memcpy(&mx, &host[1], 4);
    std::memcpy(out->histogram, &host[2], 256 * sizeof(int32_t));
    out->max_luminance = mx;
    // stats.average_luminance = pow(2.0f, float(total_log_lum / size)) (color_processing.hpp:183)
    out->average_luminance = n_pixels ? std::pow(2.0f, float(sum / double(n_pixels))) : 0.f;
    return ZB_OK;
}

// post_processor::apply_auto_exposure (color_processing.hpp:189-210)
extern "C" double zb_auto_exposure(const zb_post* post, const zb_image_stats* stats) {
    if (!post || !stats) return 0.0;
    if (stats->average_luminance <= 0.0) return double(post->exposure);
    if (!post->use_auto_exposure) return std::min(std::max(post->exposure, 0.01f), 10.0f);
    double safe = std::max(double(stats->average_luminance), 0.02);
    double raw = post->target_luminance / safe;
    double cur = raw * std::pow(2.0, double(post->exposure_compensation_stops));
    return std::min(std::max(float(cur), 0.01f), 4.0f);
}

extern "C" int zb_finish_device(zb_context* ctx, const void* mean_f4, int width, int height, const zb_post* post,
                                int pass_kind, void* out_f4, void* out_rgba8) {
    if (!ctx || !mean_f4 || !post || (!out_f4 && !out_rgba8)) return fail(ZB_ERR_INVALID, "zb_finish_device: NULL argument");
    if (width < 1 || height < 1 || pass_kind < ZB_PASS_BEAUTY || pass_kind > ZB_PASS_AUX) return fail(ZB_ERR_INVALID, "zb_finish_device: bad size / pass");
    ZB_CUDA(cudaSetDevice(ctx->device));
    const PostDev P = post_dev(post);
    const size_t pixels = size_t(width) * height;
    const float4* src = static_cast<const float4*>(mean_f4);
    const float4* overlay = nullptr;
    if (pass_kind == ZB_PASS_BEAUTY && P.bloom && P.bloom_radius <= ZB_FUSED_MAX_R && ctx->tuning.fuse_finish) {
        // bloom + sharpening + process() in one pass over shared-memory tiles
        const size_t smem = fused_smem_bytes(P.bloom_radius);
        const dim3 grid((width + ZB_FT_W - 1) / ZB_FT_W, (height + ZB_FT_H - 1) / ZB_FT_H);
        float4* o4 = static_cast<float4*>(out_f4);
        uchar4* o8 = static_cast<uchar4*>(out_rgba8);
#define ZB_FUSED_CASE(RR)                                                                                                          \
    case RR:                                                                                                                       \
        ZB_CUDA(cudaFuncSetAttribute(k_finish_fused<RR>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(fused_smem_bytes(RR)))); \
        k_finish_fused<RR><<<grid, 256, smem, ctx->stream>>>(src, width, height, P, o4, o8);                                       \
        break;
        switch (P.bloom_radius) {
            ZB_FUSED_CASE(0) ZB_FUSED_CASE(1) ZB_FUSED_CASE(2) ZB_FUSED_CASE(3) ZB_FUSED_CASE(4) ZB_FUSED_CASE(5) ZB_FUSED_CASE(6)
            default: return fail(ZB_ERR_INVALID, "zb_finish_device: bloom radius outside the fused range");
        }
#undef ZB_FUSED_CASE
        ZB_CUDA(cudaGetLastError());
        ctx->stats.launches++;
        return ZB_OK;
    }
    if (pass_kind == ZB_PASS_BEAUTY && P.bloom) {
        int rc = ensure_finish_scratch(ctx, pixels, true);
        if (rc) return rc;
        k_bloom_h<<<dim3((width + 255) / 256, height), 256, 0, ctx->stream>>>(src, ctx->fin_a, width, height, P);
        k_bloom_v<<<dim3((width + 255) / 256, (height + ZB_BLOOM_ROWS - 1) / ZB_BLOOM_ROWS), 256, 0, ctx->stream>>>(ctx->fin_a, ctx->fin_b, width, height, P);
        ZB_CUDA(cudaGetLastError());
        ctx->stats.launches += 2;
        overlay = ctx->fin_b;
    }
    k_finish<<<dim3((width + 128 * ZB_FIN_PX - 1) / (128 * ZB_FIN_PX), height), 128, 0, ctx->stream>>>(src, overlay, width, height, pass_kind, P, static_cast<float4*>(out_f4),
                                            static_cast<uchar4*>(out_rgba8));
    ZB_CUDA(cudaGetLastError());
    ctx->stats.launches++;
    return ZB_OK;
}

// Host-buffer variant: the reference's std::vector<color> in, final_framebuffer out (camera.hpp:142-207).
extern "C" int zb_finish(zb_context* ctx, const double* rgb_f64, int width, int height, const zb_post* post, int pass_kind,
                         double* out_rgb_f64) {
    if (!ctx || !rgb_f64 || !post || !out_rgb_f64) return fail(ZB_ERR_INVALID, "zb_finish: NULL argument");
    if (width < 1 || height < 1) return fail(ZB_ERR_INVALID, "zb_finish: bad size");
    ZB_CUDA(cudaSetDevice(ctx->device));
    const size_t pixels = size_t(width) * height;
    int rc = ensure_finish_scratch(ctx, pixels, true);
    if (rc) return rc;
    std::vector<float4> stage(pixels);
    for (size_t i = 0; i < pixels; i++) stage[i] = make_float4(float(rgb_f64[3 * i]), float(rgb_f64[3 * i + 1]), float(rgb_f64[3 * i + 2]), 1.f);
    ZB_CUDA(cudaMemcpyAsync(ctx->fin_in, stage.data(), pixels * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    // the display-referred result goes to fin_a unless bloom needs it; use a separate output plane then
    float4* out_dev = (pass_kind == ZB_PASS_BEAUTY && post->use_bloom) ? nullptr : ctx->fin_a;
    float4* tmp = nullptr;
    if (!out_dev) {
        ZB_CUDA(cudaMalloc(&tmp, pixels * sizeof(float4)));
        out_dev = tmp;
    }
    rc = zb_finish_device(ctx, ctx->fin_in, width, height, post, pass_kind, out_dev, nullptr);
    if (!rc) {
        cudaError_t e = cudaMemcpyAsync(stage.data(), out_dev, pixels * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = fail(ZB_ERR_CUDA, std::string("zb_finish: ") + cudaGetErrorString(e));
    }
    if (tmp) cudaFree(tmp);
    if (rc) return rc;
    for (size_t i = 0; i < pixels; i++) {
        out_rgb_f64[3 * i] = double(stage[i].x);
        out_rgb_f64[3 * i + 1] = double(stage[i].y);
        out_rgb_f64[3 * i + 2] = double(stage[i].z);
    }
    return ZB_OK;
}

// camera::save_render_pass (camera.hpp:299-343) -> process_framebuffer_to_image (:701-780): RGB / DENOISE take the
// beauty chain (the export's order of bloom, sharpening and exposure equals update_post_processing's in exact
// arithmetic), every other pass — the light passes too, unlike the preview — is clamp + gamma; 255.999 * v to 8 bits.
extern "C" int zb_save_pass(zb_context* ctx, const void* mean_f4, int width, int height, const zb_post* post, int render_pass,
                            const char* path) {
    if (!ctx || !mean_f4 || !post || !path) return fail(ZB_ERR_INVALID, "zb_save_pass: NULL argument");
    if (width < 1 || height < 1 || render_pass < ZB_RP_RGB || render_pass > ZB_RP_Z_DEPTH) return fail(ZB_ERR_INVALID, "zb_save_pass: bad size / pass");
    ZB_CUDA(cudaSetDevice(ctx->device));
    const size_t pixels = size_t(width) * height;
    uchar4* dev = nullptr;
    ZB_CUDA(cudaMalloc(&dev, pixels * sizeof(uchar4)));
    const int kind = (render_pass == ZB_RP_RGB || render_pass == ZB_RP_DENOISE) ? ZB_PASS_BEAUTY : ZB_PASS_AUX;
    int rc = zb_finish_device(ctx, mean_f4, width, height, post, kind, nullptr, dev);
    std::vector<uchar4> host(pixels);
    if (!rc) {
        cudaError_t e = cudaMemcpyAsync(host.data(), dev, pixels * sizeof(uchar4), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = fail(ZB_ERR_CUDA, std::string("zb_save_pass: ") + cudaGetErrorString(e));
    }
    cudaFree(dev);
    if (rc) return rc;
    return zb_write_png(path, width, height, reinterpret_cast<const unsigned char*>(host.data()), 4, 0);
}

// Linear radiance (float4 means on the device) as a Radiance .hdr file: what stbi_loadf reads back as an environment.
extern "C" int zb_save_hdr(zb_context* ctx, const void* mean_f4, int width, int height, const char* path) {
    if (!ctx || !mean_f4 || !path || width < 1 || height < 1) return fail(ZB_ERR_INVALID, "zb_save_hdr: bad argument");
    ZB_CUDA(cudaSetDevice(ctx->device));
    const size_t pixels = size_t(width) * height;
    std::vector<float4> host(pixels);
    ZB_CUDA(cudaMemcpyAsync(host.data(), mean_f4, pixels * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));
    return zb_write_hdr(path, width, height, reinterpret_cast<const float*>(host.data()), 4, 0);
}

// Host-buffer variant of zb_save_pass: the reference's std::vector<color> (3 doubles per pixel) in.
extern "C" int zb_save_pass_f64(zb_context* ctx, const double* rgb_f64, int width, int height, const zb_post* post,
                                int render_pass, const char* path) {
    if (!ctx || !rgb_f64 || !post || !path) return fail(ZB_ERR_INVALID, "zb_save_pass_f64: NULL argument");
    if (width < 1 || height < 1) return fail(ZB_ERR_INVALID, "zb_save_pass_f64: bad size");
    ZB_CUDA(cudaSetDevice(ctx->device));
    const size_t pixels = size_t(width) * height;
    int rc = ensure_finish_scratch(ctx, pixels, true);
    if (rc) return rc;
    std::vector<float4> stage(pixels);
    for (size_t i = 0; i < pixels; i++) stage[i] = make_float4(float(rgb_f64[3 * i]), float(rgb_f64[3 * i + 1]), float(rgb_f64[3 * i + 2]), 1.f);
    ZB_CUDA(cudaMemcpyAsync(ctx->fin_in, stage.data(), pixels * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));  // `stage` leaves scope before zb_save_pass returns otherwise
    return zb_save_pass(ctx, ctx->fin_in, width, height, post, render_pass, path);
}

// Host-buffer variant of zb_analyze_device (post_processor::analyze_framebuffer on a std::vector<color>).
extern "C" int zb_analyze_f64(zb_context* ctx, const double* rgb_f64, size_t n_pixels, zb_image_stats* out) {
    if (!ctx || !rgb_f64 || !out || !n_pixels) return fail(ZB_ERR_INVALID, "zb_analyze_f64: bad argument");
    ZB_CUDA(cudaSetDevice(ctx->device));
    int rc = ensure_finish_scratch(ctx, n_pixels, true);
    if (rc) return rc;
    std::vector<float4> stage(n_pixels);
    for (size_t i = 0; i < n_pixels; i++) stage[i] = make_float4(float(rgb_f64[3 * i]), float(rgb_f64[3 * i + 1]), float(rgb_f64[3 * i + 2]), 1.f);
    ZB_CUDA(cudaMemcpyAsync(ctx->fin_in, stage.data(), n_pixels * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));
    return zb_analyze_device(ctx, ctx->fin_in, n_pixels, out);
}
