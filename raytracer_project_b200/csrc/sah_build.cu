// sah_build.cu — GPU construction of a mesh BLAS with the surface-area heuristic.  Replaces the host
// build the reference repeats on every GUI restart (bvh.hpp:11-44 via model.hpp:95, main.cpp:1492-1500) with a tree of the
// same quality as host/bvh_build.hpp's binned-SAH tree — the linear BVH of lbvh.cu builds faster but its frames are 42 %
// slower on the 1 M triangle config.
//
// Top-down, two phases:
//  A  level-synchronous over "segments" (contiguous ranges of the primitive order) of more than 32 primitives: per level
//     one pass reduces centroid and box bounds per segment, one pass bins every primitive into 16 bins on each axis (count +
//     box per bin; shared-memory bins when a block's 2048 positions lie in one segment), one thread per segment sweeps the
//     3 x 15 planes for the cheapest split, and a STABLE partition (flags, exclusive scan, scatter) moves boxes and ids.
//  B  one warp per segment of <= 32 primitives, one primitive per lane: every lane proposes "everything before my centroid
//     goes left" on each axis and evaluates it exactly against the other lanes' boxes by shuffles — the full sweep SAH, all
//     sub-segments of the warp in the same round — then the lanes are permuted in place; <= 4 primitives become a leaf where
//     that is cheaper (same rule as the host builder).
// Nodes need no allocator: the inner node that splits [lo, hi) at position mid is staged at index mid (every boundary is
// used once), children patch their box and code into the parent's 64-byte node, and a final scan compacts the used slots
// (root first).  Everything is keyed by position, integer atomics only: the tree is the same on every run.
// The depth cap of the traversal stack (ZB_MAX_BLAS_DEPTH) is kept as the host builder keeps it: a split that would leave
// a child too large to finish with balanced halves falls back to a split by count.
//
// HBM-bound streaming passes; per level and primitive: reduce 36 B, bin 36 B, flags 36 B + 4 B, scan 12 B, scatter 44 B
// read + 40 B written.  ~20 levels for 1 M triangles.
#include <algorithm>
#include <cstdio>

#include "abi_internal.h"
#include "build_util.cuh"
#include "vecmath.cuh"

namespace {

#define SB_BINS 16
#define SB_SMALL 32
#define SB_LEAF_MAX 4
#define SB_CHUNK 2048  // positions per block of the streaming passes: 256 threads x 8

struct Seg {  // one segment of phase A
    uint32_t start, count;
    int parent, side, depth;  // staging index of the parent node (-1: root), which child of it, inner nodes above
    float cb_lo[3], k1[3];    // bin = clamp(int(k1 * (centroid - cb_lo))); k1 = 0: no extent on that axis
    float half_area;
    int axis, bin;            // chosen split: bins <= bin go left; axis 3 = by count
    uint32_t n_left;
    int child[2];             // segment index on the next level, -1 = finished by phase B
    int pad[3];
};
struct SmallSeg { uint32_t start, count; int parent, side_depth; };  // side = bit 0, depth above

ZB_DEV int bin_of(float c, float lo, float k1) { return max(0, min(SB_BINS - 1, int(k1 * (c - lo)))); }
ZB_DEV float axis_of(float3 v, int a) { return a == 0 ? v.x : a == 1 ? v.y : v.z; }
ZB_DEV float half_area(float3 lo, float3 hi) {
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return (dx < 0.f || dy < 0.f || dz < 0.f) ? 0.f : dx * dy + dy * dz + dz * dx;
}
// inner levels a balanced tree over n primitives needs (host/bvh_build.hpp balanced_levels)
ZB_DEV int balanced_levels(uint32_t n, uint32_t leaf_max = SB_LEAF_MAX) {
    int l = 0;
    while (n > leaf_max) { n = (n + 1) / 2; l++; }
    return l;
}
// child `side` of a staged node: box and code (zb_bvh_node keeps both children's boxes in the parent)
ZB_DEV void patch_child(zb_bvh_node* nd, int side, float3 lo, float3 hi, int code) {
    float* xy = side ? nd->rx_ry : nd->lx_ly;
    xy[0] = lo.x; xy[1] = hi.x; xy[2] = lo.y; xy[3] = hi.y;
    nd->lz_rz[2 * side] = lo.z;
    nd->lz_rz[2 * side + 1] = hi.z;
    (side ? nd->right : nd->left) = code;
}
ZB_DEV int leaf_code(uint32_t first, uint32_t count) { return int(~((first << 3) | (count - 1u))); }

// ---- boxes, identity order, everything in segment 0 (or in none when one warp can finish the mesh)
__global__ void __launch_bounds__(256) k_prep(const float4* __restrict__ tris, uint32_t n, float4* __restrict__ lo4, float4* __restrict__ hi4,
                                              uint32_t* __restrict__ idx, int* __restrict__ seg) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const TriBox b = tri_box(tris + 3 * size_t(i));
    lo4[i] = make_float4(b.lo.x, b.lo.y, b.lo.z, 0.f);
    hi4[i] = make_float4(b.hi.x, b.hi.y, b.hi.z, 0.f);
    idx[i] = i;
    seg[i] = n > SB_SMALL ? 0 : -1;
}

// the same for primitives given by their boxes (TLAS items: 2 float4 each, lo and hi)
__global__ void __launch_bounds__(256) k_prep_boxes(const float4* __restrict__ boxes, uint32_t n, float4* __restrict__ lo4, float4* __restrict__ hi4,
                                                    uint32_t* __restrict__ idx, int* __restrict__ seg) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    lo4[i] = boxes[2 * size_t(i)];
    hi4[i] = boxes[2 * size_t(i) + 1];
    idx[i] = i;
    seg[i] = n > SB_SMALL ? 0 : -1;
}
__global__ void __launch_bounds__(256) k_permute_pairs(const float4* __restrict__ in, const uint32_t* __restrict__ ids, uint32_t n,
                                                       float4* __restrict__ out) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per float4
    if (t >= 2 * n) return;
    out[t] = in[2 * size_t(ids[t >> 1]) + (t & 1)];
}

// ---- exclusive scan of n uint32 in place: tile sums, one block over the sums, tiles again
__global__ void __launch_bounds__(256) k_scan_tiles(const uint32_t* __restrict__ v, uint32_t n, uint32_t* __restrict__ sums) {
    __shared__ uint32_t ws[8];
    const uint32_t base = blockIdx.x * SB_CHUNK;
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const uint32_t i = base + k * 256 + threadIdx.x;
        s += i < n ? v[i] : 0u;
    }
    s = __reduce_add_sync(0xFFFFFFFFu, s);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) sums[blockIdx.x] = ws[0] + ws[1] + ws[2] + ws[3] + ws[4] + ws[5] + ws[6] + ws[7];
}
__global__ void __launch_bounds__(1024) k_scan_sums(uint32_t* __restrict__ sums, uint32_t total) {
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t start = 0; start < total; start += 1024) {
        const uint32_t i = start + threadIdx.x;
        const uint32_t v = i < total ? sums[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if ((threadIdx.x & 31) >= o) x += y;
        }
        if ((threadIdx.x & 31) == 31) warp_sum[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_sum[threadIdx.x];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, w, o);
                if (threadIdx.x >= o) w += y;
            }
            warp_sum[threadIdx.x] = w;
        }
        __syncthreads();
        const uint32_t before = carry + (threadIdx.x >= 32 ? warp_sum[(threadIdx.x >> 5) - 1] : 0u);
        if (i < total) sums[i] = before + x - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + x;
        __syncthreads();
    }
}
// a thread owns 8 CONSECUTIVE entries of its tile here (the order of the sum is what matters, not coalescing of 32 B runs)
__global__ void __launch_bounds__(256) k_scan_apply(uint32_t* __restrict__ v, uint32_t n, const uint32_t* __restrict__ sums) {
    __shared__ uint32_t ws[8];
    const uint32_t first = blockIdx.x * SB_CHUNK + threadIdx.x * 8;
    uint32_t x[8], s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        x[k] = first + k < n ? v[first + k] : 0u;
        s += x[k];
    }
    uint32_t inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if ((threadIdx.x & 31) >= o) inc += y;
    }
    if ((threadIdx.x & 31) == 31) ws[threadIdx.x >> 5] = inc;
    __syncthreads();
    uint32_t before = sums[blockIdx.x] + inc - s;
    for (int w = 0; w < int(threadIdx.x >> 5); w++) before += ws[w];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        if (first + k < n) v[first + k] = before;
        before += x[k];
    }
}

// ---- A1: centroid bounds and box bounds per segment, 12 ordered uints: centroid lo.xyz hi.xyz, box lo.xyz hi.xyz
__global__ void __launch_bounds__(256) k_seg_reduce(const float4* __restrict__ lo4, const float4* __restrict__ hi4, const int* __restrict__ seg,
                                                    uint32_t n, uint32_t* __restrict__ red) {
    __shared__ float sm[8][12];
    const uint32_t base = blockIdx.x * SB_CHUNK, last = min(base + SB_CHUNK, n) - 1;
    const int s_first = seg[base], s_last = seg[last];
    // positions of a segment are contiguous: equal ends = one segment.  (Finished positions, -1, are NOT one segment: live
    // segments may lie between two finished ones.)
    const bool uniform = s_first == s_last && s_first >= 0;
    float v[12];
#pragma unroll
    for (int j = 0; j < 12; j++) v[j] = (j % 6) < 3 ? INFINITY : -INFINITY;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const uint32_t i = base + k * 256 + threadIdx.x;
        if (i >= n) continue;
        const int s = uniform ? s_first : seg[i];
        if (s < 0) continue;
        const float3 lo = f3(__ldg(lo4 + i)), hi = f3(__ldg(hi4 + i));
        const float3 c = 0.5f * (lo + hi);
        if (uniform) {
            v[0] = fminf(v[0], c.x); v[1] = fminf(v[1], c.y); v[2] = fminf(v[2], c.z);
            v[3] = fmaxf(v[3], c.x); v[4] = fmaxf(v[4], c.y); v[5] = fmaxf(v[5], c.z);
            v[6] = fminf(v[6], lo.x); v[7] = fminf(v[7], lo.y); v[8] = fminf(v[8], lo.z);
            v[9] = fmaxf(v[9], hi.x); v[10] = fmaxf(v[10], hi.y); v[11] = fmaxf(v[11], hi.z);
        } else {
            uint32_t* r = red + 12 * size_t(s);
            atomicMin(r + 0, f2ord(c.x)); atomicMin(r + 1, f2ord(c.y)); atomicMin(r + 2, f2ord(c.z));
            atomicMax(r + 3, f2ord(c.x)); atomicMax(r + 4, f2ord(c.y)); atomicMax(r + 5, f2ord(c.z));
            atomicMin(r + 6, f2ord(lo.x)); atomicMin(r + 7, f2ord(lo.y)); atomicMin(r + 8, f2ord(lo.z));
            atomicMax(r + 9, f2ord(hi.x)); atomicMax(r + 10, f2ord(hi.y)); atomicMax(r + 11, f2ord(hi.z));
        }
    }
    if (!uniform) return;
#pragma unroll
    for (int j = 0; j < 12; j++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float y = __shfl_xor_sync(0xFFFFFFFFu, v[j], o);
            v[j] = (j % 6) < 3 ? fminf(v[j], y) : fmaxf(v[j], y);
        }
        if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5][j] = v[j];
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        const int j = threadIdx.x;
        float x = sm[0][j];
        for (int w = 1; w < 8; w++) x = (j % 6) < 3 ? fminf(x, sm[w][j]) : fmaxf(x, sm[w][j]);
        uint32_t* r = red + 12 * size_t(s_first) + j;
        if ((j % 6) < 3) atomicMin(r, f2ord(x));
        else atomicMax(r, f2ord(x));
    }
}

// ---- A1b: per segment: bin scales, node area; the segment's box goes into its parent's node
__global__ void __launch_bounds__(256) k_seg_prepare(Seg* __restrict__ segs, uint32_t m, const uint32_t* __restrict__ red,
                                                     zb_bvh_node* __restrict__ stage) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= m) return;
    const uint32_t* r = red + 12 * size_t(s);
    Seg& S = segs[s];
    for (int a = 0; a < 3; a++) {
        const float lo = ord2f(r[a]), hi = ord2f(r[3 + a]);
        const float ext = hi - lo;
        S.cb_lo[a] = lo;
        S.k1[a] = ext > 0.f ? float(SB_BINS) * (1.0f - 1e-6f) / ext : 0.f;
    }
    const float3 blo = f3(ord2f(r[6]), ord2f(r[7]), ord2f(r[8])), bhi = f3(ord2f(r[9]), ord2f(r[10]), ord2f(r[11]));
    S.half_area = half_area(blo, bhi);
    if (S.parent >= 0) {  // the code is written by k_decide, when the split position is known
        zb_bvh_node* nd = stage + S.parent;
        float* xy = S.side ? nd->rx_ry : nd->lx_ly;
        xy[0] = blo.x; xy[1] = bhi.x; xy[2] = blo.y; xy[3] = bhi.y;
        nd->lz_rz[2 * S.side] = blo.z;
        nd->lz_rz[2 * S.side + 1] = bhi.z;
    }
}

// ---- A2: bins.  Global layout [segment][axis][bin][8]: count, box lo.xyz, box hi.xyz (ordered uints), pad
__global__ void __launch_bounds__(256) k_bin(const float4* __restrict__ lo4, const float4* __restrict__ hi4, const int* __restrict__ seg,
                                             uint32_t n, const Seg* __restrict__ segs, uint32_t* __restrict__ bins) {
    __shared__ uint32_t sb[3 * SB_BINS * 7];
    const uint32_t base = blockIdx.x * SB_CHUNK, last = min(base + SB_CHUNK, n) - 1;
    const int s_first = seg[base], s_last = seg[last];
    const bool uniform = s_first == s_last && s_first >= 0;
    if (uniform) {
        for (int j = threadIdx.x; j < 3 * SB_BINS * 7; j += 256) {
            const int f = j % 7;
            sb[j] = f == 0 ? 0u : f <= 3 ? 0xFFFFFFFFu : 0u;
        }
        __syncthreads();
    }
#pragma unroll 2
    for (int k = 0; k < 8; k++) {
        const uint32_t i = base + k * 256 + threadIdx.x;
        if (i >= n) continue;
        const int s = uniform ? s_first : seg[i];
        if (s < 0) continue;
        const Seg& S = segs[s];
        const float3 lo = f3(__ldg(lo4 + i)), hi = f3(__ldg(hi4 + i));
        const float3 c = 0.5f * (lo + hi);
        const uint32_t o[6] = {f2ord(lo.x), f2ord(lo.y), f2ord(lo.z), f2ord(hi.x), f2ord(hi.y), f2ord(hi.z)};
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float k1 = S.k1[a];
            if (k1 == 0.f) continue;
            const int b = bin_of(axis_of(c, a), S.cb_lo[a], k1);
            if (uniform) {
                uint32_t* t = sb + (a * SB_BINS + b) * 7;
                atomicAdd(t, 1u);
                atomicMin(t + 1, o[0]); atomicMin(t + 2, o[1]); atomicMin(t + 3, o[2]);
                atomicMax(t + 4, o[3]); atomicMax(t + 5, o[4]); atomicMax(t + 6, o[5]);
            } else {
                uint32_t* t = bins + ((size_t(s) * 3 + a) * SB_BINS + b) * 8;
                atomicAdd(t, 1u);
                atomicMin(t + 1, o[0]); atomicMin(t + 2, o[1]); atomicMin(t + 3, o[2]);
                atomicMax(t + 4, o[3]); atomicMax(t + 5, o[4]); atomicMax(t + 6, o[5]);
            }
        }
    }
    if (!uniform) return;
    __syncthreads();
    for (int j = threadIdx.x; j < 3 * SB_BINS * 7; j += 256) {
        const int f = j % 7, ab = j / 7;
        const uint32_t v = sb[j];
        uint32_t* t = bins + (size_t(s_first) * 3 * SB_BINS + ab) * 8 + f;
        if (f == 0) { if (v) atomicAdd(t, v); }
        else if (f <= 3) { if (v != 0xFFFFFFFFu) atomicMin(t, v); }
        else if (v) atomicMax(t, v);
    }
}

__global__ void __launch_bounds__(256) k_init_level(uint32_t* __restrict__ red, uint32_t* __restrict__ bins, uint32_t m) {
    const size_t t = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t < size_t(m) * 12) red[t] = (t % 6) < 3 ? 0xFFFFFFFFu : 0u;
    if (t < size_t(m) * 3 * SB_BINS * 8) {
        const int f = int(t & 7);
        bins[t] = (f >= 1 && f <= 3) ? 0xFFFFFFFFu : 0u;
    }
}

// ---- A3: one thread per segment: the cheapest of the 3 x 15 planes (host/bvh_build.hpp split(): same sweep, same tie
// order), the staged node, the two children
__global__ void __launch_bounds__(128) k_decide(Seg* __restrict__ segs, uint32_t m, const uint32_t* __restrict__ bins,
                                                zb_bvh_node* __restrict__ stage, Seg* __restrict__ next, uint32_t* __restrict__ counters,
                                                SmallSeg* __restrict__ small, int depth_cap) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= m) return;
    Seg& S = segs[s];
    float best = INFINITY;
    int best_axis = 3, best_bin = 0;
    uint32_t best_nl = 0;
    for (int a = 0; a < 3; a++) {
        if (S.k1[a] == 0.f) continue;
        const uint32_t* B = bins + (size_t(s) * 3 + a) * SB_BINS * 8;
        float ra[SB_BINS - 1];
        uint32_t rc[SB_BINS - 1];
        float3 lo = f3(INFINITY, INFINITY, INFINITY), hi = f3(-INFINITY, -INFINITY, -INFINITY);
        uint32_t c = 0;
        for (int b = SB_BINS - 1; b > 0; b--) {
            const uint32_t* t = B + b * 8;
            if (t[0]) {
                lo = f3(fminf(lo.x, ord2f(t[1])), fminf(lo.y, ord2f(t[2])), fminf(lo.z, ord2f(t[3])));
                hi = f3(fmaxf(hi.x, ord2f(t[4])), fmaxf(hi.y, ord2f(t[5])), fmaxf(hi.z, ord2f(t[6])));
                c += t[0];
            }
            ra[b - 1] = half_area(lo, hi);
            rc[b - 1] = c;
        }
        lo = f3(INFINITY, INFINITY, INFINITY); hi = f3(-INFINITY, -INFINITY, -INFINITY);
        c = 0;
        for (int b = 0; b < SB_BINS - 1; b++) {
            const uint32_t* t = B + b * 8;
            if (t[0]) {
                lo = f3(fminf(lo.x, ord2f(t[1])), fminf(lo.y, ord2f(t[2])), fminf(lo.z, ord2f(t[3])));
                hi = f3(fmaxf(hi.x, ord2f(t[4])), fmaxf(hi.y, ord2f(t[5])), fmaxf(hi.z, ord2f(t[6])));
                c += t[0];
            }
            if (c == 0 || rc[b] == 0) continue;
            const float cost = half_area(lo, hi) * float(c) + ra[b] * float(rc[b]);
            if (cost < best) { best = cost; best_axis = a; best_bin = b; best_nl = c; }
        }
    }
    uint32_t nl = best_nl;
    if (best_axis < 3 && S.depth + 1 + max(balanced_levels(nl), balanced_levels(S.count - nl)) > depth_cap) best_axis = 3;
    if (best_axis == 3) nl = (S.count + 1) / 2;  // coincident centroids, or the depth cap: halves of the current order
    S.axis = best_axis;
    S.bin = best_bin;
    S.n_left = nl;
    const int node = int(S.start + nl);
    stage[node].pad0 = 1;
    if (S.parent >= 0) (S.side ? stage[S.parent].right : stage[S.parent].left) = node;
    else counters[2] = uint32_t(node);  // the root
    atomicMax(counters + 3, uint32_t(S.depth + 1));
    for (int k = 0; k < 2; k++) {
        const uint32_t start = k ? S.start + nl : S.start, count = k ? S.count - nl : nl;
        if (count > SB_SMALL) {
            const uint32_t j = atomicAdd(counters + 0, 1u);
            Seg N{};
            N.start = start; N.count = count; N.parent = node; N.side = k; N.depth = S.depth + 1;
            next[j] = N;
            S.child[k] = int(j);
        } else {
            const uint32_t j = atomicAdd(counters + 1, 1u);
            small[j] = SmallSeg{start, count, node, ((S.depth + 1) << 1) | k};
            S.child[k] = -1;
        }
    }
}

// ---- A4: 1 = goes left.  flags has n + 1 entries (the last stays 0) and becomes the exclusive scan
__global__ void __launch_bounds__(256) k_flags(const float4* __restrict__ lo4, const float4* __restrict__ hi4, const int* __restrict__ seg,
                                               uint32_t n, const Seg* __restrict__ segs, uint32_t* __restrict__ flags) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    uint32_t f = 0;
    const int s = i < n ? seg[i] : -1;
    if (s >= 0) {
        const Seg& S = segs[s];
        if (S.axis == 3) f = (i - S.start) < S.n_left ? 1u : 0u;
        else {
            const float c = 0.5f * (axis_of(f3(__ldg(lo4 + i)), S.axis) + axis_of(f3(__ldg(hi4 + i)), S.axis));
            f = bin_of(c, S.cb_lo[S.axis], S.k1[S.axis]) <= S.bin ? 1u : 0u;
        }
    }
    flags[i] = f;
}

// ---- A6: stable partition of every split segment; finished positions are copied through
__global__ void __launch_bounds__(256) k_scatter(const float4* __restrict__ lo4, const float4* __restrict__ hi4, const uint32_t* __restrict__ idx,
                                                 const int* __restrict__ seg, uint32_t n, const Seg* __restrict__ segs,
                                                 const uint32_t* __restrict__ pre, float4* __restrict__ lo4o, float4* __restrict__ hi4o,
                                                 uint32_t* __restrict__ idxo, int* __restrict__ sego) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int s = seg[i];
    uint32_t dest = i;
    int ns = -1;
    if (s >= 0) {
        const Seg& S = segs[s];
        const uint32_t r = pre[i] - pre[S.start];  // lefts before i in the segment
        const bool left = pre[i + 1] != pre[i];
        dest = left ? S.start + r : S.start + S.n_left + (i - S.start - r);
        ns = S.child[left ? 0 : 1];
    }
    lo4o[dest] = __ldg(lo4 + i);
    hi4o[dest] = __ldg(hi4 + i);
    idxo[dest] = idx[i];
    sego[dest] = ns;
}

// ---- B: one warp per segment of <= 32 primitives (header).  Lane state: its primitive (box, id) and the sub-segment
// [slo, shi) of lanes it currently belongs to with that sub-segment's parent node, side and depth.
ZB_DEV float3 shfl3(float3 v, int src) {
    return f3(__shfl_sync(0xFFFFFFFFu, v.x, src), __shfl_sync(0xFFFFFFFFu, v.y, src), __shfl_sync(0xFFFFFFFFu, v.z, src));
}
ZB_DEV void grow_if(bool on, float3& lo, float3& hi, float3 qlo, float3 qhi) {
    lo = f3(on ? fminf(lo.x, qlo.x) : lo.x, on ? fminf(lo.y, qlo.y) : lo.y, on ? fminf(lo.z, qlo.z) : lo.z);
    hi = f3(on ? fmaxf(hi.x, qhi.x) : hi.x, on ? fmaxf(hi.y, qhi.y) : hi.y, on ? fmaxf(hi.z, qhi.z) : hi.z);
}

__global__ void __launch_bounds__(128) k_small(const float4* __restrict__ lo4, const float4* __restrict__ hi4, const uint32_t* __restrict__ idx,
                                               const SmallSeg* __restrict__ small, uint32_t n_small, zb_bvh_node* __restrict__ stage,
                                               uint32_t tri_first, uint32_t* __restrict__ out_idx, uint32_t* __restrict__ counters,
                                               int depth_cap, float isect_cost, int leaf_max) {
    __shared__ int src_of[4][32];
    const int w = threadIdx.x >> 5, me = threadIdx.x & 31;
    const uint32_t sidx = blockIdx.x * 4 + w;
    if (sidx >= n_small) return;  // the whole warp leaves
    const SmallSeg S = small[sidx];
    const bool have = uint32_t(me) < S.count;
    const float3 inf3 = f3(INFINITY, INFINITY, INFINITY), ninf3 = f3(-INFINITY, -INFINITY, -INFINITY);
    float3 lo = inf3, hi = ninf3;
    uint32_t id = 0;
    if (have) {
        lo = f3(__ldg(lo4 + S.start + me));
        hi = f3(__ldg(hi4 + S.start + me));
        id = idx[S.start + me];
    }
    int slo = have ? 0 : me, shi = have ? int(S.count) : me + 1;
    int parent = S.parent, side = S.side_depth & 1, depth = S.side_depth >> 1;
    bool done = !have;
    int deepest = 0;
    while (__any_sync(0xFFFFFFFFu, !done)) {
        const int cnt = shi - slo;
        const float3 c = 0.5f * (lo + hi);
        const int trips = __reduce_max_sync(0xFFFFFFFFu, done ? 0 : cnt);
        // round 1: the sub-segment's box, and for "before my centroid goes left" on each axis both sides' boxes and counts
        float3 blo = inf3, bhi = ninf3;
        float3 Llo[3] = {inf3, inf3, inf3}, Lhi[3] = {ninf3, ninf3, ninf3}, Rlo[3] = {inf3, inf3, inf3}, Rhi[3] = {ninf3, ninf3, ninf3};
        int nl[3] = {0, 0, 0};
        for (int t = 0; t < trips; t++) {
            const int q = min(slo + t, 31);
            const bool act = !done && t < cnt;
            const float3 qlo = shfl3(lo, q), qhi = shfl3(hi, q);
            const float3 qc = 0.5f * (qlo + qhi);
            grow_if(act, blo, bhi, qlo, qhi);
#pragma unroll
            for (int a = 0; a < 3; a++) {
                const float cq = axis_of(qc, a), cm = axis_of(c, a);
                const bool before = cq < cm || (cq == cm && q < me);
                grow_if(act && before, Llo[a], Lhi[a], qlo, qhi);
                grow_if(act && !before, Rlo[a], Rhi[a], qlo, qhi);
                nl[a] += (act && before) ? 1 : 0;
            }
        }
        float bc = INFINITY;
        int ba = 0, bn = 0;
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float cost = nl[a] > 0 ? half_area(Llo[a], Lhi[a]) * float(nl[a]) + half_area(Rlo[a], Rhi[a]) * float(cnt - nl[a]) : INFINITY;
            if (cost < bc) { bc = cost; ba = a; bn = nl[a]; }
        }
        // round 2: the cheapest proposal of the sub-segment (first lane on ties)
        float wc = INFINITY;
        int wl = slo, wa = 0, wn = 0;
        for (int t = 0; t < trips; t++) {
            const int q = min(slo + t, 31);
            const float qc = __shfl_sync(0xFFFFFFFFu, bc, q);
            const int qa = __shfl_sync(0xFFFFFFFFu, ba, q), qn = __shfl_sync(0xFFFFFFFFu, bn, q);
            if (!done && t < cnt && qc < wc) { wc = qc; wl = q; wa = qa; wn = qn; }
        }
        const float3 wcen = shfl3(c, wl);
        const float ha = half_area(blo, bhi);
        const float split_cost = ha > 0.f ? 1.0f + isect_cost * wc / ha : INFINITY;  // one node visit + the children's expected tests
        // host rule (bvh_build.hpp): a leaf only where <= 4 primitives are cheaper tested than split, or nothing can split
        // them, or the depth cap is reached; the root always splits
        const bool can_split = wc < INFINITY;
        bool leaf = cnt == 1 || (cnt <= leaf_max && parent >= 0 && (!can_split || isect_cost * float(cnt) <= split_cost || depth >= depth_cap));
        bool by_count = !leaf && (!can_split || depth + 1 + max(balanced_levels(uint32_t(wn), leaf_max), balanced_levels(uint32_t(cnt - wn), leaf_max)) > depth_cap);
        const int n_left = by_count ? (cnt + 1) / 2 : wn;
        const float cm = axis_of(c, wa), cw = axis_of(wcen, wa);
        const bool is_left = by_count ? (me - slo) < n_left : (cm < cw || (cm == cw && me < wl));
        const int mid = slo + n_left;
        const int node = int(S.start) + mid;
        if (!done && me == slo) {
            const int code = leaf ? leaf_code(tri_first + S.start + uint32_t(slo), uint32_t(cnt)) : node;
            if (parent >= 0) patch_child(stage + parent, side, blo, bhi, code);
            else counters[2] = uint32_t(node);
            if (!leaf) stage[node].pad0 = 1;
        }
        const bool splitting = !done && !leaf;
        const uint32_t bl = __ballot_sync(0xFFFFFFFFu, splitting && is_left), br = __ballot_sync(0xFFFFFFFFu, splitting && !is_left);
        const uint32_t segmask = (shi >= 32 ? 0xFFFFFFFFu : (1u << shi) - 1u) & ~((1u << slo) - 1u);
        const uint32_t lt = (1u << me) - 1u;
        int dest = me;
        if (splitting) {
            dest = is_left ? slo + __popc(bl & segmask & lt) : mid + __popc(br & segmask & lt);
            if (is_left) shi = mid; else slo = mid;
            parent = node;
            side = is_left ? 0 : 1;
            depth += 1;
            deepest = max(deepest, depth);
        } else if (!done) {
            done = true;
        }
        src_of[w][dest] = me;
        __syncwarp();
        const int from = src_of[w][me];
        __syncwarp();
        lo = shfl3(lo, from);
        hi = shfl3(hi, from);
        id = __shfl_sync(0xFFFFFFFFu, id, from);
        slo = __shfl_sync(0xFFFFFFFFu, slo, from);
        shi = __shfl_sync(0xFFFFFFFFu, shi, from);
        parent = __shfl_sync(0xFFFFFFFFu, parent, from);
        side = __shfl_sync(0xFFFFFFFFu, side, from);
        depth = __shfl_sync(0xFFFFFFFFu, depth, from);
        done = __shfl_sync(0xFFFFFFFFu, int(done), from) != 0;
    }
    if (have) out_idx[S.start + me] = id;
    deepest = __reduce_max_sync(0xFFFFFFFFu, deepest);
    if (me == 0 && deepest > 0) atomicMax(counters + 3, uint32_t(deepest));
}

// ---- compaction of the staged nodes: used slots in index order, the root first
__global__ void __launch_bounds__(256) k_mark(const zb_bvh_node* __restrict__ stage, uint32_t n, uint32_t* __restrict__ flags) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    flags[i] = (i < n && stage[i].pad0) ? 1u : 0u;
}
__global__ void __launch_bounds__(256) k_compact(const zb_bvh_node* __restrict__ stage, uint32_t n, const uint32_t* __restrict__ pre,
                                                 uint32_t root, int node_base, zb_bvh_node* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || pre[i + 1] == pre[i]) return;
    auto remap = [&](uint32_t j) { return j == root ? 0u : pre[j] + (j < root ? 1u : 0u); };
    zb_bvh_node nd = stage[i];
    if (nd.left >= 0) nd.left = node_base + int(remap(uint32_t(nd.left)));
    if (nd.right >= 0) nd.right = node_base + int(remap(uint32_t(nd.right)));
    nd.pad0 = nd.pad1 = 0;
    out[remap(i)] = nd;
}
__global__ void k_single(const float4* __restrict__ tris, uint32_t tri_first, zb_bvh_node* __restrict__ out) {
    const TriBox b = tri_box(tris);
    zb_bvh_node nd;
    nd.lx_ly[0] = b.lo.x; nd.lx_ly[1] = b.hi.x; nd.lx_ly[2] = b.lo.y; nd.lx_ly[3] = b.hi.y;
    nd.rx_ry[0] = INFINITY; nd.rx_ry[1] = -INFINITY; nd.rx_ry[2] = INFINITY; nd.rx_ry[3] = -INFINITY;
    nd.lz_rz[0] = b.lo.z; nd.lz_rz[1] = b.hi.z; nd.lz_rz[2] = INFINITY; nd.lz_rz[3] = -INFINITY;
    nd.left = leaf_code(tri_first, 1);
    nd.right = ZB_CHILD_EMPTY;
    nd.pad0 = nd.pad1 = 0;
    out[0] = nd;
}
__global__ void __launch_bounds__(256) k_permute_tris(const float4* __restrict__ tris_in, const float4* __restrict__ nrm_in,
                                                      const uint32_t* __restrict__ ids, uint32_t n, float4* __restrict__ tris_out,
                                                      float4* __restrict__ nrm_out) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per float4
    if (t >= 3 * n) return;
    const uint32_t k = t / 3, j = t - 3 * k;
    const size_t src = 3 * size_t(ids[k]) + j;
    tris_out[t] = __ldg(tris_in + src);
    nrm_out[t] = __ldg(nrm_in + src);
}

// all scratch of a build is carved from ONE buffer that the context keeps across builds (zb_context::build_scratch)
struct Scratch {
    char* base = nullptr;
    size_t size = 0, used = 0;
    cudaError_t reserve(zb_context* ctx, size_t bytes) {
        void* p = nullptr;
        const cudaError_t e = ctx->build_scratch.get(bytes, &p);
        base = static_cast<char*>(p);
        size = e == cudaSuccess ? bytes : 0;
        return e;
    }
    template <class T> cudaError_t get(T** p, size_t count) {
        const size_t bytes = (std::max<size_t>(count, 1) * sizeof(T) + 255) & ~size_t(255);
        if (used + bytes > size) return cudaErrorMemoryAllocation;
        *p = reinterpret_cast<T*>(base + used);
        used += bytes;
        return cudaSuccess;
    }
};

int scan_in_place(uint32_t* v, uint32_t count, uint32_t* sums, cudaStream_t st) {
    const uint32_t tiles = (count + SB_CHUNK - 1) / SB_CHUNK;
    k_scan_tiles<<<tiles, 256, 0, st>>>(v, count, sums);
    k_scan_sums<<<1, 1024, 0, st>>>(sums, tiles);
    k_scan_apply<<<tiles, 256, 0, st>>>(v, count, sums);
    ZB_CUDA(cudaGetLastError());
    return ZB_OK;
}

// ---- optional structural check of a finished tree (tuning "build_check"; the GPU tests switch it on): every child index
// and leaf range is in bounds, every node but the root has exactly one parent, every primitive lies in exactly one leaf and
// inside that leaf's box, and an inner child's box equals the union of the boxes that child node stores.
// err[0] counts violations, err[1] keeps the first violating node + 1.
struct TriBoxes {  // primitive p of a mesh: its (padded) triangle box, from the permuted triangles
    const float4* tris;
    uint32_t first;
    ZB_DEV void get(uint32_t p, float3& lo, float3& hi) const {
        const TriBox b = tri_box(tris + 3 * size_t(p - first));
        lo = b.lo;
        hi = b.hi;
    }
};
struct PairBoxes {  // primitive p of a TLAS: the permuted item bounds
    const float4* boxes;
    ZB_DEV void get(uint32_t p, float3& lo, float3& hi) const {
        lo = f3(boxes[2 * size_t(p)]);
        hi = f3(boxes[2 * size_t(p) + 1]);
    }
};
template <class Boxes>
__global__ void __launch_bounds__(256) k_validate(const zb_bvh_node* __restrict__ nodes, uint32_t n_nodes, int node_base, uint32_t leaf_first,
                                                  uint32_t n, Boxes boxes, uint32_t* __restrict__ parents, uint32_t* __restrict__ cover,
                                                  uint32_t* __restrict__ err) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const zb_bvh_node nd = nodes[i];
    auto bad = [&]() { atomicAdd(err, 1u); atomicCAS(err + 1, 0u, i + 1u); };
    for (int side = 0; side < 2; side++) {
        const int code = side ? nd.right : nd.left;
        const float3 lo = side ? f3(nd.rx_ry[0], nd.rx_ry[2], nd.lz_rz[2]) : f3(nd.lx_ly[0], nd.lx_ly[2], nd.lz_rz[0]);
        const float3 hi = side ? f3(nd.rx_ry[1], nd.rx_ry[3], nd.lz_rz[3]) : f3(nd.lx_ly[1], nd.lx_ly[3], nd.lz_rz[1]);
        if (code == ZB_CHILD_EMPTY) { bad(); continue; }  // the builder never leaves a child empty (n >= 2)
        if (code >= 0) {
            const int c = code - node_base;
            if (c < 1 || uint32_t(c) >= n_nodes) { bad(); continue; }
            atomicAdd(parents + c, 1u);
            const zb_bvh_node ch = nodes[c];
            const float3 ulo = f3(fminf(ch.lx_ly[0], ch.rx_ry[0]), fminf(ch.lx_ly[2], ch.rx_ry[2]), fminf(ch.lz_rz[0], ch.lz_rz[2]));
            const float3 uhi = f3(fmaxf(ch.lx_ly[1], ch.rx_ry[1]), fmaxf(ch.lx_ly[3], ch.rx_ry[3]), fmaxf(ch.lz_rz[1], ch.lz_rz[3]));
            if (ulo.x != lo.x || ulo.y != lo.y || ulo.z != lo.z || uhi.x != hi.x || uhi.y != hi.y || uhi.z != hi.z) bad();
        } else {
            const uint32_t lc = ~uint32_t(code), first = lc >> 3, count = (lc & 7u) + 1u;
            if (first < leaf_first || first + count > leaf_first + n) { bad(); continue; }
            float3 ulo = f3(INFINITY, INFINITY, INFINITY), uhi = f3(-INFINITY, -INFINITY, -INFINITY);
            for (uint32_t p = first; p < first + count; p++) {
                atomicAdd(cover + (p - leaf_first), 1u);
                float3 plo, phi;
                boxes.get(p, plo, phi);
                ulo = f3(fminf(ulo.x, plo.x), fminf(ulo.y, plo.y), fminf(ulo.z, plo.z));
                uhi = f3(fmaxf(uhi.x, phi.x), fmaxf(uhi.y, phi.y), fmaxf(uhi.z, phi.z));
            }
            if (ulo.x != lo.x || ulo.y != lo.y || ulo.z != lo.z || uhi.x != hi.x || uhi.y != hi.y || uhi.z != hi.z) bad();
        }
    }
}
__global__ void __launch_bounds__(256) k_validate_counts(const uint32_t* __restrict__ parents, uint32_t n_nodes, const uint32_t* __restrict__ cover,
                                                         uint32_t n, uint32_t* __restrict__ err) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_nodes && parents[i] != (i == 0 ? 0u : 1u)) { atomicAdd(err, 1u); atomicCAS(err + 1, 0u, i + 1u); }
    if (i < n && cover[i] != 1u) { atomicAdd(err, 1u); atomicCAS(err + 1, 0u, 0x80000000u | i); }
}
template <class Boxes>
int validate_tree(zb_context* ctx, const zb_bvh_node* nodes, uint32_t n_nodes, uint32_t node_base, uint32_t leaf_first, uint32_t n, Boxes boxes,
                  const char* what) {
    cudaStream_t st = ctx->stream;
    uint32_t* buf = nullptr;
    const size_t words = size_t(n_nodes) + n + 2;
    ZB_CUDA(cudaMalloc(&buf, words * sizeof(uint32_t)));
    cudaMemsetAsync(buf, 0, words * sizeof(uint32_t), st);
    uint32_t *err = buf, *parents = buf + 2, *cover = buf + 2 + n_nodes;
    k_validate<<<(n_nodes + 255) / 256, 256, 0, st>>>(nodes, n_nodes, int(node_base), leaf_first, n, boxes, parents, cover, err);
    k_validate_counts<<<(std::max(n_nodes, n) + 255) / 256, 256, 0, st>>>(parents, n_nodes, cover, n, err);
    uint32_t h[2] = {0, 0};
    cudaMemcpyAsync(h, err, sizeof(h), cudaMemcpyDeviceToHost, st);
    const cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(buf);
    if (e != cudaSuccess) return fail(ZB_ERR_CUDA, std::string("device SAH build check: ") + cudaGetErrorString(e));
    if (h[0])
        return fail(ZB_ERR_CUDA, std::string("device SAH build check (") + what + "): " + std::to_string(h[0]) + " violations, first at " +
                                     ((h[1] & 0x80000000u) ? "primitive " + std::to_string(h[1] & 0x7FFFFFFFu) : "node " + std::to_string(h[1] - 1)));
    return ZB_OK;
}

struct CoreParams {
    uint32_t n;           // primitives (>= 2)
    uint32_t leaf_first;  // added to every leaf's first position (global position of primitive 0)
    uint32_t node_base;   // added to every inner child index
    int depth_cap;        // ZB_MAX_BLAS_DEPTH / ZB_MAX_TLAS_DEPTH
    float isect_cost;     // cost of one primitive test in node visits (host/bvh_build.hpp isect_cost)
    uint32_t max_nodes;   // room in nodes_out
    const char* what;
    int leaf_max = SB_LEAF_MAX;  // primitives per leaf at most (<= 8: three bits of the leaf code)
};

// The build proper.  prep(lo4, hi4, idx, seg) fills boxes / identity order / initial segment ids on the stream;
// finish(order) receives the device array order[new position] = old position while the scratch is still alive.
template <class Prep, class Finish>
int build_core(zb_context* ctx, const CoreParams& cp, zb_bvh_node* nodes_out, uint32_t* n_nodes, Prep prep, Finish finish) {
    cudaStream_t st = ctx->stream;
    const uint32_t n = cp.n;
    Scratch sc;
    const uint32_t max_segs = n / (SB_SMALL + 1) + 2;
    const unsigned nb = (n + 255) / 256, tiles = (n + SB_CHUNK - 1) / SB_CHUNK;
    float4 *lo4[2], *hi4[2];
    uint32_t *idx[2], *flags, *sums, *red, *bins, *counters, *out_idx;
    int* seg[2];
    Seg* segs[2];
    SmallSeg* small;
    zb_bvh_node* stage;
    ZB_CUDA(sc.reserve(ctx, size_t(n) * (2 * (16 + 16 + 4 + 4) + 4 + 4 + 16 + 64) + size_t(max_segs) * (2 * sizeof(Seg) + 48 + 3 * SB_BINS * 32) +
                       (size_t(n) / SB_CHUNK + 2) * 4 + 32 * 256));
    for (int k = 0; k < 2; k++) {
        ZB_CUDA(sc.get(&lo4[k], n)); ZB_CUDA(sc.get(&hi4[k], n)); ZB_CUDA(sc.get(&idx[k], n)); ZB_CUDA(sc.get(&seg[k], n));
        ZB_CUDA(sc.get(&segs[k], max_segs));
    }
    ZB_CUDA(sc.get(&flags, size_t(n) + 1)); ZB_CUDA(sc.get(&sums, size_t(n) / SB_CHUNK + 2));
    ZB_CUDA(sc.get(&red, size_t(max_segs) * 12)); ZB_CUDA(sc.get(&bins, size_t(max_segs) * 3 * SB_BINS * 8));
    ZB_CUDA(sc.get(&counters, 4)); ZB_CUDA(sc.get(&out_idx, n)); ZB_CUDA(sc.get(&small, n)); ZB_CUDA(sc.get(&stage, n));
    ZB_CUDA(cudaMemsetAsync(stage, 0, size_t(n) * sizeof(zb_bvh_node), st));
    ZB_CUDA(cudaMemsetAsync(counters, 0, 4 * sizeof(uint32_t), st));
    prep(lo4[0], hi4[0], idx[0], seg[0]);
    ZB_CUDA(cudaGetLastError());
    uint32_t m = n > SB_SMALL ? 1u : 0u;
    if (m) {
        Seg root{};
        root.count = n;
        root.parent = -1;
        ZB_CUDA(cudaMemcpyAsync(segs[0], &root, sizeof(root), cudaMemcpyHostToDevice, st));
    } else {
        const SmallSeg root{0u, n, -1, 0};
        const uint32_t one = 1;
        ZB_CUDA(cudaMemcpyAsync(small, &root, sizeof(root), cudaMemcpyHostToDevice, st));
        ZB_CUDA(cudaMemcpyAsync(counters + 1, &one, sizeof(one), cudaMemcpyHostToDevice, st));
    }
    ZB_CUDA(cudaStreamSynchronize(st));  // the two small host structs above live on this frame
    int cur = 0, launches = 2, levels = 0;
    const bool timing = ctx->tuning.build_timing != 0;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    if (timing) {
        for (auto& e : ev) cudaEventCreate(&e);
        cudaEventRecord(ev[0], st);
    }
    for (int level = 0; m > 0; level++, levels++) {
        if (level > 2 * ZB_MAX_BLAS_DEPTH) return fail(ZB_ERR_UNSUPPORTED, "device SAH build: level loop did not finish");
        const size_t init = size_t(m) * 3 * SB_BINS * 8;
        k_init_level<<<unsigned((init + 255) / 256), 256, 0, st>>>(red, bins, m);
        k_seg_reduce<<<tiles, 256, 0, st>>>(lo4[cur], hi4[cur], seg[cur], n, red);
        k_seg_prepare<<<(m + 255) / 256, 256, 0, st>>>(segs[cur], m, red, stage);
        k_bin<<<tiles, 256, 0, st>>>(lo4[cur], hi4[cur], seg[cur], n, segs[cur], bins);
        ZB_CUDA(cudaMemsetAsync(counters, 0, sizeof(uint32_t), st));
        k_decide<<<(m + 127) / 128, 128, 0, st>>>(segs[cur], m, bins, stage, segs[cur ^ 1], counters, small, cp.depth_cap);
        k_flags<<<(n + 256) / 256, 256, 0, st>>>(lo4[cur], hi4[cur], seg[cur], n, segs[cur], flags);
        ZB_CUDA(cudaGetLastError());
        if (int rc = scan_in_place(flags, n + 1, sums, st)) return rc;
        k_scatter<<<nb, 256, 0, st>>>(lo4[cur], hi4[cur], idx[cur], seg[cur], n, segs[cur], flags, lo4[cur ^ 1], hi4[cur ^ 1], idx[cur ^ 1],
                                      seg[cur ^ 1]);
        ZB_CUDA(cudaGetLastError());
        ZB_CUDA(cudaMemcpyAsync(&m, counters, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        ZB_CUDA(cudaStreamSynchronize(st));
        if (m > max_segs) return fail(ZB_ERR_CUDA, "device SAH build: segment table overflow");
        cur ^= 1;
        launches += 10;
    }
    uint32_t n_small = 0;
    ZB_CUDA(cudaMemcpyAsync(&n_small, counters + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    ZB_CUDA(cudaStreamSynchronize(st));
    if (timing) cudaEventRecord(ev[1], st);
    k_small<<<(n_small + 3) / 4, 128, 0, st>>>(lo4[cur], hi4[cur], idx[cur], small, n_small, stage, cp.leaf_first, out_idx, counters,
                                               cp.depth_cap, cp.isect_cost, cp.leaf_max);
    if (timing) cudaEventRecord(ev[2], st);
    k_mark<<<(n + 256) / 256, 256, 0, st>>>(stage, n, flags);
    ZB_CUDA(cudaGetLastError());
    if (int rc = scan_in_place(flags, n + 1, sums, st)) return rc;
    uint32_t info[4] = {0, 0, 0, 0}, total = 0;
    ZB_CUDA(cudaMemcpyAsync(info, counters, sizeof(info), cudaMemcpyDeviceToHost, st));
    ZB_CUDA(cudaMemcpyAsync(&total, flags + n, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    ZB_CUDA(cudaStreamSynchronize(st));
    if (total == 0 || total > std::max(n - 1, 1u)) return fail(ZB_ERR_CUDA, "device SAH build: inconsistent node count");
    if (total > cp.max_nodes)
        return fail(ZB_ERR_UNSUPPORTED, std::string("device SAH build: the ") + cp.what + " needs " + std::to_string(total) + " nodes, room for " +
                                            std::to_string(cp.max_nodes));
    if (int(info[3]) > cp.depth_cap)
        return fail(ZB_ERR_UNSUPPORTED, std::string("device SAH tree of a ") + cp.what + " is " + std::to_string(info[3]) + " levels deep (cap " +
                                            std::to_string(cp.depth_cap) + ")");
    k_compact<<<nb, 256, 0, st>>>(stage, n, flags, info[2], int(cp.node_base), nodes_out);
    finish(out_idx);
    ZB_CUDA(cudaGetLastError());
    if (timing) cudaEventRecord(ev[3], st);
    ZB_CUDA(cudaStreamSynchronize(st));  // the caller may free what prep / finish read
    if (timing) {
        float a = 0, b = 0, c = 0;
        cudaEventElapsedTime(&a, ev[0], ev[1]);
        cudaEventElapsedTime(&b, ev[1], ev[2]);
        cudaEventElapsedTime(&c, ev[2], ev[3]);
        std::fprintf(stderr, "[zb build] device SAH (%s): %u primitives, %d levels %.3f ms, %u warp segments %.3f ms, compact + permute %.3f ms, %u nodes, depth %u\n",
                     cp.what, n, levels, a, n_small, b, c, total, info[3]);
        for (auto& e : ev) cudaEventDestroy(e);
    }
    ctx->stats.launches += launches + 7;
    *n_nodes = total;
    return ZB_OK;
}

}  // namespace

// Same contract as zbk_build_blas_device (lbvh.cu): tris_in / nrm_in hold n triangles (3 float4 each) in any order;
// tris_out / nrm_out receive them in leaf order (global position tri_first); nodes_out has room for max(n - 1, 1) nodes,
// child indices are absolute (node_base + local), the root is nodes_out[0].
int zbk_build_blas_device_sah(zb_context* ctx, const float4* tris_in, const float4* nrm_in, uint32_t n, uint32_t tri_first,
                              uint32_t node_base, float4* tris_out, float4* nrm_out, zb_bvh_node* nodes_out, uint32_t* n_nodes) {
    cudaStream_t st = ctx->stream;
    if (n == 1) {
        k_single<<<1, 1, 0, st>>>(tris_in, tri_first, nodes_out);
        ZB_CUDA(cudaMemcpyAsync(tris_out, tris_in, 3 * sizeof(float4), cudaMemcpyDeviceToDevice, st));
        ZB_CUDA(cudaMemcpyAsync(nrm_out, nrm_in, 3 * sizeof(float4), cudaMemcpyDeviceToDevice, st));
        ZB_CUDA(cudaGetLastError());
        ctx->stats.launches += 1;
        *n_nodes = 1;
        return ZB_OK;
    }
    CoreParams cp{n, tri_first, node_base, ZB_MAX_BLAS_DEPTH, 1.0f, std::max(n - 1, 1u), "mesh"};
    // measurement knobs (tuning "blas_isect_pct", "blas_leaf"); the defaults are the host builder's
    cp.isect_cost = 0.01f * float(std::max(1, ctx->tuning.blas_isect_pct));
    cp.leaf_max = std::max(1, std::min(8, ctx->tuning.blas_leaf));
    int rc = build_core(
        ctx, cp, nodes_out, n_nodes,
        [&](float4* lo4, float4* hi4, uint32_t* idx, int* seg) { k_prep<<<(n + 255) / 256, 256, 0, st>>>(tris_in, n, lo4, hi4, idx, seg); },
        [&](const uint32_t* order) { k_permute_tris<<<(3 * n + 255) / 256, 256, 0, st>>>(tris_in, nrm_in, order, n, tris_out, nrm_out); });
    if (!rc && ctx->tuning.build_check) rc = validate_tree(ctx, nodes_out, *n_nodes, node_base, tri_first, n, TriBoxes{tris_out, tri_first}, "mesh");
    return rc;
}

// TLAS over n items from their world boxes (2 float4 each), in place: nodes_out = the scene's TLAS region (root at 0, room
// for max_nodes), items and item_bounds (2 float4 per item both) are permuted into leaf order through a temporary copy.
// Same cost model as the flattener's host build (host/flatten.hpp: isect_cost 1.5, leaves of <= 4 items, depth cap
// ZB_MAX_TLAS_DEPTH).
int zbk_build_tlas_device(zb_context* ctx, float4* items, float4* item_bounds, uint32_t n, zb_bvh_node* nodes_out, uint32_t max_nodes,
                          uint32_t* n_nodes) {
    cudaStream_t st = ctx->stream;
    if (n == 0) { *n_nodes = 0; return ZB_OK; }
    float4* tmp = nullptr;
    ZB_CUDA(cudaMalloc(&tmp, size_t(n) * 4 * sizeof(float4)));
    cudaMemcpyAsync(tmp, items, size_t(n) * 2 * sizeof(float4), cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(tmp + 2 * size_t(n), item_bounds, size_t(n) * 2 * sizeof(float4), cudaMemcpyDeviceToDevice, st);
    int rc = ZB_OK;
    if (n == 1) {  // one leaf under a root whose right child is empty
        zb_bvh_node nd{};
        float4 b[2];
        cudaMemcpyAsync(b, item_bounds, sizeof(b), cudaMemcpyDeviceToHost, st);
        cudaStreamSynchronize(st);
        nd.lx_ly[0] = b[0].x; nd.lx_ly[1] = b[1].x; nd.lx_ly[2] = b[0].y; nd.lx_ly[3] = b[1].y;
        nd.rx_ry[0] = INFINITY; nd.rx_ry[1] = -INFINITY; nd.rx_ry[2] = INFINITY; nd.rx_ry[3] = -INFINITY;
        nd.lz_rz[0] = b[0].z; nd.lz_rz[1] = b[1].z; nd.lz_rz[2] = INFINITY; nd.lz_rz[3] = -INFINITY;
        nd.left = int(~0u);  // first 0, count 1
        nd.right = ZB_CHILD_EMPTY;
        cudaMemcpyAsync(nodes_out, &nd, sizeof(nd), cudaMemcpyHostToDevice, st);
        cudaStreamSynchronize(st);
        *n_nodes = 1;
    } else {
        const CoreParams cp{n, 0u, 0u, ZB_MAX_TLAS_DEPTH, 1.5f, max_nodes, "TLAS"};
        rc = build_core(
            ctx, cp, nodes_out, n_nodes,
            [&](float4* lo4, float4* hi4, uint32_t* idx, int* seg) { k_prep_boxes<<<(n + 255) / 256, 256, 0, st>>>(item_bounds, n, lo4, hi4, idx, seg); },
            [&](const uint32_t* order) {
                k_permute_pairs<<<(2 * n + 255) / 256, 256, 0, st>>>(tmp, order, n, items);
                k_permute_pairs<<<(2 * n + 255) / 256, 256, 0, st>>>(tmp + 2 * size_t(n), order, n, item_bounds);
            });
    }
    cudaStreamSynchronize(st);
    cudaFree(tmp);
    if (!rc && n > 1 && ctx->tuning.build_check) rc = validate_tree(ctx, nodes_out, *n_nodes, 0u, 0u, n, PairBoxes{item_bounds}, "TLAS");
    return rc;
}
