// lbvh.cu — GPU construction of a mesh BLAS: replaces the host build the reference repeats on
// every GUI restart (bvh.hpp:11-44 via model.hpp:95, main.cpp:1492-1500) for meshes the flattener marked
// "build on device" (zb_mesh::node_count == 0).
//
// Linear BVH (Lauterbach 2009, Karras 2012): 30-bit Morton codes of the triangle centroids, an LSD radix sort of our own
// (k_sort_hist / k_sort_scan / k_sort_scatter below), one thread per internal node finds its key range
// and split, a bottom-up pass fits the boxes, and an emit pass writes the renderer's 64-byte nodes (both child boxes
// in the parent): subtrees of <= 4 triangles become one leaf, because a subtree of the sorted order is a contiguous
// triangle range.  Triangles and normals are permuted into sorted order.  Closest-hit results do not depend on the
// tree, so parity tests are the renderer's own (ids / hit records / radiance), run on a scene built this way.
//
// All passes stream over n elements: HBM-bound.  Bytes per triangle: bounds 48 read; morton 48 + 8; sort 4 passes x
// 16; hierarchy ~40; fit ~72; emit ~64 per alive node; permute 192 read + 192 written.
#include <algorithm>

#include "abi_internal.h"
#include "vecmath.cuh"
#include "build_util.cuh"

namespace {

#define LBVH_LEAF_MAX 4

// pass 1: bounds of all centroids (6 ordered uints: lo.xyz, hi.xyz)
__global__ void __launch_bounds__(256) k_centroid_bounds(const float4* __restrict__ tris, uint32_t n, uint32_t* __restrict__ bounds) {
    float3 lo = f3(INFINITY, INFINITY, INFINITY), hi = f3(-INFINITY, -INFINITY, -INFINITY);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4* t = tris + 3 * size_t(i);
        float3 c = (1.0f / 3.0f) * (f3(__ldg(t)) + f3(__ldg(t + 1)) + f3(__ldg(t + 2)));
        lo = f3(fminf(lo.x, c.x), fminf(lo.y, c.y), fminf(lo.z, c.z));
        hi = f3(fmaxf(hi.x, c.x), fmaxf(hi.y, c.y), fmaxf(hi.z, c.z));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo.x = fminf(lo.x, __shfl_xor_sync(0xFFFFFFFFu, lo.x, o)); lo.y = fminf(lo.y, __shfl_xor_sync(0xFFFFFFFFu, lo.y, o));
        lo.z = fminf(lo.z, __shfl_xor_sync(0xFFFFFFFFu, lo.z, o)); hi.x = fmaxf(hi.x, __shfl_xor_sync(0xFFFFFFFFu, hi.x, o));
        hi.y = fmaxf(hi.y, __shfl_xor_sync(0xFFFFFFFFu, hi.y, o)); hi.z = fmaxf(hi.z, __shfl_xor_sync(0xFFFFFFFFu, hi.z, o));
    }
    if ((threadIdx.x & 31) == 0 && lo.x <= hi.x) {
        atomicMin(bounds + 0, f2ord(lo.x)); atomicMin(bounds + 1, f2ord(lo.y)); atomicMin(bounds + 2, f2ord(lo.z));
        atomicMax(bounds + 3, f2ord(hi.x)); atomicMax(bounds + 4, f2ord(hi.y)); atomicMax(bounds + 5, f2ord(hi.z));
    }
}

// ---- own LSD radix sort of (30-bit Morton key, triangle id) pairs: four passes of 8 bits.  Per pass: per-tile digit
// histograms, one scan over (digit, tile) counts, a STABLE scatter.  A tile = 256 threads x 8 keys; inside a tile keys are
// ranked step by step (key j of every thread in step j): __match_any_sync groups the lanes of a warp with equal digits,
// shared-memory counters order the warps and carry the running count of every digit from step to step.
// HBM-bound: per pass 8 B read (histogram) + 8 B read + 8 B written (scatter) per pair.
#define SORT_THREADS 256
#define SORT_ITEMS 8
#define SORT_TILE (SORT_THREADS * SORT_ITEMS)

__global__ void __launch_bounds__(SORT_THREADS) k_sort_hist(const uint32_t* __restrict__ keys, uint32_t n, int shift, uint32_t n_tiles,
                                                            uint32_t* __restrict__ counts) {  // counts[digit * n_tiles + tile]
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * SORT_TILE;
#pragma unroll
    for (int j = 0; j < SORT_ITEMS; j++) {
        const uint32_t i = base + j * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    counts[threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan of counts[] (256 * n_tiles entries, digit-major) in place: one block, sequential chunks of 1024
__global__ void __launch_bounds__(1024) k_sort_scan(uint32_t* __restrict__ counts, uint32_t total) {
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t start = 0; start < total; start += 1024) {
        const uint32_t i = start + threadIdx.x;
        const uint32_t v = i < total ? counts[i] : 0u;
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
            warp_sum[threadIdx.x] = w;  // inclusive over warps
        }
        __syncthreads();
        const uint32_t before = carry + (threadIdx.x >= 32 ? warp_sum[(threadIdx.x >> 5) - 1] : 0u);
        if (i < total) counts[i] = before + x - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + x;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(SORT_THREADS) k_sort_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                               uint32_t n, int shift, uint32_t n_tiles, const uint32_t* __restrict__ offsets,
                                                               uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out) {
    constexpr int WARPS = SORT_THREADS / 32;
    __shared__ uint32_t running[256];          // next output slot of every digit for this tile
    __shared__ uint32_t warp_count[WARPS][256];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    running[threadIdx.x] = offsets[threadIdx.x * n_tiles + blockIdx.x];
    const uint32_t base = blockIdx.x * SORT_TILE;
    for (int j = 0; j < SORT_ITEMS; j++) {
#pragma unroll
        for (int w = 0; w < WARPS; w++) warp_count[w][threadIdx.x] = 0;
        __syncthreads();
        const uint32_t i = base + j * SORT_THREADS + threadIdx.x;
        const bool live = i < n;
        const uint32_t key = live ? keys_in[i] : 0xFFFFFFFFu;
        const uint32_t d = live ? (key >> shift) & 255u : 256u + lane;  // dead lanes match nobody
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, d);
        const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
        if (live && rank == 0) warp_count[warp][d] = __popc(peers);
        __syncthreads();
        if (live) {
            uint32_t at = running[d] + rank;
            for (uint32_t w = 0; w < warp; w++) at += warp_count[w][d];
            keys_out[at] = key;
            vals_out[at] = vals_in[i];
        }
        __syncthreads();
        uint32_t add = 0;
#pragma unroll
        for (int w = 0; w < WARPS; w++) add += warp_count[w][threadIdx.x];
        running[threadIdx.x] += add;
        __syncthreads();
    }
}

// sorts (keys, vals) by the low `bits` bits of the keys; result in (keys_out, vals_out) = the caller's second pair of
// buffers when the number of passes is odd, else back in (keys, vals): returns which.  counts: 256 * n_tiles words.
int radix_sort_pairs(uint32_t* keys, uint32_t* vals, uint32_t* keys2, uint32_t* vals2, uint32_t n, int bits, uint32_t* counts,
                     cudaStream_t st, bool* in_second) {
    const uint32_t n_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    uint32_t *ki = keys, *vi = vals, *ko = keys2, *vo = vals2;
    bool second = false;
    for (int shift = 0; shift < bits; shift += 8) {
        k_sort_hist<<<n_tiles, SORT_THREADS, 0, st>>>(ki, n, shift, n_tiles, counts);
        k_sort_scan<<<1, 1024, 0, st>>>(counts, 256u * n_tiles);
        k_sort_scatter<<<n_tiles, SORT_THREADS, 0, st>>>(ki, vi, n, shift, n_tiles, counts, ko, vo);
        ZB_CUDA(cudaGetLastError());
        std::swap(ki, ko);
        std::swap(vi, vo);
        second = !second;
    }
    *in_second = second;
    return ZB_OK;
}

ZB_DEV uint32_t expand10(uint32_t v) {  // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

// pass 2: Morton code of each centroid on a 1024^3 grid over the centroid bounds
__global__ void __launch_bounds__(256) k_morton(const float4* __restrict__ tris, uint32_t n, const uint32_t* __restrict__ bounds,
                                                uint32_t* __restrict__ keys, uint32_t* __restrict__ ids) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float3 lo = f3(ord2f(bounds[0]), ord2f(bounds[1]), ord2f(bounds[2]));
    const float3 hi = f3(ord2f(bounds[3]), ord2f(bounds[4]), ord2f(bounds[5]));
    const float4* t = tris + 3 * size_t(i);
    const float3 c = (1.0f / 3.0f) * (f3(__ldg(t)) + f3(__ldg(t + 1)) + f3(__ldg(t + 2)));
    // one scale for the three axes (a cube around the centroid bounds): on a thin mesh such as a terrain the short
    // axis then keeps its top bits constant and the tree splits along the two long ones — per-axis scaling made
    // every second level cut the sheet into an upper and a lower layer that overlap completely (measured 2.1x
    // node visits and 3.3x triangle tests per ray on the 1 M triangle heightfield)
    const float ext = fmaxf(hi.x - lo.x, fmaxf(hi.y - lo.y, hi.z - lo.z));
    const float sc = ext > 0.f ? 1024.0f / ext : 0.f;
    const uint32_t x = min(uint32_t(fmaxf((c.x - lo.x) * sc, 0.f)), 1023u), y = min(uint32_t(fmaxf((c.y - lo.y) * sc, 0.f)), 1023u),
                   z = min(uint32_t(fmaxf((c.z - lo.z) * sc, 0.f)), 1023u);
    keys[i] = (expand10(x) << 2) | (expand10(y) << 1) | expand10(z);
    ids[i] = i;
}

// length of the common prefix of the (key, index) pairs i and j; -1 outside the array (Karras 2012, section 4)
ZB_DEV int delta(const uint32_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint32_t a = keys[i], b = keys[j];
    return a != b ? __clz(a ^ b) : 32 + __clz(uint32_t(i) ^ uint32_t(j));
}

// pass 3: one thread per internal node i in [0, n-1): its range of sorted keys and its two children.
// child encoding: >= 0 internal node index, < 0 leaf ~k (sorted position k)
__global__ void __launch_bounds__(256) k_hierarchy(const uint32_t* __restrict__ keys, int n, int2* __restrict__ child,
                                                   int2* __restrict__ range, int* __restrict__ parent_int, int* __restrict__ parent_leaf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    const int left = first == gamma ? ~gamma : gamma;
    const int right = last == gamma + 1 ? ~(gamma + 1) : gamma + 1;
    child[i] = make_int2(left, right);
    range[i] = make_int2(first, last);
    if (left >= 0) parent_int[left] = i; else parent_leaf[~left] = i;
    if (right >= 0) parent_int[right] = i; else parent_leaf[~right] = i;
    if (i == 0) parent_int[0] = -1;
}

// pass 4: boxes bottom-up; the second thread to reach a node owns it (Karras 2012, section 5)
__global__ void __launch_bounds__(256) k_fit(const float4* __restrict__ tris, const uint32_t* __restrict__ ids, int n,
                                             const int2* __restrict__ child, const int* __restrict__ parent_int,
                                             const int* __restrict__ parent_leaf, float* __restrict__ leaf_box,
                                             float* __restrict__ node_box, int* __restrict__ visits, const int2* __restrict__ range,
                                             float* __restrict__ cost, int* __restrict__ collapse) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    TriBox b = tri_box(tris + 3 * size_t(ids[k]));
    float* lb = leaf_box + 6 * size_t(k);
    lb[0] = b.lo.x; lb[1] = b.lo.y; lb[2] = b.lo.z; lb[3] = b.hi.x; lb[4] = b.hi.y; lb[5] = b.hi.z;
    if (n < 2) return;
    int node = parent_leaf[k];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(visits + node, 1) == 0) return;  // the sibling subtree is not finished yet
        const int2 c = child[node];
        const float* l = c.x >= 0 ? node_box + 6 * size_t(c.x) : leaf_box + 6 * size_t(~c.x);
        const float* r = c.y >= 0 ? node_box + 6 * size_t(c.y) : leaf_box + 6 * size_t(~c.y);
        float* o = node_box + 6 * size_t(node);
        // volatile reads: the sibling's box was written by another thread before its atomicAdd
        const volatile float* vl = l;
        const volatile float* vr = r;
        o[0] = fminf(vl[0], vr[0]); o[1] = fminf(vl[1], vr[1]); o[2] = fminf(vl[2], vr[2]);
        o[3] = fmaxf(vl[3], vr[3]); o[4] = fmaxf(vl[4], vr[4]); o[5] = fmaxf(vl[5], vr[5]);
        // surface-area heuristic, bottom-up (traversal step = primitive test = 1, as the host builder counts): a subtree of
        // <= LBVH_LEAF_MAX triangles becomes ONE leaf only where that is cheaper than keeping its split.  (Collapsing every
        // such subtree, as round 1 did, doubled the triangle tests per ray against the SAH tree: 8.4 vs 4.2.)
        auto half_area = [](const volatile float* b) {
            const float dx = b[3] - b[0], dy = b[4] - b[1], dz = b[5] - b[2];
            return dx * dy + dy * dz + dz * dx;
        };
        const volatile float* vc = cost;
        const float cl = c.x >= 0 ? vc[c.x] : half_area(vl), cr = c.y >= 0 ? vc[c.y] : half_area(vr);
        const float a_node = half_area(o);
        const int2 rg = range[node];
        const int count = rg.y - rg.x + 1;
        const float split_cost = a_node + cl + cr, leaf_cost = float(count) * a_node;
        const bool leaf = count <= LBVH_LEAF_MAX && leaf_cost <= split_cost;
        cost[node] = leaf ? leaf_cost : split_cost;
        collapse[node] = leaf ? 1 : 0;
        node = parent_int[node];
    }
}

// pass 5a: an internal node survives as an inner node unless it or one of its ancestors was collapsed into a leaf
// (only subtrees of <= LBVH_LEAF_MAX triangles can be, so the walk up is at most three steps)
__global__ void __launch_bounds__(256) k_alive(const int2* __restrict__ range, const int* __restrict__ collapse,
                                               const int* __restrict__ parent_int, int n_int, int* __restrict__ alive) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_int) return;
    int dead = 0;
    for (int a = i; a > 0; a = parent_int[a]) {
        const int2 r = range[a];
        if (r.y - r.x + 1 > LBVH_LEAF_MAX) break;
        dead |= collapse[a];
    }
    alive[i] = (i == 0 || !dead) ? 1 : 0;
}

// depth of the emitted tree: surviving inner nodes above each sorted triangle, maximum over the mesh.  A Karras tree over
// 30-bit codes plus the index tie-break can be deeper than the traversal stack allows (ZB_MAX_BLAS_DEPTH); the build
// then fails loudly instead of handing the renderer a tree whose deferred children it would drop.
__global__ void __launch_bounds__(256) k_depth(const int* __restrict__ parent_int, const int* __restrict__ parent_leaf,
                                               const int* __restrict__ alive, int n, int* __restrict__ max_depth) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    int d = 0;
    if (k < n)
        for (int node = parent_leaf[k]; node >= 0; node = parent_int[node]) d += alive[node];
    d = __reduce_max_sync(0xFFFFFFFFu, d);
    if ((threadIdx.x & 31) == 0 && d > 0) atomicMax(max_depth, d);
}

// pass 5b: the renderer's 64-byte node (zb_bvh_node: both child boxes in the parent) for every surviving node
__global__ void __launch_bounds__(256) k_emit(const int2* __restrict__ child, const int2* __restrict__ range, const int* __restrict__ alive,
                                              const int* __restrict__ compact, const float* __restrict__ leaf_box,
                                              const float* __restrict__ node_box, int n_int, uint32_t tri_first, int node_base,
                                              zb_bvh_node* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_int || !alive[i]) return;
    const int2 c = child[i];
    int code[2];
    const float* box[2];
    const int cc[2] = {c.x, c.y};
#pragma unroll
    for (int k = 0; k < 2; k++) {
        if (cc[k] < 0) {
            code[k] = int(~(((tri_first + uint32_t(~cc[k])) << 3) | 0u));
            box[k] = leaf_box + 6 * size_t(~cc[k]);
        } else {
            const int2 r = range[cc[k]];
            box[k] = node_box + 6 * size_t(cc[k]);
            if (!alive[cc[k]]) code[k] = int(~(((tri_first + uint32_t(r.x)) << 3) | uint32_t(r.y - r.x)));  // collapsed: one leaf
            else code[k] = node_base + compact[cc[k]];
        }
    }
    zb_bvh_node nd;
    nd.lx_ly[0] = box[0][0]; nd.lx_ly[1] = box[0][3]; nd.lx_ly[2] = box[0][1]; nd.lx_ly[3] = box[0][4];
    nd.rx_ry[0] = box[1][0]; nd.rx_ry[1] = box[1][3]; nd.rx_ry[2] = box[1][1]; nd.rx_ry[3] = box[1][4];
    nd.lz_rz[0] = box[0][2]; nd.lz_rz[1] = box[0][5]; nd.lz_rz[2] = box[1][2]; nd.lz_rz[3] = box[1][5];
    nd.left = code[0];
    nd.right = code[1];
    nd.pad0 = nd.pad1 = 0;
    out[compact[i]] = nd;
}

// a mesh of one triangle has no internal node: one node whose left child is the leaf and whose right child is empty
__global__ void k_emit_single(const float* __restrict__ leaf_box, uint32_t tri_first, zb_bvh_node* __restrict__ out) {
    zb_bvh_node nd;
    nd.lx_ly[0] = leaf_box[0]; nd.lx_ly[1] = leaf_box[3]; nd.lx_ly[2] = leaf_box[1]; nd.lx_ly[3] = leaf_box[4];
    nd.rx_ry[0] = INFINITY; nd.rx_ry[1] = -INFINITY; nd.rx_ry[2] = INFINITY; nd.rx_ry[3] = -INFINITY;
    nd.lz_rz[0] = leaf_box[2]; nd.lz_rz[1] = leaf_box[5]; nd.lz_rz[2] = INFINITY; nd.lz_rz[3] = -INFINITY;
    nd.left = int(~((tri_first << 3) | 0u));
    nd.right = ZB_CHILD_EMPTY;
    nd.pad0 = nd.pad1 = 0;
    out[0] = nd;
}

// pass 6: triangles and their normals into sorted (leaf) order
__global__ void __launch_bounds__(256) k_permute(const float4* __restrict__ tris_in, const float4* __restrict__ nrm_in,
                                                 const uint32_t* __restrict__ ids, uint32_t n, float4* __restrict__ tris_out,
                                                 float4* __restrict__ nrm_out) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per float4
    if (t >= 3 * n) return;
    const uint32_t k = t / 3, j = t - 3 * k;
    const size_t src = 3 * size_t(ids[k]) + j;
    tris_out[t] = __ldg(tris_in + src);
    nrm_out[t] = __ldg(nrm_in + src);
}

struct Scratch {
    std::vector<void*> ptrs;
    ~Scratch() { for (void* p : ptrs) cudaFree(p); }
    template <class T> cudaError_t get(T** p, size_t count) {
        void* d = nullptr;
        cudaError_t e = cudaMalloc(&d, std::max<size_t>(count, 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(d);
        *p = static_cast<T*>(d);
        return e;
    }
};

}  // namespace

// Builds the BLAS of one mesh on the device.  tris_in / nrm_in: n triangles (3 float4 each) in any order;
// tris_out / nrm_out: the same triangles in leaf order (written at the same global position `tri_first`);
// nodes_out: room for max(n - 1, 1) nodes, child indices are absolute (node_base + local).  Returns the node count.
int zbk_build_blas_device(zb_context* ctx, const float4* tris_in, const float4* nrm_in, uint32_t n, uint32_t tri_first,
                          uint32_t node_base, float4* tris_out, float4* nrm_out, zb_bvh_node* nodes_out, uint32_t* n_nodes) {
    cudaStream_t st = ctx->stream;
    Scratch sc;
    const unsigned nb = (n + 255) / 256;
    uint32_t *bounds, *keys, *keys2, *ids, *ids2;
    float *leaf_box, *node_box;
    int2 *child, *range;
    int *parent_int, *parent_leaf, *visits, *alive, *compact, *max_depth, *collapse;
    float* cost;
    ZB_CUDA(sc.get(&bounds, 6));
    ZB_CUDA(sc.get(&keys, n)); ZB_CUDA(sc.get(&keys2, n)); ZB_CUDA(sc.get(&ids, n)); ZB_CUDA(sc.get(&ids2, n));
    ZB_CUDA(sc.get(&leaf_box, 6 * size_t(n))); ZB_CUDA(sc.get(&node_box, 6 * size_t(n)));
    ZB_CUDA(sc.get(&child, n)); ZB_CUDA(sc.get(&range, n));
    ZB_CUDA(sc.get(&parent_int, n)); ZB_CUDA(sc.get(&parent_leaf, n)); ZB_CUDA(sc.get(&visits, n));
    ZB_CUDA(sc.get(&alive, n)); ZB_CUDA(sc.get(&compact, n + 1)); ZB_CUDA(sc.get(&max_depth, 1));
    ZB_CUDA(sc.get(&collapse, n)); ZB_CUDA(sc.get(&cost, n));
    const uint32_t init[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u};
    ZB_CUDA(cudaMemcpyAsync(bounds, init, sizeof(init), cudaMemcpyHostToDevice, st));
    k_centroid_bounds<<<std::min(nb, unsigned(ctx->sm_count) * 8u), 256, 0, st>>>(tris_in, n, bounds);
    k_morton<<<nb, 256, 0, st>>>(tris_in, n, bounds, keys, ids);
    ZB_CUDA(cudaGetLastError());
    uint32_t* sort_counts;
    ZB_CUDA(sc.get(&sort_counts, 256 * size_t((n + SORT_TILE - 1) / SORT_TILE)));
    bool in_second = false;
    {
        int rc = radix_sort_pairs(keys, ids, keys2, ids2, n, 30, sort_counts, st, &in_second);
        if (rc) return rc;
    }
    if (!in_second) {  // the rest of the build reads the sorted pairs from (keys2, ids2)
        std::swap(keys, keys2);
        std::swap(ids, ids2);
    }
    k_permute<<<(3 * n + 255) / 256, 256, 0, st>>>(tris_in, nrm_in, ids2, n, tris_out, nrm_out);
    ZB_CUDA(cudaGetLastError());
    ctx->stats.launches += 4;
    if (n == 1) {
        ZB_CUDA(cudaMemsetAsync(visits, 0, sizeof(int), st));
        k_fit<<<1, 256, 0, st>>>(tris_in, ids2, 1, child, parent_int, parent_leaf, leaf_box, node_box, visits, range, cost, collapse);
        k_emit_single<<<1, 1, 0, st>>>(leaf_box, tri_first, nodes_out);
        ZB_CUDA(cudaGetLastError());
        *n_nodes = 1;
        return ZB_OK;
    }
    const int n_int = int(n) - 1;
    const unsigned nbi = (n_int + 255) / 256;
    ZB_CUDA(cudaMemsetAsync(visits, 0, size_t(n) * sizeof(int), st));
    k_hierarchy<<<nbi, 256, 0, st>>>(keys2, int(n), child, range, parent_int, parent_leaf);
    k_fit<<<nb, 256, 0, st>>>(tris_in, ids2, int(n), child, parent_int, parent_leaf, leaf_box, node_box, visits, range, cost, collapse);
    k_alive<<<nbi, 256, 0, st>>>(range, collapse, parent_int, n_int, alive);
    ZB_CUDA(cudaMemsetAsync(max_depth, 0, sizeof(int), st));
    k_depth<<<nb, 256, 0, st>>>(parent_int, parent_leaf, alive, int(n), max_depth);
    ZB_CUDA(cudaGetLastError());
    // compact[i] = number of surviving nodes before i (own exclusive scan, in place on a copy of alive[])
    ZB_CUDA(cudaMemsetAsync(alive + n_int, 0, sizeof(int), st));
    ZB_CUDA(cudaMemcpyAsync(compact, alive, size_t(n_int + 1) * sizeof(int), cudaMemcpyDeviceToDevice, st));
    k_sort_scan<<<1, 1024, 0, st>>>(reinterpret_cast<uint32_t*>(compact), uint32_t(n_int + 1));
    k_emit<<<nbi, 256, 0, st>>>(child, range, alive, compact, leaf_box, node_box, n_int, tri_first, int(node_base), nodes_out);
    ZB_CUDA(cudaGetLastError());
    ctx->stats.launches += 6;
    int total = 0, depth = 0;
    ZB_CUDA(cudaMemcpyAsync(&total, compact + n_int, sizeof(int), cudaMemcpyDeviceToHost, st));
    ZB_CUDA(cudaMemcpyAsync(&depth, max_depth, sizeof(int), cudaMemcpyDeviceToHost, st));
    ZB_CUDA(cudaStreamSynchronize(st));
    if (depth > ZB_MAX_BLAS_DEPTH)
        return fail(ZB_ERR_UNSUPPORTED, "device LBVH of a mesh is " + std::to_string(depth) + " levels deep (cap " +
                                            std::to_string(ZB_MAX_BLAS_DEPTH) + "): flatten without ZB_FLATTEN_GPU_BLAS for this scene");
    *n_nodes = uint32_t(total);
    return ZB_OK;
}
