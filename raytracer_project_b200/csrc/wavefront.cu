// wavefront.cu — wavefront form of the path tracer (alternative to the k_render megakernel, selected per render).
//
// Why: in the megakernel 22-34 % of the node-loop lane slots are lanes that hold a leaf and 17-32 % are lanes whose ray
// is finished (profiles/r01_traversal_experiments.txt), and no scheduling change inside a warp recovered them.  Here
// the path state lives in HBM / L2 and the work is cut into two kernels per bounce:
//   k_wf_trace  persistent warps pull rays from a compacted queue; a lane whose ray is finished writes the hit and takes
//               the next ray at once, so no lane waits for the slowest ray of its warp; no path state and no shading
//               code in registers (higher occupancy, smaller instruction footprint)
//   k_wf_shade  one thread per live path: hit record, emission, scatter, Russian roulette (camera.hpp:928-1004); a
//               finished path adds its radiance to the pixel's 64-bit fixed-point sum (same format as the megakernel's
//               tile sums, so the frame does not depend on the order paths finish in) and the slot takes the next
//               (pixel, sample) from a global head; live slots are appended to the next queue (warp-aggregated
//               atomics = stream compaction)
// Path order: 8x4-pixel tiles, inside a tile one sample of all 32 pixels, then the next sample — so 32 consecutive
// paths are the 32 pixels of a tile (coherent primary rays).
// Same RNG keys, estimator, termination rules and shading code (shade.cuh) as the megakernel.
// State per slot: 5 float4 = 80 B; per ray and bounce the two kernels move ~200 B, served from L2 when the slot count
// is kept below ~1 M.
#ifdef ZB_EXPERIMENTS  // rejected form, kept for A/B runs only (build.sh: ZB_EXPERIMENTS=1)
#include "shade.cuh"
#include "kernels.h"

#define WF_FIX_SCALE 4194304.0f  // = ZB_FIX_SCALE of kernels.cu
#define WF_FIX_INV (1.0 / 4194304.0)

struct WfState {
    float4* ray;    // [2*slot] o.xyz, pixel (bits); [2*slot+1] d.xyz, sample | segment << 24 (bits)
    float4* path;   // [2*slot] throughput.xyz; [2*slot+1] radiance of the current path.xyz
    float4* hit;    // t, item, prim (bits)
    unsigned long long* fix;   // 3 per pixel: fixed-point radiance sums
    unsigned long long* head;  // next path index to start
    uint32_t* queue[2];
    uint32_t* cnt;   // [3] rotating queue counts
    uint32_t* fetch; // [2] rotating fetch heads of k_wf_trace
    unsigned long long total_paths;
    uint32_t n_slots, tiles_x, tiles_y, n_s;
    uint32_t iter;   // bounce iteration: queue[iter & 1] / cnt[iter % 3] are read, queue[~iter & 1] / cnt[(iter+1) % 3] written
};

// K1 raygen: get_ray (camera.hpp:784-794), sample_square (:817), defocus disk (:822); same draws as k_render
ZB_DEV void wf_camera_ray(const RenderParams& P, const Philox& ph, int px, int py, uint32_t pixel, uint32_t s, float3& o, float3& d) {
    uint4 r0 = ph(pixel, s, 0u, 0u);
    d = P.cam.pix00_rel + (float(px) + u01(r0.x) - 0.5f) * P.cam.du + (float(py) + u01(r0.y) - 0.5f) * P.cam.dv;
    o = P.cam.center;
    if (P.cam.defocus) {
        float rr = sqrt_fast(u01(r0.z)), sn, cs;
        sincos_turn(u01(r0.w), &sn, &cs);
        float3 off = (rr * cs) * P.cam.disk_u + (rr * sn) * P.cam.disk_v;
        o += off;
        d = d - off;
    }
}

// Lanes with `need` start the next path of the frame (skipping tile pixels outside the image); returns false for a
// lane when the frame has no paths left.  Called by whole warps.
ZB_DEV bool wf_next_path(const RenderParams& P, const WfState& S, const Philox& ph, bool need, uint32_t& pixel, uint32_t& s,
                         float3& o, float3& d) {
    const unsigned lane = threadIdx.x & 31u;
    bool got = false;
    for (;;) {
        const unsigned m = __ballot_sync(0xFFFFFFFFu, need);
        if (!m) break;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(S.head, (unsigned long long)__popc(m));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (need) {
            const unsigned long long k = base + __popc(m & ((1u << lane) - 1u));
            if (k >= S.total_paths) {
                need = false;
            } else {
                const uint32_t in = uint32_t(k & 31u);
                const unsigned long long g = k >> 5;
                const uint32_t tile = uint32_t(g / S.n_s), sr = uint32_t(g - (unsigned long long)tile * S.n_s);
                const int px = int((tile % S.tiles_x) * 8u + (in & 7u)), py = int((tile / S.tiles_x) * 4u + (in >> 3));
                if (px < P.cam.width && py < P.cam.height) {
                    pixel = uint32_t(py) * uint32_t(P.cam.width) + uint32_t(px);
                    s = uint32_t(P.sample_begin) + sr;
                    wf_camera_ray(P, ph, px, py, pixel, s, o, d);
                    need = false;
                    got = true;
                }
            }
        }
    }
    return got;
}

// append `slot` to the next queue: one atomic per warp
ZB_DEV void wf_append(uint32_t* q, uint32_t* count, bool alive, uint32_t slot) {
    const unsigned m = __ballot_sync(0xFFFFFFFFu, alive);
    if (!m) return;
    const unsigned lane = threadIdx.x & 31u;
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(count, uint32_t(__popc(m)));
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (alive) q[base + __popc(m & ((1u << lane) - 1u))] = slot;
}

ZB_DEV void wf_store_path(const WfState& S, uint32_t slot, float3 o, float3 d, uint32_t pixel, uint32_t s, int seg, float3 beta, float3 L) {
    S.ray[2 * slot] = make_float4(o.x, o.y, o.z, __uint_as_float(pixel));
    S.ray[2 * slot + 1] = make_float4(d.x, d.y, d.z, __uint_as_float(s | (uint32_t(seg) << 24)));
    S.path[2 * slot] = make_float4(beta.x, beta.y, beta.z, 0.f);
    S.path[2 * slot + 1] = make_float4(L.x, L.y, L.z, 0.f);
}

// first queue: every slot starts a path (iteration 0 reads queue[0] / cnt[0])
__global__ void __launch_bounds__(128) k_wf_init(const __grid_constant__ RenderParams P, const WfState S) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    const Philox ph{P.pk[0]};
    uint32_t pixel = 0, s = 0;
    float3 o, d;
    const bool alive = wf_next_path(P, S, ph, slot < S.n_slots, pixel, s, o, d);
    if (alive) wf_store_path(S, slot, o, d, pixel, s, 0, f3(1, 1, 1), f3(0, 0, 0));
    wf_append(S.queue[0], S.cnt + 0, alive, slot);
}

// ---------------------------------------------------------------------------------- trace
#define WF_CHUNK 64u  // queue entries a warp claims per global atomic

#ifndef WF_TRACE_BLOCKS
#define WF_TRACE_BLOCKS 8  // resident blocks of 128 threads per SM the trace kernel is compiled for (<= 64 registers)
#endif

__global__ void __launch_bounds__(128, WF_TRACE_BLOCKS) k_wf_trace(const __grid_constant__ RenderParams P, const WfState S) {
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const uint32_t count = S.cnt[S.iter % 3u];
    const uint32_t* q_in = S.queue[S.iter & 1u];
    uint32_t* fetch = S.fetch + (S.iter & 1u);
    const Philox ph{P.pk[0]};
    int stack[ZB_STACK];
    Counters cnt = {0, 0, 0, 0};
    RayWalk w;
    w.cur = ZB_TRAVERSAL_DONE;
    w.sp = 0;
    w.in_blas = false;
    w.blas_item = 0;
    w.hit.t = 0.f;
    w.hit.item = 0xFFFFFFFFu;
    w.hit.prim = 0;
    bool have = false, dry = false;  // dry: the queue has nothing left for this lane
    uint32_t slot = 0, pixel = 0, sseg = 0;
    float3 o = f3(0, 0, 0), d = f3(0, 0, 1);
    uint32_t w_pos = 0, w_end = 0;   // the warp's claimed range of queue entries (warp-uniform)
    for (;;) {
        // ---- lanes without a ray take the next queue entries
        const unsigned need = __ballot_sync(0xFFFFFFFFu, !have && !dry);
        if (need) {
            const uint32_t want = uint32_t(__popc(need)), avail = w_end - w_pos;
            uint32_t base2 = 0;
            if (avail < want) {  // claim another chunk; leftovers of the old one are used first
                if (lane == 0) base2 = atomicAdd(fetch, WF_CHUNK);
                base2 = __shfl_sync(0xFFFFFFFFu, base2, 0);
            }
            if (!have && !dry) {
                const uint32_t k = uint32_t(__popc(need & lt));
                const uint32_t at = k < avail ? w_pos + k : base2 + (k - avail);
                if (at < count) {
                    slot = q_in[at];
                    const float4 ro = S.ray[2 * slot], rd = S.ray[2 * slot + 1];
                    o = f3(ro);
                    d = f3(rd);
                    pixel = __float_as_uint(ro.w);
                    sseg = __float_as_uint(rd.w);
                    w.begin(P.sc, o, d, INFINITY, stack);
                    have = true;
                } else {
                    dry = true;
                }
            }
            if (avail < want) {
                w_pos = base2 + (want - avail);
                w_end = base2 + WF_CHUNK;
            } else {
                w_pos += want;
            }
        }
        if (!__any_sync(0xFFFFFFFFu, have)) break;
        // ---- one while-while round: descend to the next leaf, test its primitives
        while (w.cur >= 0) w.node_step<false>(P.sc, 0.001f, stack, cnt);
        if (have) {
            if (w.cur != ZB_TRAVERSAL_DONE) {
                const MediumRng mr{ph, pixel, sseg & 0xFFFFFFu, (sseg >> 24) + 1u};
                w.leaf_step<false>(P.sc, o, d, 0.001f, mr, stack, cnt);
            }
            if (w.cur == ZB_TRAVERSAL_DONE) {
                S.hit[slot] = make_float4(w.hit.t, __uint_as_float(w.hit.item), __uint_as_float(w.hit.prim), 0.f);
                have = false;
            }
        }
    }
}

// ---------------------------------------------------------------------------------- shade
__global__ void __launch_bounds__(128) k_wf_shade(const __grid_constant__ RenderParams P, const WfState S) {
    const uint32_t count = S.cnt[S.iter % 3u];
    const uint32_t* q_in = S.queue[S.iter & 1u];
    uint32_t* q_out = S.queue[(S.iter & 1u) ^ 1u];
    uint32_t* n_out = S.cnt + (S.iter + 1u) % 3u;
    const Philox ph{P.pk[0]};
    if (blockIdx.x == 0 && threadIdx.x == 0) {  // counters of later iterations (nobody reads or writes them now)
        S.cnt[(S.iter + 2u) % 3u] = 0;
        S.fetch[(S.iter & 1u) ^ 1u] = 0;
    }
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; (i & ~31u) < count; i += stride) {
        bool alive = false, done = false;
        uint32_t slot = 0, pixel = 0, s = 0;
        int seg = 0;
        float3 o, d, beta, L;
        if (i < count) {
            slot = q_in[i];
            const float4 ro = S.ray[2 * slot], rd = S.ray[2 * slot + 1], bt = S.path[2 * slot], rl = S.path[2 * slot + 1];
            const float4 hh = S.hit[slot];
            o = f3(ro);
            d = f3(rd);
            beta = f3(bt);
            L = f3(rl);
            pixel = __float_as_uint(ro.w);
            s = __float_as_uint(rd.w) & 0xFFFFFFu;
            seg = int(__float_as_uint(rd.w) >> 24);
            Hit hit{hh.x, __float_as_uint(hh.y), __float_as_uint(hh.z)};
            // ---- K3 shade: one segment of ray_color (camera.hpp:928-1004), as in k_render
            if (hit.item == 0xFFFFFFFFu) {
                L += beta * background(P.sc, P.env, d);
                done = true;
            } else {
                Surface sf;
                make_surface(P.sc, hit, o, d, sf);
                L += beta * material_emitted(P.sc, sf);
                uint4 rb = ph(pixel, s, uint32_t(seg + 1), 0u);
                float3 atten = f3(1, 1, 1), no, nd;
                done = !material_scatter(P.sc, sf, d, f3(u01(rb.x), u01(rb.y), u01(rb.z)), atten, no, nd);
                if (!done) {
                    beta *= atten;
                    o = no;
                    d = nd;
                    if (seg >= 12) {  // camera.hpp:968-983
                        float pr = fminf(fmaxf(max3(beta.x, beta.y, beta.z), 0.05f), 0.95f);
                        if (dot(beta, beta) < 1e-8f || u01(rb.w) > pr) done = true;
                        beta *= 1.0f / pr;
                    }
                    if (++seg >= P.max_depth) done = true;
                }
            }
            alive = !done;
            if (done) {  // K5 accumulate: order-independent fixed-point sums
                unsigned long long* f = S.fix + 3 * size_t(pixel);
                atomicAdd(f + 0, __float2ull_rn(fmaxf(L.x, 0.f) * WF_FIX_SCALE));
                atomicAdd(f + 1, __float2ull_rn(fmaxf(L.y, 0.f) * WF_FIX_SCALE));
                atomicAdd(f + 2, __float2ull_rn(fmaxf(L.z, 0.f) * WF_FIX_SCALE));
            }
        }
        // regeneration: a finished slot starts the next path of the frame
        if (wf_next_path(P, S, ph, done, pixel, s, o, d)) {
            seg = 0;
            L = f3(0, 0, 0);
            beta = f3(1, 1, 1);
            alive = true;
        }
        if (alive) wf_store_path(S, slot, o, d, pixel, s, seg, beta, L);
        wf_append(q_out, n_out, alive, slot);
    }
}

// fixed-point sums -> the float4 accumulation buffer of the megakernel (xyz sums, w = samples)
__global__ void k_wf_resolve(const __grid_constant__ RenderParams P, const unsigned long long* fix, uint32_t n_s) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t n = size_t(P.cam.width) * P.cam.height;
    if (i >= n) return;
    float4 a = make_float4(float(double((long long)fix[3 * i]) * WF_FIX_INV), float(double((long long)fix[3 * i + 1]) * WF_FIX_INV),
                           float(double((long long)fix[3 * i + 2]) * WF_FIX_INV), float(n_s));
    if (P.accumulate) {
        const float4 old = P.accum[i];
        a.x += old.x; a.y += old.y; a.z += old.z; a.w += old.w;
    }
    P.accum[i] = a;
}

// ---------------------------------------------------------------------------------- host
struct zb_wf_scratch {
    WfState S = {};
    size_t slots = 0, pixels = 0;
    uint32_t* ctr = nullptr;   // cnt[3], fetch[2], pad, head (u64) : 8 words
    uint32_t* host_cnt = nullptr;  // pinned: 3 counts read back per batch
    cudaEvent_t ev[2] = {};
    int sm_count = 0;
};

void zbk_wavefront_free(zb_wf_scratch* w) {
    if (!w) return;
    cudaFree(w->S.ray);
    cudaFree(w->S.path);
    cudaFree(w->S.hit);
    cudaFree(w->S.fix);
    cudaFree(w->S.queue[0]);
    cudaFree(w->S.queue[1]);
    cudaFree(w->ctr);
    cudaFreeHost(w->host_cnt);
    for (auto& e : w->ev) if (e) cudaEventDestroy(e);
    delete w;
}

#define WF_TRY(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return e_; } while (0)

bool zbk_wavefront_supported(const RenderParams& P) {
    return !P.light_pass && !P.accum_aux && P.tile_ranks == 1 && P.max_depth > 0 && P.max_depth < 256 &&
           P.sample_end > P.sample_begin && P.sample_end < (1 << 24);
}

// One launch = the samples [sample_begin, sample_end) of the whole frame, like zbk_launch_render with all tiles owned.
cudaError_t zbk_launch_wavefront(const RenderParams& P, zb_wf_scratch** scratch, size_t want_slots, int* iterations, cudaStream_t stream) {
    if (!*scratch) {
        *scratch = new zb_wf_scratch();
        int dev = 0;
        WF_TRY(cudaGetDevice(&dev));
        WF_TRY(cudaDeviceGetAttribute(&(*scratch)->sm_count, cudaDevAttrMultiProcessorCount, dev));
        WF_TRY(cudaMalloc(&(*scratch)->ctr, 8 * sizeof(uint32_t)));
        WF_TRY(cudaMallocHost(&(*scratch)->host_cnt, 2 * 4 * sizeof(uint32_t)));
        for (auto& e : (*scratch)->ev) WF_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    zb_wf_scratch& w = **scratch;
    const size_t pixels = size_t(P.cam.width) * P.cam.height;
    const uint32_t tiles_x = uint32_t(P.cam.width + 7) / 8u, tiles_y = uint32_t(P.cam.height + 3) / 4u;
    const uint32_t n_s = uint32_t(P.sample_end - P.sample_begin);
    const unsigned long long total = (unsigned long long)tiles_x * tiles_y * 32ull * n_s;
    size_t slots = want_slots ? want_slots : (size_t(1) << 20);
    slots = size_t(std::min<unsigned long long>(slots, total));
    slots = (slots + 127) / 128 * 128;
    if (slots > w.slots) {
        cudaFree(w.S.ray); cudaFree(w.S.path); cudaFree(w.S.hit); cudaFree(w.S.queue[0]); cudaFree(w.S.queue[1]);
        w.S.ray = w.S.path = w.S.hit = nullptr;
        w.S.queue[0] = w.S.queue[1] = nullptr;
        w.slots = 0;
        WF_TRY(cudaMalloc(&w.S.ray, slots * 2 * sizeof(float4)));
        WF_TRY(cudaMalloc(&w.S.path, slots * 2 * sizeof(float4)));
        WF_TRY(cudaMalloc(&w.S.hit, slots * sizeof(float4)));
        WF_TRY(cudaMalloc(&w.S.queue[0], slots * sizeof(uint32_t)));
        WF_TRY(cudaMalloc(&w.S.queue[1], slots * sizeof(uint32_t)));
        w.slots = slots;
    }
    if (pixels > w.pixels) {
        cudaFree(w.S.fix);
        w.S.fix = nullptr;
        w.pixels = 0;
        WF_TRY(cudaMalloc(&w.S.fix, pixels * 3 * sizeof(unsigned long long)));
        w.pixels = pixels;
    }
    WfState S = w.S;
    S.cnt = w.ctr;
    S.fetch = w.ctr + 3;
    S.head = reinterpret_cast<unsigned long long*>(w.ctr + 6);
    S.total_paths = total;
    S.n_slots = uint32_t(slots);
    S.tiles_x = tiles_x;
    S.tiles_y = tiles_y;
    S.n_s = n_s;
    S.iter = 0;
    WF_TRY(cudaMemsetAsync(w.ctr, 0, 8 * sizeof(uint32_t), stream));
    WF_TRY(cudaMemsetAsync(S.fix, 0, pixels * 3 * sizeof(unsigned long long), stream));
    k_wf_init<<<unsigned(slots / 128), 128, 0, stream>>>(P, S);
    const unsigned trace_grid = unsigned(w.sm_count * WF_TRACE_BLOCKS), shade_grid = unsigned(w.sm_count * 8);
    // Bounce iterations are enqueued in batches; the live count of a batch is read back one batch late, so the
    // stream never drains while the host decides whether to go on.
    const int BATCH = 16;
    int iters = 0;
    for (int batch = 0;; batch++) {
        for (int k = 0; k < BATCH; k++, S.iter++, iters++) {
            k_wf_trace<<<trace_grid, 128, 0, stream>>>(P, S);
            k_wf_shade<<<shade_grid, 128, 0, stream>>>(P, S);
        }
        uint32_t* dst = w.host_cnt + 4 * (batch & 1);
        WF_TRY(cudaMemcpyAsync(dst, w.ctr, 3 * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
        WF_TRY(cudaEventRecord(w.ev[batch & 1], stream));
        if (batch == 0) continue;
        WF_TRY(cudaEventSynchronize(w.ev[(batch - 1) & 1]));
        const uint32_t* seen = w.host_cnt + 4 * ((batch - 1) & 1);
        // after iteration i the live count is in cnt[(i + 1) % 3]; S.iter at that copy was a multiple of BATCH
        const uint32_t it_then = uint32_t(batch) * BATCH;  // value of S.iter when batch-1's copy was enqueued
        if (seen[it_then % 3u] == 0) break;
    }
    k_wf_resolve<<<unsigned((pixels + 255) / 256), 256, 0, stream>>>(P, S.fix, n_s);
    if (iterations) *iterations = iters;
    return cudaGetLastError();
}

#endif  // ZB_EXPERIMENTS
