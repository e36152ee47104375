// group.cu — one frame on n GPUs behind the C ABI (the reference's counterpart is the static row split over
// std::thread inside the one blocking camera::render call, camera.hpp:404-414, :426-427, :557-573).
//
//   zb_context_create_multi   one process, n devices (peer access enabled): zb_scene_upload replicates the scene,
//                             zb_render partitions the frame and returns the whole image
//   zb_group_export / _join   one process per GPU (torchrun, MPI): buffers and events are shared through CUDA IPC
//                             handles the caller's own transport carries; zb_render is then collective
//
// Partition: sample slices (rank g renders samples [g spp/n, (g+1) spp/n) of every pixel) when each rank gets >= 8
// samples, else tile stripes (16x8 tile t belongs to rank t mod n).  Either way a rank leaves 64-bit fixed-point sums
// (2^-22 units, the path-queue kernel's native output) in its own HBM.  Exchange: rank g owns the image rows
// [g H/n, (g+1) H/n) and runs k_gather_resolve over them — one kernel that loads the n partial sums of each pixel from
// the n GPUs' buffers (its own directly, the others through peer-mapped pointers over NVLink), adds them AS INTEGERS,
// scales to means and packs RGB floats for the copy to the host, which every GPU does for its own rows over its own
// PCIe link.  Integer addition is associative, so the frame is bit-identical for every n and for both partitions.
// Ordering between ranks uses events only (no device-side spinning): rank r records render_done[r] behind its render
// kernels and every gather waits on all of them (cudaStreamWaitEvent; interprocess events across processes, where a host
// barrier in shared memory orders "recorded" before "waited on").  Buffers alternate by frame parity: a rank reaches
// frame k+2's render only after its gather k+1, which waited for every peer's render k+1, which follows that peer's
// gather k in stream order — so nobody still reads the buffer being cleared.
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <string>

#include "abi_internal.h"

namespace {

enum { PQ_BEAUTY = 0, PQ_AUX = 1, PQ_LIGHT = 2 };  // kernels.cu

struct GatherArgs {
    const unsigned long long* fix[ZB_MAX_GROUP];
    int n, mode;
    int width, height, row0, row1;
    int tile_ranks, tiles_x;  // tile stripes: only the owner of a pixel's tile holds sums for it
    float scale, aux_scale;
    float* out;               // planes x band pixels x 3 floats
    double* out64;            // not NULL: planes x band pixels x 3 DOUBLES instead (the same float values, widened): direct path
};

__device__ __forceinline__ float fixed_to_float(unsigned long long v) { return float(double((long long)v) * (1.0 / 4194304.0)); }

// One thread per pixel of the band.  Loads are plain 8-byte global loads: for a peer's buffer they travel over NVLink.
__global__ void __launch_bounds__(256) k_gather_resolve(const __grid_constant__ GatherArgs A) {
    const size_t band = size_t(A.row1 - A.row0) * A.width;
    const size_t t = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= band) return;
    const size_t pixels = size_t(A.width) * A.height, i = size_t(A.row0) * A.width + t;
    int g0 = 0, g1 = A.n;
    if (A.tile_ranks > 1) {
        const int px = int(i % A.width), py = int(i / A.width);
        g0 = ((py / ZB_TILE_H) * A.tiles_x + px / ZB_TILE_W) % A.tile_ranks;
        g1 = g0 + 1;
    }
    unsigned long long s[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    const int extra = A.mode == PQ_AUX ? 7 : A.mode == PQ_LIGHT ? 3 : 0;
    for (int g = g0; g < g1; g++) {
        const unsigned long long* f = A.fix[g];
#pragma unroll
        for (int k = 0; k < 3; k++) s[k] += f[3 * i + k];
        const unsigned long long* fa = f + 3 * pixels + size_t(extra) * i;
#pragma unroll
        for (int k = 0; k < 7; k++)
            if (k < extra) s[3 + k] += fa[k];
    }
    float v[12];
    int planes = 1;
    v[0] = fixed_to_float(s[0]) * A.scale;  // = k_pq_resolve followed by k_resolve: float(sum) * float(1 / spp)
    v[1] = fixed_to_float(s[1]) * A.scale;
    v[2] = fixed_to_float(s[2]) * A.scale;
    if (A.mode == PQ_LIGHT) {
        planes = 2;
        v[3] = fixed_to_float(s[3]) * A.scale;
        v[4] = fixed_to_float(s[4]) * A.scale;
        v[5] = fixed_to_float(s[5]) * A.scale;
    } else if (A.mode == PQ_AUX) {
        planes = 4;
#pragma unroll
        for (int k = 0; k < 6; k++) v[3 + k] = fixed_to_float(s[3 + k]) * A.aux_scale;  // albedo, normal
        v[9] = v[10] = v[11] = fixed_to_float(s[9]) * A.aux_scale;                      // z-depth
    }
#pragma unroll
    for (int p = 0; p < 4; p++) {
        if (p >= planes) break;
        if (A.out64) {  // the float the staged path would widen on the host, widened here
            double* o = A.out64 + (size_t(p) * band + t) * 3;
            o[0] = double(v[3 * p]);
            o[1] = double(v[3 * p + 1]);
            o[2] = double(v[3 * p + 2]);
        } else {
            float* o = A.out + (size_t(p) * band + t) * 3;
            o[0] = v[3 * p];
            o[1] = v[3 * p + 1];
            o[2] = v[3 * p + 2];
        }
    }
}

// ---- host barrier between the processes of a group: a sense-reversing counter in POSIX shared memory
struct ShmBarrier {
    std::atomic<uint32_t> count, generation, failed;
};

int host_barrier(zb_group* G) {
    if (!G->multi_process || G->n <= 1) return ZB_OK;
    ShmBarrier* b = static_cast<ShmBarrier*>(G->shm);
    if (b->failed.load()) return fail(ZB_ERR_CUDA, "group: a peer rank reported a failure");
    const uint32_t gen = b->generation.load(std::memory_order_acquire);
    if (b->count.fetch_add(1, std::memory_order_acq_rel) + 1 == uint32_t(G->n)) {
        b->count.store(0, std::memory_order_relaxed);
        b->generation.store(gen + 1, std::memory_order_release);
        return ZB_OK;
    }
    const auto t0 = std::chrono::steady_clock::now();
    for (unsigned spins = 0; b->generation.load(std::memory_order_acquire) == gen; spins++) {
        if ((spins & 1023u) == 1023u) {
            sched_yield();
            if (b->failed.load()) return fail(ZB_ERR_CUDA, "group: a peer rank reported a failure");
            if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(120)) {
                b->failed.store(1);
                return fail(ZB_ERR_CUDA, "group: timed out waiting for the other ranks (120 s)");
            }
        }
    }
    return ZB_OK;
}

void mark_failed(zb_group* G) {
    if (G && G->multi_process && G->shm) static_cast<ShmBarrier*>(G->shm)->failed.store(1);
}

size_t group_words(size_t pixels) { return zbk_pq_words(pixels, PQ_AUX) + 8; }  // the largest mode + queue heads

// single-process groups size their buffers on demand (sums of the mode in use + 8 queue heads); multi-process groups
// fixed theirs at zb_group_export
int ensure_local_fix(zb_group* G, size_t words) {
    if (G->fix_words >= words) return ZB_OK;
    if (G->multi_process) return fail(ZB_ERR_INVALID, "group: frame larger than the max_pixels given to zb_group_export");
    for (zb_context* m : G->local) {
        ZB_CUDA(cudaSetDevice(m->device));
        ZB_CUDA(cudaStreamSynchronize(m->stream));
        for (int b = 0; b < 2; b++) {
            cudaFree(G->fix[b][m->group_rank]);
            G->fix[b][m->group_rank] = nullptr;
            ZB_CUDA(cudaMalloc(&G->fix[b][m->group_rank], words * sizeof(unsigned long long)));
        }
    }
    G->fix_words = words;
    return ZB_OK;
}

struct Handle {  // contents of zb_group_handle
    uint32_t magic, device;
    uint64_t fix_words;
    int64_t pid;
    cudaIpcMemHandle_t mem[2];
    cudaIpcEventHandle_t event;
};
static_assert(sizeof(Handle) <= sizeof(zb_group_handle), "zb_group_handle too small");
constexpr uint32_t HANDLE_MAGIC = 0x5A424748u;  // "ZBGH"

int ensure_band_stage(zb_context* c, size_t floats) {
    if (c->rgb_pack_pixels * 3 < floats) {
        cudaFree(c->rgb_pack);
        c->rgb_pack = nullptr;
        c->rgb_pack_pixels = 0;
        ZB_CUDA(cudaMalloc(&c->rgb_pack, floats * sizeof(float)));
        c->rgb_pack_pixels = (floats + 2) / 3;
    }
    return zbr_ensure_pinned(c, floats * sizeof(float));
}

zb_group* solo_group(zb_context* ctx) {  // a plain context is a group of one (created on first use)
    if (ctx->group) return ctx->group;
    zb_group* G = new zb_group();
    G->n = 1;
    G->local.push_back(ctx);
    if (cudaEventCreateWithFlags(&G->render_done[0], cudaEventDisableTiming) != cudaSuccess) {
        delete G;
        return nullptr;
    }
    ctx->group = G;
    ctx->group_rank = 0;
    return G;
}

}  // namespace

// A plain context renders through the same code as a group of one when the request is what the path-queue kernel serves.
bool zbg_fast_path(const zb_context* ctx, const zb_render_opts* o, const zb_outputs* out) {
    if (ctx->group && ctx->group->n > 1) return true;
    return ctx->tuning.pq == 1 && !o->collect_stats && o->max_depth > 0 && !out->beauty_f4 &&
#ifdef ZB_EXPERIMENTS
           !ctx->tuning.wavefront &&
#endif
           true;
}

void zbg_release(zb_context* ctx) {
    zb_group* G = ctx->group;
    if (!G) return;
    if (G->multi_process) {
        for (int r = 0; r < G->n; r++) {
            if (!G->opened[r]) continue;
            for (int b = 0; b < 2; b++)
                if (G->fix[b][r]) cudaIpcCloseMemHandle(G->fix[b][r]);
            if (G->render_done[r]) cudaEventDestroy(G->render_done[r]);
        }
        const int me = ctx->group_rank;
        for (int b = 0; b < 2; b++) cudaFree(G->fix[b][me]);
        if (G->render_done[me]) cudaEventDestroy(G->render_done[me]);
        if (G->shm) munmap(G->shm, G->shm_bytes);
        if (me == 0 && !G->shm_name.empty()) shm_unlink(G->shm_name.c_str());
    } else {
        for (zb_context* m : G->local) {
            cudaSetDevice(m->device);
            for (int b = 0; b < 2; b++) cudaFree(G->fix[b][m->group_rank]);
            if (G->render_done[m->group_rank]) cudaEventDestroy(G->render_done[m->group_rank]);
        }
    }
    for (zb_context* m : G->local) m->group = nullptr;
    delete G;
}

extern "C" int zb_context_create_multi(const int* devices, int n_devices, zb_context** out) {
    if (!devices || !out || n_devices < 1 || n_devices > ZB_MAX_GROUP) return fail(ZB_ERR_INVALID, "zb_context_create_multi: bad argument");
    zb_group* G = new zb_group();
    G->n = n_devices;
    int rc = ZB_OK;
    for (int r = 0; r < n_devices && !rc; r++) {
        zb_context* c = nullptr;
        rc = zb_context_create(devices[r], &c);
        if (rc) break;
        c->group = G;
        c->group_rank = r;
        G->local.push_back(c);
        if (cudaEventCreateWithFlags(&G->render_done[r], cudaEventDisableTiming) != cudaSuccess) rc = fail(ZB_ERR_CUDA, "zb_context_create_multi: event");
    }
    for (int a = 0; a < int(G->local.size()) && !rc; a++) {
        cudaSetDevice(devices[a]);
        for (int b = 0; b < n_devices && !rc; b++) {
            if (a == b || devices[a] == devices[b]) continue;  // ranks may share a GPU (see the header)
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[a], devices[b]);
            if (!can) rc = fail(ZB_ERR_UNSUPPORTED, "zb_context_create_multi: devices cannot access each other's memory (no NVLink / P2P)");
            else {
                cudaError_t e = cudaDeviceEnablePeerAccess(devices[b], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else if (e != cudaSuccess) rc = fail(ZB_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
            }
        }
    }
    if (rc) {
        std::vector<zb_context*> made = G->local;
        if (!made.empty()) zbg_release(made[0]);
        else delete G;
        for (zb_context* c : made) zb_context_destroy(c);
        return rc;
    }
    *out = G->local[0];
    return ZB_OK;
}

extern "C" int zb_context_device_count(zb_context* ctx) { return ctx && ctx->group ? ctx->group->n : (ctx ? 1 : 0); }

extern "C" int zb_group_export(zb_context* ctx, size_t max_pixels, zb_group_handle* out) {
    if (!ctx || !out || max_pixels == 0) return fail(ZB_ERR_INVALID, "zb_group_export: bad argument");
    if (ctx->group && (ctx->group->n != 1 || ctx->group->multi_process)) return fail(ZB_ERR_INVALID, "zb_group_export: context already belongs to a group");
    ZB_CUDA(cudaSetDevice(ctx->device));
    if (ctx->group) {  // the group-of-one a plain render made
        ZB_CUDA(cudaStreamSynchronize(ctx->stream));
        zbg_release(ctx);
    }
    zb_group* G = new zb_group();
    G->multi_process = true;
    G->n = 0;  // set by zb_group_join
    G->local.push_back(ctx);
    G->fix_words = group_words(max_pixels);
    ctx->group = G;
    ctx->group_rank = 0;
    Handle h{};
    h.magic = HANDLE_MAGIC;
    h.device = uint32_t(ctx->device);
    h.fix_words = G->fix_words;
    h.pid = int64_t(getpid());
    cudaError_t e = cudaSuccess;
    // the own buffers are parked in slot 0 until zb_group_join knows this rank's number
    for (int b = 0; b < 2 && e == cudaSuccess; b++) {
        e = cudaMalloc(&G->fix[b][0], G->fix_words * sizeof(unsigned long long));
        if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h.mem[b], G->fix[b][0]);
    }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&G->render_done[0], cudaEventDisableTiming | cudaEventInterprocess);
    if (e == cudaSuccess) e = cudaIpcGetEventHandle(&h.event, G->render_done[0]);
    if (e != cudaSuccess) {
        for (int b = 0; b < 2; b++) cudaFree(G->fix[b][0]);
        if (G->render_done[0]) cudaEventDestroy(G->render_done[0]);
        ctx->group = nullptr;
        delete G;
        return fail(ZB_ERR_CUDA, std::string("zb_group_export: ") + cudaGetErrorString(e));
    }
    std::memset(out, 0, sizeof(*out));
    std::memcpy(out, &h, sizeof(h));
    return ZB_OK;
}

extern "C" int zb_group_join(zb_context* ctx, int rank, int n_ranks, const zb_group_handle* all, const char* shm_name) {
    if (!ctx || !all || !shm_name || n_ranks < 1 || n_ranks > ZB_MAX_GROUP || rank < 0 || rank >= n_ranks)
        return fail(ZB_ERR_INVALID, "zb_group_join: bad argument");
    zb_group* G = ctx->group;
    if (!G || !G->multi_process || G->n != 0) return fail(ZB_ERR_INVALID, "zb_group_join: call zb_group_export on this context first (once)");
    ZB_CUDA(cudaSetDevice(ctx->device));
    Handle mine;
    std::memcpy(&mine, &all[rank], sizeof(mine));
    if (mine.magic != HANDLE_MAGIC || mine.pid != int64_t(getpid()) || mine.device != uint32_t(ctx->device))
        return fail(ZB_ERR_INVALID, "zb_group_join: all[rank] is not the handle this context exported");
    // move the own buffers / event from the parking slot to this rank's slot
    unsigned long long* own[2] = {G->fix[0][0], G->fix[1][0]};
    cudaEvent_t own_ev = G->render_done[0];
    G->fix[0][0] = G->fix[1][0] = nullptr;
    G->render_done[0] = nullptr;
    G->fix[0][rank] = own[0];
    G->fix[1][rank] = own[1];
    G->render_done[rank] = own_ev;
    ctx->group_rank = rank;
    G->n = n_ranks;
    for (int r = 0; r < n_ranks; r++) {
        if (r == rank) continue;
        Handle h;
        std::memcpy(&h, &all[r], sizeof(h));
        if (h.magic != HANDLE_MAGIC || h.fix_words != G->fix_words)
            return fail(ZB_ERR_INVALID, "zb_group_join: a peer handle is malformed or was exported for another max_pixels");
        for (int b = 0; b < 2; b++) {
            void* p = nullptr;
            ZB_CUDA(cudaIpcOpenMemHandle(&p, h.mem[b], cudaIpcMemLazyEnablePeerAccess));
            G->fix[b][r] = static_cast<unsigned long long*>(p);
        }
        ZB_CUDA(cudaIpcOpenEventHandle(&G->render_done[r], h.event));
        G->opened[r] = true;
    }
    // host barrier block: a fresh POSIX shm object reads as zeros, whoever creates it
    G->shm_name = shm_name;
    G->shm_bytes = 4096;
    const int fd = shm_open(shm_name, O_CREAT | O_RDWR, 0600);
    if (fd < 0) return fail(ZB_ERR_IO, std::string("zb_group_join: shm_open ") + shm_name + " failed");
    if (ftruncate(fd, off_t(G->shm_bytes)) != 0) {
        close(fd);
        return fail(ZB_ERR_IO, "zb_group_join: ftruncate of the barrier block failed");
    }
    G->shm = mmap(nullptr, G->shm_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (G->shm == MAP_FAILED) {
        G->shm = nullptr;
        return fail(ZB_ERR_IO, "zb_group_join: mmap of the barrier block failed");
    }
    return host_barrier(G);
}

// Synchronise the context's stream, then meet the other ranks of a process group (host barrier in shared memory: a few
// microseconds, against ~0.1 ms for a collective-library barrier).  On any other context: the stream synchronisation only.
extern "C" int zb_group_barrier(zb_context* ctx) {
    if (!ctx) return fail(ZB_ERR_INVALID, "zb_group_barrier: ctx is NULL");
    ZB_CUDA(cudaSetDevice(ctx->device));
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->group && ctx->group->multi_process && ctx->group->n > 1) return host_barrier(ctx->group);
    return ZB_OK;
}

extern "C" int zb_group_leave(zb_context* ctx) {
    if (!ctx || !ctx->group) return ZB_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->group->multi_process && ctx->group->n > 1 && ctx->group->shm) host_barrier(ctx->group);  // nobody unmaps a buffer still being read
    zbg_release(ctx);
    return ZB_OK;
}

// opt-in phase timer of one frame (tuning "group_timing" / ZB_GROUP_TIMING=1): wall clock at the phase boundaries, printed by
// every rank to stderr
struct PhaseTimer {
    bool on;
    std::chrono::steady_clock::time_point t0, last;
    std::string line;
    explicit PhaseTimer(bool enabled) : on(enabled) { if (on) t0 = last = std::chrono::steady_clock::now(); }
    void mark(const char* what) {
        if (!on) return;
        const auto now = std::chrono::steady_clock::now();
        char buf[96];
        std::snprintf(buf, sizeof(buf), " %s %.3f", what, std::chrono::duration<double, std::milli>(now - last).count());
        line += buf;
        last = now;
    }
    void done(int rank) {
        if (!on) return;
        std::fprintf(stderr, "[zb group] rank %d:%s | total %.3f ms\n", rank, line.c_str(),
                     std::chrono::duration<double, std::milli>(last - t0).count());
    }
};

static int render_impl(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                       const zb_render_opts* opts, zb_outputs* out, const volatile int* cancel, zb_progress_fn progress, void* user);

// The frame of one render call on a group: beauty (+ first-hit AOVs), then — on more than one GPU — the light passes.
int zbg_render(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
               const zb_render_opts* opts, zb_outputs* out, const volatile int* cancel, zb_progress_fn progress, void* user) {
    const int rc = render_impl(ctx, scene, cam, env, opts, out, cancel, progress, user);
    if (rc != ZB_OK && rc != ZB_ERR_CANCELLED) mark_failed(ctx->group);  // the other processes of the group stop waiting for this one
    return rc;
}

static int render_impl(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                       const zb_render_opts* opts, zb_outputs* out, const volatile int* cancel, zb_progress_fn progress, void* user) {
    zb_group* G = ctx->group ? ctx->group : solo_group(ctx);
    if (!G) return fail(ZB_ERR_CUDA, "group: cannot create the render-done event");
    if (G->multi_process && G->n == 0) return fail(ZB_ERR_INVALID, "zb_render: zb_group_export without zb_group_join");
    const int n = G->n;
    if (n > 1 && (opts->collect_stats || out->beauty_f4 || opts->tile_ranks > 1 || opts->sample_begin != 0 ||
                  (opts->sample_end != 0 && opts->sample_end != opts->samples_per_pixel)))
        return fail(ZB_ERR_UNSUPPORTED, "zb_render on a multi-GPU context: collect_stats, beauty_f4 and manual sample / tile shares are single-GPU options");
    if (G->multi_process) {  // a cancel flag or a progress callback of one process cannot steer a collective call
        cancel = nullptr;
        progress = nullptr;
    }
    const int W = cam->image_width, H = cam->image_height;
    if (W < 1 || H < 1) return fail(ZB_ERR_INVALID, "camera: image size must be >= 1");
    const size_t pixels = size_t(W) * H;
    const bool aux = out->albedo || out->normal || out->z_depth;
    const bool light = n > 1 && ((opts->use_reflection && out->reflection) || (opts->use_refraction && out->refraction));
    int rc = ensure_local_fix(G, zbk_pq_words(pixels, aux ? PQ_AUX : light ? PQ_LIGHT : PQ_BEAUTY) + 8);
    if (rc) return rc;
    // partition: sample slices when every rank gets >= 8 samples, else tile stripes; opts->partition forces one
    const int spp = opts->samples_per_pixel;
    const bool by_tiles = n > 1 && (opts->partition == ZB_PARTITION_TILES || (opts->partition == ZB_PARTITION_AUTO && spp < 8 * n));

    struct Member {
        zb_context* c;
        const zb_scene* sc;
        RenderParams P;
        int row0, row1, s0, s1;
        uint64_t paths, rays;  // of the beauty pass (the light passes are not counted)
        uint32_t overflows;
    };
    std::vector<Member> M;
    for (zb_context* c : G->local) {
        Member m{};
        m.c = c;
        const int r = c->group_rank;
        m.sc = (c == scene->ctx) ? scene : nullptr;
        if (!m.sc)
            for (zb_scene* rep : scene->replicas)
                if (rep->ctx == c) m.sc = rep;
        if (!m.sc) return fail(ZB_ERR_INVALID, "zb_render: the scene was not uploaded through this multi-GPU context");
        rc = zbr_fill_params(c, m.sc, cam, env, opts, m.P);
        if (rc) return rc;
        if (!out->albedo) m.P.use_albedo = 0;
        if (!out->normal) m.P.use_normal = 0;
        if (!out->z_depth) m.P.use_z = 0;
        m.s0 = m.P.sample_begin;
        m.s1 = m.P.sample_end;
        if (n > 1) {
            if (by_tiles) {
                m.P.tile_rank = r;
                m.P.tile_ranks = n;
            } else {
                m.s0 = int((long long)r * spp / n);
                m.s1 = int((long long)(r + 1) * spp / n);
            }
        }
        m.row0 = int((long long)r * H / n);
        m.row1 = int((long long)(r + 1) * H / n);
        M.push_back(m);
    }
    const double scale = 1.0 / spp;
    const double aux_scale = 1.0 / std::min(M[0].P.aux_samples, spp);  // camera.hpp:535-541
    const bool gather_tiles = M[0].P.tile_ranks > 1;  // also a single GPU asked for one tile stripe: only the owner holds sums

    std::vector<FrameCopy> jobs;
    PhaseTimer pt(ctx->tuning.group_timing != 0);
    for (int pass = 0; pass < (light ? 2 : 1); pass++) {
        // beauty + AOVs: the queue kernel for the beauty sums, the AOV sums from a primary-ray pass of their own (k_aov_pass)
        // written as fixed point into the aux words; light passes: the queue kernel's LIGHT mode
        const int mode = pass == 1 ? PQ_LIGHT : PQ_BEAUTY;
        const bool aov_pass = pass == 0 && aux;
        const int gather_mode = aov_pass ? PQ_AUX : mode;
        const int buf = int(G->frame & 1);
        const size_t words = zbk_pq_words(pixels, gather_mode);
        int longest = 0;
        for (Member& m : M) longest = std::max(longest, m.s1 - m.s0);
        int per = opts->samples_per_launch > 0 ? opts->samples_per_launch : (cancel || progress ? 16 : std::max(longest, 1));
        per = std::max(per, 1);
        // ---- every local member: clear its sums, render its share in sample batches, record render_done
        for (Member& m : M) {
            zb_context* c = m.c;
            ZB_CUDA(cudaSetDevice(c->device));
            if (pass == 0) {
                c->stats = zb_stats{};
                ZB_CUDA(cudaMemsetAsync(c->counters, 0, 8 * sizeof(unsigned long long), c->stream));
                ZB_CUDA(cudaEventRecord(c->ev[0], c->stream));
            }
            ZB_CUDA(cudaMemsetAsync(G->fix[buf][c->group_rank], 0, (words + 8) * sizeof(unsigned long long), c->stream));
        }
        const int n_batches = std::max((longest + per - 1) / per, 1);
        for (int b = 0; b < n_batches; b++) {
            if (cancel && *cancel) return fail(ZB_ERR_CANCELLED, "render cancelled");
            for (Member& m : M) {
                zb_context* c = m.c;
                RenderParams P = m.P;
                P.sample_begin = std::min(m.s0 + b * per, m.s1);
                P.sample_end = std::min(P.sample_begin + per, m.s1);
                if (P.sample_end <= P.sample_begin) continue;
                ZB_CUDA(cudaSetDevice(c->device));
                unsigned long long* fix = G->fix[buf][c->group_rank];
                unsigned long long* head = fix + words + (b & 7);  // a fresh queue head per launch: 8 rotate, re-zeroed from the 9th on
                if (b >= 8) ZB_CUDA(cudaMemsetAsync(head, 0, sizeof(unsigned long long), c->stream));
                P.light_pass = mode == PQ_LIGHT ? 1 : 0;
                P.accum = nullptr;
                P.accum_aux = nullptr;
                ZB_CUDA(zbk_launch_render_pq(P, zbr_owned_tiles(P), mode, fix, head, c->sm_count, c->tuning, c->stream));
                c->stats.launches++;
                if (pass == 0) c->stats.kernel_pick = zbk_last_pq_pick();
                if (aov_pass && P.sample_begin < P.aux_samples) {
                    ZB_CUDA(zbk_launch_aov_pass_fixed(P, zbr_owned_tiles(P), fix + 3 * pixels, c->tuning, c->stream));
                    c->stats.launches++;
                }
            }
            if (progress || cancel) {
                for (Member& m : M) {
                    ZB_CUDA(cudaSetDevice(m.c->device));
                    ZB_CUDA(cudaStreamSynchronize(m.c->stream));
                }
                if (progress && pass == 0) progress(int((long long)H * std::min((b + 1) * per, longest) / std::max(longest, 1)), user);
            }
        }
        for (Member& m : M) {
            ZB_CUDA(cudaSetDevice(m.c->device));
            if (pass == 0) ZB_CUDA(cudaEventRecord(m.c->ev[1], m.c->stream));
            ZB_CUDA(cudaEventRecord(G->render_done[m.c->group_rank], m.c->stream));
        }
        pt.mark("launch");
        rc = host_barrier(G);  // every rank has recorded its event before anybody waits on it
        if (rc) return rc;
        pt.mark("barrier1");
        // ---- every local member: wait for all ranks' sums, gather + resolve its band, copy it home
        const int planes = gather_mode == PQ_AUX ? 4 : gather_mode == PQ_LIGHT ? 2 : 1;
        for (Member& m : M) {
            zb_context* c = m.c;
            ZB_CUDA(cudaSetDevice(c->device));
            for (int r = 0; r < n; r++)
                if (r != c->group_rank) ZB_CUDA(cudaStreamWaitEvent(c->stream, G->render_done[r], 0));
            const size_t band = size_t(m.row1 - m.row0) * W;
            if (band == 0) {  // more ranks than image rows
                if (pass == 0) ZB_CUDA(cudaEventRecord(c->ev[2], c->stream));
                continue;
            }
            double* dst[4] = {nullptr, nullptr, nullptr, nullptr};
            if (gather_mode == PQ_LIGHT) {
                dst[0] = opts->use_reflection ? out->reflection : nullptr;
                dst[1] = opts->use_refraction ? out->refraction : nullptr;
            } else {
                dst[0] = out->beauty;
                if (gather_mode == PQ_AUX) {
                    dst[1] = out->albedo;
                    dst[2] = out->normal;
                    dst[3] = out->z_depth;
                }
            }
            // direct path: every requested buffer was registered by the caller (zb_context_register_output) — the band
            // goes home as doubles in one DMA per plane, no pinned stage, no widening on the host
            bool direct = false;
            for (int p = 0; p < planes; p++)
                if (dst[p]) direct = true;
            for (int p = 0; p < planes; p++)
                if (dst[p] && !ctx->out_registered(dst[p] + size_t(m.row0) * W * 3, band * 3 * sizeof(double))) direct = false;
            if (direct) {
                const size_t need = size_t(planes) * band * 3;
                if (c->rgb_pack64_doubles < need) {
                    cudaFree(c->rgb_pack64);
                    c->rgb_pack64 = nullptr;
                    c->rgb_pack64_doubles = 0;
                    ZB_CUDA(cudaMalloc(&c->rgb_pack64, need * sizeof(double)));
                    c->rgb_pack64_doubles = need;
                }
            } else {
                rc = ensure_band_stage(c, size_t(planes) * band * 3);
                if (rc) return rc;
            }
            GatherArgs A{};
            for (int r = 0; r < n; r++) A.fix[r] = G->fix[buf][r];
            A.n = n;
            A.mode = gather_mode;
            A.width = W;
            A.height = H;
            A.row0 = m.row0;
            A.row1 = m.row1;
            A.tile_ranks = (by_tiles && n > 1) ? n : 1;
            A.tiles_x = m.P.tiles_x;
            A.scale = float(scale);
            A.aux_scale = float(aux_scale);
            A.out = c->rgb_pack;
            A.out64 = direct ? c->rgb_pack64 : nullptr;
            k_gather_resolve<<<unsigned((band + 255) / 256), 256, 0, c->stream>>>(A);
            ZB_CUDA(cudaGetLastError());
            c->stats.launches++;
            if (pass == 0) ZB_CUDA(cudaEventRecord(c->ev[2], c->stream));
            for (int p = 0; p < planes; p++) {
                if (!dst[p]) continue;
                double* home = dst[p] + size_t(m.row0) * W * 3;
                if (direct) {
                    ZB_CUDA(cudaMemcpyAsync(home, c->rgb_pack64 + size_t(p) * band * 3, band * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
                    continue;
                }
                rc = zbr_copy_issue(c, c->rgb_pack + size_t(p) * band * 3, band * 3, home, size_t(p) * band * 3, jobs);
                if (rc) return rc;
            }
        }
        (void)gather_tiles;
        G->frame++;
        pt.mark("issue");
        rc = zbr_widen_all(jobs);
        jobs.clear();
        if (rc) return rc;
        pt.mark("widen(waits for the kernel)");
        const bool last_pass = pass == (light ? 1 : 0);
        for (Member& m : M) {
            ZB_CUDA(cudaSetDevice(m.c->device));
            if (pass == 0) ZB_CUDA(cudaEventRecord(m.c->ev[3], m.c->stream));
            if (last_pass && (rc = zbr_fetch_counters_async(m.c)) != ZB_OK) return rc;  // rides on this sync: no round trip of its own
            ZB_CUDA(cudaStreamSynchronize(m.c->stream));
        }
        pt.mark("sync");
        if (pass == 0 && light) {  // keep the beauty pass's counters apart from the light passes'
            for (Member& m : M) {
                ZB_CUDA(cudaSetDevice(m.c->device));
                const uint32_t launches = m.c->stats.launches, pick = m.c->stats.kernel_pick;
                rc = zbr_read_counters(m.c);
                if (rc) return rc;
                m.paths = m.c->stats.paths;
                m.rays = m.c->stats.rays;
                m.overflows = m.c->stats.stack_overflows;
                m.c->stats.launches = launches;
                m.c->stats.kernel_pick = pick;
                ZB_CUDA(cudaMemsetAsync(m.c->counters, 0, 8 * sizeof(unsigned long long), m.c->stream));
            }
        }
        rc = host_barrier(G);  // the frame is complete in the caller's buffer of every rank
        if (rc) return rc;
        pt.mark("barrier2");
    }
    // ---- statistics: the sums over the local members (multi-process: this rank's own)
    zb_stats total{};
    PhaseTimer* ptp = &pt;
    for (Member& m : M) {
        zb_context* c = m.c;
        ZB_CUDA(cudaSetDevice(c->device));
        const uint32_t launches = c->stats.launches, pick = c->stats.kernel_pick;
        zbr_take_counters(c);  // fetched before the frame's last stream sync
        float ms = 0;
        ZB_CUDA(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
        total.kernel_ms = std::max(total.kernel_ms, double(ms));
        ZB_CUDA(cudaEventElapsedTime(&ms, c->ev[1], c->ev[2]));
        total.resolve_ms = std::max(total.resolve_ms, double(ms));
        ZB_CUDA(cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]));
        total.d2h_ms = std::max(total.d2h_ms, double(ms));
        total.paths += light ? m.paths : c->stats.paths;
        total.rays += light ? m.rays : c->stats.rays;
        total.launches += launches;
        total.stack_overflows += c->stats.stack_overflows + (light ? m.overflows : 0u);
        if (c == ctx || !total.kernel_pick) total.kernel_pick = pick;
    }
    ZB_CUDA(cudaSetDevice(ctx->device));
    total.n_gpus = uint32_t(n);
    ctx->stats = total;
    ptp->mark("stats");
    ptp->done(ctx->group_rank);
    return zbr_check_overflow(ctx, "zb_render");
}
