// noalloc.cu — zb_render_device_noalloc: the beauty pass of zb_render_device for callers that cannot allocate device
// memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// k_render_pq adds every finished path into 64-bit fixed-point per-pixel sums (3 words per pixel) and takes its work from
// a queue head; zb_render_device keeps both in a cudaMalloc'ed scratch buffer that grows with the frame.  Here the scratch
// is a file-scope __device__ array of a fixed, modest size (ZB_NOALLOC_MAX_PIXELS pixels: 24 MiB + 64 B, held by every GPU
// that loads the library until the process exits).  This entry point calls no cudaMalloc / cudaFree and no allocating
// library routine: parameter set-up on the host, two cudaMemsetAsync, two kernel launches, all on the context stream.
//
// The array is shared by every call on the device and keeps the last call's contents: each call clears the part it uses
// before its kernel runs (stream-ordered).  Two calls that overlap on one GPU (different streams / contexts) race on it.
#include <mutex>

#include "abi_internal.h"

namespace {

constexpr size_t NOALLOC_WORDS = 3 * size_t(ZB_NOALLOC_MAX_PIXELS) + 8;  // sums + queue heads
static __device__ __align__(256) unsigned long long g_noalloc_scratch[NOALLOC_WORDS];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address each device holds is asked
// for once per device, with that device current, and cached.
constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
unsigned long long* g_dev_ptr[MAX_DEVICES] = {};

int scratch_for_device(int device, unsigned long long** out) {
    if (device < 0 || device >= MAX_DEVICES) return fail(ZB_ERR_INVALID, "zb_render_device_noalloc: device index beyond 63");
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_dev_ptr[device]) {
        void* p = nullptr;
        ZB_CUDA(cudaSetDevice(device));
        ZB_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_dev_ptr[device] = static_cast<unsigned long long*>(p);
    }
    *out = g_dev_ptr[device];
    return ZB_OK;
}

}  // namespace

extern "C" size_t zb_noalloc_max_pixels(void) { return size_t(ZB_NOALLOC_MAX_PIXELS); }

extern "C" int zb_render_device_noalloc(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                                        const zb_render_opts* opts, void* accum_f4) {
    RenderParams P;
    int rc = zbr_fill_params(ctx, scene, cam, env, opts, P);
    if (rc) return rc;
    if (!accum_f4) return fail(ZB_ERR_INVALID, "zb_render_device_noalloc: accum_f4 is NULL");
    if (opts->light_pass || opts->collect_stats)
        return fail(ZB_ERR_UNSUPPORTED, "zb_render_device_noalloc: beauty pass only (no light_pass, no collect_stats)");
    if (P.max_depth <= 0 || P.sample_end <= P.sample_begin)
        return fail(ZB_ERR_INVALID, "zb_render_device_noalloc: needs max_depth >= 1 and a non-empty sample range");
    const size_t pixels = size_t(cam->image_width) * cam->image_height;
    if (pixels > size_t(ZB_NOALLOC_MAX_PIXELS))
        return fail(ZB_ERR_INVALID, "zb_render_device_noalloc: " + std::to_string(pixels) + " pixels exceed the static scratch of " +
                                        std::to_string(size_t(ZB_NOALLOC_MAX_PIXELS)) + " pixels (ZB_NOALLOC_MAX_PIXELS); use zb_render_device");
    unsigned long long* fix = nullptr;
    rc = scratch_for_device(ctx->device, &fix);
    if (rc) return rc;
    ZB_CUDA(cudaSetDevice(ctx->device));
    P.accum = static_cast<float4*>(accum_f4);
    P.accum_aux = nullptr;
    P.light_pass = 0;
    P.accumulate = opts->accumulate ? 1 : 0;
    unsigned long long* head = fix + 3 * pixels;
    ZB_CUDA(cudaMemsetAsync(ctx->counters, 0, 8 * sizeof(unsigned long long), ctx->stream));
    ZB_CUDA(zbk_pq_clear(fix, pixels, 0, head, 1, ctx->stream));  // the sums and the queue head must start at zero
    if (P.tile_ranks > 1 && !P.accumulate) ZB_CUDA(cudaMemsetAsync(P.accum, 0, pixels * sizeof(float4), ctx->stream));  // as zb_render_device
    ZB_CUDA(zbk_launch_render_pq(P, zbr_owned_tiles(P), 0, fix, head, ctx->sm_count, ctx->tuning, ctx->stream));
    ZB_CUDA(zbk_pq_resolve(P, 0, fix, 0, P.cam.height, 0.f, nullptr, ctx->stream));
    ctx->stats = zb_stats{};
    ctx->stats.launches = 2;
    return ZB_OK;
}
