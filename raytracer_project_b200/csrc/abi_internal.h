// abi_internal.h — state shared by the C-ABI translation units (abi.cu, abi_render.cu).
#pragma once
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/zenith_b200.h"
#include "kernels.h"

std::string& zb_err_slot();
inline int fail(int code, const std::string& msg) {
    zb_err_slot() = msg;
    return code;
}
#define ZB_CUDA(expr)                                                                                     \
    do {                                                                                                  \
        cudaError_t e__ = (expr);                                                                         \
        if (e__ != cudaSuccess)                                                                           \
            return fail(ZB_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));                \
    } while (0)

struct zb_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int sm_count = 0, cc_major = 0, cc_minor = 0;
    size_t l2_bytes = 0;
    unsigned long long* counters = nullptr;  // 8 x u64 on the device
    // scratch reused across renders
    float4* accum = nullptr;
    float4* accum_aux = nullptr;
    size_t accum_pixels = 0;
    float4* out_f4 = nullptr;
    float* rgb_pack = nullptr;  // packed RGB floats of the frame being copied to the host (d2h_rgb_f64)
    size_t rgb_pack_pixels = 0;
    void* pinned = nullptr;
    size_t pinned_bytes = 0;
    uint32_t* id_scratch = nullptr;
    float4* fin_a = nullptr;   // finishing chain scratch: two W*H planes (bloom) + a staging plane for host input
    float4* fin_b = nullptr;
    float4* fin_in = nullptr;
    size_t fin_pixels = 0;
    unsigned long long* fin_stats = nullptr;  // 2 + 128 u64: log sum, max, 256-bin histogram
    size_t id_pixels = 0;
    unsigned long long* pq_fix = nullptr;  // k_render_pq: fixed-point sums (3, 6 or 10 per pixel) + queue heads
    size_t pq_words = 0;
    struct zb_wf_scratch* wf = nullptr;  // wavefront render scratch (wavefront.cu)
    cudaEvent_t ev[6] = {};
    std::vector<cudaEvent_t> chunk_ev;  // one per 4 MiB chunk of the frame copy (abi_render.cu d2h_rgb_f64)
    zb_stats stats{};
    ZbTuning tuning{};  // A/B switches: environment at creation, zb_context_set_tuning afterwards
};

struct zb_scene {
    zb_context* ctx = nullptr;
    SceneDev dev{};
    std::vector<void*> allocs;
    uint32_t env_hdr_texture = ZB_NO_TEXTURE;
    size_t bytes = 0;
};


// abi_render.cu helpers shared with the other translation units
int zbr_fill_params(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                    const zb_render_opts* o, RenderParams& P);
int zbr_owned_tiles(const RenderParams& P);
