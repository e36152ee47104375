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

#define ZB_MAX_GROUP 16
struct zb_group;

struct zb_context {
    int device = 0;
    // multi-GPU group this context belongs to (group.cu): null for a plain single-GPU context until its first render.  A
    // context made by zb_context_create_multi is the leader (rank 0) and owns the other members.
    zb_group* group = nullptr;
    int group_rank = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int sm_count = 0, cc_major = 0, cc_minor = 0;
    size_t l2_bytes = 0;
    unsigned long long* counters = nullptr;  // 8 x u64 on the device
    unsigned long long* counters_host = nullptr;  // pinned copy: fetched with the frame's last copies, read after its last sync
    void* arena = nullptr;  // scene arena kept across zb_scene_destroy / zb_scene_upload (abi.cu)
    size_t arena_bytes = 0;
    bool arena_in_use = false;
    // scratch of the device BVH builds (sah_build.cu) and the triangle copies they read, kept across uploads: a dozen
    // cudaMalloc / cudaFree of hundreds of MB per upload cost more than the build.  Released by zb_context_destroy.
    struct cached_buf {
        void* p = nullptr;
        size_t bytes = 0;
        cudaError_t get(size_t need, void** out) {
            if (bytes < need) {
                cudaFree(p);
                p = nullptr;
                bytes = 0;
                const cudaError_t e = cudaMalloc(&p, need);
                if (e != cudaSuccess) return e;
                bytes = need;
            }
            *out = p;
            return cudaSuccess;
        }
        void release() { cudaFree(p); p = nullptr; bytes = 0; }
    } build_scratch, build_tmp;
    // scratch reused across renders
    float4* accum = nullptr;
    float4* accum_aux = nullptr;
    size_t accum_pixels = 0;
    float4* out_f4 = nullptr;
    // output buffers the caller registered (zb_context_register_output): page-locked, zb_render DMAs finished rows into them as doubles
    struct out_range { char* ptr; size_t bytes; };
    std::vector<out_range> registered_out;
    bool out_registered(const void* p, size_t bytes) const {
        for (const out_range& r : registered_out)
            if (static_cast<const char*>(p) >= r.ptr && static_cast<const char*>(p) + bytes <= r.ptr + r.bytes) return true;
        return false;
    }
    double* rgb_pack64 = nullptr;  // device stage of the direct path: the band as 3 doubles per pixel, per plane
    size_t rgb_pack64_doubles = 0;
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
    const float4* wide = nullptr;  // the scene's 8-wide tree on the device (null: none)
    struct refit_state {  // what zb_scene_refit needs (refit.cu); ready == false: the scene cannot be refitted
        bool ready = false;
        float4* inst_bounds = nullptr;   // object-space bounds of every instance's geometry
        float4* item_bounds = nullptr;   // world bounds of every item
        float4* node_bounds = nullptr;   // scratch: bounds of every TLAS node
        uint32_t* order = nullptr;       // TLAS nodes grouped by level, deepest first
        std::vector<uint32_t> level_start;
        std::vector<zb_instance> host_instances;
        uint32_t n_tlas_nodes = 0, n_items = 0;
        uint32_t tlas_capacity = 0;      // nodes [0, tlas_capacity) belong to the TLAS (the flattener pads it to n_items - 1)
    } refit;
    std::vector<zb_bvh_node> device_mesh_roots;  // per mesh: root node (lo / hi boxes) of a BLAS built on the device
    std::vector<zb_scene*> replicas;  // scene of a multi-device context: copies on the other members
    SceneDev dev{};
    std::vector<void*> allocs;
    void* arena = nullptr;  // one allocation holding the scene arrays (abi.cu); the context's cached arena when arena_from_ctx
    size_t arena_bytes = 0, arena_used = 0;
    bool arena_from_ctx = false;
    uint32_t env_hdr_texture = ZB_NO_TEXTURE;
    size_t bytes = 0;
};


// abi_render.cu helpers shared with the other translation units
int zbr_fill_params(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                    const zb_render_opts* o, RenderParams& P);
int zbr_owned_tiles(const RenderParams& P);

int zbr_read_counters(zb_context* ctx);
int zbr_fetch_counters_async(zb_context* ctx);  // enqueue the copy into counters_host ...
void zbr_take_counters(zb_context* ctx);        // ... and, after a stream sync, move it into ctx->stats
int zbr_check_overflow(zb_context* ctx, const char* who);

// One frame copy in flight: packed RGB floats on the device -> this context's pinned stage -> the caller's 3-double rows.
// zbr_copy_issue enqueues the chunked copies (+ one event per chunk) on the context stream; zbr_widen_all runs host
// threads over the chunks of any number of contexts as their events fire.
struct FrameCopy {
    zb_context* ctx = nullptr;
    std::vector<size_t> begin;      // chunk boundaries, in floats, relative to stage / dst
    const float* stage = nullptr;   // pinned
    double* dst = nullptr;          // caller memory (pageable or shared)
    size_t first_event = 0;
};
int zbr_copy_issue(zb_context* ctx, const float* dev_rgb, size_t n_floats, double* dst, size_t stage_offset_floats,
                   std::vector<FrameCopy>& jobs);
int zbr_widen_all(std::vector<FrameCopy>& jobs);
int zbr_ensure_pinned(zb_context* ctx, size_t bytes);

// ---- group.cu: a set of n GPUs that renders one frame together
struct zb_group {
    int n = 1;                       // ranks in the group
    bool multi_process = false;      // members are separate processes (CUDA IPC handles, host barrier in shared memory)
    std::vector<zb_context*> local;  // members owned by this process: all n (zb_context_create_multi) or one (zb_group_join)
    // per rank, double-buffered by frame parity: fixed-point sums (this process's view: own pointer or peer mapping)
    unsigned long long* fix[2][ZB_MAX_GROUP] = {};
    size_t fix_words = 0;            // capacity of every buffer, in u64 words (sums + queue heads)
    cudaEvent_t render_done[ZB_MAX_GROUP] = {};  // recorded by rank r after its render kernels (interprocess when multi_process)
    unsigned long long frame = 0;
    void* shm = nullptr;             // multi-process: host barrier block in POSIX shared memory
    size_t shm_bytes = 0;
    std::string shm_name;
    bool opened[ZB_MAX_GROUP] = {};  // which peer mappings / events this process opened (to close them)
};
int zbg_render(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
               const zb_render_opts* opts, zb_outputs* out, const volatile int* cancel, zb_progress_fn progress, void* user);
bool zbg_fast_path(const zb_context* ctx, const zb_render_opts* opts, const zb_outputs* out);
void zbg_release(zb_context* ctx);

// refit.cu
int zbr_prepare_refit(zb_scene* s, const zb_flat_scene* f, const zb_instance* instances);
