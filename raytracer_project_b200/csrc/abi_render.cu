// abi_render.cu — zb_render / zb_render_device / zb_resolve_device / zb_primary_ids / zb_get_stats.
// The drop-in for camera::render's hot part (camera.hpp:236-250 -> execute_render_threads :404-579).
#include <algorithm>
#include <atomic>
#include <thread>

#include "abi_internal.h"

void zbh_widen_f32_f64(const float* s, double* d, size_t n);  // host_glue.cpp

namespace {

struct d3 {
    double x, y, z;
};
d3 operator+(d3 a, d3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
d3 operator-(d3 a, d3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
d3 operator*(double s, d3 a) { return {s * a.x, s * a.y, s * a.z}; }
double dot(d3 a, d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
d3 cross(d3 a, d3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
d3 unit(d3 a) {
    double l = std::sqrt(dot(a, a));
    return l < 1e-8 ? d3{0, 0, 0} : (1.0 / l) * a;
}
d3 from(const double* p) { return {p[0], p[1], p[2]}; }
float3 to_f3(d3 a) { return make_float3(float(a.x), float(a.y), float(a.z)); }
double clampd(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }

// camera::initialize (camera.hpp:358-402), in double; narrowed to fp32 only at the end.
int setup_camera(const zb_camera* c, CameraDev& cam) {
    if (c->image_width < 1 || c->image_height < 1) return fail(ZB_ERR_INVALID, "camera: image size must be >= 1");
    if (double(c->image_width) * c->image_height > 2.0e9) return fail(ZB_ERR_INVALID, "camera: image too large");
    const double pi = 3.14159265358979323846;
    double aspect = double(c->image_width) / c->image_height;
    d3 lookfrom = from(c->lookfrom), lookat = from(c->lookat), vup = from(c->vup);
    double theta = c->vfov * pi / 180.0;
    double h = std::tan(theta / 2);
    double vh = 2 * h * c->focus_dist, vw = vh * aspect;
    d3 w = unit(lookfrom - lookat), u = unit(cross(vup, w)), v = cross(w, u);
    d3 vu = vw * u, vv = vh * (-1.0 * v);
    d3 du = (1.0 / c->image_width) * vu, dv = (1.0 / c->image_height) * vv;
    d3 upper_left = lookfrom - c->focus_dist * w - 0.5 * vu - 0.5 * vv;
    d3 pix00 = upper_left + 0.5 * (du + dv);
    double radius = c->focus_dist * std::tan((c->defocus_angle / 2) * pi / 180.0);
    d3 rel = pix00 - lookfrom;
    cam.center = to_f3(lookfrom);
    cam.pix00_rel = to_f3(rel);
    cam.du = to_f3(du);
    cam.dv = to_f3(dv);
    cam.disk_u = to_f3(radius * u);
    cam.disk_v = to_f3(radius * v);
    cam.u = to_f3(u);
    cam.v = to_f3(v);
    cam.w = to_f3(w);
    // the primary-ID kernel subtracts the fp32-rounded centre it really uses as the origin
    d3 cf = {double(cam.center.x), double(cam.center.y), double(cam.center.z)};
    d3 rel_id = pix00 - cf;
    cam.pix00_rel_d[0] = rel_id.x; cam.pix00_rel_d[1] = rel_id.y; cam.pix00_rel_d[2] = rel_id.z;
    cam.du_d[0] = du.x; cam.du_d[1] = du.y; cam.du_d[2] = du.z;
    cam.dv_d[0] = dv.x; cam.dv_d[1] = dv.y; cam.dv_d[2] = dv.z;
    cam.width = c->image_width;
    cam.height = c->image_height;
    cam.defocus = c->defocus_angle > 0 ? 1 : 0;
    cam.pad = 0;
    return ZB_OK;
}

// The ray-independent part of camera::get_background_color (camera.hpp:828-925), in double.
int setup_env(const zb_environment* e, const zb_scene* scene, EnvDev& env) {
    std::memset(&env, 0, sizeof(env));
    env.mode = e->mode;
    env.intensity = float(e->intensity);
    env.hdr_texture = ZB_NO_TEXTURE;
    if (e->mode == ZB_ENV_SOLID_COLOR) {
        env.background = to_f3(e->intensity * from(e->background_color));
    } else if (e->mode == ZB_ENV_HDR_MAP) {
        env.hdr_texture = e->hdr_texture != ZB_NO_TEXTURE ? e->hdr_texture : scene->env_hdr_texture;
        // yaw (Y), then tilt (X), then roll (Z): d3 = Rz * Rx * Ry * d   (camera.hpp:843-862)
        double cy = std::cos(e->hdri_rotation), sy = std::sin(e->hdri_rotation);
        double cp = std::cos(e->hdri_tilt), sp = std::sin(e->hdri_tilt);
        double cr = std::cos(e->hdri_roll), sr = std::sin(e->hdri_roll);
        double Ry[3][3] = {{cy, 0, sy}, {0, 1, 0}, {-sy, 0, cy}};
        double Rx[3][3] = {{1, 0, 0}, {0, cp, -sp}, {0, sp, cp}};
        double Rz[3][3] = {{cr, -sr, 0}, {sr, cr, 0}, {0, 0, 1}};
        double T[3][3], M[3][3];
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                T[i][j] = 0;
                for (int k = 0; k < 3; k++) T[i][j] += Rx[i][k] * Ry[k][j];
            }
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                M[i][j] = 0;
                for (int k = 0; k < 3; k++) M[i][j] += Rz[i][k] * T[k][j];
            }
        env.hdr_r0 = make_float3(float(M[0][0]), float(M[0][1]), float(M[0][2]));
        env.hdr_r1 = make_float3(float(M[1][0]), float(M[1][1]), float(M[1][2]));
        env.hdr_r2 = make_float3(float(M[2][0]), float(M[2][1]), float(M[2][2]));
    } else if (e->mode == ZB_ENV_PHYSICAL_SUN) {
        d3 sun = unit(from(e->sun_direction));
        double sun_height = sun.y, adj = sun_height - 0.05;
        double sky_exposure = clampd(adj * 8.0 + 1.4, 0.0, 1.0);
        double day = clampd(adj * 10.0 + 1.1, 0.0, 1.0);
        double sunset_int = clampd(1.0 - std::fabs(adj + 0.05) * 30.0, 0.0, 1.0);
        double sunset = adj > -0.1 ? sunset_int : 0.0;
        if (sun_height < 0) sunset *= (sun_height * 10.0 + 1.0);
        sunset = clampd(sunset, 0.0, 1.0);
        d3 zen = (1.0 - day) * d3{0.01, 0.03, 0.1} + day * d3{0.2, 0.5, 1.0};
        d3 hor = (1.0 - day) * d3{0.05, 0.02, 0.01} + day * d3{0.6, 0.8, 1.0};
        hor = (1.0 - sunset) * hor + sunset * d3{1.0, 0.35, 0.1};
        d3 s_color = (1.0 - sunset) * from(e->sun_color) + sunset * d3{1.0, 0.3, 0.1};
        double vis = clampd(sun_height * 5.0 + 1.0, 0.0, 1.0);
        env.sun_dir = to_f3(sun);
        env.zenith = to_f3(zen);
        env.horizon = to_f3(hor);
        env.sun_add = to_f3((e->sun_intensity * vis) * s_color);
        env.sky_scale = float(e->intensity * 1.5 * sky_exposure);
        env.sun_threshold = float(1.0 - e->sun_size * 0.001);
        env.sun_visible = adj > -0.1 ? 1 : 0;
    } else {
        return fail(ZB_ERR_INVALID, "environment: unknown mode");
    }
    return ZB_OK;
}

int ensure_scratch(zb_context* ctx, size_t pixels, bool aux) {
    if (ctx->accum_pixels < pixels) {
        cudaFree(ctx->accum);
        cudaFree(ctx->accum_aux);
        cudaFree(ctx->out_f4);
        ctx->accum = nullptr;
        ctx->accum_aux = nullptr;
        ctx->out_f4 = nullptr;
        ctx->accum_pixels = 0;
        ZB_CUDA(cudaMalloc(&ctx->accum, pixels * sizeof(float4)));
        ZB_CUDA(cudaMalloc(&ctx->out_f4, pixels * sizeof(float4)));
        ctx->accum_pixels = pixels;
    }
    if (aux && !ctx->accum_aux) ZB_CUDA(cudaMalloc(&ctx->accum_aux, 3 * ctx->accum_pixels * sizeof(float4)));
    return zbr_ensure_pinned(ctx, pixels * sizeof(float4));
}

int fill_params(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                const zb_render_opts* o, RenderParams& P) {
    if (!ctx || !scene || !cam || !env || !o) return fail(ZB_ERR_INVALID, "render: NULL argument");
    if (scene->ctx != ctx) return fail(ZB_ERR_INVALID, "render: scene belongs to another context");
    if (o->samples_per_pixel < 1 || o->max_depth < 0) return fail(ZB_ERR_INVALID, "render: bad spp / max_depth");
    std::memset(&P, 0, sizeof(P));
    P.sc = scene->dev;
    if (!ctx->tuning.wide) P.sc.wide = nullptr;
    P.exp_flags = ctx->tuning.pq_redeal;
    int rc = setup_camera(cam, P.cam);
    if (rc) return rc;
    rc = setup_env(env, scene, P.env);
    if (rc) return rc;
    P.counters = ctx->counters;
    P.seed_lo = uint32_t(o->seed);
    P.seed_hi = uint32_t(o->seed >> 32);
    static_assert(ZB_PHILOX_WORDS <= 24, "RenderParams::pk is too small");
    philox_expand_key(P.seed_lo, P.seed_hi, P.pk[0]);
    philox_expand_key(P.seed_lo, P.seed_hi ^ 0x4C495445u, P.pk[1]);  // the light passes draw their own stream
    int sb = o->sample_begin, se = o->sample_end;
    if (sb == 0 && se == 0) se = o->samples_per_pixel;
    if (sb < 0 || se < sb || se > o->samples_per_pixel) return fail(ZB_ERR_INVALID, "render: bad sample range");
    P.sample_begin = sb;
    P.sample_end = se;
    P.max_depth = o->max_depth;
    P.aux_samples = std::min(std::max(o->samples_per_pixel / 8, 64), 1024);  // camera.hpp:433
    P.use_albedo = o->use_albedo;
    P.use_normal = o->use_normal;
    P.use_z = o->use_z_depth;
    P.inv_z_max = o->z_depth_max_dist > 0 ? float(1.0 / o->z_depth_max_dist) : 0.f;
    P.tile_rank = o->tile_ranks > 1 ? o->tile_rank : 0;
    P.tile_ranks = o->tile_ranks > 1 ? o->tile_ranks : 1;
    if (P.tile_rank < 0 || P.tile_rank >= P.tile_ranks) return fail(ZB_ERR_INVALID, "render: bad tile rank");
    P.tiles_x = (cam->image_width + ZB_TILE_W - 1) / ZB_TILE_W;
    P.tiles_y = (cam->image_height + ZB_TILE_H - 1) / ZB_TILE_H;
    P.light_pass = o->light_pass ? 1 : 0;
    return ZB_OK;
}

constexpr int PQ_MAX_HEADS = 8;
int ensure_pq_scratch(zb_context* ctx, size_t pixels, int mode) {
    const size_t words = zbk_pq_words(pixels, mode) + PQ_MAX_HEADS;
    if (ctx->pq_words < words) {
        cudaFree(ctx->pq_fix);
        ctx->pq_fix = nullptr;
        ctx->pq_words = 0;
        ZB_CUDA(cudaMalloc(&ctx->pq_fix, words * sizeof(unsigned long long)));
        ctx->pq_words = words;
    }
    return ZB_OK;
}

int owned_tiles(const RenderParams& P) {
    int total = P.tiles_x * P.tiles_y;
    return (total - P.tile_rank + P.tile_ranks - 1) / P.tile_ranks;
}

// Launch this rank's share in batches of samples; accum is overwritten by the first batch.
int launch_share(zb_context* ctx, RenderParams P, const zb_render_opts* o, const volatile int* cancel,
                 zb_progress_fn progress, void* user, bool zero_first) {
    const int sb = P.sample_begin, se = P.sample_end;
    int per = o->samples_per_launch > 0 ? o->samples_per_launch : (cancel || progress ? 16 : se - sb);
    per = std::max(per, 1);
    const size_t pixels = size_t(P.cam.width) * P.cam.height;
    const bool keep = o->accumulate != 0;  // the caller's sums stay: progressive rendering / resume
    if (zero_first && !keep && (P.tile_ranks > 1 || se == sb)) {
        // tile stripes leave other ranks' pixels untouched: they must read as zero sums
        ZB_CUDA(cudaMemsetAsync(P.accum, 0, pixels * sizeof(float4), ctx->stream));
        if (P.accum_aux) ZB_CUDA(cudaMemsetAsync(P.accum_aux, 0, 3 * pixels * sizeof(float4), ctx->stream));
    }
    // A/B switches (ZbTuning, read once at context creation).  pq: the beauty pass runs as k_render_pq (frame-wide path
    // queue) unless 0 asks for the tile-bound k_render; 2 also selects the queue kernel's AOV and light-pass modes
    // (bit-identical to k_render's but slower: the AOV sums of a warp's lanes go to the same one or two pixels at the
    // same moment and serialise as global REDs — Weekend 64 spp: 44.4 vs 40.6 ms with AOVs, 38.2 vs 32.6 ms light
    // passes; profiles/r01_pq.txt)
    const ZbTuning& tn = ctx->tuning;
    const bool pq = tn.pq != 0, pq_all = tn.pq >= 2;
#ifdef ZB_EXPERIMENTS
    const bool wavefront = tn.wavefront && !o->collect_stats;
    const size_t wf_slots = size_t(std::max(0ll, tn.wf_slots));
#endif
    int done = 0;
    for (int s = sb; s < se; s += per) {
        if (cancel && *cancel) return fail(ZB_ERR_CANCELLED, "render cancelled");
        P.sample_begin = s;
        P.sample_end = std::min(s + per, se);
        P.accumulate = (s > sb || keep) ? 1 : 0;
#ifdef ZB_EXPERIMENTS
        if (wavefront && zbk_wavefront_supported(P)) {
            int iters = 0;
            ZB_CUDA(zbk_launch_wavefront(P, &ctx->wf, wf_slots, &iters, ctx->stream));
            ctx->stats.launches += 2 + 2 * iters;
        } else
#endif
        if (pq && !pq_all && !o->collect_stats && P.max_depth > 0 && P.sample_end > P.sample_begin && !P.light_pass && P.accum_aux &&
                   P.sc.n_nodes < 65536u) {
            // beauty + AOVs: the queue kernel for the beauty sums, the AOV planes from a primary-ray pass of their own
            // (1080p x 64 spp against k_render with AOVs: quirks 62.7 -> 52.5 ms, Weekend 40.7 -> 41.0; on the 1 M-triangle
            // mesh the second set of primary rays costs more than the queue wins, 80.3 -> 84.3, so big trees stay with k_render)
            RenderParams B = P;
            B.accum_aux = nullptr;
            int rc2 = ensure_pq_scratch(ctx, pixels, 0);
            if (rc2) return rc2;
            unsigned long long* heads = ctx->pq_fix + zbk_pq_words(pixels, 0);
            ZB_CUDA(zbk_pq_clear(ctx->pq_fix, pixels, 0, heads, 1, ctx->stream));
            ZB_CUDA(zbk_launch_render_pq(B, owned_tiles(B), 0, ctx->pq_fix, heads, ctx->sm_count, tn, ctx->stream));
            ZB_CUDA(zbk_pq_resolve(B, 0, ctx->pq_fix, 0, B.cam.height, 0.f, nullptr, ctx->stream));
            ZB_CUDA(zbk_launch_aov_pass(P, owned_tiles(P), tn, ctx->stream));
            ctx->stats.launches += 3;
        } else if (pq && !o->collect_stats && P.max_depth > 0 && P.sample_end > P.sample_begin &&
                   (pq_all || (!P.light_pass && !P.accum_aux))) {
            const int mode = P.light_pass ? 2 : P.accum_aux ? 1 : 0;
            int rc2 = ensure_pq_scratch(ctx, pixels, mode);
            if (rc2) return rc2;
            unsigned long long* heads = ctx->pq_fix + zbk_pq_words(pixels, mode);
            ZB_CUDA(zbk_pq_clear(ctx->pq_fix, pixels, mode, heads, 1, ctx->stream));
            ZB_CUDA(zbk_launch_render_pq(P, owned_tiles(P), mode, ctx->pq_fix, heads, ctx->sm_count, tn, ctx->stream));
            ZB_CUDA(zbk_pq_resolve(P, mode, ctx->pq_fix, 0, P.cam.height, 0.f, nullptr, ctx->stream));
            ctx->stats.launches += 2;
        } else {
            ZB_CUDA(zbk_launch_render(P, owned_tiles(P), o->collect_stats != 0, tn, ctx->stream));
            ctx->stats.launches++;
        }
        done = P.sample_end - sb;
        if (progress || cancel) {
            ZB_CUDA(cudaStreamSynchronize(ctx->stream));
            if (progress) progress(int((long long)P.cam.height * done / std::max(se - sb, 1)), user);
        }
    }
    return ZB_OK;
}

int d2h(zb_context* ctx, void* host, const void* dev, size_t bytes) {
    ZB_CUDA(cudaMemcpyAsync(ctx->pinned, dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));
    std::memcpy(host, ctx->pinned, bytes);
    return ZB_OK;
}

int ensure_chunk_events(zb_context* ctx, size_t n) {
    while (ctx->chunk_ev.size() < n) {
        cudaEvent_t e;
        ZB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ctx->chunk_ev.push_back(e);
    }
    return ZB_OK;
}
}  // namespace

int zbr_ensure_pinned(zb_context* ctx, size_t bytes) {
    if (ctx->pinned_bytes >= bytes) return ZB_OK;
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    ctx->pinned = nullptr;
    ctx->pinned_bytes = 0;
    ZB_CUDA(cudaMallocHost(&ctx->pinned, bytes));
    ctx->pinned_bytes = bytes;
    return ZB_OK;
}

// Packed RGB floats on the device (12 B/pixel instead of 24 as doubles) -> this context's pinned stage, in chunks of
// 3 MiB with one event each, asynchronously on the context stream.  The chunks are widened into the caller's 3-double
// layout (camera.hpp:531) by zbr_widen_all while later chunks — of this and of other GPUs — are still in flight.
int zbr_copy_issue(zb_context* ctx, const float* dev_rgb, size_t n_floats, double* dst, size_t stage_offset_floats,
                   std::vector<FrameCopy>& jobs) {
    if (!n_floats) return ZB_OK;
    if ((stage_offset_floats + n_floats) * sizeof(float) > ctx->pinned_bytes) return fail(ZB_ERR_INVALID, "frame copy: pinned stage too small");
    // 3 MiB per copy for a whole frame; a quarter of the job (>= 512 KiB) for the small bands of a multi-GPU frame, so that
    // the copy of chunk c + 1 overlaps the widening of chunk c there too (every copy costs ~10 us of set-up: not more of them)
    const size_t chunk = std::min(size_t(3) << 18, std::max(size_t(1) << 17, (n_floats / 4 + 1023) & ~size_t(1023)));
    const size_t n_chunks = (n_floats + chunk - 1) / chunk;
    size_t first = 0;
    for (const FrameCopy& j : jobs)
        if (j.ctx == ctx) first = std::max(first, j.first_event + j.begin.size() - 1);
    int rc = ensure_chunk_events(ctx, first + n_chunks);
    if (rc) return rc;
    FrameCopy job;
    job.ctx = ctx;
    job.stage = static_cast<const float*>(ctx->pinned) + stage_offset_floats;
    job.dst = dst;
    job.first_event = first;
    float* stage = static_cast<float*>(ctx->pinned) + stage_offset_floats;
    for (size_t c = 0; c < n_chunks; c++) {
        const size_t b = c * chunk, n = std::min(chunk, n_floats - b);
        job.begin.push_back(b);
        ZB_CUDA(cudaMemcpyAsync(stage + b, dev_rgb + b, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        ZB_CUDA(cudaEventRecord(ctx->chunk_ev[first + c], ctx->stream));
    }
    job.begin.push_back(n_floats);
    jobs.push_back(std::move(job));
    return ZB_OK;
}

// Host side of the frame copies: every chunk is widened from its pinned stage into the caller's doubles as soon as its
// event has fired; up to 8 host threads share the chunks of all jobs (round-robin over the jobs, so that every GPU's
// first chunk is served first).
int zbr_widen_all(std::vector<FrameCopy>& jobs) {
    // A work item is a PIECE of a chunk (<= 32 Ki floats): the chunk that arrives last is then widened by all threads
    // together instead of by one (a 3 MiB chunk takes one thread ~1 ms — that was most of what the end-to-end frame
    // added after the kernel).
    constexpr size_t PIECE = size_t(1) << 15;
    struct Item { const FrameCopy* job; size_t chunk, begin, end; };
    std::vector<Item> items;
    size_t longest = 0;
    for (const FrameCopy& j : jobs) longest = std::max(longest, j.begin.size() - 1);
    for (size_t c = 0; c < longest; c++)
        for (const FrameCopy& j : jobs)
            if (c + 1 < j.begin.size())
                for (size_t b = j.begin[c]; b < j.begin[c + 1]; b += PIECE) items.push_back({&j, c, b, std::min(b + PIECE, j.begin[c + 1])});
    if (items.empty()) return ZB_OK;
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    size_t total_floats = 0;
    for (const FrameCopy& j : jobs) total_floats += j.begin.back();
    // The widening is what the end-to-end frame waits for after the kernel (49.8 MB written for a 1080p frame: ~1 ms with 8
    // threads' streaming stores, twice the D2H time), so take the cores there are: up to 16 threads, one per 128 Ki floats of
    // work — but the ranks of a process group share the host, each takes its share of the cores
    const zb_group* g = jobs[0].ctx->group;
    const size_t sharers = (g && g->multi_process && g->n > 1) ? size_t(g->n) : 1;
    const size_t cap = std::max<size_t>(2, std::min<size_t>(16, hw / sharers));
    const size_t n_thr = std::min(std::min(cap, items.size()), total_floats / (size_t(1) << 17) + 1);
    std::atomic<size_t> next{0};
    std::atomic<int> err{0};
    auto work = [&]() {
        size_t seen_job = ~size_t(0), seen_chunk = ~size_t(0);  // the last (job, chunk) this thread has waited for
        for (;;) {
            const size_t k = next.fetch_add(1);
            if (k >= items.size()) return;
            const Item& it = items[k];
            const FrameCopy& j = *it.job;
            if (size_t(it.job - jobs.data()) != seen_job || it.chunk != seen_chunk) {
                if (cudaEventSynchronize(j.ctx->chunk_ev[j.first_event + it.chunk]) != cudaSuccess) { err.store(1); return; }
                seen_job = size_t(it.job - jobs.data());
                seen_chunk = it.chunk;
            }
            zbh_widen_f32_f64(j.stage + it.begin, j.dst + it.begin, it.end - it.begin);
        }
    };
    std::vector<std::thread> pool;
    for (size_t t = 1; t < n_thr; t++) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    if (err.load()) return fail(ZB_ERR_CUDA, "device-to-host copy of the frame failed");
    return ZB_OK;
}

namespace {
// Device float4 means -> the reference's host layout (3 doubles per pixel, camera.hpp:531): packed to RGB floats on the
// device, then zbr_copy_issue / zbr_widen_all.
int d2h_rgb_f64(zb_context* ctx, double* host, const float4* dev, size_t pixels) {
    if (ctx->rgb_pack_pixels < pixels) {
        cudaFree(ctx->rgb_pack);
        ctx->rgb_pack = nullptr;
        ctx->rgb_pack_pixels = 0;
        ZB_CUDA(cudaMalloc(&ctx->rgb_pack, pixels * 3 * sizeof(float)));
        ctx->rgb_pack_pixels = pixels;
    }
    int rc = zbr_ensure_pinned(ctx, pixels * 3 * sizeof(float));
    if (rc) return rc;
    ZB_CUDA(zbk_launch_pack_rgb(dev, pixels, ctx->rgb_pack, ctx->stream));
    ctx->stats.launches++;
    std::vector<FrameCopy> jobs;
    rc = zbr_copy_issue(ctx, ctx->rgb_pack, pixels * 3, host, 0, jobs);
    if (!rc) rc = zbr_widen_all(jobs);
    if (rc) return rc;
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));
    return ZB_OK;
}

void take_counters(zb_context* ctx, const unsigned long long* c);

int read_counters(zb_context* ctx) {
    ZB_CUDA(cudaMemcpyAsync(ctx->counters_host, ctx->counters, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));
    take_counters(ctx, ctx->counters_host);
    return ZB_OK;
}

void take_counters(zb_context* ctx, const unsigned long long* c) {
    ctx->stats.rays = c[0];
    ctx->stats.node_visits = c[1];
    ctx->stats.box_tests = 2 * c[1];
    ctx->stats.prim_tests = c[2];
    ctx->stats.instance_visits = c[3];
    ctx->stats.paths = c[4];
    ctx->stats.stack_overflows = uint32_t(std::min<unsigned long long>(c[6], 0xFFFFFFFFull));
}

// a ray that ran out of traversal stack has dropped geometry: the frame is not the scene's (csrc/trace.cuh)
int check_overflow(zb_context* ctx, const char* who) {
    if (ctx->stats.stack_overflows == 0) return ZB_OK;
    return fail(ZB_ERR_UNSUPPORTED, std::string(who) + ": " + std::to_string(ctx->stats.stack_overflows) +
                                        " traversal-stack overflows (tree deeper than the depth caps); outputs are incomplete");
}

}  // namespace

// exported to the other translation units of the library (noalloc.cu)
int zbr_fill_params(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                    const zb_render_opts* o, RenderParams& P) {
    return fill_params(ctx, scene, cam, env, o, P);
}
int zbr_owned_tiles(const RenderParams& P) { return owned_tiles(P); }
int zbr_read_counters(zb_context* ctx) { return read_counters(ctx); }
int zbr_fetch_counters_async(zb_context* ctx) {
    ZB_CUDA(cudaMemcpyAsync(ctx->counters_host, ctx->counters, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    return ZB_OK;
}
void zbr_take_counters(zb_context* ctx) { take_counters(ctx, ctx->counters_host); }
int zbr_check_overflow(zb_context* ctx, const char* who) { return check_overflow(ctx, who); }

extern "C" int zb_render_device(zb_context* ctx, const zb_scene* scene, const zb_camera* cam,
                                const zb_environment* env, const zb_render_opts* opts, void* accum_f4,
                                void* accum_aux_f4) {
    RenderParams P;
    int rc = fill_params(ctx, scene, cam, env, opts, P);
    if (rc) return rc;
    if (!accum_f4) return fail(ZB_ERR_INVALID, "zb_render_device: accum_f4 is NULL");
    if (P.light_pass && !accum_aux_f4) return fail(ZB_ERR_INVALID, "zb_render_device: light_pass needs accum_aux_f4");
    ZB_CUDA(cudaSetDevice(ctx->device));
    P.accum = static_cast<float4*>(accum_f4);
    P.accum_aux = static_cast<float4*>(accum_aux_f4);
    ZB_CUDA(cudaMemsetAsync(ctx->counters, 0, 8 * sizeof(unsigned long long), ctx->stream));
    ctx->stats = zb_stats{};
    return launch_share(ctx, P, opts, nullptr, nullptr, nullptr, true);
}

// BVH debug view (bvh.hpp:56-110 + camera.hpp:937-954): sums of opts->samples_per_pixel (or the
// sample_begin..sample_end slice) debug-coloured primary rays per pixel into accum_f4, like zb_render_device.
extern "C" int zb_render_bvh_debug_device(zb_context* ctx, const zb_scene* scene, const zb_camera* cam,
                                          const zb_environment* env, const zb_render_opts* opts, int level,
                                          float thickness, void* accum_f4) {
    RenderParams P;
    int rc = fill_params(ctx, scene, cam, env, opts, P);
    if (rc) return rc;
    if (!accum_f4) return fail(ZB_ERR_INVALID, "zb_render_bvh_debug_device: accum_f4 is NULL");
    if (level < -1 || level > 0xFFFF || !(thickness > 0.f)) return fail(ZB_ERR_INVALID, "zb_render_bvh_debug_device: bad level / thickness");
    ZB_CUDA(cudaSetDevice(ctx->device));
    P.accum = static_cast<float4*>(accum_f4);
    P.accum_aux = nullptr;
    P.accumulate = opts->accumulate ? 1 : 0;
    ZB_CUDA(zbk_launch_wireframe(P, level, thickness, ctx->stream));
    ctx->stats = zb_stats{};
    ctx->stats.launches = 1;
    return ZB_OK;
}

// Host-buffer variant in the reference's layout: what camera::render leaves in render_accumulator when
// global_settings::bvh_debug_mode is set (3 doubles per pixel, means over the samples).  Blocking.
extern "C" int zb_render_bvh_debug(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                                   const zb_render_opts* opts, int level, float thickness, double* rgb_f64) {
    if (!ctx || !cam || !opts || !rgb_f64) return fail(ZB_ERR_INVALID, "zb_render_bvh_debug: NULL argument");
    if (cam->image_width < 1 || cam->image_height < 1 || opts->samples_per_pixel < 1) return fail(ZB_ERR_INVALID, "zb_render_bvh_debug: bad size");
    ZB_CUDA(cudaSetDevice(ctx->device));
    const size_t pixels = size_t(cam->image_width) * cam->image_height;
    int rc = ensure_scratch(ctx, pixels, false);
    if (rc) return rc;
    zb_render_opts o = *opts;
    o.accumulate = 0;
    rc = zb_render_bvh_debug_device(ctx, scene, cam, env, &o, level, thickness, ctx->accum);
    if (rc) return rc;
    const int n_s = (o.sample_end > o.sample_begin ? o.sample_end - o.sample_begin : o.samples_per_pixel);
    ZB_CUDA(zbk_launch_resolve(ctx->accum, pixels, float(1.0 / std::max(n_s, 1)), ctx->out_f4, nullptr, ctx->stream));
    ctx->stats.launches++;
    return d2h_rgb_f64(ctx, rgb_f64, ctx->out_f4, pixels);
}

extern "C" int zb_resolve_device(zb_context* ctx, const void* accum_f4, size_t n_pixels, double scale, void* out_f4,
                                 void* out_rgb_f64) {
    if (!ctx || !accum_f4) return fail(ZB_ERR_INVALID, "zb_resolve_device: NULL argument");
    ZB_CUDA(cudaSetDevice(ctx->device));
    ZB_CUDA(zbk_launch_resolve(static_cast<const float4*>(accum_f4), n_pixels, float(scale), static_cast<float4*>(out_f4),
                               static_cast<double*>(out_rgb_f64), ctx->stream));
    ctx->stats.launches++;
    return ZB_OK;
}

extern "C" int zb_read_frame(zb_context* ctx, const void* mean_f4, size_t n_pixels, double* out_rgb_f64) {
    if (!ctx || !mean_f4 || !out_rgb_f64) return fail(ZB_ERR_INVALID, "zb_read_frame: NULL argument");
    ZB_CUDA(cudaSetDevice(ctx->device));
    return d2h_rgb_f64(ctx, out_rgb_f64, static_cast<const float4*>(mean_f4), n_pixels);
}

extern "C" int zb_render(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                         const zb_render_opts* opts, zb_outputs* out, const volatile int* cancel,
                         zb_progress_fn progress, void* user) {
    RenderParams P;
    int rc = fill_params(ctx, scene, cam, env, opts, P);
    if (rc) return rc;
    if (!out) return fail(ZB_ERR_INVALID, "zb_render: outputs is NULL");
    ZB_CUDA(cudaSetDevice(ctx->device));
    const size_t pixels = size_t(cam->image_width) * cam->image_height;
    const bool light = (opts->use_reflection && out->reflection) || (opts->use_refraction && out->refraction);
    const bool aux = out->albedo || out->normal || out->z_depth;
    const bool multi = ctx->group && ctx->group->n > 1;
    if (zbg_fast_path(ctx, opts, out)) {
        // The product path (group.cu): path-queue kernel into 64-bit fixed-point sums, one fused resolve (+ gather over the
        // group's GPUs) that packs the frame for the copy home.  A plain context is a group of one.  On a multi-GPU
        // context this returns all requested buffers; on a single GPU the light passes follow below (k_render<LIGHT>).
        rc = zbg_render(ctx, scene, cam, env, opts, out, cancel, progress, user);
        if (rc) return rc;
        if (multi || !light) {
            if (progress) progress(cam->image_height, user);  // camera.hpp:576-578
            return ZB_OK;
        }
        rc = ensure_scratch(ctx, pixels, true);
        if (rc) return rc;
        const zb_stats beauty_stats = ctx->stats;
        ZB_CUDA(cudaMemsetAsync(ctx->counters, 0, 8 * sizeof(unsigned long long), ctx->stream));
        ZB_CUDA(cudaEventRecord(ctx->ev[4], ctx->stream));
        P.accum = ctx->accum;
        P.accum_aux = ctx->accum_aux;
        P.light_pass = 1;
        rc = launch_share(ctx, P, opts, cancel, nullptr, nullptr, true);
        double* light_out[2] = {opts->use_reflection ? out->reflection : nullptr, opts->use_refraction ? out->refraction : nullptr};
        const float4* light_src[2] = {ctx->accum, ctx->accum_aux};
        for (int k = 0; k < 2 && !rc; k++) {
            if (!light_out[k]) continue;
            ZB_CUDA(zbk_launch_resolve(light_src[k], pixels, float(1.0 / opts->samples_per_pixel), ctx->out_f4, nullptr, ctx->stream));
            ctx->stats.launches++;
            rc = d2h_rgb_f64(ctx, light_out[k], ctx->out_f4, pixels);
        }
        if (rc) return rc;
        ZB_CUDA(cudaEventRecord(ctx->ev[5], ctx->stream));
        rc = read_counters(ctx);
        if (rc) return rc;
        float lms = 0;
        ZB_CUDA(cudaEventElapsedTime(&lms, ctx->ev[4], ctx->ev[5]));
        const uint32_t light_launches = ctx->stats.launches - beauty_stats.launches, light_overflows = ctx->stats.stack_overflows;
        ctx->stats = beauty_stats;  // paths / rays keep the beauty pass's meaning (light passes are not counted)
        ctx->stats.launches += light_launches;
        ctx->stats.stack_overflows += light_overflows;
        ctx->stats.kernel_ms += lms;
        rc = check_overflow(ctx, "zb_render");
        if (rc) return rc;
        if (progress) progress(cam->image_height, user);
        return ZB_OK;
    }
    // ---- general path: STATS kernel, tile-bound k_render (tuning pq = 0), the queue kernel's own AOV / light modes (pq = 2), beauty_f4
    P.light_pass = 0;
    rc = ensure_scratch(ctx, pixels, aux || light);
    if (rc) return rc;
    P.accum = ctx->accum;
    P.accum_aux = aux ? ctx->accum_aux : nullptr;
    if (!out->albedo) P.use_albedo = 0;
    if (!out->normal) P.use_normal = 0;
    if (!out->z_depth) P.use_z = 0;
    ctx->stats = zb_stats{};
    ZB_CUDA(cudaMemsetAsync(ctx->counters, 0, 8 * sizeof(unsigned long long), ctx->stream));
    ZB_CUDA(cudaEventRecord(ctx->ev[0], ctx->stream));
    rc = launch_share(ctx, P, opts, cancel, progress, user, true);
    if (rc) return rc;
    ZB_CUDA(cudaEventRecord(ctx->ev[1], ctx->stream));
    // K5 resolve + copy back in the reference's layout (3 doubles per pixel, camera.hpp:531)
    const double scale = 1.0 / opts->samples_per_pixel;
    const double aux_scale = 1.0 / std::min(P.aux_samples, opts->samples_per_pixel);  // camera.hpp:535-541
    if (out->beauty || out->beauty_f4) {
        ZB_CUDA(zbk_launch_resolve(ctx->accum, pixels, float(scale), ctx->out_f4, nullptr, ctx->stream));
        ctx->stats.launches++;
    }
    ZB_CUDA(cudaEventRecord(ctx->ev[2], ctx->stream));
    if (out->beauty) rc = d2h_rgb_f64(ctx, out->beauty, ctx->out_f4, pixels);
    if (!rc && out->beauty_f4) rc = d2h(ctx, out->beauty_f4, ctx->out_f4, pixels * sizeof(float4));
    double* aux_out[3] = {out->albedo, out->normal, out->z_depth};
    for (int k = 0; k < 3 && !rc; k++) {
        if (!aux_out[k]) continue;
        ZB_CUDA(zbk_launch_resolve(ctx->accum_aux + k * pixels, pixels, float(aux_scale), ctx->out_f4, nullptr, ctx->stream));
        ctx->stats.launches++;
        rc = d2h_rgb_f64(ctx, aux_out[k], ctx->out_f4, pixels);
    }
    if (rc) return rc;
    if (light) {
        // reflection / refraction light passes (camera.hpp:492-517): one more pass over the same samples, with the
        // scratch buffers of the beauty pass reused (their contents are already on the host)
        P.light_pass = 1;
        P.accum_aux = ctx->accum_aux;
        rc = launch_share(ctx, P, opts, cancel, nullptr, nullptr, true);
        double* light_out[2] = {opts->use_reflection ? out->reflection : nullptr, opts->use_refraction ? out->refraction : nullptr};
        const float4* light_src[2] = {ctx->accum, ctx->accum_aux};
        for (int k = 0; k < 2 && !rc; k++) {
            if (!light_out[k]) continue;
            ZB_CUDA(zbk_launch_resolve(light_src[k], pixels, float(scale), ctx->out_f4, nullptr, ctx->stream));
            ctx->stats.launches++;
            rc = d2h_rgb_f64(ctx, light_out[k], ctx->out_f4, pixels);
        }
        if (rc) return rc;
    }
    ZB_CUDA(cudaEventRecord(ctx->ev[3], ctx->stream));
    rc = read_counters(ctx);
    if (rc) return rc;
    float ms = 0;
    ZB_CUDA(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]));
    ctx->stats.kernel_ms = ms;
    ZB_CUDA(cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]));
    ctx->stats.resolve_ms = ms;
    ZB_CUDA(cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]));
    ctx->stats.d2h_ms = ms;
    rc = check_overflow(ctx, "zb_render");
    if (rc) return rc;
    if (progress) progress(cam->image_height, user);  // camera.hpp:576-578
    return ZB_OK;
}

extern "C" int zb_primary_ids(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, uint32_t* leaf_id,
                              uint32_t* prim_id, float* t) {
    if (!ctx || !scene || !cam) return fail(ZB_ERR_INVALID, "zb_primary_ids: NULL argument");
    ZB_CUDA(cudaSetDevice(ctx->device));
    RenderParams P;
    std::memset(&P, 0, sizeof(P));
    P.sc = scene->dev;
    if (!ctx->tuning.wide) P.sc.wide = nullptr;
    int rc = setup_camera(cam, P.cam);
    if (rc) return rc;
    const size_t pixels = size_t(cam->image_width) * cam->image_height;
    if (ctx->id_pixels < pixels) {
        cudaFree(ctx->id_scratch);
        ctx->id_scratch = nullptr;
        ctx->id_pixels = 0;
        ZB_CUDA(cudaMalloc(&ctx->id_scratch, pixels * 3 * sizeof(uint32_t)));
        ctx->id_pixels = pixels;
    }
    uint32_t* d_leaf = ctx->id_scratch;
    uint32_t* d_prim = ctx->id_scratch + pixels;
    float* d_t = reinterpret_cast<float*>(ctx->id_scratch + 2 * pixels);
    ZB_CUDA(cudaMemsetAsync(ctx->counters + 6, 0, sizeof(unsigned long long), ctx->stream));
    ZB_CUDA(zbk_launch_primary_ids(P, d_leaf, d_prim, d_t, ctx->stream));
    ctx->stats.launches++;
    if (leaf_id) ZB_CUDA(cudaMemcpyAsync(leaf_id, d_leaf, pixels * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (prim_id) ZB_CUDA(cudaMemcpyAsync(prim_id, d_prim, pixels * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (t) ZB_CUDA(cudaMemcpyAsync(t, d_t, pixels * 4, cudaMemcpyDeviceToHost, ctx->stream));
    rc = read_counters(ctx);
    if (rc) return rc;
    return check_overflow(ctx, "zb_primary_ids");
}

extern "C" int zb_probe_rays(zb_context* ctx, const zb_scene* scene, const zb_environment* env, size_t n,
                             const float* origins, const float* directions, zb_probe* out) {
    if (!ctx || !scene || !env || (n && (!origins || !directions || !out))) return fail(ZB_ERR_INVALID, "zb_probe_rays: NULL argument");
    if (scene->ctx != ctx) return fail(ZB_ERR_INVALID, "zb_probe_rays: scene belongs to another context");
    ZB_CUDA(cudaSetDevice(ctx->device));
    RenderParams P;
    std::memset(&P, 0, sizeof(P));
    P.sc = scene->dev;
    if (!ctx->tuning.wide) P.sc.wide = nullptr;
    int rc = setup_env(env, scene, P.env);
    if (rc || n == 0) return rc;
    float *d_o = nullptr, *d_d = nullptr;
    zb_probe* d_out = nullptr;
    cudaError_t e = cudaMemsetAsync(ctx->counters + 6, 0, sizeof(unsigned long long), ctx->stream);
    if (e == cudaSuccess) e = cudaMalloc(&d_o, n * 12);
    if (e == cudaSuccess) e = cudaMalloc(&d_d, n * 12);
    if (e == cudaSuccess) e = cudaMalloc(&d_out, n * sizeof(zb_probe));
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_o, origins, n * 12, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_d, directions, n * 12, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = zbk_launch_probe(P, n, d_o, d_d, d_out, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, n * sizeof(zb_probe), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_o);
    cudaFree(d_d);
    cudaFree(d_out);
    if (e != cudaSuccess) return fail(ZB_ERR_CUDA, std::string("zb_probe_rays: ") + cudaGetErrorString(e));
    ctx->stats.launches++;
    rc = read_counters(ctx);
    if (rc) return rc;
    return check_overflow(ctx, "zb_probe_rays");
}

extern "C" int zb_kat_eval(zb_context* ctx, const zb_scene* scene, const zb_environment* env, size_t n, const zb_kat_case* cases,
                           zb_kat_result* out) {
    if (!ctx || !scene || !env || (n && (!cases || !out))) return fail(ZB_ERR_INVALID, "zb_kat_eval: NULL argument");
    if (scene->ctx != ctx) return fail(ZB_ERR_INVALID, "zb_kat_eval: scene belongs to another context");
    ZB_CUDA(cudaSetDevice(ctx->device));
    RenderParams P;
    std::memset(&P, 0, sizeof(P));
    P.sc = scene->dev;
    int rc = setup_env(env, scene, P.env);
    if (rc || n == 0) return rc;
    zb_kat_case* d_in = nullptr;
    zb_kat_result* d_out = nullptr;
    cudaError_t e = cudaMalloc(&d_in, n * sizeof(zb_kat_case));
    if (e == cudaSuccess) e = cudaMalloc(&d_out, n * sizeof(zb_kat_result));
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_in, cases, n * sizeof(zb_kat_case), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = zbk_launch_kat(P, n, d_in, d_out, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, n * sizeof(zb_kat_result), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_in);
    cudaFree(d_out);
    if (e != cudaSuccess) return fail(ZB_ERR_CUDA, std::string("zb_kat_eval: ") + cudaGetErrorString(e));
    return ZB_OK;
}

extern "C" int zb_checkpoint_save(zb_context* ctx, const char* path, const void* accum_f4, int width, int height,
                                  int samples_done, uint64_t seed) {
    if (!ctx || !path || !accum_f4 || width < 1 || height < 1) return fail(ZB_ERR_INVALID, "zb_checkpoint_save: bad argument");
    ZB_CUDA(cudaSetDevice(ctx->device));
    const size_t pixels = size_t(width) * height;
    std::vector<float4> host(pixels);
    ZB_CUDA(cudaMemcpyAsync(host.data(), accum_f4, pixels * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail(ZB_ERR_IO, std::string("cannot create ") + path);
    const uint32_t hdr[4] = {uint32_t(width), uint32_t(height), uint32_t(samples_done), 0u};
    bool ok = std::fwrite("ZBCK", 1, 4, f) == 4 && std::fwrite(hdr, 4, 4, f) == 4 && std::fwrite(&seed, 8, 1, f) == 1 &&
              std::fwrite(host.data(), sizeof(float4), pixels, f) == pixels;
    ok = (std::fclose(f) == 0) && ok;
    return ok ? ZB_OK : fail(ZB_ERR_IO, std::string("short write to ") + path);
}

extern "C" int zb_checkpoint_load(zb_context* ctx, const char* path, void* accum_f4, int width, int height, int* samples_done,
                                  uint64_t* seed) {
    if (!ctx || !path || !accum_f4 || width < 1 || height < 1) return fail(ZB_ERR_INVALID, "zb_checkpoint_load: bad argument");
    FILE* f = std::fopen(path, "rb");
    if (!f) return fail(ZB_ERR_IO, std::string("cannot open ") + path);
    char magic[4];
    uint32_t hdr[4];
    uint64_t sd = 0;
    const size_t pixels = size_t(width) * height;
    std::vector<float4> host(pixels);
    bool ok = std::fread(magic, 1, 4, f) == 4 && std::memcmp(magic, "ZBCK", 4) == 0 && std::fread(hdr, 4, 4, f) == 4 &&
              std::fread(&sd, 8, 1, f) == 1;
    if (ok && (hdr[0] != uint32_t(width) || hdr[1] != uint32_t(height))) {
        std::fclose(f);
        return fail(ZB_ERR_INVALID, std::string(path) + ": checkpoint of another frame size");
    }
    ok = ok && std::fread(host.data(), sizeof(float4), pixels, f) == pixels;
    std::fclose(f);
    if (!ok) return fail(ZB_ERR_INVALID, std::string(path) + ": not a ZBCK checkpoint or truncated");
    ZB_CUDA(cudaSetDevice(ctx->device));
    ZB_CUDA(cudaMemcpyAsync(accum_f4, host.data(), pixels * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (samples_done) *samples_done = int(hdr[2]);
    if (seed) *seed = sd;
    return ZB_OK;
}

extern "C" int zb_get_stats(zb_context* ctx, zb_stats* out) {
    if (!ctx || !out) return fail(ZB_ERR_INVALID, "zb_get_stats: NULL argument");
    if (ctx->stats.n_gpus == 0) {  // a frame rendered through the group path already holds the sums over its GPUs
        int rc = read_counters(ctx);
        if (rc) return rc;
    }
    *out = ctx->stats;
    return ZB_OK;
}
