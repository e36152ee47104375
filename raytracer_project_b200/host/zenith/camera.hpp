// zenith/camera.hpp — drop-in for the reference's `camera` on its hot path (camera.hpp:26-88, :209-296).
//
// Same public fields, same buffers, same call:
//     cam.reset_accumulator();
//     cam.render(world, env, post, render_flag);          // camera.hpp:236, caller main.cpp:1520-1531
// but the body of render() up to the end of execute_render_threads (camera.hpp:238-250, :404-579) runs on the
// GPU through the C ABI (include/zenith_b200.h): the scene graph is flattened once per world object
// (flatten.hpp), uploaded, and zb_render fills the caller-visible std::vector<color> buffers in the reference's
// layout (row-major, top row first, double RGB), reports progress through `lines_rendered` and honours
// `render_flag` between sample batches.  The tail of render() runs too: auto-exposure (camera.hpp:258-267) and
// update_post_processing into `final_framebuffer` (:294, :142-207) through the finishing-chain kernels; OIDN
// (:270-291) is absent here, as in SURVEY 8(f).  save_render_pass (:299-343) writes "output/<file>" PNGs, and
// global_settings::bvh_debug_mode switches render() to the BVH debug view (bvh.hpp:56-110).
//
// Header-only C++20 over the C ABI; link with libzenith_b200.so.  There is no CPU fallback: without a B200 the
// call fails and `last_error` says why.
#pragma once
#include <atomic>
#include <filesystem>
#include <memory>
#include <string>
#include <vector>

#include "../../../include/zenith_b200.h"
#include "../flatten.hpp"

namespace zenith {

// common.hpp:110-118
enum class render_pass : int { RGB = 0, DENOISE, ALBEDO, NORMALS, REFLECTIONS, REFRACTIONS, Z_DEPTH };

// common.hpp:120-124
namespace global_settings {
inline bool bvh_debug_mode = false;
inline float bvh_thickness = 0.01f;
inline int debug_bvh_level = -1;
}  // namespace global_settings

// color_processing.hpp:9-28
struct image_statistics {
    float average_luminance = 0.0f;
    float max_luminance = 0.0f;
    int histogram[256] = {0};
};

// color_processing.hpp:30-42
struct debug_flags {
    bool red = false, green = false, blue = false, luminance = false, bvh = false;
    bool any_active() const { return red || green || blue || luminance || bvh; }
};

// The fields of the reference's post_processor (color_processing.hpp:44-74).  Its arithmetic lives in the finishing
// kernels (csrc/finish.cu); this is the parameter block the caller edits.
struct post_processor {
    mutable float exposure = 0.5f;
    float saturation = 1.0f;
    float contrast = 1.0f;
    float hue_shift = 0.0f;
    float vignette_intensity = 1.0f;
    vec3 color_balance = vec3(1.0, 1.0, 1.0);
    float z_depth_max_dist = 1.0f;
    float exposure_compensation_stops = 0.0f;
    bool use_aces_tone_mapping = false;
    bool use_auto_exposure = false;
    float target_luminance = 0.12f;
    mutable debug_flags debug;
    bool use_bloom = false;
    float bloom_threshold = 1.0f;
    float bloom_intensity = 0.3f;
    int bloom_radius = 4;
    bool needs_update = true;
    mutable image_statistics last_stats;
    bool use_sharpening = false;
    double sharpen_amount = 0.2;

    zb_post to_abi() const {
        zb_post p{};
        p.exposure = exposure;
        p.saturation = saturation;
        p.contrast = contrast;
        p.hue_shift = hue_shift;
        p.vignette_intensity = vignette_intensity;
        for (int k = 0; k < 3; k++) p.color_balance[k] = float(color_balance[k]);
        p.exposure_compensation_stops = exposure_compensation_stops;
        p.target_luminance = target_luminance;
        p.use_aces_tone_mapping = use_aces_tone_mapping;
        p.use_auto_exposure = use_auto_exposure;
        p.debug_red = debug.red;
        p.debug_green = debug.green;
        p.debug_blue = debug.blue;
        p.debug_luminance = debug.luminance;
        p.debug_bvh = debug.bvh;
        p.use_bloom = use_bloom;
        p.bloom_threshold = bloom_threshold;
        p.bloom_intensity = bloom_intensity;
        p.bloom_radius = bloom_radius;
        p.use_sharpening = use_sharpening;
        p.sharpen_amount = sharpen_amount;
        return p;
    }
};

class camera {
public:
    // camera.hpp:26-42
    double aspect_ratio = 16.0 / 9.0;
    int image_width = 400;
    int image_height = 225;
    int samples_per_pixel = 30;
    int max_depth = 10;
    double vfov = 30;
    point3 lookfrom = point3(10.0, 1.5, 0.0);
    point3 lookat = point3(0, 0, 0);
    vec3 vup = vec3(0, 1, 0);
    double defocus_angle = 0.5;
    double focus_dist = 10.0;
    // camera.hpp:53-57
    bool use_albedo_buffer = false, use_normal_buffer = false, use_z_depth_buffer = false;
    bool use_reflection = false, use_refraction = false;
    // camera.hpp:63, :80-85, :88
    std::vector<color> render_accumulator, albedo_buffer, normal_buffer, z_depth_buffer, reflection_buffer, refraction_buffer;
    std::vector<color> final_framebuffer;  // camera.hpp:64: what the GUI shows, written by update_post_processing
    std::atomic<int> lines_rendered{0};
    // camera.hpp:66-78
    static inline const char* pass_names[] = {"RGB", "Denoise", "Albedo", "Normals", "Reflections", "Refractions", "Z-Depth"};
    render_pass current_display_pass = render_pass::RGB;

    // backend knobs (no counterpart in the reference)
    int device = 0;
    // more than one entry: the frame is rendered by all of these GPUs (zb_context_create_multi); render() stays one
    // blocking call that fills the same buffers — the reference's thread fan-out (camera.hpp:426-427) by other means
    std::vector<int> devices;
    double* registered[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // output buffers page-locked for direct copies
    size_t registered_bytes[6] = {0, 0, 0, 0, 0, 0};
    // mesh BLASes are built on the device at upload (binned SAH, csrc/sah_build.cu: 6.5 ms for 1 M triangles against 2.8 s
    // for the same tree on the host); false = host build (host/bvh_build.hpp).  The reference rebuilds geometry and BVH on
    // the host at every restart (main.cpp:1492-1500, bvh.hpp:11-44 via model.hpp:95)
    bool device_blas = true;
    uint64_t seed = 0x5EED0001ull;
    int rc = ZB_OK;            // status of the last render()
    std::string last_error;

    camera() = default;
    camera(const camera&) = delete;
    camera& operator=(const camera&) = delete;
    ~camera() { release(); }

    // camera.hpp:209-233: sizes every buffer to W*H and clears it; render() never resizes
    void reset_accumulator() {
        const size_t n = size_t(std::max(image_width, 1)) * size_t(std::max(image_height, 1));
        unregister_outputs();  // the vectors may move
        for (auto* b : {&render_accumulator, &albedo_buffer, &normal_buffer, &z_depth_buffer, &reflection_buffer, &refraction_buffer})
            b->assign(n, color(0, 0, 0));
        lines_rendered.store(0);
    }

    // zb_context_register_output for the buffers a frame writes; re-done when a vector has moved or grown
    void register_outputs(const zb_outputs& out) {
        const size_t bytes = size_t(std::max(image_width, 1)) * size_t(std::max(image_height, 1)) * sizeof(color);
        double* want[6] = {out.beauty, out.albedo, out.normal, out.z_depth, out.reflection, out.refraction};
        for (int k = 0; k < 6; k++) {
            if (want[k] == registered[k] && (!want[k] || registered_bytes[k] == bytes)) continue;
            if (registered[k]) zb_context_unregister_output(ctx, registered[k]);
            registered[k] = nullptr;
            if (want[k] && zb_context_register_output(ctx, want[k], bytes) == ZB_OK) {
                registered[k] = want[k];
                registered_bytes[k] = bytes;
            }
        }
    }
    void unregister_outputs() {
        for (int k = 0; k < 6; k++) {
            if (registered[k] && ctx) zb_context_unregister_output(ctx, registered[k]);
            registered[k] = nullptr;
        }
    }

    // The uploaded scene is re-used while the world object AND a content stamp of its graph stay the same: the stamp
    // (FNV-1a over every object's kind, bounding box, wrapper matrices, material pointer and parameters; a mesh counts
    // by pointer, triangle count and box) catches in-place edits and a new world allocated at an old address — the
    // reference's GUI replaces bvh_world on every restart (main.cpp:1492-1500).  invalidate_scene() forces a re-flatten.
    void invalidate_scene() { cached_world = nullptr; }

    static void stamp_mix(uint64_t& h, const void* p, size_t n) {
        const unsigned char* b = static_cast<const unsigned char*>(p);
        for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    }
    static void stamp_material(uint64_t& h, const material* m) {
        stamp_mix(h, &m, sizeof(m));
        if (!m) return;
        const int k = int(m->kind());
        const void* t = m->tex.get();
        const void* b = m->bump.get();
        stamp_mix(h, &k, sizeof(k));
        stamp_mix(h, &t, sizeof(t));
        stamp_mix(h, &b, sizeof(b));
        stamp_mix(h, &m->bump_strength, sizeof(double));
        stamp_mix(h, &m->param, sizeof(double));
        stamp_mix(h, &m->tint, sizeof(color));
    }
    static void stamp_object(uint64_t& h, const hittable& o) {
        using K = hittable::kind_t;
        const int k = int(o.kind());
        const aabb box = o.bounding_box();
        const double bb[6] = {box.x.min, box.x.max, box.y.min, box.y.max, box.z.min, box.z.max};
        stamp_mix(h, &k, sizeof(k));
        stamp_mix(h, bb, sizeof(bb));
        switch (o.kind()) {
            case K::list: for (auto& c : static_cast<const hittable_list&>(o).objects) stamp_object(h, *c); break;
            case K::bvh: for (auto& c : static_cast<const bvh_node&>(o).objects) stamp_object(h, *c); break;
            case K::sphere: stamp_material(h, static_cast<const sphere&>(o).mat.get()); break;
            case K::triangle: stamp_material(h, static_cast<const triangle&>(o).mat.get()); break;
            case K::cube: stamp_material(h, static_cast<const cube&>(o).mat.get()); break;
            case K::translate: case K::rotate_x: case K::rotate_y: case K::rotate_z: case K::scale: {
                auto& w = static_cast<const transform_wrapper&>(o);
                const affine m = w.object_to_world();
                stamp_mix(h, m.m, sizeof(m.m));
                stamp_object(h, *w.ptr);
            } break;
            case K::material_instance: {
                auto& mi = static_cast<const material_instance&>(o);
                stamp_material(h, mi.new_material.get());
                stamp_object(h, *mi.object);
            } break;
            case K::constant_medium: {
                auto& cm = static_cast<const constant_medium&>(o);
                stamp_mix(h, &cm.neg_inv_density, sizeof(double));
                stamp_material(h, cm.phase_function.get());
                stamp_object(h, *cm.boundary);
            } break;
            case K::model: {
                auto& md = static_cast<const model&>(o);
                const void* p = &md;
                const size_t nt = md.tris.size();
                stamp_mix(h, &p, sizeof(p));
                stamp_mix(h, &nt, sizeof(nt));
                stamp_material(h, md.mat.get());
            } break;
        }
    }
    static uint64_t scene_stamp(const hittable& world) {
        uint64_t h = 1469598103934665603ull;
        stamp_object(h, world);
        return h;
    }

    int create_context() {
        if (devices.size() > 1) return zb_context_create_multi(devices.data(), int(devices.size()), &ctx);
        return zb_context_create(devices.empty() ? device : devices[0], &ctx);
    }

    void render(const hittable& world, const EnvironmentSettings& env, const post_processor& post, std::atomic<bool>& render_flag) {
        last_error.clear();
        if (image_width < 1) image_width = 1;  // camera::initialize (camera.hpp:359-366)
        if (image_height < 1) image_height = 1;
        aspect_ratio = double(image_width) / image_height;
        lines_rendered.store(0);
        const size_t n = size_t(image_width) * image_height;
        if (render_accumulator.size() < n) { set_error(ZB_ERR_INVALID, "camera::render: call reset_accumulator() first (camera.hpp:209)"); return; }
        if (!ctx && (rc = create_context()) != ZB_OK) { set_error(rc, zb_last_error()); return; }
        const uint64_t stamp = scene_stamp(world);
        if (cached_world != &world || cached_stamp != stamp || cached_hdr != env.hdr_texture.get()) {
            if (scene) zb_scene_destroy(scene);
            scene = nullptr;
            flattener fl;
            fl.gpu_blas = device_blas;
            std::unique_ptr<flat_scene> flat = fl.run(world, &env);
            hdr_texture_id = flat->env_hdr_texture;
            if ((rc = zb_scene_upload(ctx, &flat->view, &scene)) != ZB_OK) { set_error(rc, zb_last_error()); return; }
            cached_world = &world;
            cached_stamp = stamp;
            cached_hdr = env.hdr_texture.get();
        }
        zb_camera c{};
        c.image_width = image_width;
        c.image_height = image_height;
        c.vfov = vfov;
        for (int k = 0; k < 3; k++) { c.lookfrom[k] = lookfrom[k]; c.lookat[k] = lookat[k]; c.vup[k] = vup[k]; }
        c.defocus_angle = defocus_angle;
        c.focus_dist = focus_dist;
        zb_environment e{};
        e.mode = env.mode() == EnvironmentSettings::PHYSICAL_SUN ? ZB_ENV_PHYSICAL_SUN
               : env.mode() == EnvironmentSettings::HDR_MAP    ? ZB_ENV_HDR_MAP : ZB_ENV_SOLID_COLOR;
        e.hdr_texture = hdr_texture_id;
        for (int k = 0; k < 3; k++) { e.background_color[k] = env.background_color[k]; e.sun_direction[k] = env.sun_direction[k]; e.sun_color[k] = env.sun_color[k]; }
        e.intensity = env.intensity;
        e.hdri_rotation = env.hdri_rotation;
        e.hdri_tilt = env.hdri_tilt;
        e.hdri_roll = env.hdri_roll;
        e.sun_intensity = env.sun_intensity;
        e.sun_size = env.sun_size;
        zb_render_opts o{};
        o.samples_per_pixel = samples_per_pixel;
        o.max_depth = max_depth;
        o.seed = seed++;  // the reference never repeats a noise pattern either (random_device, common.hpp:31)
        o.sample_begin = 0;
        o.sample_end = samples_per_pixel;
        o.tile_ranks = 1;
        o.use_albedo = use_albedo_buffer;
        o.use_normal = use_normal_buffer;
        o.use_z_depth = use_z_depth_buffer;
        o.use_reflection = use_reflection;
        o.use_refraction = use_refraction;
        o.z_depth_max_dist = post.z_depth_max_dist;
        static_assert(sizeof(color) == 3 * sizeof(double), "color must be 3 packed doubles (vec3.hpp:10)");
        zb_outputs out{};
        out.beauty = reinterpret_cast<double*>(render_accumulator.data());
        out.albedo = use_albedo_buffer ? reinterpret_cast<double*>(albedo_buffer.data()) : nullptr;
        out.normal = use_normal_buffer ? reinterpret_cast<double*>(normal_buffer.data()) : nullptr;
        out.z_depth = use_z_depth_buffer ? reinterpret_cast<double*>(z_depth_buffer.data()) : nullptr;
        out.reflection = use_reflection ? reinterpret_cast<double*>(reflection_buffer.data()) : nullptr;
        out.refraction = use_refraction ? reinterpret_cast<double*>(refraction_buffer.data()) : nullptr;
        // several GPUs: let every GPU DMA its rows straight into these buffers (they are sized once and reused,
        // camera.hpp:209-233) instead of widening eight bands with host threads
        if (devices.size() > 1) register_outputs(out);
        watch w{this, &render_flag, 0};
        if (!render_flag.load()) return;  // camera.hpp:441-443
        if (global_settings::bvh_debug_mode)  // bvh.hpp:46-48: the traversal itself switches to the debug view
            rc = zb_render_bvh_debug(ctx, scene, &c, &e, &o, global_settings::debug_bvh_level, global_settings::bvh_thickness,
                                     reinterpret_cast<double*>(render_accumulator.data()));
        else
            rc = zb_render(ctx, scene, &c, &e, &o, &out, &w.cancel, &camera::on_progress, &w);
        if (rc == ZB_ERR_CANCELLED) return;  // partial buffers stay, like the reference's early return (camera.hpp:252-254)
        if (rc != ZB_OK) { set_error(rc, zb_last_error()); return; }
        lines_rendered.store(image_height);  // camera.hpp:576-578
        if (!render_flag.load()) return;     // camera.hpp:252-254
        if (post.use_auto_exposure) {        // camera.hpp:258-267
            zb_image_stats st{};
            if ((rc = zb_analyze_f64(ctx, reinterpret_cast<const double*>(render_accumulator.data()), n, &st)) != ZB_OK) { set_error(rc, zb_last_error()); return; }
            const zb_post p = post.to_abi();
            post.exposure = float(zb_auto_exposure(&p, &st));
        }
        update_post_processing(post, image_width, image_height);  // camera.hpp:294
    }

    // camera.hpp:90-113
    const std::vector<color>& get_active_buffer() const {
        switch (current_display_pass) {
            case render_pass::ALBEDO: return albedo_buffer;
            case render_pass::NORMALS: return normal_buffer;
            case render_pass::REFLECTIONS: return reflection_buffer;
            case render_pass::REFRACTIONS: return refraction_buffer;
            case render_pass::Z_DEPTH: return z_depth_buffer;
            default: return render_accumulator;  // RGB, and DENOISE (no denoiser here)
        }
    }

    // camera.hpp:142-207: the displayed pass through exposure, bloom, sharpening and process() into final_framebuffer
    void update_post_processing(const post_processor& post, int w, int h) {
        if (render_accumulator.empty()) return;
        const std::vector<color>& source = get_active_buffer();
        if (source.size() < size_t(w) * size_t(h) || w < 1 || h < 1) return;
        final_framebuffer.resize(source.size());
        if (!ctx && (rc = create_context()) != ZB_OK) { set_error(rc, zb_last_error()); return; }
        const bool beauty = current_display_pass == render_pass::RGB || current_display_pass == render_pass::DENOISE;
        const bool light = current_display_pass == render_pass::REFLECTIONS || current_display_pass == render_pass::REFRACTIONS;
        const zb_post p = post.to_abi();
        rc = zb_finish(ctx, reinterpret_cast<const double*>(source.data()), w, h, &p, beauty ? ZB_PASS_BEAUTY : light ? ZB_PASS_LIGHT : ZB_PASS_AUX,
                       reinterpret_cast<double*>(final_framebuffer.data()));
        if (rc != ZB_OK) set_error(rc, zb_last_error());
    }

    // camera.hpp:299-343: "output/<filename>", created on demand
    void save_render_pass(render_pass pass, const std::string& filename, const post_processor& pp) {
        std::error_code ec;
        std::filesystem::create_directory("output", ec);
        const std::vector<color>* src = &render_accumulator;
        switch (pass) {
            case render_pass::ALBEDO: src = &albedo_buffer; break;
            case render_pass::NORMALS: src = &normal_buffer; break;
            case render_pass::REFLECTIONS: src = &reflection_buffer; break;
            case render_pass::REFRACTIONS: src = &refraction_buffer; break;
            case render_pass::Z_DEPTH: src = &z_depth_buffer; break;
            default: break;
        }
        if (src->size() < size_t(image_width) * size_t(image_height)) { set_error(ZB_ERR_INVALID, "save_render_pass: the pass buffer is empty"); return; }
        if (!ctx && (rc = create_context()) != ZB_OK) { set_error(rc, zb_last_error()); return; }
        const zb_post p = pp.to_abi();
        rc = zb_save_pass_f64(ctx, reinterpret_cast<const double*>(src->data()), image_width, image_height, &p, int(pass), ("output/" + filename).c_str());
        if (rc != ZB_OK) set_error(rc, zb_last_error());
    }

    void release() {
        unregister_outputs();
        if (scene) zb_scene_destroy(scene);
        if (ctx) zb_context_destroy(ctx);
        scene = nullptr;
        ctx = nullptr;
        cached_world = nullptr;
    }

private:
    struct watch {
        camera* self;
        std::atomic<bool>* flag;
        volatile int cancel;
    };
    // called by zb_render between sample batches on the calling thread
    static void on_progress(int lines, void* user) {
        watch* w = static_cast<watch*>(user);
        w->self->lines_rendered.store(lines);
        if (!w->flag->load()) w->cancel = 1;
    }
    void set_error(int code, const char* msg) {
        rc = code;
        last_error = msg ? msg : "";
    }
    zb_context* ctx = nullptr;
    zb_scene* scene = nullptr;
    const hittable* cached_world = nullptr;
    uint64_t cached_stamp = 0;
    const void* cached_hdr = nullptr;
    uint32_t hdr_texture_id = ZB_NO_TEXTURE;
};

}  // namespace zenith
