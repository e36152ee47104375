// zenith/camera.hpp — drop-in for the reference's `camera` on its hot path (camera.hpp:26-88, :209-296).
//
// Same public fields, same buffers, same call:
//     cam.reset_accumulator();
//     cam.render(world, env, post, render_flag);          // camera.hpp:236, caller main.cpp:1520-1531
// but the body of render() up to the end of execute_render_threads (camera.hpp:238-250, :404-579) runs on the
// GPU through the C ABI (include/zenith_b200.h): the scene graph is flattened once per world object
// (flatten.hpp), uploaded, and zb_render fills the caller-visible std::vector<color> buffers in the reference's
// layout (row-major, top row first, double RGB), reports progress through `lines_rendered` and honours
// `render_flag` between sample batches.  The reference's tail of render() — auto-exposure, OIDN,
// post-processing (camera.hpp:258-295) — is host/UI code and stays with the caller.
//
// Header-only C++20 over the C ABI; link with libzenith_b200.so.  There is no CPU fallback: without a B200 the
// call fails and `last_error` says why.
#pragma once
#include <atomic>
#include <memory>
#include <string>
#include <vector>

#include "../../../include/zenith_b200.h"
#include "../flatten.hpp"

namespace zenith {

// Only z_depth_max_dist reaches the hot path (camera.hpp:247); the other post-processing fields are UI state.
struct post_processor {
    double z_depth_max_dist = 20.0;
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
    std::atomic<int> lines_rendered{0};

    // backend knobs (no counterpart in the reference)
    int device = 0;
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
        for (auto* b : {&render_accumulator, &albedo_buffer, &normal_buffer, &z_depth_buffer, &reflection_buffer, &refraction_buffer})
            b->assign(n, color(0, 0, 0));
        lines_rendered.store(0);
    }

    // Call after editing the scene graph in place (the reference rebuilds its BVH at main.cpp:1492-1500).
    void invalidate_scene() { cached_world = nullptr; }

    void render(const hittable& world, const EnvironmentSettings& env, const post_processor& post, std::atomic<bool>& render_flag) {
        last_error.clear();
        if (image_width < 1) image_width = 1;  // camera::initialize (camera.hpp:359-366)
        if (image_height < 1) image_height = 1;
        aspect_ratio = double(image_width) / image_height;
        lines_rendered.store(0);
        const size_t n = size_t(image_width) * image_height;
        if (render_accumulator.size() < n) { set_error(ZB_ERR_INVALID, "camera::render: call reset_accumulator() first (camera.hpp:209)"); return; }
        if (!ctx && (rc = zb_context_create(device, &ctx)) != ZB_OK) { set_error(rc, zb_last_error()); return; }
        if (cached_world != &world || cached_hdr != env.hdr_texture.get()) {
            if (scene) zb_scene_destroy(scene);
            scene = nullptr;
            flattener fl;
            std::unique_ptr<flat_scene> flat = fl.run(world, &env);
            hdr_texture_id = flat->env_hdr_texture;
            if ((rc = zb_scene_upload(ctx, &flat->view, &scene)) != ZB_OK) { set_error(rc, zb_last_error()); return; }
            cached_world = &world;
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
        watch w{this, &render_flag, 0};
        if (!render_flag.load()) return;  // camera.hpp:441-443
        rc = zb_render(ctx, scene, &c, &e, &o, &out, &w.cancel, &camera::on_progress, &w);
        if (rc == ZB_ERR_CANCELLED) return;  // partial buffers stay, like the reference's early return (camera.hpp:252-254)
        if (rc != ZB_OK) { set_error(rc, zb_last_error()); return; }
        lines_rendered.store(image_height);  // camera.hpp:576-578
    }

    void release() {
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
    const void* cached_hdr = nullptr;
    uint32_t hdr_texture_id = ZB_NO_TEXTURE;
};

}  // namespace zenith
