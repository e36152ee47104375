// Drop-in check of zenith::camera::render (raytracer_project_b200/host/zenith/camera.hpp) — scene and camera
// code written the way the reference's own call sites write it (scene_management.hpp:104-235, main.cpp:1492-1531),
// compiled against this backend's host headers with `using namespace zenith;`.
//
//   drop_in_render SCENE.zscene OUT.bin W H SPP DEPTH SEED
// renders the zscene world through camera::render and writes the six buffers (W*H*3 doubles each); then renders a
// small hand-built scene and exercises lines_rendered and render_flag.  Exit code 0 = all checks passed.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>

#include "../../raytracer_project_b200/host/zenith/camera.hpp"
using namespace zenith;
#include "../../raytracer_project_b200/host/zscene.hpp"

static bool finite_nonzero(const std::vector<color>& b) {
    double s = 0;
    for (const color& c : b) {
        if (!std::isfinite(c.x() + c.y() + c.z())) return false;
        s += c.x() + c.y() + c.z();
    }
    return s > 0;
}

int main(int argc, char** argv) {
    if (argc < 8) { std::fprintf(stderr, "usage: drop_in_render SCENE OUT W H SPP DEPTH SEED [DEVICES]\n"); return 2; }
    // ---- 1. a zscene world through camera::render
    std::ifstream in(argv[1]);
    if (!in) return 2;
    zscene::loader ld;
    ld.parse(in);
    shared_ptr<hittable> world = ld.build_world();
    EnvironmentSettings env;
    env.set_mode(EnvironmentSettings::SOLID_COLOR);
    env.background_color = color(0.7, 0.8, 1.0);
    post_processor post;
    post.z_depth_max_dist = 20.0f;  // the value the ABI-side render of the test uses (the reference's default is 1)
    camera cam;
    cam.image_width = std::atoi(argv[3]);
    cam.image_height = std::atoi(argv[4]);
    cam.samples_per_pixel = std::atoi(argv[5]);
    cam.max_depth = std::atoi(argv[6]);
    cam.seed = std::strtoull(argv[7], nullptr, 10);
    cam.vfov = 20;
    cam.lookfrom = point3(13, 2, 3);
    cam.lookat = point3(0, 0, 0);
    cam.vup = vec3(0, 1, 0);
    cam.defocus_angle = 0.6;
    cam.focus_dist = 10.0;
    if (argc > 8)  // "0,1,..": render on several GPUs through the same call (camera::devices)
        for (const char* p = argv[8]; *p;) {
            cam.devices.push_back(int(std::strtol(p, const_cast<char**>(&p), 10)));
            if (*p == ',') p++;
        }
    cam.use_albedo_buffer = cam.use_normal_buffer = cam.use_z_depth_buffer = true;
    cam.use_reflection = cam.use_refraction = true;
    std::atomic<bool> is_rendering{true};
    cam.reset_accumulator();
    cam.render(*world, env, post, is_rendering);
    if (cam.rc != ZB_OK) { std::fprintf(stderr, "render failed: %s\n", cam.last_error.c_str()); return 1; }
    if (cam.lines_rendered.load() != cam.image_height) { std::fprintf(stderr, "lines_rendered %d\n", cam.lines_rendered.load()); return 1; }
    FILE* f = std::fopen(argv[2], "wb");
    if (!f) return 2;
    for (auto* b : {&cam.render_accumulator, &cam.albedo_buffer, &cam.normal_buffer, &cam.z_depth_buffer, &cam.reflection_buffer, &cam.refraction_buffer})
        std::fwrite(b->data(), sizeof(color), b->size(), f);
    std::fclose(f);

    // ---- 2. reference-style scene code (cf. scene_management.hpp:104-203)
    hittable_list scene;
    auto ground = make_shared<lambertian>(make_shared<checker_texture>(0.5, color(0.1, 0.1, 0.1), color(0.9, 0.9, 0.9)));
    scene.add(make_shared<sphere>(point3(0, -1000, 0), 1000, ground));
    scene.add(make_shared<sphere>(point3(0, 1, 0), 1.0, make_shared<dielectric>(1.5)));
    scene.add(make_shared<sphere>(point3(-4, 1, 0), 1.0, make_shared<lambertian>(color(0.4, 0.2, 0.1))));
    scene.add(make_shared<sphere>(point3(4, 1, 0), 1.0, make_shared<metal>(color(0.7, 0.6, 0.5), 0.0)));
    shared_ptr<hittable> box = make_shared<cube>(point3(-0.5, -0.5, -0.5), point3(0.5, 0.5, 0.5), make_shared<metal>(color(0.8, 0.8, 0.9), 0.1));
    box = make_shared<rotate_y>(box, 0.6);            // radians (rotate_y.hpp:9-13)
    box = make_shared<translate>(box, vec3(2, 0.5, 2));
    scene.add(box);
    scene.add(make_shared<sphere>(point3(0, 6, 0), 1.5, make_shared<diffuse_light>(color(6, 6, 6))));
    auto bvh_world = make_shared<bvh_node>(scene);    // main.cpp:204
    camera cam2;
    cam2.image_width = 64;
    cam2.image_height = 36;
    cam2.samples_per_pixel = 32;
    cam2.max_depth = 20;
    cam2.lookfrom = point3(13, 2, 3);
    cam2.vfov = 25;
    cam2.defocus_angle = 0;
    EnvironmentSettings sun;  // PHYSICAL_SUN is the reference's default mode
    cam2.reset_accumulator();
    cam2.render(*bvh_world, sun, post, is_rendering);
    if (cam2.rc != ZB_OK || !finite_nonzero(cam2.render_accumulator) || cam2.lines_rendered.load() != 36) {
        std::fprintf(stderr, "hand-built scene failed: rc %d %s\n", cam2.rc, cam2.last_error.c_str());
        return 1;
    }
    // ---- 2b. the tail of render(): final_framebuffer (camera.hpp:294), auto-exposure stored back (:263), export (:299)
    if (cam2.final_framebuffer.size() != cam2.render_accumulator.size() || !finite_nonzero(cam2.final_framebuffer)) { std::fprintf(stderr, "final_framebuffer missing\n"); return 1; }
    for (const color& c : cam2.final_framebuffer)
        if (c.x() < 0 || c.x() > 1 || c.y() < 0 || c.y() > 1 || c.z() < 0 || c.z() > 1) { std::fprintf(stderr, "final_framebuffer not display-referred\n"); return 1; }
    post_processor auto_post;
    auto_post.use_auto_exposure = true;
    auto_post.use_bloom = true;
    const float before = auto_post.exposure;
    cam2.render(*bvh_world, sun, auto_post, is_rendering);
    if (cam2.rc != ZB_OK || auto_post.exposure == before) { std::fprintf(stderr, "auto-exposure not applied: %s\n", cam2.last_error.c_str()); return 1; }
    cam2.save_render_pass(render_pass::RGB, std::string("render_") + camera::pass_names[0] + ".png", auto_post);  // main.cpp:1317
    if (cam2.rc != ZB_OK) { std::fprintf(stderr, "save_render_pass: %s\n", cam2.last_error.c_str()); return 1; }
    {
        std::ifstream png("output/render_RGB.png", std::ios::binary);
        char sig[8] = {0};
        png.read(sig, 8);
        if (!png || sig[1] != 'P' || sig[2] != 'N' || sig[3] != 'G') { std::fprintf(stderr, "output/render_RGB.png missing\n"); return 1; }
    }
    // ---- 2c. BVH debug view through the reference's global switch (common.hpp:121-123, bvh.hpp:46-48)
    global_settings::bvh_debug_mode = true;
    global_settings::debug_bvh_level = 0;
    global_settings::bvh_thickness = 0.05f;
    cam2.render(*bvh_world, sun, post, is_rendering);
    global_settings::bvh_debug_mode = false;
    {
        bool frame = false;
        for (const color& c : cam2.render_accumulator) frame = frame || c.z() > 1.0;
        if (cam2.rc != ZB_OK || !frame) { std::fprintf(stderr, "debug view drew no frame: %s\n", cam2.last_error.c_str()); return 1; }
    }
    // ---- 3. render_flag already false: returns at once and leaves the buffers untouched (camera.hpp:441-443)
    cam2.reset_accumulator();
    is_rendering.store(false);
    cam2.render(*bvh_world, sun, post, is_rendering);
    if (finite_nonzero(cam2.render_accumulator) || cam2.lines_rendered.load() != 0) { std::fprintf(stderr, "cancel ignored\n"); return 1; }
    std::printf("{\"drop_in\": \"ok\", \"width\": %d, \"height\": %d}\n", cam.image_width, cam.image_height);
    return 0;
}
