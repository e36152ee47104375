// TEST INFRASTRUCTURE — oracle harness around the UNMODIFIED reference (jarek1992/raytracer_project).
//
// One translation unit that includes the reference's own headers from /root/reference (nothing is
// copied into this repository) in the order main.cpp:7-13 uses, and drives the reference's classes:
//   ids     centre-of-pixel primary hits through camera::get_center_ray (camera.hpp:809-814) and
//           world.hit(interval(0.001, inf)) (camera.hpp:459), with every leaf wrapped in an id recorder
//   render  camera::initialize + camera::execute_render_threads (camera.hpp:358, :404), timed
//   stats   N independent renders -> per-pixel mean and variance of the mean (for the 3-sigma test)
//   probe   per-ray hit_record / albedo / emitted / background for arbitrary rays (known-answer vectors)
//   decode  stb_image decode of a texture into a .zbtex raw dump (identical texels on both sides)
//   kat     known-answer values of single reference functions (tests/golden/kat.json)
// Built by oracle/Makefile into oracle/_ref/ref_harness with -fno-access-control so private members
// (get_center_ray, execute_render_threads, bvh_node::left/right, triangle::v0..) are reachable
// without touching the reference.  Only tests/, bench.py (cpu_baseline / --impl reference) and
// __graft_entry__.smoke() may execute it; the product never does.
//
// Never run with cwd = /root/reference: camera::initialize creates ./assets/hdr_maps (camera.hpp:123-140).
#include "common.hpp"
#include "camera.hpp"
#include "bvh.hpp"
#include "environment.hpp"
#include "sphere.hpp"
#include "triangle.hpp"
#include "cube.hpp"
#include "translate.hpp"
#include "rotate_x.hpp"
#include "rotate_y.hpp"
#include "rotate_z.hpp"
#include "scale.hpp"
#include "material_instance.hpp"
#include "constant_medium.hpp"
#include "model.hpp"
#include "material_library.hpp"

#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <unordered_map>

// ------------------------------------------------------------------ leaf id recorder
static thread_local int tl_leaf = -1, tl_tri = -1, tl_cur_tri = -1;

class id_tag : public hittable {
public:
    id_tag(shared_ptr<hittable> o, int id, bool is_tri) : obj(o), id(id), is_tri(is_tri) {}
    bool hit(const ray& r, interval ray_t, hit_record& rec, int depth = 0, bool debug_wire = false) const override {
        if (!is_tri) tl_cur_tri = -1;
        if (!obj->hit(r, ray_t, rec, depth, debug_wire)) return false;
        if (is_tri) {
            tl_cur_tri = id;
        } else {  // accepted inside the shrinking interval => currently the closest
            tl_leaf = id;
            tl_tri = tl_cur_tri;
        }
        return true;
    }
    aabb bounding_box() const override { return obj->bounding_box(); }
    shared_ptr<hittable> obj;
    int id;
    bool is_tri;
};

// every triangle of every model gets a face index = its position in the OBJ's triangulated face
// order; found by re-running the same TinyObjLoader call and matching the exact vertex doubles.
struct tri_key {
    double v[9];
    bool operator==(const tri_key& o) const { return std::memcmp(v, o.v, sizeof(v)) == 0; }
};
struct tri_key_hash {
    size_t operator()(const tri_key& k) const {
        uint64_t h = 1469598103934665603ull;
        const unsigned char* p = reinterpret_cast<const unsigned char*>(k.v);
        for (size_t i = 0; i < sizeof(k.v); i++) h = (h ^ p[i]) * 1099511628211ull;
        return size_t(h);
    }
};

static void tag_mesh_tree(shared_ptr<hittable>& slot, const std::unordered_map<tri_key, int, tri_key_hash>& faces,
                          std::map<hittable*, shared_ptr<hittable>>& done) {
    auto seen = done.find(slot.get());
    if (seen != done.end()) {  // span-1 leaves hold the same object as left and right (bvh.hpp:25-26)
        slot = seen->second;
        return;
    }
    if (auto node = std::dynamic_pointer_cast<bvh_node>(slot)) {
        tag_mesh_tree(node->left, faces, done);
        tag_mesh_tree(node->right, faces, done);
        return;
    }
    if (auto tri = std::dynamic_pointer_cast<triangle>(slot)) {
        tri_key k;
        for (int c = 0; c < 3; c++) { k.v[c] = tri->v0[c]; k.v[3 + c] = tri->v1[c]; k.v[6 + c] = tri->v2[c]; }
        auto f = faces.find(k);
        int face = f == faces.end() ? -2 : f->second;
        hittable* raw = slot.get();
        shared_ptr<hittable> tagged = make_shared<id_tag>(slot, face, true);
        done[raw] = tagged;
        slot = tagged;
    }
}

static void tag_model(const std::string& path, double scl, model& m) {
    tinyobj::attrib_t attrib;
    std::vector<tinyobj::shape_t> shapes;
    std::vector<tinyobj::material_t> mats;
    std::string warn, err;
    if (!tinyobj::LoadObj(&attrib, &shapes, &mats, &warn, &err, path.c_str())) return;
    double lo[3] = {infinity, infinity, infinity}, hi[3] = {-infinity, -infinity, -infinity};
    for (size_t i = 0; i + 2 < attrib.vertices.size(); i += 3)
        for (int k = 0; k < 3; k++) { lo[k] = fmin(lo[k], attrib.vertices[i + k]); hi[k] = fmax(hi[k], attrib.vertices[i + k]); }
    const double off[3] = {(lo[0] + hi[0]) / 2.0, lo[1], (lo[2] + hi[2]) / 2.0};
    std::unordered_map<tri_key, int, tri_key_hash> faces;
    int face = 0;
    for (const auto& sh : shapes) {
        size_t at = 0;
        for (size_t f = 0; f < sh.mesh.num_face_vertices.size(); f++, at += 3, face++) {
            tri_key k;
            for (int c = 0; c < 3; c++) {
                int vi = sh.mesh.indices[at + c].vertex_index;
                for (int a = 0; a < 3; a++) k.v[3 * c + a] = (attrib.vertices[3 * vi + a] - off[a]) * scl;
            }
            faces.emplace(k, face);
        }
    }
    std::map<hittable*, shared_ptr<hittable>> done;
    shared_ptr<hittable> root = m.mesh_bvh;
    tag_mesh_tree(root, faces, done);
}

// the loader tags every container-add site; models are additionally tagged per triangle
#define ZSCENE_TAG(obj, id) std::static_pointer_cast<hittable>(make_shared<id_tag>((obj), (id), false))
#include "../raytracer_project_b200/host/zscene.hpp"

// ------------------------------------------------------------------ render configuration ("zcam")
struct config {
    std::map<std::string, std::vector<std::string>> kv;
    bool has(const std::string& k) const { return kv.count(k) != 0; }
    double num(const std::string& k, double dflt, size_t i = 0) const {
        auto it = kv.find(k);
        return it == kv.end() || it->second.size() <= i ? dflt : std::strtod(it->second[i].c_str(), nullptr);
    }
    std::string str(const std::string& k, const std::string& dflt) const {
        auto it = kv.find(k);
        return it == kv.end() || it->second.empty() ? dflt : it->second[0];
    }
    vec3 v3(const std::string& k, vec3 d) const { return vec3(num(k, d.x(), 0), num(k, d.y(), 1), num(k, d.z(), 2)); }
    static config load(const std::string& path) {
        config c;
        std::ifstream in(path);
        if (!in) { std::fprintf(stderr, "cannot open config %s\n", path.c_str()); std::exit(2); }
        std::string line;
        while (std::getline(in, line)) {
            size_t hash = line.find('#');
            if (hash != std::string::npos) line.resize(hash);
            std::istringstream ss(line);
            std::string k, t;
            if (!(ss >> k)) continue;
            while (ss >> t) c.kv[k].push_back(t);
        }
        return c;
    }
};

static void apply_camera(const config& c, camera& cam) {
    cam.image_width = int(c.num("image_width", 400));
    cam.image_height = int(c.num("image_height", 225));
    cam.samples_per_pixel = int(c.num("samples_per_pixel", 10));
    cam.max_depth = int(c.num("max_depth", 50));
    cam.vfov = c.num("vfov", 30);
    cam.lookfrom = c.v3("lookfrom", vec3(10, 1.5, 0));
    cam.lookat = c.v3("lookat", vec3(0, 0, 0));
    cam.vup = c.v3("vup", vec3(0, 1, 0));
    cam.defocus_angle = c.num("defocus_angle", 0.0);
    cam.focus_dist = c.num("focus_dist", 10.0);
    cam.use_albedo_buffer = c.num("use_albedo", 0) != 0;
    cam.use_normal_buffer = c.num("use_normal", 0) != 0;
    cam.use_z_depth_buffer = c.num("use_z_depth", 0) != 0;
    cam.use_reflection = c.num("use_reflection", 0) != 0;
    cam.use_refraction = c.num("use_refraction", 0) != 0;
}

static void apply_env(const config& c, const std::string& base_dir, EnvironmentSettings& env) {
    std::string mode = c.str("env_mode", "solid");
    env.intensity = c.num("env_intensity", 1.0);
    if (mode == "solid") {
        env.set_mode(EnvironmentSettings::SOLID_COLOR);
        env.background_color = c.v3("background", vec3(0, 0, 0));
    } else if (mode == "hdr") {
        std::string p = c.str("hdr_path", "");
        if (!p.empty() && p[0] != '/' && !base_dir.empty()) p = base_dir + "/" + p;
        env.load_hdr(p);
        env.hdri_rotation = c.num("hdri_rotation", 0.0);
        env.hdri_tilt = c.num("hdri_tilt", 0.0);
        env.hdri_roll = c.num("hdri_roll", 0.0);
    } else {
        env.set_mode(EnvironmentSettings::PHYSICAL_SUN);
        env.sun_direction = c.v3("sun_direction", env.sun_direction);
        env.sun_color = c.v3("sun_color", env.sun_color);
        env.sun_intensity = c.num("sun_intensity", 1.0);
        env.sun_size = c.num("sun_size", 1.0);
    }
}

static std::string dir_of(const std::string& p) {
    size_t s = p.find_last_of('/');
    return s == std::string::npos ? std::string() : p.substr(0, s);
}

struct scene_bundle {
    shared_ptr<hittable> world;
    int n_leaves = 0;
};

static scene_bundle load_scene(const std::string& path, bool tag_triangles) {
    std::ifstream in(path);
    if (!in) { std::fprintf(stderr, "cannot open scene %s\n", path.c_str()); std::exit(2); }
    zscene::loader ld;
    ld.base_dir = dir_of(path);
    ld.parse(in);
    scene_bundle b;
    b.world = ld.build_world();
    b.n_leaves = ld.next_leaf_id;
    if (tag_triangles) {  // loader internals are reachable because of -fno-access-control
        for (auto& kv : ld.built) {
            auto m = std::dynamic_pointer_cast<model>(kv.second);
            if (!m) continue;
            const zscene::record& r = ld.objects.at(kv.first);
            tag_model(ld.path_of(r.args[0]), std::strtod(r.args[2].c_str(), nullptr), *m);
        }
    }
    return b;
}

// ------------------------------------------------------------------ commands
static void write_header(FILE* f, int w, int h) {
    int32_t hdr[2] = {w, h};
    std::fwrite(hdr, 4, 2, f);
}

static int cmd_ids(const std::string& scene, const std::string& cfg, const std::string& out) {
    scene_bundle sb = load_scene(scene, true);
    config c = config::load(cfg);
    camera cam;
    apply_camera(c, cam);
    cam.defocus_angle = 0.0;  // "jitter and DOF disabled"
    cam.initialize();
    const int W = cam.image_width, H = cam.image_height;
    std::vector<uint32_t> leaf(size_t(W) * H), tri(size_t(W) * H);
    std::vector<double> t(size_t(W) * H);
#pragma omp parallel for schedule(dynamic, 4)
    for (int j = 0; j < H; j++)
        for (int i = 0; i < W; i++) {
            ray r = cam.get_center_ray(i, j);
            hit_record rec;
            tl_leaf = -1;
            tl_tri = -1;
            size_t idx = size_t(j) * W + i;
            if (sb.world->hit(r, interval(0.001, infinity), rec)) {
                leaf[idx] = uint32_t(tl_leaf);
                tri[idx] = tl_tri >= 0 ? uint32_t(tl_tri) : 0u;
                t[idx] = rec.t;
            } else {
                leaf[idx] = 0xFFFFFFFFu;
                tri[idx] = 0;
                t[idx] = infinity;
            }
        }
    FILE* f = std::fopen(out.c_str(), "wb");
    if (!f) return 2;
    write_header(f, W, H);
    std::fwrite(leaf.data(), 4, leaf.size(), f);
    std::fwrite(tri.data(), 4, tri.size(), f);
    std::fwrite(t.data(), 8, t.size(), f);
    std::fclose(f);
    std::printf("{\"cmd\": \"ids\", \"width\": %d, \"height\": %d, \"leaves\": %d}\n", W, H, sb.n_leaves);
    return 0;
}

// counts world.hit calls = rays (one per path segment, camera.hpp:459, :937)
static std::atomic<unsigned long long> g_rays{0};
class ray_counter : public hittable {
public:
    explicit ray_counter(shared_ptr<hittable> o) : obj(o) {}
    bool hit(const ray& r, interval ray_t, hit_record& rec, int depth = 0, bool debug_wire = false) const override {
        static thread_local unsigned local = 0;
        if (++local == 1024) { g_rays.fetch_add(1024, std::memory_order_relaxed); local = 0; }
        return obj->hit(r, ray_t, rec, depth, debug_wire);
    }
    aabb bounding_box() const override { return obj->bounding_box(); }
    shared_ptr<hittable> obj;
};

struct render_result {
    double seconds = 0;
    int threads = 0;
};

static render_result run_render(camera& cam, const hittable& world, const EnvironmentSettings& env, double z_max) {
    cam.reset_accumulator();  // sizes the buffers; render() itself never does (camera.hpp:209-233)
    std::atomic<bool> flag{true};
    auto t0 = std::chrono::steady_clock::now();
    cam.initialize();
    cam.execute_render_threads(world, env, cam.render_accumulator, cam.albedo_buffer, cam.normal_buffer,
                               cam.z_depth_buffer, cam.reflection_buffer, cam.refraction_buffer, z_max, flag);
    auto t1 = std::chrono::steady_clock::now();
    render_result r;
    r.seconds = std::chrono::duration<double>(t1 - t0).count();
    r.threads = int(std::max(1u, std::thread::hardware_concurrency()));
    return r;
}

static void write_colors(FILE* f, const std::vector<color>& buf) {
    for (const color& c : buf) {
        double v[3] = {c.x(), c.y(), c.z()};
        std::fwrite(v, 8, 3, f);
    }
}

static int cmd_render(const std::string& scene, const std::string& cfg, const std::string& out, bool count_rays) {
    scene_bundle sb = load_scene(scene, false);
    config c = config::load(cfg);
    camera cam;
    apply_camera(c, cam);
    EnvironmentSettings env;
    apply_env(c, dir_of(cfg), env);
    shared_ptr<hittable> world = sb.world;
    if (count_rays) world = make_shared<ray_counter>(world);
    if (c.num("bvh_debug", 0) != 0) {  // BVH debug view (bvh.hpp:56-110, camera.hpp:937-954; GUI: main.cpp:1039-1083)
        global_settings::bvh_debug_mode = true;
        global_settings::debug_bvh_level = int(c.num("debug_bvh_level", -1));
        global_settings::bvh_thickness = float(c.num("bvh_thickness", 0.01));
    }
    render_result r = run_render(cam, *world, env, c.num("z_depth_max_dist", 20.0));
    if (out != "-") {
        FILE* f = std::fopen(out.c_str(), "wb");
        if (!f) return 2;
        write_header(f, cam.image_width, cam.image_height);
        write_colors(f, cam.render_accumulator);
        write_colors(f, cam.albedo_buffer);
        write_colors(f, cam.normal_buffer);
        write_colors(f, cam.z_depth_buffer);
        write_colors(f, cam.reflection_buffer);
        write_colors(f, cam.refraction_buffer);
        std::fclose(f);
    }
    double paths = double(cam.image_width) * cam.image_height * cam.samples_per_pixel;
    std::printf("{\"cmd\": \"render\", \"width\": %d, \"height\": %d, \"spp\": %d, \"seconds\": %.6f, \"threads\": %d, "
                "\"paths\": %.0f, \"mpaths_per_s\": %.6f, \"rays\": %llu}\n",
                cam.image_width, cam.image_height, cam.samples_per_pixel, r.seconds, r.threads, paths,
                paths / r.seconds * 1e-6, count_rays ? (unsigned long long)g_rays.load() : 0ull);
    return 0;
}

// N independent renders -> mean and variance of the mean per pixel channel (float32)
static int cmd_stats(const std::string& scene, const std::string& cfg, const std::string& out, int batches) {
    scene_bundle sb = load_scene(scene, false);
    config c = config::load(cfg);
    camera cam;
    apply_camera(c, cam);
    EnvironmentSettings env;
    apply_env(c, dir_of(cfg), env);
    const size_t n = size_t(cam.image_width) * cam.image_height * 3;
    std::vector<double> mean(n, 0.0), m2(n, 0.0);
    double secs = 0;
    for (int b = 0; b < batches; b++) {
        render_result r = run_render(cam, *sb.world, env, c.num("z_depth_max_dist", 20.0));
        secs += r.seconds;
        for (size_t i = 0; i < n; i++) {
            double x = cam.render_accumulator[i / 3][int(i % 3)];
            double d = x - mean[i];
            mean[i] += d / (b + 1);
            m2[i] += d * (x - mean[i]);
        }
    }
    std::vector<float> fm(n), fv(n);
    for (size_t i = 0; i < n; i++) {
        fm[i] = float(mean[i]);
        fv[i] = batches > 1 ? float(m2[i] / (batches - 1) / batches) : 0.f;
    }
    FILE* f = std::fopen(out.c_str(), "wb");
    if (!f) return 2;
    write_header(f, cam.image_width, cam.image_height);
    std::fwrite(fm.data(), 4, n, f);
    std::fwrite(fv.data(), 4, n, f);
    std::fclose(f);
    std::printf("{\"cmd\": \"stats\", \"width\": %d, \"height\": %d, \"spp\": %d, \"batches\": %d, \"seconds\": %.3f}\n",
                cam.image_width, cam.image_height, cam.samples_per_pixel, batches, secs);
    return 0;
}

// per-ray hit records of the reference for arbitrary rays (rays file: int32 n, n x 3 float32 origins,
// n x 3 float32 directions; output: n x 19 float64 = t, leaf, tri, front_face, p, normal, u, v, albedo, emitted)
static int cmd_probe(const std::string& scene, const std::string& cfg, const std::string& rays, const std::string& out) {
    scene_bundle sb = load_scene(scene, true);
    config c = config::load(cfg);
    camera cam;
    apply_camera(c, cam);
    cam.initialize();
    EnvironmentSettings env;
    apply_env(c, dir_of(cfg), env);
    FILE* f = std::fopen(rays.c_str(), "rb");
    if (!f) return 2;
    int32_t n = 0;
    if (std::fread(&n, 4, 1, f) != 1 || n < 0) return 2;
    std::vector<float> o(size_t(n) * 3), d(size_t(n) * 3);
    if (std::fread(o.data(), 4, o.size(), f) != o.size() || std::fread(d.data(), 4, d.size(), f) != d.size()) return 2;
    std::fclose(f);
    std::vector<double> rec_out(size_t(n) * 19, 0.0);
#pragma omp parallel for schedule(dynamic, 64)
    for (int i = 0; i < n; i++) {
        ray r(point3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), vec3(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
        hit_record rec;
        tl_leaf = -1;
        tl_tri = -1;
        double* q = &rec_out[size_t(i) * 19];
        if (sb.world->hit(r, interval(0.001, infinity), rec)) {
            q[0] = rec.t;
            q[1] = double(tl_leaf);
            q[2] = tl_tri >= 0 ? double(tl_tri) : 0.0;
            q[3] = rec.front_face ? 1.0 : 0.0;
            for (int k = 0; k < 3; k++) { q[4 + k] = rec.p[k]; q[7 + k] = rec.normal[k]; }
            q[10] = rec.u;
            q[11] = rec.v;
            color a = rec.mat->get_albedo(rec), e = rec.mat->emitted(rec.u, rec.v, rec.p);
            for (int k = 0; k < 3; k++) { q[12 + k] = a[k]; q[15 + k] = e[k]; }
        } else {
            q[0] = infinity;
            q[1] = -1.0;
            color b = cam.get_background_color(r, env);
            for (int k = 0; k < 3; k++) q[12 + k] = b[k];
        }
    }
    f = std::fopen(out.c_str(), "wb");
    if (!f) return 2;
    std::fwrite(&n, 4, 1, f);
    std::fwrite(rec_out.data(), 8, rec_out.size(), f);
    std::fclose(f);
    std::printf("{\"cmd\": \"probe\", \"rays\": %d}\n", n);
    return 0;
}


static void load_post(const config& c, post_processor& post) {
    post.exposure = float(c.num("exposure", 0.5));
    post.saturation = float(c.num("saturation", 1.0));
    post.contrast = float(c.num("contrast", 1.0));
    post.hue_shift = float(c.num("hue_shift", 0.0));
    post.vignette_intensity = float(c.num("vignette_intensity", 1.0));
    post.color_balance = c.v3("color_balance", vec3(1, 1, 1));
    post.exposure_compensation_stops = float(c.num("exposure_compensation_stops", 0.0));
    post.target_luminance = float(c.num("target_luminance", 0.12));
    post.use_aces_tone_mapping = c.num("use_aces", 0) != 0;
    post.use_auto_exposure = c.num("use_auto_exposure", 0) != 0;
    post.debug.red = c.num("debug_red", 0) != 0;
    post.debug.green = c.num("debug_green", 0) != 0;
    post.debug.blue = c.num("debug_blue", 0) != 0;
    post.debug.luminance = c.num("debug_luminance", 0) != 0;
    post.debug.bvh = c.num("debug_bvh", 0) != 0;
    post.use_bloom = c.num("use_bloom", 0) != 0;
    post.bloom_threshold = float(c.num("bloom_threshold", 1.0));
    post.bloom_intensity = float(c.num("bloom_intensity", 0.3));
    post.bloom_radius = int(c.num("bloom_radius", 4));
    post.use_sharpening = c.num("use_sharpening", 0) != 0;
    post.sharpen_amount = c.num("sharpen_amount", 0.2);
}

// Finishing chain of the reference: analyze_framebuffer + apply_auto_exposure
// (color_processing.hpp:153-210, camera.hpp:258-267) and camera::update_post_processing (camera.hpp:142-207) on a raw
// image (int32 w, h, then w*h*3 doubles).  Output: int32 w, h; final_framebuffer (w*h*3 doubles); average luminance,
// max luminance, exposure after auto-exposure (3 doubles); histogram (256 int32).
static int cmd_finish(const std::string& in, const std::string& cfg, const std::string& out) {
    FILE* f = std::fopen(in.c_str(), "rb");
    if (!f) return 2;
    int32_t wh[2];
    if (std::fread(wh, 4, 2, f) != 2) return 2;
    const size_t n = size_t(wh[0]) * wh[1];
    std::vector<color> img(n);
    for (size_t i = 0; i < n; i++) {
        double v[3];
        if (std::fread(v, 8, 3, f) != 3) return 2;
        img[i] = color(v[0], v[1], v[2]);
    }
    std::fclose(f);
    config c = config::load(cfg);
    post_processor post;
    load_post(c, post);
    const int pass = int(c.num("pass", 0));
    image_statistics stats = post.analyze_framebuffer(img);
    if (post.use_auto_exposure) post.exposure = static_cast<float>(post.apply_auto_exposure(stats));  // camera.hpp:259-263
    camera cam;
    cam.image_width = wh[0];
    cam.image_height = wh[1];
    cam.render_accumulator = img;
    cam.reflection_buffer = img;
    cam.albedo_buffer = img;
    cam.current_display_pass = pass == 0 ? render_pass::RGB : pass == 1 ? render_pass::REFLECTIONS : render_pass::ALBEDO;
    cam.update_post_processing(post, wh[0], wh[1]);
    f = std::fopen(out.c_str(), "wb");
    if (!f) return 2;
    std::fwrite(wh, 4, 2, f);
    write_colors(f, cam.final_framebuffer);
    double s3[3] = {stats.average_luminance, stats.max_luminance, post.exposure};
    std::fwrite(s3, 8, 3, f);
    std::fwrite(stats.histogram, 4, 256, f);
    std::fclose(f);
    std::printf("{\"cmd\": \"finish\", \"width\": %d, \"height\": %d, \"exposure\": %.9g}\n", wh[0], wh[1], double(post.exposure));
    return 0;
}


// Image export of the reference: camera::save_render_pass (camera.hpp:299-343) ->
// process_framebuffer_to_image (:701-780) -> stbi_write_png into ./output/.  Every pass buffer is set to the input image;
// cfg `render_pass` is the reference's enum value (common.hpp:110-118).  The PNG is read back with stbi_load and
// written as int32 w, h + w*h*3 bytes.
static int cmd_export(const std::string& in, const std::string& cfg, const std::string& out) {
    FILE* f = std::fopen(in.c_str(), "rb");
    if (!f) return 2;
    int32_t wh[2];
    if (std::fread(wh, 4, 2, f) != 2) return 2;
    const size_t n = size_t(wh[0]) * wh[1];
    std::vector<color> img(n);
    for (size_t i = 0; i < n; i++) {
        double v[3];
        if (std::fread(v, 8, 3, f) != 3) return 2;
        img[i] = color(v[0], v[1], v[2]);
    }
    std::fclose(f);
    config c = config::load(cfg);
    post_processor post;
    load_post(c, post);
    const int rp = int(c.num("render_pass", 0));
    camera cam;
    cam.image_width = wh[0];
    cam.image_height = wh[1];
    cam.render_accumulator = cam.denoise_buffer = cam.albedo_buffer = cam.normal_buffer = cam.reflection_buffer =
        cam.refraction_buffer = cam.z_depth_buffer = img;
    const std::string name = std::string("render_") + camera::pass_names[rp] + ".png";  // main.cpp:1317
    cam.save_render_pass(static_cast<render_pass>(rp), name, post);
    int w = 0, h = 0, comp = 0;
    unsigned char* d = stbi_load(("output/" + name).c_str(), &w, &h, &comp, 3);
    if (!d) return 3;
    f = std::fopen(out.c_str(), "wb");
    if (!f) return 2;
    int32_t wh2[2] = {w, h};
    std::fwrite(wh2, 4, 2, f);
    std::fwrite(d, 1, size_t(w) * h * 3, f);
    std::fclose(f);
    stbi_image_free(d);
    std::printf("{\"cmd\": \"export\", \"file\": \"output/%s\", \"width\": %d, \"height\": %d, \"channels\": %d}\n", name.c_str(), w, h, comp);
    return 0;
}

static int cmd_decode(const std::string& in, const std::string& out, bool is_hdr) {
    int w = 0, h = 0, comp = 3;
    FILE* f = nullptr;
    uint32_t hdr[4];
    std::memcpy(&hdr[0], "ZBTX", 4);
    if (is_hdr) {
        float* d = stbi_loadf(in.c_str(), &w, &h, &comp, 3);  // texture.hpp:23
        if (!d) return 3;
        f = std::fopen(out.c_str(), "wb");
        hdr[1] = uint32_t(w); hdr[2] = uint32_t(h); hdr[3] = 1;
        std::fwrite(hdr, 4, 4, f);
        std::fwrite(d, 4, size_t(w) * h * 3, f);
        stbi_image_free(d);
    } else {
        unsigned char* d = stbi_load(in.c_str(), &w, &h, &comp, 3);  // texture.hpp:31
        if (!d) return 3;
        f = std::fopen(out.c_str(), "wb");
        hdr[1] = uint32_t(w); hdr[2] = uint32_t(h); hdr[3] = 0;
        std::fwrite(hdr, 4, 4, f);
        std::fwrite(d, 1, size_t(w) * h * 3, f);
        stbi_image_free(d);
    }
    std::fclose(f);
    std::printf("{\"cmd\": \"decode\", \"width\": %d, \"height\": %d}\n", w, h);
    return 0;
}

int cmd_kat(const std::string& scene, const std::string& cfg_path, const std::string& out);  // ref_kat.inc

#include "ref_kat.inc"

int main(int argc, char** argv) {
    std::vector<std::string> a(argv + 1, argv + argc);
    if (a.size() >= 4 && a[0] == "ids") return cmd_ids(a[1], a[2], a[3]);
    if (a.size() >= 4 && a[0] == "render") return cmd_render(a[1], a[2], a[3], a.size() > 4 && a[4] == "--count-rays");
    if (a.size() >= 5 && a[0] == "stats") return cmd_stats(a[1], a[2], a[3], std::atoi(a[4].c_str()));
    if (a.size() >= 5 && a[0] == "probe") return cmd_probe(a[1], a[2], a[3], a[4]);
    if (a.size() >= 4 && a[0] == "finish") return cmd_finish(a[1], a[2], a[3]);
    if (a.size() >= 4 && a[0] == "export") return cmd_export(a[1], a[2], a[3]);
    if (a.size() >= 4 && a[0] == "decode") return cmd_decode(a[1], a[2], a[3] == "1");
    if (a.size() >= 4 && a[0] == "kat") return cmd_kat(a[1], a[2], a[3]);
    std::fprintf(stderr,
                 "usage: ref_harness ids SCENE CFG OUT | render SCENE CFG OUT|- [--count-rays] | stats SCENE CFG OUT N |\n"
                 "       decode IN OUT.zbtex ISHDR | kat SCENE CFG OUT.json\n");
    return 1;
}
