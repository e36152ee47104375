// host_glue.cpp — C handles over the C++20 host side: zscene text -> scene API -> flatten().
// This is the piece a Zenith maintainer would NOT need (they already hold a hittable graph and call
// zenith::flattener directly); it exists so that language bindings and tests can
// drive the same host code through the C ABI.
#include <algorithm>
#include <cstdlib>
#include <fstream>
#include <vector>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>

#include "../../include/zenith_b200.h"
#include "../host/flatten.hpp"

using namespace zenith;
#include "../host/zscene.hpp"

std::string& zb_err_slot();
static int fail_host(int code, const std::string& msg) {
    zb_err_slot() = msg;
    return code;
}

namespace {
std::mutex g_mu;
std::map<const zb_flat_scene*, std::unique_ptr<flat_scene>> g_live;

int load(std::istream& in, const std::string& base_dir, const zb_flat_scene** out, uint32_t flags = 0) {
    if (!out) return fail_host(ZB_ERR_INVALID, "zb_flat_scene_from_zscene: out is NULL");
    try {
        zscene::loader ld;
        ld.base_dir = base_dir;
        ld.parse(in);
        std::shared_ptr<hittable> world = ld.build_world();
        flattener fl;
        fl.gpu_blas = (flags & ZB_FLATTEN_GPU_BLAS) != 0;
        if (const char* e = std::getenv("ZB_TLAS_LEAF")) fl.tlas_leaf = std::max(1, std::min(8, std::atoi(e)));  // tuning knobs
        if (const char* e = std::getenv("ZB_BLAS_LEAF")) fl.blas_leaf = std::max(1, std::min(8, std::atoi(e)));
        EnvironmentSettings env;
        if (!ld.env_hdr_path.empty()) env.load_hdr(ld.path_of(ld.env_hdr_path));  // environment.hpp:46-69
        std::unique_ptr<flat_scene> fs = fl.run(*world, &env);
        const zb_flat_scene* view = &fs->view;
        std::lock_guard<std::mutex> lock(g_mu);
        g_live[view] = std::move(fs);
        *out = view;
        return ZB_OK;
    } catch (const std::exception& e) {
        return fail_host(ZB_ERR_INVALID, e.what());
    }
}
}  // namespace

extern "C" int zb_flat_scene_from_zscene(const char* path, const zb_flat_scene** out) {
    if (!path) return fail_host(ZB_ERR_INVALID, "path is NULL");
    std::ifstream in(path);
    if (!in) return fail_host(ZB_ERR_IO, std::string("cannot open ") + path);
    std::string p(path), dir;
    size_t slash = p.find_last_of('/');
    if (slash != std::string::npos) dir = p.substr(0, slash);
    return load(in, dir, out);
}

extern "C" int zb_flat_scene_from_zscene_text(const char* text, size_t len, const zb_flat_scene** out) {
    if (!text) return fail_host(ZB_ERR_INVALID, "text is NULL");
    std::istringstream in(std::string(text, len));
    return load(in, "", out);
}

extern "C" int zb_flat_scene_from_zscene_ex(const char* path, const char* text, size_t len, uint32_t flags, const zb_flat_scene** out) {
    if (path) {
        std::ifstream in(path);
        if (!in) return fail_host(ZB_ERR_IO, std::string("cannot open ") + path);
        std::string p(path), dir;
        size_t slash = p.find_last_of('/');
        if (slash != std::string::npos) dir = p.substr(0, slash);
        return load(in, dir, out, flags);
    }
    if (!text) return fail_host(ZB_ERR_INVALID, "path and text are both NULL");
    std::istringstream in(std::string(text, len));
    return load(in, "", out, flags);
}


// ---------------------------------------------------------------------------------- flat scene on disk
// "ZBFS" container: a flattened scene written once (rank 0, or an offline tool) and loaded by every rank / every
// later run without re-parsing OBJ files and re-building the BVH.  Little-endian PODs exactly as in zenith_b200.h:
//   char magic[4] = "ZBFS"; u32 abi_version; u32 n_tlas_nodes, n_leaves, env_hdr_texture, reserved;
//   then 13 sections, each = u64 byte count + payload padded to 16 bytes, in the order of zb_flat_scene's pointers.
namespace {
template <class T> bool put(std::ofstream& f, const std::vector<T>& v) {
    const uint64_t bytes = uint64_t(v.size()) * sizeof(T);
    f.write(reinterpret_cast<const char*>(&bytes), 8);
    if (bytes) f.write(reinterpret_cast<const char*>(v.data()), std::streamsize(bytes));
    static const char zero[16] = {0};
    if (bytes % 16) f.write(zero, std::streamsize(16 - bytes % 16));
    return bool(f);
}
template <class T> bool get(std::ifstream& f, std::vector<T>& v) {
    uint64_t bytes = 0;
    f.read(reinterpret_cast<char*>(&bytes), 8);
    if (!f || bytes % sizeof(T) || bytes > (uint64_t(1) << 40)) return false;
    v.resize(size_t(bytes / sizeof(T)));
    if (bytes) f.read(reinterpret_cast<char*>(v.data()), std::streamsize(bytes));
    if (bytes % 16) f.ignore(std::streamsize(16 - bytes % 16));
    return bool(f);
}
template <class T> std::vector<T> span_of(const T* p, uint64_t n) { return p && n ? std::vector<T>(p, p + n) : std::vector<T>(); }
}  // namespace

extern "C" int zb_flat_scene_save(const zb_flat_scene* v, const char* path) {
    if (!v || !path) return fail_host(ZB_ERR_INVALID, "zb_flat_scene_save: NULL argument");
    std::ofstream f(path, std::ios::binary);
    if (!f) return fail_host(ZB_ERR_IO, std::string("cannot create ") + path);
    const uint32_t hdr[5] = {v->abi_version, v->n_tlas_nodes, v->n_leaves, v->pad, 0u};
    f.write("ZBFS", 4);
    f.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
    bool ok = put(f, span_of(v->nodes, v->n_nodes)) && put(f, span_of(v->items, v->n_items)) && put(f, span_of(v->cubes, v->n_cubes)) &&
              put(f, span_of(v->tris, v->n_tris)) && put(f, span_of(v->tri_normals, v->n_tris)) &&
              put(f, span_of(v->instances, v->n_instances)) && put(f, span_of(v->meshes, v->n_meshes)) &&
              put(f, span_of(v->media, v->n_media)) && put(f, span_of(v->materials, v->n_materials)) &&
              put(f, span_of(v->textures, v->n_textures)) && put(f, span_of(v->spheres_obj, uint64_t(v->n_spheres_obj) * 4)) &&
              put(f, span_of(v->texels_u8, v->n_texels_u8 * 4)) && put(f, span_of(v->texels_f32, v->n_texels_f32 * 4));
    f.close();
    if (!ok || !f) return fail_host(ZB_ERR_IO, std::string("short write to ") + path);
    return ZB_OK;
}

extern "C" int zb_flat_scene_load(const char* path, const zb_flat_scene** out) {
    if (!path || !out) return fail_host(ZB_ERR_INVALID, "zb_flat_scene_load: NULL argument");
    std::ifstream f(path, std::ios::binary);
    if (!f) return fail_host(ZB_ERR_IO, std::string("cannot open ") + path);
    char magic[4];
    uint32_t hdr[5];
    f.read(magic, 4);
    f.read(reinterpret_cast<char*>(hdr), sizeof(hdr));
    if (!f || std::string(magic, 4) != "ZBFS") return fail_host(ZB_ERR_INVALID, std::string(path) + " is not a ZBFS flat scene");
    if (hdr[0] != ZB_ABI_VERSION) return fail_host(ZB_ERR_INVALID, std::string(path) + ": flat scene written for another ABI version");
    auto fs = std::make_unique<flat_scene>();
    fs->n_tlas_nodes = hdr[1];
    fs->n_leaves = hdr[2];
    fs->env_hdr_texture = hdr[3];
    bool ok = get(f, fs->nodes) && get(f, fs->items) && get(f, fs->cubes) && get(f, fs->tris) && get(f, fs->tri_normals) &&
              get(f, fs->instances) && get(f, fs->meshes) && get(f, fs->media) && get(f, fs->materials) && get(f, fs->textures) &&
              get(f, fs->spheres_obj) && get(f, fs->texels_u8) && get(f, fs->texels_f32);
    if (!ok || fs->tris.size() != fs->tri_normals.size()) return fail_host(ZB_ERR_INVALID, std::string(path) + ": truncated or malformed flat scene");
    const zb_flat_scene* view = fs->finish();  // zb_scene_upload validates every index before anything is used
    std::lock_guard<std::mutex> lock(g_mu);
    g_live[view] = std::move(fs);
    *out = view;
    return ZB_OK;
}

extern "C" int zb_flat_scene_free(const zb_flat_scene* flat) {
    std::lock_guard<std::mutex> lock(g_mu);
    g_live.erase(flat);
    return ZB_OK;
}
