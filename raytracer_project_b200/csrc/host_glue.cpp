// host_glue.cpp — C handles over the C++20 host side: zscene text -> scene API -> flatten().
// This is the piece a Zenith maintainer would NOT need (they already hold a hittable graph and call
// zenith::flattener directly); it exists so that language bindings and tests can
// drive the same host code through the C ABI.
#include <algorithm>
#include <cstdlib>
#include <fstream>
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

extern "C" int zb_flat_scene_free(const zb_flat_scene* flat) {
    std::lock_guard<std::mutex> lock(g_mu);
    g_live.erase(flat);
    return ZB_OK;
}
