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
extern "C" void zbi_pin_flat_scene(const zb_flat_scene* f, int pin);  // abi.cu: page-lock the large arrays for fast uploads

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
        if (flags & ZB_FLATTEN_WIDE) fl.wide = 1;
        if (flags & ZB_FLATTEN_WIDE_ALL) { fl.wide = 1; fl.wide_all = true; }
        if (const char* e = std::getenv("ZB_WIDE_BUILD")) {  // A/B runs: 1 = for scenes with meshes, 2 = for every scene
            fl.wide = std::atoi(e) > 0 ? 1 : fl.wide;
            fl.wide_all = std::atoi(e) > 1;
        }
        if (const char* e = std::getenv("ZB_TLAS_LEAF")) fl.tlas_leaf = std::max(1, std::min(8, std::atoi(e)));  // tuning knobs
        if (const char* e = std::getenv("ZB_BLAS_LEAF")) fl.blas_leaf = std::max(1, std::min(8, std::atoi(e)));
        EnvironmentSettings env;
        if (!ld.env_hdr_path.empty()) env.load_hdr(ld.path_of(ld.env_hdr_path));  // environment.hpp:46-69
        std::unique_ptr<flat_scene> fs = fl.run(*world, &env);
        const zb_flat_scene* view = &fs->view;
        zbi_pin_flat_scene(view, 1);
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
//   (reserved = wide_stack_need), then 15 sections, each = u64 byte count + payload padded to 16 bytes, in the order of
//   zb_flat_scene's pointers (the 8-wide nodes and the item bounds last).
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

extern "C" int zb_flat_scene_features(const zb_flat_scene* f, uint32_t* features) {
    if (!f || !features) return fail_host(ZB_ERR_INVALID, "zb_flat_scene_features: NULL argument");
    uint32_t feat = 0;
    for (uint32_t i = 0; i < f->n_items; i++) {
        const uint32_t kind = f->items[i].kind_flags & 0xFFu;
        feat |= kind == ZB_ITEM_INSTANCE ? ZB_F_INST : kind == ZB_ITEM_CUBE ? ZB_F_CUBE : kind == ZB_ITEM_TRIANGLE ? ZB_F_TRI
              : kind == ZB_ITEM_MEDIUM ? ZB_F_MEDIA : 0u;
    }
    for (uint32_t i = 0; i < f->n_instances; i++) {
        const uint32_t g = f->instances[i].geom_kind;
        feat |= g == ZB_GEOM_MESH ? uint32_t(ZB_F_MESH | ZB_F_TRI) : g == ZB_GEOM_CUBE ? ZB_F_CUBE : g == ZB_GEOM_TRIANGLE ? ZB_F_TRI : 0u;
    }
    for (uint32_t i = 0; i < f->n_media; i++) {
        const uint32_t k = f->media[i].boundary_kind;
        feat |= ZB_F_MEDIA | (k == ZB_ITEM_CUBE ? ZB_F_CUBE : k == ZB_ITEM_INSTANCE ? ZB_F_INST : 0u);
    }
    for (uint32_t i = 0; i < f->n_materials; i++)
        if (f->materials[i].albedo_tex != ZB_NO_TEXTURE || f->materials[i].bump_tex != ZB_NO_TEXTURE) feat |= ZB_F_TEX;
    *features = feat;
    return ZB_OK;
}

extern "C" int zb_flat_scene_save(const zb_flat_scene* v, const char* path) {
    if (!v || !path) return fail_host(ZB_ERR_INVALID, "zb_flat_scene_save: NULL argument");
    std::ofstream f(path, std::ios::binary);
    if (!f) return fail_host(ZB_ERR_IO, std::string("cannot create ") + path);
    const uint32_t hdr[5] = {v->abi_version, v->n_tlas_nodes, v->n_leaves, v->pad, v->wide_stack_need};
    f.write("ZBFS", 4);
    f.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
    bool ok = put(f, span_of(v->nodes, v->n_nodes)) && put(f, span_of(v->items, v->n_items)) && put(f, span_of(v->cubes, v->n_cubes)) &&
              put(f, span_of(v->tris, v->n_tris)) && put(f, span_of(v->tri_normals, v->n_tris)) &&
              put(f, span_of(v->instances, v->n_instances)) && put(f, span_of(v->meshes, v->n_meshes)) &&
              put(f, span_of(v->media, v->n_media)) && put(f, span_of(v->materials, v->n_materials)) &&
              put(f, span_of(v->textures, v->n_textures)) && put(f, span_of(v->spheres_obj, uint64_t(v->n_spheres_obj) * 4)) &&
              put(f, span_of(v->texels_u8, v->n_texels_u8 * 4)) && put(f, span_of(v->texels_f32, v->n_texels_f32 * 4)) &&
              put(f, span_of(v->wide_nodes, v->n_wide_nodes)) &&
              put(f, span_of(v->item_bounds, v->item_bounds ? uint64_t(v->n_items) * 8 : 0));
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
    fs->wide_stack_need = hdr[4];
    bool ok = get(f, fs->nodes) && get(f, fs->items) && get(f, fs->cubes) && get(f, fs->tris) && get(f, fs->tri_normals) &&
              get(f, fs->instances) && get(f, fs->meshes) && get(f, fs->media) && get(f, fs->materials) && get(f, fs->textures) &&
              get(f, fs->spheres_obj) && get(f, fs->texels_u8) && get(f, fs->texels_f32) && get(f, fs->wide_nodes) && get(f, fs->item_bounds);
    if (!ok || fs->tris.size() != fs->tri_normals.size()) return fail_host(ZB_ERR_INVALID, std::string(path) + ": truncated or malformed flat scene");
    const zb_flat_scene* view = fs->finish();  // zb_scene_upload validates every index before anything is used
    zbi_pin_flat_scene(view, 1);
    std::lock_guard<std::mutex> lock(g_mu);
    g_live[view] = std::move(fs);
    *out = view;
    return ZB_OK;
}

extern "C" int zb_flat_scene_free(const zb_flat_scene* flat) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (g_live.count(flat)) zbi_pin_flat_scene(flat, 0);
    g_live.erase(flat);
    return ZB_OK;
}

// ------------------------------------------------------------------ image files
// The reference writes its passes with stbi_write_png (camera.hpp:779) and reads HDR environments with stbi_loadf
// (texture.hpp:23).  These writers are written from the PNG (RFC 2083) and Radiance RGBE format definitions: 8-bit
// RGB PNG, filter Sub on every row, one zlib stream; flat (not run-length encoded) RGBE.
#include <zlib.h>

#include <cmath>
#include <cstdio>
#include <cstring>

namespace {
void put_be32(std::vector<unsigned char>& v, uint32_t x) {
    v.push_back(uint8_t(x >> 24));
    v.push_back(uint8_t(x >> 16));
    v.push_back(uint8_t(x >> 8));
    v.push_back(uint8_t(x));
}
void png_chunk(std::vector<unsigned char>& out, const char* type, const unsigned char* data, size_t n) {
    put_be32(out, uint32_t(n));
    const size_t at = out.size();
    out.insert(out.end(), type, type + 4);
    if (n) out.insert(out.end(), data, data + n);
    put_be32(out, uint32_t(crc32(0L, out.data() + at, uInt(n + 4))));
}
}  // namespace

extern "C" int zb_write_png(const char* path, int width, int height, const unsigned char* pixels, int channels, size_t row_stride) {
    if (!path || !pixels || width < 1 || height < 1 || (channels != 3 && channels != 4))
        return fail_host(ZB_ERR_INVALID, "zb_write_png: bad argument");
    if (!row_stride) row_stride = size_t(width) * channels;
    const size_t line = size_t(width) * 3 + 1;
    std::vector<unsigned char> raw(line * height);
    for (int y = 0; y < height; y++) {
        const unsigned char* src = pixels + size_t(y) * row_stride;
        unsigned char* dst = raw.data() + size_t(y) * line;
        dst[0] = 1;  // filter Sub: byte - byte of the pixel to the left
        unsigned char prev[3] = {0, 0, 0};
        for (int x = 0; x < width; x++)
            for (int k = 0; k < 3; k++) {
                const unsigned char v = src[size_t(x) * channels + k];
                dst[1 + size_t(x) * 3 + k] = uint8_t(v - prev[k]);
                prev[k] = v;
            }
    }
    uLongf zlen = compressBound(uLong(raw.size()));
    std::vector<unsigned char> z(zlen);
    if (compress2(z.data(), &zlen, raw.data(), uLong(raw.size()), 4) != Z_OK) return fail_host(ZB_ERR_INVALID, "zb_write_png: deflate failed");
    std::vector<unsigned char> out = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
    std::vector<unsigned char> ihdr;
    put_be32(ihdr, uint32_t(width));
    put_be32(ihdr, uint32_t(height));
    const unsigned char tail[5] = {8, 2, 0, 0, 0};  // 8 bits, colour type 2 (RGB), deflate, adaptive filtering, no interlace
    ihdr.insert(ihdr.end(), tail, tail + 5);
    png_chunk(out, "IHDR", ihdr.data(), ihdr.size());
    png_chunk(out, "IDAT", z.data(), zlen);
    png_chunk(out, "IEND", nullptr, 0);
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail_host(ZB_ERR_IO, std::string("zb_write_png: cannot open ") + path);
    const bool ok = std::fwrite(out.data(), 1, out.size(), f) == out.size();
    std::fclose(f);
    return ok ? ZB_OK : fail_host(ZB_ERR_IO, std::string("zb_write_png: short write to ") + path);
}

extern "C" int zb_write_hdr(const char* path, int width, int height, const float* pixels, int channels, size_t row_stride_floats) {
    if (!path || !pixels || width < 1 || height < 1 || channels < 3) return fail_host(ZB_ERR_INVALID, "zb_write_hdr: bad argument");
    if (!row_stride_floats) row_stride_floats = size_t(width) * channels;
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail_host(ZB_ERR_IO, std::string("zb_write_hdr: cannot open ") + path);
    std::fprintf(f, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n", height, width);
    std::vector<unsigned char> row(size_t(width) * 4);
    bool ok = true;
    for (int y = 0; y < height && ok; y++) {
        const float* src = pixels + size_t(y) * row_stride_floats;
        for (int x = 0; x < width; x++) {
            float r = src[size_t(x) * channels], g = src[size_t(x) * channels + 1], b = src[size_t(x) * channels + 2];
            r = std::isfinite(r) && r > 0.f ? r : 0.f;
            g = std::isfinite(g) && g > 0.f ? g : 0.f;
            b = std::isfinite(b) && b > 0.f ? b : 0.f;
            const float m = std::max(r, std::max(g, b));
            unsigned char* q = row.data() + size_t(x) * 4;
            if (m < 1e-32f) {
                q[0] = q[1] = q[2] = q[3] = 0;
            } else {
                int e = 0;
                const float scale = std::frexp(m, &e) * 256.0f / m;  // shared exponent of the largest channel
                q[0] = uint8_t(r * scale);
                q[1] = uint8_t(g * scale);
                q[2] = uint8_t(b * scale);
                q[3] = uint8_t(e + 128);
            }
        }
        ok = std::fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    std::fclose(f);
    return ok ? ZB_OK : fail_host(ZB_ERR_IO, std::string("zb_write_hdr: short write to ") + path);
}

// camera::pass_names (camera.hpp:66-74) and the export file name of main.cpp:1317
extern "C" const char* zb_pass_file_name(int render_pass) {
    static const char* names[] = {"render_RGB.png", "render_Denoise.png", "render_Albedo.png", "render_Normals.png",
                                  "render_Reflections.png", "render_Refractions.png", "render_Z-Depth.png"};
    return render_pass >= 0 && render_pass < 7 ? names[render_pass] : nullptr;
}

// ------------------------------------------------------------------ host side of the frame copy
// float -> double widening of a chunk of the frame into the caller's buffer (abi_render.cu widen_chunks).  The AVX2
// form converts 8 floats per step and writes with streaming stores: the destination is written once and never read, so
// skipping the read-for-ownership halves the memory traffic of this step.
#if defined(__x86_64__)
#include <immintrin.h>
__attribute__((target("avx2"))) static void widen_avx2(const float* s, double* d, size_t n) {
    size_t i = 0;
    while (i < n && (reinterpret_cast<uintptr_t>(d + i) & 31u)) { d[i] = double(s[i]); i++; }
    for (; i + 8 <= n; i += 8) {
        const __m256 v = _mm256_loadu_ps(s + i);
        _mm256_stream_pd(d + i, _mm256_cvtps_pd(_mm256_castps256_ps128(v)));
        _mm256_stream_pd(d + i + 4, _mm256_cvtps_pd(_mm256_extractf128_ps(v, 1)));
    }
    for (; i < n; i++) d[i] = double(s[i]);
    _mm_sfence();
}
#endif

void zbh_widen_f32_f64(const float* s, double* d, size_t n) {
#if defined(__x86_64__)
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2) { widen_avx2(s, d, n); return; }
#endif
    for (size_t i = 0; i < n; i++) d[i] = double(s[i]);
}
