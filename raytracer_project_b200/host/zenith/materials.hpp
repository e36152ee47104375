// zenith/materials.hpp — textures, materials and the material library of the host scene API.
// Same class names and constructor signatures as the reference (texture.hpp:12-133,
// material.hpp:58-279, constant_medium.hpp:9-22, material_library.hpp:11-67) but these are pure
// DESCRIPTIONS: scatter()/emitted()/value() are evaluated by the CUDA shade code from the flattened
// tagged unions (zb_material / zb_texture), never on the host.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <map>
#include <string>
#include <vector>

#include "math.hpp"

// The reference's own decoder (third-party, public domain; host/third_party/README.md): static copy per translation unit,
// so header-only users (tests/cpp, oracle/zoracle.cpp) need no extra object file.
#if defined(__GNUC__)
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Wunused-function"
#pragma GCC diagnostic ignored "-Wunused-but-set-variable"
#pragma GCC diagnostic ignored "-Wsign-compare"
#endif
#ifndef STBI_INCLUDE_STB_IMAGE_H
#define STB_IMAGE_STATIC
#define STB_IMAGE_IMPLEMENTATION
#include "../third_party/stb_image.h"
#endif
#if defined(__GNUC__)
#pragma GCC diagnostic pop
#endif

namespace zenith {

class texture {
public:
    enum class kind_t { solid, checker, image };
    virtual ~texture() = default;
    virtual kind_t kind() const = 0;
};

class solid_color : public texture {
public:
    solid_color(const color& albedo) : albedo(albedo) {}
    solid_color(double r, double g, double b) : albedo(r, g, b) {}
    kind_t kind() const override { return kind_t::solid; }
    color albedo;
};

class checker_texture : public texture {
public:
    checker_texture(double scale, shared_ptr<texture> odd, shared_ptr<texture> even)
        : inv_scale(1.0 / scale), odd(odd), even(even) {}
    checker_texture(double scale, color c1, color c2)
        : inv_scale(1.0 / scale), odd(make_shared<solid_color>(c1)), even(make_shared<solid_color>(c2)) {}
    kind_t kind() const override { return kind_t::checker; }
    double inv_scale;
    shared_ptr<texture> odd, even;
};

// image_texture(filename, is_hdr) texture.hpp:14-43.  Decoded like the reference does, with the stb_image it vendors:
// stbi_loadf(.., 3) when is_hdr, else stbi_load(.., 3) (texture.hpp:23, :31) — JPEG (baseline and progressive), PNG, HDR
// and the rest of stb's formats, texel for texel.  A ".zbtex" raw dump ("ZBTX" u32 width u32 height u32 is_float, then
// RGB8 or RGB32F rows, top row first) is read when the file itself is one, or as a side-car "NAME.zbtex" cache when NAME
// cannot be decoded.  A file that cannot be read gives the reference's "missing" texture (cyan, texture.hpp:52-54).
class image_texture : public texture {
public:
    image_texture(const char* filename, bool is_hdr = false) : is_float(is_hdr) { load(filename); }
    image_texture(int w, int h, const uint8_t* rgb8) : width(w), height(h), is_float(false), u8(rgb8, rgb8 + size_t(w) * h * 3) {}
    image_texture(int w, int h, const float* rgbf) : width(w), height(h), is_float(true), f32(rgbf, rgbf + size_t(w) * h * 3) {}
    kind_t kind() const override { return kind_t::image; }
    bool missing() const { return width <= 0 || height <= 0; }
    int width = 0, height = 0;
    bool is_float = false;
    std::vector<uint8_t> u8;
    std::vector<float> f32;

private:
    void load(const char* filename) {
        std::string path(filename);
        if (read_zbtex(path) || read_stb(path) || read_zbtex(path + ".zbtex") || read_radiance_hdr(path)) return;
        std::cerr << (is_float ? "ERROR: Could not load HDR: " : "ERROR: Could not load texture: ") << filename << "\n";  // texture.hpp:25, :34
        width = height = 0;
    }

    bool read_stb(const std::string& path) {
        int w = 0, h = 0, comp = 3;
        if (is_float) {
            float* d = stbi_loadf(path.c_str(), &w, &h, &comp, 3);
            if (!d) return false;
            f32.assign(d, d + size_t(w) * h * 3);
            stbi_image_free(d);
        } else {
            unsigned char* d = stbi_load(path.c_str(), &w, &h, &comp, 3);
            if (!d) return false;
            u8.assign(d, d + size_t(w) * h * 3);
            stbi_image_free(d);
        }
        width = w;
        height = h;
        return true;
    }

    bool read_zbtex(const std::string& path) {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) return false;
        uint32_t hdr[4] = {0, 0, 0, 0};
        bool ok = std::fread(hdr, 4, 4, f) == 4 && std::memcmp(&hdr[0], "ZBTX", 4) == 0 && hdr[1] > 0 && hdr[2] > 0;
        if (ok) {
            width = int(hdr[1]);
            height = int(hdr[2]);
            is_float = hdr[3] != 0;
            size_t n = size_t(width) * height * 3;
            if (is_float) {
                f32.resize(n);
                ok = std::fread(f32.data(), 4, n, f) == n;
            } else {
                u8.resize(n);
                ok = std::fread(u8.data(), 1, n, f) == n;
            }
        }
        std::fclose(f);
        if (!ok) { width = height = 0; u8.clear(); f32.clear(); }
        return ok;
    }

    bool read_radiance_hdr(const std::string& path) {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) return false;
        auto line = [&](std::string& out) {
            out.clear();
            int c;
            while ((c = std::fgetc(f)) != EOF && c != '\n') out.push_back(char(c));
            return c != EOF || !out.empty();
        };
        std::string l;
        bool ok = line(l) && (l == "#?RADIANCE" || l == "#?RGBE"), fmt = false;
        while (ok && line(l) && !l.empty())
            if (l == "FORMAT=32-bit_rle_rgbe") fmt = true;
        int w = 0, h = 0;
        ok = ok && fmt && line(l) && std::sscanf(l.c_str(), "-Y %d +X %d", &h, &w) == 2 && w > 0 && h > 0;
        std::vector<uint8_t> rgbe;
        if (ok) {
            rgbe.resize(size_t(w) * h * 4);
            std::vector<uint8_t> rest;
            int c;
            while ((c = std::fgetc(f)) != EOF) rest.push_back(uint8_t(c));
            size_t at = 0;
            auto need = [&](size_t n) { return at + n <= rest.size(); };
            bool flat = w < 8 || w >= 32768;
            for (int y = 0; ok && y < h && !flat; y++) {
                if (!need(4)) { ok = false; break; }
                if (rest[at] != 2 || rest[at + 1] != 2 || (rest[at + 2] & 0x80)) {
                    if (y == 0) flat = true; else ok = false;
                    break;
                }
                if (((int(rest[at + 2]) << 8) | rest[at + 3]) != w) { ok = false; break; }
                at += 4;
                for (int k = 0; ok && k < 4; k++) {
                    int x = 0;
                    while (ok && x < w) {
                        if (!need(1)) { ok = false; break; }
                        int count = rest[at++];
                        if (count > 128) {
                            count -= 128;
                            if (!need(1) || x + count > w) { ok = false; break; }
                            uint8_t v = rest[at++];
                            for (int z = 0; z < count; z++) rgbe[(size_t(y) * w + x++) * 4 + k] = v;
                        } else {
                            if (count == 0 || !need(size_t(count)) || x + count > w) { ok = false; break; }
                            for (int z = 0; z < count; z++) rgbe[(size_t(y) * w + x++) * 4 + k] = rest[at++];
                        }
                    }
                }
            }
            if (ok && flat) {
                ok = rest.size() >= rgbe.size();
                if (ok) std::memcpy(rgbe.data(), rest.data(), rgbe.size());
            }
        }
        std::fclose(f);
        if (!ok) return false;
        width = w;
        height = h;
        is_float = true;
        f32.resize(size_t(w) * h * 3);
        for (size_t i = 0; i < size_t(w) * h; i++) {
            const uint8_t* px = &rgbe[4 * i];
            float s = px[3] ? std::ldexp(1.0f, int(px[3]) - 136) : 0.0f;
            for (int k = 0; k < 3; k++) f32[3 * i + k] = px[3] ? float(px[k]) * s : 0.0f;
        }
        return true;
    }
};

class material {
public:
    enum class kind_t { lambertian, metal, dielectric, diffuse_light, isovolumetric };
    virtual ~material() = default;
    virtual kind_t kind() const = 0;
    shared_ptr<texture> tex;       // albedo / emission / phase albedo
    shared_ptr<texture> bump;      // my_bump_texture (material.hpp:105, :160, :233)
    double bump_strength = 1.0;
    double param = 0.0;            // metal fuzz / dielectric refraction_index
    color tint = color(1, 1, 1);   // dielectric albedo (material.hpp:232)
};

class lambertian : public material {
public:
    lambertian(const color& albedo, shared_ptr<texture> b = nullptr, double strength = 1.0) {
        tex = make_shared<solid_color>(albedo); bump = b; bump_strength = strength;
    }
    lambertian(shared_ptr<texture> t, shared_ptr<texture> b = nullptr, double strength = 1.0) {
        tex = t; bump = b; bump_strength = strength;
    }
    kind_t kind() const override { return kind_t::lambertian; }
};

class metal : public material {
public:
    metal(shared_ptr<texture> a, double f, shared_ptr<texture> b = nullptr, double strength = 1.0) {
        tex = a; param = f < 1 ? f : 1; bump = b; bump_strength = strength;
    }
    metal(const color& a, double f, shared_ptr<texture> b = nullptr, double strength = 1.0) {
        tex = make_shared<solid_color>(a); param = f < 1 ? f : 1; bump = b; bump_strength = strength;
    }
    kind_t kind() const override { return kind_t::metal; }
};

class dielectric : public material {
public:
    dielectric(double ri, const color& a = color(1.0, 1.0, 1.0)) { param = ri; tint = a; }
    dielectric(double ri, const color& a, shared_ptr<texture> b, double strength) {
        param = ri; tint = a; bump = b; bump_strength = strength;
    }
    dielectric(double ri, shared_ptr<texture> b, double strength) { param = ri; bump = b; bump_strength = strength; }
    kind_t kind() const override { return kind_t::dielectric; }
};

class diffuse_light : public material {
public:
    diffuse_light(shared_ptr<texture> a) { tex = a; }
    diffuse_light(color c) { tex = make_shared<solid_color>(c); }
    kind_t kind() const override { return kind_t::diffuse_light; }
};

class isovolumetric : public material {
public:
    isovolumetric(color c) { tex = make_shared<solid_color>(c); }
    isovolumetric(shared_ptr<texture> t) { tex = t; }
    kind_t kind() const override { return kind_t::isovolumetric; }
};

// material_library.hpp:11-67
class MaterialLibrary {
public:
    void add(const std::string& name, shared_ptr<material> m) { library[name] = m; }
    shared_ptr<material> get(const std::string& name) const {
        auto it = library.find(name);
        if (it != library.end()) return it->second;
        std::cerr << "Material '" << name << "' not found in library. Returning nullptr." << std::endl;
        return nullptr;
    }
    std::vector<std::string> get_material_names() const {
        std::vector<std::string> names;
        for (const auto& kv : library) names.push_back(kv.first);
        return names;
    }
    std::vector<std::string> get_emissive_names() const {
        std::vector<std::string> out;
        for (const auto& kv : library)
            if (is_light_name(kv.first)) out.push_back(kv.first);
        return out;
    }
    std::vector<std::string> get_regular_names() const {
        std::vector<std::string> out;
        for (const auto& kv : library)
            if (!is_light_name(kv.first) && kv.first != "reflective_checker_mat") out.push_back(kv.first);
        return out;
    }

private:
    static bool is_light_name(const std::string& n) {
        return n.find("neon") != std::string::npos || n.find("emissive") != std::string::npos;
    }
    std::map<std::string, shared_ptr<material>> library;
};

}  // namespace zenith
