// zscene.hpp — neutral scene description ("zscene" text) -> scene graph, through the scene API.
//
// This loader is written ONLY against the public scene API the reference and this host share
// (sphere / cube / triangle / translate / rotate_* / scale / material_instance /
// constant_medium / model / hittable_list / bvh_node, the material and texture constructors), with
// unqualified names.  It is compiled twice: into the product library against raytracer_project_b200/
// host/zenith/*.hpp (after `using namespace zenith;`), and into the oracle harness against the
// reference's own headers — which is also the proof that the host API is a drop-in for scene code.
//
// Format, one record per line ('#' starts a comment; names are single tokens):
//   tex  NAME solid R G B | checker SCALE ODDTEX EVENTEX | image PATH ISHDR
//   mat  NAME lambertian (color R G B | tex TEX) [bump TEX STRENGTH]
//   mat  NAME metal      (color R G B | tex TEX) FUZZ [bump TEX STRENGTH]
//   mat  NAME dielectric IOR [tint R G B] [bump TEX STRENGTH]
//   mat  NAME light      (color R G B | tex TEX)
//   obj  NAME sphere CX CY CZ R MAT
//   obj  NAME cube X0 Y0 Z0 X1 Y1 Z1 MAT          obj NAME cube1 CX CY CZ MAT
//   obj  NAME triangle (9 vertex numbers) (9 normal numbers) MAT
//   obj  NAME translate CHILD DX DY DZ            obj NAME scale CHILD SX SY SZ
//   obj  NAME rotate_x|rotate_y|rotate_z CHILD ANGLE     (x,z degrees; y radians — as the reference)
//   obj  NAME matinst CHILD MAT                   (MAT may be `null`)
//   obj  NAME model PATH MAT SCALE
//   obj  NAME medium CHILD DENSITY (color R G B | tex TEX)
//   obj  NAME list CHILD...                       obj NAME bvh CHILD...
//   world NAME
//   envmap PATH                                   (lat-long HDRI for EnvironmentSettings::load_hdr)
// Objects are built on demand from `world` in DFS order; non-container objects are built once and
// shared (DAG).  ZSCENE_TAG(object, leaf_id) is applied at every site where a non-container object
// is added to a container — the oracle harness uses it to wrap the object in an id recorder; the
// product leaves it as the identity, and the flattener hands out the same ids in the same order.
#pragma once
#include <cstdlib>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#ifndef ZSCENE_TAG
#define ZSCENE_TAG(obj, id) ((void)(id), (obj))
#endif

namespace zscene {

struct record {
    std::string type;
    std::vector<std::string> args;
};

class loader {
public:
    std::map<std::string, std::shared_ptr<texture>> textures;
    std::map<std::string, std::shared_ptr<material>> materials;
    std::string base_dir;  // prefix for relative asset paths
    std::string env_hdr_path;  // `envmap PATH`: the HDRI handed to EnvironmentSettings::load_hdr by the caller
    int next_leaf_id = 0;

    void parse(std::istream& in) {
        std::string line;
        int lineno = 0;
        while (std::getline(in, line)) {
            lineno++;
            size_t hash = line.find('#');
            if (hash != std::string::npos) line.resize(hash);
            std::istringstream ss(line);
            std::vector<std::string> tok;
            std::string t;
            while (ss >> t) tok.push_back(t);
            if (tok.empty()) continue;
            if (tok[0] == "world" && tok.size() == 2) {
                world_name = tok[1];
            } else if (tok[0] == "envmap" && tok.size() == 2) {
                env_hdr_path = tok[1];
            } else if ((tok[0] == "tex" || tok[0] == "mat" || tok[0] == "obj") && tok.size() >= 3) {
                record r;
                r.type = tok[2];
                r.args.assign(tok.begin() + 3, tok.end());
                if (tok[0] == "tex") make_texture(tok[1], r);
                else if (tok[0] == "mat") make_material(tok[1], r);
                else objects[tok[1]] = r;
            } else {
                throw std::runtime_error("zscene: bad record at line " + std::to_string(lineno));
            }
        }
    }

    std::string path_of(const std::string& p) const {
        if (p.empty() || p[0] == '/' || base_dir.empty()) return p;
        return base_dir + "/" + p;
    }

    std::shared_ptr<hittable> build_world() {
        if (world_name.empty()) throw std::runtime_error("zscene: no world record");
        return build(world_name);
    }

private:
    std::map<std::string, record> objects;
    std::map<std::string, std::shared_ptr<hittable>> built;
    std::string world_name;

    static double num(const std::string& s) {
        char* e = nullptr;
        double v = std::strtod(s.c_str(), &e);
        if (e == s.c_str()) throw std::runtime_error("zscene: expected a number, got '" + s + "'");
        return v;
    }
    static const std::string& arg(const record& r, size_t i) {
        if (i >= r.args.size()) throw std::runtime_error("zscene: too few arguments for " + r.type);
        return r.args[i];
    }
    static vec3 v3(const record& r, size_t i) { return vec3(num(arg(r, i)), num(arg(r, i + 1)), num(arg(r, i + 2))); }
    std::shared_ptr<texture> tex(const std::string& n) const {
        auto it = textures.find(n);
        if (it == textures.end()) throw std::runtime_error("zscene: unknown texture " + n);
        return it->second;
    }
    std::shared_ptr<material> mat(const std::string& n) const {
        if (n == "null") return nullptr;
        auto it = materials.find(n);
        if (it == materials.end()) throw std::runtime_error("zscene: unknown material " + n);
        return it->second;
    }

    void make_texture(const std::string& name, const record& r) {
        if (r.type == "solid") textures[name] = std::make_shared<solid_color>(v3(r, 0));
        else if (r.type == "checker") textures[name] = std::make_shared<checker_texture>(num(arg(r, 0)), tex(arg(r, 1)), tex(arg(r, 2)));
        else if (r.type == "image") textures[name] = std::make_shared<image_texture>(path_of(arg(r, 0)).c_str(), num(arg(r, 1)) != 0);
        else throw std::runtime_error("zscene: unknown texture type " + r.type);
    }

    // (color R G B | tex TEX) at position i; returns the number of tokens consumed
    size_t albedo(const record& r, size_t i, std::shared_ptr<texture>& t, vec3& c, bool& is_tex) const {
        if (arg(r, i) == "color") { c = v3(r, i + 1); is_tex = false; return 4; }
        if (arg(r, i) == "tex") { t = tex(arg(r, i + 1)); is_tex = true; return 2; }
        throw std::runtime_error("zscene: expected color|tex");
    }
    bool bump(const record& r, size_t i, std::shared_ptr<texture>& b, double& strength) const {
        if (i < r.args.size() && r.args[i] == "bump") { b = tex(arg(r, i + 1)); strength = num(arg(r, i + 2)); return true; }
        return false;
    }

    void make_material(const std::string& name, const record& r) {
        std::shared_ptr<texture> t, b;
        vec3 c(0, 0, 0);
        bool is_tex = false;
        double strength = 1.0;
        if (r.type == "lambertian") {
            size_t i = albedo(r, 0, t, c, is_tex);
            bump(r, i, b, strength);
            materials[name] = is_tex ? std::make_shared<lambertian>(t, b, strength) : std::make_shared<lambertian>(c, b, strength);
        } else if (r.type == "metal") {
            size_t i = albedo(r, 0, t, c, is_tex);
            double fuzz = num(arg(r, i));
            bump(r, i + 1, b, strength);
            materials[name] = is_tex ? std::make_shared<metal>(t, fuzz, b, strength) : std::make_shared<metal>(c, fuzz, b, strength);
        } else if (r.type == "dielectric") {
            double ior = num(arg(r, 0));
            size_t i = 1;
            vec3 tint(1, 1, 1);
            if (i < r.args.size() && r.args[i] == "tint") { tint = v3(r, i + 1); i += 4; }
            if (bump(r, i, b, strength)) materials[name] = std::make_shared<dielectric>(ior, tint, b, strength);
            else materials[name] = std::make_shared<dielectric>(ior, tint);
        } else if (r.type == "light") {
            albedo(r, 0, t, c, is_tex);
            materials[name] = is_tex ? std::make_shared<diffuse_light>(t) : std::make_shared<diffuse_light>(c);
        } else {
            throw std::runtime_error("zscene: unknown material type " + r.type);
        }
    }

    bool is_container(const std::string& name) const {
        auto it = objects.find(name);
        return it != objects.end() && (it->second.type == "list" || it->second.type == "bvh");
    }

    std::shared_ptr<hittable> build(const std::string& name) {
        auto done = built.find(name);
        if (done != built.end()) return done->second;
        auto it = objects.find(name);
        if (it == objects.end()) throw std::runtime_error("zscene: unknown object " + name);
        const record& r = it->second;
        std::shared_ptr<hittable> h;
        if (r.type == "list" || r.type == "bvh") {
            hittable_list l;
            for (const auto& c : r.args) {
                std::shared_ptr<hittable> child = build(c);
                if (!is_container(c)) {
                    int id = next_leaf_id++;
                    child = ZSCENE_TAG(child, id);
                }
                l.add(child);
            }
            if (r.type == "list") return std::make_shared<hittable_list>(l);  // containers are not shared
            return std::make_shared<bvh_node>(l);
        } else if (r.type == "sphere") {
            h = std::make_shared<sphere>(v3(r, 0), num(arg(r, 3)), mat(arg(r, 4)));
        } else if (r.type == "cube") {
            h = std::make_shared<cube>(v3(r, 0), v3(r, 3), mat(arg(r, 6)));
        } else if (r.type == "cube1") {
            h = std::make_shared<cube>(v3(r, 0), mat(arg(r, 3)));
        } else if (r.type == "triangle") {
            h = std::make_shared<triangle>(v3(r, 0), v3(r, 3), v3(r, 6), v3(r, 9), v3(r, 12), v3(r, 15), mat(arg(r, 18)));
        } else if (r.type == "translate") {
            h = std::make_shared<translate>(build(arg(r, 0)), v3(r, 1));
        } else if (r.type == "scale") {
            h = std::make_shared<scale>(build(arg(r, 0)), v3(r, 1));
        } else if (r.type == "rotate_x") {
            h = std::make_shared<rotate_x>(build(arg(r, 0)), num(arg(r, 1)));
        } else if (r.type == "rotate_y") {
            h = std::make_shared<rotate_y>(build(arg(r, 0)), num(arg(r, 1)));
        } else if (r.type == "rotate_z") {
            h = std::make_shared<rotate_z>(build(arg(r, 0)), num(arg(r, 1)));
        } else if (r.type == "matinst") {
            h = std::make_shared<material_instance>(build(arg(r, 0)), mat(arg(r, 1)));
        } else if (r.type == "model") {
            h = std::make_shared<model>(path_of(arg(r, 0)), mat(arg(r, 1)), num(arg(r, 2)));
        } else if (r.type == "medium") {
            std::shared_ptr<texture> t;
            vec3 c(0, 0, 0);
            bool is_tex = false;
            albedo(r, 2, t, c, is_tex);
            if (is_tex) h = std::make_shared<constant_medium>(build(arg(r, 0)), num(arg(r, 1)), t);
            else h = std::make_shared<constant_medium>(build(arg(r, 0)), num(arg(r, 1)), c);
        } else {
            throw std::runtime_error("zscene: unknown object type " + r.type);
        }
        built[name] = h;
        return h;
    }
};

}  // namespace zscene
