// zenith/model.hpp — OBJ mesh object of the host scene API: model(path, material, scale) model.hpp:12.
// Own OBJ reader (v / vn / f, negative indices, quads split on the shorter diagonal like
// TinyObjLoader v2.0.0 tiny_obj_loader.h:1521-1612, larger polygons as fans), positions parsed to
// float as TinyObjLoader does (real_t = float, tiny_obj_loader.h:150), then recentred in XZ, dropped
// to y = 0 and scaled in double exactly as model.hpp:24-53.  The triangles are kept as plain arrays
// (face order) for the flattener, which builds the mesh BLAS.
#pragma once
#include <cstdlib>
#include <fstream>
#include <sstream>

#include "scene.hpp"

namespace zenith {

class model : public hittable {
public:
    struct tri {
        point3 v[3];
        vec3 n[3];
    };

    model(const std::string& filename, shared_ptr<material> mat, double scale = 1.0) : mat(mat) {
        std::vector<float> pos, nrm;
        struct corner { int v, n; };
        std::vector<corner> corners;  // 3 per triangle, face order
        std::ifstream in(filename);
        if (!in) {
            std::cerr << "Cannot load the model: " << filename << std::endl;
            return;  // the reference would later dereference a null BVH (model.hpp:18-21,109-111)
        }
        std::string line;
        std::vector<corner> poly;
        while (std::getline(in, line)) {
            const char* s = line.c_str();
            while (*s == ' ' || *s == '\t') s++;
            if (s[0] == 'v' && (s[1] == ' ' || s[1] == '\t')) {
                char* e = const_cast<char*>(s + 1);
                for (int k = 0; k < 3; k++) pos.push_back(float(std::strtod(e, &e)));
            } else if (s[0] == 'v' && s[1] == 'n' && (s[2] == ' ' || s[2] == '\t')) {
                char* e = const_cast<char*>(s + 2);
                for (int k = 0; k < 3; k++) nrm.push_back(float(std::strtod(e, &e)));
            } else if (s[0] == 'f' && (s[1] == ' ' || s[1] == '\t')) {
                poly.clear();
                const char* p = s + 1;
                while (*p) {
                    while (*p == ' ' || *p == '\t' || *p == '\r') p++;
                    if (!*p) break;
                    char* e;
                    long vi = std::strtol(p, &e, 10), ni = 0;
                    bool has_n = false;
                    p = e;
                    if (*p == '/') {
                        p++;
                        if (*p != '/') { std::strtol(p, &e, 10); p = e; }  // vt ignored
                        if (*p == '/') {
                            p++;
                            ni = std::strtol(p, &e, 10);
                            has_n = e != p;
                            p = e;
                        }
                    }
                    int nv = int(pos.size() / 3), nn = int(nrm.size() / 3);
                    corner c;
                    c.v = vi > 0 ? int(vi) - 1 : nv + int(vi);
                    c.n = has_n ? (ni > 0 ? int(ni) - 1 : nn + int(ni)) : -1;
                    poly.push_back(c);
                }
                if (poly.size() == 3) {
                    corners.insert(corners.end(), poly.begin(), poly.end());
                } else if (poly.size() == 4) {
                    auto P = [&](int c, int k) { return pos[3 * poly[c].v + k]; };
                    float s02 = 0, s13 = 0;
                    for (int k = 0; k < 3; k++) {
                        float a = P(2, k) - P(0, k), b = P(3, k) - P(1, k);
                        s02 += a * a;
                        s13 += b * b;
                    }
                    const int split_a[6] = {0, 1, 2, 0, 2, 3}, split_b[6] = {0, 1, 3, 1, 2, 3};
                    const int* sp = s02 < s13 ? split_a : split_b;
                    for (int k = 0; k < 6; k++) corners.push_back(poly[sp[k]]);
                } else if (poly.size() > 4) {
                    for (size_t k = 1; k + 1 < poly.size(); k++) {
                        corners.push_back(poly[0]);
                        corners.push_back(poly[k]);
                        corners.push_back(poly[k + 1]);
                    }
                }
            }
        }
        // model.hpp:24-42: bounds over ALL parsed vertices, offset = (midX, minY, midZ)
        double lo[3] = {infinity, infinity, infinity}, hi[3] = {-infinity, -infinity, -infinity};
        for (size_t i = 0; i + 2 < pos.size(); i += 3)
            for (int k = 0; k < 3; k++) {
                lo[k] = std::fmin(lo[k], pos[i + k]);
                hi[k] = std::fmax(hi[k], pos[i + k]);
            }
        vec3 off((lo[0] + hi[0]) / 2.0, lo[1], (lo[2] + hi[2]) / 2.0);
        auto get_v = [&](int vi) {
            return point3((pos[3 * vi + 0] - off.x()) * scale, (pos[3 * vi + 1] - off.y()) * scale,
                          (pos[3 * vi + 2] - off.z()) * scale);
        };
        tris.reserve(corners.size() / 3);
        for (size_t i = 0; i + 2 < corners.size(); i += 3) {
            tri t;
            for (int k = 0; k < 3; k++) t.v[k] = get_v(corners[i + k].v);
            if (corners[i].n >= 0) {  // model.hpp:71-80 decides on the first corner's normal index
                for (int k = 0; k < 3; k++) {
                    int ni = corners[i + k].n;
                    t.n[k] = ni >= 0 ? vec3(nrm[3 * ni], nrm[3 * ni + 1], nrm[3 * ni + 2]) : vec3(0, 0, 0);
                }
            } else {
                vec3 flat = unit_vector(cross(t.v[1] - t.v[0], t.v[2] - t.v[0]));
                t.n[0] = t.n[1] = t.n[2] = flat;
            }
            tris.push_back(t);
            bbox = aabb(bbox, triangle::box_of(t.v[0], t.v[1], t.v[2]));
        }
        std::cout << "Model: " << filename << " loaded (" << tris.size() << " triangles)." << std::endl;
    }
    kind_t kind() const override { return kind_t::model; }
    aabb bounding_box() const override { return bbox; }
    void set_material(shared_ptr<material> m) { mat = m; }
    std::vector<tri> tris;
    shared_ptr<material> mat;

private:
    aabb bbox;
};

}  // namespace zenith
