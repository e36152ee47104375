// TEST INFRASTRUCTURE — CPU restatement ("port") of the reference's path-tracing hot path, in double.
//
// Restates, function by function, what jarek1992/raytracer_project computes on the path behind
// camera::render (file:line citations are relative to /root/reference).  It evaluates the SAME scene
// graph the product's flattener consumes (raytracer_project_b200/host/zenith/*.hpp, a pure data model
// without intersection code) with the reference's recursive closest-hit semantics, so it checks the
// flattener and the CUDA kernels together.  Parity of this restatement is PINNED on outputs of the
// reference itself: tests/test_oracle_cpu.py compares it with the committed fixtures under tests/golden/
// (per-ray hit records, primary-hit ids, converged radiance) that oracle/_ref/ref_harness — the
// unmodified reference — produced.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build or run
// this file.  The product never links or calls it; there is no CPU fallback in the product.
//
//   zoracle ids    SCENE CFG OUT          centre-of-pixel primary hits (camera.hpp:809-814 + :459)
//   zoracle probe  SCENE CFG RAYS OUT     per-ray hit_record / albedo / emitted / background
//   zoracle render SCENE CFG OUT|- [SEED] render_rows (camera.hpp:430-554), OpenMP over rows
//   zoracle finish IMAGE CFG OUT          finishing chain (camera.hpp:142-207, color_processing.hpp, bloom.hpp)
//   zoracle export IMAGE CFG OUT          image export up to the 8-bit rows (camera.hpp:299-343, :701-780)
// File formats are those of oracle/ref_harness.cpp, so the same test code reads both.
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <random>
#include <sstream>
#include <unordered_map>

#ifdef _OPENMP
#include <omp.h>
#endif

#include "../raytracer_project_b200/host/zenith/environment.hpp"
#include "../raytracer_project_b200/host/zenith/model.hpp"
#include "../raytracer_project_b200/host/zenith/scene.hpp"

using namespace zenith;

// Leaf-id recorder: the zscene loader wraps every non-container object at the site where it is added to a
// container (same numbering as the product's flattener and as ref_harness's id_tag).
struct otag : material_instance {
    int id;
    otag(shared_ptr<hittable> o, int id) : material_instance(o, nullptr), id(id) {}
};
#define ZSCENE_TAG(obj, id) std::static_pointer_cast<hittable>(std::make_shared<otag>((obj), (id)))
#include "../raytracer_project_b200/host/zscene.hpp"

namespace {

constexpr double ray_epsilon = 0.0001;  // common.hpp:16

struct ray {
    point3 o;
    vec3 d;
    point3 at(double t) const { return o + t * d; }
};

// hit_record (hittable.hpp:9-27) + the ids the parity tests compare
struct hit_rec {
    point3 p;
    vec3 normal, tangent, bitangent;
    const material* mat = nullptr;
    bool front_face = false;
    double t = 0, u = 0, v = 0;
    int leaf = -1, tri = -1;
    // hittable.hpp:22-25
    void set_face_normal(const ray& r, const vec3& outward) {
        front_face = dot(r.d, outward) < 0;
        normal = front_face ? outward : -outward;
    }
};

inline bool surrounds(const interval& i, double x) { return i.min < x && x < i.max; }   // interval.hpp
inline bool contains(const interval& i, double x) { return i.min <= x && x <= i.max; }

// ------------------------------------------------------------------ RNG (common.hpp:29-45, vec3.hpp:174-191)
// The reference shares one unseedable mt19937; the restatement keeps one seeded generator per image row,
// which leaves every distribution unchanged and makes runs reproducible.
thread_local std::mt19937_64 tl_gen(1);
inline double random_double() { return std::uniform_real_distribution<double>(0.0, 1.0)(tl_gen); }
inline double random_double(double a, double b) { return a + (b - a) * random_double(); }
inline vec3 random_in_unit_disk() {
    for (;;) {
        vec3 p(random_double(-1, 1), random_double(-1, 1), 0);
        if (p.length_squared() < 1) return p;
    }
}
inline vec3 random_unit_vector() {
    for (;;) {
        vec3 p(random_double(-1, 1), random_double(-1, 1), random_double(-1, 1));
        double l2 = p.length_squared();
        if (1e-160 < l2 && l2 <= 1) return p / std::sqrt(l2);
    }
}
inline vec3 reflect(const vec3& v, const vec3& n) { return v - 2 * dot(v, n) * n; }  // vec3.hpp:204-206
inline vec3 refract(const vec3& uv, const vec3& n, double eta) {                       // vec3.hpp:209-214
    double cos_theta = std::fmin(dot(-uv, n), 1.0);
    vec3 perp = eta * (uv + cos_theta * n);
    vec3 par = -std::sqrt(std::fabs(1.0 - perp.length_squared())) * n;
    return perp + par;
}

// ------------------------------------------------------------------ textures (texture.hpp:50-126)
color texture_value(const texture* t, double u, double v, const point3& p) {
    switch (t->kind()) {
        case texture::kind_t::solid: return static_cast<const solid_color*>(t)->albedo;  // texture.hpp:96-98
        case texture::kind_t::checker: {                                                 // texture.hpp:118-126
            auto c = static_cast<const checker_texture*>(t);
            int xi = int(std::floor(c->inv_scale * p.x())), yi = int(std::floor(c->inv_scale * p.y())),
                zi = int(std::floor(c->inv_scale * p.z()));
            bool even = (xi + yi + zi) % 2 == 0;  // C++ %: a negative odd sum gives -1, i.e. "odd"
            return texture_value(even ? c->even.get() : c->odd.get(), u, v, p);
        }
        case texture::kind_t::image: {                                                   // texture.hpp:50-78
            auto im = static_cast<const image_texture*>(t);
            if (im->missing()) return color(0.0, 1.0, 1.0);
            u = u - std::floor(u);  // u wraps, v is only clamped
            int i = int(u * im->width), j = int(v * im->height);
            i = std::clamp(i, 0, im->width - 1);
            j = std::clamp(j, 0, im->height - 1);
            size_t at = (size_t(j) * im->width + i) * 3;
            if (im->is_float) return color(im->f32[at], im->f32[at + 1], im->f32[at + 2]);
            const double s = 1.0 / 255.0;
            return color(s * im->u8[at], s * im->u8[at + 1], s * im->u8[at + 2]);
        }
    }
    return color(0, 0, 0);
}

// ------------------------------------------------------------------ materials (material.hpp, constant_medium.hpp:14-18)
const lambertian g_error_mat(color(1, 0, 1));  // material_instance.hpp:22-25

// material::get_bumped_normal (material.hpp:35-54)
vec3 bumped_normal(const material& m, const hit_rec& rec) {
    if (!m.bump) return rec.normal;
    const double du = 1.0 / 1024.0, dv = 1.0 / 1024.0;
    double hc = texture_value(m.bump.get(), rec.u, rec.v, rec.p).x();
    double hu = texture_value(m.bump.get(), rec.u + du, rec.v, rec.p).x();
    double hv = texture_value(m.bump.get(), rec.u, rec.v + dv, rec.p).x();
    double fu = (hu - hc) * m.bump_strength, fv = (hv - hc) * m.bump_strength;
    return unit_vector(rec.normal - (fu * rec.tangent) - (fv * rec.bitangent));
}

color material_emitted(const material& m, const hit_rec& rec) {  // material.hpp:12-14, :261-263
    if (m.kind() != material::kind_t::diffuse_light) return color(0, 0, 0);
    return texture_value(m.tex.get(), rec.u, rec.v, rec.p);
}

color material_albedo(const material& m, const hit_rec& rec) {  // get_albedo overrides
    switch (m.kind()) {
        case material::kind_t::dielectric: return color(1, 1, 1);      // material.hpp:226-228
        case material::kind_t::isovolumetric: return color(0, 0, 0);   // material.hpp:28-30 (default)
        case material::kind_t::diffuse_light: {                        // material.hpp:265-272
            color c = texture_value(m.tex.get(), rec.u, rec.v, rec.p);
            return color(std::fmin(c.x(), 1.0), std::fmin(c.y(), 1.0), std::fmin(c.z(), 1.0));
        }
        default: return texture_value(m.tex.get(), rec.u, rec.v, rec.p);
    }
}

bool material_scatter(const material& m, const ray& r_in, const hit_rec& rec, color& attenuation, ray& scattered) {
    switch (m.kind()) {
        case material::kind_t::lambertian: {  // material.hpp:74-96
            vec3 wn = bumped_normal(m, rec);
            vec3 dir = wn + random_unit_vector();
            if (std::fabs(dir[0]) < 1e-8 && std::fabs(dir[1]) < 1e-8 && std::fabs(dir[2]) < 1e-8) dir = wn;
            scattered = ray{rec.p + rec.normal * ray_epsilon, dir};
            attenuation = texture_value(m.tex.get(), rec.u, rec.v, rec.p);
            return true;
        }
        case material::kind_t::metal: {  // material.hpp:129-151 (RNG drawn even for fuzz 0)
            vec3 wn = bumped_normal(m, rec);
            vec3 refl = reflect(unit_vector(r_in.d), wn);
            vec3 dir = unit_vector(refl + (m.param * random_unit_vector()));
            scattered = ray{rec.p + ray_epsilon * rec.normal, dir};
            attenuation = texture_value(m.tex.get(), rec.u, rec.v, rec.p);
            return dot(dir, rec.normal) > 0;
        }
        case material::kind_t::dielectric: {  // material.hpp:192-224, :237-241
            attenuation = m.tint;
            vec3 wn = bumped_normal(m, rec);
            double ri = rec.front_face ? (1.0 / m.param) : m.param;
            vec3 ud = unit_vector(r_in.d);
            double cos_theta = std::fmin(dot(-ud, wn), 1.0);
            double sin_theta = std::sqrt(1.0 - cos_theta * cos_theta);
            bool cannot_refract = ri * sin_theta > 1.0;
            double r0 = (1 - ri) / (1 + ri);
            r0 = r0 * r0;
            vec3 dir;
            if (cannot_refract || r0 + (1 - r0) * std::pow(1 - cos_theta, 5) > random_double()) dir = reflect(ud, wn);
            else dir = refract(ud, wn, ri);
            vec3 off = dot(dir, rec.normal) > 0 ? ray_epsilon * rec.normal : -ray_epsilon * rec.normal;
            scattered = ray{rec.p + off, dir};
            return true;
        }
        case material::kind_t::isovolumetric:  // constant_medium.hpp:14-18
            scattered = ray{rec.p, random_unit_vector()};
            attenuation = texture_value(m.tex.get(), rec.u, rec.v, rec.p);
            return true;
        case material::kind_t::diffuse_light: return false;  // material.hpp:255-259
    }
    return false;
}

// ------------------------------------------------------------------ aabb::hit (aabb.hpp:44-66)
bool box_hit(const aabb& b, const ray& r, interval t) {
    for (int a = 0; a < 3; a++) {
        const interval& ax = b.axis(a);
        const double inv = 1.0 / r.d[a];
        double t0 = (ax.min - r.o[a]) * inv, t1 = (ax.max - r.o[a]) * inv;
        if (inv < 0) std::swap(t0, t1);
        if (t0 > t.min) t.min = t0;
        if (t1 < t.max) t.max = t1;
        if (t.max <= t.min) return false;
    }
    return true;
}

// ------------------------------------------------------------------ accelerator for bvh_node / model
// The reference's tree (bvh.hpp:15-44) splits on a RANDOM axis at the median; closest-hit answers do not
// depend on the tree (exact ties aside), so this restatement splits at the median of the longest axis.
// Traversal is the reference's: box test, left, shrink t_max, right (bvh.hpp:112-118).
struct accel {
    struct node {
        aabb box;
        int left = -1, right = -1;  // children, or (first, count) of a leaf when left < 0
        int first = 0, count = 0;
    };
    std::vector<node> nodes;
    std::vector<int> order;
    template <class BoxOf>
    void build(int n, BoxOf&& box_of) {
        order.resize(n);
        for (int i = 0; i < n; i++) order[i] = i;
        std::vector<aabb> boxes(n);
        for (int i = 0; i < n; i++) boxes[i] = box_of(i);
        nodes.clear();
        if (n > 0) split(0, n, boxes);
    }
    int split(int lo, int hi, const std::vector<aabb>& boxes) {
        int me = int(nodes.size());
        nodes.emplace_back();
        aabb b;
        for (int i = lo; i < hi; i++) b = aabb(b, boxes[order[i]]);
        nodes[me].box = b;
        if (hi - lo <= 2) {
            nodes[me].first = lo;
            nodes[me].count = hi - lo;
            return me;
        }
        int axis = 0;
        if (b.y.size() > b.axis(axis).size()) axis = 1;
        if (b.z.size() > b.axis(axis).size()) axis = 2;
        int mid = (lo + hi) / 2;
        std::nth_element(order.begin() + lo, order.begin() + mid, order.begin() + hi, [&](int a, int c) {
            return boxes[a].axis(axis).min + boxes[a].axis(axis).max < boxes[c].axis(axis).min + boxes[c].axis(axis).max;
        });
        int l = split(lo, mid, boxes), r = split(mid, hi, boxes);
        nodes[me].left = l;
        nodes[me].right = r;
        return me;
    }
    template <class Leaf>
    bool hit(int at, const ray& r, interval t, hit_rec& rec, Leaf&& leaf) const {
        const node& n = nodes[at];
        if (!box_hit(n.box, r, t)) return false;
        if (n.left < 0) {
            bool any = false;
            for (int i = n.first; i < n.first + n.count; i++)
                if (leaf(order[i], r, t, rec)) { any = true; t.max = rec.t; }
            return any;
        }
        bool hl = hit(n.left, r, t, rec, leaf);
        bool hr = hit(n.right, r, interval(t.min, hl ? rec.t : t.max), rec, leaf);
        return hl || hr;
    }
};
std::unordered_map<const hittable*, accel> g_accels;  // built before rendering starts (prepare())

bool hit_any(const hittable& h, const ray& r, interval t, hit_rec& rec);

// sphere::hit (sphere.hpp:18-64) + get_sphere_uv (:70-79)
bool hit_sphere(const sphere& s, const ray& r, interval t, hit_rec& rec) {
    vec3 oc = s.center - r.o;
    double a = r.d.length_squared(), h = dot(r.d, oc), c = oc.length_squared() - s.radius * s.radius;
    double disc = h * h - a * c;
    if (disc < 0) return false;
    double sq = std::sqrt(disc);
    double root = (h - sq) / a;
    if (!surrounds(t, root)) {
        root = (h + sq) / a;
        if (!surrounds(t, root)) return false;
    }
    rec.t = root;
    rec.p = r.at(root);
    vec3 outward = (rec.p - s.center) / s.radius;
    rec.set_face_normal(r, outward);
    rec.u = (std::atan2(-outward.z(), outward.x()) + pi) / (2 * pi);
    rec.v = std::acos(-outward.y()) / pi;
    rec.tangent = cross(vec3(0, 1, 0), rec.normal);  // frame from the FLIPPED normal
    if (rec.tangent.length_squared() < 0.001) rec.tangent = cross(vec3(0, 0, 1), rec.normal);
    rec.tangent = unit_vector(rec.tangent);
    rec.bitangent = cross(rec.normal, rec.tangent);
    rec.mat = s.mat.get();
    rec.tri = -1;
    return true;
}

// triangle::hit (triangle.hpp:17-82): u, v, tangent, bitangent are NOT written (stale values leak through)
bool hit_tri(const point3& v0, const point3& v1, const point3& v2, const vec3& n0, const vec3& n1, const vec3& n2,
             const ray& r, interval t, hit_rec& rec) {
    vec3 normal = cross(v1 - v0, v2 - v0);
    double len = normal.length();
    if (len < 1e-8) return false;
    vec3 un = normal / len;
    double nd = dot(un, r.d);
    if (std::fabs(nd) < 1e-8) return false;
    double tt = (dot(un, v0) - dot(un, r.o)) / nd;
    if (!contains(t, tt)) return false;
    point3 p = r.at(tt);
    vec3 c0 = cross(v1 - v0, p - v0), c1 = cross(v2 - v1, p - v1), c2 = cross(v0 - v2, p - v2);
    if (dot(normal, c0) < 0 || dot(normal, c1) < 0 || dot(normal, c2) < 0) return false;
    double area2 = dot(normal, normal);
    double u = dot(normal, c2) / area2, v = dot(normal, c0) / area2, w = 1.0 - u - v;
    rec.t = tt;
    rec.p = p;
    rec.set_face_normal(r, unit_vector(w * n0 + u * n1 + v * n2));
    return true;
}

// cube::hit (cube.hpp:44-86) + set_cube_hit_data (:100-142): slabs about the ORIGIN of the current frame,
// face lookup relative to the true centre, t = entry distance or ray_t.min when the origin is inside
bool hit_cube(const cube& c, const ray& r, interval t, hit_rec& rec) {
    double tmin = t.min, tmax = t.max;
    for (int i = 0; i < 3; i++) {
        double inv = 1.0 / r.d[i];
        double t0 = (-c.half_extents[i] - r.o[i]) * inv, t1 = (c.half_extents[i] - r.o[i]) * inv;
        if (inv < 0.0) std::swap(t0, t1);
        tmin = std::fmax(t0, tmin);
        tmax = std::fmin(t1, tmax);
        if (tmax < tmin) return false;
    }
    rec.t = tmin;
    rec.p = r.at(tmin);
    const vec3 p = rec.p - c.center, h = c.half_extents;
    const double EPS = 1e-3;
    vec3 n;
    if (std::fabs(p.x() + h.x()) < EPS) { n = vec3(-1, 0, 0); rec.u = (p.z() + h.z()) / (2 * h.z()); rec.v = (p.y() + h.y()) / (2 * h.y()); rec.tangent = vec3(0, 0, 1); }
    else if (std::fabs(p.x() - h.x()) < EPS) { n = vec3(1, 0, 0); rec.u = (p.z() + h.z()) / (2 * h.z()); rec.v = (p.y() + h.y()) / (2 * h.y()); rec.tangent = vec3(0, 0, -1); }
    else if (std::fabs(p.y() + h.y()) < EPS) { n = vec3(0, -1, 0); rec.u = (p.x() + h.x()) / (2 * h.x()); rec.v = (p.z() + h.z()) / (2 * h.z()); rec.tangent = vec3(1, 0, 0); }
    else if (std::fabs(p.y() - h.y()) < EPS) { n = vec3(0, 1, 0); rec.u = (p.x() + h.x()) / (2 * h.x()); rec.v = (p.z() + h.z()) / (2 * h.z()); rec.tangent = vec3(-1, 0, 0); }
    else if (std::fabs(p.z() + h.z()) < EPS) { n = vec3(0, 0, -1); rec.u = (h.x() - p.x()) / (2 * h.x()); rec.v = (p.y() + h.y()) / (2 * h.y()); rec.tangent = vec3(-1, 0, 0); }
    else { n = vec3(0, 0, 1); rec.u = (p.x() + h.x()) / (2 * h.x()); rec.v = (p.y() + h.y()) / (2 * h.y()); rec.tangent = vec3(1, 0, 0); }
    rec.bitangent = cross(n, rec.tangent);
    rec.mat = c.mat.get();
    rec.set_face_normal(r, n);
    rec.tri = -1;
    return true;
}

// constant_medium::hit (constant_medium.hpp:39-77): one free-flight draw per call
bool hit_medium(const constant_medium& m, const ray& r, interval t, hit_rec& rec) {
    hit_rec r1, r2;
    if (!hit_any(*m.boundary, r, interval(-infinity, infinity), r1)) return false;
    if (!hit_any(*m.boundary, r, interval(r1.t + 0.0001, infinity), r2)) return false;
    if (r1.t < t.min) r1.t = t.min;
    if (r2.t > t.max) r2.t = t.max;
    if (r1.t >= r2.t) return false;
    if (r1.t < 0) r1.t = 0;
    double len = r.d.length();
    double inside = (r2.t - r1.t) * len;
    double dist = m.neg_inv_density * std::log(random_double());
    if (dist > inside) return false;
    rec.t = r1.t + dist / len;
    rec.p = r.at(rec.t);
    rec.normal = vec3(1, 0, 0);
    rec.front_face = true;
    rec.mat = m.phase_function.get();
    rec.tri = -1;
    return true;
}

// rotation of components (a, b) by (c, s):  a' = c a + s b,  b' = -s a + c b   (world -> object)
inline void rot_fwd(double c, double s, double& a, double& b) { double x = c * a + s * b, y = -s * a + c * b; a = x; b = y; }
// and back:                                 a' = c a - s b,  b' =  s a + c b   (object -> world)
inline void rot_back(double c, double s, double& a, double& b) { double x = c * a - s * b, y = s * a + c * b; a = x; b = y; }

bool hit_any(const hittable& h, const ray& r, interval t, hit_rec& rec) {
    using K = hittable::kind_t;
    switch (h.kind()) {
        case K::list: {  // hittable_list::hit (hittable_list.hpp:28-41)
            bool any = false;
            double closest = t.max;
            hit_rec tmp;  // one temp record for the whole loop, as the reference: stale u/v/tangent leak into triangle hits
            for (auto& o : static_cast<const hittable_list&>(h).objects) {
                if (hit_any(*o, r, interval(t.min, closest), tmp)) { any = true; closest = tmp.t; rec = tmp; }
            }
            return any;
        }
        case K::bvh: {  // bvh_node::hit normal path (bvh.hpp:46-54, :112-118)
            auto& b = static_cast<const bvh_node&>(h);
            const accel& a = g_accels.at(&h);
            if (a.nodes.empty()) return false;
            return a.hit(0, r, t, rec, [&](int i, const ray& rr, interval tt, hit_rec& rc) { return hit_any(*b.objects[i], rr, tt, rc); });
        }
        case K::sphere: return hit_sphere(static_cast<const sphere&>(h), r, t, rec);
        case K::cube: return hit_cube(static_cast<const cube&>(h), r, t, rec);
        case K::triangle: {
            auto& tr = static_cast<const triangle&>(h);
            if (!hit_tri(tr.v0, tr.v1, tr.v2, tr.n0, tr.n1, tr.n2, r, t, rec)) return false;
            rec.mat = tr.mat.get();
            rec.tri = -1;
            return true;
        }
        case K::model: {  // model::hit forwards to its own bvh_node of triangles (model.hpp:95, :109-111)
            auto& m = static_cast<const model&>(h);
            const accel& a = g_accels.at(&h);
            if (a.nodes.empty()) return false;
            return a.hit(0, r, t, rec, [&](int i, const ray& rr, interval tt, hit_rec& rc) {
                const model::tri& q = m.tris[i];
                if (!hit_tri(q.v[0], q.v[1], q.v[2], q.n[0], q.n[1], q.n[2], rr, tt, rc)) return false;
                rc.mat = m.mat.get();
                rc.tri = i;
                return true;
            });
        }
        case K::translate: {  // translate.hpp:15-32: set_face_normal on the already flipped normal => front_face forced
            auto& w = static_cast<const translate&>(h);
            ray lr{r.o - w.offset, r.d};
            if (!hit_any(*w.ptr, lr, t, rec)) return false;
            rec.p += w.offset;
            rec.set_face_normal(r, rec.normal);
            return true;
        }
        case K::rotate_y: {  // rotate_y.hpp:41-73 (radians), set_face_normal re-run as in translate
            auto& w = static_cast<const rotate_y&>(h);
            ray lr = r;
            rot_fwd(w.c, w.s, lr.o[0], lr.o[2]);
            rot_fwd(w.c, w.s, lr.d[0], lr.d[2]);
            if (!hit_any(*w.ptr, lr, t, rec)) return false;
            vec3 n = rec.normal;
            rot_back(w.c, w.s, rec.p[0], rec.p[2]);
            rot_back(w.c, w.s, n[0], n[2]);
            rec.set_face_normal(r, n);
            return true;
        }
        case K::rotate_x: {  // rotate_x.hpp:42-70 (degrees): normal rotated, front_face kept
            auto& w = static_cast<const rotate_x&>(h);
            ray lr = r;
            rot_fwd(w.c, w.s, lr.o[1], lr.o[2]);
            rot_fwd(w.c, w.s, lr.d[1], lr.d[2]);
            if (!hit_any(*w.ptr, lr, t, rec)) return false;
            rot_back(w.c, w.s, rec.p[1], rec.p[2]);
            rot_back(w.c, w.s, rec.normal[1], rec.normal[2]);
            return true;
        }
        case K::rotate_z: {  // rotate_z.hpp:41-67 (degrees): normal rotated, front_face kept
            auto& w = static_cast<const rotate_z&>(h);
            ray lr = r;
            rot_fwd(w.c, w.s, lr.o[0], lr.o[1]);
            rot_fwd(w.c, w.s, lr.d[0], lr.d[1]);
            if (!hit_any(*w.ptr, lr, t, rec)) return false;
            rot_back(w.c, w.s, rec.p[0], rec.p[1]);
            rot_back(w.c, w.s, rec.normal[0], rec.normal[1]);
            return true;
        }
        case K::scale: {  // scale.hpp:20-36
            auto& w = static_cast<const scale&>(h);
            ray lr{point3(r.o[0] / w.s[0], r.o[1] / w.s[1], r.o[2] / w.s[2]), vec3(r.d[0] / w.s[0], r.d[1] / w.s[1], r.d[2] / w.s[2])};
            if (!hit_any(*w.ptr, lr, t, rec)) return false;
            rec.p = point3(rec.p[0] * w.s[0], rec.p[1] * w.s[1], rec.p[2] * w.s[2]);
            rec.normal = unit_vector(vec3(rec.normal[0] / w.s[0], rec.normal[1] / w.s[1], rec.normal[2] / w.s[2]));
            return true;
        }
        case K::material_instance: {  // material_instance.hpp:12-28, or the harness's leaf-id recorder
            auto& mi = static_cast<const material_instance&>(h);
            if (!hit_any(*mi.object, r, t, rec)) return false;
            if (auto tag = dynamic_cast<const otag*>(&h)) rec.leaf = tag->id;
            else rec.mat = mi.new_material ? mi.new_material.get() : &g_error_mat;
            return true;
        }
        case K::constant_medium: return hit_medium(static_cast<const constant_medium&>(h), r, t, rec);
    }
    return false;
}

// build the accelerators of every bvh_node / model reachable from h (single-threaded, before rendering)
void prepare(const hittable& h) {
    using K = hittable::kind_t;
    switch (h.kind()) {
        case K::list: for (auto& o : static_cast<const hittable_list&>(h).objects) prepare(*o); return;
        case K::bvh: {
            auto& b = static_cast<const bvh_node&>(h);
            if (g_accels.count(&h)) return;
            g_accels[&h].build(int(b.objects.size()), [&](int i) { return b.objects[i]->bounding_box(); });
            for (auto& o : b.objects) prepare(*o);
            return;
        }
        case K::model: {
            auto& m = static_cast<const model&>(h);
            if (g_accels.count(&h)) return;
            g_accels[&h].build(int(m.tris.size()), [&](int i) { return triangle::box_of(m.tris[i].v[0], m.tris[i].v[1], m.tris[i].v[2]); });
            return;
        }
        case K::translate: case K::rotate_x: case K::rotate_y: case K::rotate_z: case K::scale:
            prepare(*static_cast<const transform_wrapper&>(h).ptr); return;
        case K::material_instance: prepare(*static_cast<const material_instance&>(h).object); return;
        case K::constant_medium: prepare(*static_cast<const constant_medium&>(h).boundary); return;
        default: return;
    }
}

// ------------------------------------------------------------------ camera (camera.hpp:358-402, :784-825)
struct config {
    std::map<std::string, std::vector<std::string>> kv;
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

struct cam_t {
    int W = 400, H = 225, spp = 10, max_depth = 50;
    double vfov = 30, defocus_angle = 0, focus_dist = 10, z_max = 20;
    point3 lookfrom = point3(10, 1.5, 0), lookat = point3(0, 0, 0);
    vec3 vup = vec3(0, 1, 0);
    bool use_albedo = false, use_normal = false, use_z = false, use_reflection = false, use_refraction = false;
    // derived (camera.hpp:346-353)
    point3 center, pixel00;
    vec3 du, dv, u, v, w, disk_u, disk_v;
    void load(const config& c) {
        W = std::max(1, int(c.num("image_width", 400)));
        H = std::max(1, int(c.num("image_height", 225)));
        spp = int(c.num("samples_per_pixel", 10));
        max_depth = int(c.num("max_depth", 50));
        vfov = c.num("vfov", 30);
        lookfrom = c.v3("lookfrom", lookfrom);
        lookat = c.v3("lookat", lookat);
        vup = c.v3("vup", vup);
        defocus_angle = c.num("defocus_angle", 0.0);
        focus_dist = c.num("focus_dist", 10.0);
        use_albedo = c.num("use_albedo", 0) != 0;
        use_normal = c.num("use_normal", 0) != 0;
        use_z = c.num("use_z_depth", 0) != 0;
        z_max = c.num("z_depth_max_dist", 20.0);
        use_reflection = c.num("use_reflection", 0) != 0;
        use_refraction = c.num("use_refraction", 0) != 0;
    }
    void initialize() {  // camera.hpp:358-402
        double aspect = double(W) / H;
        center = lookfrom;
        double h = std::tan(degrees_to_radians(vfov) / 2);
        double vh = 2 * h * focus_dist, vw = vh * aspect;
        w = unit_vector(lookfrom - lookat);
        u = unit_vector(cross(vup, w));
        v = cross(w, u);
        vec3 vu = vw * u, vv = vh * -v;
        du = vu / W;
        dv = vv / H;
        point3 ul = center - (focus_dist * w) - vu / 2 - vv / 2;
        pixel00 = ul + 0.5 * (du + dv);
        double rad = focus_dist * std::tan(degrees_to_radians(defocus_angle / 2));
        disk_u = u * rad;
        disk_v = v * rad;
    }
    ray get_center_ray(int i, int j) const {  // camera.hpp:809-814
        point3 pc = pixel00 + (i * du) + (j * dv);
        return ray{center, pc - center};
    }
    ray get_ray(int i, int j) const {  // camera.hpp:784-794, :817-825
        double ox = random_double() - 0.5, oy = random_double() - 0.5;
        point3 ps = pixel00 + ((i + ox) * du) + ((j + oy) * dv);
        point3 o = center;
        if (defocus_angle > 0) {
            vec3 p = random_in_unit_disk();
            o = center + (p[0] * disk_u) + (p[1] * disk_v);
        }
        return ray{o, ps - o};
    }
};

// ------------------------------------------------------------------ environment (camera.hpp:828-925)
void load_env(const config& c, const std::string& base_dir, EnvironmentSettings& env) {
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

inline double clamp01(double x) { return std::clamp(x, 0.0, 1.0); }

color background(const ray& r, const EnvironmentSettings& env) {
    vec3 ud = unit_vector(r.d);
    if (env.mode() == EnvironmentSettings::SOLID_COLOR) return env.background_color * env.intensity;  // :832-834
    if (env.mode() == EnvironmentSettings::HDR_MAP) {                                                   // :837-870
        if (!env.hdr_texture) return color(0, 0, 0);
        vec3 d = ud;
        double cy = std::cos(env.hdri_rotation), sy = std::sin(env.hdri_rotation);
        d = vec3(cy * d.x() + sy * d.z(), d.y(), -sy * d.x() + cy * d.z());
        double cp = std::cos(env.hdri_tilt), sp = std::sin(env.hdri_tilt);
        d = vec3(d.x(), cp * d.y() - sp * d.z(), sp * d.y() + cp * d.z());
        double cr = std::cos(env.hdri_roll), sr = std::sin(env.hdri_roll);
        d = vec3(cr * d.x() - sr * d.y(), sr * d.x() + cr * d.y(), d.z());
        double phi = std::atan2(d.z(), d.x()) + pi, theta = std::acos(std::clamp(d.y(), -1.0, 1.0));
        return texture_value(env.hdr_texture.get(), phi / (2 * pi), theta / pi, point3(0, 0, 0)) * env.intensity;
    }
    // PHYSICAL_SUN (:871-924)
    vec3 sun = unit_vector(env.sun_direction);
    double sun_h = sun.y(), adj = sun_h - 0.05;
    double sky_exposure = clamp01(adj * 8.0 + 1.4), day = clamp01(adj * 10.0 + 1.1);
    double sunset_i = clamp01(1.0 - std::fabs(adj + 0.05) * 30.0);
    double sunset = adj > -0.1 ? sunset_i : 0.0;
    if (sun_h < 0) sunset *= (sun_h * 10.0 + 1.0);
    sunset = clamp01(sunset);
    color zenit = color(0.01, 0.03, 0.1) * (1.0 - day) + color(0.2, 0.5, 1.0) * day;
    color horizon = color(0.05, 0.02, 0.01) * (1.0 - day) + color(0.6, 0.8, 1.0) * day;
    horizon = horizon * (1.0 - sunset) + color(1.0, 0.35, 0.1) * sunset;
    double a = ud.y();
    color sky = a > 0.0 ? (1.0 - a) * horizon + a * zenit : horizon * 0.1;
    color out = sky * (env.intensity * 1.5) * sky_exposure;
    double focus = dot(ud, sun), thr = 1.0 - (env.sun_size * 0.001);
    if (focus > thr && adj > -0.1) {
        color sc = env.sun_color * (1.0 - sunset) + color(1.0, 0.3, 0.1) * sunset;
        double vis = clamp01(sun_h * 5.0 + 1.0);
        double x = clamp01((focus - thr) / ((thr + 0.0002) - thr));  // smoothstep (common.hpp:87-91)
        out += sc * env.sun_intensity * vis * (x * x * (3 - 2 * x));
    }
    return out;
}

// ------------------------------------------------------------------ ray_color (camera.hpp:928-1004)
const interval k_ray_t(0.001, infinity);  // camera.hpp:459, :937

color ray_color(ray cur, const hittable& world, int depth, const EnvironmentSettings& env, unsigned long long& rays) {
    color light(0, 0, 0), atten(1, 1, 1);
    for (int i = 0; i < depth; i++) {
        hit_rec rec;
        rays++;
        if (!hit_any(world, cur, k_ray_t, rec)) return light + atten * background(cur, env);
        light += atten * material_emitted(*rec.mat, rec);
        ray scattered;
        color a;
        if (!material_scatter(*rec.mat, cur, rec, a, scattered)) break;
        atten = atten * a;
        cur = scattered;
        if (i > 10 && atten.length() < 0.0001) break;  // :968
        if (i > 10) {                                   // Russian roulette :976-983
            double p = std::clamp(std::max({atten.x(), atten.y(), atten.z()}), 0.05, 0.95);
            if (random_double() > p) break;
            atten /= p;
        }
    }
    return light;  // hard stop at `depth` segments keeps the light gathered so far
}

color ray_color_from_hit(const ray& r, const hit_rec& first, const hittable& world, int depth, const EnvironmentSettings& env,
                         unsigned long long& rays) {
    color light = material_emitted(*first.mat, first);
    ray scattered;
    color a;
    if (material_scatter(*first.mat, r, first, a, scattered)) return light + a * ray_color(scattered, world, depth - 1, env, rays);
    return light;
}

// ------------------------------------------------------------------ commands
std::string dir_of(const std::string& p) {
    size_t s = p.find_last_of('/');
    return s == std::string::npos ? std::string() : p.substr(0, s);
}

shared_ptr<hittable> load_scene(const std::string& path) {
    std::ifstream in(path);
    if (!in) { std::fprintf(stderr, "cannot open scene %s\n", path.c_str()); std::exit(2); }
    zscene::loader ld;
    ld.base_dir = dir_of(path);
    ld.parse(in);
    shared_ptr<hittable> world = ld.build_world();
    prepare(*world);
    return world;
}

void write_header(FILE* f, int w, int h) {
    int32_t hdr[2] = {w, h};
    std::fwrite(hdr, 4, 2, f);
}

int cmd_ids(const std::string& scene, const std::string& cfg, const std::string& out) {
    auto world = load_scene(scene);
    cam_t cam;
    cam.load(config::load(cfg));
    cam.defocus_angle = 0.0;  // "jitter and DOF disabled"
    cam.initialize();
    const int W = cam.W, H = cam.H;
    std::vector<uint32_t> leaf(size_t(W) * H), tri(size_t(W) * H);
    std::vector<double> t(size_t(W) * H);
#pragma omp parallel for schedule(dynamic, 4)
    for (int j = 0; j < H; j++) {
        tl_gen.seed(0x1D5u + j);  // only media draw random numbers on this path
        for (int i = 0; i < W; i++) {
            hit_rec rec;
            size_t idx = size_t(j) * W + i;
            if (hit_any(*world, cam.get_center_ray(i, j), k_ray_t, rec)) {
                leaf[idx] = uint32_t(rec.leaf);
                tri[idx] = rec.tri >= 0 ? uint32_t(rec.tri) : 0u;
                t[idx] = rec.t;
            } else {
                leaf[idx] = 0xFFFFFFFFu;
                tri[idx] = 0;
                t[idx] = infinity;
            }
        }
    }
    FILE* f = std::fopen(out.c_str(), "wb");
    if (!f) return 2;
    write_header(f, W, H);
    std::fwrite(leaf.data(), 4, leaf.size(), f);
    std::fwrite(tri.data(), 4, tri.size(), f);
    std::fwrite(t.data(), 8, t.size(), f);
    std::fclose(f);
    std::printf("{\"cmd\": \"ids\", \"width\": %d, \"height\": %d}\n", W, H);
    return 0;
}

// rays file: int32 n, n x 3 float32 origins, n x 3 float32 directions; output: int32 n, n x 19 float64 =
// t, leaf, tri, front_face, p, normal, u, v, albedo, emitted   (background in the albedo slot on a miss)
int cmd_probe(const std::string& scene, const std::string& cfg, const std::string& rays, const std::string& out) {
    auto world = load_scene(scene);
    config c = config::load(cfg);
    EnvironmentSettings env;
    load_env(c, dir_of(cfg), env);
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
        tl_gen.seed(0x9E37u + i);
        ray r{point3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), vec3(d[3 * i], d[3 * i + 1], d[3 * i + 2])};
        hit_rec rec;
        double* q = &rec_out[size_t(i) * 19];
        if (hit_any(*world, r, k_ray_t, rec)) {
            q[0] = rec.t;
            q[1] = double(rec.leaf);
            q[2] = rec.tri >= 0 ? double(rec.tri) : 0.0;
            q[3] = rec.front_face ? 1.0 : 0.0;
            for (int k = 0; k < 3; k++) { q[4 + k] = rec.p[k]; q[7 + k] = rec.normal[k]; }
            q[10] = rec.u;
            q[11] = rec.v;
            color a = material_albedo(*rec.mat, rec), e = material_emitted(*rec.mat, rec);
            for (int k = 0; k < 3; k++) { q[12 + k] = a[k]; q[15 + k] = e[k]; }
        } else {
            q[0] = infinity;
            q[1] = -1.0;
            color b = background(r, env);
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

// render_rows (camera.hpp:430-554); rows are independent and seeded by (seed, row)
int cmd_render(const std::string& scene, const std::string& cfg, const std::string& out, unsigned long long seed) {
    auto world = load_scene(scene);
    config c = config::load(cfg);
    cam_t cam;
    cam.load(c);
    EnvironmentSettings env;
    load_env(c, dir_of(cfg), env);
    const int W = cam.W, H = cam.H, spp = cam.spp;
    const size_t n = size_t(W) * H;
    std::vector<color> beauty(n), albedo(n), normal(n), zdepth(n), reflection(n), refraction(n);
    unsigned long long rays = 0;
    auto t0 = std::chrono::steady_clock::now();
    cam.initialize();
    const int aux_sample = std::clamp(spp / 8, 64, 1024);  // camera.hpp:433
    const double light_scale = 1.0 / spp, aux_scale = 1.0 / std::min(aux_sample, spp);
#pragma omp parallel for schedule(dynamic, 2) reduction(+ : rays)
    for (int j = 0; j < H; j++) {
        tl_gen.seed(seed * 0x9E3779B97F4A7C15ull + j);
        for (int i = 0; i < W; i++) {
            color pc(0, 0, 0), pa(0, 0, 0), pn(0, 0, 0), pz(0, 0, 0), prefl(0, 0, 0), prefr(0, 0, 0);
            for (int s = 0; s < spp; s++) {
                ray r = cam.get_ray(i, j);
                hit_rec rec;
                rays++;
                if (hit_any(*world, r, k_ray_t, rec)) {
                    pc += ray_color_from_hit(r, rec, *world, cam.max_depth, env, rays);
                    if (s < aux_sample) {  // camera.hpp:464-488
                        if (cam.use_albedo) pa += material_albedo(*rec.mat, rec);
                        if (cam.use_normal) {
                            vec3 nn = unit_vector(rec.normal);
                            pn += color((dot(nn, cam.u) + 1.0) * 0.5, (dot(nn, cam.v) + 1.0) * 0.5, (dot(nn, cam.w) + 1.0) * 0.5);
                        }
                        if (cam.use_z) {
                            double z = 1.0 - std::clamp(rec.t / cam.z_max, 0.0, 1.0);
                            pz += color(z, z, z);
                        }
                    }
                    if (cam.use_reflection || cam.use_refraction) {  // light passes, camera.hpp:492-517
                        ray scattered;
                        color attenuation;
                        if (material_scatter(*rec.mat, r, rec, attenuation, scattered)) {
                            unsigned long long unused = 0;  // light-pass rays are not counted
                            color sc = ray_color(scattered, *world, cam.max_depth - 1, env, unused);
                            double luma = 0.2126 * sc.length();  // vector length, not luminance
                            if (luma > 2.0) sc *= (2.0 / luma);
                            vec3 mirror = reflect(unit_vector(r.d), unit_vector(rec.normal));
                            if (dot(unit_vector(scattered.d), mirror) > 0.9) prefl += attenuation * sc;
                            else if (dot(scattered.d, rec.normal) < 0) prefr += attenuation * sc;
                        }
                    }
                } else {
                    pc += background(r, env);
                    if (s < aux_sample && cam.use_normal) pn += color(0.5, 0.5, 1.0);  // camera.hpp:518-526
                }
            }
            size_t idx = size_t(j) * W + i;
            beauty[idx] = pc * light_scale;
            albedo[idx] = pa * aux_scale;
            normal[idx] = pn * aux_scale;
            zdepth[idx] = pz * aux_scale;
            reflection[idx] = prefl * light_scale;
            refraction[idx] = prefr * light_scale;
        }
    }
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (out != "-") {
        FILE* f = std::fopen(out.c_str(), "wb");
        if (!f) return 2;
        write_header(f, W, H);
        for (auto* buf : {&beauty, &albedo, &normal, &zdepth, &reflection, &refraction})
            for (const color& v : *buf) {
                double e[3] = {v.x(), v.y(), v.z()};
                std::fwrite(e, 8, 3, f);
            }
        std::fclose(f);
    }
    int threads = 1;
#ifdef _OPENMP
    threads = omp_get_max_threads();
#endif
    double paths = double(n) * spp;
    std::printf("{\"cmd\": \"render\", \"width\": %d, \"height\": %d, \"spp\": %d, \"seconds\": %.6f, \"threads\": %d, "
                "\"paths\": %.0f, \"mpaths_per_s\": %.6f, \"rays\": %llu}\n",
                W, H, spp, secs, threads, paths, paths / secs * 1e-6, rays);
    return 0;
}


// ------------------------------------------------------------------ finishing chain
// analyze_framebuffer / apply_auto_exposure (color_processing.hpp:153-210), update_post_processing
// (camera.hpp:142-207), bloom (bloom.hpp:18-68), sharpening (color_processing.hpp:212-232), process (:76-150).
// Same file formats as ref_harness's `finish`.
struct post_t {
    float exposure = 0.5f, saturation = 1, contrast = 1, hue_shift = 0, vignette = 1, stops = 0, target = 0.12f;
    vec3 balance = vec3(1, 1, 1);
    bool aces = false, auto_exp = false, dr = false, dg = false, db = false, dl = false, dbvh = false, bloom = false, sharpen = false;
    float bloom_threshold = 1, bloom_intensity = 0.3f;
    int bloom_radius = 4;
    double sharpen_amount = 0.2;
};
inline double lum709(const color& c) { return 0.2126 * c[0] + 0.7152 * c[1] + 0.0722 * c[2]; }  // vec3.hpp:106-108
inline double gamma22(double v) { return v > 0 ? std::pow(v, 1.0 / 2.2) : 0; }                    // common.hpp:67-72
inline color clamp_gamma(const color& c) { return color(gamma22(clamp01(c[0])), gamma22(clamp01(c[1])), gamma22(clamp01(c[2]))); }
inline double aces1(double v) {                                                                    // common.hpp:48-64
    if (std::isnan(v) || std::isinf(v)) return 0.0;
    double x = std::max(0.0, v);
    return (x * (2.51 * x + 0.03)) / (x * (2.43 * x + 0.59) + 0.14);
}
color post_process(color c, float u, float v, const post_t& P) {  // color_processing.hpp:76-150, beauty pass
    c = c * double(P.exposure);
    c = c * P.balance;
    if (std::abs(P.contrast - 1.0f) > 0.001f)
        c = color(std::max(0.0, (c[0] - 0.18) * P.contrast + 0.18), std::max(0.0, (c[1] - 0.18) * P.contrast + 0.18),
                  std::max(0.0, (c[2] - 0.18) * P.contrast + 0.18));
    if (P.vignette > 0.0f) {
        float dist = std::sqrt((u - 0.5f) * (u - 0.5f) + (v - 0.5f) * (v - 0.5f));
        c = c * double(std::clamp(1.0f - dist * P.vignette, 0.0f, 1.0f));
    }
    if (std::abs(P.saturation - 1.0f) > 0.001f || std::abs(P.hue_shift) > 0.001f) {
        double luma = lum709(c);
        if (luma > 0.0001) {
            color nrm = c / luma;
            float r = float(nrm[0]), g = float(nrm[1]), b = float(nrm[2]);  // rgb_to_hsv works in float (:280-306)
            float mx = std::max({r, g, b}), mn = std::min({r, g, b}), d = mx - mn, h = 0.0f;
            float sat = mx < 1e-6f ? 0.0f : d / mx;
            if (mx != mn) {
                if (mx == r) h = (g - b) / d + (g < b ? 6.0f : 0.0f);
                else if (mx == g) h = (b - r) / d + 2.0f;
                else h = (r - g) / d + 4.0f;
                h /= 6.0f;
            }
            double hue = std::fmod(double(h * 360.0f) + P.hue_shift, 360.0);
            if (hue < 0) hue += 360.0;
            float s2 = std::clamp(float(double(sat) * P.saturation), 0.0f, 1.0f);
            float hh = float(hue) / 360.0f, vv = mx;  // hsv_to_rgb (:308-343)
            int i = int(hh * 6.0f);
            float f = hh * 6.0f - float(i), p = vv * (1.0f - s2), q = vv * (1.0f - f * s2), t = vv * (1.0f - (1.0f - f) * s2);
            color rgb;
            switch (i % 6) {
                case 0: rgb = color(vv, t, p); break;
                case 1: rgb = color(q, vv, p); break;
                case 2: rgb = color(p, vv, t); break;
                case 3: rgb = color(p, q, vv); break;
                case 4: rgb = color(t, p, vv); break;
                case 5: rgb = color(vv, p, q); break;
                default: rgb = color(0, 0, 0);
            }
            c = rgb * luma;
        }
    }
    if (P.aces) c = color(aces1(c[0]), aces1(c[1]), aces1(c[2]));
    if (P.dr || P.dg || P.db || P.dl || P.dbvh) {  // apply_debug_view (:246-278)
        if (P.dl) {
            double l = lum709(c);
            c = l >= 1.0 ? color(1, 1, 1) : l > 0.95 ? color(1, 0, 0) : l > 0.70 ? color(1, 1, 0) : l > 0.40 ? color(.5, .5, .5)
              : l > 0.10 ? color(0, .5, 0) : l > 0.02 ? color(0, 0, 1) : color(0.1, 0.0, 0.2);
        } else if (!P.dbvh) {
            c = color(P.dr ? c[0] : 0.0, P.dg ? c[1] : 0.0, P.db ? c[2] : 0.0);
        }
    }
    return clamp_gamma(c);
}

void blur_pass(const std::vector<color>& in, std::vector<color>& out, int W, int H, int r, bool horizontal) {  // bloom.hpp:48-67
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            color sum(0, 0, 0);
            float total = 0.0f;
            for (int o = -r; o <= r; o++) {
                int sx = x + (horizontal ? o : 0), sy = y + (horizontal ? 0 : o);
                if (sx < 0 || sx >= W || sy < 0 || sy >= H) continue;
                float w = 1.0f - (std::abs(o) / float(r + 1));
                sum += in[size_t(sy) * W + sx] * double(w);
                total += w;
            }
            out[size_t(y) * W + x] = total > 0 ? sum / double(total) : color(0, 0, 0);
        }
}

bool load_image(const std::string& in, int& W, int& H, std::vector<color>& img) {
    FILE* f = std::fopen(in.c_str(), "rb");
    if (!f) return false;
    int32_t wh[2];
    if (std::fread(wh, 4, 2, f) != 2) return false;
    W = wh[0];
    H = wh[1];
    img.resize(size_t(W) * H);
    for (color& p : img) {
        double v[3];
        if (std::fread(v, 8, 3, f) != 3) return false;
        p = color(v[0], v[1], v[2]);
    }
    std::fclose(f);
    return true;
}

post_t load_post(const config& c) {
    post_t P;
    P.exposure = float(c.num("exposure", 0.5));
    P.saturation = float(c.num("saturation", 1.0));
    P.contrast = float(c.num("contrast", 1.0));
    P.hue_shift = float(c.num("hue_shift", 0.0));
    P.vignette = float(c.num("vignette_intensity", 1.0));
    P.balance = c.v3("color_balance", vec3(1, 1, 1));
    P.stops = float(c.num("exposure_compensation_stops", 0.0));
    P.target = float(c.num("target_luminance", 0.12));
    P.aces = c.num("use_aces", 0) != 0;
    P.auto_exp = c.num("use_auto_exposure", 0) != 0;
    P.dr = c.num("debug_red", 0) != 0; P.dg = c.num("debug_green", 0) != 0; P.db = c.num("debug_blue", 0) != 0;
    P.dl = c.num("debug_luminance", 0) != 0; P.dbvh = c.num("debug_bvh", 0) != 0;
    P.bloom = c.num("use_bloom", 0) != 0;
    P.bloom_threshold = float(c.num("bloom_threshold", 1.0));
    P.bloom_intensity = float(c.num("bloom_intensity", 0.3));
    P.bloom_radius = int(c.num("bloom_radius", 4));
    P.sharpen = c.num("use_sharpening", 0) != 0;
    P.sharpen_amount = c.num("sharpen_amount", 0.2);
    return P;
}

// bloom_filter::generate_bloom_overlay (bloom.hpp:18-46) with its exposure argument 1.0f
std::vector<color> bloom_overlay(const std::vector<color>& src, int W, int H, const post_t& P) {
    const size_t n = src.size();
    std::vector<color> bright(n, color(0, 0, 0)), tmp(n), overlay(n);
    for (size_t i = 0; i < n; i++) {  // bright pass, bloom.hpp:27-38
        color e = src[i] * 1.0;
        float lum = float(lum709(e));
        if (lum > P.bloom_threshold) bright[i] = e * double((lum - P.bloom_threshold) * P.bloom_intensity / std::max(lum, 0.0001f));
    }
    blur_pass(bright, tmp, W, H, P.bloom_radius, true);
    blur_pass(tmp, overlay, W, H, P.bloom_radius, false);
    return overlay;
}

void sharpen(std::vector<color>& fin, int W, int H, double amount) {  // color_processing.hpp:212-232
    if (amount <= 0.0) return;
    std::vector<color> orig = fin;
    for (int y = 1; y < H - 1; y++)
        for (int x = 1; x < W - 1; x++) {
            size_t i = size_t(y) * W + x;
            color sum = orig[i] * 5.0 - orig[i - W] - orig[i + W] - orig[i - 1] - orig[i + 1];
            fin[i] = (orig[i] * (1.0 - amount)) + (sum * amount);
        }
}

int cmd_finish(const std::string& in, const std::string& cfg, const std::string& out) {
    int W = 0, H = 0;
    std::vector<color> img;
    if (!load_image(in, W, H, img)) return 2;
    const int32_t wh[2] = {W, H};
    const size_t n = img.size();
    FILE* f = nullptr;
    config c = config::load(cfg);
    post_t P = load_post(c);
    const int pass = int(c.num("pass", 0));
    // analyze_framebuffer (color_processing.hpp:153-186)
    int32_t hist[256] = {0};
    double total_log = 0.0;
    float max_lum = 0.0f;
    for (const color& p : img) {
        float lum = float(lum709(p));
        if (lum > max_lum) max_lum = lum;
        float cl = std::max(0.0001f, lum), l2 = std::log2(cl);
        total_log += l2;
        hist[std::clamp(int(((l2 - (-10.0f)) / 20.0f) * 255.0f), 0, 255)]++;
    }
    float avg = std::pow(2.0f, float(total_log / double(n)));
    // apply_auto_exposure (color_processing.hpp:189-210), stored back by render() (camera.hpp:259-263)
    if (P.auto_exp && avg > 0.0) {
        double cur = (P.target / std::max(double(avg), 0.02)) * std::pow(2.0, double(P.stops));
        P.exposure = std::clamp(float(cur), 0.01f, 4.0f);
    }
    // update_post_processing (camera.hpp:142-207)
    std::vector<color> fin = img;
    const double ev = std::pow(2.0, P.exposure);
    if (pass == 0) {
        for (color& p : fin) p *= ev;
        if (P.bloom) {
            std::vector<color> overlay = bloom_overlay(fin, W, H, P);
            for (size_t i = 0; i < n; i++) fin[i] += overlay[i];
        }
        if (P.sharpen) sharpen(fin, W, H, P.sharpen_amount);
    }
    for (size_t i = 0; i < n; i++) {
        float u = float(i % W) / (W - 1), v = float(i / W) / (H - 1);
        if (pass == 0) fin[i] = post_process(fin[i], u, v, P);
        else if (pass == 1) fin[i] = post_process(fin[i] * ev, u, v, P);
        else fin[i] = clamp_gamma(fin[i]);
    }
    f = std::fopen(out.c_str(), "wb");
    if (!f) return 2;
    std::fwrite(wh, 4, 2, f);
    for (const color& p : fin) {
        double e[3] = {p[0], p[1], p[2]};
        std::fwrite(e, 8, 3, f);
    }
    double s3[3] = {avg, max_lum, P.exposure};
    std::fwrite(s3, 8, 3, f);
    std::fwrite(hist, 4, 256, f);
    std::fclose(f);
    std::printf("{\"cmd\": \"finish\", \"width\": %d, \"height\": %d, \"exposure\": %.9g}\n", W, H, double(P.exposure));
    return 0;
}

// Image export: camera::save_render_pass (camera.hpp:299-343) -> process_framebuffer_to_image (:701-780), up to the
// 8-bit RGB rows handed to the PNG encoder.  cfg `render_pass` is the reference's enum (common.hpp:110-118): RGB and
// DENOISE take bloom (found on the exposed image, added back unexposed), sharpening, exposure and process(); every
// other pass (the light passes too) is clamp + gamma.  Output: int32 w, h; w*h*3 bytes.
int cmd_export(const std::string& in, const std::string& cfg, const std::string& out) {
    int W = 0, H = 0;
    std::vector<color> img;
    if (!load_image(in, W, H, img)) return 2;
    const size_t n = img.size();
    config c = config::load(cfg);
    const post_t P = load_post(c);
    const int rp = int(c.num("render_pass", 0));
    const bool is_data = rp != 0 && rp != 1;
    const double ev = std::pow(2.0, double(P.exposure));
    std::vector<color> buf = img;
    if (!is_data && P.bloom) {
        std::vector<color> exposed = img;
        for (color& p : exposed) p *= ev;
        std::vector<color> overlay = bloom_overlay(exposed, W, H, P);
        const double inv_ev = 1.0 / ev;
        for (size_t i = 0; i < n; i++) buf[i] = img[i] + overlay[i] * inv_ev;
    }
    if (!is_data && P.sharpen) sharpen(buf, W, H, P.sharpen_amount);
    std::vector<unsigned char> rgb(n * 3);
    for (size_t i = 0; i < n; i++) {
        color px = buf[i];
        if (!is_data) {
            float u = float(i % W) / (W - 1), v = float(i / W) / (H - 1);
            px = post_process(px * ev, u, v, P);
        } else {
            px = clamp_gamma(px);
        }
        for (int k = 0; k < 3; k++) rgb[3 * i + k] = static_cast<unsigned char>(255.999 * px[k]);
    }
    FILE* f = std::fopen(out.c_str(), "wb");
    if (!f) return 2;
    const int32_t wh[2] = {W, H};
    std::fwrite(wh, 4, 2, f);
    std::fwrite(rgb.data(), 1, rgb.size(), f);
    std::fclose(f);
    std::printf("{\"cmd\": \"export\", \"width\": %d, \"height\": %d, \"render_pass\": %d}\n", W, H, rp);
    return 0;
}

}  // namespace

int main(int argc, char** argv) {
    std::vector<std::string> a(argv + 1, argv + argc);
    std::cout.setstate(std::ios::failbit);  // silence the host model loader's progress line; results go through printf
    if (a.size() >= 4 && a[0] == "ids") return cmd_ids(a[1], a[2], a[3]);
    if (a.size() >= 5 && a[0] == "probe") return cmd_probe(a[1], a[2], a[3], a[4]);
    if (a.size() >= 4 && a[0] == "finish") return cmd_finish(a[1], a[2], a[3]);
    if (a.size() >= 4 && a[0] == "export") return cmd_export(a[1], a[2], a[3]);
    if (a.size() >= 4 && a[0] == "render") return cmd_render(a[1], a[2], a[3], a.size() > 4 ? std::strtoull(a[4].c_str(), nullptr, 10) : 1ull);
    std::fprintf(stderr, "usage: zoracle ids SCENE CFG OUT | probe SCENE CFG RAYS OUT | render SCENE CFG OUT|- [SEED]\n");
    return 1;
}
