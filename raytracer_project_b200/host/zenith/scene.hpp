// zenith/scene.hpp — the reference-shaped scene graph (C++20 host side of the boundary).
// Class names and constructor signatures follow hittable.hpp:29-36, hittable_list.hpp:9-25,
// bvh.hpp:11-15, sphere.hpp:7, triangle.hpp:7, cube.hpp:11,24, translate.hpp:8, rotate_x.hpp:9,
// rotate_y.hpp:9, rotate_z.hpp:9, scale.hpp:7, material_instance.hpp:7, constant_medium.hpp:27,33,
// model.hpp:12.  Objects only DESCRIBE the scene and report bounding boxes;
// ray intersection happens on the GPU after flatten() (flatten.hpp).  There is no host hit().
#pragma once
#include <string>
#include <vector>

#include "materials.hpp"

namespace zenith {

class hittable {
public:
    enum class kind_t { list, bvh, sphere, triangle, cube, translate, rotate_x, rotate_y, rotate_z, scale,
                        material_instance, constant_medium, model };
    virtual ~hittable() = default;
    virtual kind_t kind() const = 0;
    virtual aabb bounding_box() const = 0;
};

class hittable_list : public hittable {
public:
    std::vector<shared_ptr<hittable>> objects;
    hittable_list() {}
    hittable_list(shared_ptr<hittable> object) { add(object); }
    void clear() { objects.clear(); bbox = aabb(); }
    void add(shared_ptr<hittable> object) {
        objects.push_back(object);
        bbox = aabb(bbox, object->bounding_box());
    }
    kind_t kind() const override { return kind_t::list; }
    aabb bounding_box() const override { return bbox; }

private:
    aabb bbox;
};

// bvh_node(hittable_list) bvh.hpp:11.  The reference builds a random-axis median-split pointer tree
// here; this host only records the member list — the flattener builds the real (binned-SAH, flat)
// BVH for the GPU, so construction is O(N).
class bvh_node : public hittable {
public:
    bvh_node(hittable_list list) : objects(std::move(list.objects)) { compute_box(); }
    bvh_node(std::vector<shared_ptr<hittable>>& objs, size_t start, size_t end)
        : objects(objs.begin() + start, objs.begin() + end) { compute_box(); }
    kind_t kind() const override { return kind_t::bvh; }
    aabb bounding_box() const override { return bbox; }
    std::vector<shared_ptr<hittable>> objects;

private:
    void compute_box() { for (auto& o : objects) bbox = aabb(bbox, o->bounding_box()); }
    aabb bbox;
};

class sphere : public hittable {
public:
    sphere(const point3& center, double radius, shared_ptr<material> mat)
        : center(center), radius(std::fmax(0, radius)), mat(mat) {
        vec3 r(radius, radius, radius);  // sphere.hpp:13-14 uses the unclamped radius for the box
        bbox = aabb(center - r, center + r);
    }
    kind_t kind() const override { return kind_t::sphere; }
    aabb bounding_box() const override { return bbox; }
    void set_material(shared_ptr<material> m) { mat = m; }
    point3 center;
    double radius;
    shared_ptr<material> mat;

private:
    aabb bbox;
};

class triangle : public hittable {
public:
    triangle(const point3& a, const point3& b, const point3& c, const vec3& n0, const vec3& n1, const vec3& n2,
             shared_ptr<material> m)
        : v0(a), v1(b), v2(c), n0(n0), n1(n1), n2(n2), mat(m) {}
    kind_t kind() const override { return kind_t::triangle; }
    aabb bounding_box() const override { return box_of(v0, v1, v2); }
    // triangle.hpp:84-101: axes thinner than 1e-4 are padded by 1e-4 on both sides
    static aabb box_of(const point3& a, const point3& b, const point3& c) {
        double lo[3], hi[3];
        for (int k = 0; k < 3; k++) {
            lo[k] = std::fmin(a[k], std::fmin(b[k], c[k]));
            hi[k] = std::fmax(a[k], std::fmax(b[k], c[k]));
            if (hi[k] - lo[k] < 0.0001) { lo[k] -= 0.0001; hi[k] += 0.0001; }
        }
        return aabb(point3(lo[0], lo[1], lo[2]), point3(hi[0], hi[1], hi[2]));
    }
    void set_material(shared_ptr<material> m) { mat = m; }
    point3 v0, v1, v2;
    vec3 n0, n1, n2;
    shared_ptr<material> mat;
};

class cube : public hittable {
public:
    cube(const point3& min_corner, const point3& max_corner, shared_ptr<material> mat)
        : mat(mat), min_p(min_corner), max_p(max_corner) {
        half_extents = 0.5 * (max_corner - min_corner);
        center = min_corner + half_extents;
    }
    cube(const point3& center_pos, shared_ptr<material> mat) : center(center_pos), mat(mat) {
        half_extents = vec3(1.0, 1.0, 1.0);
        min_p = center_pos - half_extents;
        max_p = center_pos + half_extents;
    }
    kind_t kind() const override { return kind_t::cube; }
    aabb bounding_box() const override {  // cube.hpp:34-41
        double d = 0.0001;
        return aabb(interval(min_p.x(), max_p.x()).expand(d), interval(min_p.y(), max_p.y()).expand(d),
                    interval(min_p.z(), max_p.z()).expand(d));
    }
    void set_material(shared_ptr<material> m) { mat = m; }
    vec3 half_extents;
    point3 center;
    shared_ptr<material> mat;
    point3 min_p, max_p;
};

// Shared base of the five transform wrappers: each knows its world->object and object->world maps.
class transform_wrapper : public hittable {
public:
    shared_ptr<hittable> ptr;
    aabb bounding_box() const override { return bbox; }
    virtual affine world_to_object() const = 0;
    virtual affine object_to_world() const = 0;
    // translate / rotate_y re-run set_face_normal on the flipped normal => front_face forced true
    // (translate.hpp:29, rotate_y.hpp:70)
    virtual bool forces_front_face() const { return false; }

protected:
    aabb bbox;
    void box_from_corners() {  // rotate_*.hpp ctor: transform the 8 corners of the child's box
        aabb in = ptr->bounding_box();
        affine m = object_to_world();
        point3 lo(infinity, infinity, infinity), hi(-infinity, -infinity, -infinity);
        for (int i = 0; i < 2; i++)
            for (int j = 0; j < 2; j++)
                for (int k = 0; k < 2; k++) {
                    vec3 c = m.point(vec3(i ? in.x.max : in.x.min, j ? in.y.max : in.y.min, k ? in.z.max : in.z.min));
                    for (int a = 0; a < 3; a++) { lo[a] = std::fmin(lo[a], c[a]); hi[a] = std::fmax(hi[a], c[a]); }
                }
        bbox = aabb(lo, hi);
    }
};

class translate : public transform_wrapper {
public:
    translate(shared_ptr<hittable> p, const vec3& displacement) : offset(displacement) {
        ptr = p;
        bbox = ptr->bounding_box() + offset;
    }
    kind_t kind() const override { return kind_t::translate; }
    affine world_to_object() const override { return {{1, 0, 0, -offset[0], 0, 1, 0, -offset[1], 0, 0, 1, -offset[2]}}; }
    affine object_to_world() const override { return {{1, 0, 0, offset[0], 0, 1, 0, offset[1], 0, 0, 1, offset[2]}}; }
    bool forces_front_face() const override { return true; }
    vec3 offset;
};

// rotate_x / rotate_z take DEGREES, rotate_y takes RADIANS (rotate_x.hpp:9-10, rotate_z.hpp:9-10,
// rotate_y.hpp:9-13) — kept verbatim.
class rotate_x : public transform_wrapper {
public:
    rotate_x(shared_ptr<hittable> p, double angle_degrees) {
        ptr = p;
        s = std::sin(degrees_to_radians(angle_degrees));
        c = std::cos(degrees_to_radians(angle_degrees));
        box_from_corners();
    }
    kind_t kind() const override { return kind_t::rotate_x; }
    affine world_to_object() const override { return {{1, 0, 0, 0, 0, c, s, 0, 0, -s, c, 0}}; }   // rotate_x.hpp:46-50
    affine object_to_world() const override { return {{1, 0, 0, 0, 0, c, -s, 0, 0, s, c, 0}}; }   // rotate_x.hpp:57-59
    double s, c;
};

class rotate_y : public transform_wrapper {
public:
    rotate_y(shared_ptr<hittable> p, double angle_rad) {
        ptr = p;
        s = std::sin(angle_rad);
        c = std::cos(angle_rad);
        box_from_corners();
    }
    kind_t kind() const override { return kind_t::rotate_y; }
    affine world_to_object() const override { return {{c, 0, s, 0, 0, 1, 0, 0, -s, 0, c, 0}}; }   // rotate_y.hpp:44-49
    affine object_to_world() const override { return {{c, 0, -s, 0, 0, 1, 0, 0, s, 0, c, 0}}; }   // rotate_y.hpp:56-62
    bool forces_front_face() const override { return true; }
    double s, c;
};

class rotate_z : public transform_wrapper {
public:
    rotate_z(shared_ptr<hittable> p, double angle_degrees) {
        ptr = p;
        s = std::sin(degrees_to_radians(angle_degrees));
        c = std::cos(degrees_to_radians(angle_degrees));
        box_from_corners();
    }
    kind_t kind() const override { return kind_t::rotate_z; }
    affine world_to_object() const override { return {{c, s, 0, 0, -s, c, 0, 0, 0, 0, 1, 0}}; }   // rotate_z.hpp:45-49
    affine object_to_world() const override { return {{c, -s, 0, 0, s, c, 0, 0, 0, 0, 1, 0}}; }   // rotate_z.hpp:56-58
    double s, c;
};

class scale : public transform_wrapper {
public:
    scale(shared_ptr<hittable> object, const vec3& factors) : s(factors) {
        ptr = object;
        aabb in = ptr->bounding_box();  // scale.hpp:12-17
        bbox = aabb(point3(in.x.min * s.x(), in.y.min * s.y(), in.z.min * s.z()),
                    point3(in.x.max * s.x(), in.y.max * s.y(), in.z.max * s.z()));
    }
    kind_t kind() const override { return kind_t::scale; }
    affine world_to_object() const override { return {{1 / s[0], 0, 0, 0, 0, 1 / s[1], 0, 0, 0, 0, 1 / s[2], 0}}; }
    affine object_to_world() const override { return {{s[0], 0, 0, 0, 0, s[1], 0, 0, 0, 0, s[2], 0}}; }
    vec3 s;
};

class material_instance : public hittable {
public:
    material_instance(shared_ptr<hittable> obj, shared_ptr<material> mat) : object(obj), new_material(mat) {}
    kind_t kind() const override { return kind_t::material_instance; }
    aabb bounding_box() const override { return object->bounding_box(); }
    void set_material(shared_ptr<material> m) { new_material = m; }
    shared_ptr<hittable> object;
    shared_ptr<material> new_material;  // nullptr => magenta lambertian (material_instance.hpp:22-25)
};

class constant_medium : public hittable {
public:
    constant_medium(shared_ptr<hittable> boundary, double density, shared_ptr<texture> tex)
        : boundary(boundary), neg_inv_density(-1.0 / density), phase_function(make_shared<isovolumetric>(tex)) {}
    constant_medium(shared_ptr<hittable> boundary, double density, color c)
        : boundary(boundary), neg_inv_density(-1.0 / density), phase_function(make_shared<isovolumetric>(c)) {}
    kind_t kind() const override { return kind_t::constant_medium; }
    aabb bounding_box() const override { return boundary->bounding_box(); }
    shared_ptr<hittable> boundary;
    double neg_inv_density;
    shared_ptr<material> phase_function;
};

}  // namespace zenith
