// zenith/math.hpp — double-precision vec3 / interval / aabb used by the host scene API.
// API shape follows the reference (vec3.hpp:7-213, interval.hpp:9-59, aabb.hpp:5-94) so scene
// authoring code written against Zenith compiles unchanged; the implementation is new.
// Host-only: no intersection code lives here (the hot path is CUDA, there is no CPU fallback).
#pragma once
#include <algorithm>
#include <cmath>
#include <limits>
#include <memory>
#include <random>

namespace zenith {

using std::make_shared;
using std::shared_ptr;

constexpr double infinity = std::numeric_limits<double>::infinity();
constexpr double pi = 3.14159265358979323846;

inline double degrees_to_radians(double d) { return d * pi / 180.0; }

// common.hpp:29-45: the scene-construction randomness of the reference's API (random_device-seeded mt19937; scene code such
// as load_materials() and build_geometry() calls these; the GPU render never does — it draws from Philox)
inline double random_double() {
    static thread_local std::mt19937 gen{std::random_device{}()};
    static thread_local std::uniform_real_distribution<double> dist(0.0, 1.0);
    return dist(gen);
}
inline double random_double(double lo, double hi) { return lo + (hi - lo) * random_double(); }
inline int random_int(int lo, int hi) { return static_cast<int>(random_double(lo, hi + 1)); }

struct vec3 {
    double e[3];
    constexpr vec3() : e{0, 0, 0} {}
    constexpr vec3(double x, double y, double z) : e{x, y, z} {}
    constexpr double x() const { return e[0]; }
    constexpr double y() const { return e[1]; }
    constexpr double z() const { return e[2]; }
    constexpr double operator[](int i) const { return e[i]; }
    double& operator[](int i) { return e[i]; }
    constexpr vec3 operator-() const { return {-e[0], -e[1], -e[2]}; }
    vec3& operator+=(const vec3& o) { e[0] += o.e[0]; e[1] += o.e[1]; e[2] += o.e[2]; return *this; }
    vec3& operator*=(double t) { e[0] *= t; e[1] *= t; e[2] *= t; return *this; }
    vec3& operator/=(double t) { return *this *= 1.0 / t; }
    double length_squared() const { return e[0] * e[0] + e[1] * e[1] + e[2] * e[2]; }
    double length() const { return std::sqrt(length_squared()); }
    static vec3 random() { return vec3(random_double(), random_double(), random_double()); }  // vec3.hpp:98-103
    static vec3 random(double lo, double hi) { return vec3(random_double(lo, hi), random_double(lo, hi), random_double(lo, hi)); }
};
using point3 = vec3;
using color = vec3;

constexpr vec3 operator+(const vec3& a, const vec3& b) { return {a.e[0] + b.e[0], a.e[1] + b.e[1], a.e[2] + b.e[2]}; }
constexpr vec3 operator-(const vec3& a, const vec3& b) { return {a.e[0] - b.e[0], a.e[1] - b.e[1], a.e[2] - b.e[2]}; }
constexpr vec3 operator*(const vec3& a, const vec3& b) { return {a.e[0] * b.e[0], a.e[1] * b.e[1], a.e[2] * b.e[2]}; }
constexpr vec3 operator*(double t, const vec3& a) { return {t * a.e[0], t * a.e[1], t * a.e[2]}; }
constexpr vec3 operator*(const vec3& a, double t) { return t * a; }
constexpr vec3 operator/(const vec3& a, double t) { return (1.0 / t) * a; }
constexpr double dot(const vec3& a, const vec3& b) { return a.e[0] * b.e[0] + a.e[1] * b.e[1] + a.e[2] * b.e[2]; }
constexpr vec3 cross(const vec3& a, const vec3& b) {
    return {a.e[1] * b.e[2] - a.e[2] * b.e[1], a.e[2] * b.e[0] - a.e[0] * b.e[2], a.e[0] * b.e[1] - a.e[1] * b.e[0]};
}
// vec3.hpp:165-171 — zero vector for near-zero length
inline vec3 unit_vector(const vec3& v) {
    double len = v.length();
    return len < 1e-8 ? vec3(0, 0, 0) : v / len;
}

struct interval {
    double min, max;
    interval() : min(+infinity), max(-infinity) {}
    interval(double lo, double hi) : min(lo), max(hi) {}
    interval(const interval& a, const interval& b) : min(std::fmin(a.min, b.min)), max(std::fmax(a.max, b.max)) {}
    interval expand(double delta) const { return {min - delta / 2, max + delta / 2}; }
    double size() const { return max - min; }
};
inline interval operator+(const interval& i, double d) { return {i.min + d, i.max + d}; }

struct aabb {
    interval x, y, z;
    aabb() = default;
    aabb(const interval& ix, const interval& iy, const interval& iz) : x(ix), y(iy), z(iz) {}
    aabb(const point3& a, const point3& b)
        : x(std::fmin(a[0], b[0]), std::fmax(a[0], b[0])),
          y(std::fmin(a[1], b[1]), std::fmax(a[1], b[1])),
          z(std::fmin(a[2], b[2]), std::fmax(a[2], b[2])) {}
    aabb(const aabb& a, const aabb& b) : x(a.x, b.x), y(a.y, b.y), z(a.z, b.z) {}
    const interval& axis(int n) const { return n == 1 ? y : (n == 2 ? z : x); }
    bool empty() const { return x.max < x.min || y.max < y.min || z.max < z.min; }
};
inline aabb operator+(const aabb& b, const vec3& o) { return {b.x + o.x(), b.y + o.y(), b.z + o.z()}; }

// 3x4 affine in double, used by the flattener to collapse wrapper chains
struct affine {
    double m[12];  // row-major: row r = m[4r..4r+3], last column = translation
    static affine identity() { return {{1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0}}; }
    vec3 point(const vec3& p) const {
        return {m[0] * p[0] + m[1] * p[1] + m[2] * p[2] + m[3], m[4] * p[0] + m[5] * p[1] + m[6] * p[2] + m[7],
                m[8] * p[0] + m[9] * p[1] + m[10] * p[2] + m[11]};
    }
    // this ∘ b  (apply b first)
    affine after(const affine& b) const {
        affine r{};
        for (int i = 0; i < 3; i++) {
            for (int j = 0; j < 4; j++) {
                double s = 0;
                for (int k = 0; k < 3; k++) s += m[4 * i + k] * b.m[4 * k + j];
                r.m[4 * i + j] = s + (j == 3 ? m[4 * i + 3] : 0.0);
            }
        }
        return r;
    }
};

}  // namespace zenith
