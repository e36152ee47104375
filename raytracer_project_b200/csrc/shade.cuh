// shade.cuh — K3: hit-record reconstruction, textures, tagged-union materials, environment.
// Restates hit_record filling of sphere.hpp:41-62 / triangle.hpp:68-80 / cube.hpp:72-142, the wrapper
// post-processing of translate.hpp:28-29, rotate_*.hpp, scale.hpp:31-34, material::scatter/emitted of
// material.hpp:58-279 + constant_medium.hpp:14-18, texture::value of texture.hpp:50-126 and
// camera::get_background_color of camera.hpp:828-925 — in fp32, on the winning primitive only.
#pragma once
#include "trace.cuh"

struct Surface {
    float3 p, n;            // world hit point, shading normal opposing the ray (hit_record::normal)
    float3 tangent, bitangent;
    float u, v;
    bool front_face;
    uint32_t material;
    uint32_t leaf_id, prim_id;
    float4 m0, m1;          // the zb_material record of the hit (type, textures, bump | rgb, param)
};

ZB_DEV float4 load_mat0(const SceneDev& sc, uint32_t m) { return __ldg(sc.materials + 2 * size_t(m)); }
ZB_DEV float4 load_mat1(const SceneDev& sc, uint32_t m) { return __ldg(sc.materials + 2 * size_t(m) + 1); }

// sphere.hpp:41-58 (outward normal, face flip, uv, tangent frame from the FLIPPED normal)
ZB_DEV void sphere_surface(float4 s, float3 p, float3 d, Surface& sf, bool want_uv) {
    float3 on = __fdividef(1.0f, s.w) * f3(p.x - s.x, p.y - s.y, p.z - s.z);
    sf.front_face = dot(d, on) < 0.f;
    sf.n = sf.front_face ? on : -on;
    if (want_uv) {
        float theta = acosf(fminf(fmaxf(-on.y, -1.f), 1.f));
        float phi = atan2f(-on.z, on.x) + 3.14159265358979f;
        sf.u = phi * (1.0f / 6.28318530717959f);
        sf.v = theta * (1.0f / 3.14159265358979f);
        float3 tg = cross(f3(0.f, 1.f, 0.f), sf.n);
        if (dot(tg, tg) < 0.001f) tg = cross(f3(0.f, 0.f, 1.f), sf.n);
        sf.tangent = unit_vector(tg);
        sf.bitangent = cross(sf.n, sf.tangent);
    }
}

// cube.hpp:72-84, :100-142 (face by 1e-3 proximity cascade X -> Y -> Z, default +Z)
ZB_DEV void cube_surface(const SceneDev& sc, uint32_t cube, float3 p, float3 d, Surface& sf) {
    float3 h = f3(__ldg(sc.cubes + 2 * size_t(cube)));
    float3 c = f3(__ldg(sc.cubes + 2 * size_t(cube) + 1));
    float3 q = p - c;
    const float EPS = 1e-3f;
    float3 n, tg;
    if (fabsf(q.x + h.x) < EPS) {
        n = f3(-1, 0, 0); sf.u = (q.z + h.z) / (2 * h.z); sf.v = (q.y + h.y) / (2 * h.y); tg = f3(0, 0, 1);
    } else if (fabsf(q.x - h.x) < EPS) {
        n = f3(1, 0, 0); sf.u = (q.z + h.z) / (2 * h.z); sf.v = (q.y + h.y) / (2 * h.y); tg = f3(0, 0, -1);
    } else if (fabsf(q.y + h.y) < EPS) {
        n = f3(0, -1, 0); sf.u = (q.x + h.x) / (2 * h.x); sf.v = (q.z + h.z) / (2 * h.z); tg = f3(1, 0, 0);
    } else if (fabsf(q.y - h.y) < EPS) {
        n = f3(0, 1, 0); sf.u = (q.x + h.x) / (2 * h.x); sf.v = (q.z + h.z) / (2 * h.z); tg = f3(-1, 0, 0);
    } else if (fabsf(q.z + h.z) < EPS) {
        n = f3(0, 0, -1); sf.u = (h.x - q.x) / (2 * h.x); sf.v = (q.y + h.y) / (2 * h.y); tg = f3(-1, 0, 0);
    } else {
        n = f3(0, 0, 1); sf.u = (q.x + h.x) / (2 * h.x); sf.v = (q.y + h.y) / (2 * h.y); tg = f3(1, 0, 0);
    }
    sf.tangent = tg;
    sf.bitangent = cross(n, tg);  // from the unflipped normal (cube.hpp:141 runs before set_face_normal)
    sf.front_face = dot(d, n) < 0.f;
    sf.n = sf.front_face ? n : -n;
}

// triangle.hpp:52-80: barycentric smooth normal; u, v, tangent, bitangent are NOT written by the
// reference (stale) — defined here as zero.
ZB_DEV void tri_surface(const SceneDev& sc, uint32_t tri, float3 o, float3 d, Surface& sf) {
    const float4* tp = sc.tris + 3 * size_t(tri);
    float4 a = __ldg(tp);
    float3 v0 = f3(a), v1 = f3(__ldg(tp + 1)), v2 = f3(__ldg(tp + 2));
    float3 e1 = v1 - v0, e2 = v2 - v0, n = cross(e1, e2);
    float inv = 1.0f / dot(n, d);
    float3 tv = v0 - o, q = cross(tv, d);
    float bu = dot(q, e2) * inv, bv = -dot(q, e1) * inv, bw = 1.0f - bu - bv;
    const float4* np = sc.tri_normals + 3 * size_t(tri);
    float3 sn = unit_vector(bw * f3(__ldg(np)) + bu * f3(__ldg(np + 1)) + bv * f3(__ldg(np + 2)));
    sf.front_face = dot(d, sn) < 0.f;
    sf.n = sf.front_face ? sn : -sn;
    sf.u = sf.v = 0.f;
    sf.tangent = sf.bitangent = f3(0, 0, 0);
    sf.prim_id = __float_as_uint(a.w);
}

// Rebuild the reference's final hit_record for the winning item.
// F: the scene features compiled in (ZB_F_*, include/zenith_b200.h)
template <uint32_t F = ZB_F_ALL>
ZB_DEV void make_surface(const SceneDev& sc, const Hit& hit, float3 o, float3 d, Surface& sf, bool force_uv = false) {
    float4 g = __ldg(sc.items + 2 * size_t(hit.item));
    float4 mf = __ldg(sc.items + 2 * size_t(hit.item) + 1);
    uint32_t kf = __float_as_uint(mf.x), kind = kf & 0xFFu, aux = __float_as_uint(mf.w);
    sf.material = __float_as_uint(mf.y);
    sf.leaf_id = __float_as_uint(mf.z);
    sf.prim_id = 0;
    sf.u = sf.v = 0.f;
    sf.tangent = sf.bitangent = f3(0, 0, 0);
    sf.m0 = load_mat0(sc, sf.material);
    sf.m1 = load_mat1(sc, sf.material);
    // sphere uv / tangent frame (sphere.hpp:44-58) feed only texture lookups and bump mapping
    const bool want_uv = (F & ZB_F_TEX) && (force_uv || __float_as_uint(sf.m0.y) != ZB_NO_TEXTURE || __float_as_uint(sf.m0.z) != ZB_NO_TEXTURE);
    if (!(F & ZB_F_GEOM) || kind == ZB_ITEM_SPHERE) {
        sf.p = fma3(hit.t, d, o);
        sphere_surface(g, sf.p, d, sf, want_uv);
        if (kf & ZB_FLAG_FORCE_FRONT) sf.front_face = true;
    } else if ((F & ZB_F_CUBE) && kind == ZB_ITEM_CUBE) {
        sf.p = fma3(hit.t, d, o);
        cube_surface(sc, aux, sf.p, d, sf);
    } else if ((F & ZB_F_TRI) && kind == ZB_ITEM_TRIANGLE) {
        sf.p = fma3(hit.t, d, o);
        tri_surface(sc, aux, o, d, sf);
    } else if ((F & ZB_F_MEDIA) && kind == ZB_ITEM_MEDIUM) {
        sf.p = fma3(hit.t, d, o);
        sf.n = f3(1, 0, 0);
        sf.front_face = true;
    } else if (F & ZB_F_INST) {  // instance: evaluate in object space, then bring p and the normal back
        const float4* ip = sc.instances + 8 * size_t(aux);
        InstanceHead ih = load_instance_head(sc, aux);
        Affine o2w{__ldg(ip + 3), __ldg(ip + 4), __ldg(ip + 5)};
        float3 lo = xform_point(ih.w2o, o), ld = xform_vector(ih.w2o, d);
        float3 lp = fma3(hit.t, ld, lo);
        if (ih.geom_kind == ZB_GEOM_SPHERE) sphere_surface(__ldg(sc.spheres_obj + ih.geom_index), lp, ld, sf, want_uv);
        else if ((F & ZB_F_CUBE) && ih.geom_kind == ZB_GEOM_CUBE) cube_surface(sc, ih.geom_index, lp, ld, sf);
        else if (F & ZB_F_TRI) tri_surface(sc, hit.prim, lo, ld, sf);
        sf.p = xform_point(o2w, lp);
        // rotate_*: n rotated; scale: unit(n / s) (scale.hpp:32-33) == normalise(W2O^T n)
        float3 wn = xform_vector_t(ih.w2o, sf.n);
        sf.n = rsqrtf(fmaxf(dot(wn, wn), 1e-30f)) * wn;
        if (ih.flags & ZB_FLAG_FORCE_FRONT) sf.front_face = true;  // translate.hpp:29 / rotate_y.hpp:70
    }
}

// ---------------------------------------------------------------------------------- textures

ZB_DEV float3 texture_value(const SceneDev& sc, uint32_t tex, float u, float v, float3 p) {
#pragma unroll 1
    for (int guard = 0; guard < 8; guard++) {
        float4 t0 = __ldg(sc.textures + 2 * size_t(tex));
        float4 t1 = __ldg(sc.textures + 2 * size_t(tex) + 1);
        uint32_t kind = __float_as_uint(t0.x), a = __float_as_uint(t0.y), b = __float_as_uint(t0.z), c = __float_as_uint(t0.w);
        if (kind == ZB_TEX_SOLID) return f3(t1);
        if (kind == ZB_TEX_CHECKER) {  // texture.hpp:118-126 (C++ % : negative odd sums are "odd")
            int s = int(floorf(t1.x * p.x)) + int(floorf(t1.x * p.y)) + int(floorf(t1.x * p.z));
            tex = (s % 2 == 0) ? b : a;
            continue;
        }
        if (kind == ZB_TEX_MISSING) return f3(0.f, 1.f, 1.f);  // texture.hpp:52-54
        // image_texture::value (texture.hpp:56-77): u wrapped, v only clamped, nearest texel
        u = u - floorf(u);
        int i = int(u * float(a)), j = int(v * float(b));
        i = min(max(i, 0), int(a) - 1);
        j = min(max(j, 0), int(b) - 1);
        size_t at = size_t(c) + size_t(j) * a + i;
        if (kind == ZB_TEX_IMAGE_F32) return f3(__ldg(sc.texels_f32 + at));
        uchar4 px = __ldg(sc.texels_u8 + at);
        return f3(px.x * (1.0f / 255.0f), px.y * (1.0f / 255.0f), px.z * (1.0f / 255.0f));
    }
    return f3(0, 0, 0);
}

ZB_DEV float3 material_albedo_tex(const SceneDev& sc, float4 m0, float4 m1, const Surface& sf) {
    uint32_t tex = __float_as_uint(m0.y);
    if (tex == ZB_NO_TEXTURE) return f3(m1.x, m1.y, m1.z);
    return texture_value(sc, tex, sf.u, sf.v, sf.p);
}

// material::get_bumped_normal (material.hpp:35-54)
ZB_DEV float3 bumped_normal(const SceneDev& sc, float4 m0, const Surface& sf) {
    uint32_t bump = __float_as_uint(m0.z);
    if (bump == ZB_NO_TEXTURE) return sf.n;
    const float du = 1.0f / 1024.0f;
    float hc = texture_value(sc, bump, sf.u, sf.v, sf.p).x;
    float hu = texture_value(sc, bump, sf.u + du, sf.v, sf.p).x;
    float hv = texture_value(sc, bump, sf.u, sf.v + du, sf.p).x;
    float fu = (hu - hc) * m0.w, fv = (hv - hc) * m0.w;
    return unit_vector(sf.n - fu * sf.tangent - fv * sf.bitangent);
}

// random_unit_vector (vec3.hpp:184-191): uniform on the sphere; direct map instead of rejection
ZB_DEV float3 random_unit_vector(float x1, float x2) {
    float z = 1.0f - 2.0f * x1;
    float r = sqrt_fast(fmaxf(0.f, 1.0f - z * z));
    float s, c;
    sincos_turn(x2, &s, &c);
    return f3(r * c, r * s, z);
}

// material::emitted (material.hpp:12-14, :261-263)
template <uint32_t F = ZB_F_ALL>
ZB_DEV float3 material_emitted(const SceneDev& sc, const Surface& sf) {
    if (__float_as_uint(sf.m0.x) != ZB_MAT_LIGHT) return f3(0, 0, 0);
    if (!(F & ZB_F_TEX)) return f3(sf.m1.x, sf.m1.y, sf.m1.z);
    return material_albedo_tex(sc, sf.m0, sf.m1, sf);
}

// material::get_albedo (material.hpp:28-30, :98-100, :153-155, :226-228, :265-272)
template <uint32_t F = ZB_F_ALL>
ZB_DEV float3 material_get_albedo(const SceneDev& sc, const Surface& sf) {
    const float4 m0 = sf.m0, m1 = sf.m1;
    uint32_t type = __float_as_uint(m0.x);
    if (type == ZB_MAT_DIELECTRIC) return f3(1, 1, 1);
    if (type == ZB_MAT_ISOTROPIC) return f3(0, 0, 0);
    float3 a = (F & ZB_F_TEX) ? material_albedo_tex(sc, m0, m1, sf) : f3(m1.x, m1.y, m1.z);
    if (type == ZB_MAT_LIGHT) return f3(fminf(a.x, 1.f), fminf(a.y, 1.f), fminf(a.z, 1.f));
    return a;
}

// reflect (vec3.hpp:204-206), refract (vec3.hpp:209-214) and dielectric::reflectance (Schlick, material.hpp:237-241) as the
// scatter code below evaluates them; also reached one by one through zb_kat_eval (known-answer tests)
ZB_DEV float3 reflect_dir(float3 v, float3 n) { return fma3(-2.0f * dot(v, n), n, v); }
ZB_DEV float3 refract_dir(float3 ud, float3 wn, float ri, float cos_t) {
    float3 perp = ri * fma3(cos_t, wn, ud);
    float3 par = -sqrtf(fabsf(1.0f - dot(perp, perp))) * wn;
    return perp + par;
}
ZB_DEV float schlick_reflectance(float cos_t, float ri) {
    float r0 = (1.f - ri) / (1.f + ri);
    r0 *= r0;
    float om = 1.f - cos_t, om2 = om * om;
    return r0 + (1.f - r0) * om2 * om2 * om;
}

// material::scatter.  xi = (x, y: unit vector; z: dielectric choice).  Returns false when the path
// ends here (light, absorbed metal sample).
template <uint32_t F = ZB_F_ALL>
ZB_DEV bool material_scatter(const SceneDev& sc, const Surface& sf, float3 d_in, float3 xi, float3& atten,
                             float3& o_out, float3& d_out) {
    const float4 m0 = sf.m0, m1 = sf.m1;
    uint32_t type = __float_as_uint(m0.x);
    const float ray_epsilon = 0.0001f;  // common.hpp:16
    if (type == ZB_MAT_LAMBERTIAN) {  // material.hpp:74-96
        float3 wn = (F & ZB_F_TEX) ? bumped_normal(sc, m0, sf) : sf.n;
        float3 dir = wn + random_unit_vector(xi.x, xi.y);
        if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = wn;
        o_out = fma3(ray_epsilon, sf.n, sf.p);
        d_out = dir;
        atten = (F & ZB_F_TEX) ? material_albedo_tex(sc, m0, m1, sf) : f3(m1.x, m1.y, m1.z);
        return true;
    }
    if (type == ZB_MAT_METAL) {  // material.hpp:129-151
        float3 wn = (F & ZB_F_TEX) ? bumped_normal(sc, m0, sf) : sf.n;
        float3 v = unit_vector(d_in);
        float3 refl = fma3(-2.0f * dot(v, wn), wn, v);
        float3 dir = unit_vector(fma3(m1.w, random_unit_vector(xi.x, xi.y), refl));
        o_out = fma3(ray_epsilon, sf.n, sf.p);
        d_out = dir;
        atten = (F & ZB_F_TEX) ? material_albedo_tex(sc, m0, m1, sf) : f3(m1.x, m1.y, m1.z);
        return dot(dir, sf.n) > 0.f;
    }
    if (type == ZB_MAT_DIELECTRIC) {  // material.hpp:192-224, :237-241; vec3.hpp:204-214
        atten = f3(m1.x, m1.y, m1.z);
        float3 wn = (F & ZB_F_TEX) ? bumped_normal(sc, m0, sf) : sf.n;
        float ri = sf.front_face ? 1.0f / m1.w : m1.w;
        float3 ud = unit_vector(d_in);
        float cos_t = fminf(dot(-ud, wn), 1.0f);
        float sin_t = sqrtf(fmaxf(0.f, 1.0f - cos_t * cos_t));
        float refl = schlick_reflectance(cos_t, ri);
        float3 dir;
        if (ri * sin_t > 1.0f || refl > xi.z) dir = reflect_dir(ud, wn);
        else dir = refract_dir(ud, wn, ri, cos_t);
        float side = dot(dir, sf.n) > 0.f ? ray_epsilon : -ray_epsilon;
        o_out = fma3(side, sf.n, sf.p);
        d_out = dir;
        return true;
    }
    if ((F & ZB_F_MEDIA) && type == ZB_MAT_ISOTROPIC) {  // constant_medium.hpp:14-18
        o_out = sf.p;
        d_out = random_unit_vector(xi.x, xi.y);
        atten = (F & ZB_F_TEX) ? material_albedo_tex(sc, m0, m1, sf) : f3(m1.x, m1.y, m1.z);
        return true;
    }
    return false;  // diffuse_light never scatters (material.hpp:255-259)
}

// ---------------------------------------------------------------------------------- environment

ZB_DEV float clamp01(float x) { return fminf(fmaxf(x, 0.f), 1.f); }

// camera::get_background_color (camera.hpp:828-925)
ZB_DEV float3 background(const SceneDev& sc, const EnvDev& env, float3 d) {
    if (env.mode == ZB_ENV_SOLID_COLOR) return env.background;
    float3 ud = unit_vector(d);
    if (env.mode == ZB_ENV_HDR_MAP) {
        if (env.hdr_texture == ZB_NO_TEXTURE) return f3(0, 0, 0);
        float3 r = f3(dot(env.hdr_r0, ud), dot(env.hdr_r1, ud), dot(env.hdr_r2, ud));
        float phi = atan2f(r.z, r.x) + 3.14159265358979f;
        float theta = acosf(fminf(fmaxf(r.y, -1.f), 1.f));
        return env.intensity *
               texture_value(sc, env.hdr_texture, phi * (1.0f / 6.28318530717959f), theta * (1.0f / 3.14159265358979f), f3(0, 0, 0));
    }
    // PHYSICAL_SUN (camera.hpp:871-924): gradient sky + smoothstep sun disc
    float a = ud.y;
    float3 sky = a > 0.f ? (1.0f - a) * env.horizon + a * env.zenith : 0.1f * env.horizon;
    float3 col = env.sky_scale * sky;
    float focus = dot(ud, env.sun_dir);
    if (focus > env.sun_threshold && env.sun_visible) {
        float x = clamp01((focus - env.sun_threshold) / 0.0002f);
        col += (x * x * (3.f - 2.f * x)) * env.sun_add;
    }
    return col;
}
