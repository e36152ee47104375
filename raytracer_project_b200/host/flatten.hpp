// flatten.hpp — "flatten once": reference-shaped scene graph -> GPU structure-of-arrays (zb_flat_scene).
// Replaces the shared_ptr<hittable>/shared_ptr<material> object graph walked by virtual dispatch
// (hittable.hpp:29-36, material.hpp:7-31, call stack in SURVEY.md §3.3) by:
//   * a two-level flat BVH (TLAS over leaf objects, one BLAS per distinct model),
//   * wrapper chains translate/rotate_*/scale/material_instance collapsed into one affine instance
//     with the reference's per-wrapper normal rules kept as flags,
//   * tagged-union materials and texture descriptors.
// Leaf ids are handed out in DFS order over list/bvh containers; the oracle harness numbers the
// same sites the same way, which is what makes the primary-hit ID buffers comparable.
#pragma once
#include <stdexcept>
#include <unordered_map>

#include "bvh_build.hpp"
#include "wide_bvh.hpp"
#include "zenith/environment.hpp"
#include "zenith/model.hpp"
#include "zenith/scene.hpp"

namespace zenith {

struct flat_scene {
    std::vector<zb_bvh_node> nodes;
    std::vector<zb_wide_node> wide_nodes;
    uint32_t wide_stack_need = 0;
    std::vector<float> item_bounds;  // 8 floats per item, items[] order
    std::vector<zb_item> items;
    std::vector<zb_cube> cubes;
    std::vector<zb_tri> tris;
    std::vector<zb_tri_normals> tri_normals;
    std::vector<zb_instance> instances;
    std::vector<zb_mesh> meshes;
    std::vector<zb_medium> media;
    std::vector<zb_material> materials;
    std::vector<zb_texture> textures;
    std::vector<float> spheres_obj;
    std::vector<uint8_t> texels_u8;
    std::vector<float> texels_f32;
    uint32_t n_tlas_nodes = 0, n_leaves = 0;
    uint32_t env_hdr_texture = ZB_NO_TEXTURE;
    zb_flat_scene view{};

    const zb_flat_scene* finish() {
        view.abi_version = ZB_ABI_VERSION;
        view.n_nodes = uint32_t(nodes.size());
        view.n_tlas_nodes = n_tlas_nodes;
        view.n_items = uint32_t(items.size());
        view.n_cubes = uint32_t(cubes.size());
        view.n_tris = uint32_t(tris.size());
        view.n_instances = uint32_t(instances.size());
        view.n_meshes = uint32_t(meshes.size());
        view.n_media = uint32_t(media.size());
        view.n_materials = uint32_t(materials.size());
        view.n_textures = uint32_t(textures.size());
        view.n_spheres_obj = uint32_t(spheres_obj.size() / 4);
        view.n_texels_u8 = texels_u8.size() / 4;
        view.n_texels_f32 = texels_f32.size() / 4;
        view.nodes = nodes.data();
        view.items = items.data();
        view.cubes = cubes.data();
        view.tris = tris.data();
        view.tri_normals = tri_normals.data();
        view.instances = instances.data();
        view.meshes = meshes.data();
        view.media = media.data();
        view.materials = materials.data();
        view.textures = textures.data();
        view.spheres_obj = spheres_obj.data();
        view.texels_u8 = texels_u8.data();
        view.texels_f32 = texels_f32.data();
        view.n_leaves = n_leaves;
        view.pad = env_hdr_texture;
        view.n_wide_nodes = uint32_t(wide_nodes.size());
        view.wide_stack_need = wide_stack_need;
        view.wide_nodes = wide_nodes.empty() ? nullptr : wide_nodes.data();
        view.item_bounds = item_bounds.size() == items.size() * 8 ? item_bounds.data() : nullptr;
        return &view;
    }
};

class flattener {
public:
    int tlas_leaf = 4, blas_leaf = 4;
    // true: leave every mesh BLAS to the device builder: triangles are emitted in face order, the mesh
    // carries node_count == 0 and zb_scene_upload builds its tree on the GPU — binned SAH (csrc/sah_build.cu) unless the
    // context's tuning "gpu_builder" is 0 (linear BVH, csrc/lbvh.cu)
    bool gpu_blas = false;
    // 8-wide compressed tree beside the binary one (host/wide_bvh.hpp).  OFF by default: measured 2.2x slower than the
    // binary tree on the 1 M and 10 M triangle configs (profiles/r02_wide_bvh.txt: these kernels are issue-bound, not
    // DRAM-bound).  1 = build it for scenes with host-built mesh BLASes, wide_all = for every scene (A/B runs), -1 = never.
    int wide = 0;
    bool wide_all = false;

    std::unique_ptr<flat_scene> run(const hittable& world, const EnvironmentSettings* env = nullptr) {
        out = std::make_unique<flat_scene>();
        item_boxes.clear();
        walk(world);
        if (env && env->hdr_texture) out->env_hdr_texture = texture_id(env->hdr_texture);
        // TLAS over the leaf items; items are permuted into leaf order
        bvh_builder b;
        std::vector<zb_bvh_node> tl;
        std::vector<uint32_t> order;
        // a TLAS item costs more than a node visit (instances: a transform and an object-space test): leaves of one
        // item wherever a split is possible.  Measured against 1.0: exterior -3.3 %, quirks -2.8 %, Weekend +-0.2 %;
        // an exact sweep SAH instead of 16 bins changed nothing (profiles/r01_pq.txt)
        b.isect_cost = 1.5;
        b.build(item_boxes, tlas_leaf, 0, tl, order, ZB_MAX_TLAS_DEPTH);
        std::vector<zb_wide_node> tlw;
        int tlas_wide_depth = 0;
        const bool want_wide = wide >= 0 && tlas_leaf <= 4 && (wide_all || (!out->meshes.empty() && wide > 0));
        if (want_wide) {
            std::vector<uint32_t> perm;
            wide_builder wb;
            if (wb.build(tl, 0, uint32_t(order.size()), tlw, perm, tlas_wide_depth)) {
                std::vector<uint32_t> o2(order.size());
                for (size_t i = 0; i < order.size(); i++) o2[i] = order[perm[i]];
                order.swap(o2);
            } else {
                tlw.clear();
            }
        }
        std::vector<zb_item> sorted(order.size());
        for (size_t i = 0; i < order.size(); i++) sorted[i] = out->items[order[i]];
        out->items.swap(sorted);
        out->item_bounds.resize(order.size() * 8);
        for (size_t i = 0; i < order.size(); i++) {
            const build_box& bb = item_boxes[order[i]];
            float* q = &out->item_bounds[8 * i];
            for (int k = 0; k < 3; k++) { q[k] = round_down(bb.lo[k]); q[4 + k] = round_up(bb.hi[k]); }
            q[3] = q[7] = 0.f;
        }
        out->n_tlas_nodes = uint32_t(tl.size());
        // BLAS nodes were collected with mesh-local child indices; append them after the TLAS
        std::vector<zb_bvh_node> all = tl;
        // room for any TLAS the device may build over these items later (zb_scene_rebuild_tlas): a binary tree over n items
        // has at most n - 1 inner nodes.  The spare nodes are unreachable and empty.
        {
            zb_bvh_node spare{};
            spare.lx_ly[0] = spare.lx_ly[2] = spare.rx_ry[0] = spare.rx_ry[2] = spare.lz_rz[0] = spare.lz_rz[2] = INFINITY;
            spare.lx_ly[1] = spare.lx_ly[3] = spare.rx_ry[1] = spare.rx_ry[3] = spare.lz_rz[1] = spare.lz_rz[3] = -INFINITY;
            spare.left = spare.right = ZB_CHILD_EMPTY;
            while (all.size() + 1 < order.size()) all.push_back(spare);
        }
        for (size_t m = 0; m < out->meshes.size(); m++) {
            uint32_t root = uint32_t(all.size());
            for (zb_bvh_node nd : blas_nodes[m]) {
                if (nd.left >= 0) nd.left += int32_t(root);
                if (nd.right >= 0) nd.right += int32_t(root);
                all.push_back(nd);
            }
            out->meshes[m].node_root = blas_nodes[m].empty() ? 0u : root;
            out->meshes[m].node_count = uint32_t(blas_nodes[m].size());
        }
        for (auto& inst : out->instances)
            if (inst.geom_kind == ZB_GEOM_MESH) inst.pad[0] = out->meshes[inst.geom_index].node_root;
        // the wide trees: TLAS first (root 0), every mesh's BLAS after it; roots into zb_instance::pad[1]
        if (!tlw.empty()) {
            std::vector<zb_wide_node> allw = tlw;
            std::vector<uint32_t> wroot(out->meshes.size(), 0);
            bool complete = true;
            for (size_t m = 0; m < out->meshes.size(); m++) {
                if (blas_wide[m].empty()) { complete = false; break; }
                wroot[m] = uint32_t(allw.size());
                for (zb_wide_node nd : blas_wide[m]) {
                    nd.child_base += wroot[m];
                    allw.push_back(nd);
                }
            }
            if (complete) {
                for (auto& inst : out->instances)
                    if (inst.geom_kind == ZB_GEOM_MESH) inst.pad[1] = wroot[inst.geom_index];
                out->wide_nodes.swap(allw);
                out->wide_stack_need = uint32_t(tlas_wide_depth + blas_wide_depth + 4);
            }
        }
        blas_wide.clear();
        out->nodes.swap(all);
        out->finish();
        blas_nodes.clear();
        return std::move(out);
    }

private:
    std::unique_ptr<flat_scene> out;
    std::vector<build_box> item_boxes;
    std::vector<std::vector<zb_bvh_node>> blas_nodes;
    std::vector<std::vector<zb_wide_node>> blas_wide;
    int blas_wide_depth = 0;
    std::unordered_map<const material*, uint32_t> mat_ids;
    std::unordered_map<const texture*, uint32_t> tex_ids;
    std::unordered_map<const hittable*, uint32_t> mesh_ids, sphere_obj_ids, cube_ids;
    uint32_t magenta_id = ZB_NO_TEXTURE;

    struct chain {
        affine w2o = affine::identity(), o2w = affine::identity();
        bool force_front = false, has_linear = false, has_any = false;
        shared_ptr<material> override_mat;
        bool has_override = false;
    };

    void walk(const hittable& h) {
        if (h.kind() == hittable::kind_t::list) {
            for (auto& c : static_cast<const hittable_list&>(h).objects) walk_child(c);
        } else if (h.kind() == hittable::kind_t::bvh) {
            for (auto& c : static_cast<const bvh_node&>(h).objects) walk_child(c);
        } else {
            // a bare object used as the world
            leaf(h, out->n_leaves++, chain(), h.bounding_box());
        }
    }
    void walk_child(const shared_ptr<hittable>& c) {
        if (c->kind() == hittable::kind_t::list || c->kind() == hittable::kind_t::bvh) walk(*c);
        else leaf(*c, out->n_leaves++, chain(), c->bounding_box());
    }

    // `outer_box` is the reference's bounding_box() of the outermost object of this leaf — the box
    // the reference's own BVH would hold for it.
    void leaf(const hittable& h, uint32_t leaf_id, chain ch, const aabb& outer_box) {
        using K = hittable::kind_t;
        switch (h.kind()) {
            case K::translate: case K::rotate_x: case K::rotate_y: case K::rotate_z: case K::scale: {
                auto& w = static_cast<const transform_wrapper&>(h);
                ch.w2o = w.world_to_object().after(ch.w2o);
                ch.o2w = ch.o2w.after(w.object_to_world());
                ch.force_front = ch.force_front || w.forces_front_face();
                ch.has_any = true;
                if (h.kind() != K::translate) ch.has_linear = true;
                leaf(*w.ptr, leaf_id, ch, outer_box);
                return;
            }
            case K::material_instance: {
                auto& mi = static_cast<const material_instance&>(h);
                if (!ch.has_override) { ch.has_override = true; ch.override_mat = mi.new_material; }
                leaf(*mi.object, leaf_id, ch, outer_box);
                return;
            }
            case K::list:
                for (auto& c : static_cast<const hittable_list&>(h).objects) leaf(*c, leaf_id, ch, transformed_box(*c, ch));
                return;
            case K::bvh:
                for (auto& c : static_cast<const bvh_node&>(h).objects) leaf(*c, leaf_id, ch, transformed_box(*c, ch));
                return;
            case K::sphere: add_sphere(static_cast<const sphere&>(h), leaf_id, ch, outer_box); return;
            case K::cube: add_cube(static_cast<const cube&>(h), leaf_id, ch, outer_box); return;
            case K::triangle: add_triangle(static_cast<const triangle&>(h), leaf_id, ch, outer_box); return;
            case K::model: add_model(static_cast<const model&>(h), leaf_id, ch, outer_box); return;
            case K::constant_medium: add_medium(static_cast<const constant_medium&>(h), leaf_id, ch, outer_box); return;
        }
    }

    static aabb transformed_box(const hittable& c, const chain& ch) {
        aabb in = c.bounding_box();
        point3 lo(infinity, infinity, infinity), hi(-infinity, -infinity, -infinity);
        for (int i = 0; i < 8; i++) {
            vec3 p = ch.o2w.point(vec3(i & 1 ? in.x.max : in.x.min, i & 2 ? in.y.max : in.y.min, i & 4 ? in.z.max : in.z.min));
            for (int a = 0; a < 3; a++) { lo[a] = std::fmin(lo[a], p[a]); hi[a] = std::fmax(hi[a], p[a]); }
        }
        return aabb(lo, hi);
    }

    uint32_t resolve_material(const chain& ch, const shared_ptr<material>& own) {
        if (ch.has_override) return ch.override_mat ? material_id(ch.override_mat) : magenta();
        return own ? material_id(own) : magenta();
    }
    uint32_t magenta() {  // material_instance.hpp:22-25
        if (magenta_id == ZB_NO_TEXTURE) {
            static auto err = make_shared<lambertian>(color(1, 0, 1));
            magenta_id = material_id(err);
        }
        return magenta_id;
    }

    uint32_t texture_id(const shared_ptr<texture>& t) {
        auto it = tex_ids.find(t.get());
        if (it != tex_ids.end()) return it->second;
        zb_texture z{};
        switch (t->kind()) {
            case texture::kind_t::solid: {
                auto& s = static_cast<const solid_color&>(*t);
                z.kind = ZB_TEX_SOLID;
                for (int k = 0; k < 3; k++) z.f[k] = float(s.albedo[k]);
                break;
            }
            case texture::kind_t::checker: {
                auto& c = static_cast<const checker_texture&>(*t);
                z.kind = ZB_TEX_CHECKER;
                z.f[0] = float(c.inv_scale);
                z.a = texture_id(c.odd);
                z.b = texture_id(c.even);
                break;
            }
            case texture::kind_t::image: {
                auto& im = static_cast<const image_texture&>(*t);
                if (im.missing()) { z.kind = ZB_TEX_MISSING; break; }
                z.a = uint32_t(im.width);
                z.b = uint32_t(im.height);
                size_t n = size_t(im.width) * im.height;
                if (im.is_float) {
                    z.kind = ZB_TEX_IMAGE_F32;
                    z.c = uint32_t(out->texels_f32.size() / 4);
                    for (size_t i = 0; i < n; i++) {
                        for (int k = 0; k < 3; k++) out->texels_f32.push_back(im.f32[3 * i + k]);
                        out->texels_f32.push_back(1.0f);
                    }
                } else {
                    z.kind = ZB_TEX_IMAGE_U8;
                    z.c = uint32_t(out->texels_u8.size() / 4);
                    for (size_t i = 0; i < n; i++) {
                        for (int k = 0; k < 3; k++) out->texels_u8.push_back(im.u8[3 * i + k]);
                        out->texels_u8.push_back(255);
                    }
                }
                break;
            }
        }
        uint32_t id = uint32_t(out->textures.size());
        out->textures.push_back(z);
        tex_ids[t.get()] = id;
        return id;
    }

    uint32_t material_id(const shared_ptr<material>& m) {
        auto it = mat_ids.find(m.get());
        if (it != mat_ids.end()) return it->second;
        zb_material z{};
        z.albedo_tex = ZB_NO_TEXTURE;
        z.bump_tex = m->bump ? texture_id(m->bump) : ZB_NO_TEXTURE;
        z.bump_strength = float(m->bump_strength);
        z.param = float(m->param);
        switch (m->kind()) {
            case material::kind_t::lambertian: z.type = ZB_MAT_LAMBERTIAN; break;
            case material::kind_t::metal: z.type = ZB_MAT_METAL; break;
            case material::kind_t::dielectric: z.type = ZB_MAT_DIELECTRIC; break;
            case material::kind_t::diffuse_light: z.type = ZB_MAT_LIGHT; break;
            case material::kind_t::isovolumetric: z.type = ZB_MAT_ISOTROPIC; break;
        }
        if (m->kind() == material::kind_t::dielectric) {
            for (int k = 0; k < 3; k++) z.rgb[k] = float(m->tint[k]);
        } else if (m->tex && m->tex->kind() == texture::kind_t::solid) {
            auto& s = static_cast<const solid_color&>(*m->tex);
            for (int k = 0; k < 3; k++) z.rgb[k] = float(s.albedo[k]);
        } else if (m->tex) {
            z.albedo_tex = texture_id(m->tex);
        }
        uint32_t id = uint32_t(out->materials.size());
        out->materials.push_back(z);
        mat_ids[m.get()] = id;
        return id;
    }

    void push_item(zb_item it, const aabb& box) {
        out->items.push_back(it);
        item_boxes.push_back(build_box::from(box));
    }

    uint32_t add_instance(const chain& ch, uint32_t geom_kind, uint32_t geom_index) {
        zb_instance in{};
        for (int k = 0; k < 12; k++) { in.w2o[k] = float(ch.w2o.m[k]); in.o2w[k] = float(ch.o2w.m[k]); }
        in.geom_kind = geom_kind;
        in.geom_index = geom_index;
        in.flags = ch.force_front ? ZB_FLAG_FORCE_FRONT : 0u;
        out->instances.push_back(in);
        return uint32_t(out->instances.size()) - 1;
    }

    void add_sphere(const sphere& s, uint32_t leaf_id, const chain& ch, const aabb& box) {
        zb_item it{};
        it.material = resolve_material(ch, s.mat);
        it.leaf_id = leaf_id;
        if (!ch.has_linear) {  // identity or pure translation: bake into a world-space sphere
            point3 c = ch.o2w.point(s.center);
            it.geom[0] = float(c[0]); it.geom[1] = float(c[1]); it.geom[2] = float(c[2]); it.geom[3] = float(s.radius);
            it.kind_flags = ZB_ITEM_SPHERE | (ch.force_front ? ZB_FLAG_FORCE_FRONT : 0u);
        } else {
            auto f = sphere_obj_ids.find(&s);
            uint32_t gi;
            if (f == sphere_obj_ids.end()) {
                gi = uint32_t(out->spheres_obj.size() / 4);
                for (int k = 0; k < 3; k++) out->spheres_obj.push_back(float(s.center[k]));
                out->spheres_obj.push_back(float(s.radius));
                sphere_obj_ids[&s] = gi;
            } else gi = f->second;
            it.kind_flags = ZB_ITEM_INSTANCE;
            it.aux = add_instance(ch, ZB_GEOM_SPHERE, gi);
        }
        push_item(it, box);
    }

    uint32_t cube_index(const cube& c) {
        auto f = cube_ids.find(&c);
        if (f != cube_ids.end()) return f->second;
        zb_cube z{};
        for (int k = 0; k < 3; k++) { z.half[k] = float(c.half_extents[k]); z.centre[k] = float(c.center[k]); }
        out->cubes.push_back(z);
        return cube_ids[&c] = uint32_t(out->cubes.size()) - 1;
    }

    void add_cube(const cube& c, uint32_t leaf_id, const chain& ch, const aabb& box) {
        zb_item it{};
        it.material = resolve_material(ch, c.mat);
        it.leaf_id = leaf_id;
        if (!ch.has_any) {
            it.kind_flags = ZB_ITEM_CUBE;
            it.aux = cube_index(c);
        } else {
            it.kind_flags = ZB_ITEM_INSTANCE;
            it.aux = add_instance(ch, ZB_GEOM_CUBE, cube_index(c));
        }
        push_item(it, box);
    }

    uint32_t push_tri(const point3* v, const vec3* n, uint32_t face) {
        zb_tri t{};
        zb_tri_normals tn{};
        for (int k = 0; k < 3; k++) {
            t.v0[k] = float(v[0][k]); t.v1[k] = float(v[1][k]); t.v2[k] = float(v[2][k]);
            tn.n0[k] = float(n[0][k]); tn.n1[k] = float(n[1][k]); tn.n2[k] = float(n[2][k]);
        }
        std::memcpy(&t.v0[3], &face, 4);
        out->tris.push_back(t);
        out->tri_normals.push_back(tn);
        return uint32_t(out->tris.size()) - 1;
    }

    void add_triangle(const triangle& t, uint32_t leaf_id, const chain& ch, const aabb& box) {
        zb_item it{};
        it.material = resolve_material(ch, t.mat);
        it.leaf_id = leaf_id;
        point3 v[3] = {t.v0, t.v1, t.v2};
        vec3 n[3] = {t.n0, t.n1, t.n2};
        uint32_t ti = push_tri(v, n, 0);
        if (!ch.has_any) {
            it.kind_flags = ZB_ITEM_TRIANGLE;
            it.aux = ti;
        } else {
            it.kind_flags = ZB_ITEM_INSTANCE;
            it.aux = add_instance(ch, ZB_GEOM_TRIANGLE, ti);
        }
        push_item(it, box);
    }

    void add_model(const model& m, uint32_t leaf_id, const chain& ch, const aabb& box) {
        if (m.tris.empty()) return;  // the reference would crash on a null mesh BVH
        uint32_t mi;
        auto f = mesh_ids.find(&m);
        if (f == mesh_ids.end()) {
            std::vector<build_box> boxes(m.tris.size());
            for (size_t i = 0; i < m.tris.size(); i++)
                boxes[i] = build_box::from(triangle::box_of(m.tris[i].v[0], m.tris[i].v[1], m.tris[i].v[2]));
            uint32_t first = uint32_t(out->tris.size());
            std::vector<zb_bvh_node> nd;
            if (gpu_blas) {
                for (uint32_t slot = 0; slot < uint32_t(m.tris.size()); slot++) push_tri(m.tris[slot].v, m.tris[slot].n, slot);
            } else {
                bvh_builder b;
                std::vector<uint32_t> order;
                b.build(boxes, blas_leaf, first, nd, order, ZB_MAX_BLAS_DEPTH);
                std::vector<zb_wide_node> wn;
                std::vector<uint32_t> perm;
                int wdepth = 0;
                wide_builder wb;
                if (wide >= 0 && blas_leaf <= 4 && wb.build(nd, first, uint32_t(order.size()), wn, perm, wdepth)) {
                    // triangles go out in the wide tree's order (both trees' leaves were remapped to it)
                    for (uint32_t k : perm) push_tri(m.tris[order[k]].v, m.tris[order[k]].n, order[k]);
                    blas_wide_depth = std::max(blas_wide_depth, wdepth);
                } else {
                    if (wide >= 0) wide = -1;  // one mesh without a wide tree: the scene ships none
                    wn.clear();
                    for (uint32_t slot : order) push_tri(m.tris[slot].v, m.tris[slot].n, slot);
                }
                blas_wide.push_back(std::move(wn));
            }
            if (gpu_blas) { blas_wide.emplace_back(); wide = -1; }
            zb_mesh zm{0, 0, first, uint32_t(m.tris.size())};
            out->meshes.push_back(zm);
            blas_nodes.push_back(std::move(nd));
            mi = mesh_ids[&m] = uint32_t(out->meshes.size()) - 1;
        } else mi = f->second;
        zb_item it{};
        it.material = resolve_material(ch, m.mat);
        it.leaf_id = leaf_id;
        it.kind_flags = ZB_ITEM_INSTANCE;
        it.aux = add_instance(ch, ZB_GEOM_MESH, mi);
        push_item(it, box);
    }

    void add_medium(const constant_medium& cm, uint32_t leaf_id, const chain& outer, const aabb& box) {
        if (outer.has_any) throw std::runtime_error("flatten: transform wrappers above constant_medium are unsupported");
        zb_medium z{};
        z.neg_inv_density = float(cm.neg_inv_density);
        // peel the boundary's own wrapper chain
        chain ch;
        const hittable* h = cm.boundary.get();
        for (;;) {
            using K = hittable::kind_t;
            K k = h->kind();
            if (k == K::translate || k == K::rotate_x || k == K::rotate_y || k == K::rotate_z || k == K::scale) {
                auto& w = static_cast<const transform_wrapper&>(*h);
                ch.w2o = w.world_to_object().after(ch.w2o);
                ch.o2w = ch.o2w.after(w.object_to_world());
                ch.force_front = ch.force_front || w.forces_front_face();
                ch.has_any = true;
                if (k != K::translate) ch.has_linear = true;
                h = w.ptr.get();
            } else if (k == K::material_instance) {
                h = static_cast<const material_instance&>(*h).object.get();
            } else break;
        }
        if (h->kind() == hittable::kind_t::sphere) {
            auto& s = static_cast<const sphere&>(*h);
            if (!ch.has_linear) {
                point3 c = ch.o2w.point(s.center);
                z.boundary_kind = ZB_ITEM_SPHERE;
                z.sphere[0] = float(c[0]); z.sphere[1] = float(c[1]); z.sphere[2] = float(c[2]); z.sphere[3] = float(s.radius);
            } else {
                uint32_t gi = uint32_t(out->spheres_obj.size() / 4);
                for (int k = 0; k < 3; k++) out->spheres_obj.push_back(float(s.center[k]));
                out->spheres_obj.push_back(float(s.radius));
                z.boundary_kind = ZB_ITEM_INSTANCE;
                z.boundary_index = add_instance(ch, ZB_GEOM_SPHERE, gi);
            }
        } else if (h->kind() == hittable::kind_t::cube) {
            auto& c = static_cast<const cube&>(*h);
            if (!ch.has_any) {
                z.boundary_kind = ZB_ITEM_CUBE;
                z.boundary_index = cube_index(c);
            } else {
                z.boundary_kind = ZB_ITEM_INSTANCE;
                z.boundary_index = add_instance(ch, ZB_GEOM_CUBE, cube_index(c));
            }
        } else {
            throw std::runtime_error("flatten: constant_medium boundary must be a sphere or cube (optionally transformed)");
        }
        out->media.push_back(z);
        zb_item it{};
        it.kind_flags = ZB_ITEM_MEDIUM;
        it.material = material_id(cm.phase_function);
        it.leaf_id = leaf_id;
        it.aux = uint32_t(out->media.size()) - 1;
        push_item(it, box);
    }
};

}  // namespace zenith
