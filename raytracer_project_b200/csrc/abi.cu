// abi.cu — implementation of the C ABI declared in include/zenith_b200.h (device side).
// Context = one GPU + one stream; scene = device copies of the flat SoA arrays; render = camera and
// environment set-up in double exactly as camera::initialize (camera.hpp:358-402) followed by the
// kernels of kernels.cu.  No CPU fallback: without a usable sm_100 device every entry point fails.
#include <chrono>
#include <cstdio>

#include "abi_internal.h"

extern "C" int zb_abi_version(void) { return ZB_ABI_VERSION; }
std::string& zb_err_slot() {
    static thread_local std::string err;
    return err;
}
extern "C" const char* zb_last_error(void) { return zb_err_slot().c_str(); }

extern "C" int zb_context_create(int device, zb_context** out) {
    if (!out) return fail(ZB_ERR_INVALID, "zb_context_create: out is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(ZB_ERR_CUDA, std::string("no CUDA device (this backend has no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= n) return fail(ZB_ERR_INVALID, "zb_context_create: device index out of range");
    ZB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    ZB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(ZB_ERR_CUDA, "device is not sm_100 class; kernels are built for sm_100a only");
    zb_context* c = new zb_context();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    c->cc_major = prop.major;
    c->cc_minor = prop.minor;
    c->l2_bytes = size_t(prop.l2CacheSize);
    c->tuning = zbk_tuning_from_env();
    ZB_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    c->own_stream = true;
    ZB_CUDA(cudaMalloc(&c->counters, 8 * sizeof(unsigned long long)));
    ZB_CUDA(cudaMallocHost(&c->counters_host, 8 * sizeof(unsigned long long)));
    for (auto& ev : c->ev) ZB_CUDA(cudaEventCreate(&ev));
    *out = c;
    return ZB_OK;
}

extern "C" int zb_context_set_stream(zb_context* ctx, void* stream) {
    if (!ctx) return fail(ZB_ERR_INVALID, "ctx is NULL");
    if (ctx->own_stream) {
        cudaStreamSynchronize(ctx->stream);
        cudaStreamDestroy(ctx->stream);
        ctx->own_stream = false;
    }
    ctx->stream = static_cast<cudaStream_t>(stream);
    return ZB_OK;
}

extern "C" int zb_build_flags(void) {
#ifdef ZB_EXPERIMENTS
    return ZB_BUILD_EXPERIMENTS;
#else
    return 0;
#endif
}

extern "C" int zb_context_set_tuning(zb_context* ctx, const char* key, long long value) {
    if (!ctx || !key) return fail(ZB_ERR_INVALID, "zb_context_set_tuning: NULL argument");
    const std::string k = key;
    ZbTuning& t = ctx->tuning;
    if (k == "pq") t.pq = int(value);
    else if (k == "no_lean") t.no_lean = value != 0;
    else if (k == "pq_blocks") t.pq_blocks = int(value);
    else if (k == "gather_p2p") t.gather_p2p = int(value);
    else if (k == "gpu_builder") t.gpu_builder = int(value);
    else if (k == "build_timing") t.build_timing = int(value);
    else if (k == "group_timing") t.group_timing = int(value);
    else if (k == "build_check") t.build_check = int(value);
    else if (k == "blas_isect_pct") t.blas_isect_pct = int(value);
    else if (k == "blas_leaf") t.blas_leaf = int(value);
    else if (k == "wide") t.wide = value != 0;
    else if (k == "fuse_finish") t.fuse_finish = value != 0;
#ifdef ZB_EXPERIMENTS
    else if (k == "pq_resume") t.pq_resume = int(value);
    else if (k == "pq_redeal") t.pq_redeal = int(value);
    else if (k == "wavefront") t.wavefront = value != 0;
    else if (k == "wf_slots") t.wf_slots = value;
#endif
    else return fail(ZB_ERR_INVALID, "zb_context_set_tuning: unknown key '" + k + "'");
    return ZB_OK;
}

extern "C" int zb_context_destroy(zb_context* ctx) {
    if (!ctx) return ZB_OK;
    if (ctx->group) {
        // the leader of a single-process group owns the other members
        std::vector<zb_context*> others;
        if (!ctx->group->multi_process)
            for (zb_context* m : ctx->group->local)
                if (m != ctx) others.push_back(m);
        for (zb_context* m : ctx->group->local) {
            cudaSetDevice(m->device);
            cudaStreamSynchronize(m->stream);
        }
        if (ctx->group->multi_process) zb_group_leave(ctx);
        else zbg_release(ctx);
        for (zb_context* m : others) zb_context_destroy(m);
    }
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaFree(ctx->counters);
    cudaFreeHost(ctx->counters_host);
    for (const auto& r : ctx->registered_out) cudaHostUnregister(r.ptr);
    cudaGetLastError();
    cudaFree(ctx->rgb_pack64);
    cudaFree(ctx->arena);
    ctx->build_scratch.release();
    ctx->build_tmp.release();
    cudaFree(ctx->accum);
    cudaFree(ctx->accum_aux);
    cudaFree(ctx->out_f4);
    cudaFree(ctx->id_scratch);
    cudaFree(ctx->fin_a);
    cudaFree(ctx->fin_b);
    cudaFree(ctx->fin_in);
    cudaFree(ctx->fin_stats);
#ifdef ZB_EXPERIMENTS
    zbk_wavefront_free(ctx->wf);
#endif
    cudaFree(ctx->pq_fix);
    cudaFree(ctx->rgb_pack);
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    for (auto& ev : ctx->ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->chunk_ev) cudaEventDestroy(ev);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return ZB_OK;
}

extern "C" int zb_device_info(zb_context* ctx, int* sm_count, int* cc_major, int* cc_minor, size_t* l2_bytes) {
    if (!ctx) return fail(ZB_ERR_INVALID, "ctx is NULL");
    if (sm_count) *sm_count = ctx->sm_count;
    if (cc_major) *cc_major = ctx->cc_major;
    if (cc_minor) *cc_minor = ctx->cc_minor;
    if (l2_bytes) *l2_bytes = ctx->l2_bytes;
    return ZB_OK;
}

// ------------------------------------------------------------------------------------ scene

// Scene arrays live in ONE device allocation (the arena): a render step that re-uploads its scene (bench e2e, GUI edits)
// pays one cudaMalloc / cudaFree at most — none when the context's cached arena fits (zb_context::arena) — instead of
// a dozen of each, which cost more than the copy of a small scene.
static size_t arena_round(size_t bytes) { return (bytes + 255) & ~size_t(255); }

template <class T>
static int upload(zb_scene* s, const void* host, size_t count, size_t elem, const T** dev) {
    *dev = nullptr;
    if (count == 0) return ZB_OK;
    if (!host) return fail(ZB_ERR_INVALID, "zb_scene_upload: array pointer is NULL but its count is not 0");
    const size_t bytes = count * elem;
    void* d = nullptr;
    if (s->arena && s->arena_used + arena_round(bytes) <= s->arena_bytes) {
        d = static_cast<char*>(s->arena) + s->arena_used;
        s->arena_used += arena_round(bytes);
    } else {
        ZB_CUDA(cudaMalloc(&d, bytes));
        s->allocs.push_back(d);
    }
    ZB_CUDA(cudaMemcpyAsync(d, host, bytes, cudaMemcpyHostToDevice, s->ctx->stream));
    s->bytes += bytes;
    *dev = static_cast<const T*>(d);
    return ZB_OK;
}

// take the context's cached arena if it is free and large enough, else allocate one of `bytes`
static int arena_acquire(zb_scene* s, size_t bytes) {
    zb_context* c = s->ctx;
    if (c->arena && !c->arena_in_use && c->arena_bytes >= bytes) {
        s->arena = c->arena;
        s->arena_bytes = c->arena_bytes;
        s->arena_from_ctx = true;
        c->arena_in_use = true;
        return ZB_OK;
    }
    ZB_CUDA(cudaMalloc(&s->arena, bytes));
    s->arena_bytes = bytes;
    if (!c->arena_in_use) {  // becomes the context's cached arena (the old, smaller one is released)
        if (c->arena) cudaFree(c->arena);
        c->arena = s->arena;
        c->arena_bytes = bytes;
        c->arena_in_use = true;
        s->arena_from_ctx = true;
    }
    return ZB_OK;
}

static int validate(const zb_flat_scene* f) {
    if (f->abi_version != ZB_ABI_VERSION) return fail(ZB_ERR_INVALID, "flat scene: ABI version mismatch");
    if (f->n_tlas_nodes > f->n_nodes) return fail(ZB_ERR_INVALID, "flat scene: n_tlas_nodes > n_nodes");
    if (f->n_items && !f->n_nodes) return fail(ZB_ERR_INVALID, "flat scene: items without BVH nodes");
    for (uint32_t i = 0; i < f->n_nodes; i++) {
        const int32_t ch[2] = {f->nodes[i].left, f->nodes[i].right};
        const bool tlas = i < f->n_tlas_nodes;
        for (int32_t c : ch) {
            if (c == ZB_CHILD_EMPTY) continue;
            if (c >= 0) {
                if (uint32_t(c) >= f->n_nodes) return fail(ZB_ERR_INVALID, "flat scene: node child index out of range");
            } else {
                uint32_t code = ~uint32_t(c), first = code >> 3, end = first + (code & 7u) + 1u;
                if (end > (tlas ? f->n_items : f->n_tris)) return fail(ZB_ERR_INVALID, "flat scene: leaf range out of bounds");
            }
        }
    }
    // tree depth (inner nodes on a root-to-leaf path) of the TLAS and of every host-built BLAS against the caps the
    // traversal stack is sized for; also rejects cycles (a node reached at a depth beyond the cap ends the walk)
    {
        std::vector<std::pair<uint32_t, int>> todo;
        auto depth_of = [&](uint32_t root, int cap) -> int {
            int deepest = 0;
            todo.clear();
            todo.push_back({root, 1});
            while (!todo.empty()) {
                auto [n, d] = todo.back();
                todo.pop_back();
                deepest = std::max(deepest, d);
                if (d > cap) return d;
                for (int32_t c : {f->nodes[n].left, f->nodes[n].right})
                    if (c >= 0) todo.push_back({uint32_t(c), d + 1});
            }
            return deepest;
        };
        if (f->n_tlas_nodes && depth_of(0, ZB_MAX_TLAS_DEPTH) > ZB_MAX_TLAS_DEPTH)
            return fail(ZB_ERR_INVALID, "flat scene: TLAS deeper than ZB_MAX_TLAS_DEPTH (or cyclic)");
        for (uint32_t m = 0; m < f->n_meshes; m++) {
            const zb_mesh& zm = f->meshes[m];
            if (zm.node_count == 0) continue;
            if (zm.node_root >= f->n_nodes) return fail(ZB_ERR_INVALID, "flat scene: mesh BLAS root out of range");
            if (depth_of(zm.node_root, ZB_MAX_BLAS_DEPTH) > ZB_MAX_BLAS_DEPTH)
                return fail(ZB_ERR_INVALID, "flat scene: mesh BLAS deeper than ZB_MAX_BLAS_DEPTH (or cyclic)");
        }
    }
    // the optional 8-wide tree: every index a traversal can form stays inside the arrays, and the stack need fits
    if (f->n_wide_nodes) {
        if (!f->wide_nodes) return fail(ZB_ERR_INVALID, "flat scene: wide_nodes is NULL but n_wide_nodes is not 0");
        if (f->wide_stack_need == 0 || f->wide_stack_need > 31) return fail(ZB_ERR_INVALID, "flat scene: 8-wide tree deeper than the traversal stack (wide_stack_need > 31)");
        for (uint32_t i = 0; i < f->n_wide_nodes; i++) {
            const zb_wide_node& w = f->wide_nodes[i];
            uint32_t n_inner = 0, max_prim = 0;
            for (int s2 = 0; s2 < 8; s2++) {
                const uint32_t cnt = w.meta[s2] >> 5, idx = w.meta[s2] & 31u;
                if (cnt == 0) continue;
                if (idx >= 24) {
                    if (cnt != 1 || idx != 24u + uint32_t(s2) || !((w.imask >> s2) & 1)) return fail(ZB_ERR_INVALID, "flat scene: malformed inner child in an 8-wide node");
                    n_inner++;
                } else {
                    if (cnt > 4 || idx + cnt > 24 || ((w.imask >> s2) & 1)) return fail(ZB_ERR_INVALID, "flat scene: malformed leaf child in an 8-wide node");
                    max_prim = std::max(max_prim, idx + cnt);
                }
            }
            if (uint32_t(__builtin_popcount(w.imask)) != n_inner) return fail(ZB_ERR_INVALID, "flat scene: 8-wide imask does not match its children");
            if (n_inner && (w.child_base <= i || uint64_t(w.child_base) + n_inner > f->n_wide_nodes))
                return fail(ZB_ERR_INVALID, "flat scene: 8-wide child range out of bounds (children must follow their parent)");
            if (uint64_t(w.prim_base) + max_prim > std::max(f->n_items, f->n_tris)) return fail(ZB_ERR_INVALID, "flat scene: 8-wide primitive range out of bounds");
            if (w.e[0] > 239 || w.e[1] > 239 || w.e[2] > 239) return fail(ZB_ERR_INVALID, "flat scene: 8-wide node exponent out of range");
        }
        for (uint32_t i = 0; i < f->n_instances; i++)
            if (f->instances[i].geom_kind == ZB_GEOM_MESH && f->instances[i].pad[1] >= f->n_wide_nodes)
                return fail(ZB_ERR_INVALID, "flat scene: instance 8-wide BLAS root out of range");
    }
    for (uint32_t i = 0; i < f->n_items; i++) {
        const zb_item& it = f->items[i];
        uint32_t kind = it.kind_flags & 0xFFu;
        if (it.material >= f->n_materials) return fail(ZB_ERR_INVALID, "flat scene: item material out of range");
        uint32_t lim = kind == ZB_ITEM_SPHERE ? 0xFFFFFFFFu
                       : kind == ZB_ITEM_CUBE ? f->n_cubes
                       : kind == ZB_ITEM_TRIANGLE ? f->n_tris
                       : kind == ZB_ITEM_INSTANCE ? f->n_instances
                       : kind == ZB_ITEM_MEDIUM ? f->n_media : 0u;
        if (lim == 0u || (kind != ZB_ITEM_SPHERE && it.aux >= lim)) return fail(ZB_ERR_INVALID, "flat scene: item reference out of range");
    }
    for (uint32_t i = 0; i < f->n_instances; i++) {
        const zb_instance& in = f->instances[i];
        uint32_t lim = in.geom_kind == ZB_GEOM_SPHERE ? f->n_spheres_obj
                       : in.geom_kind == ZB_GEOM_CUBE ? f->n_cubes
                       : in.geom_kind == ZB_GEOM_MESH ? f->n_meshes
                       : in.geom_kind == ZB_GEOM_TRIANGLE ? f->n_tris : 0u;
        if (in.geom_index >= lim) return fail(ZB_ERR_INVALID, "flat scene: instance geometry out of range");
        if (in.geom_kind == ZB_GEOM_MESH && in.pad[0] != f->meshes[in.geom_index].node_root)
            return fail(ZB_ERR_INVALID, "flat scene: instance BLAS root does not match its mesh");
    }
    for (uint32_t i = 0; i < f->n_materials; i++) {
        const zb_material& m = f->materials[i];
        if (m.type > ZB_MAT_ISOTROPIC) return fail(ZB_ERR_INVALID, "flat scene: unknown material type");
        if ((m.albedo_tex != ZB_NO_TEXTURE && m.albedo_tex >= f->n_textures) ||
            (m.bump_tex != ZB_NO_TEXTURE && m.bump_tex >= f->n_textures))
            return fail(ZB_ERR_INVALID, "flat scene: material texture out of range");
    }
    for (uint32_t i = 0; i < f->n_textures; i++) {
        const zb_texture& t = f->textures[i];
        if (t.kind == ZB_TEX_CHECKER && (t.a >= f->n_textures || t.b >= f->n_textures))
            return fail(ZB_ERR_INVALID, "flat scene: checker sub-texture out of range");
        if (t.kind == ZB_TEX_IMAGE_U8 && uint64_t(t.c) + uint64_t(t.a) * t.b > f->n_texels_u8)
            return fail(ZB_ERR_INVALID, "flat scene: u8 texture outside the texel pool");
        if (t.kind == ZB_TEX_IMAGE_F32 && uint64_t(t.c) + uint64_t(t.a) * t.b > f->n_texels_f32)
            return fail(ZB_ERR_INVALID, "flat scene: f32 texture outside the texel pool");
    }
    return ZB_OK;
}

// csrc/lbvh.cu
int zbk_build_blas_device(zb_context* ctx, const float4* tris_in, const float4* nrm_in, uint32_t n, uint32_t tri_first,
                          uint32_t node_base, float4* tris_out, float4* nrm_out, zb_bvh_node* nodes_out, uint32_t* n_nodes);
// csrc/sah_build.cu
int zbk_build_blas_device_sah(zb_context* ctx, const float4* tris_in, const float4* nrm_in, uint32_t n, uint32_t tri_first,
                              uint32_t node_base, float4* tris_out, float4* nrm_out, zb_bvh_node* nodes_out, uint32_t* n_nodes);

// Meshes the flattener left without a BLAS (node_count == 0): build linear BVHs on the device.
// d.tris / d.tri_normals already hold the triangles in face order; nodes are appended after the host nodes and
// the BLAS roots are written into `instances` (host copy, uploaded by the caller afterwards).
static int build_device_blas(zb_scene* s, const zb_flat_scene* f, std::vector<zb_instance>& instances) {
    SceneDev& d = s->dev;
    size_t extra = 0, largest = 0;
    for (uint32_t m = 0; m < f->n_meshes; m++)
        if (f->meshes[m].node_count == 0 && f->meshes[m].tri_count > 0) {
            extra += std::max<size_t>(f->meshes[m].tri_count - 1, 1);
            largest = std::max<size_t>(largest, f->meshes[m].tri_count);
        }
    zb_bvh_node* nodes = nullptr;
    ZB_CUDA(cudaMalloc(&nodes, (size_t(f->n_nodes) + extra) * sizeof(zb_bvh_node)));
    s->allocs.push_back(nodes);
    if (f->n_nodes)
        ZB_CUDA(cudaMemcpyAsync(nodes, f->nodes, size_t(f->n_nodes) * sizeof(zb_bvh_node), cudaMemcpyHostToDevice, s->ctx->stream));
    float4 *tmp_t = nullptr, *tmp_n = nullptr;
    {
        void* tmp = nullptr;
        if (s->ctx->build_tmp.get(2 * largest * 3 * sizeof(float4), &tmp) != cudaSuccess) {
            cudaGetLastError();
            return fail(ZB_ERR_CUDA, "zb_scene_upload: out of device memory for the BLAS build");
        }
        tmp_t = static_cast<float4*>(tmp);
        tmp_n = tmp_t + largest * 3;
    }
    uint32_t next = f->n_nodes;
    std::vector<uint32_t> roots(f->n_meshes, 0);
    int rc = ZB_OK;
    for (uint32_t m = 0; m < f->n_meshes && !rc; m++) {
        const zb_mesh& zm = f->meshes[m];
        roots[m] = zm.node_root;
        if (zm.node_count != 0 || zm.tri_count == 0) continue;
        float4* t = const_cast<float4*>(d.tris) + 3 * size_t(zm.tri_first);
        float4* n = const_cast<float4*>(d.tri_normals) + 3 * size_t(zm.tri_first);
        const size_t bytes = size_t(zm.tri_count) * 3 * sizeof(float4);
        cudaMemcpyAsync(tmp_t, t, bytes, cudaMemcpyDeviceToDevice, s->ctx->stream);
        cudaMemcpyAsync(tmp_n, n, bytes, cudaMemcpyDeviceToDevice, s->ctx->stream);
        uint32_t count = 0;
        const bool timing = s->ctx->tuning.build_timing != 0;
        std::chrono::steady_clock::time_point t0;
        if (timing) {
            cudaStreamSynchronize(s->ctx->stream);
            t0 = std::chrono::steady_clock::now();
        }
        rc = (s->ctx->tuning.gpu_builder ? zbk_build_blas_device_sah : zbk_build_blas_device)(s->ctx, tmp_t, tmp_n, zm.tri_count, zm.tri_first, next, t,
                                                                                             n, nodes + next, &count);
        if (timing) {
            cudaStreamSynchronize(s->ctx->stream);
            std::fprintf(stderr, "[zb build] mesh %u: %u triangles, %s: %.3f ms wall (scratch allocation included), %u nodes\n", m, zm.tri_count,
                         s->ctx->tuning.gpu_builder ? "binned SAH" : "linear BVH",
                         std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(), count);
        }
        roots[m] = next;
        next += count;
    }
    // root nodes of the device-built trees (still lo / hi boxes): zb_scene_refit needs every instance's object-space bounds
    s->device_mesh_roots.assign(f->n_meshes, zb_bvh_node{});
    for (uint32_t m = 0; m < f->n_meshes && !rc; m++)
        if (f->meshes[m].node_count == 0 && f->meshes[m].tri_count > 0)
            cudaMemcpyAsync(&s->device_mesh_roots[m], nodes + roots[m], sizeof(zb_bvh_node), cudaMemcpyDeviceToHost, s->ctx->stream);
    cudaStreamSynchronize(s->ctx->stream);
    if (rc) return rc;
    for (auto& in : instances)
        if (in.geom_kind == ZB_GEOM_MESH) in.pad[0] = roots[in.geom_index];
    d.nodes = reinterpret_cast<const float4*>(nodes);
    d.n_nodes = next;
    s->bytes += size_t(next) * sizeof(zb_bvh_node);
    return ZB_OK;
}

static int upload_one(zb_context* ctx, const zb_flat_scene* f, zb_scene** out);

extern "C" int zb_context_register_output(zb_context* ctx, void* ptr, size_t bytes) {
    if (!ctx || !ptr || !bytes) return fail(ZB_ERR_INVALID, "zb_context_register_output: NULL argument");
    if (ctx->out_registered(ptr, bytes)) return ZB_OK;
    if (ctx->registered_out.size() >= 8) return fail(ZB_ERR_UNSUPPORTED, "zb_context_register_output: at most 8 ranges per context");
    ZB_CUDA(cudaSetDevice(ctx->device));
    const cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(ZB_ERR_UNSUPPORTED, std::string("zb_context_register_output: the range cannot be page-locked (") + cudaGetErrorString(e) +
                                            "); zb_render keeps using its staged copy");
    }
    ctx->registered_out.push_back({static_cast<char*>(ptr), bytes});
    return ZB_OK;
}

extern "C" int zb_context_unregister_output(zb_context* ctx, void* ptr) {
    if (!ctx) return fail(ZB_ERR_INVALID, "zb_context_unregister_output: NULL context");
    for (size_t i = 0; i < ctx->registered_out.size(); i++)
        if (ctx->registered_out[i].ptr == static_cast<char*>(ptr)) {
            cudaSetDevice(ctx->device);
            cudaStreamSynchronize(ctx->stream);
            cudaHostUnregister(ptr);
            cudaGetLastError();
            ctx->registered_out.erase(ctx->registered_out.begin() + long(i));
            return ZB_OK;
        }
    return ZB_OK;  // not registered (or its registration had failed): nothing to do
}

extern "C" int zb_scene_upload(zb_context* ctx, const zb_flat_scene* f, zb_scene** out) {
    if (!ctx || !f || !out) return fail(ZB_ERR_INVALID, "zb_scene_upload: NULL argument");
    int rc = validate(f);
    if (rc) return rc;
    zb_scene* s = nullptr;
    rc = upload_one(ctx, f, &s);
    if (rc) return rc;
    // a multi-GPU context keeps a replica on every member (the scene is read-only and small next to 180 GB of HBM)
    if (ctx->group && !ctx->group->multi_process && ctx->group->n > 1 && ctx->group->local[0] == ctx) {
        for (zb_context* m : ctx->group->local) {
            if (m == ctx) continue;
            zb_scene* rep = nullptr;
            rc = upload_one(m, f, &rep);
            if (rc) {
                zb_scene_destroy(s);
                return rc;
            }
            s->replicas.push_back(rep);
        }
        cudaSetDevice(ctx->device);
    }
    *out = s;
    return ZB_OK;
}

static int upload_one(zb_context* ctx, const zb_flat_scene* f, zb_scene** out) {
    int rc = ZB_OK;
    ZB_CUDA(cudaSetDevice(ctx->device));
    zb_scene* s = new zb_scene();
    s->ctx = ctx;
    SceneDev& d = s->dev;
    bool device_blas = false;
    for (uint32_t m = 0; m < f->n_meshes; m++) device_blas |= f->meshes[m].node_count == 0 && f->meshes[m].tri_count > 0;
    {
        size_t need = 0;
        if (!device_blas) need += arena_round(size_t(f->n_nodes) * sizeof(zb_bvh_node)) + arena_round(size_t(f->n_wide_nodes) * sizeof(zb_wide_node));
        need += arena_round(size_t(f->n_items) * sizeof(zb_item)) + arena_round(size_t(f->n_cubes) * sizeof(zb_cube)) +
                arena_round(size_t(f->n_tris) * sizeof(zb_tri)) + arena_round(size_t(f->n_tris) * sizeof(zb_tri_normals)) +
                arena_round(size_t(f->n_instances) * sizeof(zb_instance)) + arena_round(size_t(f->n_media) * sizeof(zb_medium)) +
                arena_round(size_t(f->n_materials) * sizeof(zb_material)) + arena_round(size_t(f->n_textures) * sizeof(zb_texture)) +
                arena_round(size_t(f->n_spheres_obj) * 16) + arena_round(size_t(f->n_texels_u8) * 4) + arena_round(size_t(f->n_texels_f32) * 16);
        if (need) rc = arena_acquire(s, need);
        if (rc) {
            delete s;
            return rc;
        }
    }
    if (!device_blas) rc = upload(s, f->nodes, f->n_nodes, sizeof(zb_bvh_node), &d.nodes);
    if (!rc && !device_blas) rc = upload(s, f->wide_nodes, f->n_wide_nodes, sizeof(zb_wide_node), &s->wide);
    if (!rc) rc = upload(s, f->items, f->n_items, sizeof(zb_item), &d.items);
    if (!rc) rc = upload(s, f->cubes, f->n_cubes, sizeof(zb_cube), &d.cubes);
    if (!rc) rc = upload(s, f->tris, f->n_tris, sizeof(zb_tri), &d.tris);
    if (!rc) rc = upload(s, f->tri_normals, f->n_tris, sizeof(zb_tri_normals), &d.tri_normals);
    std::vector<zb_instance> patched;
    const zb_instance* instances = f->instances;
    d.n_nodes = f->n_nodes;
    if (!rc && device_blas) {
        patched.assign(f->instances, f->instances + f->n_instances);
        rc = build_device_blas(s, f, patched);
        instances = patched.data();
    }
    if (!rc) rc = upload(s, instances, f->n_instances, sizeof(zb_instance), &d.instances);
    if (!rc) rc = upload(s, f->media, f->n_media, sizeof(zb_medium), &d.media);
    if (!rc) rc = upload(s, f->materials, f->n_materials, sizeof(zb_material), &d.materials);
    if (!rc) rc = upload(s, f->textures, f->n_textures, sizeof(zb_texture), &d.textures);
    if (!rc) rc = upload(s, f->spheres_obj, f->n_spheres_obj, 16, &d.spheres_obj);
    if (!rc) rc = upload(s, f->texels_u8, f->n_texels_u8, 4, &d.texels_u8);
    if (!rc) rc = upload(s, f->texels_f32, f->n_texels_f32, 16, &d.texels_f32);
    if (rc) {
        zb_scene_destroy(s);
        return rc;
    }
    if (!rc) {
        cudaError_t e = zbk_convert_nodes(const_cast<float4*>(d.nodes), d.n_nodes, ctx->stream);
        if (e != cudaSuccess) rc = fail(ZB_ERR_CUDA, std::string("zb_scene_upload: node re-encoding: ") + cudaGetErrorString(e));
    }
    if (rc) {
        zb_scene_destroy(s);
        return rc;
    }
    rc = zbr_prepare_refit(s, f, instances);
    if (rc) {
        zb_scene_destroy(s);
        return rc;
    }
    d.n_items = f->n_items;
    d.n_media = f->n_media;
    d.overflow = ctx->counters + 6;
    d.wide = s->wide;  // fill_params drops it when tuning wide = 0
    // scene features (ZB_F_*): which code a render kernel has to carry for this scene
    d.features = 0;
    zb_flat_scene_features(f, &d.features);
    s->env_hdr_texture = f->pad;
    ZB_CUDA(cudaStreamSynchronize(ctx->stream));  // host arrays may be freed by the caller on return
    *out = s;
    return ZB_OK;
}

// Page-lock (or release) the large arrays of a host flat scene so that zb_scene_upload's copies run at PCIe speed instead
// of through the driver's staging of pageable memory (measured on the 1.28 GB scene of config C5: 264-342 ms per upload
// from pageable memory).  Called by the host glue after a scene is flattened or loaded and before it is freed; arrays
// under 8 MiB are left alone; failures are ignored (no CUDA device, memory already registered).
extern "C" void zbi_pin_flat_scene(const zb_flat_scene* f, int pin) {
    if (!f) return;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) { cudaGetLastError(); return; }
    const std::pair<const void*, size_t> arrays[] = {
        {f->nodes, size_t(f->n_nodes) * sizeof(zb_bvh_node)},       {f->tris, size_t(f->n_tris) * sizeof(zb_tri)},
        {f->tri_normals, size_t(f->n_tris) * sizeof(zb_tri_normals)}, {f->texels_u8, size_t(f->n_texels_u8) * 4},
        {f->texels_f32, size_t(f->n_texels_f32) * 16},               {f->wide_nodes, size_t(f->n_wide_nodes) * sizeof(zb_wide_node)}};
    for (const auto& a : arrays) {
        if (!a.first || a.second < (size_t(8) << 20)) continue;
        if (pin) cudaHostRegister(const_cast<void*>(a.first), a.second, cudaHostRegisterPortable | cudaHostRegisterReadOnly);
        else cudaHostUnregister(const_cast<void*>(a.first));
        cudaGetLastError();
    }
}

extern "C" int zb_scene_destroy(zb_scene* s) {
    if (!s) return ZB_OK;
    for (zb_scene* rep : s->replicas) zb_scene_destroy(rep);
    s->replicas.clear();
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
    for (void* p : s->allocs) cudaFree(p);
    if (s->arena) {
        if (s->arena_from_ctx && s->ctx->arena == s->arena) s->ctx->arena_in_use = false;  // stays cached for the next upload
        else cudaFree(s->arena);
    }
    delete s;
    return ZB_OK;
}
