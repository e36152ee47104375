// refit.cu — zb_scene_refit: transform-only scene edits without re-flattening.  The reference rebuilds
// its geometry and its whole pointer BVH on every GUI restart (main.cpp:1492-1500, bvh.hpp:11-44); here a moved, rotated or
// rescaled instance costs one small upload and ONE kernel: new world boxes of the items that reference instances (eight
// corners of the geometry's object-space bounds through the new object->world matrix), then the TLAS boxes bottom-up with
// the tree's topology unchanged.  BLASes live in object space and are untouched.
//
// k_refit_tlas runs as ONE block (a TLAS has hundreds to a few thousand nodes): phase 1 over the items, then one phase per
// tree level from the deepest up, __syncthreads between phases.  Nodes were emitted in DFS order by the host builder; their
// levels are computed once at upload (zbr_prepare_refit).  HBM traffic is a few hundred KB: launch-latency-bound, ~20 us.
#include <algorithm>

#include "abi_internal.h"

namespace {

struct RefitArgs {
    float4* nodes;             // TLAS nodes (first n_tlas of the scene's node array), centre / half-extent or lo / hi encoding
    const float4* items;       // 2 x float4 per item
    const float4* instances;   // 8 x float4 per instance (already the new table)
    const float4* media;       // 2 x float4 per medium
    const float4* inst_bounds; // 2 x float4 per instance: object-space lo, hi of its geometry
    float4* item_bounds;       // 2 x float4 per item: world lo, hi (rewritten for instance items)
    float4* node_bounds;       // scratch, 2 x float4 per TLAS node
    const uint32_t* order;     // TLAS node indices grouped by level, deepest level first
    uint32_t level_start[ZB_MAX_TLAS_DEPTH + 3];  // order[level_start[k] .. level_start[k + 1]) = k-th group
    uint32_t n_levels, n_items;
    int center_half;           // node encoding (ZB_BOX_CH)
};

__device__ void grow(float3& lo, float3& hi, float3 a, float3 b) {
    lo = make_float3(fminf(lo.x, a.x), fminf(lo.y, a.y), fminf(lo.z, a.z));
    hi = make_float3(fmaxf(hi.x, b.x), fmaxf(hi.y, b.y), fmaxf(hi.z, b.z));
}

// (lo, hi) of one axis -> the node's stored pair, rounded outwards (as k_convert_nodes does at upload)
__device__ float2 encode_axis(float lo, float hi, int center_half) {
    if (!center_half) return make_float2(lo, hi);
    if (!(lo <= hi)) return make_float2(0.f, -INFINITY);
    const float c = 0.5f * lo + 0.5f * hi;
    return make_float2(c, fmaxf(__fsub_ru(hi, c), __fsub_ru(c, lo)));
}

// world box of item i if its geometry sits under an instance transform (other items keep the flattener's box)
__device__ void item_world_bounds(const RefitArgs& A, uint32_t i) {
        const float4 mf = A.items[2 * size_t(i) + 1];
        const uint32_t kind = __float_as_uint(mf.x) & 0xFFu, aux = __float_as_uint(mf.w);
        uint32_t inst = 0xFFFFFFFFu;
        if (kind == ZB_ITEM_INSTANCE) inst = aux;
        else if (kind == ZB_ITEM_MEDIUM) {
            const float4 m0 = A.media[2 * size_t(aux)];
            if (__float_as_uint(m0.x) == ZB_ITEM_INSTANCE) inst = __float_as_uint(m0.y);
        }
        if (inst == 0xFFFFFFFFu) return;
        const float4* ip = A.instances + 8 * size_t(inst);
        const float4 r0 = ip[3], r1 = ip[4], r2 = ip[5];  // object -> world, row-major 3x4
        const float4 blo = A.inst_bounds[2 * size_t(inst)], bhi = A.inst_bounds[2 * size_t(inst) + 1];
        float3 lo = make_float3(INFINITY, INFINITY, INFINITY), hi = make_float3(-INFINITY, -INFINITY, -INFINITY);
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const float x = (c & 1) ? bhi.x : blo.x, y = (c & 2) ? bhi.y : blo.y, z = (c & 4) ? bhi.z : blo.z;
            const float3 w = make_float3(fmaf(r0.x, x, fmaf(r0.y, y, fmaf(r0.z, z, r0.w))), fmaf(r1.x, x, fmaf(r1.y, y, fmaf(r1.z, z, r1.w))),
                                         fmaf(r2.x, x, fmaf(r2.y, y, fmaf(r2.z, z, r2.w))));
            grow(lo, hi, w, w);
        }
        // a few ulp of slack for the fp32 matrix product (the flattener's boxes come from double arithmetic)
        const float3 pad = make_float3(1e-6f * fmaxf(fabsf(lo.x), fabsf(hi.x)) + 1e-7f, 1e-6f * fmaxf(fabsf(lo.y), fabsf(hi.y)) + 1e-7f,
                                       1e-6f * fmaxf(fabsf(lo.z), fabsf(hi.z)) + 1e-7f);
        A.item_bounds[2 * size_t(i)] = make_float4(lo.x - pad.x, lo.y - pad.y, lo.z - pad.z, 0.f);
        A.item_bounds[2 * size_t(i) + 1] = make_float4(hi.x + pad.x, hi.y + pad.y, hi.z + pad.z, 0.f);
}

// the same as a kernel of its own: zb_scene_rebuild_tlas needs the boxes before it builds a new tree
__global__ void __launch_bounds__(256) k_item_bounds(const __grid_constant__ RefitArgs A) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < A.n_items; i += gridDim.x * blockDim.x) item_world_bounds(A, i);
}

__global__ void __launch_bounds__(1024) k_refit_tlas(const __grid_constant__ RefitArgs A) {
    // ---- phase 1: world boxes of the items whose geometry sits under an instance transform
    for (uint32_t i = threadIdx.x; i < A.n_items; i += blockDim.x) item_world_bounds(A, i);
    __syncthreads();
    // ---- phase 2: node boxes, deepest level first
    for (uint32_t g = 0; g < A.n_levels; g++) {
        for (uint32_t k = A.level_start[g] + threadIdx.x; k < A.level_start[g + 1]; k += blockDim.x) {
            const uint32_t node = A.order[k];
            float4* np = A.nodes + 4 * size_t(node);
            const float4 codes = np[3];
            const int child[2] = {__float_as_int(codes.x), __float_as_int(codes.y)};
            float3 lo[2], hi[2];
#pragma unroll
            for (int s = 0; s < 2; s++) {
                lo[s] = make_float3(INFINITY, INFINITY, INFINITY);
                hi[s] = make_float3(-INFINITY, -INFINITY, -INFINITY);
                if (child[s] == ZB_CHILD_EMPTY) continue;
                if (child[s] >= 0) {
                    const float4 a = A.node_bounds[2 * size_t(child[s])], b = A.node_bounds[2 * size_t(child[s]) + 1];
                    lo[s] = make_float3(a.x, a.y, a.z);
                    hi[s] = make_float3(b.x, b.y, b.z);
                } else {
                    const uint32_t code = ~uint32_t(child[s]), first = code >> 3, end = first + (code & 7u) + 1u;
                    for (uint32_t i = first; i < end; i++) {
                        const float4 a = A.item_bounds[2 * size_t(i)], b = A.item_bounds[2 * size_t(i) + 1];
                        grow(lo[s], hi[s], make_float3(a.x, a.y, a.z), make_float3(b.x, b.y, b.z));
                    }
                }
            }
            const float2 lx = encode_axis(lo[0].x, hi[0].x, A.center_half), ly = encode_axis(lo[0].y, hi[0].y, A.center_half),
                         lz = encode_axis(lo[0].z, hi[0].z, A.center_half), rx = encode_axis(lo[1].x, hi[1].x, A.center_half),
                         ry = encode_axis(lo[1].y, hi[1].y, A.center_half), rz = encode_axis(lo[1].z, hi[1].z, A.center_half);
            np[0] = make_float4(lx.x, lx.y, ly.x, ly.y);
            np[1] = make_float4(rx.x, rx.y, ry.x, ry.y);
            np[2] = make_float4(lz.x, lz.y, rz.x, rz.y);
            float3 nlo = lo[0], nhi = hi[0];
            grow(nlo, nhi, lo[1], hi[1]);
            A.node_bounds[2 * size_t(node)] = make_float4(nlo.x, nlo.y, nlo.z, 0.f);
            A.node_bounds[2 * size_t(node) + 1] = make_float4(nhi.x, nhi.y, nhi.z, 0.f);
        }
        __syncthreads();
    }
}

}  // namespace

// TLAS nodes grouped by level, deepest level first (order[level_start[k] .. level_start[k + 1]) = k-th group); false when
// the tree is deeper than RefitArgs::level_start has room for
static bool tlas_levels(const zb_bvh_node* nodes, uint32_t n_tlas, std::vector<uint32_t>& order, std::vector<uint32_t>& level_start) {
    std::vector<int> level(n_tlas, -1);
    std::vector<uint32_t> todo{0u};
    level[0] = 0;
    int deepest = 0;
    while (!todo.empty()) {
        const uint32_t n = todo.back();
        todo.pop_back();
        for (int32_t c : {nodes[n].left, nodes[n].right})
            if (c >= 0 && uint32_t(c) < n_tlas && level[c] < 0) {
                level[c] = level[n] + 1;
                deepest = std::max(deepest, level[c]);
                todo.push_back(uint32_t(c));
            }
    }
    if (deepest + 1 > ZB_MAX_TLAS_DEPTH + 2) return false;
    order.clear();
    level_start.clear();
    for (int l = deepest; l >= 0; l--) {
        level_start.push_back(uint32_t(order.size()));
        for (uint32_t n = 0; n < n_tlas; n++)
            if (level[n] == l) order.push_back(n);
    }
    level_start.push_back(uint32_t(order.size()));
    return true;
}

// Called by zb_scene_upload after the arrays are on the device: everything a later refit needs.  Leaves s->refit empty
// (refit unsupported for this scene) when the flat scene carries no item bounds or a mesh without a host BLAS.
int zbr_prepare_refit(zb_scene* s, const zb_flat_scene* f, const zb_instance* instances) {
    s->refit = zb_scene::refit_state{};
    if (!f->item_bounds || !f->n_tlas_nodes || !f->n_items) return ZB_OK;
    zb_scene::refit_state R;
    // object-space bounds of every instance's geometry
    std::vector<float> ib(size_t(f->n_instances) * 8, 0.f);
    for (uint32_t i = 0; i < f->n_instances; i++) {
        const zb_instance& in = instances[i];
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        auto add = [&](const float* a, const float* b) {
            for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], a[k]); hi[k] = std::max(hi[k], b[k]); }
        };
        if (in.geom_kind == ZB_GEOM_SPHERE) {
            const float* sp = f->spheres_obj + 4 * size_t(in.geom_index);
            const float a[3] = {sp[0] - sp[3], sp[1] - sp[3], sp[2] - sp[3]}, b[3] = {sp[0] + sp[3], sp[1] + sp[3], sp[2] + sp[3]};
            add(a, b);
        } else if (in.geom_kind == ZB_GEOM_CUBE) {
            const zb_cube& c = f->cubes[in.geom_index];  // cube::bounding_box = centre -+ half (cube.hpp:34-41)
            const float a[3] = {c.centre[0] - c.half[0], c.centre[1] - c.half[1], c.centre[2] - c.half[2]},
                        b[3] = {c.centre[0] + c.half[0], c.centre[1] + c.half[1], c.centre[2] + c.half[2]};
            add(a, b);
            // the hit test uses slabs about the ORIGIN of the local frame (cube.hpp:44-70, SURVEY A.q4): cover those too
            const float a2[3] = {-c.half[0], -c.half[1], -c.half[2]}, b2[3] = {c.half[0], c.half[1], c.half[2]};
            add(a2, b2);
        } else if (in.geom_kind == ZB_GEOM_TRIANGLE) {
            const zb_tri& t = f->tris[in.geom_index];
            add(t.v0, t.v0); add(t.v1, t.v1); add(t.v2, t.v2);
            for (int k = 0; k < 3; k++) { lo[k] -= 1e-4f; hi[k] += 1e-4f; }  // triangle::bounding_box pads thin axes (triangle.hpp:84-101)
        } else if (in.geom_kind == ZB_GEOM_MESH) {
            const zb_mesh& m = f->meshes[in.geom_index];
            if (m.node_count == 0 && in.geom_index >= s->device_mesh_roots.size()) return ZB_OK;
            const zb_bvh_node& r = m.node_count == 0 ? s->device_mesh_roots[in.geom_index]  // built on the device at upload
                                                     : f->nodes[m.node_root];
            const float la[3] = {r.lx_ly[0], r.lx_ly[2], r.lz_rz[0]}, lb[3] = {r.lx_ly[1], r.lx_ly[3], r.lz_rz[1]};
            const float ra[3] = {r.rx_ry[0], r.rx_ry[2], r.lz_rz[2]}, rb[3] = {r.rx_ry[1], r.rx_ry[3], r.lz_rz[3]};
            if (r.left != ZB_CHILD_EMPTY) add(la, lb);
            if (r.right != ZB_CHILD_EMPTY) add(ra, rb);
        }
        for (int k = 0; k < 3; k++) { ib[8 * size_t(i) + k] = lo[k]; ib[8 * size_t(i) + 4 + k] = hi[k]; }
    }
    // TLAS nodes grouped by level, deepest first
    std::vector<uint32_t> order;
    if (!tlas_levels(f->nodes, f->n_tlas_nodes, order, R.level_start)) return ZB_OK;
    // nodes [0, capacity) belong to the TLAS: the flattener pads the region to n_items - 1 so that any tree over these
    // items fits (zb_scene_rebuild_tlas); an older flat scene has capacity = its node count
    uint32_t capacity = f->n_nodes;
    for (uint32_t m = 0; m < f->n_meshes; m++)
        if (f->meshes[m].node_count) capacity = std::min(capacity, f->meshes[m].node_root);
    capacity = std::max(capacity, f->n_tlas_nodes);
    R.tlas_capacity = capacity;
    order.resize(capacity, 0u);  // the device array has room for the largest tree
    cudaStream_t st = s->ctx->stream;
    auto put = [&](const void* host, size_t bytes, void** dev) -> int {
        ZB_CUDA(cudaMalloc(dev, std::max<size_t>(bytes, 16)));
        s->allocs.push_back(*dev);
        if (bytes) ZB_CUDA(cudaMemcpyAsync(*dev, host, bytes, cudaMemcpyHostToDevice, st));
        return ZB_OK;
    };
    int rc = put(ib.data(), ib.size() * 4, reinterpret_cast<void**>(&R.inst_bounds));
    if (!rc) rc = put(f->item_bounds, size_t(f->n_items) * 32, reinterpret_cast<void**>(&R.item_bounds));
    if (!rc) rc = put(order.data(), order.size() * 4, reinterpret_cast<void**>(&R.order));
    if (rc) return rc;
    {
        void* nb = nullptr;
        ZB_CUDA(cudaMalloc(&nb, size_t(capacity) * 32));
        s->allocs.push_back(nb);
        R.node_bounds = static_cast<float4*>(nb);
    }
    ZB_CUDA(cudaStreamSynchronize(st));  // ib / order are locals
    R.host_instances.assign(instances, instances + f->n_instances);
    R.n_tlas_nodes = f->n_tlas_nodes;
    R.n_items = f->n_items;
    R.ready = true;
    s->refit = std::move(R);
    return ZB_OK;
}

// csrc/sah_build.cu
int zbk_build_tlas_device(zb_context* ctx, float4* items, float4* item_bounds, uint32_t n, zb_bvh_node* nodes_out, uint32_t max_nodes,
                          uint32_t* n_nodes);

// rebuild = false: boxes of the uploaded topology refitted bottom-up (one kernel).  true: new item boxes, then a new TLAS
// built over them on the device (binned SAH, sah_build.cu) in the same node region; items are permuted into its leaf order.
static int refit_one(zb_scene* s, const zb_instance* instances, uint32_t n, bool rebuild) {
    zb_scene::refit_state& R = s->refit;
    const char* who = rebuild ? "zb_scene_rebuild_tlas" : "zb_scene_refit";
    if (!R.ready) return fail(ZB_ERR_UNSUPPORTED, std::string(who) + ": the scene was uploaded without item bounds");
    if (n != R.host_instances.size()) return fail(ZB_ERR_INVALID, std::string(who) + ": instance count differs from the uploaded scene");
    std::vector<zb_instance> next(instances, instances + n);
    for (uint32_t i = 0; i < n; i++) {
        const zb_instance& old = R.host_instances[i];
        if (next[i].geom_kind != old.geom_kind || next[i].geom_index != old.geom_index)
            return fail(ZB_ERR_INVALID, std::string(who) + ": instance " + std::to_string(i) + " references other geometry (transform-only edits; re-flatten for anything else)");
        next[i].pad[0] = old.pad[0];  // BLAS roots stay the uploaded ones
        next[i].pad[1] = old.pad[1];
    }
    zb_context* ctx = s->ctx;
    ZB_CUDA(cudaSetDevice(ctx->device));
    if (n) ZB_CUDA(cudaMemcpyAsync(const_cast<float4*>(s->dev.instances), next.data(), size_t(n) * sizeof(zb_instance), cudaMemcpyHostToDevice, ctx->stream));
    RefitArgs A{};
    A.nodes = const_cast<float4*>(s->dev.nodes);
    A.items = s->dev.items;
    A.instances = s->dev.instances;
    A.media = s->dev.media;
    A.inst_bounds = R.inst_bounds;
    A.item_bounds = R.item_bounds;
    A.node_bounds = R.node_bounds;
    A.order = R.order;
    A.n_items = R.n_items;
    A.center_half = zbk_nodes_center_half() ? 1 : 0;
    if (!rebuild) {
        A.n_levels = uint32_t(R.level_start.size()) - 1;
        for (size_t k = 0; k < R.level_start.size(); k++) A.level_start[k] = R.level_start[k];
        k_refit_tlas<<<1, 1024, 0, ctx->stream>>>(A);
        ZB_CUDA(cudaGetLastError());
        ZB_CUDA(cudaStreamSynchronize(ctx->stream));  // `next` is a local; the caller may render at once
        ctx->stats.launches++;
    } else {
        k_item_bounds<<<std::max(1u, std::min((R.n_items + 255) / 256, 1024u)), 256, 0, ctx->stream>>>(A);
        ZB_CUDA(cudaGetLastError());
        ctx->stats.launches++;
        zb_bvh_node* tlas = reinterpret_cast<zb_bvh_node*>(A.nodes);
        uint32_t n_new = 0;
        int rc = zbk_build_tlas_device(ctx, const_cast<float4*>(s->dev.items), R.item_bounds, R.n_items, tlas, R.tlas_capacity, &n_new);
        if (rc) return rc;  // nothing but scratch and the item boxes was touched before the build's checks passed
        cudaError_t e = zbk_convert_nodes(A.nodes, n_new, ctx->stream);
        if (e != cudaSuccess) return fail(ZB_ERR_CUDA, std::string("zb_scene_rebuild_tlas: node re-encoding: ") + cudaGetErrorString(e));
        // the new topology's levels, for later refits
        std::vector<zb_bvh_node> host_nodes(n_new);
        ZB_CUDA(cudaMemcpyAsync(host_nodes.data(), tlas, size_t(n_new) * sizeof(zb_bvh_node), cudaMemcpyDeviceToHost, ctx->stream));
        ZB_CUDA(cudaStreamSynchronize(ctx->stream));
        std::vector<uint32_t> order;
        if (!tlas_levels(host_nodes.data(), n_new, order, R.level_start)) return fail(ZB_ERR_CUDA, "zb_scene_rebuild_tlas: rebuilt TLAS deeper than the cap");
        ZB_CUDA(cudaMemcpyAsync(R.order, order.data(), order.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
        ZB_CUDA(cudaStreamSynchronize(ctx->stream));
        R.n_tlas_nodes = n_new;
    }
    R.host_instances.swap(next);
    s->wide = nullptr;  // the 8-wide tree (an opt-in) is not refitted
    s->dev.wide = nullptr;
    return ZB_OK;
}

extern "C" int zb_scene_refit(zb_scene* scene, const zb_instance* instances, uint32_t n_instances) {
    if (!scene || (n_instances && !instances)) return fail(ZB_ERR_INVALID, "zb_scene_refit: NULL argument");
    int rc = refit_one(scene, instances, n_instances, false);
    for (zb_scene* rep : scene->replicas)
        if (!rc) rc = refit_one(rep, instances, n_instances, false);
    if (scene->ctx) cudaSetDevice(scene->ctx->device);
    return rc;
}

extern "C" int zb_scene_rebuild_tlas(zb_scene* scene, const zb_instance* instances, uint32_t n_instances) {
    if (!scene || (n_instances && !instances)) return fail(ZB_ERR_INVALID, "zb_scene_rebuild_tlas: NULL argument");
    int rc = refit_one(scene, instances, n_instances, true);
    for (zb_scene* rep : scene->replicas)
        if (!rc) rc = refit_one(rep, instances, n_instances, true);
    if (scene->ctx) cudaSetDevice(scene->ctx->device);
    return rc;
}
