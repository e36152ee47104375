// kernels.h — launch entry points of kernels.cu, used by the C-ABI implementation (abi.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/zenith_b200.h"
#include "scene_dev.cuh"

// A/B switches of a context: read ONCE from the environment by zb_context_create (ZB_PQ, ZB_NO_LEAN, ZB_PQ_BLOCKS, and with
// -DZB_EXPERIMENTS ZB_PQ_RESUME, ZB_WAVEFRONT, ZB_WF_SLOTS), changed afterwards only through zb_context_set_tuning
struct ZbTuning {
    int pq = 1;          // 0: tile-bound k_render for everything; 1: k_render_pq beauty pass (default); 2: its AOV / light modes too
    int no_lean = 0;     // 1: general kernels instead of the per-feature-set instantiations
    int pq_blocks = 0;   // 6..8: resident blocks per SM of the specialised instantiations; 0 = measured default
    int pq_resume = 0;   // experiments build: k_render_pqr threshold
    int wavefront = 0;   // experiments build: wavefront kernels for the beauty pass
    long long wf_slots = 0;
    int pq_redeal = 0;   // experiments build: 1 = block-synchronous segments in k_render_pq, 2 = + rays re-dealt by direction octant
    int fuse_finish = 1; // finishing chain with bloom: 1 = one fused pass (k_finish_fused, radius <= 6), 0 = three kernels (A/B)
    int wide = 1;        // 1: walk the scene's 8-wide compressed tree where it has one; 0: always the binary tree
    int gpu_builder = 1; // device BLAS build (ZB_FLATTEN_GPU_BLAS): 1 = binned SAH (sah_build.cu), 0 = linear BVH (lbvh.cu)
    int blas_isect_pct = 100, blas_leaf = 4;  // device SAH builder: cost of a triangle test in percent of a node visit, leaf size
    int build_check = 0;  // 1 = every device SAH build is followed by a structural check of the tree (tests)
    int group_timing = 0; // 1 = zb_render prints the wall-clock phases of every frame to stderr
    int build_timing = 0; // 1 = device BLAS builders print their phase times to stderr
    int gather_p2p = 1;  // multi-GPU combine: 1 = fused peer-memory resolve-gather kernel, 0 = staged copies (A/B)
};
ZbTuning zbk_tuning_from_env();

cudaError_t zbk_launch_render(const RenderParams& P, int n_tiles_owned, bool stats, const ZbTuning& tn, cudaStream_t stream);
// render with a frame-wide path queue (k_render_pq).  mode: 0 beauty, 1 beauty + first-hit AOVs, 2 light passes.
// fix = zbk_pq_words(pixels, mode) u64 fixed-point sums, head = one u64 queue head per launch; clear, launch (one or
// several disjoint tile ranges), then resolve rows into P.accum / P.accum_aux or (beauty) into scaled means
size_t zbk_pq_words(size_t pixels, int mode);
// the instantiation the calling thread's last zbk_launch_render_pq ran: (blocks per SM << 16) | (mode << 8) | feature set
uint32_t zbk_last_pq_pick();
cudaError_t zbk_pq_clear(unsigned long long* fix, size_t pixels, int mode, unsigned long long* heads, int n_heads, cudaStream_t stream);
cudaError_t zbk_launch_render_pq(const RenderParams& P, int n_tiles_owned, int mode, unsigned long long* fix, unsigned long long* head,
                                 int sm_count, const ZbTuning& tn, cudaStream_t stream);
cudaError_t zbk_pq_resolve(const RenderParams& P, int mode, const unsigned long long* fix, int row0, int row1, float scale,
                           float4* mean_out, cudaStream_t stream);
// first-hit AOV planes of this rank's tiles into P.accum_aux, as a pass beside the beauty queue kernel
cudaError_t zbk_launch_aov_pass(const RenderParams& P, int n_tiles_owned, const ZbTuning& tn, cudaStream_t stream);
// the same pass adding its 64-bit fixed-point sums (7 words per pixel) to fix_aux instead of writing float planes
cudaError_t zbk_launch_aov_pass_fixed(const RenderParams& P, int n_tiles_owned, unsigned long long* fix_aux, const ZbTuning& tn,
                                      cudaStream_t stream);
cudaError_t zbk_launch_primary_ids(const RenderParams& P, uint32_t* leaf_id, uint32_t* prim_id, float* t,
                                   cudaStream_t stream);
cudaError_t zbk_launch_resolve(const float4* accum, size_t n, float scale, float4* out_f4, double* out_f64,
                               cudaStream_t stream);
cudaError_t zbk_launch_pack_rgb(const float4* in, size_t n, float* out, cudaStream_t stream);
cudaError_t zbk_launch_probe(const RenderParams& P, size_t n, const float* org, const float* dir, zb_probe* out,
                             cudaStream_t stream);
cudaError_t zbk_launch_kat(const RenderParams& P, size_t n, const zb_kat_case* in, zb_kat_result* out, cudaStream_t stream);
// nodes (lo, hi) -> (centre, half extent) in place, when the traversal is compiled for that encoding (ZB_BOX_CH)
cudaError_t zbk_convert_nodes(float4* nodes, uint32_t n, cudaStream_t stream);
bool zbk_nodes_center_half();

// wavefront form of the render (wavefront.cu, built only with -DZB_EXPERIMENTS: measured 1.5x slower than the megakernel,
// profiles/r01_wavefront.txt): trace / shade kernels per bounce over a compacted path queue
struct zb_wf_scratch;
bool zbk_wavefront_supported(const RenderParams& P);
cudaError_t zbk_launch_wavefront(const RenderParams& P, zb_wf_scratch** scratch, size_t want_slots, int* iterations,
                                 cudaStream_t stream);
void zbk_wavefront_free(zb_wf_scratch* w);

// BVH debug view (wireframe.cu): frames of the tree level `level` (-1 = leaves), sums into P.accum
cudaError_t zbk_launch_wireframe(const RenderParams& P, int level, float thickness, cudaStream_t stream);
