// scene_dev.cuh — device views of the flattened scene and the per-launch parameter block.
// All arrays are read-only for the whole render and are fetched through the non-coherent path
// (__ldg, 16-byte vectors).  Layouts are the zb_* PODs of include/zenith_b200.h viewed as float4s.
#pragma once
#include "../../include/zenith_b200.h"
#include "vecmath.cuh"

// ZB_F_* scene-feature bits: include/zenith_b200.h
struct SceneDev {
    const float4* nodes;        // 4 x float4 per zb_bvh_node (64 B)
    const float4* items;        // 2 x float4 per zb_item (32 B)
    const float4* cubes;        // 2 x float4 per zb_cube
    const float4* tris;         // 3 x float4 per zb_tri
    const float4* tri_normals;  // 3 x float4 per zb_tri_normals
    const float4* instances;    // 8 x float4 per zb_instance (128 B)
    const float4* media;        // 2 x float4 per zb_medium
    const float4* materials;    // 2 x float4 per zb_material
    const float4* textures;     // 2 x float4 per zb_texture
    const float4* spheres_obj;  // 1 x float4 per object-space sphere
    const uchar4* texels_u8;
    const float4* texels_f32;
    const float4* wide;         // 5 x float4 per zb_wide_node (80 B); null: walk the binary tree (no wide tree, or tuning wide = 0)
    uint32_t n_nodes, n_items, n_media;
    uint32_t features;  // ZB_F_* bits the scene needs, set by zb_scene_upload
    // traversal-stack overflow events (trace.cuh): one u64 in the owning context's counter block, never null.  The
    // builders bound the tree depth so that this stays 0 (ZB_MAX_TLAS_DEPTH / ZB_MAX_BLAS_DEPTH); a non-zero count
    // turns the render into ZB_ERR_UNSUPPORTED instead of a frame with missing geometry.
    unsigned long long* overflow;
};

// camera::initialize (camera.hpp:358-402) evaluated in double on the host, then narrowed.
struct CameraDev {
    float3 center;         // lookfrom
    float3 pix00_rel;      // pixel00_loc - center (subtracted in double: no cancellation on the device)
    float3 du, dv;         // pixel_delta_u / pixel_delta_v
    float3 disk_u, disk_v; // defocus_disk_u / defocus_disk_v
    float3 u, v, w;        // camera basis (view-space normal AOV, camera.hpp:471-481)
    double pix00_rel_d[3], du_d[3], dv_d[3];  // double copies for the primary-ID kernel
    int width, height;
    int defocus;           // defocus_angle > 0
    int pad;
};

struct EnvDev {
    int mode;              // ZB_ENV_*
    uint32_t hdr_texture;
    float intensity;
    float pad0;
    float3 background;     // background_color * intensity for SOLID_COLOR
    // HDR_MAP: rotation matrices folded into one 3x3 (yaw, then tilt, then roll; camera.hpp:843-862)
    float3 hdr_r0, hdr_r1, hdr_r2;
    // PHYSICAL_SUN: everything that does not depend on the ray direction is precomputed on the host
    float3 sun_dir, zenith, horizon, sun_add;  // sun_add = s_color * sun_intensity * visibility
    float sky_scale;       // intensity * 1.5 * sky_exposure
    float sun_threshold;
    int sun_visible;       // adjusted_height > -0.1
    float pad1;
};

struct RenderParams {
    SceneDev sc;
    CameraDev cam;
    EnvDev env;
    float4* accum;      // W*H radiance sums (xyz) + sample count (w)
    float4* accum_aux;  // 3 planes of W*H: albedo, normal, z-depth sums; may be null
    unsigned long long* counters;  // [0] rays [1] node visits [2] box tests [3] prim tests [4] instance visits [5] paths
    uint32_t seed_lo, seed_hi;
    uint32_t pk[2][24];            // Philox round keys of the seed (vecmath.cuh): [0] beauty, [1] light passes (seed_hi ^ "LITE")
    int sample_begin, sample_end;  // global sample indices rendered by this launch
    int max_depth;
    int aux_samples;               // clamp(spp/8, 64, 1024) (camera.hpp:433)
    int use_albedo, use_normal, use_z;
    float inv_z_max;
    int tile_rank, tile_ranks;     // tile stripes: block b renders tile b*tile_ranks + tile_rank
    int tiles_x, tiles_y;
    int accumulate;                // 0: overwrite accum, 1: add to it
    int light_pass;                // 1: reflection / refraction passes (camera.hpp:492-517) instead of beauty
    int exp_flags;                 // experiments build only (tuning "pq_redeal"): 1 = block-synchronous segments, 2 = + re-deal rays by octant
};

#ifndef ZB_TILE_W
#define ZB_TILE_W 16  // block tile in pixels: (ZB_TILE_W/8) x (ZB_TILE_H/4) warps of 8x4 pixels each
#define ZB_TILE_H 8
#endif
#define ZB_BLOCK_THREADS (ZB_TILE_W * ZB_TILE_H)
#ifndef ZB_MIN_BLOCKS
#define ZB_MIN_BLOCKS 6  // resident blocks per SM the render kernel is compiled for (80 registers; measured best of 3..12)
#endif
