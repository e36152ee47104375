/*
 * zenith_b200.h — C ABI of the B200-native backend for Zenith's path-tracing hot path.
 *
 * The reference (jarek1992/raytracer_project) has no FFI layer; its seam is the C++ member
 *   void camera::render(const hittable& world, const EnvironmentSettings& env,
 *                       const post_processor& post, std::atomic<bool>& render_flag)   camera.hpp:236
 * whose hot body is camera::execute_render_threads (camera.hpp:404-579).  Every entry point below
 * cites the piece of that interface it replaces.  Plain pointers and sizes only; no C++ or torch
 * types cross this boundary.  All functions return ZB_OK (0) or a negative zb_status; the message
 * for the last failure on the calling thread is available from zb_last_error().
 *
 * There is no CPU fallback behind this ABI: every compute entry point fails with ZB_ERR_CUDA when
 * no sm_100 device is usable.
 */
#ifndef ZENITH_B200_H
#define ZENITH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ZB_ABI_VERSION 4

typedef enum zb_status {
    ZB_OK = 0,
    ZB_ERR_INVALID = -1,   /* bad argument / malformed flat scene                     */
    ZB_ERR_CUDA = -2,      /* CUDA runtime failure or no usable device                */
    ZB_ERR_IO = -3,        /* scene / asset file could not be read                    */
    ZB_ERR_CANCELLED = -4, /* cancel flag observed; outputs hold the finished batches */
    ZB_ERR_UNSUPPORTED = -5
} zb_status;

/* ------------------------------------------------------------------------------------------
 * Flat scene: structure-of-arrays, POD, produced once by the C++20 host flattener
 * (raytracer_project_b200/host/flatten.hpp) from the reference-shaped scene graph
 * (hittable_list / bvh_node / sphere / triangle / cube / translate / rotate_* / scale /
 * material_instance / constant_medium / model; hittable.hpp:29-36 and Appendix C of SURVEY.md).
 * ---------------------------------------------------------------------------------------- */

/* BVH2 node, 64 B, 32-byte aligned: the boxes of BOTH children live in the parent so one node
 * fetch (4 x LDG.128) decides both slab tests.  Replaces bvh_node {left,right,bbox} bvh.hpp:119-122.
 * child >= 0 : absolute index of an inner node in nodes[]
 * child <  0 : leaf, code = ~child, first = code >> 3, count = (code & 7) + 1 (items or triangles)
 * child == ZB_CHILD_EMPTY : no child (box is inverted and never hit). */
typedef struct zb_bvh_node {
    float lx_ly[4]; /* left.lo.x, left.hi.x, left.lo.y, left.hi.y   */
    float rx_ry[4]; /* right.lo.x, right.hi.x, right.lo.y, right.hi.y */
    float lz_rz[4]; /* left.lo.z, left.hi.z, right.lo.z, right.hi.z */
    int32_t left, right;
    uint32_t pad0, pad1;
} zb_bvh_node;
#define ZB_CHILD_EMPTY ((int32_t)0x80000000)
/* Depth caps (inner nodes on a root-to-leaf path) all builders keep — the host and the device binned-SAH builders fall back
 * to splits by count where SAH would exceed them, the device linear BVH is measured after the build — and zb_scene_upload verifies: the
 * per-ray traversal stack (csrc/trace.cuh) holds one deferred child per level.  The reference's recursive traversal
 * (bvh.hpp:46-54) has no such bound; its median-split tree is log2(N) deep by construction (bvh.hpp:15-44). */
#define ZB_MAX_TLAS_DEPTH 22
#define ZB_MAX_BLAS_DEPTH 38

/* 8-wide compressed BVH node, 80 B (five 16-byte loads), built by the host for scenes with meshes beside the binary tree
 * (host/wide_bvh.hpp; layout after Ylitie, Karras, Laine 2017).  Child i: lo = p + qlo[axis][i] * 2^(e[axis] - 127),
 * hi likewise, rounded outwards.  meta[i]: high 3 bits = primitives of a leaf child (1..4; 1 for an inner child, 0 = empty
 * slot), low 5 bits = 24 + i for an inner child, else the offset of the leaf's first primitive from prim_base.  Inner
 * children are the consecutive nodes child_base + popcount(imask below slot i); a node's leaf primitives (TLAS items or
 * BLAS triangles, <= 24) are consecutive from prim_base.  Both trees address the same primitive arrays. */
typedef struct zb_wide_node {
    float p[3];
    uint8_t e[3];
    uint8_t imask;
    uint32_t child_base;
    uint32_t prim_base;
    uint8_t meta[8];
    uint8_t qlo[3][8];
    uint8_t qhi[3][8];
} zb_wide_node;

/* Top-level leaf item, 32 B.  One per leaf object of the world (DFS order of the scene graph
 * gives `leaf_id`; items are stored in BVH leaf order).  Spheres that sit directly in the world
 * (no non-trivial wrapper) are inlined so a test costs one 32-byte load (sphere.hpp:18-64). */
enum {
    ZB_ITEM_SPHERE = 0,   /* geom = centre.xyz, radius; world space                        */
    ZB_ITEM_CUBE = 1,     /* aux = index into cubes[]; world space (cube.hpp:44-86)        */
    ZB_ITEM_TRIANGLE = 2, /* aux = index into tris[] (a loose triangle in the world)       */
    ZB_ITEM_INSTANCE = 3, /* aux = index into instances[]                                  */
    ZB_ITEM_MEDIUM = 4    /* aux = index into media[] (constant_medium.hpp:39-77)          */
};
#define ZB_FLAG_FORCE_FRONT 0x100u /* translate / rotate_y above the primitive: front_face forced
                                      to true (translate.hpp:29, rotate_y.hpp:70, SURVEY A.q1) */
typedef struct zb_item {
    float geom[4];
    uint32_t kind_flags; /* low 8 bits kind, then ZB_FLAG_*                                 */
    uint32_t material;   /* index into materials[]                                          */
    uint32_t leaf_id;    /* DFS index of the leaf object in the scene graph                 */
    uint32_t aux;
} zb_item;

/* Origin-slab cube (cube.hpp:44-86): slabs are +-half about the WORLD origin of the local frame,
 * face classification uses p - centre (cube.hpp:75, :100-142).  32 B. */
typedef struct zb_cube {
    float half[4];   /* half extents, w unused */
    float centre[4]; /* cube::center, w unused */
} zb_cube;

/* Triangle, 48 B for the test + 48 B normals fetched only at the accepted hit (triangle.hpp:17-82).
 * v0.w carries the face id (bit pattern of a uint32). */
typedef struct zb_tri {
    float v0[4], v1[4], v2[4];
} zb_tri;
typedef struct zb_tri_normals {
    float n0[4], n1[4], n2[4];
} zb_tri_normals;

/* Instance = collapsed wrapper chain translate/rotate_x/rotate_y/rotate_z/scale (+ material_instance)
 * over one shared geometry (translate.hpp:15-32, rotate_*.hpp, scale.hpp:20-36).  128 B.
 * Ray directions are never renormalised, so t is invariant under the transform (SURVEY A.q7). */
enum { ZB_GEOM_SPHERE = 0, ZB_GEOM_CUBE = 1, ZB_GEOM_MESH = 2, ZB_GEOM_TRIANGLE = 3 };
typedef struct zb_instance {
    float w2o[12]; /* world -> object, row-major 3x4 */
    float o2w[12]; /* object -> world, row-major 3x4 */
    uint32_t geom_kind;
    uint32_t geom_index; /* spheres_obj[] / cubes[] / meshes[] / tris[] */
    uint32_t flags;      /* ZB_FLAG_FORCE_FRONT */
    uint32_t pad[5];     /* mesh instances: pad[0] = BLAS root in nodes[], pad[1] = BLAS root in wide_nodes[] */
} zb_instance;

typedef struct zb_mesh {
    uint32_t node_root; /* first node of this mesh's BLAS inside nodes[]        */
    uint32_t node_count; /* 0 with tri_count > 0: no host BLAS, zb_scene_upload builds one on the device */
    uint32_t tri_first; /* first triangle of this mesh inside tris[] (BLAS leaf order) */
    uint32_t tri_count;
} zb_mesh;

/* constant_medium (constant_medium.hpp:27-77): boundary is a sphere or cube, optionally under an
 * instance transform.  One free-flight draw per (ray, medium) (SURVEY A.q6). */
typedef struct zb_medium {
    uint32_t boundary_kind;  /* ZB_ITEM_SPHERE (geom in sphere[]), ZB_ITEM_CUBE, ZB_ITEM_INSTANCE */
    uint32_t boundary_index; /* cubes[] / instances[] index; unused for an inline sphere          */
    float neg_inv_density;
    uint32_t pad;
    float sphere[4];         /* inline boundary sphere                                            */
} zb_medium;

enum { ZB_MAT_LAMBERTIAN = 0, ZB_MAT_METAL = 1, ZB_MAT_DIELECTRIC = 2, ZB_MAT_LIGHT = 3, ZB_MAT_ISOTROPIC = 4 };
#define ZB_NO_TEXTURE 0xFFFFFFFFu
/* Tagged-union material replacing shared_ptr<material> virtual dispatch (material.hpp:7-279,
 * constant_medium.hpp:9-22).  32 B. */
typedef struct zb_material {
    uint32_t type;
    uint32_t albedo_tex; /* ZB_NO_TEXTURE: use the inline colour (solid_color fast path)          */
    uint32_t bump_tex;   /* ZB_NO_TEXTURE: none (material.hpp:35-54)                              */
    float bump_strength;
    float rgb[3];        /* albedo / tint / emission when albedo_tex == ZB_NO_TEXTURE             */
    float param;         /* metal: fuzz (clamped <=1, material.hpp:116); dielectric: ior          */
} zb_material;

enum { ZB_TEX_SOLID = 0, ZB_TEX_CHECKER = 1, ZB_TEX_IMAGE_U8 = 2, ZB_TEX_IMAGE_F32 = 3, ZB_TEX_MISSING = 4 };
/* texture.hpp:12-133.  32 B.  Images: a = width, b = height, c = first texel (uchar4 / float4 units)
 * in texels_u8 / texels_f32.  Checker: f[0] = 1/scale, a = odd texture, b = even texture. */
typedef struct zb_texture {
    uint32_t kind, a, b, c;
    float f[4];
} zb_texture;

typedef struct zb_flat_scene {
    uint32_t abi_version; /* ZB_ABI_VERSION */
    uint32_t n_nodes;     /* TLAS nodes come first (root = node 0); mesh BLASes follow */
    uint32_t n_tlas_nodes;
    uint32_t n_items, n_cubes, n_tris, n_instances, n_meshes, n_media, n_materials, n_textures;
    uint32_t n_spheres_obj; /* object-space spheres referenced by instances                  */
    uint64_t n_texels_u8;   /* uchar4 units                                                  */
    uint64_t n_texels_f32;  /* float4 units                                                  */
    const zb_bvh_node* nodes;
    const zb_item* items;
    const zb_cube* cubes;
    const zb_tri* tris;
    const zb_tri_normals* tri_normals;
    const zb_instance* instances;
    const zb_mesh* meshes;
    const zb_medium* media;
    const zb_material* materials;
    const zb_texture* textures;
    const float* spheres_obj;  /* 4 floats each */
    const uint8_t* texels_u8;  /* RGBA8 */
    const float* texels_f32;   /* RGBA32F */
    uint32_t n_leaves;         /* number of leaf ids handed out */
    uint32_t pad;
    /* optional 8-wide tree over the same primitives (n_wide_nodes == 0, the default: none; the renderer walks the binary tree):
     * TLAS root = wide_nodes[0], BLAS roots in zb_instance::pad[1].  wide_stack_need = traversal-stack entries a ray can
     * hold: deepest TLAS path + deepest BLAS path + 4. */
    uint32_t n_wide_nodes;
    uint32_t wide_stack_need;
    const zb_wide_node* wide_nodes;
    /* optional: world-space bounds of every item, 8 floats each (lo.xyz, 0, hi.xyz, 0), in items[] order — the
     * reference's bounding_box() of the leaf object.  Needed by zb_scene_refit only; NULL = the scene cannot be refitted. */
    const float* item_bounds;
} zb_flat_scene;

/* The 13 public camera parameters (camera.hpp:26-42); derived quantities are computed in double by
 * the library exactly as camera::initialize does (camera.hpp:358-402). */
typedef struct zb_camera {
    int32_t image_width, image_height;
    double vfov;
    double lookfrom[3], lookat[3], vup[3];
    double defocus_angle, focus_dist;
} zb_camera;

/* EnvironmentSettings (environment.hpp:8-77) as consumed by camera::get_background_color
 * (camera.hpp:828-925). */
enum { ZB_ENV_PHYSICAL_SUN = 0, ZB_ENV_HDR_MAP = 1, ZB_ENV_SOLID_COLOR = 2 };
typedef struct zb_environment {
    int32_t mode;
    uint32_t hdr_texture; /* index into the scene's textures[] or ZB_NO_TEXTURE */
    double background_color[3];
    double intensity;
    double hdri_rotation, hdri_tilt, hdri_roll;
    double sun_direction[3], sun_color[3];
    double sun_intensity, sun_size;
} zb_environment;

/* how a multi-GPU context splits a frame: AUTO = sample slices when every GPU gets >= 8 samples per pixel,
 * else tile stripes; both give the same bits (integer sums) */
enum { ZB_PARTITION_AUTO = 0, ZB_PARTITION_SAMPLES = 1, ZB_PARTITION_TILES = 2 };
/* camera.hpp:28-31, :53-57 and post_processor::z_depth_max_dist (camera.hpp:247). */
typedef struct zb_render_opts {
    int32_t samples_per_pixel;
    int32_t max_depth;
    uint64_t seed;
    /* this rank's share of the frame (multi-GPU) */
    int32_t sample_begin, sample_end; /* [begin,end) of the global sample index; (0,spp) or a zeroed (0,0) = all; an EMPTY
                                         share is written as begin == end > 0 (e.g. spp,spp), never as (0,0)            */
    int32_t tile_rank, tile_ranks;    /* tile stripes: this rank owns tiles t % ranks == rank  */
    int32_t use_albedo, use_normal, use_z_depth;
    double z_depth_max_dist;
    int32_t collect_stats;            /* 1: STATS kernel variant (node visits / primitive tests) */
    int32_t samples_per_launch;       /* 0 = library default; cancel is polled between launches   */
    /* camera::use_reflection / use_refraction (camera.hpp:56-57): the two light passes of render_rows
     * (camera.hpp:492-517).  zb_render fills zb_outputs::reflection / refraction when set; for
     * zb_render_device, light_pass = 1 renders the light passes INSTEAD of the beauty pass: reflection
     * sums go to accum_f4 and refraction sums to the first W*H float4 of accum_aux_f4. */
    int32_t use_reflection, use_refraction;
    int32_t light_pass;
    int32_t accumulate;               /* zb_render_device: 1 = add this call's samples to the sums already in accum_f4
                                         (progressive rendering, resume from a checkpoint) instead of overwriting */
    int32_t partition;                /* multi-GPU contexts: ZB_PARTITION_*; ignored on one GPU                          */
    int32_t reserved;
} zb_render_opts;

/* Host-side outputs in the reference's own layout: row-major, top row first, 3 doubles per pixel
 * (std::vector<color>, camera.hpp:63,80-85).  Any pointer may be NULL. */
typedef struct zb_outputs {
    double* beauty;  /* render_accumulator */
    double* albedo;  /* albedo_buffer      */
    double* normal;  /* normal_buffer      */
    double* z_depth; /* z_depth_buffer     */
    float* beauty_f4; /* optional: RGBA32F mean radiance, 4 floats per pixel */
    double* reflection; /* reflection_buffer (needs opts.use_reflection) */
    double* refraction; /* refraction_buffer (needs opts.use_refraction) */
} zb_outputs;

typedef struct zb_stats {
    uint64_t paths, rays, node_visits, box_tests, prim_tests, instance_visits;
    double kernel_ms;  /* device time of the path-tracing launches of the last render   */
    double resolve_ms; /* device time of the resolve kernel                             */
    double h2d_ms, d2h_ms;
    uint32_t launches; /* kernels launched by the last render                           */
    uint32_t stack_overflows; /* traversal-stack overflow events of the last render (saturating): must be 0; zb_render
                                 returns ZB_ERR_UNSUPPORTED otherwise                                                 */
    uint32_t kernel_pick;     /* k_render_pq instantiation of the last render: (blocks per SM << 16) | (mode << 8) | ZB_F_* set;
                                 ties measured instruction counts (profiles/) to the code that ran                  */
    uint32_t n_gpus;          /* GPUs that rendered the last frame                                                   */
} zb_stats;

typedef struct zb_context zb_context;
typedef struct zb_scene zb_scene;
typedef void (*zb_progress_fn)(int lines_rendered, void* user); /* camera::lines_rendered camera.hpp:88 */

int zb_abi_version(void);
const char* zb_last_error(void);

/* one context per process/GPU; `stream` is a cudaStream_t (or NULL for the default stream) */
int zb_context_create(int device, zb_context** out);
/* ---- multi-GPU.  In the reference ALL parallelism lives inside the one blocking camera::render call
 * (thread fan-out camera.hpp:426-427, static row split :557-566, joins :569-573); so it does here.
 * zb_context_create_multi: one process drives n GPUs of one node (peer access enabled between all pairs).  The returned
 * context is used like any other: zb_scene_upload replicates the scene on every GPU, zb_render partitions the frame
 * (zb_render_opts::partition), every GPU gathers + resolves its own band of rows from all GPUs' 64-bit integer sums over
 * NVLink and copies that band home over its own PCIe link.  The frame is bit-identical to the one-GPU frame.
 * zb_primary_ids, zb_probe_rays, the *_device entry points and the finishing chain run on the first device.  A device
 * may be listed more than once (its ranks then share that GPU): of no use in production, it lets the whole path run
 * where a single GPU is present. */
int zb_context_create_multi(const int* device_ids, int n_devices, zb_context** out);
int zb_context_device_count(zb_context* ctx);
/* One process per GPU (torchrun / MPI launchers): every rank creates a plain context, exports a handle (CUDA IPC handles of
 * its sum buffers and its render-done event, sized for frames of up to max_pixels), the caller's own transport gathers the
 * n handles, and every rank joins with the same array and the same shared-memory name (a POSIX shm object holding the
 * host barrier, e.g. "/zb_<job id>").  From then on zb_render on these contexts is COLLECTIVE: all ranks call it with the
 * same arguments; each writes its band of rows into its `out` buffers — which may be one buffer in shared memory mapped by
 * all ranks, so that the whole frame lands in one place without a gather on the host.  cancel / progress are ignored. */
typedef struct zb_group_handle { unsigned char opaque[256]; } zb_group_handle;
int zb_group_export(zb_context* ctx, size_t max_pixels, zb_group_handle* out);
int zb_group_join(zb_context* ctx, int rank, int n_ranks, const zb_group_handle* all, const char* shm_name);
/* stream synchronise + barrier over the ranks of the group (shared-memory host barrier; plain stream sync elsewhere) */
int zb_group_barrier(zb_context* ctx);
int zb_group_leave(zb_context* ctx);
int zb_context_set_stream(zb_context* ctx, void* stream);
int zb_context_destroy(zb_context* ctx);
/* Optional: [ptr, ptr + bytes) is an output buffer the caller hands to zb_render again and again — the reference's
 * render_accumulator is sized once by reset_accumulator and reused (camera.hpp:209-233) — and stays valid until
 * zb_context_unregister_output or zb_context_destroy.  The range is page-locked (cudaHostRegister) and zb_render then sends
 * every finished band of rows straight into it as doubles, one DMA per output plane, instead of staging packed floats and
 * widening them with host threads: the same values bit for bit, no host work after the kernel (what matters when 8 ranks
 * share one host's cores).  Applies when EVERY buffer requested in zb_outputs lies inside registered ranges; at most 8
 * ranges per context; ZB_ERR_UNSUPPORTED (and the staged path stays in use) when the range cannot be page-locked.
 * No counterpart in the reference (its workers write the accumulator directly, camera.hpp:531). */
int zb_context_register_output(zb_context* ctx, void* ptr, size_t bytes);
int zb_context_unregister_output(zb_context* ctx, void* ptr);
/* A/B switches for measurements and equality tests; product defaults need none of them.  Read once from the environment at
 * context creation (ZB_PQ, ZB_NO_LEAN, ZB_PQ_BLOCKS, ZB_GATHER_P2P), changed afterwards only here.  Keys: "pq" (0 tile-bound
 * k_render, 1 path-queue beauty kernel, 2 also its AOV / light modes), "no_lean" (1: general kernels), "pq_blocks" (6..8),
 * "gather_p2p" (reserved), "gpu_builder" (device BLAS build: 1 binned SAH, 0 linear BVH), "build_timing" (1: the device
 * builders print their phase times to stderr), "wide" (1: walk the 8-wide tree where a scene has one, 0: the binary tree); a library built with -DZB_EXPERIMENTS
 * (zb_build_flags() & ZB_BUILD_EXPERIMENTS) adds "pq_resume", "wavefront", "wf_slots" for the rejected kernel forms. */
int zb_context_set_tuning(zb_context* ctx, const char* key, long long value);
#define ZB_BUILD_EXPERIMENTS 1
int zb_build_flags(void);
int zb_device_info(zb_context* ctx, int* sm_count, int* cc_major, int* cc_minor, size_t* l2_bytes);

/* copies the flat scene to the GPU; the caller keeps ownership of the host arrays */
int zb_scene_upload(zb_context* ctx, const zb_flat_scene* flat, zb_scene** out);
int zb_scene_destroy(zb_scene* scene);
/* Transform-only scene edit without re-flattening (the reference rebuilds geometry and BVH on every GUI
 * restart, main.cpp:1492-1500): `instances` = the scene's instance table with NEW transforms (same count, same geometry
 * references — e.g. the table of a re-run flattener, or the uploaded one with w2o / o2w edited).  The table is copied to the
 * GPU(s), the world boxes of the items that reference instances are recomputed from the object-space bounds of their
 * geometry, and the TLAS is refitted bottom-up ON THE DEVICE with its topology unchanged; BLASes, materials and textures
 * are untouched.  One small kernel (k_refit_tlas): well under a millisecond for thousands of items.  Closest hits equal
 * those of a freshly flattened scene; a heavily rearranged scene traverses faster after zb_scene_rebuild_tlas (SAH
 * topology).  Needs zb_flat_scene::item_bounds at upload; an 8-wide tree, if any, is dropped (not refitted). */
int zb_scene_refit(zb_scene* scene, const zb_instance* instances, uint32_t n_instances);
/* Same edit, but the TLAS TOPOLOGY is rebuilt as well, on the device: new world boxes of the items, then a binned-SAH tree
 * over them (csrc/sah_build.cu, the builder of the device BLASes, with the flattener's TLAS cost model) written into the
 * scene's TLAS node region; the items are permuted into the new leaf order.  The flattener pads the TLAS region to
 * n_items - 1 nodes so that any tree fits; a flat scene without that room fails with ZB_ERR_UNSUPPORTED when the new tree
 * needs more nodes than the uploaded one had.  Closest hits equal those of a freshly flattened scene; later
 * zb_scene_refit calls refit the new topology.  Launch-bound: ~1 ms for a few thousand items.  If either call fails after
 * its argument checks (CUDA error, no room for the new tree), the instance table is already the new one while the tree is
 * the old one: refit again or upload the scene again before rendering. */
int zb_scene_rebuild_tlas(zb_scene* scene, const zb_instance* instances, uint32_t n_instances);

/* Drop-in for camera::render's hot part (camera.hpp:236-250 -> :404-579): blocking; fills host
 * buffers.  `cancel` mirrors render_flag inverted (non-zero = stop, camera.hpp:441). */
int zb_render(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
              const zb_render_opts* opts, zb_outputs* out, const volatile int* cancel,
              zb_progress_fn progress, void* user);

/* Device-resident variant used for multi-GPU and kernel-only timing: accumulates radiance SUMS of
 * this rank's share into `accum_f4` (W*H float4 on the device, overwritten) on the context stream
 * without synchronising; aux sums (albedo, normal, z) go to accum_aux_f4 (3*W*H float4) if non-NULL. */
int zb_render_device(zb_context* ctx, const zb_scene* scene, const zb_camera* cam,
                     const zb_environment* env, const zb_render_opts* opts, void* accum_f4,
                     void* accum_aux_f4);
/* zb_render_device's beauty pass WITHOUT any device allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The kernel's scratch (64-bit fixed-point per-pixel sums + queue head) lives in a static __device__ array
 * sized for ZB_NOALLOC_MAX_PIXELS pixels (24 MiB, held by every GPU that loads the library until exit); larger frames are
 * rejected with ZB_ERR_INVALID.  Beauty only (no aux planes, light_pass, collect_stats); sample ranges, tile stripes and
 * `accumulate` work as in zb_render_device and the sums written to accum_f4 are bit-identical to its.  Asynchronous on the
 * context stream.  All calls on one GPU share the array: each call clears what it uses first, and calls that OVERLAP on
 * the GPU (other streams / contexts) race. */
#define ZB_NOALLOC_MAX_PIXELS (1u << 20)
int zb_render_device_noalloc(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                             const zb_render_opts* opts, void* accum_f4);
size_t zb_noalloc_max_pixels(void);
/* K5 resolve: out_f4[i] = accum_f4[i] * scale (device -> device), or to the reference's double RGB */
int zb_resolve_device(zb_context* ctx, const void* accum_f4, size_t n_pixels, double scale, void* out_f4,
                      void* out_rgb_f64);

/* Device float4 means (the out_f4 of zb_resolve_device) -> caller-owned host buffer in the reference's layout
 * (row-major, 3 doubles per pixel, std::vector<color> camera.hpp:63).  Blocking; the frame crosses PCIe as
 * float4 in chunks and is widened by host threads while later chunks are in flight. */
int zb_read_frame(zb_context* ctx, const void* mean_f4, size_t n_pixels, double* out_rgb_f64);

/* K7 parity artefact: centre-of-pixel pinhole primary hit (camera::get_center_ray camera.hpp:809-814
 * + world.hit with interval(0.001, inf) camera.hpp:459).  leaf_id = 0xFFFFFFFF on a miss;
 * prim_id = face index inside a mesh, else 0.  Host pointers, W*H entries each; any may be NULL. */
int zb_primary_ids(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, uint32_t* leaf_id,
                   uint32_t* prim_id, float* t);

/* Per-ray known-answer probe (parity tool, and a plain ray-query API): closest hit of arbitrary rays with
 * the reference's interval (0.001, inf) and the hit_record the reference would build for it
 * (hittable.hpp:9-26: t, p, normal, front_face, u, v) plus material::get_albedo / emitted at the hit, or
 * camera::get_background_color for a miss (stored in `albedo`).  Host pointers; origins / directions are
 * n x 3 floats. */
typedef struct zb_probe {
    float t;             /* INFINITY on a miss */
    uint32_t leaf_id;    /* 0xFFFFFFFF on a miss */
    uint32_t prim_id;
    uint32_t front_face;
    float p[3], normal[3];
    float u, v;
    float albedo[3];     /* get_albedo(rec), or the background colour on a miss */
    float emitted[3];
} zb_probe;
int zb_probe_rays(zb_context* ctx, const zb_scene* scene, const zb_environment* env, size_t n,
                  const float* origins, const float* directions, zb_probe* out);

/* Known-answer probe (parity tool): ONE device function of the shading code per case, on explicit inputs, so that each can
 * be compared with the reference function it restates — refract / reflect (vec3.hpp:204-214), dielectric::reflectance
 * (material.hpp:237-241), texture::value (texture.hpp:50-78, :96-98, :118-126; index = textures[] index),
 * material::get_bumped_normal (material.hpp:35-54; index = materials[] index), sphere::get_sphere_uv (sphere.hpp:70-79),
 * camera::get_background_color (camera.hpp:828-925).  Argument layouts: csrc/kernels.cu k_kat.  Host pointers; blocking. */
enum { ZB_KAT_REFRACT = 0, ZB_KAT_REFLECT = 1, ZB_KAT_REFLECTANCE = 2, ZB_KAT_TEXTURE = 3, ZB_KAT_BUMP = 4,
       ZB_KAT_SPHERE_UV = 5, ZB_KAT_BACKGROUND = 6 };
typedef struct zb_kat_case {
    uint32_t op, index;
    float a[14];
} zb_kat_case;
typedef struct zb_kat_result {
    float r[4]; /* r[3] < 0: unknown op */
} zb_kat_result;
int zb_kat_eval(zb_context* ctx, const zb_scene* scene, const zb_environment* env, size_t n, const zb_kat_case* cases,
                zb_kat_result* out);

/* Accumulation checkpoint: the float4 radiance sums + how many samples of which seed they hold, so a
 * progressive render can stop and resume (the Philox counter carries the global sample index, so samples
 * [0, k) + [k, n) give the frame of [0, n)).  "ZBCK" file: magic, u32 width, height, samples_done, reserved, u64 seed,
 * then W*H float4. */
int zb_checkpoint_save(zb_context* ctx, const char* path, const void* accum_f4, int width, int height, int samples_done,
                       uint64_t seed);
int zb_checkpoint_load(zb_context* ctx, const char* path, void* accum_f4, int width, int height, int* samples_done,
                       uint64_t* seed);

int zb_get_stats(zb_context* ctx, zb_stats* out);

/* ------------------------------------------------------------------------------------------------
 * Finishing chain on the device-resident frame: what camera::render does after the
 * path tracer — post_processor::analyze_framebuffer + apply_auto_exposure (color_processing.hpp:153-210,
 * camera.hpp:258-267) and camera::update_post_processing (camera.hpp:142-207): exposure, bloom
 * (bloom.hpp:18-68), 5-tap sharpening (color_processing.hpp:212-232) and post_processor::process
 * (color_processing.hpp:76-150: balance, contrast, vignette, HSV, ACES, debug views, gamma).
 * The fields mirror post_processor (color_processing.hpp:44-74). */
typedef struct zb_post {
    float exposure, saturation, contrast, hue_shift, vignette_intensity;
    float color_balance[3];
    float exposure_compensation_stops, target_luminance;
    int32_t use_aces_tone_mapping, use_auto_exposure;
    int32_t debug_red, debug_green, debug_blue, debug_luminance, debug_bvh;
    int32_t use_bloom;
    float bloom_threshold, bloom_intensity;
    int32_t bloom_radius;
    int32_t use_sharpening;
    double sharpen_amount;
} zb_post;

/* image_statistics (color_processing.hpp:9-28) */
typedef struct zb_image_stats {
    float average_luminance; /* 2^(mean log2 max(lum, 1e-4)) */
    float max_luminance;
    int32_t histogram[256];  /* log2 luminance in [-10, 10] mapped to 256 bins */
} zb_image_stats;

enum { ZB_PASS_BEAUTY = 0, ZB_PASS_LIGHT = 1, ZB_PASS_AUX = 2 }; /* RGB/DENOISE | REFLECTIONS/REFRACTIONS | ALBEDO/NORMALS/Z_DEPTH */

/* post_processor::analyze_framebuffer over W*H float4 means on the device (blocking, small D2H of the result) */
int zb_analyze_device(zb_context* ctx, const void* mean_f4, size_t n_pixels, zb_image_stats* out);
/* host-buffer variant (W*H*3 doubles, the reference's std::vector<color>); blocking */
int zb_analyze_f64(zb_context* ctx, const double* rgb_f64, size_t n_pixels, zb_image_stats* out);
/* post_processor::apply_auto_exposure (host arithmetic, no device work): the exposure render() stores back */
double zb_auto_exposure(const zb_post* post, const zb_image_stats* stats);
/* camera::update_post_processing for one pass: mean_f4 (W*H float4, device) -> display-referred colour in
 * out_f4 (float4, device, may be NULL) and/or out_rgba8 (uchar4, device, may be NULL).  Asynchronous on the
 * context stream.  Bloom needs two W*H float4 scratch planes, owned by the context. */
int zb_finish_device(zb_context* ctx, const void* mean_f4, int width, int height, const zb_post* post,
                     int pass_kind, void* out_f4, void* out_rgba8);
/* Host-buffer variant in the reference's layout (3 doubles per pixel in, final_framebuffer out); blocking. */
int zb_finish(zb_context* ctx, const double* rgb_f64, int width, int height, const zb_post* post, int pass_kind,
              double* out_rgb_f64);

/* ---- BVH debug view: bvh_node::hit with global_settings::bvh_debug_mode (bvh.hpp:46-110,
 * camera.hpp:937-954).  `level` = global_settings::debug_bvh_level (-1: leaves), `thickness` = bvh_thickness
 * (common.hpp:121-123, GUI range 0.001 .. 0.05).  Shows the product's own tree (see csrc/wireframe.cu for what that
 * changes); level 0 is tree-independent.  A primary miss shows the regular background.  Writes per-pixel SUMS + sample count into accum_f4 (device, W*H float4),
 * asynchronous on the context stream; zb_resolve_device / zb_finish_device / zb_save_pass take it from there. ---- */
int zb_render_bvh_debug_device(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                               const zb_render_opts* opts, int level, float thickness, void* accum_f4);
/* host-buffer variant: means in the reference's layout (W*H*3 doubles), what camera::render leaves in
 * render_accumulator when global_settings::bvh_debug_mode is set.  Blocking. */
int zb_render_bvh_debug(zb_context* ctx, const zb_scene* scene, const zb_camera* cam, const zb_environment* env,
                        const zb_render_opts* opts, int level, float thickness, double* rgb_f64);

/* ---- image export: camera::save_render_pass (camera.hpp:299-343), process_framebuffer_to_image
 * (:701-780), file names "render_<Pass>.png" (main.cpp:1317) ---- */
enum { ZB_RP_RGB = 0, ZB_RP_DENOISE = 1, ZB_RP_ALBEDO = 2, ZB_RP_NORMALS = 3, ZB_RP_REFLECTIONS = 4, ZB_RP_REFRACTIONS = 5,
       ZB_RP_Z_DEPTH = 6 }; /* render_pass, common.hpp:110-118 */
/* "render_RGB.png" ... "render_Z-Depth.png"; NULL for a value outside the enum */
const char* zb_pass_file_name(int render_pass);
/* mean_f4 (W*H float4 means, device) -> finishing chain of the pass -> 8-bit RGB PNG at `path`.  RGB / DENOISE: bloom,
 * sharpening, exposure, process(); every other pass: clamp + gamma (the export treats the light passes as data). Blocking. */
int zb_save_pass(zb_context* ctx, const void* mean_f4, int width, int height, const zb_post* post, int render_pass,
                 const char* path);
/* host-buffer variant: the reference's std::vector<color> (W*H*3 doubles) in */
int zb_save_pass_f64(zb_context* ctx, const double* rgb_f64, int width, int height, const zb_post* post, int render_pass,
                     const char* path);
/* linear radiance as a flat Radiance RGBE file (readable by stbi_loadf, texture.hpp:23).  Blocking. */
int zb_save_hdr(zb_context* ctx, const void* mean_f4, int width, int height, const char* path);
/* host-side encoders behind the two calls above: 8-bit RGB PNG from RGB or RGBA rows (alpha dropped), RGBE from float
 * rows of `channels` >= 3 floats per pixel; a stride of 0 means tightly packed rows */
int zb_write_png(const char* path, int width, int height, const unsigned char* pixels, int channels, size_t row_stride);
int zb_write_hdr(const char* path, int width, int height, const float* pixels, int channels, size_t row_stride_floats);

/* Host side of the boundary (C++20 scene API + flattener behind a C handle): parse a neutral scene
 * description ("zscene" text, raytracer_project_b200/host/zscene.hpp) through the reference-shaped
 * scene API and flatten it.  The returned zb_flat_scene stays valid until zb_flat_scene_free. */
int zb_flat_scene_from_zscene(const char* path, const zb_flat_scene** out);
int zb_flat_scene_from_zscene_text(const char* text, size_t len, const zb_flat_scene** out);
/* Same with options.  ZB_FLATTEN_GPU_BLAS: skip the host SAH build of every mesh BLAS (bvh.hpp:11-44 equivalent) and
 * let zb_scene_upload build them on the device: binned SAH (csrc/sah_build.cu; the tree the host builder would make —
 * same traversal work per ray — in 6.5 ms for 1 M triangles and 37 ms for 10 M, against 2.8 s and 33 s on the host), or
 * with tuning "gpu_builder" = 0 the linear BVH of csrc/lbvh.cu (same build time, frames 42 % slower).  Exactly one of
 * path / text is given. */
#define ZB_FLATTEN_GPU_BLAS 1u
#define ZB_FLATTEN_WIDE 2u     /* also build the 8-wide compressed tree (zb_wide_node) for scenes with host-built mesh BLASes.
                                  Off by default: measured slower than the binary tree (profiles/r02_wide_bvh.txt) */
#define ZB_FLATTEN_WIDE_ALL 4u /* ... and for scenes without meshes too (A/B measurements) */
int zb_flat_scene_from_zscene_ex(const char* path, const char* text, size_t len, uint32_t flags, const zb_flat_scene** out);
/* Scene features (SceneDev::features; template parameter F of the render kernels, csrc/trace.cuh, shade.cuh, kernels.cu).
 * The render kernel is compiled for several feature sets: code for item kinds, textures and environments a scene does
 * not use is left out, which shrinks the kernel (instruction-cache misses were 1.07 stall cycles per issue in the
 * 5.8 k-instruction general kernel) and its register count.  A launch picks the smallest compiled superset of the
 * scene's bits.  No counterpart in the reference (virtual dispatch over shared_ptr<hittable> / <material>). */
enum {
    ZB_F_INST = 1,   /* instance items (translate / rotate / scale chains)                   */
    ZB_F_MESH = 2,   /* mesh instances: BLAS traversal in object space                       */
    ZB_F_CUBE = 4,   /* cubes (items, instanced, or medium boundaries)                       */
    ZB_F_TRI = 8,    /* triangles (loose items, instanced, or mesh faces)                    */
    ZB_F_MEDIA = 16, /* constant_medium items, isotropic phase function                      */
    ZB_F_TEX = 32,   /* textured or bump-mapped materials                                    */
    ZB_F_ENV = 64,   /* HDR-map or physical-sun environment (0: solid colour); per render    */
    ZB_F_GEOM = 1 | 2 | 4 | 8 | 16, /* none of these: every item is an inline sphere         */
    ZB_F_ALL = 127
};
/* The ZB_F_* bits a flattened scene needs (all but ZB_F_ENV, which depends on the render's environment). Host only. */
int zb_flat_scene_features(const zb_flat_scene* flat, uint32_t* features);
/* Flat scene on disk ("ZBFS" container): written once, loaded by every rank / later run without
 * re-parsing assets or re-building the BVH.  A loaded scene is released with zb_flat_scene_free. */
int zb_flat_scene_save(const zb_flat_scene* flat, const char* path);
int zb_flat_scene_load(const char* path, const zb_flat_scene** out);
int zb_flat_scene_free(const zb_flat_scene* flat);

#ifdef __cplusplus
}
#endif
#endif /* ZENITH_B200_H */
