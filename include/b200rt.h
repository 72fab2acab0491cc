/*
 * b200rt.h - C ABI of the B200-native path-trace hot path (libb200rt.so).
 *
 * Drop-in seam: the two `extern "C"` functions the reference's host code binds
 * (declared at RayTracing/RayTracing.cpp:10-14, defined at RayTracing/main_raytracing.cu:202 and
 * RayTracing/Random.cu:10) are exported under the SAME names and argument lists. Everything else
 * is the extended surface (`rt_*`) the headless host (cuda-raytracing_b200/host), tests and
 * bench drive. Plain pointers and sizes only; all `int` returns are 0 on success, a negative RT_E_*
 * otherwise (the reference prints and exits, utils/CUDAHelper.h:7-22 - this layer never exits or
 * throws across the boundary). There is NO CPU fallback: compute entry points fail with
 * RT_E_CUDA when no device is present.
 *
 * Ownership (as in the reference): the caller owns surfaces, RNG states and the
 * reference-layout scene arrays; the callee allocates nothing per frame. The B200 re-layout of the
 * BVH (RtAccel) and the env cubemap (RtEnvMap) are explicit objects created/destroyed by the caller.
 */
#ifndef B200RT_H
#define B200RT_H

#include "rt_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#define RT_OK 0
#define RT_E_INVALID (-1)   /* bad argument / inconsistent arrays */
#define RT_E_CUDA (-2)      /* CUDA runtime error (rt_last_error() has the text) */
#define RT_E_NOACCEL (-3)   /* raytracing_process(): no RtAccel bound AND the scene's device arrays cannot be sized to build one */
#define RT_E_IO (-4)        /* file missing / malformed */
#define RT_E_DEPTH (-5)     /* BVH deeper than the traversal stack */

#define RT_MISS_ID 0xFFFFFFFFu
#define RT_SPHERE_ID_BIT 0x80000000u  /* hit id = RT_SPHERE_ID_BIT | sphere index */

typedef struct RtAccel RtAccel;    /* device re-layout of one BVH (wide nodes + pre-gathered triangles) */
typedef struct RtEnvMap RtEnvMap;  /* cubemap mipmapped array + texture object */

const char* rt_last_error(void);
int rt_device_count(void);

/* ---- reference-named entry points --------------------------------------------------------- */

/* RayTracing/main_raytracing.cu:202. One progressive frame: 5 spp x <=6 bounces, running-mean blend
 * of `surface_last_frame` into `surface` (alpha=1), default stream, asynchronous. `scene` is a HOST
 * struct whose pointer members are device pointers. The B200 re-layout of its BVH is the one bound with rt_accel_bind()
 * (Scene::Upload does); an UNMODIFIED host never binds one, so on the first call for a node array it is built from the scene's own
 * device arrays (sizes from their allocations and a walk of the tree; ~30 ms for the bunny scene, once) and kept by the library.
 * A binding is tied to the CUDA buffer id of the node allocation: arrays freed and re-uploaded - even at the same address - get a
 * new re-layout; a host that OVERWRITES its arrays in place calls rt_accel_unbind(scene->gpu_bvh_nodes). Errors are printed, as
 * the reference does. */
void raytracing_process(void* surface, void* surface_last_frame, int width, int height, size_t pitch,
                        int frame_index, RtGPUScene* scene);

/* RayTracing/Random.cu:10. states[tid] = curand_init(seed, tid, 0) for tid < count*size. */
void init_rng(uint32_t thread_block_count, uint32_t thread_block_size, RtRngState* rngStates, const unsigned int seed);

/* ---- RNG ---------------------------------------------------------------------------------- */

/* states[i] = curand_init(seed, first_subsequence + i, 0), bit-exact XORWOW (RayTracing/Random.cu:7). */
int rt_init_rng(RtRngState* d_states, uint64_t count, uint64_t seed, uint64_t first_subsequence, void* stream);
/* test hook: n raw 32-bit draws (curand()) from each of `count` states; d_out[i*n + k]; states advance */
int rt_rng_draw(RtRngState* d_states, uint64_t count, uint32_t n, uint32_t* d_out, void* stream);

/* ---- acceleration structure --------------------------------------------------------------- */

/* Re-lay out the reference BVH (HOST arrays: RayTracing/BVH.cpp output + Scene's faces/vertices) into
 * the device format. The LEAVES are preserved: same leaves, same leaf contents, every leaf with its exact reference
 * box; leaves larger than `fat_leaf_threshold` (0 = default 8) get a private sub-tree over their own primitives.
 * The tree ABOVE the leaves is the library's own (surface-area heuristic + one insertion-based optimisation pass):
 * hits depend on leaf boxes and triangles only (min t, ties to the reference's visiting order; DESIGN.md sections 2, 3),
 * so every entry point returns what it returned on BVH.cpp's midpoint tree, ~16 % faster. Scenes whose reference boxes
 * are not finite or do not nest keep the reference's topology. Host cost: ~0.5 s for the bunny scene's 69 376 leaves,
 * ~2-5 s for 7 M leaves (OpenMP). Environment switches for A/B: csrc/accel_build.cpp header. */
int rt_accel_create(const RtBVHNode* nodes, uint32_t node_count, const uint32_t* face_indices,
                    const RtFace* faces, uint32_t face_count, const RtVertex* vertices, uint32_t vertex_count,
                    uint32_t fat_leaf_threshold, RtAccel** out);
int rt_accel_destroy(RtAccel* accel);
/* host only, no device: builds the same re-layout and checks its structure (every triangle slot reachable exactly once from the root
 * through the byte-offset references, ranges, depth, leaf bookkeeping). info4 = inner nodes, product leaves, max depth, arena bytes. */
int rt_accel_check_host(const RtBVHNode* nodes, uint32_t node_count, const uint32_t* face_indices,
                        const RtFace* faces, uint32_t face_count, const RtVertex* vertices, uint32_t vertex_count,
                        uint32_t fat_leaf_threshold, uint64_t* info4);
/* info[0..7] = wide nodes, triangle slots, reference leaves, fat leaves, max depth (wide), device bytes lo, hi, reference depth */
int rt_accel_info(const RtAccel* accel, uint64_t* info8);
/* associate `accel` with scenes whose gpu_bvh_nodes == d_nodes_key (lookup used by raytracing_process; valid while that allocation
 * lives). rt_accel_unbind also releases a re-layout the library built by itself for that key. */
int rt_accel_bind(const void* d_nodes_key, RtAccel* accel);
int rt_accel_unbind(const void* d_nodes_key);

/* ---- environment map ---------------------------------------------------------------------- */

/* Legacy-DDS RGBA32F cubemap (FourCC 116; utils/image/DDSLoader.cpp:135, utils/CUDATexture.cpp:112-220)
 * -> mipmapped cubemap array + linear/wrap/seamless texture object. */
int rt_envmap_create_from_file(const char* dds_path, RtEnvMap** out);
int rt_envmap_create_from_memory(const void* dds_bytes, size_t size, RtEnvMap** out);
uint64_t rt_envmap_texture(const RtEnvMap* env);  /* cudaTextureObject_t */
int rt_envmap_destroy(RtEnvMap* env);
/* General DDS containers (utils/image/DDSLoader.cpp:135-371, utils/CUDATexture.cpp:9-75,187-220): legacy and DX10
 * headers, 1D/2D/3D, cubemaps, mip chains, plain / normalised / block-compressed (BC1-BC7) texels -> mipmapped array + texture
 * object (linear, wrap, normalised coordinates; seamless for cubemaps). rt_envmap_create_* accepts every CUBEMAP format of this
 * path as the environment map (e.g. RGBA16F, BC6H). The handle type is shared with the environment map. */
int rt_texture_create_from_file(const char* dds_path, int srgb, RtEnvMap** out);
int rt_texture_create_from_memory(const void* dds_bytes, size_t size, int srgb, RtEnvMap** out);
/* texel fetch probe: d_coords = 3 floats per sample (cubemap: direction; else u, v, w), d_rgba = 4 floats per sample */
int rt_texture_fetch(const RtEnvMap* tex, int count, const float* d_coords, float lod, float* d_rgba, void* stream);
/* host-side parse only (no device). info[0..11] = width, height, depth, mips, array size (x6 for cubemaps), is_cubemap, dimensions,
 * DXGI format, bits per pixel, block bytes, has a CUDA array format, subresource count; then {offset, row pitch, slice pitch} per
 * subresource in file order (array element major) while they fit `capacity`. Returns the subresource count or < 0. */
int rt_dds_info(const void* dds_bytes, size_t size, uint64_t* info, int capacity);
/* host-side parse only (no device): level-0 texels [6][size][size][4] floats; returns size or <0 */
int rt_dds_read_level0(const char* dds_path, float* out_texels, size_t capacity_floats);

/* ---- closest-hit queries (parity + ray-throughput configs) -------------------------------- */

/* Pixel-centre primary rays of scene->camera (GPUCamera::GetRay, RayTracing/GPUScene.h:13).
 * d_id/d_t: width*height. with_spheres: also test scene->gpu_spheres first (GetRayHit semantics).
 * variant: 0 = product kernel, 1 = reference-order exact kernel (by-construction parity). */
int rt_trace_primary(const RtAccel* accel, const RtGPUScene* scene, int width, int height, int with_spheres,
                     int variant, uint32_t* d_id, float* d_t, unsigned long long* d_ray_counter, void* stream);
/* Arbitrary rays: 6 floats each (origin, direction as given; normalised inside like BVHRayHit does). */
int rt_trace_rays(const RtAccel* accel, const RtGPUScene* scene, const float* d_rays, uint64_t ray_count,
                  int with_spheres, int variant, uint32_t* d_id, float* d_t, unsigned long long* d_ray_counter, void* stream);

/* ---- path tracing ------------------------------------------------------------------------- */

typedef struct RtProcessParams
{
    void* surface;              /* float4 rows, `pitch` bytes apart (device) */
    void* surface_last_frame;   /* may equal `surface` when blend_mode != 0 */
    int32_t width, height;
    uint64_t pitch;
    int32_t frame_index;
    int32_t samples_per_launch; /* reference: 5  (main_raytracing.cu:167) */
    int32_t max_bounces;        /* reference: 6  (main_raytracing.cu:115) */
    int32_t blend_mode;         /* 0 = reference running mean (lerp last->surface, weight 1/(frame+1));
                                   1 = surface += sum of this launch's samples (multi-GPU sum buffers) */
    int32_t variant;            /* 0 = megakernel (regenerating), 1 = regrouping persistent kernel (in-register wavefront:
                                   finished rays are parked, shaded and refilled together by warp ballot; bit-identical image),
                                   2 = measurement: reference traversal order + d_work_counters (nodes popped, triangle tests),
                                   3 = measurement: the megakernel with d_work_counters (wide-node steps, triangle tests executed),
                                   (4..9, 11, 12: measured-and-rejected kernels that live as patches under profiles/experiments/),
                                   10 = queue-based wavefront: one stage per ray depth, path state in HBM, ballot / prefix-sum
                                        compaction into camera + direction-octant queues between stages; needs the workspace of
                                        rt_process_workspace_bytes(width, height, 10); bit-identical, measured 2x slower */
    int32_t write_last_frame;   /* 1 = also store the result into surface_last_frame (folds the reference's D2D copy) */
    const RtGPUScene* scene;    /* HOST struct, device pointers inside; rng_state must hold width*height states */
    const RtAccel* accel;
    unsigned long long* d_ray_counter;  /* optional: += closest-hit queries issued */
    void* stream;
    void* workspace;            /* wavefront only: device scratch of rt_process_workspace_bytes() */
    uint64_t workspace_bytes;
    unsigned long long* d_work_counters; /* optional, variants 2 / 3: [0] += nodes popped (wide-node steps), [1] += triangle tests */
} RtProcessParams;

int rt_process_ex(const RtProcessParams* params);
uint64_t rt_process_workspace_bytes(int width, int height, int variant);
/* No counterpart in the reference. rt_process_ex (variant 0 only) for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the tile-schedule scratch that
 * `workspace` would name lives in a file-scope device array of rt_process_static_scratch_bytes() = 1 MiB per GPU that loads the
 * library (held until exit), so `workspace` / `workspace_bytes` of `params` are ignored and the call allocates nothing, on the host
 * side of CUDA or the device. Limit: frames of at most 87 378 tiles of 16 x 8 pixels (11.1 M pixels; 4K = 64 800 tiles) -
 * larger frames return RT_E_INVALID with the sizes in rt_last_error(). Same image, RNG states and ray count as rt_process_ex
 * (the schedule only orders independent pixels). Every call on a device shares the array: a call with frame_index == 0 drops the
 * schedule history, later frames reuse the previous launch's measured tile costs; calls that OVERLAP on one GPU (two streams,
 * two threads) race on it - serialise them, or give all but one a workspace of their own through rt_process_ex. */
int rt_process_ex_static_scratch(const RtProcessParams* params);
uint64_t rt_process_static_scratch_bytes(void);

/* AO stress (BASELINE config 3): primary hit, then `ao_rays` diffuse-lobe rays normalize(n + S) from p + 0.01 n;
 * d_ao[pixel] = unoccluded fraction. Consumes 2 draws per AO ray from scene->rng_state. */
int rt_ambient_occlusion(const RtAccel* accel, const RtGPUScene* scene, int width, int height, int ao_rays,
                         float* d_ao, unsigned long long* d_ray_counter, void* stream);

/* Resolve of a blend_mode-1 sum buffer (rgb = summed radiance, alpha = sample count; after a multi-GPU reduce: of all ranks):
 * out = (rgb / alpha, 1) - the surface contract of main_raytracing.cu:195-199. In place (out == sum) is allowed. */
int rt_resolve_sum(const void* sum_surface, void* out_surface, int width, int height, uint64_t pitch_in, uint64_t pitch_out, void* stream);

/* ---- display transform (main.cpp:78-94 HLSL restated as CUDA) ---------------- */
/* exposure 0.5, ACES fit, sRGB encode, vertical flip (row 0 of the surface is the bottom row) -> RGBA8 */
int rt_tonemap_rgba8(const void* surface, int width, int height, uint64_t pitch, float divide_by, uint8_t* d_rgba8, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200RT_H */
