/*
 * b200rt_host.h - C handles over the C++ host mirror of the reference's classes
 * (cuda-raytracing_b200/host: Scene / Camera / BVH / CUDARayTracer, same names as
 * RayTracing/Scene.h, BVH.h, RayTracing.h). For callers that cannot link C++ (tests, bench.py via
 * ctypes). All functions return 0 or a negative RT_E_* unless stated; C++ exceptions are caught at
 * this boundary and reported through rt_last_error().
 */
#ifndef B200RT_HOST_H
#define B200RT_HOST_H

#include "b200rt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct RtHostScene RtHostScene;    /* RayTracing::Scene */
typedef struct RtHostTracer RtHostTracer;  /* RayTracing::CUDARayTracer */
typedef struct RtHostMesh RtHostMesh;      /* Loader::LoadedMesh */

/* ---- mesh ingest (utils/AssimpLoader.cpp:29 ImportScene stand-in) ---- */
int rt_host_mesh_load_obj(const char* obj_path, RtHostMesh** out);
/* k rounds of Loop subdivision of the OBJ's triangles, then the smooth-normal pass (BASELINE config 4a) */
int rt_host_mesh_load_obj_subdivided(const char* obj_path, int rounds, RtHostMesh** out);
void rt_host_mesh_destroy(RtHostMesh* mesh);
int rt_host_mesh_counts(const RtHostMesh* mesh, uint32_t* vertex_count, uint32_t* face_count);
int rt_host_mesh_copy(const RtHostMesh* mesh, float* positions, float* normals, uint32_t* indices, float* transform16);

/* ---- Scene (RayTracing/Scene.h:87-129) ---- */
RtHostScene* rt_host_scene_create(const char* env_dds_path /* may be NULL */);
void rt_host_scene_destroy(RtHostScene* scene);
int rt_host_scene_setup_cornell_box(RtHostScene* scene);                       /* RayTracing.cpp:79-203 */
int rt_host_scene_setup_stanford_bunny(RtHostScene* scene, const RtHostMesh* mesh);  /* RayTracing.cpp:33-69 */
int rt_host_scene_add_material(RtHostScene* scene, const RtMaterial* material);      /* returns index >= 0 */
int rt_host_scene_add_sphere(RtHostScene* scene, const float* position3, float radius, int material);
int rt_host_scene_add_triangle(RtHostScene* scene, const float* a3, const float* b3, const float* c3, int material);
int rt_host_scene_add_loaded_scene(RtHostScene* scene, const RtHostMesh* mesh, const float* transform16_column_major, int material);
int rt_host_scene_build_bvh(RtHostScene* scene);                               /* host only: BVH::Calculate */
/* counts6 = vertices, faces, spheres, materials, bvh nodes, face indices */
int rt_host_scene_counts(const RtHostScene* scene, uint32_t* counts6);
int rt_host_scene_copy(const RtHostScene* scene, RtVertex* vertices, RtFace* faces, RtSphere* spheres, RtMaterial* materials,
                       RtBVHNode* nodes, uint32_t* face_indices);
/* Camera (Scene.h:33-71): set pose + viewport, run Camera::Update, read back the 60-B GPUCamera */
int rt_host_scene_set_camera(RtHostScene* scene, const float* position3, float angle_x_deg, float angle_y_deg, float width, float height);
int rt_host_scene_get_camera(const RtHostScene* scene, RtCamera* out, float* angles2 /* x, y; may be NULL */);
int rt_host_scene_add_dirty_flags(RtHostScene* scene, uint32_t flags);
/* device: Scene::Upload (Scene.cpp:182). Afterwards rt_host_scene_gpu() is the GPUScene to pass to the C ABI. */
int rt_host_scene_upload(RtHostScene* scene, RtRngState* d_rng_state);
const RtGPUScene* rt_host_scene_gpu(const RtHostScene* scene);
RtAccel* rt_host_scene_accel(const RtHostScene* scene);

/* ---- CUDARayTracer (RayTracing/RayTracing.h:12-45) ---- */
typedef struct RtHostTracerOptions
{
    const char* bunny_obj;      /* NULL = no bunny */
    const char* env_dds;        /* NULL = no environment map */
    int32_t cornell_box;
    int32_t samples_per_launch; /* 0 = 5 */
    int32_t max_bounces;        /* 0 = 6 */
    uint32_t seed;              /* 0 = wall clock (reference behaviour) */
    int32_t kernel_variant;
} RtHostTracerOptions;

int rt_host_tracer_create(void* surface, void* last_frame_surface, int width, int height, uint64_t pitch,
                          const RtHostTracerOptions* options, RtHostTracer** out);
void rt_host_tracer_destroy(RtHostTracer* tracer);
int rt_host_tracer_process(RtHostTracer* tracer);                 /* CUDARayTracer::Process */
int rt_host_tracer_on_resize(RtHostTracer* tracer, void* surface, void* last_frame_surface, int width, int height, uint64_t pitch);
/* Headless end-to-end call: restart accumulation, `frames` Process() frames, then the float4 surface -> HOST memory
 * (replaces the D3D11 present, main.cpp:683-716). Synchronous. *h2d_bytes (optional) = host->device bytes moved. */
int rt_host_tracer_render_to_host(RtHostTracer* tracer, int frames, void* host_pixels, uint64_t host_pitch, uint64_t* h2d_bytes);
/* Progressive-state checkpoint (the reference keeps this state in RayTracing.cpp:213-234 and main.cpp:634-646 and
 * loses it on exit): {frame index, both float4 surfaces, all XORWOW states}. A tracer of the same size over the same scene that
 * loads it continues the accumulation bit for bit. */
int rt_host_tracer_save_checkpoint(RtHostTracer* tracer, const char* path);
int rt_host_tracer_load_checkpoint(RtHostTracer* tracer, const char* path);
RtHostScene* rt_host_tracer_scene(RtHostTracer* tracer);          /* borrowed */
int rt_host_tracer_frame_index(const RtHostTracer* tracer);
int rt_host_tracer_set_sampling(RtHostTracer* tracer, int samples_per_launch, int max_bounces, int kernel_variant);
unsigned long long rt_host_tracer_ray_count(const RtHostTracer* tracer);

/* ---- multi-GPU renderer (north-star item 5; the reference is single-GPU: what is replicated per GPU is its per-frame
 * protocol, RayTracing/RayTracing.cpp:205-235) ----
 * One process, one host thread per device. Scene, re-layout and env map replicated; device r renders total_spp / n samples of every
 * pixel with cuRAND subsequences [r * states, (r + 1) * states) (device 0 == the single-GPU stream) into an fp32 sum buffer, ONE
 * ncclReduce(sum, fp32) of the frame to device 0 (libnccl.so.2 bound at run time, only when n > 1), divide by the sample count,
 * float4 frame -> HOST memory. `options` as for rt_host_tracer_create (samples_per_launch and kernel_variant unused). */
typedef struct RtHostMultiTracer RtHostMultiTracer;
int rt_host_multi_tracer_create(int width, int height, const RtHostTracerOptions* options, const int* devices, int device_count /* 0 = all */,
                                RtHostMultiTracer** out);
void rt_host_multi_tracer_destroy(RtHostMultiTracer* tracer);
int rt_host_multi_tracer_device_count(const RtHostMultiTracer* tracer);
int rt_host_multi_tracer_set_camera(RtHostMultiTracer* tracer, const float* position3, float angle_x_deg, float angle_y_deg);
int rt_host_multi_tracer_reseed(RtHostMultiTracer* tracer, uint32_t seed);
/* Synchronous. timing_ms3 (optional) = slowest device's launches, the reduce, wall clock of the whole call. */
int rt_host_multi_tracer_render(RtHostMultiTracer* tracer, int total_spp, int spp_per_launch, void* host_pixels, uint64_t host_pitch,
                                double* timing_ms3);
unsigned long long rt_host_multi_tracer_ray_count(const RtHostMultiTracer* tracer);

/* headless sink: binary PPM (P6) from RGBA8 rows, top row first */
int rt_host_write_ppm(const char* path, const uint8_t* rgba8, int width, int height);

/* extra query used by bench/tests: per-pixel (nodes popped, triangles tested) under the reference traversal order */
int rt_count_reference_work(const RtAccel* accel, const RtGPUScene* scene, int width, int height, int with_spheres,
                            uint32_t* d_counts, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200RT_HOST_H */
