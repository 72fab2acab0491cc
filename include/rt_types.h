/*
 * rt_types.h - device data model of the path-trace hot path, byte-for-byte compatible with the
 * reference's RayTracing/GPUScene.h (sizes/offsets: SURVEY.md §8a "[probe]" figures, re-checked
 * against the reference structs by tests/test_layout.py through oracle/_ref).
 *
 * Plain C (no GLM, no cuRAND types) so the structs can cross the C ABI of include/b200rt.h.
 *
 *   RtCamera   <-> GPUCamera        RayTracing/GPUScene.h:11-23   60 B
 *   RtVertex   <-> GPUVertex        RayTracing/GPUScene.h:25-30   32 B
 *   RtFace     <-> GPUFace          RayTracing/GPUScene.h:32-38   16 B   memory order {v0, v2, v1, material}
 *   RtBVHNode  <-> GPUBVHNode       RayTracing/GPUScene.h:40-50   32 B
 *   RtSphere   <-> GeometrySphere   RayTracing/GPUScene.h:59-64   32 B
 *   RtMaterial <-> GPUMaterial      RayTracing/GPUScene.h:66-74   64 B
 *   RtGPUScene <-> GPUScene         RayTracing/GPUScene.h:76-96  136 B
 *   RtRngState <-> curandStateXORWOW  curand_kernel.h:150-156     48 B
 */
#ifndef B200RT_TYPES_H
#define B200RT_TYPES_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
#define RT_ALIGNAS(n) alignas(n)
#define RT_STATIC_ASSERT(c, m) static_assert(c, m)
#else
#define RT_ALIGNAS(n) _Alignas(n)
#define RT_STATIC_ASSERT(c, m) _Static_assert(c, m)
#endif

typedef struct RtVec2 { float x, y; } RtVec2;
typedef struct RtVec3 { float x, y, z; } RtVec3;
typedef struct RtVec4 { float x, y, z, w; } RtVec4;

typedef struct RtCamera
{
    RtVec3 origin;                    /* @0  */
    RtVec2 viewport_worldspace_size;  /* @12 */
    float aspect;                     /* @20 */
    RtVec3 horizontal;                /* @24 */
    RtVec3 vertical;                  /* @36 */
    RtVec3 lower_left_corner;         /* @48 */
} RtCamera;

typedef struct RT_ALIGNAS(16) RtVertex
{
    RtVec3 position;  /* @0  */
    RtVec3 normal;    /* @12 */
    RtVec2 uv;        /* @24 */
} RtVertex;

typedef struct RT_ALIGNAS(16) RtFace
{
    uint32_t v0;        /* @0  */
    uint32_t v2;        /* @4  (sic - the reference declares v2 before v1) */
    uint32_t v1;        /* @8  */
    uint32_t material;  /* @12 */
} RtFace;

typedef struct RT_ALIGNAS(32) RtBVHNode
{
    RtVec3 bounds_min;     /* @0  */
    RtVec3 bounds_max;     /* @12 */
    uint32_t first_index;  /* @24 leaf: first slot in face_indices; inner: left child (right = +1) */
    uint32_t prim_count;   /* @28 >0 => leaf */
} RtBVHNode;

typedef struct RT_ALIGNAS(16) RtSphere
{
    RtVec3 position;  /* @0  */
    float radius;     /* @12 */
    int32_t material; /* @16 */
} RtSphere;

typedef struct RT_ALIGNAS(16) RtMaterial
{
    RtVec4 albedo;           /* @0  */
    RtVec4 emissive;         /* @16 */
    RtVec4 specular;         /* @32 */
    float roughness;         /* @48 */
    float specular_percent;  /* @52 */
    float IOR;               /* @56 */
} RtMaterial;

typedef struct RtRngState  /* curandStateXORWOW */
{
    uint32_t d;
    uint32_t v[5];
    int32_t boxmuller_flag;
    int32_t boxmuller_flag_double;
    float boxmuller_extra;
    double boxmuller_extra_double;
} RtRngState;

typedef struct RtGPUScene
{
    const RtSphere* gpu_spheres;           /* @0   device pointers, caller-owned */
    const RtMaterial* gpu_materials;       /* @8   */
    const RtBVHNode* gpu_bvh_nodes;        /* @16  */
    const uint32_t* gpu_bvh_face_indices;  /* @24  */
    const RtVertex* gpu_vertices;          /* @32  */
    const RtFace* gpu_faces;               /* @40  */
    int32_t sphere_count;                  /* @48  */
    int32_t material_count;                /* @52  */
    RtRngState* rng_state;                 /* @56  */
    uint64_t environment_cubemap_tex;      /* @64  cudaTextureObject_t */
    RtCamera camera;                       /* @72  */
} RtGPUScene;

RT_STATIC_ASSERT(sizeof(RtCamera) == 60, "GPUCamera is 60 B");
RT_STATIC_ASSERT(offsetof(RtCamera, horizontal) == 24 && offsetof(RtCamera, lower_left_corner) == 48, "GPUCamera offsets");
RT_STATIC_ASSERT(sizeof(RtVertex) == 32 && offsetof(RtVertex, normal) == 12 && offsetof(RtVertex, uv) == 24, "GPUVertex");
RT_STATIC_ASSERT(sizeof(RtFace) == 16 && offsetof(RtFace, v2) == 4 && offsetof(RtFace, v1) == 8, "GPUFace order v0,v2,v1");
RT_STATIC_ASSERT(sizeof(RtBVHNode) == 32 && offsetof(RtBVHNode, first_index) == 24 && offsetof(RtBVHNode, prim_count) == 28, "GPUBVHNode");
RT_STATIC_ASSERT(sizeof(RtSphere) == 32 && offsetof(RtSphere, material) == 16, "GeometrySphere");
RT_STATIC_ASSERT(sizeof(RtMaterial) == 64 && offsetof(RtMaterial, roughness) == 48, "GPUMaterial");
RT_STATIC_ASSERT(sizeof(RtRngState) == 48, "curandState is 48 B");
RT_STATIC_ASSERT(sizeof(RtGPUScene) == 136 && offsetof(RtGPUScene, rng_state) == 56 && offsetof(RtGPUScene, camera) == 72, "GPUScene");

#endif /* B200RT_TYPES_H */
