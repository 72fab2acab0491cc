// TEST INFRASTRUCTURE (oracle/_ref). C API over the REFERENCE'S OWN code compiled host-side:
//   * BVH::Calculate                      /root/reference/RayTracing/BVH.cpp:8        (verbatim TU)
//   * BVHRayHit / GetRayHit / ray_color / raytracing_kernel_main
//                                         /root/reference/RayTracing/main_raytracing.cu:33-200
//                                         (textually included below; one mechanical edit, build_ref.sh)
//   * curand_init / curand / curand_uniform  toolkit curand_kernel.h (host-qualified)
// What is RESTATED here with the reference's own GLM (because Scene.cpp / RayTracing.cpp need
// <windows.h>, Assimp, ImGui - SURVEY.md §8(c)):
//   * scene assembly   RayTracing/RayTracing.cpp:33-69,79-203 ; RayTracing/Scene.cpp:46-73,75-139
//   * camera           RayTracing/Scene.cpp:15-36 ; RayTracing/Math.h:63-70
//   * Assimp root pre-transform  utils/AssimpLoader.cpp:8-13,46-48
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs load this.
#include "host_pre.h"
#include RT_HOST_SRC
#include "RayTracing/BVH.h"
#include "glm/gtc/matrix_transform.hpp"
#include <omp.h>

using namespace RayTracing;
using namespace Math;

#define REF_API extern "C" __attribute__((visibility("default")))

// ---------------------------------------------------------------------------------------------
// Layout facts (SURVEY.md §8a "[probe]" numbers) - checked by tests against include/rt_types.h
// ---------------------------------------------------------------------------------------------
namespace { struct CameraPeek : GPUCamera { static size_t off(int i) {
    switch (i) { case 0: return offsetof(CameraPeek, origin); case 1: return offsetof(CameraPeek, viewport_worldspace_size);
    case 2: return offsetof(CameraPeek, aspect); case 3: return offsetof(CameraPeek, horizontal);
    case 4: return offsetof(CameraPeek, vertical); default: return offsetof(CameraPeek, lower_left_corner); } } }; }

REF_API uint32_t ref_sizeof(int which)
{
    switch (which)
    {
    case 0: return sizeof(GPUVertex);
    case 1: return sizeof(GPUFace);
    case 2: return sizeof(GPUBVHNode);
    case 3: return sizeof(GPUMaterial);
    case 4: return sizeof(GeometrySphere);
    case 5: return sizeof(GPUCamera);
    case 6: return sizeof(GPUScene);
    case 7: return sizeof(curandState);
    case 8: return offsetof(GPUScene, camera);
    case 9: return offsetof(GPUScene, rng_state);
    case 10: return offsetof(GPUScene, environment_cubemap_tex);
    case 11: return offsetof(GPUFace, v2);
    case 12: return offsetof(GPUFace, v1);
    case 13: return offsetof(GPUBVHNode, first_index);
    case 14: return offsetof(GPUMaterial, roughness);
    case 15: return offsetof(GeometrySphere, material);
    case 16: return (uint32_t)CameraPeek::off(3);
    case 17: return (uint32_t)CameraPeek::off(5);
    default: return 0;
    }
}

REF_API int ref_num_threads() { return omp_get_max_threads(); }
// torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU-baseline legs ask for all host cores explicitly
REF_API void ref_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

// ---------------------------------------------------------------------------------------------
// BVH: the reference builder, verbatim
// ---------------------------------------------------------------------------------------------
REF_API int ref_bvh_build(const void* vertices, uint32_t vertex_count, const void* faces, uint32_t face_count,
                          void* out_nodes /* 32 B x (2*face_count-1) */, uint32_t* out_nodes_used,
                          uint32_t* out_face_indices /* face_count */)
{
    BVH bvh;
    bvh.Calculate(std::span<const GPUVertex>((const GPUVertex*)vertices, vertex_count),
                  std::span<const GPUFace>((const GPUFace*)faces, face_count));
    const auto nodes = bvh.GetGPUBVHNodes();
    memcpy(out_nodes, nodes.data(), nodes.size_bytes());
    *out_nodes_used = (uint32_t)nodes.size();
    memcpy(out_face_indices, bvh.GetFaceIndices().data(), bvh.GetFaceIndices().size_bytes());
    return 0;
}

// DebugDraw symbols BVH.cpp links against (UI overlay - out of scope)
namespace DebugDraw
{
    void DrawLine(glm::vec3, glm::vec3, uint32_t, float) {}
    void DrawAABB(const Math::AABB&, uint32_t, float) {}
}

// ---------------------------------------------------------------------------------------------
// Scene assembly, restated with the reference's GLM (host vectors of the reference PODs)
// ---------------------------------------------------------------------------------------------
namespace
{
    struct Material : GPUMaterial  // RayTracing/Scene.h:74-85
    {
        Material(vec3 albedo_, vec3 emissive_) { albedo = vec4(albedo_, 1); emissive = vec4(emissive_, 1); }
        Material(vec3 albedo_) : Material(albedo_, vec3(0)) {}
        Material() : Material(vec3(0)) {}
    };

    struct RefScene
    {
        std::vector<GeometrySphere> spheres;
        std::vector<GPUMaterial> materials;
        std::vector<GPUFace> faces;
        std::vector<GPUVertex> vertices;

        void AddTriangle(vec3 a, vec3 b, vec3 c, int material = 0)  // Scene.cpp:46-67
        {
            const vec3 normal = glm::normalize(glm::cross(c - b, a - b));
            const uint32_t index0 = (uint32_t)vertices.size();
            GPUVertex va; va.position = a; va.normal = normal; va.uv = vec2(0);
            GPUVertex vb; vb.position = b; vb.normal = normal; vb.uv = vec2(0);
            GPUVertex vc; vc.position = c; vc.normal = normal; vc.uv = vec2(0);
            vertices.push_back(va); vertices.push_back(vb); vertices.push_back(vc);
            faces.push_back(GPUFace{ index0, index0 + 1, index0 + 2, (uint32_t)material });
        }
        void AddQuad(vec3 a, vec3 b, vec3 c, vec3 d, int material = 0) { AddTriangle(a, b, c, material); AddTriangle(c, d, a, material); }
        void AddSphere(vec3 position, float radius, int material = 0) { spheres.push_back(GeometrySphere{ position, radius, material }); }
        uint32_t AddMaterial(Material m) { materials.push_back(m); return (uint32_t)materials.size() - 1; }

        // Scene.cpp:75-132 with the aiMesh replaced by plain arrays
        void AddLoadedMesh(const float* pos, const float* nrm, uint32_t nv, const uint32_t* idx, uint32_t nf,
                           const mat4& transform, const mat4& node_transform, int default_material)
        {
            const uint32_t index_offset = (uint32_t)vertices.size();
            auto mesh_transform = transform * node_transform;
            for (uint32_t v = 0; v < nv; v++)
            {
                GPUVertex vertex;
                vertex.position = mesh_transform * glm::vec4(pos[3 * v], pos[3 * v + 1], pos[3 * v + 2], 1.0f);
                vertex.normal = mesh_transform * glm::vec4(nrm[3 * v], nrm[3 * v + 1], nrm[3 * v + 2], 0.0f);
                vertex.uv = glm::vec2(0.0f, 0.0f);
                vertices.push_back(vertex);
            }
            for (uint32_t f = 0; f < nf; f++)
            {
                GPUFace gpu_face;
                gpu_face.v0 = idx[3 * f] + index_offset;
                gpu_face.v1 = idx[3 * f + 1] + index_offset;
                gpu_face.v2 = idx[3 * f + 2] + index_offset;
                gpu_face.material = default_material;
                faces.push_back(gpu_face);
                const float* p0 = pos + 3 * idx[3 * f]; const float* p1 = pos + 3 * idx[3 * f + 1]; const float* p2 = pos + 3 * idx[3 * f + 2];
                glm::vec3 v0 = mesh_transform * glm::vec4(p0[0], p0[1], p0[2], 1);
                glm::vec3 v1 = mesh_transform * glm::vec4(p1[0], p1[1], p1[2], 1);
                glm::vec3 v2 = mesh_transform * glm::vec4(p2[0], p2[1], p2[2], 1);
                AddTriangle(v0, v1, v2, default_material);
            }
        }
    };

    // utils/AssimpLoader.cpp:46-48 + :10-13 : aiMatrix4x4::RotationX(-(float)M_PI/2) (Assimp 5.2.4
    // matrix4x4.inl: b2=c3=cos(a), c2=sin(a), b3=-sin(a), row-major) reinterpreted + transposed into GLM.
    mat4 AssimpRootTransform()
    {
        const float a = -(float)M_PI / 2;
        float m[16] = { 1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1 };  // row-major a1..d4
        m[5] = m[10] = std::cos(a);
        m[9] = std::sin(a);
        m[6] = -m[9];
        glm::mat4 result; memcpy(&result, m, sizeof(m));
        return glm::transpose(result);
    }
}

REF_API void* ref_scene_new() { return new RefScene(); }
REF_API void ref_scene_free(void* s) { delete (RefScene*)s; }

REF_API void ref_scene_setup_cornell(void* sp)  // RayTracing.cpp:79-203 (camera yaw 180 is the caller's business)
{
    RefScene* scene = (RefScene*)sp;
    struct Q { vec3 A, B, C, D; vec3 albedo, emissive; };
    const Q quads[6] = {
        { vec3(-12.6f, -12.6f, 25.0f), vec3(12.6f, -12.6f, 25.0f), vec3(12.6f, 12.6f, 25.0f), vec3(-12.6f, 12.6f, 25.0f), vec3(0.7f, 0.7f, 0.7f), vec3(0.0f) },
        { vec3(-12.6f, -12.45f, 25.0f), vec3(12.6f, -12.45f, 25.0f), vec3(12.6f, -12.45f, 15.0f), vec3(-12.6f, -12.45f, 15.0f), vec3(0.7f, 0.7f, 0.7f), vec3(0.0f) },
        { vec3(-12.6f, 12.5f, 25.0f), vec3(12.6f, 12.5f, 25.0f), vec3(12.6f, 12.5f, 15.0f), vec3(-12.6f, 12.5f, 15.0f), vec3(0.7f, 0.7f, 0.7f), vec3(0.0f) },
        { vec3(-12.5f, -12.6f, 25.0f), vec3(-12.5f, -12.6f, 15.0f), vec3(-12.5f, 12.6f, 15.0f), vec3(-12.5f, 12.6f, 25.0f), vec3(0.1f, 0.7f, 0.1f), vec3(0.0f) },
        { vec3(12.5f, -12.6f, 25.0f), vec3(12.5f, -12.6f, 15.0f), vec3(12.5f, 12.6f, 15.0f), vec3(12.5f, 12.6f, 25.0f), vec3(0.7f, 0.1f, 0.1f), vec3(0.0f) },
        { vec3(-5.0f, 12.4f, 22.5f), vec3(5.0f, 12.4f, 22.5f), vec3(5.0f, 12.4f, 17.5f), vec3(-5.0f, 12.4f, 17.5f), vec3(0), vec3(1.0f, 0.9f, 0.7f) * 20.0f },
    };
    for (const Q& q : quads)
        scene->AddQuad(q.A, q.B, q.C, q.D, scene->AddMaterial({ q.albedo, q.emissive }));

    auto ball = [&](vec3 albedo, float pct, vec4 spec, float rough, vec3 pos, float radius) {
        Material m(albedo); m.specular_percent = pct; m.specular = spec; m.roughness = rough;
        scene->AddSphere(pos, radius, scene->AddMaterial(m));
    };
    ball(vec3(0.9f, 0.9f, 0.50f), 0.5f, vec4(0.9f, 0.9f, 0.9f, 0), 0.2f, vec3(-9.0f, -9.5f, 20.0f), 3);
    ball(vec3(0.9f, 0.5f, 0.90f), 0.3f, vec4(0.9f, 0.9f, 0.9f, 0), 0.2f, vec3(0.0f, -9.5f, 20.0f), 3);
    ball(vec3(0.0f, 0.0f, 1.0f), 0.5f, vec4(1.0f, 0.0f, 0.0f, 0), 0.4f, vec3(9.0f, -9.5f, 20.0f), 3);
    const float rough[5] = { 0.0f, 0.25f, 0.5f, 0.75f, 0.97f };
    const float xs[5] = { -10.0f, -5.0f, 0.0f, 5.0f, 10.0f };
    for (int i = 0; i < 5; i++)
        ball(vec3(1), 1.0f, vec4(0.3f, 1.0f, 0.3f, 0.0f), rough[i], vec3(xs[i], 0.0f, 23.0f), 1.75f);
}

REF_API void ref_scene_setup_bunny(void* sp, const float* pos, const float* nrm, uint32_t nv, const uint32_t* idx, uint32_t nf)
{   // RayTracing.cpp:33-69
    RefScene* scene = (RefScene*)sp;
    auto m = glm::translate(mat4(1), vec3(30, -18, 20));
    m = glm::rotate(m, -(float)M_PI, vec3(0, 1, 0));
    m = glm::rotate(m, (float)M_PI / 2, vec3(1, 0, 0));
    m = glm::scale(m, vec3(150.0f));
    Material material(vec3(1));
    material.specular_percent = 0.5f;
    material.roughness = 0.8f;
    material.specular = vec4(0.3f, 1.0f, 0.3f, 0.0f);
    scene->AddLoadedMesh(pos, nrm, nv, idx, nf, m, AssimpRootTransform(), scene->AddMaterial(material));
    {
        const vec3 offset(20, 0, 0);
        const vec3 scale(50, 1, 50);
        vec3 A = scale * vec3(-1.0f, -12.45f, 1.0f) + offset;
        vec3 B = scale * vec3(1.0f, -12.45f, 1.0f) + offset;
        vec3 C = scale * vec3(1.0f, -12.45f, -1.0f) + offset;
        vec3 D = scale * vec3(-1.0f, -12.45f, -1.0f) + offset;
        auto material2 = scene->AddMaterial({ vec3(0.7f, 0.7f, 0.7f), vec3(0.0f, 0.0f, 0.0f) });
        scene->AddQuad(A, B, C, D, material2);
    }
    {
        auto material3 = scene->AddMaterial({ vec3(0), vec3(0.3f, 0.9f, 0.7f) * 10.0f });
        scene->AddSphere(vec3(30, 10, 40), 8, material3);
    }
}

// generic mesh with a caller-supplied column-major transform (scaled-mesh configs)
REF_API void ref_scene_add_mesh(void* sp, const float* pos, const float* nrm, uint32_t nv, const uint32_t* idx, uint32_t nf,
                                const float* transform16, int material)
{
    mat4 t; memcpy(&t, transform16, 64);
    ((RefScene*)sp)->AddLoadedMesh(pos, nrm, nv, idx, nf, t, AssimpRootTransform(), material);
}

REF_API void ref_scene_counts(void* sp, uint32_t* out4)
{
    RefScene* s = (RefScene*)sp;
    out4[0] = (uint32_t)s->vertices.size(); out4[1] = (uint32_t)s->faces.size();
    out4[2] = (uint32_t)s->spheres.size(); out4[3] = (uint32_t)s->materials.size();
}

REF_API void ref_scene_copy(void* sp, void* vertices, void* faces, void* spheres, void* materials)
{
    RefScene* s = (RefScene*)sp;
    memcpy(vertices, s->vertices.data(), s->vertices.size() * sizeof(GPUVertex));
    memcpy(faces, s->faces.data(), s->faces.size() * sizeof(GPUFace));
    memcpy(spheres, s->spheres.data(), s->spheres.size() * sizeof(GeometrySphere));
    memcpy(materials, s->materials.data(), s->materials.size() * sizeof(GPUMaterial));
}

// ---------------------------------------------------------------------------------------------
// Camera (Scene.cpp:15-36, Math.h:63-70), restated with GLM
// ---------------------------------------------------------------------------------------------
namespace
{
    struct RefCamera : GPUCamera
    {
        void Update(vec3 position, float angle_x, float angle_y, vec2 viewport_size, float fov_y)
        {
            origin = position;
            aspect = viewport_size.x / viewport_size.y;
            mat4 transform = Math::ComposeMatrix(origin, quat(vec3(glm::radians(angle_x), glm::radians(angle_y), 0)), vec3(1));
            mat4 projection = glm::perspectiveRH(glm::radians(fov_y), aspect, 1.0f, 1000.0f);
            auto inv_proj = glm::inverse(projection);
            auto lower_left_corner4 = inv_proj * vec4(-1, -1, -1, 1);
            auto upper_right_corner4 = inv_proj * vec4(1, 1, -1, 1);
            lower_left_corner = lower_left_corner4 / lower_left_corner4.w;
            vec3 upper_right_corner = upper_right_corner4 / upper_right_corner4.w;
            viewport_worldspace_size = upper_right_corner - lower_left_corner;
            horizontal = transform * vec4(viewport_worldspace_size.x, 0, 0, 0);
            vertical = transform * vec4(0, viewport_worldspace_size.y, 0, 0);
            lower_left_corner = transform * vec4(lower_left_corner, 1);
        }
    };
}

REF_API void ref_camera_update(const float* position3, float angle_x_deg, float angle_y_deg, float width, float height,
                               void* out_gpu_camera /* 60 B */)
{
    RefCamera cam;
    cam.Update(vec3(position3[0], position3[1], position3[2]), angle_x_deg, angle_y_deg, vec2(width, height), 90.0f);
    memcpy(out_gpu_camera, static_cast<GPUCamera*>(&cam), sizeof(GPUCamera));
}

// ---------------------------------------------------------------------------------------------
// Tracing through the UNMODIFIED BVHRayHit / GetRayHit (main_raytracing.cu:33-109)
// ---------------------------------------------------------------------------------------------
// The reference records no primitive id. It does copy `*scene.GetMaterial(face.material)` into the
// hit (main_raytracing.cu:67,100), so the id query runs on a COPY of the faces whose material field
// is the face index and a material table whose albedo encodes that index: the unmodified function
// then reports which face won (first-visited on exact ties, main_raytracing.cu:62).
namespace
{
    struct IdScene
    {
        std::vector<GPUFace> faces;
        std::vector<GeometrySphere> spheres;
        std::vector<GPUMaterial> materials;
        GPUScene scene;
        uint32_t face_count = 0;

        IdScene(const void* vertices, const void* faces_in, uint32_t nf, const void* nodes, const uint32_t* face_indices,
                const void* spheres_in, uint32_t ns)
        {
            face_count = nf;
            faces.assign((const GPUFace*)faces_in, (const GPUFace*)faces_in + nf);
            materials.resize(size_t(nf) + ns);
            for (uint32_t i = 0; i < nf + ns; i++)
                materials[i].albedo = vec4(float(i & 0xFFFFu), float(i >> 16), 0, 0);
            for (uint32_t i = 0; i < nf; i++) faces[i].material = i;
            if (ns) spheres.assign((const GeometrySphere*)spheres_in, (const GeometrySphere*)spheres_in + ns);
            for (uint32_t i = 0; i < ns; i++) spheres[i].material = int(nf + i);
            scene.gpu_spheres = spheres.data();
            scene.gpu_materials = materials.data();
            scene.gpu_bvh_nodes = (const GPUBVHNode*)nodes;
            scene.gpu_bvh_face_indices = face_indices;
            scene.gpu_vertices = (const GPUVertex*)vertices;
            scene.gpu_faces = faces.data();
            scene.sphere_count = (int)ns;
            scene.material_count = (int)materials.size();
        }

        // id: face index, 0x80000000|sphere index, or 0xFFFFFFFF on miss
        void Query(const Ray& ray, bool with_spheres, uint32_t& id, float& t)
        {
            HitData hit;
            bool found;
            if (with_spheres) found = GetRayHit(ray, scene, hit);
            else { hit.distance = 1e30f; found = BVHRayHit(ray, scene, scene.gpu_bvh_nodes, hit); }
            if (!found) { id = 0xFFFFFFFFu; t = 1e30f; return; }
            const uint32_t m = uint32_t(hit.material.albedo.x) | (uint32_t(hit.material.albedo.y) << 16);
            id = m < face_count ? m : (0x80000000u | (m - face_count));
            t = hit.distance;
        }
    };
}

// Primary rays at pixel centres: uv = (x+0.5, y+0.5)/(W,H), ray = GPUCamera::GetRay.
REF_API int ref_primary_hits(const void* vertices, const void* faces, uint32_t face_count, const void* nodes,
                             const uint32_t* face_indices, const void* spheres, uint32_t sphere_count,
                             const void* gpu_camera60, int width, int height, int with_spheres,
                             uint32_t* out_id, float* out_t)
{
    IdScene ids(vertices, faces, face_count, nodes, face_indices, spheres, with_spheres ? sphere_count : 0);
    GPUCamera camera; memcpy(&camera, gpu_camera60, sizeof(GPUCamera));
#pragma omp parallel for schedule(dynamic, 8)
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++)
        {
            const vec2 pixel_coords(x, y);
            const auto uv = (pixel_coords + vec2(0.5f, 0.5f)) / vec2(width, height);
            Ray ray = camera.GetRay(uv);
            ids.Query(ray, with_spheres != 0, out_id[size_t(y) * width + x], out_t[size_t(y) * width + x]);
        }
    return 0;
}

// Arbitrary rays (6 floats each: origin, direction - direction used as given, as ray_color does)
REF_API int ref_trace_rays(const void* vertices, const void* faces, uint32_t face_count, const void* nodes,
                           const uint32_t* face_indices, const void* spheres, uint32_t sphere_count,
                           const float* rays, uint64_t ray_count, int with_spheres, uint32_t* out_id, float* out_t)
{
    IdScene ids(vertices, faces, face_count, nodes, face_indices, spheres, with_spheres ? sphere_count : 0);
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < (int64_t)ray_count; i++)
    {
        Ray ray(vec3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]), vec3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]));
        ids.Query(ray, with_spheres != 0, out_id[i], out_t[i]);
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// RNG: the toolkit's XORWOW, host-side (this IS the reference generator, Random.cu:3-8, Random.h:19)
// ---------------------------------------------------------------------------------------------
REF_API void ref_rng_init(void* states48, uint64_t count, uint32_t seed)  // initRNG: curand_init(seed, tid, 0)
{
    curandState* s = (curandState*)states48;
#pragma omp parallel for schedule(static)
    for (int64_t tid = 0; tid < (int64_t)count; tid++) curand_init(seed, (unsigned long long)tid, 0, &s[tid]);
}
REF_API void ref_rng_init_one(void* state48, uint64_t seed, uint64_t subsequence, uint64_t offset)
{
    curand_init(seed, subsequence, offset, (curandState*)state48);
}
REF_API void ref_rng_draw_u32(void* state48, uint32_t n, uint32_t* out) { for (uint32_t i = 0; i < n; i++) out[i] = curand((curandState*)state48); }
REF_API void ref_rng_draw_uniform(void* state48, uint32_t n, float* out) { for (uint32_t i = 0; i < n; i++) out[i] = curand_uniform((curandState*)state48); }
REF_API void ref_rng_point_on_sphere(void* state48, float* out3)
{
    RNG rng((curandState*)state48);
    vec3 p = rng.GetRandomPointOnSphere();
    out3[0] = p.x; out3[1] = p.y; out3[2] = p.z;
}

// ---------------------------------------------------------------------------------------------
// Full frame: the reference kernel body run once per pixel on the host
// (raytracing_kernel_main, main_raytracing.cu:162-200: 5 spp x <=6 bounces, running-mean blend)
// ---------------------------------------------------------------------------------------------
REF_API int ref_render_frame(const void* vertices, const void* faces, const void* nodes, const uint32_t* face_indices,
                             const void* spheres, uint32_t sphere_count, const void* materials, uint32_t material_count,
                             const void* gpu_camera60, const float* cubemap_level0, int cubemap_size,
                             void* rng_states48, void* surface, void* surface_last_frame,
                             int width, int height, uint64_t pitch, int frame_index, int row_begin, int row_end)
{
    OracleCubemap cm; cm.texels = cubemap_level0; cm.size = cubemap_size;
    RenderData rd{};
    rd.surface = (unsigned char*)surface; rd.surface_last_frame = (unsigned char*)surface_last_frame;
    rd.width = width; rd.height = height; rd.pitch = pitch; rd.frame_index = frame_index;
    rd.scene.gpu_spheres = (const GeometrySphere*)spheres; rd.scene.gpu_materials = (const GPUMaterial*)materials;
    rd.scene.gpu_bvh_nodes = (const GPUBVHNode*)nodes; rd.scene.gpu_bvh_face_indices = face_indices;
    rd.scene.gpu_vertices = (const GPUVertex*)vertices; rd.scene.gpu_faces = (const GPUFace*)faces;
    rd.scene.sphere_count = (int)sphere_count; rd.scene.material_count = (int)material_count;
    rd.scene.rng_state = (curandState*)rng_states48;
    rd.scene.environment_cubemap_tex = (cudaTextureObject_t)(uintptr_t)&cm;
    memcpy(&rd.scene.camera, gpu_camera60, sizeof(GPUCamera));
    if (row_end <= 0 || row_end > height) row_end = height;
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = row_begin; y < row_end; y++)
    {
        blockDim = dim3(16, 16, 1); gridDim = dim3((width + 15) / 16, (height + 15) / 16, 1);
        for (int x = 0; x < width; x++)
        {
            blockIdx.x = x / 16; blockIdx.y = y / 16; blockIdx.z = 0;
            threadIdx.x = x % 16; threadIdx.y = y % 16; threadIdx.z = 0;
            raytracing_kernel_main(rd);
        }
    }
    return 0;
}
