// Host-side scene: the reference's Scene / Camera / Material surface (RayTracing/Scene.h:19-129),
// kept by name so code written against the reference compiles against this with the same meaning.
// Differences: AddLoadedScene takes a plain Loader::LoadedMesh instead of an
// AssimpScene; Update(dt) has no keyboard polling (Win32); Upload() additionally builds the
// B200 re-layout (RtAccel) next to the reference-layout device arrays.
#pragma once
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include "Math.h"
#include "MeshLoader.h"
#include "rt_types.h"

struct RtAccel;   // include/b200rt.h
struct RtEnvMap;  // include/b200rt.h

namespace RayTracing
{
    using namespace Math;

    enum class DirtyFlagValue : uint32_t { Samples = 1 << 0, SceneMemory = 1 << 1, BVH = 1 << 2 };  // Scene.h:19-24
    typedef uint32_t DirtyFlags;

    struct Scene;
    class BVH;

    struct Camera : public RtCamera  // Scene.h:33-71
    {
        Camera(Scene& scene_) : RtCamera{}, scene(scene_) { viewport_worldspace_size = RtVec2{ 1, 1 }; aspect = 1.0f; }
        vec2 GetViewportSize() const { return viewport_size; }
        void SetViewportSize(vec2 value) { viewport_size = value; }
        vec3 GetPosition() const { return vec3(origin); }
        void SetPosition(vec3 value) { origin = value; }
        float GetXAngle() const { return angle_x; }
        void SetXAndle(float value) { angle_x = value; }   // (sic) reference spelling, Scene.h:44
        float GetYAngle() const { return angle_y; }
        void SetYAndle(float value) { angle_y = value; }   // (sic) Scene.h:47
        void Update();                                     // Scene.cpp:15-36
        vec3 GetForward() const { return -transform[2].xyz(); }
        vec3 GetBackward() const { return transform[2].xyz(); }
        vec3 GetRight() const { return transform[0].xyz(); }
        vec3 GetLeft() const { return -transform[0].xyz(); }
        vec3 GetUp() const { return transform[1].xyz(); }
        vec3 GetDown() const { return -transform[1].xyz(); }
        const mat4& GetProjectionMatrix() const { return projection; }
        const mat4& GetTransform() const { return transform; }
    private:
        Scene& scene;
        vec2 viewport_size = vec2(1.0f);
        float angle_x = 0, angle_y = 0, fov_y = 90;
        mat4 transform = mat4(1.0f), projection = mat4(1.0f);
    };

    struct Material : public RtMaterial  // Scene.h:74-85 + GPUScene.h:66-74 defaults
    {
        Material(vec3 albedo_, vec3 emissive_)
        {
            albedo = vec4(albedo_, 1); emissive = vec4(emissive_, 1); specular = vec4(0.0f);
            roughness = 0.9f; specular_percent = 0.0f; IOR = 1.0f;
        }
        Material(vec3 albedo_) : Material(albedo_, vec3(0.0f)) {}
        Material() : Material(vec3(0.0f)) {}
    };

    struct Scene : public RtGPUScene  // Scene.h:87-129
    {
        // env_dds_path: the reference hard-codes L"data/sunset_uncompressed.dds" (Scene.cpp:40); empty = no env map.
        explicit Scene(const std::string& env_dds_path = std::string());
        ~Scene();
        Scene(const Scene&) = delete;
        Scene& operator=(const Scene&) = delete;

        void AddSphere(vec3 position, float radius, int material = 0);
        void AddTriangle(vec3 a, vec3 b, vec3 c, int material = 0);
        void AddQuad(vec3 a, vec3 b, vec3 c, vec3 d, int material = 0) { AddTriangle(a, b, c, material); AddTriangle(c, d, a, material); }
        uint32_t AddMaterial(Material material);
        void Update(float dt);
        void Upload(RtRngState* rng_state);  // device side; defined in SceneUpload.cpp (needs the CUDA library)
        void AddLoadedScene(const Loader::LoadedMesh& mesh, const mat4& transform = mat4(1.0f), int default_material = 0);
        void BuildBVH();                      // host part of Upload()'s BVH-dirty branch (Scene.cpp:189-191)
        // Re-copies the per-frame mutable state (spheres, materials) into the existing device arrays; returns bytes copied.
        // (The camera travels by value inside GPUScene, main_raytracing.cu:208-211.) Headless e2e path only.
        size_t RefreshDynamic();

        Camera& GetCamera() { return camera; }
        void AddDirtyFlags(DirtyFlags flags = ~0u) { dirty_flags |= flags; if (flags & static_cast<DirtyFlags>(DirtyFlagValue::BVH)) bvh_current = false; }
        void AddDirtyFlag(DirtyFlagValue flag) { AddDirtyFlags(static_cast<DirtyFlags>(flag)); }
        DirtyFlags GetDirtyFlags() const { return dirty_flags; }
        bool IsFlagDirty(DirtyFlagValue flag) const { return dirty_flags & static_cast<DirtyFlags>(flag); }

        // host arrays (read-only views for tests / exporters)
        const std::vector<RtSphere>& Spheres() const { return spheres; }
        const std::vector<Material>& Materials() const { return materials; }
        const std::vector<RtFace>& Faces() const { return faces; }
        const std::vector<RtVertex>& Vertices() const { return vertices; }
        const BVH& GetBVH() const { return *bvh; }
        RtAccel* GetAccel() const { return accel; }
        const std::string& EnvPath() const { return env_path; }

    private:
        Camera camera;
        std::vector<RtSphere> spheres;
        std::vector<Material> materials;
        std::vector<RtFace> faces;
        std::vector<RtVertex> vertices;
        std::unique_ptr<BVH> bvh;
        DirtyFlags dirty_flags = ~0u;
        bool bvh_current = false;  // BuildBVH() ran after the last geometry change (Upload() then does not build again)
        std::string env_path;
        // device side (owned; released in ~Scene through the C ABI)
        void* d_spheres = nullptr; void* d_materials = nullptr; void* d_nodes = nullptr;
        void* d_face_indices = nullptr; void* d_vertices = nullptr; void* d_faces = nullptr;
        RtAccel* accel = nullptr;
        RtEnvMap* env = nullptr;
        void ReleaseDevice();
    };
}
