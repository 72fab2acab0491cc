// Device side of Scene: construction (env map), Upload() and teardown.
// Follows RayTracing/Scene.cpp:38-44 (ctor), :182-234 (Upload): blocking copies on the default
// stream, once per scene change. New here: after the reference-layout arrays are uploaded, the
// B200 re-layout (RtAccel) is built from the same host arrays and bound to the node pointer so the
// reference-named raytracing_process() finds it.
#include <cuda_runtime.h>
#include "BVH.h"
#include "Scene.h"
#include "b200rt.h"

namespace RayTracing
{
    namespace
    {
        void Check(cudaError_t e, const char* what)
        {
            if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
        }
        template <class T> void UploadArray(void*& d, const T* src, size_t count, const char* what)
        {
            if (d) { cudaFree(d); d = nullptr; }
            const size_t bytes = count * sizeof(T);
            if (cudaMalloc(&d, bytes ? bytes : 16) != cudaSuccess) throw std::runtime_error("allocation failed");  // utils/CUDAHelper.h:120-126
            if (bytes) Check(cudaMemcpy(d, src, bytes, cudaMemcpyHostToDevice), what);
        }
    }

    Scene::Scene(const std::string& env_dds_path) : RtGPUScene{}, camera(*this), env_path(env_dds_path)
    {
        bvh = std::make_unique<BVH>();
    }

    Scene::~Scene() { ReleaseDevice(); }

    void Scene::ReleaseDevice()
    {
        if (accel) { rt_accel_destroy(accel); accel = nullptr; }
        if (env) { rt_envmap_destroy(env); env = nullptr; }
        void** arrays[6] = { &d_spheres, &d_materials, &d_nodes, &d_face_indices, &d_vertices, &d_faces };
        for (void** a : arrays) if (*a) { cudaFree(*a); *a = nullptr; }
    }

    void Scene::Upload(RtRngState* rng_state_)
    {
        RtGPUScene::camera = camera;  // slice Camera -> GPUCamera (Scene.cpp:184)
        rng_state = rng_state_;
        if (!env && !env_path.empty())
        {
            if (rt_envmap_create_from_file(env_path.c_str(), &env) != RT_OK) throw std::runtime_error(rt_last_error());
        }
        environment_cubemap_tex = env ? rt_envmap_texture(env) : 0;

        const bool bvh_dirty = IsFlagDirty(DirtyFlagValue::BVH), memory_dirty = IsFlagDirty(DirtyFlagValue::SceneMemory);
        if (bvh_dirty)
        {
            BuildBVH();
            if (d_nodes) rt_accel_unbind(d_nodes);
            UploadArray(d_nodes, bvh->GetGPUBVHNodes().data(), bvh->GetGPUBVHNodes().size(), "upload bvh nodes");
            UploadArray(d_face_indices, bvh->GetFaceIndices().data(), bvh->GetFaceIndices().size(), "upload face indices");
            gpu_bvh_nodes = static_cast<const RtBVHNode*>(d_nodes);
            gpu_bvh_face_indices = static_cast<const uint32_t*>(d_face_indices);
        }
        if (memory_dirty)
        {
            UploadArray(d_spheres, spheres.data(), spheres.size(), "upload spheres");
            UploadArray(d_materials, static_cast<const RtMaterial*>(materials.data()), materials.size(), "upload materials");
            UploadArray(d_vertices, vertices.data(), vertices.size(), "upload vertices");
            UploadArray(d_faces, faces.data(), faces.size(), "upload faces");
            sphere_count = (int)spheres.size(); material_count = (int)materials.size();
            gpu_spheres = static_cast<const RtSphere*>(d_spheres); gpu_materials = static_cast<const RtMaterial*>(d_materials);
            gpu_vertices = static_cast<const RtVertex*>(d_vertices); gpu_faces = static_cast<const RtFace*>(d_faces);
        }
        // the re-layout depends on nodes, faces and vertices only, and every geometry change raises the BVH flag (Scene.cpp AddTriangle /
        // AddLoadedScene): a material or sphere edit (SceneMemory alone) does not rebuild it
        if (bvh_dirty && !faces.empty())
        {
            if (accel) { rt_accel_destroy(accel); accel = nullptr; }
            const auto nodes = bvh->GetGPUBVHNodes();
            if (rt_accel_create(nodes.data(), (uint32_t)nodes.size(), bvh->GetFaceIndices().data(), faces.data(), (uint32_t)faces.size(),
                                vertices.data(), (uint32_t)vertices.size(), 0, &accel) != RT_OK)
                throw std::runtime_error(rt_last_error());
            rt_accel_bind(d_nodes, accel);
        }
        dirty_flags = 0;
    }

    size_t Scene::RefreshDynamic()
    {
        if (!d_spheres || !d_materials || sphere_count != (int)spheres.size() || material_count != (int)materials.size())
        {
            AddDirtyFlag(DirtyFlagValue::SceneMemory);
            return 0;
        }
        const size_t bs = spheres.size() * sizeof(RtSphere), bm = materials.size() * sizeof(RtMaterial);
        if (bs) Check(cudaMemcpyAsync(d_spheres, spheres.data(), bs, cudaMemcpyHostToDevice, nullptr), "refresh spheres");
        if (bm) Check(cudaMemcpyAsync(d_materials, static_cast<const RtMaterial*>(materials.data()), bm, cudaMemcpyHostToDevice, nullptr), "refresh materials");
        return bs + bm;
    }
}
