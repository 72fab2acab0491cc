// True drop-in support for raytracing_process (RayTracing/main_raytracing.cu:202): an UNMODIFIED host only ever hands over a
// GPUScene whose six pointers are device arrays it uploaded itself (RayTracing/Scene.cpp:189-231). When no re-layout (RtAccel) is
// bound for scene->gpu_bvh_nodes, one is built here from those device arrays:
//   extents: the allocation each pointer lies in (cuMemGetAddressRange) bounds what may be read; the tree itself says how much is
//   meant - node count = highest node reachable from node 0 + 1, face count = sum of prim_count over the leaves (every face sits in
//   exactly one leaf, BVH.cpp:34-35), vertex count = highest vertex index + 1;
//   then D2H copies, the same BuildAccelHost + upload rt_accel_create does, and a binding owned by the library.
// Staleness: a binding remembers the CUDA buffer id (CU_POINTER_ATTRIBUTE_BUFFER_ID: unique per allocation for the life of the
// process) of the node array it was made for. A host that frees its arrays and gets the same address back from cudaMalloc has a
// NEW buffer id: the old binding is dropped and rebuilt instead of rendering a stale tree. (Overwriting the same allocation in
// place is not detectable this way: such a host calls rt_accel_unbind.)
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include "capi_internal.h"

namespace rt_internal
{
    namespace
    {
        typedef CUresult (*PointerGetAttributeFn)(void*, CUpointer_attribute, CUdeviceptr);
        typedef CUresult (*MemGetAddressRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);

        void* DriverEntry(const char* name)
        {
            void* fn = nullptr;
            cudaDriverEntryPointQueryResult status;
            if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &status) != cudaSuccess || status != cudaDriverEntryPointSuccess) return nullptr;
            return fn;
        }

        // bytes readable from `p` to the end of the allocation it lies in (0 = not a device allocation this process can query)
        size_t BytesToAllocationEnd(const void* p)
        {
            static MemGetAddressRangeFn fn = reinterpret_cast<MemGetAddressRangeFn>(DriverEntry("cuMemGetAddressRange"));
            if (!fn || !p) return 0;
            CUdeviceptr base = 0; size_t size = 0;
            if (fn(&base, &size, (CUdeviceptr)(uintptr_t)p) != CUDA_SUCCESS) return 0;
            const size_t offset = (size_t)((uintptr_t)p - (uintptr_t)base);
            return offset < size ? size - offset : 0;
        }
    }

    uint64_t DeviceBufferId(const void* p)
    {
        static PointerGetAttributeFn fn = reinterpret_cast<PointerGetAttributeFn>(DriverEntry("cuPointerGetAttribute"));
        if (!fn || !p) return 0;
        unsigned long long id = 0;
        if (fn(&id, CU_POINTER_ATTRIBUTE_BUFFER_ID, (CUdeviceptr)(uintptr_t)p) != CUDA_SUCCESS) return 0;
        return id;
    }

    int BuildAccelFromDeviceScene(const RtGPUScene* scene, RtAccel** out)
    {
        *out = nullptr;
        if (!scene->gpu_bvh_nodes || !scene->gpu_bvh_face_indices || !scene->gpu_faces || !scene->gpu_vertices)
            return Fail(RT_E_NOACCEL, "raytracing_process: the GPUScene has no BVH / face / vertex arrays");
        const size_t node_cap = BytesToAllocationEnd(scene->gpu_bvh_nodes) / sizeof(RtBVHNode);
        const size_t index_cap = BytesToAllocationEnd(scene->gpu_bvh_face_indices) / sizeof(uint32_t);
        const size_t face_cap = BytesToAllocationEnd(scene->gpu_faces) / sizeof(RtFace);
        const size_t vertex_cap = BytesToAllocationEnd(scene->gpu_vertices) / sizeof(RtVertex);
        if (!node_cap || !index_cap || !face_cap || !vertex_cap)
            return Fail(RT_E_NOACCEL, "raytracing_process: cannot size the GPUScene's device arrays (not cudaMalloc allocations of this process); bind a re-layout with rt_accel_bind");
        std::vector<RtBVHNode> nodes(node_cap);
        RT_CUDA(cudaMemcpy(nodes.data(), scene->gpu_bvh_nodes, node_cap * sizeof(RtBVHNode), cudaMemcpyDeviceToHost));
        // walk from the root (main_raytracing.cu:37-78 visits exactly these): node count, face count
        size_t node_count = 0, face_count = 0, visited = 0;
        {
            std::vector<uint32_t> stack; stack.push_back(0);
            while (!stack.empty())
            {
                const uint32_t n = stack.back(); stack.pop_back();
                if (n >= node_cap || ++visited > node_cap) return Fail(RT_E_INVALID, "raytracing_process: the GPUScene's BVH leaves its node allocation (or is not a tree)");
                node_count = std::max<size_t>(node_count, (size_t)n + 1);
                const RtBVHNode& node = nodes[n];
                if (node.prim_count > 0) face_count = std::max<size_t>(face_count, (size_t)node.first_index + node.prim_count);
                else { stack.push_back(node.first_index); stack.push_back(node.first_index + 1); }
            }
        }
        if (!face_count || face_count > index_cap || face_count > face_cap) return Fail(RT_E_INVALID, "raytracing_process: the GPUScene's BVH names more faces than its arrays hold");
        nodes.resize(node_count);
        std::vector<uint32_t> face_indices(face_count);
        std::vector<RtFace> faces(face_count);
        RT_CUDA(cudaMemcpy(face_indices.data(), scene->gpu_bvh_face_indices, face_count * sizeof(uint32_t), cudaMemcpyDeviceToHost));
        RT_CUDA(cudaMemcpy(faces.data(), scene->gpu_faces, face_count * sizeof(RtFace), cudaMemcpyDeviceToHost));
        size_t vertex_count = 0;
        for (const RtFace& f : faces) vertex_count = std::max<size_t>(vertex_count, (size_t)std::max(f.v0, std::max(f.v1, f.v2)) + 1);
        if (vertex_count > vertex_cap) return Fail(RT_E_INVALID, "raytracing_process: the GPUScene's faces name vertices beyond the vertex allocation");
        std::vector<RtVertex> vertices(vertex_count);
        RT_CUDA(cudaMemcpy(vertices.data(), scene->gpu_vertices, vertex_count * sizeof(RtVertex), cudaMemcpyDeviceToHost));
        return rt_accel_create(nodes.data(), (uint32_t)node_count, face_indices.data(), faces.data(), (uint32_t)face_count,
                               vertices.data(), (uint32_t)vertex_count, 0, out);
    }
}
