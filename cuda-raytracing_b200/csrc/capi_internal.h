// Shared between the C-ABI translation units (capi.cu, capi_render.cu).
#pragma once
#include <cuda_runtime.h>
#include <map>
#include <mutex>
#include <string>
#include "accel.h"
#include "b200rt.h"
#include "kernels.h"

namespace rt_internal
{
    inline thread_local std::string g_error;
    inline std::mutex g_mutex;
    // gpu_bvh_nodes -> re-layout (raytracing_process lookup). buffer_id = CUDA buffer id of the node allocation at bind time (0 = unknown):
    // a host that freed its arrays and got the same address back has another id, and the binding is dropped instead of used (dropin.cpp).
    // owned = built by the library itself for an unmodified host (destroyed when replaced, unbound or found stale).
    struct BoundAccel { RtAccel* accel; uint64_t buffer_id; bool owned; };
    inline std::map<const void*, BoundAccel> g_bound;
    uint64_t DeviceBufferId(const void* device_pointer);                       // dropin.cpp
    int BuildAccelFromDeviceScene(const RtGPUScene* scene, RtAccel** out);     // dropin.cpp

    inline int Fail(int code, const std::string& msg) { g_error = msg; return code; }
    inline int CudaFail(cudaError_t e, const char* what) { return Fail(RT_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e)); }

    inline SceneView MakeSceneView(const RtGPUScene* s)
    {
        SceneView v{};
        if (!s) return v;
        v.spheres = s->gpu_spheres; v.materials = s->gpu_materials;
        v.sphere_count = s->gpu_spheres ? s->sphere_count : 0; v.material_count = s->material_count;
        v.rng_state = s->rng_state; v.env_tex = (cudaTextureObject_t)s->environment_cubemap_tex;
        v.camera = s->camera;
        return v;
    }
}
#define RT_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return rt_internal::CudaFail(e_, #call); } while (0)
