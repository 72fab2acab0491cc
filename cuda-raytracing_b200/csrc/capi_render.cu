// C ABI, part 2: environment map, closest-hit queries, path tracing, display transform.
#include <cstdio>
#include <cstring>
#include <vector>
#include "capi_internal.h"
#include "dds.h"

using namespace rt_internal;

struct RtEnvMap
{
    cudaMipmappedArray_t array = nullptr;
    cudaTextureObject_t tex = 0;
};

namespace
{
    // utils/CUDATexture.cpp:112-172 for the cubemap / RGBA32F case
    int CreateEnvMap(const DdsCubemap& dds, RtEnvMap** out)
    {
        RtEnvMap* env = new RtEnvMap();
        const cudaChannelFormatDesc desc = cudaCreateChannelDesc<float4>();
        const cudaExtent extent = make_cudaExtent(dds.size, dds.size, 6);
        cudaError_t e = cudaMallocMipmappedArray(&env->array, &desc, extent, dds.mip_count, cudaArrayCubemap);
        for (uint32_t m = 0; e == cudaSuccess && m < dds.mip_count; m++)
        {
            cudaArray_t level;
            e = cudaGetMipmappedArrayLevel(&level, env->array, m);
            if (e != cudaSuccess) break;
            const uint32_t w = dds.size >> m ? dds.size >> m : 1;
            cudaMemcpy3DParms p = {};
            p.srcPtr = make_cudaPitchedPtr((void*)dds.levels[m].data(), (size_t)w * 16, w, w);
            p.dstArray = level;
            p.extent = make_cudaExtent(w, w, 6);
            p.kind = cudaMemcpyHostToDevice;
            e = cudaMemcpy3D(&p);
        }
        if (e == cudaSuccess)
        {
            cudaResourceDesc res; memset(&res, 0, sizeof(res));
            res.resType = cudaResourceTypeMipmappedArray;
            res.res.mipmap.mipmap = env->array;
            cudaTextureDesc td; memset(&td, 0, sizeof(td));
            td.normalizedCoords = 1;
            td.filterMode = cudaFilterModeLinear;
            td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeWrap;
            td.readMode = cudaReadModeElementType;
            td.sRGB = 1;  // as the reference sets it; no effect on float texels
            td.seamlessCubemap = 1;
            td.maxMipmapLevelClamp = (float)(dds.mip_count - 1);
            e = cudaCreateTextureObject(&env->tex, &res, &td, nullptr);
        }
        if (e != cudaSuccess)
        {
            if (env->array) cudaFreeMipmappedArray(env->array);
            delete env;
            return CudaFail(e, "rt_envmap_create");
        }
        *out = env;
        return RT_OK;
    }

    // Work counters of the persistent kernels: a ring of device words per device, one slot per launch (launches in flight on
    // different streams never share a slot unless more than kCounterSlots are outstanding).
    constexpr int kCounterSlots = 256;
    std::map<int, unsigned int*> g_counters;
    unsigned g_counter_next = 0;
    int NextCounter(unsigned int** out)
    {
        int dev = 0;
        RT_CUDA(cudaGetDevice(&dev));
        std::lock_guard<std::mutex> lock(g_mutex);
        auto it = g_counters.find(dev);
        if (it == g_counters.end())
        {
            unsigned int* d = nullptr;
            RT_CUDA(cudaMalloc(&d, kCounterSlots * sizeof(unsigned int)));
            it = g_counters.emplace(dev, d).first;
        }
        *out = it->second + (g_counter_next++ % kCounterSlots);
        return RT_OK;
    }

    const RtAccel* LookupAccel(const RtGPUScene* scene)
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        auto it = g_bound.find(scene->gpu_bvh_nodes);
        return it == g_bound.end() ? nullptr : it->second;
    }
}

extern "C" {

int rt_envmap_create_from_memory(const void* dds_bytes, size_t size, RtEnvMap** out)
{
    if (!dds_bytes || !out) return Fail(RT_E_INVALID, "rt_envmap_create_from_memory: null");
    DdsCubemap dds; std::string err;
    if (int r = ParseDdsCubemap(dds_bytes, size, dds, err)) return Fail(r, err);
    return CreateEnvMap(dds, out);
}

int rt_envmap_create_from_file(const char* dds_path, RtEnvMap** out)
{
    if (!dds_path || !out) return Fail(RT_E_INVALID, "rt_envmap_create_from_file: null");
    std::vector<uint8_t> bytes;
    if (LoadFileBytes(dds_path, bytes)) return Fail(RT_E_IO, std::string("cannot read ") + dds_path);
    return rt_envmap_create_from_memory(bytes.data(), bytes.size(), out);
}

uint64_t rt_envmap_texture(const RtEnvMap* env) { return env ? (uint64_t)env->tex : 0; }

int rt_envmap_destroy(RtEnvMap* env)
{
    if (!env) return RT_OK;
    if (env->tex) cudaDestroyTextureObject(env->tex);
    if (env->array) cudaFreeMipmappedArray(env->array);
    delete env;
    return RT_OK;
}

int rt_dds_read_level0(const char* dds_path, float* out_texels, size_t capacity_floats)
{
    if (!dds_path) return Fail(RT_E_INVALID, "rt_dds_read_level0: null path");
    std::vector<uint8_t> bytes;
    if (LoadFileBytes(dds_path, bytes)) return Fail(RT_E_IO, std::string("cannot read ") + dds_path);
    DdsCubemap dds; std::string err;
    if (int r = ParseDdsCubemap(bytes.data(), bytes.size(), dds, err)) return Fail(r, err);
    if (out_texels)
    {
        if (capacity_floats < dds.levels[0].size()) return Fail(RT_E_INVALID, "rt_dds_read_level0: buffer too small");
        memcpy(out_texels, dds.levels[0].data(), dds.levels[0].size() * sizeof(float));
    }
    return (int)dds.size;
}

// ---- closest-hit queries --------------------------------------------------------------------
int rt_trace_primary(const RtAccel* accel, const RtGPUScene* scene, int width, int height, int with_spheres,
                     int variant, uint32_t* d_id, float* d_t, unsigned long long* d_ray_counter, void* stream)
{
    if (!accel || !scene || !d_id || !d_t || width <= 0 || height <= 0) return Fail(RT_E_INVALID, "rt_trace_primary: bad argument");
    if (variant < 0 || variant > 1) return Fail(RT_E_INVALID, "rt_trace_primary: unknown variant");
    RT_CUDA(LaunchTracePrimary(accel->view, MakeSceneView(scene), width, height, with_spheres != 0, variant, d_id, d_t, d_ray_counter, (cudaStream_t)stream));
    return RT_OK;
}

int rt_trace_rays(const RtAccel* accel, const RtGPUScene* scene, const float* d_rays, uint64_t ray_count,
                  int with_spheres, int variant, uint32_t* d_id, float* d_t, unsigned long long* d_ray_counter, void* stream)
{
    if (!accel || !scene || (ray_count && (!d_rays || !d_id || !d_t))) return Fail(RT_E_INVALID, "rt_trace_rays: bad argument");
    if (variant < 0 || variant > 2) return Fail(RT_E_INVALID, "rt_trace_rays: unknown variant");
    if (variant == 2)
    {
        unsigned int* counter = nullptr;
        if (int r = NextCounter(&counter)) return r;
        RT_CUDA(LaunchExtendRays(accel->view, MakeSceneView(scene), d_rays, ray_count, with_spheres != 0, d_id, d_t, d_ray_counter, counter, (cudaStream_t)stream));
        return RT_OK;
    }
    RT_CUDA(LaunchTraceRays(accel->view, MakeSceneView(scene), d_rays, ray_count, with_spheres != 0, variant, d_id, d_t, d_ray_counter, (cudaStream_t)stream));
    return RT_OK;
}

int rt_count_reference_work(const RtAccel* accel, const RtGPUScene* scene, int width, int height, int with_spheres,
                            uint32_t* d_counts, void* stream)
{
    if (!accel || !scene || !d_counts || width <= 0 || height <= 0) return Fail(RT_E_INVALID, "rt_count_reference_work: bad argument");
    RT_CUDA(LaunchCountWork(accel->view, MakeSceneView(scene), width, height, with_spheres != 0, d_counts, (cudaStream_t)stream));
    return RT_OK;
}

// ---- path tracing ---------------------------------------------------------------------------
uint64_t rt_process_workspace_bytes(int width, int height, int variant)
{
    return variant == 0 && width > 0 && height > 0 ? PathTraceWorkspaceBytes(width, height) : 0;
}

int rt_process_ex(const RtProcessParams* p)
{
    if (!p || !p->surface || !p->scene || !p->accel || p->width <= 0 || p->height <= 0) return Fail(RT_E_INVALID, "rt_process_ex: bad argument");
    if (p->blend_mode == 0 && !p->surface_last_frame) return Fail(RT_E_INVALID, "rt_process_ex: running-mean blend needs surface_last_frame");
    if (p->samples_per_launch <= 0 || p->max_bounces <= 0) return Fail(RT_E_INVALID, "rt_process_ex: spp and bounces must be positive");
    if (p->pitch < (uint64_t)p->width * 16) return Fail(RT_E_INVALID, "rt_process_ex: pitch smaller than a row");
    if (!p->scene->rng_state) return Fail(RT_E_INVALID, "rt_process_ex: scene has no rng_state");
    if (p->variant < 0 || p->variant > 2) return Fail(RT_E_INVALID, "rt_process_ex: unknown variant");
    if (p->variant == 1)
    {
        unsigned int* counter = nullptr;
        if (int r = NextCounter(&counter)) return r;
        RT_CUDA(LaunchPathTraceRegroup(p->accel->view, MakeSceneView(p->scene), *p, counter, (cudaStream_t)p->stream));
        return RT_OK;
    }
    RT_CUDA(LaunchPathTrace(p->accel->view, MakeSceneView(p->scene), *p, (cudaStream_t)p->stream));
    return RT_OK;
}

void raytracing_process(void* surface, void* surface_last_frame, int width, int height, size_t pitch, int frame_index, RtGPUScene* scene)
{
    RtProcessParams p; memset(&p, 0, sizeof(p));
    p.surface = surface; p.surface_last_frame = surface_last_frame; p.width = width; p.height = height; p.pitch = pitch;
    p.frame_index = frame_index; p.samples_per_launch = 5; p.max_bounces = 6; p.blend_mode = 0; p.variant = 0;
    p.scene = scene;
    p.accel = scene ? LookupAccel(scene) : nullptr;
    int r = p.accel ? rt_process_ex(&p) : Fail(RT_E_NOACCEL, "raytracing_process: no RtAccel bound for scene->gpu_bvh_nodes (rt_accel_bind)");
    if (r != RT_OK) printf("(raytracing_kernel_main) failed to launch error = %d (%s)\n", r, rt_last_error());
}

int rt_ambient_occlusion(const RtAccel* accel, const RtGPUScene* scene, int width, int height, int ao_rays,
                         float* d_ao, unsigned long long* d_ray_counter, void* stream)
{
    if (!accel || !scene || !d_ao || width <= 0 || height <= 0 || ao_rays < 0 || !scene->rng_state) return Fail(RT_E_INVALID, "rt_ambient_occlusion: bad argument");
    RT_CUDA(LaunchAmbientOcclusion(accel->view, MakeSceneView(scene), width, height, ao_rays, d_ao, d_ray_counter, (cudaStream_t)stream));
    return RT_OK;
}

int rt_tonemap_rgba8(const void* surface, int width, int height, uint64_t pitch, float divide_by, uint8_t* d_rgba8, void* stream)
{
    if (!surface || !d_rgba8 || width <= 0 || height <= 0 || !(divide_by > 0)) return Fail(RT_E_INVALID, "rt_tonemap_rgba8: bad argument");
    RT_CUDA(LaunchTonemap(surface, width, height, pitch, divide_by, d_rgba8, (cudaStream_t)stream));
    return RT_OK;
}

}  // extern "C"
