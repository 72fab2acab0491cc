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
    int dims = 2;          // 1 / 2 / 3
    bool is_cube = true;
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

    // utils/CUDATexture.cpp:112-172 + :187-220 for every container / format the general parser accepts and CUDA has an array
    // format for: 1D/2D, cubemap, 3D; plain, normalised and block-compressed texels; full mip chains.
    // Differences from the reference, which never exercised these paths (its only texture is the RGBA32F cubemap):
    //  * block-compressed arrays take TEXEL extents in cudaMallocMipmappedArray and cudaMemcpy3D (measured on the B200 box,
    //    benchmarks/probe_bc.cu); the reference passes width/4 x height/4 and sizes the copy with 8+8+8+8 "bits per pixel", which
    //    its own size check rejects for every BC format;
    //  * a 2D texture is copied with depth 1 (the reference's TextureDesc::depth stays 0 for non-cubes, an empty copy);
    //  * R16G16B16A16_UINT / _SINT are 4 x 16-bit channels (the reference maps them to uint4 / short2).
    int CreateTexture(const uint8_t* bytes, const DdsImage& img, bool srgb, RtEnvMap** out)
    {
        const DxgiFormatInfo& f = img.format;
        if (f.cuda_kind < 0) return Fail(RT_E_INVALID, std::string("DDS: no CUDA array format for DXGI ") + f.name);
        const bool bc = f.block_bytes >= 8;
        if (bc && ((img.width & 3u) || (img.height & 3u))) return Fail(RT_E_INVALID, "DDS: block-compressed size must be a multiple of 4");  // CUDATexture.cpp:122-124
        if (img.is_cubemap ? (img.array_size != 6 || img.width != img.height) : (img.array_size != 1))
            return Fail(RT_E_INVALID, "DDS: texture arrays (other than one cubemap) are parsed but have no texture path");
        RtEnvMap* env = new RtEnvMap();
        env->dims = img.num_dimensions; env->is_cube = img.is_cubemap;
        const cudaChannelFormatDesc desc = cudaCreateChannelDesc(f.cx, f.cy, f.cz, f.cw, (cudaChannelFormatKind)f.cuda_kind);
        const cudaExtent extent = img.is_cubemap ? make_cudaExtent(img.width, img.height, 6)
                                : img.num_dimensions == 3 ? make_cudaExtent(img.width, img.height, img.depth)
                                : make_cudaExtent(img.width, img.num_dimensions == 1 ? 0 : img.height, 0);
        cudaError_t e = cudaMallocMipmappedArray(&env->array, &desc, extent, img.mip_count, img.is_cubemap ? cudaArrayCubemap : 0);
        std::vector<uint8_t> staging;
        for (uint32_t m = 0; e == cudaSuccess && m < img.mip_count; m++)
        {
            cudaArray_t level;
            e = cudaGetMipmappedArrayLevel(&level, env->array, m);
            if (e != cudaSuccess) break;
            const DdsSubresource& s0 = img.subresources[m];
            // mip-major re-pack: all array elements (cube faces) of this level back to back (CUDATexture.cpp:195-217)
            const size_t element_bytes = s0.slice_pitch * s0.depth;
            staging.resize(element_bytes * img.array_size);
            for (uint32_t a = 0; a < img.array_size; a++)
                memcpy(staging.data() + a * element_bytes, bytes + img.subresources[(size_t)a * img.mip_count + m].offset, element_bytes);
            cudaMemcpy3DParms p = {};
            const size_t rows = bc ? (s0.height + 3) / 4 : s0.height;  // rows of the source as laid out in memory
            p.srcPtr = make_cudaPitchedPtr(staging.data(), s0.row_pitch, s0.width, rows);
            p.dstArray = level;
            p.extent = make_cudaExtent(s0.width, img.num_dimensions == 1 ? 1 : s0.height, img.is_cubemap ? 6 : s0.depth);
            p.kind = cudaMemcpyHostToDevice;
            if (bc)
            {   // source rows are rows of 4x4 blocks: the pitched pointer describes texels, so its height is the texel height
                p.srcPtr = make_cudaPitchedPtr(staging.data(), s0.row_pitch, s0.width < 4 ? 4 : s0.width, s0.height < 4 ? 4 : s0.height);
                p.extent = make_cudaExtent(s0.width < 4 ? 4 : s0.width, s0.height < 4 ? 4 : s0.height, img.is_cubemap ? 6 : s0.depth);
            }
            e = cudaMemcpy3D(&p);
        }
        if (e == cudaSuccess)
        {
            cudaResourceDesc res; memset(&res, 0, sizeof(res));
            res.resType = cudaResourceTypeMipmappedArray;
            res.res.mipmap.mipmap = env->array;
            cudaTextureDesc td; memset(&td, 0, sizeof(td));
            td.normalizedCoords = 1;
            const bool integer_texels = !f.normalized && f.cuda_kind != 2 && !bc;  // plain UINT / SINT: no filtering possible
            td.filterMode = integer_texels ? cudaFilterModePoint : cudaFilterModeLinear;
            td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeWrap;
            td.readMode = f.normalized ? cudaReadModeNormalizedFloat : cudaReadModeElementType;
            td.sRGB = srgb ? 1 : 0;
            td.seamlessCubemap = img.is_cubemap ? 1 : 0;
            td.maxMipmapLevelClamp = (float)(img.mip_count - 1);
            e = cudaCreateTextureObject(&env->tex, &res, &td, nullptr);
        }
        if (e != cudaSuccess)
        {
            if (env->array) cudaFreeMipmappedArray(env->array);
            delete env;
            return CudaFail(e, "rt_texture_create");
        }
        *out = env;
        return RT_OK;
    }

    // one texel fetch per thread: coords = 3 floats (direction for cubemaps; u, v, w otherwise)
    __global__ void texture_fetch_kernel(cudaTextureObject_t tex, int dims, bool cube, int n, const float* __restrict__ coords, float lod, float4* __restrict__ out)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n) return;
        const float a = coords[3 * i], b = coords[3 * i + 1], c = coords[3 * i + 2];
        out[i] = cube ? texCubemapLod<float4>(tex, a, b, c, lod)
               : dims == 3 ? tex3DLod<float4>(tex, a, b, c, lod)
               : dims == 1 ? tex1DLod<float4>(tex, a, lod) : tex2DLod<float4>(tex, a, b, lod);
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

    // The re-layout for scene->gpu_bvh_nodes: the bound one if it still belongs to that allocation, else one built here from the
    // GPUScene's own device arrays (an unmodified host never calls rt_accel_bind; csrc/dropin.cpp).
    int ResolveAccel(const RtGPUScene* scene, const RtAccel** out)
    {
        *out = nullptr;
        const void* key = scene->gpu_bvh_nodes;
        if (!key) return Fail(RT_E_NOACCEL, "raytracing_process: scene->gpu_bvh_nodes is null");
        const uint64_t id = DeviceBufferId(key);
        std::vector<RtAccel*> dead;
        {
            std::lock_guard<std::mutex> lock(g_mutex);
            auto it = g_bound.find(key);
            if (it != g_bound.end())
            {
                if (it->second.buffer_id == 0 || id == 0 || it->second.buffer_id == id) { *out = it->second.accel; return RT_OK; }
                if (it->second.owned) dead.push_back(it->second.accel);  // same address, another allocation: stale
                g_bound.erase(it);
            }
            // library-built re-layouts whose node arrays are gone (freed, or the address re-used) go with them
            for (auto jt = g_bound.begin(); jt != g_bound.end();)
                if (jt->second.owned && jt->second.buffer_id && DeviceBufferId(jt->first) != jt->second.buffer_id) { dead.push_back(jt->second.accel); jt = g_bound.erase(jt); }
                else ++jt;
        }
        for (RtAccel* a : dead) rt_accel_destroy(a);
        RtAccel* built = nullptr;
        if (int r = BuildAccelFromDeviceScene(scene, &built)) return r;
        {
            std::lock_guard<std::mutex> lock(g_mutex);
            g_bound[key] = BoundAccel{ built, id, true };
        }
        *out = built;
        return RT_OK;
    }
}

extern "C" {

int rt_envmap_create_from_memory(const void* dds_bytes, size_t size, RtEnvMap** out)
try
{
    if (!dds_bytes || !out) return Fail(RT_E_INVALID, "rt_envmap_create_from_memory: null");
    DdsImage img; std::string err;
    if (int r = ParseDds(dds_bytes, size, img, err)) return Fail(r, err);
    if (!img.is_cubemap) return Fail(RT_E_IO, "DDS: not a six-face cubemap");
    if (img.format.dxgi == 2 && img.array_size == 6)
    {   // the asset's own layout (RGBA32F): the dedicated reader
        DdsCubemap dds;
        if (int r = ParseDdsCubemap(dds_bytes, size, dds, err)) return Fail(r, err);
        return CreateEnvMap(dds, out);
    }
    return CreateTexture(static_cast<const uint8_t*>(dds_bytes), img, true, out);  // any other cubemap format CUDA can sample as float4
}
catch (const std::exception& e) { return rt_internal::Fail(RT_E_INVALID, std::string("rt_envmap_create_from_memory: ") + e.what()); }
catch (...) { return rt_internal::Fail(RT_E_INVALID, "rt_envmap_create_from_memory: unknown C++ exception"); }

int rt_texture_create_from_memory(const void* dds_bytes, size_t size, int srgb, RtEnvMap** out)
try
{
    if (!dds_bytes || !out) return Fail(RT_E_INVALID, "rt_texture_create_from_memory: null");
    DdsImage img; std::string err;
    if (int r = ParseDds(dds_bytes, size, img, err)) return Fail(r, err);
    return CreateTexture(static_cast<const uint8_t*>(dds_bytes), img, srgb != 0, out);
}
catch (const std::exception& e) { return rt_internal::Fail(RT_E_INVALID, std::string("rt_texture_create_from_memory: ") + e.what()); }
catch (...) { return rt_internal::Fail(RT_E_INVALID, "rt_texture_create_from_memory: unknown C++ exception"); }

int rt_texture_create_from_file(const char* dds_path, int srgb, RtEnvMap** out)
try
{
    if (!dds_path || !out) return Fail(RT_E_INVALID, "rt_texture_create_from_file: null");
    std::vector<uint8_t> bytes;
    if (LoadFileBytes(dds_path, bytes)) return Fail(RT_E_IO, std::string("cannot read ") + dds_path);
    return rt_texture_create_from_memory(bytes.data(), bytes.size(), srgb, out);
}
catch (const std::exception& e) { return rt_internal::Fail(RT_E_INVALID, std::string("rt_texture_create_from_file: ") + e.what()); }
catch (...) { return rt_internal::Fail(RT_E_INVALID, "rt_texture_create_from_file: unknown C++ exception"); }

int rt_texture_fetch(const RtEnvMap* tex, int count, const float* d_coords, float lod, float* d_rgba, void* stream)
{
    if (!tex || !tex->tex || count < 0 || (count && (!d_coords || !d_rgba))) return Fail(RT_E_INVALID, "rt_texture_fetch: bad argument");
    if (!count) return RT_OK;
    texture_fetch_kernel<<<(count + 127) / 128, 128, 0, (cudaStream_t)stream>>>(tex->tex, tex->dims, tex->is_cube, count, d_coords, lod, (float4*)d_rgba);
    RT_CUDA(cudaGetLastError());
    return RT_OK;
}

int rt_dds_info(const void* dds_bytes, size_t size, uint64_t* info, int capacity)
try
{
    if (!dds_bytes || !info || capacity < 12) return Fail(RT_E_INVALID, "rt_dds_info: bad argument");
    DdsImage img; std::string err;
    if (int r = ParseDds(dds_bytes, size, img, err)) return Fail(r, err);
    info[0] = img.width; info[1] = img.height; info[2] = img.depth; info[3] = img.mip_count; info[4] = img.array_size;
    info[5] = img.is_cubemap; info[6] = (uint64_t)img.num_dimensions; info[7] = img.format.dxgi; info[8] = img.format.bits_per_pixel;
    info[9] = img.format.block_bytes; info[10] = img.format.cuda_kind >= 0; info[11] = img.subresources.size();
    const int n = (int)img.subresources.size();
    for (int i = 0; i < n && 12 + 3 * i + 2 < capacity; i++)
    {
        info[12 + 3 * i] = img.subresources[i].offset; info[13 + 3 * i] = img.subresources[i].row_pitch; info[14 + 3 * i] = img.subresources[i].slice_pitch;
    }
    return n;
}
catch (const std::exception& e) { return rt_internal::Fail(RT_E_INVALID, std::string("rt_dds_info: ") + e.what()); }
catch (...) { return rt_internal::Fail(RT_E_INVALID, "rt_dds_info: unknown C++ exception"); }

int rt_envmap_create_from_file(const char* dds_path, RtEnvMap** out)
try
{
    if (!dds_path || !out) return Fail(RT_E_INVALID, "rt_envmap_create_from_file: null");
    std::vector<uint8_t> bytes;
    if (LoadFileBytes(dds_path, bytes)) return Fail(RT_E_IO, std::string("cannot read ") + dds_path);
    return rt_envmap_create_from_memory(bytes.data(), bytes.size(), out);
}
catch (const std::exception& e) { return rt_internal::Fail(RT_E_INVALID, std::string("rt_envmap_create_from_file: ") + e.what()); }
catch (...) { return rt_internal::Fail(RT_E_INVALID, "rt_envmap_create_from_file: unknown C++ exception"); }

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
try
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
catch (const std::exception& e) { return rt_internal::Fail(RT_E_INVALID, std::string("rt_dds_read_level0: ") + e.what()); }
catch (...) { return rt_internal::Fail(RT_E_INVALID, "rt_dds_read_level0: unknown C++ exception"); }

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
    if (variant < 0 || variant > 1) return Fail(RT_E_INVALID, "rt_trace_rays: unknown variant");
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
    if (width <= 0 || height <= 0) return 0;
    if (variant == 10) return PathTraceWavefrontWorkspaceBytes(width, height);
    return variant == 0 ? PathTraceWorkspaceBytes(width, height) : 0;
}

int rt_process_ex(const RtProcessParams* p)
{
    if (!p || !p->surface || !p->scene || !p->accel || p->width <= 0 || p->height <= 0) return Fail(RT_E_INVALID, "rt_process_ex: bad argument");
    if (p->blend_mode == 0 && !p->surface_last_frame) return Fail(RT_E_INVALID, "rt_process_ex: running-mean blend needs surface_last_frame");
    if (p->samples_per_launch <= 0 || p->max_bounces <= 0) return Fail(RT_E_INVALID, "rt_process_ex: spp and bounces must be positive");
    if (p->pitch < (uint64_t)p->width * 16) return Fail(RT_E_INVALID, "rt_process_ex: pitch smaller than a row");
    if (!p->scene->rng_state) return Fail(RT_E_INVALID, "rt_process_ex: scene has no rng_state");
    if (p->variant < 0 || p->variant > 10 || (p->variant >= 4 && p->variant <= 9)) return Fail(RT_E_INVALID, "rt_process_ex: unknown variant");
    if (p->variant == 10)
    {
        if (!p->workspace || p->workspace_bytes < PathTraceWavefrontWorkspaceBytes(p->width, p->height) || (int64_t)p->samples_per_launch * p->max_bounces > 4096)
            return Fail(RT_E_INVALID, "rt_process_ex: the wavefront variant needs a workspace of rt_process_workspace_bytes(width, height, 10) and spp x bounces <= 4096");
        RT_CUDA(LaunchPathTraceWavefront(p->accel->view, MakeSceneView(p->scene), *p, (cudaStream_t)p->stream));
        return RT_OK;
    }
    if (p->write_last_frame && !p->surface_last_frame) return Fail(RT_E_INVALID, "rt_process_ex: write_last_frame needs surface_last_frame");
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
    int r = scene ? ResolveAccel(scene, &p.accel) : Fail(RT_E_INVALID, "raytracing_process: null scene");
    if (r == RT_OK) r = rt_process_ex(&p);
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

int rt_resolve_sum(const void* sum_surface, void* out_surface, int width, int height, uint64_t pitch_in, uint64_t pitch_out, void* stream)
{
    if (!sum_surface || !out_surface || width <= 0 || height <= 0 || pitch_in < (uint64_t)width * 16 || pitch_out < (uint64_t)width * 16)
        return Fail(RT_E_INVALID, "rt_resolve_sum: bad argument");
    RT_CUDA(LaunchResolveSum(sum_surface, out_surface, width, height, pitch_in, pitch_out, (cudaStream_t)stream));
    return RT_OK;
}

}  // extern "C"
