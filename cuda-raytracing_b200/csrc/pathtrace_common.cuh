// Shared by the path-trace kernels (pathtrace.cu: megakernel + batched kernel; pathtrace_wavefront.cu: queue-based wavefront):
// frame parameters, the camera ray (GPUCamera::GetRay, RayTracing/GPUScene.h:13) and the final pixel write
// (RayTracing/main_raytracing.cu:195-199). Compiled with -fmad=false.
#pragma once
#include <cmath>
#include "shade.cuh"

#ifndef RT_PT_BLOCK
#define RT_PT_BLOCK 128  // threads per CTA of the megakernel: 256 / 128 / 64 = 57.0 / 56.0 / 56.8 ms (finer-grained tail of the last wave)
#endif
#define RT_PT_WARP_W 8     // warp footprint 8 x 4 pixels (4 x 8 / 8 x 4 / 16 x 2 = 52.8 / 52.9 / 53.3 ms)

__device__ __forceinline__ void camera_ray_pt(const RtCamera& cam, float u, float v, Vec3& o, Vec3& d)
{
    o = mk3(cam.origin.x, cam.origin.y, cam.origin.z);
    const Vec3 h = mk3(cam.horizontal.x, cam.horizontal.y, cam.horizontal.z), w = mk3(cam.vertical.x, cam.vertical.y, cam.vertical.z);
    const Vec3 llc = mk3(cam.lower_left_corner.x, cam.lower_left_corner.y, cam.lower_left_corner.z);
    d = ((llc + u * h) + v * w) - o;
}

struct FrameParams
{
    unsigned char* surface;
    unsigned char* surface_last;
    int width, height;
    uint64_t pitch;
    int frame_index, spp, max_bounces, blend_mode, write_last;
    const uint32_t* tile_order;  // cost-ordered scheduling (workspace): tile processed by the k-th CTA, or nullptr = grid order
    uint32_t* tile_cost;         // per tile: SM cycles / 16 its slowest warp took in this launch (input of the next launch's order)
    unsigned long long* ray_counter;
    unsigned long long* work_counters;  // variants 2 / 3: [0] += nodes popped / wide-node steps, [1] += triangle tests
    PathConsts consts;
};

// Final pixel write: main_raytracing.cu:195-199 (blend_mode 0) or sum accumulation (blend_mode 1).
__device__ __forceinline__ void write_pixel(const FrameParams& P, int x, int y, Vec3 sum)
{
    float4* px = reinterpret_cast<float4*>(P.surface + (size_t)y * P.pitch) + x;
    float4 out;
    if (P.blend_mode == 0)
    {
        const float n = (float)P.spp;
        const Vec3 mean = mk3(sum.x / n, sum.y / n, sum.z / n);
        const float a = P.frame_index > 0 ? 1.0f / (float)(P.frame_index + 1) : 1.0f;
        const float4 last = *(reinterpret_cast<const float4*>(P.surface_last + (size_t)y * P.pitch) + x);
        out = make_float4(mix1(last.x, mean.x, a), mix1(last.y, mean.y, a), mix1(last.z, mean.z, a), 1.0f);
    }
    else
    {
        const float4 cur = *px;
        out = make_float4(cur.x + sum.x, cur.y + sum.y, cur.z + sum.z, cur.w + (float)P.spp);
    }
    *px = out;
    if (P.write_last) *(reinterpret_cast<float4*>(P.surface_last + (size_t)y * P.pitch) + x) = out;
}


// quat(vec3(0, PI, 0)) of the miss shader (main_raytracing.cu:151; glm type_quat.inl:204-213), in host fp32 like the oracle
inline PathConsts MakePathConsts()
{
    PathConsts c;
    const float PI = 3.1415926536f;
    const float cx = std::cos(0.0f * 0.5f), cy = std::cos(PI * 0.5f), cz = std::cos(0.0f * 0.5f);
    const float sx = std::sin(0.0f * 0.5f), sy = std::sin(PI * 0.5f), sz = std::sin(0.0f * 0.5f);
    c.qw = cx * cy * cz + sx * sy * sz;
    c.qx = sx * cy * cz - cx * sy * sz;
    c.qy = cx * sy * cz + sx * cy * sz;
    c.qz = cx * cy * sz - sx * sy * cz;
    return c;
}
