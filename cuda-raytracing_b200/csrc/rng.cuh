// cuRAND-XORWOW-compatible per-pixel generator held in registers (north-star item 4).
// Restates curand_kernel.h:863-874 (curand), curand_uniform.h:69-72 (_curand_uniform) and
// RayTracing/Random.h:19-31 (GetFloat01 / GetRandomPointOnSphere). The reference keeps the state in
// global memory and does a 24-B read-modify-write per draw; here the six live words are loaded once
// per launch and stored once (the Box-Muller fields of curandState are never touched by this path).
#pragma once
#include <cstdint>
#include "rt_types.h"
#include "traverse.cuh"

struct Xorwow
{
    uint32_t d, v0, v1, v2, v3, v4;

    __device__ __forceinline__ void load(const RtRngState* s)
    {
        const uint4 a = *reinterpret_cast<const uint4*>(s);            // d v0 v1 v2
        const uint2 b = *reinterpret_cast<const uint2*>(&s->v[3]);     // v3 v4
        d = a.x; v0 = a.y; v1 = a.z; v2 = a.w; v3 = b.x; v4 = b.y;
    }
    __device__ __forceinline__ void store(RtRngState* s) const
    {
        *reinterpret_cast<uint4*>(s) = make_uint4(d, v0, v1, v2);
        *reinterpret_cast<uint2*>(&s->v[3]) = make_uint2(v3, v4);
    }
    __device__ __forceinline__ uint32_t next()
    {
        const uint32_t t = v0 ^ (v0 >> 2);
        v0 = v1; v1 = v2; v2 = v3; v3 = v4;
        v4 = (v4 ^ (v4 << 4)) ^ (t ^ (t << 1));
        d += 362437u;
        return v4 + d;
    }
    // x * 2^-32 + 2^-33 : the product is exact, so fused or not gives the same float
    __device__ __forceinline__ float uniform() { return (float)next() * 2.3283064e-10f + (2.3283064e-10f / 2.0f); }

    __device__ __forceinline__ Vec3 point_on_sphere()  // Random.h:23-31
    {
        const float z = uniform() * 2.0f - 1.0f;
        const float a = uniform() * 3.14159265358979323846f * 2.0f;
        const float r = sqrtf(1.0f - z * z);
        const float x = r * cosf(a);
        const float y = r * sinf(a);
        return mk3(x, y, z);
    }
};
