// Shading step of the path tracer: restates ray_color's hit / miss branches
// (RayTracing/main_raytracing.cu:111-160) and the hit attributes BVHRayHit / GetRayHit produce
// (:65-69, :98-100). Compiled with -fmad=false; GLM expression order kept (mix = x*(1-a) + y*a,
// reflect = I - N*dot(N,I)*2, quat*vec = v + ((uv*w) + uuv)*2).
#pragma once
#include "kernels.h"
#include "rng.cuh"
#include "traverse.cuh"

struct PathConsts
{
    // quat(vec3(0, PI, 0)) with PI = 3.1415926536f (main_raytracing.cu:7,151), evaluated on the host
    float qx, qy, qz, qw;
};

__device__ __forceinline__ float mix1(float x, float y, float a) { return x * (1.0f - a) + y * a; }
__device__ __forceinline__ Vec3 mix3(Vec3 x, Vec3 y, float a) { return x * (1.0f - a) + y * a; }  // per component mix1: x*(1-a) + y*a (packed multiplies)
__device__ __forceinline__ float glm_max(float a, float b) { return (a < b) ? b : a; }
__device__ __forceinline__ float glm_min(float a, float b) { return (b < a) ? b : a; }

struct PathState
{
    Vec3 o, d;        // current ray (d unnormalised on bounce 0, as the reference passes it on)
    Vec3 throughput;
    Vec3 radiance;
};

// Returns true if the path continues with (ps.o, ps.d) updated.
__device__ __forceinline__ bool shade_hit(const AccelView& A, const SceneView& S, const HitRecord& h, PathState& ps, Xorwow& rng)
{
    const Vec3 nd = normalize3(ps.d);
    const Vec3 position = ps.o + nd * h.t;
    Vec3 normal; uint32_t material_index;
    if (h.id & RT_SPHERE_ID_BIT)
    {   // main_raytracing.cu:98-100 - no back-face flip for spheres
        const float4 s = ldg4(&S.spheres[h.slot]);
        normal = (position - mk3(s.x, s.y, s.z)) / s.w;
        material_index = (uint32_t)S.spheres[h.slot].material;
    }
    else
    {   // main_raytracing.cu:63-69 - weights (bary.x -> v0, bary.y -> v1, 1-x-y -> v2) as the reference assigns them
        const ShadeRecord* sr = A.shade + h.slot;
        const float4 n0 = ldg4(sr), n1 = ldg4(reinterpret_cast<const char*>(sr) + 16), n2 = ldg4(reinterpret_cast<const char*>(sr) + 32);
        const float bz = 1.0f - h.bu - h.bv;
        normal = normalize3((mk3(n0.x, n0.y, n0.z) * h.bu + mk3(n1.x, n1.y, n1.z) * h.bv) + mk3(n2.x, n2.y, n2.z) * bz);
        material_index = __float_as_uint(n0.w);
        if (dot3(nd, normal) >= 0.0f) normal = -normal;
    }
    const RtMaterial* m = S.materials + material_index;
    const float4 albedo = ldg4(m), emissive = ldg4(reinterpret_cast<const char*>(m) + 16), specular = ldg4(reinterpret_cast<const char*>(m) + 32);
    const float4 params = ldg4(reinterpret_cast<const char*>(m) + 48);  // roughness, specular_percent, IOR

    const float do_specular = (rng.uniform() < params.y) ? 1.0f : 0.0f;                                   // :124
    ps.radiance = ps.radiance + ps.throughput * mk3(emissive.x, emissive.y, emissive.z);                   // :126
    ps.throughput = ps.throughput * mix3(mk3(albedo.x, albedo.y, albedo.z), mk3(specular.x, specular.y, specular.z), do_specular);  // :127

    const Vec3 diffuse_dir = normalize3(normal + rng.point_on_sphere());                                  // :129
    const Vec3 refl = ps.d - (normal * dot3(normal, ps.d)) * 2.0f;                                        // glm::reflect, unnormalised d on bounce 0
    Vec3 specular_dir = normalize3(refl);                                                                 // :130
    specular_dir = normalize3(mix3(specular_dir, diffuse_dir, params.x * params.x));                      // :132
    const Vec3 new_dir = normalize3(mix3(diffuse_dir, specular_dir, do_specular));                        // :133

    ps.o = position + normal * 0.01f;                                                                     // :135
    ps.d = new_dir;

    const float p = glm_max(ps.throughput.x, glm_max(ps.throughput.y, ps.throughput.z));                  // :141
    if (rng.uniform() > p) return false;                                                                  // :142
    ps.throughput = ps.throughput * (1.0f / p);                                                           // :146
    return true;
}

__device__ __forceinline__ void shade_miss(const SceneView& S, const PathConsts& C, PathState& ps)
{   // main_raytracing.cu:151-153
    const Vec3 qv = mk3(C.qx, C.qy, C.qz);
    const Vec3 uv = cross3(qv, ps.d);
    const Vec3 uuv = cross3(qv, uv);
    const Vec3 r = ps.d + ((uv * C.qw) + uuv) * 2.0f;
    const float4 c = texCubemapLod<float4>(S.env_tex, r.x, r.y, r.z, 0.0f);
    const Vec3 e = mk3(glm_min(glm_max(c.x, 0.0f), 50.0f), glm_min(glm_max(c.y, 0.0f), 50.0f), glm_min(glm_max(c.z, 0.0f), 50.0f));
    ps.radiance = ps.radiance + ps.throughput * e;
}
