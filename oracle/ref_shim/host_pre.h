// TEST INFRASTRUCTURE (oracle/_ref build). Prelude that lets the reference's
// RayTracing/main_raytracing.cu compile as plain host C++ (g++), per SURVEY.md §8(c):
//  * cuRAND's device API is header-only; overriding QUALIFIERS makes it host-callable
//    (/usr/local/cuda/include/curand_kernel.h:60-62 guards the definition),
//  * the CUDA built-in index variables become thread_local globals so rows can run under OpenMP,
//  * texCubemapLod<float4> becomes the CPU cubemap sampler of oracle/cubemap_cpu.h.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <memory>
#include <span>
#include <stdexcept>
#include <vector>
#ifndef QUALIFIERS
#define QUALIFIERS static inline
#endif
#include <cuda_runtime.h>
#include <curand_kernel.h>
#include "../cubemap_cpu.h"

// curand_mtgp32_kernel.h:123-124 declares `extern const` blockDim/threadIdx for host builds, so the
// reference TU's uses are renamed (after the toolkit headers) onto thread_local stand-ins.
static thread_local uint3 ref_threadIdx, ref_blockIdx;
static thread_local dim3 ref_blockDim, ref_gridDim;
#define threadIdx ref_threadIdx
#define blockIdx ref_blockIdx
#define blockDim ref_blockDim
#define gridDim ref_gridDim

// launcher tail of raytracing_process (main_raytracing.cu:215) - no device, nothing to query
#define cudaGetLastError() cudaSuccess

template <class T> T texCubemapLod(cudaTextureObject_t tex, float x, float y, float z, float lod);
template <> inline float4 texCubemapLod<float4>(cudaTextureObject_t tex, float x, float y, float z, float)
{
    const OracleCubemap* cm = reinterpret_cast<const OracleCubemap*>(static_cast<uintptr_t>(tex));
    const OracleFloat4 v = oracle_tex_cubemap_lod0(*cm, x, y, z);
    return make_float4(v.x, v.y, v.z, v.w);
}

// C++ leaves the evaluation order of `vec2(rng.GetFloat01(), rng.GetFloat01())` (main_raytracing.cu:190) unspecified.
// nvcc's device code draws the x jitter first (PTX evidence: DESIGN.md "jitter order"); g++ would draw y first. The host
// build pins the device order through this helper (build_ref.sh substitutes the one expression).
template <class V, class R> static inline V ref_jitter_device_order(R& rng)
{
    const float jx = rng.GetFloat01();
    const float jy = rng.GetFloat01();
    return V(jx, jy);
}
