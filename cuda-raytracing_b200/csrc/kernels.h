// Launchers implemented in kernels.cu / pathtrace.cu (device code compiled with -fmad=false).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "accel.h"
#include "b200rt.h"

// scene pieces a kernel needs, by value (no struct copy of the 136-B GPUScene per thread)
struct SceneView
{
    const RtSphere* spheres;
    const RtMaterial* materials;
    int sphere_count;
    int material_count;
    RtRngState* rng_state;
    cudaTextureObject_t env_tex;
    RtCamera camera;
};

cudaError_t LaunchRngInit(RtRngState* states, uint64_t count, uint64_t seed, uint64_t first_subsequence,
                          const uint32_t* d_jump, int levels, cudaStream_t stream);
cudaError_t LaunchRngDraw(RtRngState* states, uint64_t count, uint32_t n, uint32_t* out, cudaStream_t stream);

cudaError_t LaunchTracePrimary(const AccelView& accel, const SceneView& scene, int width, int height, bool with_spheres,
                               int variant, uint32_t* d_id, float* d_t, unsigned long long* d_ray_counter, cudaStream_t stream);
cudaError_t LaunchTraceRays(const AccelView& accel, const SceneView& scene, const float* d_rays, uint64_t ray_count,
                            bool with_spheres, int variant, uint32_t* d_id, float* d_t, unsigned long long* d_ray_counter,
                            cudaStream_t stream);
// the count-work launcher additionally fills per-ray work counters (nodes popped, triangles tested)
// under the reference traversal order: d_counts[2*i], d_counts[2*i+1]
cudaError_t LaunchCountWork(const AccelView& accel, const SceneView& scene, int width, int height, bool with_spheres,
                            uint32_t* d_counts, cudaStream_t stream);

cudaError_t LaunchPathTrace(const AccelView& accel, const SceneView& scene, const RtProcessParams& p, cudaStream_t stream);
uint64_t PathTraceWorkspaceBytes(int width, int height);  // cost-ordered tile scheduling state (pathtrace.cu), zero-filled by the caller
// variant 1: regrouping persistent kernel (pathtrace_regroup.cu); d_pixel_counter = one device uint owned by the caller for this launch
cudaError_t LaunchPathTraceRegroup(const AccelView& accel, const SceneView& scene, const RtProcessParams& p, unsigned int* d_pixel_counter,
                                   cudaStream_t stream);
cudaError_t LaunchAmbientOcclusion(const AccelView& accel, const SceneView& scene, int width, int height, int ao_rays,
                                   float* d_ao, unsigned long long* d_ray_counter, cudaStream_t stream);
cudaError_t LaunchTonemap(const void* surface, int width, int height, uint64_t pitch, float divide_by, uint8_t* d_rgba8,
                          cudaStream_t stream);
cudaError_t LaunchResolveSum(const void* sum, void* out, int width, int height, uint64_t pitch_in, uint64_t pitch_out, cudaStream_t stream);
// variant 10: queue-based wavefront (pathtrace_wavefront.cu); workspace = PathTraceWavefrontWorkspaceBytes, any contents
cudaError_t LaunchPathTraceWavefront(const AccelView& accel, const SceneView& scene, const RtProcessParams& p, cudaStream_t stream);
uint64_t PathTraceWavefrontWorkspaceBytes(int width, int height);
