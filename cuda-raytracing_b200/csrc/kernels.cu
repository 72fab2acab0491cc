// RNG init, closest-hit query kernels, tone-map. Compiled with -fmad=false (see traverse.cuh).
#include <cstdlib>
#include "kernels.h"
#include "rng.cuh"
#include "traverse.cuh"

// ------------------------------------------------------------------------------------------------
// RNG: curand_init(seed, first_subsequence + i, 0)   (RayTracing/Random.cu:3-8)
// ------------------------------------------------------------------------------------------------
// Skip-ahead = product of jump matrices selected by the bits of the subsequence number. All matrices
// are powers of one matrix, so they commute: the bits shared by a whole 32-aligned warp (>= 5) are
// applied FIRST, while every lane still holds the same seed state - the data-dependent row selects
// are then warp-uniform and the loads broadcast. Only the low five bits run divergent.
__device__ __forceinline__ void jump_apply(uint32_t v[5], const uint32_t* __restrict__ m)
{
    uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0;
#pragma unroll 1
    for (int i = 0; i < 5; i++)
    {
        uint32_t bits = v[i];
        while (bits)
        {
            const int j = __ffs(bits) - 1;
            bits &= bits - 1;
            const uint32_t* row = m + (i * 32 + j) * 5;
            r0 ^= __ldg(row); r1 ^= __ldg(row + 1); r2 ^= __ldg(row + 2); r3 ^= __ldg(row + 3); r4 ^= __ldg(row + 4);
        }
    }
    v[0] = r0; v[1] = r1; v[2] = r2; v[3] = r3; v[4] = r4;
}

__global__ void __launch_bounds__(128) rng_init_kernel(RtRngState* __restrict__ states, uint64_t count, uint64_t seed,
                                                       uint64_t first_subsequence, const uint32_t* __restrict__ jump, int levels)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    // curand_kernel.h:772-790
    const uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
    const uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
    const uint32_t t0 = 1099087573u * s0;
    const uint32_t t1 = 2591861531u * s1;
    uint32_t v[5] = { 123456789u + t0, 362436069u ^ t0, 521288629u + t1, 88675123u ^ t1, 5783321u + t0 };
    const uint64_t n = first_subsequence + i;
    for (int k = levels - 1; k >= 0; k--)
        if ((n >> k) & 1ull) jump_apply(v, jump + (size_t)k * 800);
    RtRngState s;
    s.d = 6615241u + t1 + t0;  // a subsequence is a multiple of 2^32 draws: d is unchanged (curand_kernel.h:746)
    s.v[0] = v[0]; s.v[1] = v[1]; s.v[2] = v[2]; s.v[3] = v[3]; s.v[4] = v[4];
    s.boxmuller_flag = 0; s.boxmuller_flag_double = 0; s.boxmuller_extra = 0.f; s.boxmuller_extra_double = 0.;
    states[i] = s;
}

__global__ void rng_draw_kernel(RtRngState* states, uint64_t count, uint32_t n, uint32_t* out)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    Xorwow g; g.load(states + i);
    for (uint32_t k = 0; k < n; k++) out[i * n + k] = g.next();
    g.store(states + i);
}

cudaError_t LaunchRngInit(RtRngState* states, uint64_t count, uint64_t seed, uint64_t first_subsequence,
                          const uint32_t* d_jump, int levels, cudaStream_t stream)
{
    if (!count) return cudaSuccess;
    const unsigned blocks = (unsigned)((count + 127) / 128);
    rng_init_kernel<<<blocks, 128, 0, stream>>>(states, count, seed, first_subsequence, d_jump, levels);
    return cudaGetLastError();
}

cudaError_t LaunchRngDraw(RtRngState* states, uint64_t count, uint32_t n, uint32_t* out, cudaStream_t stream)
{
    if (!count) return cudaSuccess;
    rng_draw_kernel<<<(unsigned)((count + 127) / 128), 128, 0, stream>>>(states, count, n, out);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Closest-hit queries
// ------------------------------------------------------------------------------------------------
// GPUCamera::GetRay (RayTracing/GPUScene.h:13): direction NOT normalised.
__device__ __forceinline__ void camera_ray(const RtCamera& cam, float u, float v, Vec3& o, Vec3& d)
{
    o = mk3(cam.origin.x, cam.origin.y, cam.origin.z);
    const Vec3 h = mk3(cam.horizontal.x, cam.horizontal.y, cam.horizontal.z), w = mk3(cam.vertical.x, cam.vertical.y, cam.vertical.z);
    const Vec3 llc = mk3(cam.lower_left_corner.x, cam.lower_left_corner.y, cam.lower_left_corner.z);
    d = ((llc + u * h) + v * w) - o;
}

template <int kVariant>
__device__ __forceinline__ void closest_hit(const AccelView& A, const SceneView& S, bool with_spheres, Vec3 o, Vec3 d, HitRecord& h)
{
    const Vec3 nd = normalize3(d);
    auto reinit = [&](HitRecord& r)
    {
        r.t = 1e30f; r.id = RT_MISS_ID; r.slot = 0; r.key = RT_MISS_ID; r.bu = 0; r.bv = 0;
        if (with_spheres) sphere_tests(S.spheres, S.sphere_count, o, nd, r);
    };
    reinit(h);
    if (kVariant == 0 && dot3(d, d) >= RT_SHORT_DIR2) trace_wide<true>(A, o, d, nd, h, reinit);
    else trace_reference_order<true>(A, o, d, nd, h);
}

// One warp covers an 8x4 pixel tile (coherent primary rays share nodes -> broadcast loads).
template <int kVariant>
__global__ void __launch_bounds__(256) trace_primary_kernel(AccelView A, SceneView S, int width, int height, bool with_spheres,
                                                            uint32_t* __restrict__ out_id, float* __restrict__ out_t,
                                                            unsigned long long* ray_counter)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int y = blockIdx.y * 16 + (warp >> 1) * 4 + (lane >> 3);
    if (x >= width || y >= height) return;
    const float u = ((float)x + 0.5f) / (float)width, v = ((float)y + 0.5f) / (float)height;
    Vec3 o, d; camera_ray(S.camera, u, v, o, d);
    HitRecord h; closest_hit<kVariant>(A, S, with_spheres, o, d, h);
    const size_t i = (size_t)y * width + x;
    out_id[i] = h.id; out_t[i] = h.t;
    if (ray_counter && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) atomicAdd(ray_counter, (unsigned long long)width * height);
}

template <int kVariant>
__global__ void __launch_bounds__(256) trace_rays_kernel(AccelView A, SceneView S, const float* __restrict__ rays, uint64_t ray_count,
                                                         bool with_spheres, uint32_t* __restrict__ out_id, float* __restrict__ out_t,
                                                         unsigned long long* ray_counter)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ray_count) return;
    const Vec3 o = mk3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]), d = mk3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
    HitRecord h; closest_hit<kVariant>(A, S, with_spheres, o, d, h);
    out_id[i] = h.id; out_t[i] = h.t;
    if (ray_counter && i == 0) atomicAdd(ray_counter, (unsigned long long)ray_count);
}

__global__ void __launch_bounds__(256) count_work_kernel(AccelView A, SceneView S, int width, int height, bool with_spheres, uint32_t* __restrict__ counts)
{
    const int x = blockIdx.x * 16 + (threadIdx.x & 15), y = blockIdx.y * 16 + (threadIdx.x >> 4);
    if (x >= width || y >= height) return;
    const float u = ((float)x + 0.5f) / (float)width, v = ((float)y + 0.5f) / (float)height;
    Vec3 o, d; camera_ray(S.camera, u, v, o, d);
    const Vec3 nd = normalize3(d);
    HitRecord h; h.t = 1e30f; h.id = RT_MISS_ID; h.slot = 0; h.key = RT_MISS_ID; h.bu = 0; h.bv = 0;
    if (with_spheres) sphere_tests(S.spheres, S.sphere_count, o, nd, h);
    uint32_t nodes = 0, tris = 0;
    trace_reference_order<true>(A, o, d, nd, h, &nodes, &tris);
    const size_t i = (size_t)y * width + x;
    counts[2 * i] = nodes; counts[2 * i + 1] = tris;
}

cudaError_t LaunchTracePrimary(const AccelView& accel, const SceneView& scene, int width, int height, bool with_spheres,
                               int variant, uint32_t* d_id, float* d_t, unsigned long long* d_ray_counter, cudaStream_t stream)
{
    const dim3 grid((width + 15) / 16, (height + 15) / 16), block(256);
    if (variant == 0) trace_primary_kernel<0><<<grid, block, 0, stream>>>(accel, scene, width, height, with_spheres, d_id, d_t, d_ray_counter);
    else trace_primary_kernel<1><<<grid, block, 0, stream>>>(accel, scene, width, height, with_spheres, d_id, d_t, d_ray_counter);
    return cudaGetLastError();
}

cudaError_t LaunchTraceRays(const AccelView& accel, const SceneView& scene, const float* d_rays, uint64_t ray_count,
                            bool with_spheres, int variant, uint32_t* d_id, float* d_t, unsigned long long* d_ray_counter,
                            cudaStream_t stream)
{
    if (!ray_count) return cudaSuccess;
    const unsigned blocks = (unsigned)((ray_count + 255) / 256);
    if (variant == 0) trace_rays_kernel<0><<<blocks, 256, 0, stream>>>(accel, scene, d_rays, ray_count, with_spheres, d_id, d_t, d_ray_counter);
    else trace_rays_kernel<1><<<blocks, 256, 0, stream>>>(accel, scene, d_rays, ray_count, with_spheres, d_id, d_t, d_ray_counter);
    return cudaGetLastError();
}

cudaError_t LaunchCountWork(const AccelView& accel, const SceneView& scene, int width, int height, bool with_spheres,
                            uint32_t* d_counts, cudaStream_t stream)
{
    const dim3 grid((width + 15) / 16, (height + 15) / 16), block(256);
    count_work_kernel<<<grid, block, 0, stream>>>(accel, scene, width, height, with_spheres, d_counts);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Display transform: main.cpp:78-94 (HLSL PS: exposure 0.5, ACES fit) + sRGB encode of the
// R8G8B8A8_UNORM_SRGB swap-chain (main.cpp:438) + vertical flip (row 0 of the surface = bottom).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float aces(float x)
{
    const float a = 2.51f, b = 0.03f, c = 2.43f, d = 0.59f, e = 0.14f;
    return __saturatef((x * (a * x + b)) / (x * (c * x + d) + e));
}
__device__ __forceinline__ float srgb_encode(float c) { return c <= 0.0031308f ? 12.92f * c : 1.055f * powf(c, 1.0f / 2.4f) - 0.055f; }

__global__ void tonemap_kernel(const unsigned char* __restrict__ surface, int width, int height, uint64_t pitch, float inv_div, uint8_t* __restrict__ out)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= width || y >= height) return;
    const float4 p = *reinterpret_cast<const float4*>(surface + (size_t)y * pitch + (size_t)x * 16);
    const float r = srgb_encode(aces(p.x * inv_div * 0.5f)), g = srgb_encode(aces(p.y * inv_div * 0.5f)), b = srgb_encode(aces(p.z * inv_div * 0.5f));
    uchar4 q; q.x = (unsigned char)(r * 255.0f + 0.5f); q.y = (unsigned char)(g * 255.0f + 0.5f); q.z = (unsigned char)(b * 255.0f + 0.5f); q.w = 255;
    reinterpret_cast<uchar4*>(out)[(size_t)(height - 1 - y) * width + x] = q;
}

cudaError_t LaunchTonemap(const void* surface, int width, int height, uint64_t pitch, float divide_by, uint8_t* d_rgba8, cudaStream_t stream)
{
    const dim3 block(32, 8), grid((width + 31) / 32, (height + 7) / 8);
    tonemap_kernel<<<grid, block, 0, stream>>>((const unsigned char*)surface, width, height, pitch, 1.0f / divide_by, d_rgba8);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Resolve of a sum buffer (blend_mode 1: rgb = sum of sample radiance, alpha = sample count; after the multi-GPU reduce the sums
// of all ranks): mean = rgb / alpha, alpha = 1 - the surface contract of main_raytracing.cu:195-199.
// ------------------------------------------------------------------------------------------------
__global__ void resolve_sum_kernel(const unsigned char* __restrict__ sum, unsigned char* __restrict__ out, int width, int height, uint64_t pitch_in, uint64_t pitch_out)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= width || y >= height) return;
    const float4 s = *reinterpret_cast<const float4*>(sum + (size_t)y * pitch_in + (size_t)x * 16);
    const float n = s.w > 0.0f ? s.w : 1.0f;
    *reinterpret_cast<float4*>(out + (size_t)y * pitch_out + (size_t)x * 16) = make_float4(s.x / n, s.y / n, s.z / n, 1.0f);
}

cudaError_t LaunchResolveSum(const void* sum, void* out, int width, int height, uint64_t pitch_in, uint64_t pitch_out, cudaStream_t stream)
{
    const dim3 block(32, 8), grid((width + 31) / 32, (height + 7) / 8);
    resolve_sum_kernel<<<grid, block, 0, stream>>>((const unsigned char*)sum, (unsigned char*)out, width, height, pitch_in, pitch_out);
    return cudaGetLastError();
}
