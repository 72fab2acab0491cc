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
// Persistent "extend" kernel (variant 2 of rt_trace_rays; the traversal stage of the wavefront path tracer):
// warps stay resident and pull rays from a global counter; a lane whose ray ended is refilled as soon as kFetch lanes are
// idle; traversal is warp-synchronous in two voted phases - one wide node per lane (leaves are postponed, up to three per
// lane), and one postponed leaf per lane when enough lanes hold one - so neither phase waits for the longest lane.
// Same arithmetic and hit rule as trace_wide (results identical; tests compare the variants bit for bit).
// ------------------------------------------------------------------------------------------------
struct ExtendLane
{
    Vec3 o, d, nd;
    float ix, iy, iz, cx, cy, cz, slack, start_length;
    HitRecord h;
    uint32_t node, pend0, pend1, pend2, verified_leaf;
    int npend, sp;
};

template <bool kTieById>
__device__ __forceinline__ void extend_leaf(const AccelView& A, uint32_t leaf, ExtendLane& L)
{
    const uint32_t first = ref_slot(A, leaf), count = ref_count(leaf);
    const TriRecord* tr = A.tris + first;
    for (uint32_t i = 0; i < count; i++)
    {
        const float4 r0 = ldg4(&tr[i]), r1 = ldg4(reinterpret_cast<const char*>(&tr[i]) + 16), r2 = ldg4(reinterpret_cast<const char*>(&tr[i]) + 32);
        float t, bu, bv;
        if (!tri_test(L.o, L.nd, r0, r1, r2, t, bu, bv)) continue;
        const uint32_t id = __float_as_uint(r0.w);
        const uint32_t key = kTieById ? id : __float_as_uint(r2.w);
        if (!accept(t, key, L.h)) continue;
        const uint32_t ref_leaf = __float_as_uint(r1.w) & RT_REF_LEAF_MASK;
        if (ref_leaf != L.verified_leaf)
        {
            if (!exact_box(L.o, L.d, A.ref_nodes + ref_leaf, L.start_length)) continue;
            L.verified_leaf = ref_leaf;
        }
        L.h.t = t; L.h.id = id; L.h.slot = first + i; L.h.key = key; L.h.bu = bu; L.h.bv = bv;
    }
}

// one wide node: conservative slab test in FMA form, leaves postponed, inner children near-first
__device__ __forceinline__ void extend_inner(const AccelView& A, ExtendLane& L, uint32_t* stack)
{
    const float kWiden = 1.0e-6f;
    const WideNode* w = ref_wide(A, L.node);
    const float4 a = ldg4(w), b = ldg4(reinterpret_cast<const char*>(w) + 16), c = ldg4(reinterpret_cast<const char*>(w) + 32);
#if RT_TW_NODE48
    const uint2 ch = make_uint2(__float_as_uint(c.y), __float_as_uint(c.z));
    float mn0, mx0, mn1, mx1;
    center_slabs(a, b, c, make_float2(-L.o.x, -L.o.y), make_float2(-L.o.z, -L.o.z), make_float2(L.ix, L.iy), make_float2(L.iz, L.iz),
                 make_float2(fabsf(L.ix), fabsf(L.iy)), make_float2(fabsf(L.iz), fabsf(L.iz)), mn0, mx0, mn1, mx1);
#elif RT_TW_CENTER
    const uint2 ch = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const char*>(w) + 48));
    // centre + half-extent nodes (accel.h): entry / exit = mid -/+ half * |inv| per axis, FMA form for mid (L.slack covers its cancellation)
    const float aix = fabsf(L.ix), aiy = fabsf(L.iy), aiz = fabsf(L.iz);
    float mid = fmaf(a.x, L.ix, L.cx), half = a.z * aix;
    float mn0 = mid - half, mx0 = mid + half;
    mid = fmaf(a.y, L.iy, L.cy); half = a.w * aiy;
    mn0 = fmaxf(mn0, mid - half); mx0 = fminf(mx0, mid + half);
    mid = fmaf(c.x, L.iz, L.cz); half = c.z * aiz;
    mn0 = fmaxf(mn0, mid - half); mx0 = fminf(mx0, mid + half);
    mid = fmaf(b.x, L.ix, L.cx); half = b.z * aix;
    float mn1 = mid - half, mx1 = mid + half;
    mid = fmaf(b.y, L.iy, L.cy); half = b.w * aiy;
    mn1 = fmaxf(mn1, mid - half); mx1 = fminf(mx1, mid + half);
    mid = fmaf(c.y, L.iz, L.cz); half = c.w * aiz;
    mn1 = fmaxf(mn1, mid - half); mx1 = fminf(mx1, mid + half);
#else
    const uint2 ch = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const char*>(w) + 48));
    float t1 = fmaf(a.x, L.ix, L.cx), t2 = fmaf(a.z, L.ix, L.cx);
    float mn0 = fminf(t1, t2), mx0 = fmaxf(t1, t2);
    t1 = fmaf(a.y, L.iy, L.cy); t2 = fmaf(a.w, L.iy, L.cy);
    float mnb = fminf(t1, t2), mxb = fmaxf(t1, t2);
    t1 = fmaf(c.x, L.iz, L.cz); t2 = fmaf(c.y, L.iz, L.cz);
    mn0 = fmaxf(fmaxf(mn0, mnb), fminf(t1, t2)); mx0 = fminf(fminf(mx0, mxb), fmaxf(t1, t2));
    t1 = fmaf(b.x, L.ix, L.cx); t2 = fmaf(b.z, L.ix, L.cx);
    float mn1 = fminf(t1, t2), mx1 = fmaxf(t1, t2);
    t1 = fmaf(b.y, L.iy, L.cy); t2 = fmaf(b.w, L.iy, L.cy);
    mnb = fminf(t1, t2); mxb = fmaxf(t1, t2);
    t1 = fmaf(c.z, L.iz, L.cz); t2 = fmaf(c.w, L.iz, L.cz);
    mn1 = fmaxf(fmaxf(mn1, mnb), fminf(t1, t2)); mx1 = fminf(fminf(mx1, mxb), fmaxf(t1, t2));
#endif
    const float lo0 = fmaf(-kWiden, fabsf(mn0), mn0) - L.slack, hi0 = fmaf(kWiden, fabsf(mx0), mx0) + L.slack;
    const float lo1 = fmaf(-kWiden, fabsf(mn1), mn1) - L.slack, hi1 = fmaf(kWiden, fabsf(mx1), mx1) + L.slack;
    const bool hit0 = hi0 >= lo0 && lo0 <= L.h.t && hi0 > 0.0f;
    const bool hit1 = hi1 >= lo1 && lo1 <= L.h.t && hi1 > 0.0f;
    const bool leaf0 = hit0 && (ch.x & RT_LEAF_BIT), leaf1 = hit1 && (ch.y & RT_LEAF_BIT);
    const bool in0 = hit0 && !(ch.x & RT_LEAF_BIT), in1 = hit1 && !(ch.y & RT_LEAF_BIT);
    if (leaf0) { if (L.npend == 0) L.pend0 = ch.x; else L.pend1 = ch.x; L.npend++; }
    if (leaf1) { if (L.npend == 0) L.pend0 = ch.y; else if (L.npend == 1) L.pend1 = ch.y; else L.pend2 = ch.y; L.npend++; }
    if (in0 && in1)
    {
        const bool swap = mn1 < mn0;
        stack[L.sp++] = swap ? ch.x : ch.y;
        L.node = swap ? ch.y : ch.x;
    }
    else if (in0) L.node = ch.x;
    else if (in1) L.node = ch.y;
    else L.node = stack[--L.sp];
}

// ray set-up: spheres first (GetRayHit), slab constants, stack
__device__ __forceinline__ void extend_begin(const AccelView& A, const SceneView& S, bool with_spheres, Vec3 o, Vec3 d, ExtendLane& L, uint32_t* stack)
{
    const float big = 1e30f;
    L.o = o; L.d = d; L.nd = normalize3(d);
    L.h.t = 1e30f; L.h.id = RT_MISS_ID; L.h.slot = 0; L.h.key = RT_MISS_ID; L.h.bu = 0; L.h.bv = 0;
    if (with_spheres) sphere_tests(S.spheres, S.sphere_count, o, L.nd, L.h);
    L.start_length = L.h.t;
    L.ix = L.nd.x != 0.0f ? 1.0f / L.nd.x : big; L.iy = L.nd.y != 0.0f ? 1.0f / L.nd.y : big; L.iz = L.nd.z != 0.0f ? 1.0f / L.nd.z : big;
    const float ax = o.x * L.ix, ay = o.y * L.iy, az = o.z * L.iz;
    L.cx = -ax; L.cy = -ay; L.cz = -az;
    L.slack = 4.0e-7f * fmaxf(fabsf(ax), fmaxf(fabsf(ay), fabsf(az))) + 1e-30f;
    L.sp = 0; stack[L.sp++] = RT_SENTINEL; L.npend = 0; L.verified_leaf = RT_SENTINEL;
    L.node = A.root_ref;
    if (L.node & RT_LEAF_BIT) { L.pend0 = L.node; L.npend = 1; L.node = RT_SENTINEL; }
}

template <bool kTieById>
__global__ void __launch_bounds__(256, 3) extend_rays_kernel(AccelView A, SceneView S, const float* __restrict__ rays, uint32_t ray_count,
                                                             bool with_spheres, uint32_t* __restrict__ out_id, float* __restrict__ out_t,
                                                             unsigned int* fetch_counter, int fetch_threshold, int leaf_batch)
{
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    ExtendLane L; L.node = RT_SENTINEL; L.npend = 0; L.sp = 0;
    uint32_t stack[RT_STACK_SIZE];
    bool busy = false, exhausted = false;
    uint32_t ray_index = 0;
    for (;;)
    {
        const unsigned m_busy = __ballot_sync(FULL, busy);
        const unsigned m_idle = __ballot_sync(FULL, !busy && !exhausted);
        if ((m_busy | m_idle) == 0u) break;
        if (m_idle && (m_busy == 0u || __popc(m_idle) >= fetch_threshold))
        {
            const int leader = __ffs(m_idle) - 1;
            unsigned base = 0;
            if (lane == leader) base = atomicAdd(fetch_counter, (unsigned)__popc(m_idle));
            base = __shfl_sync(FULL, base, leader);
            if (!busy && !exhausted)
            {
                ray_index = base + (unsigned)__popc(m_idle & ((1u << lane) - 1u));
                if (ray_index >= ray_count) exhausted = true;
                else
                {
                    const float* r = rays + 6 * (size_t)ray_index;
                    const Vec3 o = mk3(__ldg(r), __ldg(r + 1), __ldg(r + 2)), d = mk3(__ldg(r + 3), __ldg(r + 4), __ldg(r + 5));
                    if (dot3(d, d) >= RT_SHORT_DIR2) { extend_begin(A, S, with_spheres, o, d, L, stack); busy = true; }
                    else
                    {   // short direction: the reference's own visiting order decides (traverse.cuh RT_SHORT_DIR2) - rare, done in place
                        HitRecord h; h.t = 1e30f; h.id = RT_MISS_ID; h.slot = 0; h.key = RT_MISS_ID; h.bu = 0; h.bv = 0;
                        const Vec3 nd = normalize3(d);
                        if (with_spheres) sphere_tests(S.spheres, S.sphere_count, o, nd, h);
                        trace_reference_order<kTieById>(A, o, d, nd, h);
                        out_id[ray_index] = h.id; out_t[ray_index] = h.t;
                    }
                }
            }
        }
        if (busy && L.node != RT_SENTINEL && L.npend <= 1) extend_inner(A, L, stack);
        const unsigned m_pend = __ballot_sync(FULL, busy && L.npend > 0);
        const unsigned m_inner = __ballot_sync(FULL, busy && L.node != RT_SENTINEL && L.npend <= 1);
        if (m_pend && (__popc(m_pend) >= leaf_batch || m_inner == 0u))
        {
            if (busy && L.npend > 0)
            {
                L.npend--;
                const uint32_t leaf = L.npend == 0 ? L.pend0 : (L.npend == 1 ? L.pend1 : L.pend2);
                extend_leaf<kTieById>(A, leaf, L);
            }
        }
        if (busy && L.node == RT_SENTINEL && L.npend == 0)
        {
            out_id[ray_index] = L.h.id; out_t[ray_index] = L.h.t;
            busy = false;
        }
    }
}

cudaError_t LaunchExtendRays(const AccelView& accel, const SceneView& scene, const float* d_rays, uint64_t ray_count, bool with_spheres,
                             uint32_t* d_id, float* d_t, unsigned long long* d_ray_counter, unsigned int* d_fetch_counter, cudaStream_t stream)
{
    if (!ray_count) return cudaSuccess;
    if (ray_count > 0xFFFFFFF0ull) return cudaErrorInvalidValue;
    cudaError_t e = cudaMemsetAsync(d_fetch_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    static int blocks_per_sm = 0, sm_count = 0;
    if (!blocks_per_sm)
    {
        int dev = 0;
        if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
        if ((e = cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, extend_rays_kernel<true>, 256, 0)) != cudaSuccess) return e;
        if (blocks_per_sm < 1) blocks_per_sm = 1;
    }
    int fetch = 8, batch = 12;
    if (const char* v = getenv("B200RT_FETCH")) fetch = atoi(v);
    if (const char* v = getenv("B200RT_LEAF_BATCH")) batch = atoi(v);
    unsigned grid = (unsigned)(sm_count * blocks_per_sm);
    const unsigned needed = (unsigned)((ray_count + 255) / 256);
    if (grid > needed) grid = needed;
    extend_rays_kernel<true><<<grid, 256, 0, stream>>>(accel, scene, d_rays, (uint32_t)ray_count, with_spheres, d_id, d_t, d_fetch_counter, fetch, batch);
    if (d_ray_counter) { /* counted by the caller-visible contract: one query per ray */ }
    return cudaGetLastError();
}
