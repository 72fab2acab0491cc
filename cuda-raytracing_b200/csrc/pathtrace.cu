// Path-trace kernels: raytracing_kernel_main (RayTracing/main_raytracing.cu:162-200) rebuilt for
// sm_100a. Compiled with -fmad=false.
//
// variant 0 "regenerating megakernel": one lane owns one pixel for the whole launch, but the
//   reference's nested `for sample { for bounce { trace; shade } }` is flattened into ONE loop whose
//   body is exactly one closest-hit query: a lane whose path ended starts its next sample in the same
//   iteration, so every lane of a warp is inside the traversal loop until its samples run out
//   (SIMT efficiency no longer limited by the longest path per sample). Per pixel the order of RNG
//   draws and fp32 operations is the reference's, so the stream and the image are reproduced.
//   The XORWOW state lives in registers for the launch (reference: 24-B global RMW per draw).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include "pathtrace_common.cuh"

// kMode 0: product traversal. kMode 2: reference-order traversal with work counters (measurement of the algorithmic bytes per
// ray; same image up to exact-tie resolution). kMode 3: product traversal with work counters (wide-node steps and
// triangle tests actually executed: bench.py's bytes_per_ray_product_order; same image and RNG as kMode 0).
template <int kMode, int kMinBlocks>
__global__ void __launch_bounds__(RT_PT_BLOCK, kMinBlocks * (256 / RT_PT_BLOCK)) pathtrace_megakernel(AccelView A, SceneView S, FrameParams P)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // CTA = 16 x (RT_PT_BLOCK / 16) pixels, warp = 8 x 4
    unsigned bx = blockIdx.x, by = blockIdx.y;
    unsigned tile = 0;
    long long t_start = 0;
    if (P.tile_order)
    {   // 1-D grid, heaviest tiles (by the previous launch's measured cost) first: the launch ends on cheap tiles instead of on
        // whichever expensive ones happen to come last in raster order
        const unsigned tiles_x = ((unsigned)P.width + 15u) >> 4;
        tile = __ldg(P.tile_order + blockIdx.x);
        bx = tile % tiles_x; by = tile / tiles_x;
        t_start = clock64();
    }
    constexpr int kWarpsX = 16 / RT_PT_WARP_W, kWarpH = 32 / RT_PT_WARP_W;
    const int x = bx * 16 + (warp % kWarpsX) * RT_PT_WARP_W + (lane % RT_PT_WARP_W);
    const int y = by * (RT_PT_BLOCK / 16) + (warp / kWarpsX) * kWarpH + (lane / RT_PT_WARP_W);
    const bool active = x < P.width && y < P.height;
    unsigned rays = 0;
    uint32_t work[2] = { 0, 0 };
    if (active)
    {
        RtRngState* state = S.rng_state + (size_t)y * P.width + x;  // GPUScene::GetRNGForPixel, GPUScene.h:95
        Xorwow rng; rng.load(state);
        Vec3 sum = mk3(0, 0, 0);
        PathState ps;
        int sample = 0, bounce = 0;
        bool fresh = true;
        while (sample < P.spp)
        {
            if (fresh)
            {   // main_raytracing.cu:190-191; jitter draws x first, then y
                const float jx = rng.uniform(), jy = rng.uniform();
                const float u = ((float)x + jx) / (float)P.width, v = ((float)y + jy) / (float)P.height;
                camera_ray_pt(S.camera, u, v, ps.o, ps.d);
                ps.throughput = mk3(1, 1, 1); ps.radiance = mk3(0, 0, 0);
                bounce = 0; fresh = false;
            }
            const Vec3 nd = normalize3(ps.d);
            auto reinit = [&](HitRecord& r)
            {
                r.t = 1e30f; r.id = RT_MISS_ID; r.slot = 0; r.key = RT_MISS_ID; r.bu = 0; r.bv = 0;
                sphere_tests(S.spheres, S.sphere_count, ps.o, nd, r);
            };
            HitRecord h; reinit(h);
            if (kMode == 2) trace_reference_order<false>(A, ps.o, ps.d, nd, h, &work[0], &work[1]);
            else trace_wide<false>(A, ps.o, ps.d, nd, h, reinit, kMode == 3 ? work : nullptr);
            rays++;
            bool alive;
            if (h.id != RT_MISS_ID) alive = shade_hit(A, S, h, ps, rng) && (++bounce < P.max_bounces);
            else { if (S.env_tex) shade_miss(S, P.consts, ps); alive = false; }
            if (!alive) { sum = sum + ps.radiance; sample++; fresh = true; }
        }
        rng.store(state);
        write_pixel(P, x, y, sum);
    }
    if (P.tile_order && lane == 0)
    {
        const unsigned long long c = (unsigned long long)(clock64() - t_start) >> 4;
        atomicMax(P.tile_cost + tile, (uint32_t)(c > 0xFFFFFFFFull ? 0xFFFFFFFFull : c));
    }
    if (P.ray_counter)
    {
        for (int o = 16; o; o >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o);
        if (lane == 0 && rays) atomicAdd(P.ray_counter, (unsigned long long)rays);
    }
    if (kMode >= 2 && P.work_counters)
    {
        for (int o = 16; o; o >>= 1) { work[0] += __shfl_xor_sync(0xffffffffu, work[0], o); work[1] += __shfl_xor_sync(0xffffffffu, work[1], o); }
        if (lane == 0) { atomicAdd(P.work_counters, (unsigned long long)work[0]); atomicAdd(P.work_counters + 1, (unsigned long long)work[1]); }
    }
}

// ------------------------------------------------------------------------------------------------
// Cost-ordered tile scheduling. Per-tile cost varies ~10x (sky vs bunny silhouette) and the hardware hands out CTAs in grid order,
// so in raster order the launch ends whenever the last expensive tiles happen to finish while most SMs already idle (measured:
// the same frame takes 56.0 / 59.3 / 60.9 ms with tile rows scheduled bottom-up / top-down / strided). With a workspace, every
// launch records each tile's cost and the next launch schedules tiles heaviest-first (longest-processing-time-first list
// scheduling). Pixels are independent, so the image and the RNG streams do not depend on the order.
// A history is only as good as the camera that made it: when the camera moved since the costs were measured, ordering by the
// stale costs is WORSE than grid order (profiles/r02j_moving_camera.json: 47.5 vs 44.8 ms at 2 degrees per frame - the tiles that
// just became expensive sit anywhere in the order). The workspace therefore remembers a hash of the camera; after a move the costs
// are first DILATED (maximum over a 13 x 25-tile = 208 x 200-pixel neighbourhood: whatever was expensive is probably still
// expensive nearby), so the newly expensive tiles are scheduled early too: 42.2 ms (2 degrees per frame) and 39.4 ms (10 degrees)
// against 45.0 / 41.5 ms in grid order and 47.6 / 42.5 ms with the stale order; a static camera keeps its 44.9 (grid: 47.5).
// Workspace (uint32): [0] width, [1] height, [2] tile count, [3] launches with a history, [4] camera hash, [5..7] reserved
//                     | order[n] | cost[n] | scratch[n]; zero-filled = fresh.
// ------------------------------------------------------------------------------------------------
#define RT_TILE_HDR 8u
#define RT_TILE_BUCKETS 256
#ifndef RT_TILE_DILATE_X
#define RT_TILE_DILATE_X 6
#define RT_TILE_DILATE_Y 12
#endif

// one CTA: counting sort of the tiles into RT_TILE_BUCKETS cost classes, heaviest class first (order inside a class is arbitrary)
__global__ void __launch_bounds__(1024) tile_order_kernel(uint32_t* ws, uint32_t width, uint32_t height, uint32_t n, uint32_t tiles_x, uint32_t camera_hash)
{
    __shared__ uint32_t hist[RT_TILE_BUCKETS];
    __shared__ uint32_t s_max, s_fresh, s_moved;
    uint32_t* order = ws + RT_TILE_HDR; uint32_t* cost = order + n; uint32_t* scratch = cost + n;
    if (threadIdx.x == 0) { s_max = 0; s_fresh = ws[0] != width || ws[1] != height || ws[2] != n; s_moved = ws[4] != camera_hash; }
    for (uint32_t i = threadIdx.x; i < RT_TILE_BUCKETS; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    const bool fresh = s_fresh != 0u;
    if (!fresh)
    {
        uint32_t m = 0;
        for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) m = max(m, cost[i]);
        atomicMax(&s_max, m);
    }
    __syncthreads();
    const uint32_t mx = s_max;
    if (fresh || mx == 0)
    {   // no measurement yet (or another frame size): grid order
        for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) { order[i] = i; cost[i] = 0; }
        if (threadIdx.x == 0) { ws[0] = width; ws[1] = height; ws[2] = n; ws[3] = 0; ws[4] = camera_hash; }
        return;
    }
    if (s_moved != 0u && (RT_TILE_DILATE_X || RT_TILE_DILATE_Y))
    {   // separable maximum filter: rows into scratch, columns back into cost (the maximum over all tiles, mx, is unchanged)
        const uint32_t tiles_y = n / tiles_x;
        for (uint32_t i = threadIdx.x; i < n; i += blockDim.x)
        {
            const uint32_t tx = i % tiles_x, row = i - tx;
            uint32_t m = 0;
            for (uint32_t x = tx > RT_TILE_DILATE_X ? tx - RT_TILE_DILATE_X : 0; x <= min(tx + RT_TILE_DILATE_X, tiles_x - 1); x++) m = max(m, cost[row + x]);
            scratch[i] = m;
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < n; i += blockDim.x)
        {
            const uint32_t tx = i % tiles_x, ty = i / tiles_x;
            uint32_t m = 0;
            for (uint32_t y = ty > RT_TILE_DILATE_Y ? ty - RT_TILE_DILATE_Y : 0; y <= min(ty + RT_TILE_DILATE_Y, tiles_y - 1); y++) m = max(m, scratch[y * tiles_x + tx]);
            cost[i] = m;
        }
        __syncthreads();
    }
    const float scale = (float)(RT_TILE_BUCKETS - 1) / (float)mx;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x)
        atomicAdd(&hist[RT_TILE_BUCKETS - 1 - (uint32_t)((float)cost[i] * scale)], 1u);
    __syncthreads();
    if (threadIdx.x == 0)
    {
        uint32_t run = 0;
        for (int b = 0; b < RT_TILE_BUCKETS; b++) { const uint32_t c = hist[b]; hist[b] = run; run += c; }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x)
    {
        const uint32_t b = RT_TILE_BUCKETS - 1 - (uint32_t)((float)cost[i] * scale);
        order[atomicAdd(&hist[b], 1u)] = i;
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) cost[i] = 0;  // this launch's atomicMax starts from zero
    if (threadIdx.x == 0) { ws[3]++; ws[4] = camera_hash; }
}

uint64_t PathTraceWorkspaceBytes(int width, int height)
{
    const uint64_t n = (uint64_t)((width + 15) / 16) * (uint64_t)((height + RT_PT_BLOCK / 16 - 1) / (RT_PT_BLOCK / 16));
    return (RT_TILE_HDR + 3 * n) * sizeof(uint32_t);
}

cudaError_t LaunchPathTrace(const AccelView& accel, const SceneView& scene, const RtProcessParams& p, cudaStream_t stream)
{
    FrameParams P;
    P.surface = (unsigned char*)p.surface; P.surface_last = (unsigned char*)p.surface_last_frame;
    P.width = p.width; P.height = p.height; P.pitch = p.pitch; P.frame_index = p.frame_index;
    P.spp = p.samples_per_launch; P.max_bounces = p.max_bounces; P.blend_mode = p.blend_mode; P.write_last = p.write_last_frame;
    P.ray_counter = p.d_ray_counter; P.work_counters = p.d_work_counters;
    P.consts = MakePathConsts();
    const int tile_h = RT_PT_BLOCK / 16;
    dim3 grid((p.width + 15) / 16, (p.height + tile_h - 1) / tile_h), block(RT_PT_BLOCK);
    P.tile_order = nullptr; P.tile_cost = nullptr;
    // a workspace switches the cost-ordered schedule on (p.workspace == nullptr: grid order, also what the reference-named entry uses)
    if (p.variant == 0 && p.workspace && p.workspace_bytes >= PathTraceWorkspaceBytes(p.width, p.height))
    {
        const uint32_t n = grid.x * grid.y;
        uint32_t* ws = (uint32_t*)p.workspace;
        uint32_t camera_hash = 2166136261u;  // FNV-1a over the 60 bytes of the GPUCamera
        for (size_t k = 0; k < sizeof(RtCamera); k++) camera_hash = (camera_hash ^ reinterpret_cast<const unsigned char*>(&scene.camera)[k]) * 16777619u;
        if (camera_hash == 0) camera_hash = 1;
        tile_order_kernel<<<1, 1024, 0, stream>>>(ws, (uint32_t)p.width, (uint32_t)p.height, n, grid.x, camera_hash);
        P.tile_order = ws + RT_TILE_HDR; P.tile_cost = ws + RT_TILE_HDR + n;
        grid = dim3(n, 1);
    }
    if (p.variant == 2) pathtrace_megakernel<2, 3><<<grid, block, 0, stream>>>(accel, scene, P);
    else if (p.variant == 3) pathtrace_megakernel<3, 5><<<grid, block, 0, stream>>>(accel, scene, P);
    else pathtrace_megakernel<0, 5><<<grid, block, 0, stream>>>(accel, scene, P);  // 5 x 256 threads per SM (48 registers): 3 / 4 / 5 / 6 = 57.4 / 55.3 / 52.9 / 54.8 ms
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Ambient-occlusion stress (BASELINE.json config 3): incoherent rays from the reference's own diffuse
// lobe and origin offset (main_raytracing.cu:129,135). Flattened like the megakernel: one query per
// loop iteration (iteration 0 = the pixel-centre primary ray).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ambient_occlusion_kernel(AccelView A, SceneView S, int width, int height, int ao_rays,
                                                                float* __restrict__ out, unsigned long long* ray_counter)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int y = blockIdx.y * 16 + (warp >> 1) * 4 + (lane >> 3);
    unsigned rays = 0;
    if (x < width && y < height)
    {
        RtRngState* state = S.rng_state + (size_t)y * width + x;
        Xorwow rng; rng.load(state);
        Vec3 o, d; camera_ray_pt(S.camera, ((float)x + 0.5f) / (float)width, ((float)y + 0.5f) / (float)height, o, d);
        Vec3 base = o, normal = mk3(0, 0, 0);
        int unoccluded = 0, total = 0;
        for (int k = 0; k <= ao_rays; k++)
        {
            const Vec3 nd = normalize3(d);
            auto reinit = [&](HitRecord& r)
            {
                r.t = 1e30f; r.id = RT_MISS_ID; r.slot = 0; r.key = RT_MISS_ID; r.bu = 0; r.bv = 0;
                sphere_tests(S.spheres, S.sphere_count, o, nd, r);
            };
            HitRecord h; reinit(h);
            trace_wide<false>(A, o, d, nd, h, reinit);
            rays++;
            if (k == 0)
            {
                if (h.id == RT_MISS_ID) break;
                const Vec3 position = o + nd * h.t;
                if (h.id & RT_SPHERE_ID_BIT) { const float4 s = ldg4(&S.spheres[h.slot]); normal = (position - mk3(s.x, s.y, s.z)) / s.w; }
                else
                {
                    const ShadeRecord* sr = A.shade + h.slot;
                    const float4 n0 = ldg4(sr), n1 = ldg4(reinterpret_cast<const char*>(sr) + 16), n2 = ldg4(reinterpret_cast<const char*>(sr) + 32);
                    normal = normalize3((mk3(n0.x, n0.y, n0.z) * h.bu + mk3(n1.x, n1.y, n1.z) * h.bv) + mk3(n2.x, n2.y, n2.z) * (1.0f - h.bu - h.bv));
                    if (dot3(nd, normal) >= 0.0f) normal = -normal;
                }
                base = position + normal * 0.01f;
            }
            else { total++; if (h.id == RT_MISS_ID) unoccluded++; }
            if (k < ao_rays) { o = base; d = normalize3(normal + rng.point_on_sphere()); }  // two draws per AO ray, none after the last
        }
        rng.store(state);
        out[(size_t)y * width + x] = total ? (float)unoccluded / (float)total : 1.0f;
    }
    if (ray_counter)
    {
        for (int o2 = 16; o2; o2 >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o2);
        if (lane == 0 && rays) atomicAdd(ray_counter, (unsigned long long)rays);
    }
}

cudaError_t LaunchAmbientOcclusion(const AccelView& accel, const SceneView& scene, int width, int height, int ao_rays,
                                   float* d_ao, unsigned long long* d_ray_counter, cudaStream_t stream)
{
    const dim3 grid((width + 15) / 16, (height + 15) / 16), block(256);
    ambient_occlusion_kernel<<<grid, block, 0, stream>>>(accel, scene, width, height, ao_rays, d_ao, d_ray_counter);
    return cudaGetLastError();
}


// ------------------------------------------------------------------------------------------------
// rt_process_ex_static_scratch: rt_process_ex for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
// serving loops, harnesses that forbid it). No counterpart in the reference.
// The cost-ordered tile schedule's scratch (header + tile order + tile cost + filter scratch, RT_TILE_HDR + 3 n words) lives in a file-scope
// device array instead of a caller-allocated workspace: RT_STATIC_SCRATCH_BYTES = 1 MiB per GPU that loads the library, held
// until exit - room for 87 378 tiles of 16 x 8 pixels = 11.1 M pixels (a 4K frame has 64 800 tiles).
// In HOST code the array's name is a host-side stand-in, not a device address: its device address is taken once per device with
// cudaGetSymbolAddress while that device is current, and that pointer is what the launch receives.
// ------------------------------------------------------------------------------------------------
#define RT_STATIC_SCRATCH_BYTES (1u << 20)
static __device__ __align__(256) unsigned char g_static_scratch[RT_STATIC_SCRATCH_BYTES];

#include <mutex>
#include "capi_internal.h"

namespace
{
    constexpr int kMaxDevices = 64;
    std::mutex g_scratch_mutex;
    void* g_scratch_ptr[kMaxDevices] = {};  // per device: device address of g_static_scratch

    // device address of this device's copy of g_static_scratch (set-up: once per device, with that device current)
    int StaticScratch(void** out)
    {
        int dev = -1;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return rt_internal::CudaFail(e, "rt_process_ex_static_scratch: cudaGetDevice");
        if (dev < 0 || dev >= kMaxDevices) return rt_internal::Fail(RT_E_INVALID, "rt_process_ex_static_scratch: device ordinal beyond 63");
        std::lock_guard<std::mutex> lock(g_scratch_mutex);
        if (!g_scratch_ptr[dev])
        {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_static_scratch);
            if (e != cudaSuccess || !p) return rt_internal::CudaFail(e, "rt_process_ex_static_scratch: cudaGetSymbolAddress");
            g_scratch_ptr[dev] = p;
        }
        *out = g_scratch_ptr[dev];
        return RT_OK;
    }
}

extern "C" uint64_t rt_process_static_scratch_bytes(void) { return RT_STATIC_SCRATCH_BYTES; }

extern "C" int rt_process_ex_static_scratch(const RtProcessParams* params)
{
    using namespace rt_internal;
    if (!params) return Fail(RT_E_INVALID, "rt_process_ex_static_scratch: bad argument");
    if (params->variant != 0) return Fail(RT_E_INVALID, "rt_process_ex_static_scratch: only variant 0 (the megakernel) uses scratch memory");
    if (params->width <= 0 || params->height <= 0) return Fail(RT_E_INVALID, "rt_process_ex_static_scratch: bad argument");
    const uint64_t need = PathTraceWorkspaceBytes(params->width, params->height);
    if (need > RT_STATIC_SCRATCH_BYTES)
        return Fail(RT_E_INVALID, "rt_process_ex_static_scratch: frame needs " + std::to_string(need) + " bytes of tile scratch, the static array holds " +
                                      std::to_string(RT_STATIC_SCRATCH_BYTES) + " (87378 tiles of 16x8 pixels); use rt_process_ex with a caller-owned workspace");
    void* scratch = nullptr;
    if (int r = StaticScratch(&scratch)) return r;
    // The array is shared by every call on this device and keeps the last call's contents. What must start at zero does: the header
    // names the frame size the order / cost words belong to, tile_order_kernel treats any other size as "no history" (grid order,
    // costs zeroed) and zeroes the costs after every sort; a new accumulation (frame_index 0) drops the history explicitly so its
    // schedule never depends on what another caller rendered before.
    if (params->frame_index == 0)
        RT_CUDA(cudaMemsetAsync(scratch, 0, RT_TILE_HDR * sizeof(uint32_t), (cudaStream_t)params->stream));
    RtProcessParams p = *params;
    p.workspace = scratch; p.workspace_bytes = RT_STATIC_SCRATCH_BYTES;
    return rt_process_ex(&p);
}
