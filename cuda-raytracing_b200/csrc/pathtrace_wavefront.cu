// Queue-based wavefront path tracer (north-star item 3; rt_process_ex variant 10): the bounce loop of
// RayTracing/main_raytracing.cu:118-158 split into STAGES with the path state in HBM between them.
//
//   stage = one closest-hit query + its shading for every live pixel (the body of the megakernel's flattened loop);
//   between stages the surviving pixels are COMPACTED by warp ballot / prefix sum (match-any + popc + one atomicAdd per warp
//   and bin) into nine queues - fresh camera rays, and the eight direction octants of the bounce rays - so that the warps of
//   the next stage hold rays of one kind and one octant, packed 32 to a warp (no lane idles because its neighbour's path ended).
//
// A pixel's samples cannot run side by side: the reference's per-pixel XORWOW stream is sequential with a path-dependent draw
// count (Random.h, main_raytracing.cu:124-146,190), so the wave is over PIXELS and a stage advances every live pixel by one ray;
// a pixel whose path ended starts its next sample in the next stage (regeneration). A 64-spp x 4-bounce launch is at most 256
// stages; the stage count is not known on the host, so all of them are enqueued and the late ones find empty queues.
// Per pixel the sequence of draws and fp32 operations is the megakernel's: image, RNG states and ray count are bit-identical.
//
// HBM layout (workspace): WavePath 96 B per pixel (ray, throughput, radiance, sum, XORWOW words, sample / bounce), two sets of
// nine queues of pixel indices (capacity = pixels each), per-stage queue counters. Traffic per ray: 96 B read + 96 B written
// + 8 B of queue = 200 B, i.e. ~45 GB per 1080p 64-spp frame.
#include <algorithm>
#include "pathtrace_common.cuh"

namespace
{
    constexpr int kBins = 9;            // 0 = fresh camera rays, 1 + octant of the bounce direction
    constexpr int kCounterStride = 16;  // words per stage: kBins counts, [9] chunk cursor
    constexpr int kBlock = 128;

    struct alignas(16) WavePath
    {
        float ox, oy, oz, dx;
        float dy, dz, tx, ty;
        float tz, rx, ry, rz;
        float sx, sy, sz; int32_t sample;
        uint32_t rng_d, rng_v0, rng_v1, rng_v2;
        uint32_t rng_v3, rng_v4; int32_t bounce; uint32_t fresh;
    };
    static_assert(sizeof(WavePath) == 96, "WavePath layout");

    struct WaveParams
    {
        WavePath* paths;
        uint32_t* queues;      // [2][kBins][pixels]
        uint32_t* counters;    // [stages + 1][kCounterStride]
        uint32_t pixels;
    };

    __device__ __forceinline__ uint32_t* queue_of(const WaveParams& Q, int set, int bin) { return Q.queues + ((size_t)set * kBins + bin) * Q.pixels; }

    // every pixel starts in the camera-ray queue of stage 0, in 8 x 4 tile order (a warp = a compact pixel footprint)
    __global__ void __launch_bounds__(kBlock) wavefront_init_kernel(SceneView S, FrameParams P, WaveParams Q)
    {
        const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i == 0) Q.counters[0] = Q.pixels;  // stage 0: every slot of the tile order sits in the camera-ray queue
        if (i >= Q.pixels) return;
        const uint32_t tiles_x = ((uint32_t)P.width + 7u) >> 3;
        const uint32_t tile = i >> 5, in = i & 31u;
        const uint32_t x = (tile % tiles_x) * 8u + (in & 7u), y = (tile / tiles_x) * 4u + (in >> 3);
        // tiles cover a padded frame: pixels outside it are skipped by giving their queue slot the "none" marker
        uint32_t* q0 = queue_of(Q, 0, 0);
        const bool inside = x < (uint32_t)P.width && y < (uint32_t)P.height;
        q0[i] = inside ? y * (uint32_t)P.width + x : 0xFFFFFFFFu;
        if (!inside) return;
        const uint32_t pix = y * (uint32_t)P.width + x;
        Xorwow rng; rng.load(S.rng_state + pix);
        WavePath w{};
        w.sample = 0; w.bounce = 0; w.fresh = 1u;
        w.rng_d = rng.d; w.rng_v0 = rng.v0; w.rng_v1 = rng.v1; w.rng_v2 = rng.v2; w.rng_v3 = rng.v3; w.rng_v4 = rng.v4;
        Q.paths[pix] = w;
    }

    template <int kMinBlocks>
    __global__ void __launch_bounds__(kBlock, kMinBlocks) wavefront_stage_kernel(AccelView A, SceneView S, FrameParams P, WaveParams Q, int stage)
    {
        __shared__ uint32_t s_chunk;
        const uint32_t* cin = Q.counters + (size_t)stage * kCounterStride;
        uint32_t* cout = Q.counters + (size_t)(stage + 1) * kCounterStride;
        uint32_t* cursor = Q.counters + (size_t)stage * kCounterStride + kBins;
        const int set_in = stage & 1, set_out = set_in ^ 1;
        // bins padded to whole warps: a warp never mixes two bins
        uint32_t start[kBins + 1];
        start[0] = 0;
#pragma unroll
        for (int b = 0; b < kBins; b++) start[b + 1] = start[b] + ((cin[b] + 31u) & ~31u);
        const uint32_t total = start[kBins];
        const int lane = threadIdx.x & 31;
        unsigned rays = 0;
        for (;;)
        {
            if (threadIdx.x == 0) s_chunk = atomicAdd(cursor, 1u);
            __syncthreads();
            const uint32_t chunk = s_chunk;
            __syncthreads();
            if ((uint64_t)chunk * kBlock >= total) break;
            const uint32_t g = chunk * kBlock + threadIdx.x;
            int bin = 0;
#pragma unroll
            for (int b = 1; b < kBins; b++) bin += g >= start[b] ? 1 : 0;
            const uint32_t idx = g - start[bin];
            uint32_t pix = 0xFFFFFFFFu;
            if (g < total && idx < cin[bin]) pix = queue_of(Q, set_in, bin)[idx];
            int next_bin = -1;
            if (pix != 0xFFFFFFFFu)
            {
                WavePath* wp = Q.paths + pix;
                const float4 a0 = reinterpret_cast<const float4*>(wp)[0], a1 = reinterpret_cast<const float4*>(wp)[1], a2 = reinterpret_cast<const float4*>(wp)[2];
                const float4 a3 = reinterpret_cast<const float4*>(wp)[3];
                const uint4 a4 = reinterpret_cast<const uint4*>(wp)[4], a5 = reinterpret_cast<const uint4*>(wp)[5];
                PathState ps;
                ps.o = mk3(a0.x, a0.y, a0.z); ps.d = mk3(a0.w, a1.x, a1.y); ps.throughput = mk3(a1.z, a1.w, a2.x); ps.radiance = mk3(a2.y, a2.z, a2.w);
                Vec3 sum = mk3(a3.x, a3.y, a3.z);
                int sample = __float_as_int(a3.w), bounce = (int)a5.z;
                bool fresh = a5.w != 0u;
                Xorwow rng; rng.d = a4.x; rng.v0 = a4.y; rng.v1 = a4.z; rng.v2 = a4.w; rng.v3 = a5.x; rng.v4 = a5.y;
                const int x = (int)(pix % (uint32_t)P.width), y = (int)(pix / (uint32_t)P.width);
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
                trace_wide<false>(A, ps.o, ps.d, nd, h, reinit);
                rays++;
                bool alive;
                if (h.id != RT_MISS_ID) alive = shade_hit(A, S, h, ps, rng) && (++bounce < P.max_bounces);
                else { if (S.env_tex) shade_miss(S, P.consts, ps); alive = false; }
                if (!alive) { sum = sum + ps.radiance; sample++; fresh = true; }
                if (sample >= P.spp)
                {   // the pixel is done: its stream position goes back to the caller's states, its mean to the surface
                    rng.store(S.rng_state + pix);
                    write_pixel(P, x, y, sum);
                }
                else
                {
                    reinterpret_cast<float4*>(wp)[0] = make_float4(ps.o.x, ps.o.y, ps.o.z, ps.d.x);
                    reinterpret_cast<float4*>(wp)[1] = make_float4(ps.d.y, ps.d.z, ps.throughput.x, ps.throughput.y);
                    reinterpret_cast<float4*>(wp)[2] = make_float4(ps.throughput.z, ps.radiance.x, ps.radiance.y, ps.radiance.z);
                    reinterpret_cast<float4*>(wp)[3] = make_float4(sum.x, sum.y, sum.z, __int_as_float(sample));
                    reinterpret_cast<uint4*>(wp)[4] = make_uint4(rng.d, rng.v0, rng.v1, rng.v2);
                    reinterpret_cast<uint4*>(wp)[5] = make_uint4(rng.v3, rng.v4, (uint32_t)bounce, fresh ? 1u : 0u);
                    next_bin = fresh ? 0 : 1 + ((ps.d.x < 0.0f ? 1 : 0) | (ps.d.y < 0.0f ? 2 : 0) | (ps.d.z < 0.0f ? 4 : 0));
                }
            }
            // compaction: the lanes of this warp that go on, grouped by bin - one atomicAdd per warp and bin, ranks by popc
            const unsigned going = __ballot_sync(0xffffffffu, next_bin >= 0);
            if (next_bin >= 0)
            {
                const unsigned peers = __match_any_sync(going, next_bin);
                const int leader = __ffs(peers) - 1;
                uint32_t base = 0;
                if (lane == leader) base = atomicAdd(cout + next_bin, (uint32_t)__popc(peers));
                base = __shfl_sync(peers, base, leader);
                queue_of(Q, set_out, next_bin)[base + (uint32_t)__popc(peers & ((1u << lane) - 1u))] = pix;
            }
        }
        if (P.ray_counter)
        {
            for (int o = 16; o; o >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o);
            if (lane == 0 && rays) atomicAdd(P.ray_counter, (unsigned long long)rays);
        }
    }
}

constexpr int kMaxStages = 4096;  // spp x max_bounces of one launch (the counters are sized for it)

uint64_t PathTraceWavefrontWorkspaceBytes(int width, int height)
{
    const uint64_t tiles = (uint64_t)((width + 7) / 8) * ((height + 3) / 4);
    const uint64_t slots = tiles * 32;                       // queue capacity (stage 0 holds the padded tile order)
    const uint64_t pixels = (uint64_t)width * height;
    const uint64_t stages = kMaxStages + 1;
    return pixels * sizeof(WavePath) + 2ull * kBins * slots * 4 + stages * kCounterStride * 4 + 256;
}

cudaError_t LaunchPathTraceWavefront(const AccelView& accel, const SceneView& scene, const RtProcessParams& p, cudaStream_t stream)
{
    FrameParams P{};
    P.surface = (unsigned char*)p.surface; P.surface_last = (unsigned char*)p.surface_last_frame;
    P.width = p.width; P.height = p.height; P.pitch = p.pitch; P.frame_index = p.frame_index;
    P.spp = p.samples_per_launch; P.max_bounces = p.max_bounces; P.blend_mode = p.blend_mode; P.write_last = p.write_last_frame;
    P.ray_counter = p.d_ray_counter; P.work_counters = nullptr; P.tile_order = nullptr; P.tile_cost = nullptr;
    P.consts = MakePathConsts();
    const uint64_t tiles = (uint64_t)((p.width + 7) / 8) * ((p.height + 3) / 4);
    const uint64_t slots = tiles * 32, pixels = (uint64_t)p.width * p.height;
    const int stages = p.samples_per_launch * p.max_bounces;
    if (!p.workspace || p.workspace_bytes < PathTraceWavefrontWorkspaceBytes(p.width, p.height) || slots > 0xFFFFFFF0ull || stages > kMaxStages)
        return cudaErrorInvalidValue;
    WaveParams Q;
    unsigned char* ws = (unsigned char*)p.workspace;
    Q.paths = reinterpret_cast<WavePath*>(ws); ws += pixels * sizeof(WavePath);
    Q.queues = reinterpret_cast<uint32_t*>(ws); ws += 2ull * kBins * slots * 4;
    Q.counters = reinterpret_cast<uint32_t*>(ws);
    Q.pixels = (uint32_t)slots;
    cudaError_t e = cudaMemsetAsync(Q.counters, 0, (size_t)(stages + 1) * kCounterStride * 4, stream);
    if (e != cudaSuccess) return e;
    wavefront_init_kernel<<<(unsigned)((slots + kBlock - 1) / kBlock), kBlock, 0, stream>>>(scene, P, Q);
    static int sm_count = 0;
    if (!sm_count)
    {
        int dev = 0;
        if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
        if ((e = cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    }
    const unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)sm_count * 10, (slots + kBlock - 1) / kBlock);
    for (int s = 0; s < stages; s++)
        wavefront_stage_kernel<10><<<grid, kBlock, 0, stream>>>(accel, scene, P, Q, s);
    return cudaGetLastError();
}
