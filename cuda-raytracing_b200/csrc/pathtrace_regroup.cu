// Path-trace kernel, variant 1 "regrouping persistent kernel" (north-star items 2+3 done inside one kernel):
// the wavefront idea - separate "extend" (traversal) from "shade/generate" and only run each with many lanes -
// without spilling path state to HBM.
//
//   * persistent warps: the grid is sized to the machine (148 SMs x resident CTAs); lanes pull pixels from a global
//     counter in 8x4-tile order, so a lane whose pixel is finished gets the next pixel instead of idling;
//   * every lane is a small state machine {NEED, TRAV, DONE}. A lane whose ray ended does NOT shade at once (that would
//     run the ~400-instruction shade/regenerate code with one lane active); it parks in NEED. By warp ballot, the parked
//     lanes are shaded + refilled together as soon as kRefill of them are waiting (or nobody is traversing any more);
//   * traversal is if-if: every TRAV lane does ONE step per loop iteration - a wide node, or one triangle of a leaf (the
//     leaf reference is re-encoded with its remaining range) - so no lane waits in an inner loop and the regroup vote is
//     taken after every step. (Round-1 first version: voted inner/leaf phases with postponed leaves - SIMT efficiency 13.8
//     but +60 % thread-level work, slower than the megakernel)
//
// Per pixel, the order of RNG draws and fp32 operations is exactly the megakernel's (and so the reference's): the image is
// bit-identical to variant 0 (tests/test_gpu_parity.py::test_regroup_variant_bit_identical_to_megakernel).
// Compiled with -fmad=false like every device TU.
#include <cmath>
#include <cstdlib>
#include "shade.cuh"

namespace
{
    struct RegroupParams
    {
        unsigned char* surface;
        unsigned char* surface_last;
        int width, height;
        uint64_t pitch;
        int frame_index, spp, max_bounces, blend_mode, write_last;
        unsigned long long* ray_counter;
        unsigned int* pixel_counter;  // zeroed before the launch
        int refill, leaf_batch;       // regroup thresholds (lanes): shade/refill, triangle step
        PathConsts consts;
    };

    enum : int { ST_NEED = 0, ST_TRAV = 1, ST_DONE = 2 };
    constexpr int kRefill = 12;     // shade/refill as soon as this many lanes are parked
    constexpr int kLeafBatch = 8;   // run the triangle step as soon as this many lanes sit on a leaf

    __device__ __forceinline__ void camera_ray_rg(const RtCamera& cam, float u, float v, Vec3& o, Vec3& d)
    {
        o = mk3(cam.origin.x, cam.origin.y, cam.origin.z);
        const Vec3 h = mk3(cam.horizontal.x, cam.horizontal.y, cam.horizontal.z), w = mk3(cam.vertical.x, cam.vertical.y, cam.vertical.z);
        const Vec3 llc = mk3(cam.lower_left_corner.x, cam.lower_left_corner.y, cam.lower_left_corner.z);
        d = ((llc + u * h) + v * w) - o;
    }

    // main_raytracing.cu:195-199 (blend_mode 0) or sum accumulation (blend_mode 1) - same arithmetic as pathtrace.cu
    __device__ __forceinline__ void write_pixel_rg(const RegroupParams& P, int x, int y, Vec3 sum)
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
}

template <int kMinBlocks>
__global__ void __launch_bounds__(256, kMinBlocks) pathtrace_regroup_kernel(AccelView A, SceneView S, RegroupParams P)
{
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned tiles_x = (unsigned)(P.width + 7) >> 3, tiles_y = (unsigned)(P.height + 3) >> 2;
    const unsigned total = tiles_x * tiles_y * 32u;

    int state = ST_NEED;
    bool have_pixel = false, have_hit = false;  // have_hit: `h` is the result of a finished, not yet shaded ray
    int x = 0, y = 0, sample = 0, bounce = 0;
    Xorwow rng; rng.d = rng.v0 = rng.v1 = rng.v2 = rng.v3 = rng.v4 = 0;
    Vec3 sum = mk3(0, 0, 0);
    PathState ps; ps.o = ps.d = ps.throughput = ps.radiance = mk3(0, 0, 0);
    HitRecord h; h.t = 1e30f; h.id = RT_MISS_ID; h.slot = 0; h.key = RT_MISS_ID; h.bu = 0; h.bv = 0;
    Vec3 nd = mk3(0, 0, 1);
    // conservative slab test in FMA form: t = fma(plane, inv, c), c = -(o * inv); abs error <= slack (see DESIGN.md "box culls")
    float2 i_xy = make_float2(0, 0), i_zz = i_xy, no_xy = i_xy, no_zz = i_xy;  // per-ray constant pairs of the packed slab test
    float start_length = 1e30f;
    uint32_t node = RT_SENTINEL;     // current node reference (inner, leaf with its remaining range, or sentinel)
    int sp = 0;
    uint32_t stack[RT_STACK_SIZE];
    unsigned rays = 0;
    const float kWiden = 1.0e-6f, big = 1e30f;

    for (;;)
    {
        const unsigned m_trav = __ballot_sync(FULL, state == ST_TRAV);
        const unsigned m_need = __ballot_sync(FULL, state == ST_NEED);
        if ((m_trav | m_need) == 0u) break;

        if (m_trav == 0u || __popc(m_need) >= P.refill)
        {   // ---------------- regroup point: every parked lane shades + refills together ----------------
            const bool need = state == ST_NEED;
            bool fresh = false;
            if (need)
            {
                if (have_hit)
                {
                    bool alive;
                    if (h.id != RT_MISS_ID) alive = shade_hit(A, S, h, ps, rng) && (++bounce < P.max_bounces);
                    else { if (S.env_tex) shade_miss(S, P.consts, ps); alive = false; }
                    if (!alive) { sum = sum + ps.radiance; sample++; fresh = true; }
                    have_hit = false;
                }
                else fresh = true;
            }
            const bool want = need && fresh && (!have_pixel || sample >= P.spp);
            if (want && have_pixel)
            {
                rng.store(S.rng_state + (size_t)y * P.width + x);
                write_pixel_rg(P, x, y, sum);
                have_pixel = false;
            }
            const unsigned m_want = __ballot_sync(FULL, want);
            if (m_want)
            {   // warp-aggregated fetch of the next pixels (8x4-tile order keeps camera rays coherent)
                const int leader = __ffs(m_want) - 1;
                unsigned base = 0;
                if (lane == leader) base = atomicAdd(P.pixel_counter, (unsigned)__popc(m_want));
                base = __shfl_sync(FULL, base, leader);
                if (want)
                {
                    const unsigned idx = base + (unsigned)__popc(m_want & ((1u << lane) - 1u));
                    if (idx >= total) state = ST_DONE;
                    else
                    {
                        const unsigned tile = idx >> 5, in = idx & 31u;
                        x = (int)((tile % tiles_x) * 8u + (in & 7u)); y = (int)((tile / tiles_x) * 4u + (in >> 3));
                        if (x < P.width && y < P.height)
                        {
                            rng.load(S.rng_state + (size_t)y * P.width + x);  // GPUScene::GetRNGForPixel, GPUScene.h:95
                            sum = mk3(0, 0, 0); sample = 0; have_pixel = true;
                        }   // else: a slot of a partial edge tile - the lane fetches again at the next regroup point
                    }
                }
            }
            if (need && have_pixel && state != ST_DONE)
            {
                if (fresh)
                {   // main_raytracing.cu:190-191; jitter draws x first, then y
                    const float jx = rng.uniform(), jy = rng.uniform();
                    const float u = ((float)x + jx) / (float)P.width, v = ((float)y + jy) / (float)P.height;
                    camera_ray_rg(S.camera, u, v, ps.o, ps.d);
                    ps.throughput = mk3(1, 1, 1); ps.radiance = mk3(0, 0, 0);
                    bounce = 0;
                }
                nd = normalize3(ps.d);
                h.t = 1e30f; h.id = RT_MISS_ID; h.slot = 0; h.key = RT_MISS_ID; h.bu = 0; h.bv = 0;
                sphere_tests(S.spheres, S.sphere_count, ps.o, nd, h);
                start_length = h.t;
                const float ix = nd.x != 0.0f ? 1.0f / nd.x : big, iy = nd.y != 0.0f ? 1.0f / nd.y : big, iz = nd.z != 0.0f ? 1.0f / nd.z : big;
                i_xy = make_float2(ix, iy); i_zz = make_float2(iz, iz);
                no_xy = make_float2(-ps.o.x, -ps.o.y); no_zz = make_float2(-ps.o.z, -ps.o.z);
                sp = 0; stack[sp++] = RT_SENTINEL;
                node = A.root_ref;
                state = ST_TRAV; rays++;
            }
        }

        // ---------------- traversal: ONE step per lane and iteration (a wide node, or one triangle of a leaf) ----------------
        // Nothing speculative: the per-ray work is the megakernel's, node for node. Node steps run every iteration; the triangle
        // step is voted: lanes that sit on a leaf wait until leaf_batch of them do (or nobody has a node step left). Waiting is
        // free here - unlike in the megakernel, no other lane's critical path depends on it, the parked capacity is refilled.
        const bool trav = state == ST_TRAV;
        const bool on_leaf = trav && (node & RT_LEAF_BIT);
        const unsigned m_leaf = __ballot_sync(FULL, on_leaf), m_inner = __ballot_sync(FULL, trav && !on_leaf);
        if (trav && !on_leaf)
        {
            const WideNode* w = reinterpret_cast<const WideNode*>(A.arena + (node & RT_REF_BYTE_MASK));
            const float4 a = ldg4(w), b = ldg4(reinterpret_cast<const char*>(w) + 16), c = ldg4(reinterpret_cast<const char*>(w) + 32);
            const uint2 ch = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const char*>(w) + 48));
            // centre + half-extent node test, three-compare conservative cull (traverse.cuh)
            float mn0, mx0, mn1, mx1;
            center_slabs(a, b, c, no_xy, no_zz, i_xy, i_zz, make_float2(fabsf(i_xy.x), fabsf(i_xy.y)), make_float2(fabsf(i_zz.x), fabsf(i_zz.y)), mn0, mx0, mn1, mx1);
            const float lo0 = fmaf(-kWiden, fabsf(mn0), mn0), hi0 = fmaf(kWiden, fabsf(mx0), mx0);
            const float lo1 = fmaf(-kWiden, fabsf(mn1), mn1), hi1 = fmaf(kWiden, fabsf(mx1), mx1);
            const bool hit0 = hi0 >= lo0 && lo0 <= h.t && hi0 > 0.0f;
            const bool hit1 = hi1 >= lo1 && lo1 <= h.t && hi1 > 0.0f;
            if (hit0 && hit1)
            {
                const bool swap = mn1 < mn0;
                stack[sp++] = swap ? ch.x : ch.y;
                node = swap ? ch.y : ch.x;
            }
            else if (hit0) node = ch.x;
            else if (hit1) node = ch.y;
            else node = stack[--sp];
        }
        if (on_leaf && (__popc(m_leaf) >= P.leaf_batch || m_inner == 0u))
        {   // one triangle; the leaf reference is re-encoded with the remaining range. No per-candidate leaf-box check (deferred).
            leaf_tests<false, false>(A, ref_slot(A, node), 1u, ps.o, ps.d, nd, start_length, h);
            if (ref_has_more(node)) node = ref_next_tri(node);
            else node = stack[--sp];
        }
        if (trav && node == RT_SENTINEL)
        {   // ray finished: the reference only counts a triangle whose leaf box passes (main_raytracing.cu:46) - checked once, on the
            // winner; if it fails (grazing within rounding, rare) the ray is redone with the per-candidate check
            if (h.id != RT_MISS_ID && !(h.id & RT_SPHERE_ID_BIT) && !exact_box(ps.o, ps.d, A.ref_nodes + (__ldg(&A.tris[h.slot].ref_leaf) & RT_REF_LEAF_MASK), start_length))
            {
                h.t = 1e30f; h.id = RT_MISS_ID; h.slot = 0; h.key = RT_MISS_ID; h.bu = 0; h.bv = 0;
                sphere_tests(S.spheres, S.sphere_count, ps.o, nd, h);
                trace_wide_checked<false>(A, ps.o, ps.d, nd, h);
            }
            state = ST_NEED; have_hit = true;
        }
    }

    if (P.ray_counter)
    {
        for (int o2 = 16; o2; o2 >>= 1) rays += __shfl_xor_sync(FULL, rays, o2);
        if (lane == 0 && rays) atomicAdd(P.ray_counter, (unsigned long long)rays);
    }
}


cudaError_t LaunchPathTraceRegroup(const AccelView& accel, const SceneView& scene, const RtProcessParams& p, unsigned int* d_pixel_counter,
                                   cudaStream_t stream)
{
    RegroupParams P;
    P.surface = (unsigned char*)p.surface; P.surface_last = (unsigned char*)p.surface_last_frame;
    P.width = p.width; P.height = p.height; P.pitch = p.pitch; P.frame_index = p.frame_index;
    P.spp = p.samples_per_launch; P.max_bounces = p.max_bounces; P.blend_mode = p.blend_mode; P.write_last = p.write_last_frame;
    P.ray_counter = p.d_ray_counter; P.pixel_counter = d_pixel_counter;
    P.refill = kRefill; P.leaf_batch = kLeafBatch;
    if (const char* v = getenv("B200RT_REFILL")) P.refill = atoi(v);        // tuning hooks (profiles/ sweeps)
    if (const char* v = getenv("B200RT_LEAF_BATCH")) P.leaf_batch = atoi(v);
    {   // glm quat(vec3(0, PI, 0)), type_quat.inl:204-213, in host fp32 like the oracle
        const float PI = 3.1415926536f;
        const float cx = std::cos(0.0f * 0.5f), cy = std::cos(PI * 0.5f), cz = std::cos(0.0f * 0.5f);
        const float sx = std::sin(0.0f * 0.5f), sy = std::sin(PI * 0.5f), sz = std::sin(0.0f * 0.5f);
        P.consts.qw = cx * cy * cz + sx * sy * sz;
        P.consts.qx = sx * cy * cz - cx * sy * sz;
        P.consts.qy = cx * sy * cz + sx * cy * sz;
        P.consts.qz = cx * cy * sz - sx * sy * cz;
    }
    cudaError_t e = cudaMemsetAsync(d_pixel_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    static int blocks_per_sm = 0, sm_count = 0, min_blocks = 0;  // persistent grid: one resident wave
    if (!blocks_per_sm)
    {
        int dev = 0;
        if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
        if ((e = cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
        const char* v = getenv("B200RT_RG_MINBLOCKS");  // occupancy sweep hook (profiles/)
        min_blocks = v ? atoi(v) : 3;
        if (min_blocks == 4) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, pathtrace_regroup_kernel<4>, 256, 0);
        else if (min_blocks == 5) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, pathtrace_regroup_kernel<5>, 256, 0);
        else e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, pathtrace_regroup_kernel<3>, 256, 0);
        if (e != cudaSuccess) return e;
        if (blocks_per_sm < 1) blocks_per_sm = 1;
    }
    const unsigned tiles = (unsigned)((p.width + 7) / 8) * (unsigned)((p.height + 3) / 4);
    unsigned grid = (unsigned)(sm_count * blocks_per_sm);
    const unsigned needed = (tiles + 7) / 8;  // 8 warps per block
    if (grid > needed) grid = needed;
    if (grid < 1) grid = 1;
    if (min_blocks == 4) pathtrace_regroup_kernel<4><<<grid, 256, 0, stream>>>(accel, scene, P);
    else if (min_blocks == 5) pathtrace_regroup_kernel<5><<<grid, 256, 0, stream>>>(accel, scene, P);
    else pathtrace_regroup_kernel<3><<<grid, 256, 0, stream>>>(accel, scene, P);
    return cudaGetLastError();
}
