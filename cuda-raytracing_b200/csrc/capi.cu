// C ABI of include/b200rt.h: argument checking, object lifetime, launch glue. No compute on the host.
#include <cstdio>
#include <cstring>
#include <map>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>
#include <vector>
#include "accel.h"
#include "b200rt.h"
#include "dds.h"
#include "kernels.h"
#include "rng_host.h"
#include "capi_internal.h"

using namespace rt_internal;

namespace
{
    std::map<int, uint32_t*> g_jump;                  // device -> XORWOW jump matrices
    constexpr int kJumpLevels = 40;                   // subsequence numbers up to 2^40

    int JumpMatrices(const uint32_t** out)
    {
        int dev = 0;
        RT_CUDA(cudaGetDevice(&dev));
        std::lock_guard<std::mutex> lock(g_mutex);
        auto it = g_jump.find(dev);
        if (it == g_jump.end())
        {
            std::vector<uint32_t> host((size_t)kJumpLevels * 800);
            XorwowSequenceJumpMatrices(host.data(), kJumpLevels);
            uint32_t* d = nullptr;
            RT_CUDA(cudaMalloc(&d, host.size() * 4));
            RT_CUDA(cudaMemcpy(d, host.data(), host.size() * 4, cudaMemcpyHostToDevice));
            it = g_jump.emplace(dev, d).first;
        }
        *out = it->second;
        return RT_OK;
    }

}

extern "C" {

const char* rt_last_error(void) { return g_error.c_str(); }

int rt_device_count(void)
{
    int n = 0;
    return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}

// ---- RNG ------------------------------------------------------------------------------------
int rt_init_rng(RtRngState* d_states, uint64_t count, uint64_t seed, uint64_t first_subsequence, void* stream)
{
    if (!d_states && count) return Fail(RT_E_INVALID, "rt_init_rng: null states");
    if (first_subsequence + count >= (1ull << kJumpLevels)) return Fail(RT_E_INVALID, "rt_init_rng: subsequence out of range");
    const uint32_t* jump = nullptr;
    if (int r = JumpMatrices(&jump)) return r;
    RT_CUDA(LaunchRngInit(d_states, count, seed, first_subsequence, jump, kJumpLevels, (cudaStream_t)stream));
    return RT_OK;
}

void init_rng(uint32_t thread_block_count, uint32_t thread_block_size, RtRngState* rngStates, const unsigned int seed)
{
    if (int r = rt_init_rng(rngStates, (uint64_t)thread_block_count * thread_block_size, seed, 0, nullptr))
        printf("(initRNG) failed to launch error = %d (%s)\n", r, rt_last_error());
}

int rt_rng_draw(RtRngState* d_states, uint64_t count, uint32_t n, uint32_t* d_out, void* stream)
{
    if ((!d_states || !d_out) && count) return Fail(RT_E_INVALID, "rt_rng_draw: null pointer");
    RT_CUDA(LaunchRngDraw(d_states, count, n, d_out, (cudaStream_t)stream));
    return RT_OK;
}

// ---- accel ----------------------------------------------------------------------------------
int rt_accel_create(const RtBVHNode* nodes, uint32_t node_count, const uint32_t* face_indices,
                    const RtFace* faces, uint32_t face_count, const RtVertex* vertices, uint32_t vertex_count,
                    uint32_t fat_leaf_threshold, RtAccel** out)
try
{
    if (!out) return Fail(RT_E_INVALID, "rt_accel_create: null out");
    *out = nullptr;
    AccelHost host;
    if (int r = BuildAccelHost(nodes, node_count, face_indices, faces, face_count, vertices, vertex_count, fat_leaf_threshold, host))
        return Fail(r, r == RT_E_DEPTH ? "rt_accel_create: BVH deeper than the traversal stack" : "rt_accel_create: inconsistent BVH/scene arrays");
    const bool timing = getenv("B200RT_BVH_TIMING") != nullptr;
    const auto t_upload = std::chrono::steady_clock::now();
    RtAccel* a = new RtAccel();
    WideNode* d_wide = nullptr; TriRecord* d_tris = nullptr; ShadeRecord* d_shade = nullptr; RtBVHNode* d_ref = nullptr;
    const size_t bw = host.wide.size() * sizeof(WideNode), bt = host.tris.size() * sizeof(TriRecord);
    const size_t bs = host.shade.size() * sizeof(ShadeRecord), br = (size_t)node_count * sizeof(RtBVHNode);
    // one arena: wide nodes, then triangle records (references are 16-byte unit offsets into it, accel.h)
    cudaError_t e = cudaMalloc(&d_wide, bw + bt + 64);  // + pad: the unified fetch of trace_wide reads 56 B from the last record
    if (e == cudaSuccess) d_tris = reinterpret_cast<TriRecord*>(reinterpret_cast<char*>(d_wide) + bw);
    if (e == cudaSuccess) e = cudaMalloc(&d_shade, bs);
    if (e == cudaSuccess) e = cudaMalloc(&d_ref, br);
    if (e == cudaSuccess) e = cudaMemcpy(d_wide, host.wide.data(), bw, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_tris, host.tris.data(), bt, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_shade, host.shade.data(), bs, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_ref, nodes, br, cudaMemcpyHostToDevice);
    if (e != cudaSuccess)
    {
        cudaFree(d_wide); cudaFree(d_shade); cudaFree(d_ref); delete a;
        return CudaFail(e, "rt_accel_create upload");
    }
    if (timing) fprintf(stderr, "[accel] %-20s %8.1f ms (%.0f MB)\n", "upload", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_upload).count(), (bw + bt + bs + br) / 1e6);
    a->view.arena = reinterpret_cast<const char*>(d_wide); a->view.wide = d_wide; a->view.tris = d_tris; a->view.shade = d_shade; a->view.ref_nodes = d_ref;
    a->view.root_ref = host.root_ref; a->view.node_count = node_count; a->view.tri_byte0 = host.tri_byte0;
    a->device_bytes = bw + bt + bs + br;
    a->wide_count = (uint32_t)host.wide.size(); a->slot_count = (uint32_t)host.tris.size();
    a->ref_leaf_count = host.ref_leaf_count; a->fat_leaf_count = host.fat_leaf_count;
    a->wide_depth = host.wide_depth; a->ref_depth = host.ref_depth;
    *out = a;
    return RT_OK;
}
catch (const std::exception& e) { return rt_internal::Fail(RT_E_INVALID, std::string("rt_accel_create: ") + e.what()); }
catch (...) { return rt_internal::Fail(RT_E_INVALID, "rt_accel_create: unknown C++ exception"); }

int rt_accel_check_host(const RtBVHNode* nodes, uint32_t node_count, const uint32_t* face_indices,
                        const RtFace* faces, uint32_t face_count, const RtVertex* vertices, uint32_t vertex_count,
                        uint32_t fat_leaf_threshold, uint64_t* info4)
try
{
    AccelHost host;
    if (int r = BuildAccelHost(nodes, node_count, face_indices, faces, face_count, vertices, vertex_count, fat_leaf_threshold, host))
        return Fail(r, r == RT_E_DEPTH ? "rt_accel_check_host: BVH deeper than the traversal stack" : "rt_accel_check_host: inconsistent BVH/scene arrays");
    const char* why = "";
    if (int r = CheckAccelHost(host, nodes, node_count, info4, &why)) return Fail(r, std::string("rt_accel_check_host: ") + why);
    return RT_OK;
}
catch (const std::exception& e) { return rt_internal::Fail(RT_E_INVALID, std::string("rt_accel_check_host: ") + e.what()); }
catch (...) { return rt_internal::Fail(RT_E_INVALID, "rt_accel_check_host: unknown C++ exception"); }

int rt_accel_destroy(RtAccel* a)
{
    if (!a) return RT_OK;
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        for (auto it = g_bound.begin(); it != g_bound.end();) it = it->second.accel == a ? g_bound.erase(it) : std::next(it);
    }
    cudaFree((void*)a->view.arena); cudaFree((void*)a->view.shade); cudaFree((void*)a->view.ref_nodes);
    delete a;
    return RT_OK;
}

int rt_accel_info(const RtAccel* a, uint64_t* info8)
{
    if (!a || !info8) return Fail(RT_E_INVALID, "rt_accel_info: null");
    info8[0] = a->wide_count; info8[1] = a->slot_count; info8[2] = a->ref_leaf_count; info8[3] = a->fat_leaf_count;
    info8[4] = a->wide_depth; info8[5] = a->device_bytes & 0xFFFFFFFFull; info8[6] = a->device_bytes >> 32; info8[7] = a->ref_depth;
    return RT_OK;
}

int rt_accel_bind(const void* d_nodes_key, RtAccel* a)
{
    if (!d_nodes_key || !a) return Fail(RT_E_INVALID, "rt_accel_bind: null");
    RtAccel* replaced = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        auto it = g_bound.find(d_nodes_key);
        if (it != g_bound.end() && it->second.owned && it->second.accel != a) replaced = it->second.accel;
        g_bound[d_nodes_key] = BoundAccel{ a, DeviceBufferId(d_nodes_key), false };
    }
    if (replaced) rt_accel_destroy(replaced);  // a library-built re-layout gives way to the caller's
    return RT_OK;
}

int rt_accel_unbind(const void* d_nodes_key)
{
    RtAccel* owned = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        auto it = g_bound.find(d_nodes_key);
        if (it != g_bound.end()) { if (it->second.owned) owned = it->second.accel; g_bound.erase(it); }
    }
    if (owned) rt_accel_destroy(owned);
    return RT_OK;
}

}  // extern "C"
