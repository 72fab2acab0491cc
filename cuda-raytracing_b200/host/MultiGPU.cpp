// MultiGPUTracer: spp split over the devices of one node, one ncclReduce per frame.
// NCCL is bound at run time (dlopen of libnccl.so.2 on first use with more than one device): libb200rt.so has no link-time
// dependency on it, a single-GPU host never needs it, and a process that already holds a libnccl (torch's) shares that one.
#include "MultiGPU.h"
#include <chrono>
#include <cstring>
#include <dlfcn.h>
#include <functional>
#include <mutex>
#include <stdexcept>
#include <thread>
#include <cuda_runtime.h>
#include <nccl.h>
#include "b200rt.h"

namespace RayTracing
{
    namespace
    {
        void Check(cudaError_t e, const char* what)
        {
            if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
        }

        struct Nccl
        {
            void* handle = nullptr;
            ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
            ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
            ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
            const char* (*GetErrorString)(ncclResult_t) = nullptr;
        };
        Nccl& LoadNccl()
        {
            static Nccl n;
            static std::once_flag once;
            std::call_once(once, [] {
                n.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
                if (!n.handle) return;
                n.CommInitAll = reinterpret_cast<decltype(n.CommInitAll)>(dlsym(n.handle, "ncclCommInitAll"));
                n.CommDestroy = reinterpret_cast<decltype(n.CommDestroy)>(dlsym(n.handle, "ncclCommDestroy"));
                n.Reduce = reinterpret_cast<decltype(n.Reduce)>(dlsym(n.handle, "ncclReduce"));
                n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(dlsym(n.handle, "ncclGetErrorString"));
            });
            if (!n.handle || !n.CommInitAll || !n.CommDestroy || !n.Reduce || !n.GetErrorString)
                throw std::runtime_error("MultiGPUTracer: libnccl.so.2 not loadable (needed for more than one device)");
            return n;
        }
        void CheckNccl(ncclResult_t r, const char* what)
        {
            if (r != ncclSuccess) throw std::runtime_error(std::string(what) + ": " + LoadNccl().GetErrorString(r));
        }
        // the calling thread's current device is the caller's business: whatever this class does on it is undone
        struct DeviceGuard
        {
            int saved = -1;
            DeviceGuard() { if (cudaGetDevice(&saved) != cudaSuccess) saved = -1; }
            ~DeviceGuard() { if (saved >= 0) cudaSetDevice(saved); }
        };
        uint64_t PaddedStates(int width, int height) { return ((uint64_t)width * height + 127) / 128 * 128; }  // RayTracing.cpp:215-218
    }

    struct MultiGPUTracer::Replica
    {
        int rank = 0, device = 0;
        cudaStream_t stream = nullptr;
        cudaEvent_t ev[3] = {};
        std::unique_ptr<Scene> scene;
        RtRngState* rng_state = nullptr;
        void* sum = nullptr;            // float4 W x H, pitch W * 16: rgb = summed radiance, alpha = samples
        void* workspace = nullptr; uint64_t workspace_bytes = 0;
        unsigned long long* d_rays = nullptr;
        ncclComm_t comm = nullptr;
        float render_ms = 0, reduce_ms = 0;
    };

    void MultiGPUTracer::ForEachDevice(const std::function<void(Replica&)>& fn)
    {
        std::vector<std::thread> threads;
        std::vector<std::string> errors(replicas.size());
        for (size_t i = 0; i < replicas.size(); i++)
            threads.emplace_back([&, i] {
                try
                {
                    Check(cudaSetDevice(replicas[i]->device), "cudaSetDevice");
                    fn(*replicas[i]);
                }
                catch (const std::exception& e) { errors[i] = e.what(); if (errors[i].empty()) errors[i] = "error"; }
            });
        for (auto& t : threads) t.join();
        for (size_t i = 0; i < errors.size(); i++)
            if (!errors[i].empty()) throw std::runtime_error("device " + std::to_string(replicas[i]->device) + ": " + errors[i]);
    }

    MultiGPUTracer::MultiGPUTracer(const Options& options_) : options(options_)
    {
        if (options.width <= 0 || options.height <= 0) throw std::runtime_error("MultiGPUTracer: bad frame size");
        int visible = 0;
        Check(cudaGetDeviceCount(&visible), "cudaGetDeviceCount");
        DeviceGuard guard;  // ncclCommInitAll may leave another device current
        std::vector<int> devices = options.devices;
        if (devices.empty()) for (int i = 0; i < visible; i++) devices.push_back(i);
        if (devices.empty()) throw std::runtime_error("MultiGPUTracer: no CUDA device (there is no CPU fallback)");
        for (int d : devices) if (d < 0 || d >= visible) throw std::runtime_error("MultiGPUTracer: device ordinal out of range");
        seed = options.tracer.seed;
        if (seed == 0)
        {   // RayTracing.cpp:219-220, drawn once so that all devices share it
            const uint32_t ms = (uint32_t)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::system_clock::now().time_since_epoch()).count();
            seed = 0xDEADBEEFu * ms;
            if (seed == 0) seed = 1;
        }
        for (size_t i = 0; i < devices.size(); i++)
        {
            replicas.push_back(std::make_unique<Replica>());
            replicas.back()->rank = (int)i; replicas.back()->device = devices[i];
        }
        if (devices.size() > 1)
        {
            std::vector<ncclComm_t> comms(devices.size());
            CheckNccl(LoadNccl().CommInitAll(comms.data(), (int)devices.size(), devices.data()), "ncclCommInitAll");
            for (size_t i = 0; i < devices.size(); i++) replicas[i]->comm = comms[i];
        }
        const uint64_t states = PaddedStates(options.width, options.height);
        const size_t frame_bytes = (size_t)options.width * options.height * 16;
        ForEachDevice([&](Replica& r) {
            Check(cudaStreamCreateWithFlags(&r.stream, cudaStreamNonBlocking), "cudaStreamCreate");
            for (auto& e : r.ev) Check(cudaEventCreate(&e), "cudaEventCreate");
            r.scene = std::make_unique<Scene>(options.tracer.env_dds);                                   // RayTracing.cpp:22-25, per device
            if (options.tracer.cornell_box) CUDARayTracer::SetupCornellBox(*r.scene);
            if (options.tracer.stanford_bunny) CUDARayTracer::SetupStanfordBunny(*r.scene, options.tracer.bunny_obj);
            Check(cudaMalloc(&r.rng_state, states * sizeof(RtRngState)), "allocation failed");
            Check(cudaMalloc(&r.sum, frame_bytes), "allocation failed");
            Check(cudaMalloc(&r.d_rays, sizeof(unsigned long long)), "allocation failed");
            Check(cudaMemset(r.d_rays, 0, sizeof(unsigned long long)), "memset");
            r.workspace_bytes = rt_process_workspace_bytes(options.width, options.height, 0);
            if (r.workspace_bytes)
            {
                Check(cudaMalloc(&r.workspace, r.workspace_bytes), "allocation failed");
                Check(cudaMemset(r.workspace, 0, r.workspace_bytes), "memset");
            }
            if (rt_init_rng(r.rng_state, states, seed, (uint64_t)r.rank * states, r.stream) != RT_OK) throw std::runtime_error(rt_last_error());
            r.scene->GetCamera().SetViewportSize(vec2((float)options.width, (float)options.height));
            r.scene->Update(0.0f);
            r.scene->Upload(r.rng_state);                                                                  // :231 (cold: arrays + re-layout)
            Check(cudaStreamSynchronize(r.stream), "rng init");
        });
    }

    MultiGPUTracer::~MultiGPUTracer()
    {
        DeviceGuard guard;
        for (auto& rp : replicas)
        {
            Replica& r = *rp;
            cudaSetDevice(r.device);
            if (r.stream) cudaStreamSynchronize(r.stream);
            if (r.comm) LoadNccl().CommDestroy(r.comm);
            r.scene.reset();
            cudaFree(r.rng_state); cudaFree(r.sum); cudaFree(r.d_rays); cudaFree(r.workspace);
            for (auto& e : r.ev) if (e) cudaEventDestroy(e);
            if (r.stream) cudaStreamDestroy(r.stream);
        }
    }

    void MultiGPUTracer::SetCamera(const float* p, float angle_x, float angle_y)
    {
        for (auto& r : replicas)
        {
            Camera& c = r->scene->GetCamera();
            if (p) c.SetPosition(vec3(p[0], p[1], p[2]));
            c.SetXAndle(angle_x); c.SetYAndle(angle_y);
            r->scene->AddDirtyFlag(DirtyFlagValue::Samples);
        }
    }

    void MultiGPUTracer::Reseed(uint32_t seed_)
    {
        seed = seed_;
        const uint64_t states = PaddedStates(options.width, options.height);
        ForEachDevice([&](Replica& r) {
            if (rt_init_rng(r.rng_state, states, seed, (uint64_t)r.rank * states, r.stream) != RT_OK) throw std::runtime_error(rt_last_error());
            Check(cudaStreamSynchronize(r.stream), "rng init");
        });
    }

    void MultiGPUTracer::Render(int total_spp, int spp_per_launch, void* host_pixels, size_t host_pitch)
    {
        const int W = options.width, H = options.height, n = (int)replicas.size();
        if (total_spp <= 0 || spp_per_launch <= 0 || !host_pixels || host_pitch < (size_t)W * 16) throw std::runtime_error("MultiGPUTracer::Render: bad argument");
        const auto t0 = std::chrono::steady_clock::now();
        ForEachDevice([&](Replica& r) {
            const int mine = total_spp / n + (r.rank < total_spp % n ? 1 : 0);
            Scene& scene = *r.scene;
            scene.GetCamera().SetViewportSize(vec2((float)W, (float)H));                                   // RayTracing.cpp:223
            scene.Update(0.0f);                                                                           // :224
            scene.Upload(r.rng_state);                                                                    // :231
            Check(cudaMemsetAsync(r.sum, 0, (size_t)W * H * 16, r.stream), "clear sum buffer");
            Check(cudaEventRecord(r.ev[0], r.stream), "event");
            for (int done = 0, k = 0; done < mine; k++)
            {
                RtProcessParams p{};
                p.surface = r.sum; p.surface_last_frame = nullptr; p.width = W; p.height = H; p.pitch = (uint64_t)W * 16;
                p.frame_index = k; p.samples_per_launch = std::min(spp_per_launch, mine - done); p.max_bounces = options.tracer.max_bounces;
                p.blend_mode = 1; p.variant = 0; p.write_last_frame = 0;
                p.scene = &scene; p.accel = scene.GetAccel(); p.d_ray_counter = r.d_rays; p.stream = r.stream;
                p.workspace = r.workspace; p.workspace_bytes = r.workspace_bytes;
                if (rt_process_ex(&p) != RT_OK) throw std::runtime_error(rt_last_error());                // :232
                done += p.samples_per_launch;
            }
            Check(cudaEventRecord(r.ev[1], r.stream), "event");
            if (n > 1)
                CheckNccl(LoadNccl().Reduce(r.sum, r.sum, (size_t)W * H * 4, ncclFloat, ncclSum, 0, r.comm, r.stream), "ncclReduce");
            Check(cudaEventRecord(r.ev[2], r.stream), "event");
            if (r.rank == 0)
            {
                if (rt_resolve_sum(r.sum, r.sum, W, H, (uint64_t)W * 16, (uint64_t)W * 16, r.stream) != RT_OK) throw std::runtime_error(rt_last_error());
                Check(cudaMemcpy2DAsync(host_pixels, host_pitch, r.sum, (size_t)W * 16, (size_t)W * 16, (size_t)H, cudaMemcpyDeviceToHost, r.stream), "frame to host");
            }
            Check(cudaStreamSynchronize(r.stream), "render");
            Check(cudaEventElapsedTime(&r.render_ms, r.ev[0], r.ev[1]), "event");
            Check(cudaEventElapsedTime(&r.reduce_ms, r.ev[1], r.ev[2]), "event");
        });
        timing = Timing();
        timing.reduce_ms = 1e30;  // the device that arrives last at the reduce does not wait in it: its figure is the reduce itself
        for (auto& r : replicas) { timing.render_ms = std::max<double>(timing.render_ms, r->render_ms); timing.reduce_ms = std::min<double>(timing.reduce_ms, r->reduce_ms); }
        timing.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }

    unsigned long long MultiGPUTracer::RayCount() const
    {
        unsigned long long total = 0;
        DeviceGuard guard;
        for (auto& r : replicas)
        {
            unsigned long long v = 0;
            cudaSetDevice(r->device);
            cudaMemcpy(&v, r->d_rays, sizeof(v), cudaMemcpyDeviceToHost);
            total += v;
        }
        return total;
    }
}
