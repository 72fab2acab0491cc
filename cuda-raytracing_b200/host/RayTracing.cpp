// CUDARayTracer: per-frame protocol of RayTracing/RayTracing.cpp:205-235 over the C ABI.
#include "RayTracing.h"
#include <chrono>
#include <cuda_runtime.h>
#include "b200rt.h"

namespace RayTracing
{
    CUDARayTracer::CUDARayTracer(SurfaceData surface_, Options options_) : options(options_), surface(surface_)
    {
        scene = std::make_unique<Scene>(options.env_dds);      // RayTracing.cpp:22
        if (options.cornell_box) SetupCornellBox(*scene);       // :24
        if (options.stanford_bunny) SetupStanfordBunny(*scene, options.bunny_obj);  // :25
    }

    CUDARayTracer::~CUDARayTracer()
    {
        if (rng_state) cudaFree(rng_state);
        if (d_ray_counter) cudaFree(d_ray_counter);
    }

    void CUDARayTracer::OnResize(SurfaceData surface_)  // RayTracing.cpp:72-77
    {
        surface = surface_;
        if (rng_state) { cudaFree(rng_state); rng_state = nullptr; }
        scene->AddDirtyFlag(DirtyFlagValue::Samples);
    }

    void CUDARayTracer::Process()
    {
        if (!rng_state)
        {   // RayTracing.cpp:213-221
            const auto pixel_count = surface.width * surface.height;
            const uint32_t thread_block_size = 128;
            const uint32_t thread_block_count = (pixel_count + thread_block_size - 1) / thread_block_size;
            if (cudaMalloc(&rng_state, sizeof(RtRngState) * thread_block_count * thread_block_size) != cudaSuccess)
                throw std::runtime_error("allocation failed");
            uint32_t seed = options.seed;
            if (seed == 0)
            {
                const uint32_t ms = (uint32_t)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::system_clock::now().time_since_epoch()).count();
                seed = 0xDEADBEEFu * ms;
            }
            init_rng(thread_block_count, thread_block_size, rng_state, seed);
        }
        if (!d_ray_counter)
        {
            if (cudaMalloc(&d_ray_counter, sizeof(unsigned long long)) != cudaSuccess) throw std::runtime_error("allocation failed");
            cudaMemset(d_ray_counter, 0, sizeof(unsigned long long));
        }

        scene->GetCamera().SetViewportSize(vec2((float)surface.width, (float)surface.height));  // :223
        scene->Update(0.0f);                                                                   // :224
        if (scene->IsFlagDirty(DirtyFlagValue::Samples)) frame_index = 0;                       // :226-229
        scene->Upload(rng_state);                                                               // :231

        // :232-233 - the reference launches raytracing_process() and then copies the whole surface into
        // the last-frame buffer with a blocking D2D cudaMemcpy; here the kernel stores both (write_last_frame).
        RtProcessParams p{};
        p.surface = surface.surface; p.surface_last_frame = surface.last_frame_surface;
        p.width = surface.width; p.height = surface.height; p.pitch = surface.pitch;
        p.frame_index = frame_index; p.samples_per_launch = options.samples_per_launch; p.max_bounces = options.max_bounces;
        p.blend_mode = 0; p.variant = options.kernel_variant; p.write_last_frame = 1;
        p.scene = scene.get(); p.accel = scene->GetAccel(); p.d_ray_counter = d_ray_counter; p.stream = nullptr;
        if (rt_process_ex(&p) != RT_OK) throw std::runtime_error(rt_last_error());
        frame_index++;                                                                          // :234
    }

    size_t CUDARayTracer::RenderToHost(int frames, void* host_pixels, size_t host_pitch)
    {
        if (frames <= 0 || !host_pixels || host_pitch < (size_t)surface.width * 16) throw std::runtime_error("RenderToHost: bad argument");
        size_t h2d = scene->RefreshDynamic();
        scene->AddDirtyFlag(DirtyFlagValue::Samples);
        for (int f = 0; f < frames; f++) { Process(); h2d += sizeof(RtGPUScene); }
        const cudaError_t e = cudaMemcpy2D(host_pixels, host_pitch, surface.surface, surface.pitch, (size_t)surface.width * 16,
                                           (size_t)surface.height, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) throw std::runtime_error(std::string("RenderToHost: ") + cudaGetErrorString(e));
        return h2d;
    }

    unsigned long long CUDARayTracer::GetRayCount() const
    {
        unsigned long long n = 0;
        if (d_ray_counter) cudaMemcpy(&n, d_ray_counter, sizeof(n), cudaMemcpyDeviceToHost);
        return n;
    }
}
