// CUDARayTracer: per-frame protocol of RayTracing/RayTracing.cpp:205-235 over the C ABI.
#include "RayTracing.h"
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <vector>
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
        if (tile_workspace) cudaFree(tile_workspace);
    }

    void CUDARayTracer::OnResize(SurfaceData surface_)  // RayTracing.cpp:72-77
    {
        surface = surface_;
        if (rng_state) { cudaFree(rng_state); rng_state = nullptr; }
        if (tile_workspace) { cudaFree(tile_workspace); tile_workspace = nullptr; tile_workspace_bytes = 0; }
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

        if (!tile_workspace)
        {   // per-tile cost of the previous frame -> heaviest-first tile order of the next one (csrc/pathtrace.cu); zero = no history
            tile_workspace_bytes = rt_process_workspace_bytes((int)surface.width, (int)surface.height, 0);
            if (tile_workspace_bytes)
            {
                if (cudaMalloc(&tile_workspace, tile_workspace_bytes) != cudaSuccess) throw std::runtime_error("allocation failed");
                cudaMemset(tile_workspace, 0, tile_workspace_bytes);
            }
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
        p.workspace = tile_workspace; p.workspace_bytes = tile_workspace_bytes;
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

    namespace
    {
        struct CheckpointHeader
        {
            char magic[8];  // "B200RTCK"
            uint32_t version, width, height, frame_index, samples_per_launch, max_bounces, rng_state_count, rng_state_bytes;
        };
        struct File
        {
            FILE* f;
            File(const std::string& path, const char* mode) : f(std::fopen(path.c_str(), mode)) { if (!f) throw std::runtime_error("checkpoint: cannot open " + path); }
            ~File() { if (f) std::fclose(f); }
        };
        uint32_t PaddedStateCount(int width, int height) { return (uint32_t)(((size_t)width * height + 127) / 128 * 128); }
    }

    void CUDARayTracer::SaveCheckpoint(const std::string& path)
    {
        if (!rng_state) throw std::runtime_error("checkpoint: nothing rendered yet");
        const size_t row = (size_t)surface.width * 16, rows = (size_t)surface.height;
        const uint32_t count = PaddedStateCount(surface.width, surface.height);
        std::vector<unsigned char> host(std::max(row * rows, (size_t)count * sizeof(RtRngState)));
        CheckpointHeader h{ { 'B', '2', '0', '0', 'R', 'T', 'C', 'K' }, 1u, (uint32_t)surface.width, (uint32_t)surface.height, (uint32_t)frame_index,
                            (uint32_t)options.samples_per_launch, (uint32_t)options.max_bounces, count, (uint32_t)sizeof(RtRngState) };
        File out(path, "wb");
        bool ok = std::fwrite(&h, sizeof(h), 1, out.f) == 1;
        for (void* s : { surface.surface, surface.last_frame_surface })
        {   // cudaMemcpy2D on the legacy stream: ordered after every Process() launched so far
            if (cudaMemcpy2D(host.data(), row, s, surface.pitch, row, rows, cudaMemcpyDeviceToHost) != cudaSuccess) throw std::runtime_error("checkpoint: surface read failed");
            ok = ok && std::fwrite(host.data(), row, rows, out.f) == rows;
        }
        if (cudaMemcpy(host.data(), rng_state, (size_t)count * sizeof(RtRngState), cudaMemcpyDeviceToHost) != cudaSuccess) throw std::runtime_error("checkpoint: RNG state read failed");
        ok = ok && std::fwrite(host.data(), sizeof(RtRngState), count, out.f) == count;
        if (!ok) throw std::runtime_error("checkpoint: short write to " + path);
    }

    void CUDARayTracer::LoadCheckpoint(const std::string& path)
    {
        File in(path, "rb");
        CheckpointHeader h{};
        if (std::fread(&h, sizeof(h), 1, in.f) != 1 || std::memcmp(h.magic, "B200RTCK", 8) != 0 || h.version != 1u) throw std::runtime_error("checkpoint: not a checkpoint file: " + path);
        if ((int)h.width != surface.width || (int)h.height != surface.height) throw std::runtime_error("checkpoint: surface size differs");
        if (h.rng_state_bytes != sizeof(RtRngState) || h.rng_state_count != PaddedStateCount(surface.width, surface.height)) throw std::runtime_error("checkpoint: RNG state layout differs");
        const size_t row = (size_t)surface.width * 16, rows = (size_t)surface.height;
        std::vector<unsigned char> host(std::max(row * rows, (size_t)h.rng_state_count * sizeof(RtRngState)));
        for (void* s : { surface.surface, surface.last_frame_surface })
        {
            if (std::fread(host.data(), row, rows, in.f) != rows) throw std::runtime_error("checkpoint: truncated file");
            if (cudaMemcpy2D(s, surface.pitch, host.data(), row, row, rows, cudaMemcpyHostToDevice) != cudaSuccess) throw std::runtime_error("checkpoint: surface write failed");
        }
        if (std::fread(host.data(), sizeof(RtRngState), h.rng_state_count, in.f) != h.rng_state_count) throw std::runtime_error("checkpoint: truncated file");
        if (!rng_state && cudaMalloc(&rng_state, (size_t)h.rng_state_count * sizeof(RtRngState)) != cudaSuccess) throw std::runtime_error("allocation failed");
        if (cudaMemcpy(rng_state, host.data(), (size_t)h.rng_state_count * sizeof(RtRngState), cudaMemcpyHostToDevice) != cudaSuccess) throw std::runtime_error("checkpoint: RNG state write failed");
        options.samples_per_launch = (int)h.samples_per_launch; options.max_bounces = (int)h.max_bounces;
        // bring the scene to its clean uploaded state first, so the Samples flag of a freshly built scene does not restart the run
        scene->GetCamera().SetViewportSize(vec2((float)surface.width, (float)surface.height));
        scene->Update(0.0f);
        scene->Upload(rng_state);
        frame_index = (int)h.frame_index;
    }

    unsigned long long CUDARayTracer::GetRayCount() const
    {
        unsigned long long n = 0;
        if (d_ray_counter) cudaMemcpy(&n, d_ray_counter, sizeof(n), cudaMemcpyDeviceToHost);
        return n;
    }
}
