// Renderer facade: the reference's CUDARayTracer (RayTracing/RayTracing.h:12-45) kept by name.
// Process() = one progressive frame through the C ABI of include/b200rt.h. The ImGui/QPC lines of
// the reference's Process() are out of scope (UI); asset paths and the sampling constants the
// reference hard-codes (5 spp, 6 bounces, wall-clock seed) are runtime options defaulting to them.
#pragma once
#include <memory>
#include <string>
#include "Scene.h"

namespace RayTracing
{
    class CUDARayTracer
    {
    public:
        struct SurfaceData  // RayTracing.h:15-22 - caller-owned cudaMallocPitch float4 surfaces
        {
            void* surface = nullptr;
            void* last_frame_surface = nullptr;
            int width = 0;
            int height = 0;
            size_t pitch = 0;
        };
        struct Options
        {
            std::string bunny_obj = "data/stanford-bunny.obj";        // RayTracing.cpp:35
            std::string env_dds = "data/sunset_uncompressed.dds";     // Scene.cpp:40
            bool cornell_box = true, stanford_bunny = true;           // RayTracing.cpp:24-25
            int samples_per_launch = 5;                               // main_raytracing.cu:167
            int max_bounces = 6;                                      // main_raytracing.cu:115
            uint32_t seed = 0;                                        // 0 => wall clock like RayTracing.cpp:219-220
            int kernel_variant = 0;                                   // see RtProcessParams::variant
        };

        CUDARayTracer(SurfaceData surface, Options options);
        explicit CUDARayTracer(SurfaceData surface) : CUDARayTracer(surface, Options()) {}
        ~CUDARayTracer();

        void OnResize(SurfaceData surface);  // RayTracing.cpp:72-77
        Scene& GetScene() { return *scene; }
        void Process();                      // RayTracing.cpp:205-235
        // Headless sink: restart accumulation, render `frames` Process() frames and copy the float4 surface into HOST
        // memory (host_pitch bytes per row). Replaces the D3D11 present of main.cpp:683-716. Synchronous.
        // Returns the host->device bytes this call moved (GPUScene by value per launch + refreshed spheres/materials).
        size_t RenderToHost(int frames, void* host_pixels, size_t host_pitch);
        // Progressive-state checkpoint: {frame index, both float4 surfaces, every pixel's XORWOW state}. After
        // LoadCheckpoint() the next Process() continues the accumulation bit for bit as if the run had never stopped
        // (the scene itself is not stored: the caller rebuilds it the way it built it before). Both calls synchronise.
        void SaveCheckpoint(const std::string& path);
        void LoadCheckpoint(const std::string& path);
        int GetFrameIndex() const { return frame_index; }
        Options& GetOptions() { return options; }
        unsigned long long GetRayCount() const;  // closest-hit queries since construction (device counter, synchronises)

        static void SetupCornellBox(Scene& scene);                                 // RayTracing.cpp:79-203
        static void SetupStanfordBunny(Scene& scene, const std::string& obj_path); // RayTracing.cpp:33-69
        static void SetupStanfordBunny(Scene& scene, const Loader::LoadedMesh& mesh);

    private:
        Options options;
        int frame_index = 0;
        std::unique_ptr<Scene> scene;
        RtRngState* rng_state = nullptr;  // device, 48 B x ceil(W*H/128)*128
        unsigned long long* d_ray_counter = nullptr;
        void* tile_workspace = nullptr;   // device: tile order + per-tile cost (rt_process_workspace_bytes), reset on resize
        uint64_t tile_workspace_bytes = 0;
        SurfaceData surface;
    };
}
