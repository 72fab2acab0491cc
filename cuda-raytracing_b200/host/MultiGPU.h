// Multi-GPU renderer (north-star item 5). The reference has no multi-GPU code; what is replicated per GPU is its
// per-frame protocol (RayTracing/RayTracing.cpp:205-235): camera update, Scene::Upload, launch. One host thread per device;
// scene, re-layout and environment map replicated on every device; device r renders its share of the samples for EVERY pixel
// with cuRAND subsequences [r * n, (r + 1) * n) (n = padded states per frame; device 0 == the single-GPU stream) into an fp32 sum
// buffer (blend_mode 1); one ncclReduce(sum, fp32) of the frame to device 0 over NVLink, a divide by the sample count, and the
// float4 frame goes to HOST memory. No other exchange exists on this path.
#pragma once
#include <cstdint>
#include <functional>
#include <memory>
#include <string>
#include <vector>
#include "RayTracing.h"

namespace RayTracing
{
    class MultiGPUTracer
    {
    public:
        struct Options
        {
            CUDARayTracer::Options tracer;      // assets, bounce limit, seed (0 = wall clock, drawn once and shared by all devices)
            int width = 0, height = 0;
            std::vector<int> devices;           // CUDA ordinals; empty = every visible device
        };
        struct Timing { double render_ms = 0, reduce_ms = 0, total_ms = 0; };  // per call: slowest device's launches and the reduce as the last device to arrive saw it (CUDA events); wall clock of the call

        explicit MultiGPUTracer(const Options& options);
        ~MultiGPUTracer();
        MultiGPUTracer(const MultiGPUTracer&) = delete;
        MultiGPUTracer& operator=(const MultiGPUTracer&) = delete;

        int DeviceCount() const { return (int)replicas.size(); }
        // camera of every replica (Camera::SetPosition / SetXAndle / SetYAndle); takes effect with the next Render
        void SetCamera(const float* position3, float angle_x_deg, float angle_y_deg);
        // Restart the per-pixel streams (curand_init on every device) and the accumulation.
        void Reseed(uint32_t seed);
        // total_spp samples per pixel, split evenly over the devices (the remainder goes to the first devices), in launches of at most
        // spp_per_launch; reduce; resolve; frame -> host_pixels (float4 rows, host_pitch bytes apart). Synchronous. The per-pixel
        // streams continue from call to call (progressive rendering: every call is a fresh estimate from fresh samples).
        void Render(int total_spp, int spp_per_launch, void* host_pixels, size_t host_pitch);
        unsigned long long RayCount() const;   // closest-hit queries of all devices since construction
        const Timing& LastTiming() const { return timing; }

    private:
        struct Replica;
        std::vector<std::unique_ptr<Replica>> replicas;
        Options options;
        uint32_t seed = 0;
        Timing timing;
        void ForEachDevice(const std::function<void(Replica&)>& fn);  // one host thread per device; rethrows the first failure
    };
}
