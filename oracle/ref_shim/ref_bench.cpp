// TEST INFRASTRUCTURE (oracle/_ref). CPU-baseline build of the reference kernel body for bench.py's
// `cpu_baseline` / `--impl reference` legs: /root/reference/RayTracing/main_raytracing.cu compiled
// host-side like ref_api.cpp, with the mechanical edits oracle/build_ref.sh lists (the two hard-coded
// constants `num_bounces` :115 and `sample_count` :167 read from variables so BASELINE.json's
// 4-bounce config can be timed, and a counter of GetRayHit calls, the unit of the Mrays/s metric).
// The parity oracle (ref_api.cpp) keeps the unedited body.
#include "host_pre.h"
#include "Math.h"
#include "GPUScene.h"
#include <chrono>
#include <omp.h>

static int refbench_bounces = 6, refbench_spp = 5;
static thread_local unsigned long long refbench_rays = 0;

namespace refbench
{
#include RT_BENCH_SRC
}

// Renders rows row_begin, row_begin+row_stride, ... (< height) of one frame with all host threads.
// Returns seconds; *out_rays = GetRayHit calls.
extern "C" __attribute__((visibility("default")))
double ref_bench_render(const void* vertices, const void* faces, const void* nodes, const uint32_t* face_indices,
                        const void* spheres, uint32_t sphere_count, const void* materials, uint32_t material_count,
                        const void* gpu_camera60, const float* cubemap_level0, int cubemap_size,
                        void* rng_states48, void* surface, void* surface_last_frame,
                        int width, int height, uint64_t pitch, int frame_index, int row_begin, int row_stride,
                        int spp, int bounces, unsigned long long* out_rays, int* out_threads)
{
    using namespace refbench;
    refbench_bounces = bounces; refbench_spp = spp;
    OracleCubemap cm; cm.texels = cubemap_level0; cm.size = cubemap_size;
    RenderData rd{};
    rd.surface = (unsigned char*)surface; rd.surface_last_frame = (unsigned char*)surface_last_frame;
    rd.width = width; rd.height = height; rd.pitch = pitch; rd.frame_index = frame_index;
    rd.scene.gpu_spheres = (const GeometrySphere*)spheres; rd.scene.gpu_materials = (const GPUMaterial*)materials;
    rd.scene.gpu_bvh_nodes = (const GPUBVHNode*)nodes; rd.scene.gpu_bvh_face_indices = face_indices;
    rd.scene.gpu_vertices = (const GPUVertex*)vertices; rd.scene.gpu_faces = (const GPUFace*)faces;
    rd.scene.sphere_count = (int)sphere_count; rd.scene.material_count = (int)material_count;
    rd.scene.rng_state = (curandState*)rng_states48;
    rd.scene.environment_cubemap_tex = (cudaTextureObject_t)(uintptr_t)&cm;
    memcpy(&rd.scene.camera, gpu_camera60, sizeof(GPUCamera));
    if (row_stride <= 0) row_stride = 1;
    const int rows = (height - row_begin + row_stride - 1) / row_stride;
    unsigned long long rays = 0;
    const auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : rays)
    for (int k = 0; k < rows; k++)
    {
        const int y = row_begin + k * row_stride;
        refbench_rays = 0;
        blockDim = dim3(16, 16, 1); gridDim = dim3((width + 15) / 16, (height + 15) / 16, 1);
        for (int x = 0; x < width; x++)
        {
            blockIdx.x = x / 16; blockIdx.y = y / 16; blockIdx.z = 0;
            threadIdx.x = x % 16; threadIdx.y = y % 16; threadIdx.z = 0;
            raytracing_kernel_main(rd);
        }
        rays += refbench_rays;
    }
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (out_rays) *out_rays = rays;
    if (out_threads) *out_threads = omp_get_max_threads();
    return s;
}
