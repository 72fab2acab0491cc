// C handles over the C++ host mirror (include/b200rt_host.h). Exceptions stop here.
#include <cstdio>
#include <cstring>
#include <string>
#include "BVH.h"
#include "MeshLoader.h"
#include "MultiGPU.h"
#include "RayTracing.h"
#include "Scene.h"
#include "b200rt_host.h"
#include "../csrc/capi_internal.h"

using namespace RayTracing;

struct RtHostMesh { Loader::LoadedMesh mesh; };
struct RtHostScene { Scene* scene; bool owned; };
struct RtHostTracer { CUDARayTracer* tracer; RtHostScene scene_handle; };
struct RtHostMultiTracer { MultiGPUTracer* tracer; };

namespace
{
    template <class F> int Guard(F&& f)
    {
        try { return f(); }
        catch (const std::exception& e) { return rt_internal::Fail(RT_E_INVALID, e.what()); }
        catch (...) { return rt_internal::Fail(RT_E_INVALID, "unknown C++ exception"); }
    }
    Math::mat4 FromColumnMajor(const float* m)
    {
        Math::mat4 r(1.0f);
        if (m) for (int c = 0; c < 4; c++) for (int k = 0; k < 4; k++) r[c][k] = m[4 * c + k];
        return r;
    }
}

extern "C" {

int rt_host_mesh_load_obj(const char* obj_path, RtHostMesh** out) { return rt_host_mesh_load_obj_subdivided(obj_path, 0, out); }

int rt_host_mesh_load_obj_subdivided(const char* obj_path, int rounds, RtHostMesh** out)
{
    if (!obj_path || !out || rounds < 0 || rounds > 4) return rt_internal::Fail(RT_E_INVALID, "rt_host_mesh_load_obj: bad argument");
    return Guard([&] {
        std::vector<float> positions; std::vector<uint32_t> indices;
        Loader::ReadObj(obj_path, positions, indices);
        for (int r = 0; r < rounds; r++) Loader::SubdivideLoop(positions, indices);
        *out = new RtHostMesh{ Loader::BuildSmoothMesh(positions, indices, 100.0f) };
        return RT_OK;
    });
}

void rt_host_mesh_destroy(RtHostMesh* mesh) { delete mesh; }

int rt_host_mesh_counts(const RtHostMesh* mesh, uint32_t* vertex_count, uint32_t* face_count)
{
    if (!mesh) return rt_internal::Fail(RT_E_INVALID, "null mesh");
    if (vertex_count) *vertex_count = mesh->mesh.VertexCount();
    if (face_count) *face_count = mesh->mesh.FaceCount();
    return RT_OK;
}

int rt_host_mesh_copy(const RtHostMesh* mesh, float* positions, float* normals, uint32_t* indices, float* transform16)
{
    if (!mesh) return rt_internal::Fail(RT_E_INVALID, "null mesh");
    const auto& m = mesh->mesh;
    if (positions) memcpy(positions, m.positions.data(), m.positions.size() * 4);
    if (normals) memcpy(normals, m.normals.data(), m.normals.size() * 4);
    if (indices) memcpy(indices, m.indices.data(), m.indices.size() * 4);
    if (transform16) for (int c = 0; c < 4; c++) for (int k = 0; k < 4; k++) transform16[4 * c + k] = m.transform[c][k];
    return RT_OK;
}

RtHostScene* rt_host_scene_create(const char* env_dds_path)
{
    try { return new RtHostScene{ new Scene(env_dds_path ? env_dds_path : ""), true }; }
    catch (...) { return nullptr; }
}

void rt_host_scene_destroy(RtHostScene* s)
{
    if (!s) return;
    if (s->owned) delete s->scene;
    delete s;
}

int rt_host_scene_setup_cornell_box(RtHostScene* s)
{
    if (!s) return rt_internal::Fail(RT_E_INVALID, "null scene");
    return Guard([&] { CUDARayTracer::SetupCornellBox(*s->scene); return RT_OK; });
}

int rt_host_scene_setup_stanford_bunny(RtHostScene* s, const RtHostMesh* mesh)
{
    if (!s || !mesh) return rt_internal::Fail(RT_E_INVALID, "null scene/mesh");
    return Guard([&] { CUDARayTracer::SetupStanfordBunny(*s->scene, mesh->mesh); return RT_OK; });
}

int rt_host_scene_add_material(RtHostScene* s, const RtMaterial* material)
{
    if (!s || !material) return rt_internal::Fail(RT_E_INVALID, "null scene/material");
    return Guard([&] { Material m; static_cast<RtMaterial&>(m) = *material; return (int)s->scene->AddMaterial(m); });
}

int rt_host_scene_add_sphere(RtHostScene* s, const float* p, float radius, int material)
{
    if (!s || !p) return rt_internal::Fail(RT_E_INVALID, "null scene/position");
    return Guard([&] { s->scene->AddSphere(Math::vec3(p[0], p[1], p[2]), radius, material); return RT_OK; });
}

int rt_host_scene_add_triangle(RtHostScene* s, const float* a, const float* b, const float* c, int material)
{
    if (!s || !a || !b || !c) return rt_internal::Fail(RT_E_INVALID, "null scene/vertex");
    return Guard([&] { s->scene->AddTriangle(Math::vec3(a[0], a[1], a[2]), Math::vec3(b[0], b[1], b[2]), Math::vec3(c[0], c[1], c[2]), material); return RT_OK; });
}

int rt_host_scene_add_loaded_scene(RtHostScene* s, const RtHostMesh* mesh, const float* transform16, int material)
{
    if (!s || !mesh) return rt_internal::Fail(RT_E_INVALID, "null scene/mesh");
    return Guard([&] { s->scene->AddLoadedScene(mesh->mesh, FromColumnMajor(transform16), material); return RT_OK; });
}

int rt_host_scene_build_bvh(RtHostScene* s)
{
    if (!s) return rt_internal::Fail(RT_E_INVALID, "null scene");
    return Guard([&] { s->scene->BuildBVH(); return RT_OK; });
}

int rt_host_scene_counts(const RtHostScene* s, uint32_t* c)
{
    if (!s || !c) return rt_internal::Fail(RT_E_INVALID, "null scene/out");
    const Scene& sc = *s->scene;
    c[0] = (uint32_t)sc.Vertices().size(); c[1] = (uint32_t)sc.Faces().size(); c[2] = (uint32_t)sc.Spheres().size();
    c[3] = (uint32_t)sc.Materials().size(); c[4] = (uint32_t)sc.GetBVH().GetGPUBVHNodes().size(); c[5] = (uint32_t)sc.GetBVH().GetFaceIndices().size();
    return RT_OK;
}

int rt_host_scene_copy(const RtHostScene* s, RtVertex* vertices, RtFace* faces, RtSphere* spheres, RtMaterial* materials,
                       RtBVHNode* nodes, uint32_t* face_indices)
{
    if (!s) return rt_internal::Fail(RT_E_INVALID, "null scene");
    const Scene& sc = *s->scene;
    if (vertices) memcpy(vertices, sc.Vertices().data(), sc.Vertices().size() * sizeof(RtVertex));
    if (faces) memcpy(faces, sc.Faces().data(), sc.Faces().size() * sizeof(RtFace));
    if (spheres) memcpy(spheres, sc.Spheres().data(), sc.Spheres().size() * sizeof(RtSphere));
    if (materials) for (size_t i = 0; i < sc.Materials().size(); i++) materials[i] = sc.Materials()[i];
    if (nodes) memcpy(nodes, sc.GetBVH().GetGPUBVHNodes().data(), sc.GetBVH().GetGPUBVHNodes().size_bytes());
    if (face_indices) memcpy(face_indices, sc.GetBVH().GetFaceIndices().data(), sc.GetBVH().GetFaceIndices().size_bytes());
    return RT_OK;
}

int rt_host_scene_set_camera(RtHostScene* s, const float* p, float ax, float ay, float width, float height)
{
    if (!s) return rt_internal::Fail(RT_E_INVALID, "null scene");
    Camera& cam = s->scene->GetCamera();
    if (p) cam.SetPosition(Math::vec3(p[0], p[1], p[2]));
    cam.SetXAndle(ax); cam.SetYAndle(ay);
    cam.SetViewportSize(Math::vec2(width, height));
    cam.Update();
    s->scene->AddDirtyFlag(DirtyFlagValue::Samples);
    return RT_OK;
}

int rt_host_scene_get_camera(const RtHostScene* s, RtCamera* out, float* angles2)
{
    if (!s || !out) return rt_internal::Fail(RT_E_INVALID, "null scene/out");
    Camera& cam = s->scene->GetCamera();
    *out = static_cast<const RtCamera&>(cam);
    if (angles2) { angles2[0] = cam.GetXAngle(); angles2[1] = cam.GetYAngle(); }
    return RT_OK;
}

int rt_host_scene_add_dirty_flags(RtHostScene* s, uint32_t flags)
{
    if (!s) return rt_internal::Fail(RT_E_INVALID, "null scene");
    s->scene->AddDirtyFlags(flags);
    return RT_OK;
}

int rt_host_scene_upload(RtHostScene* s, RtRngState* d_rng_state)
{
    if (!s) return rt_internal::Fail(RT_E_INVALID, "null scene");
    if (rt_device_count() <= 0) return rt_internal::Fail(RT_E_CUDA, "rt_host_scene_upload: no CUDA device (there is no CPU fallback)");
    return Guard([&] { s->scene->Upload(d_rng_state); return RT_OK; });
}

const RtGPUScene* rt_host_scene_gpu(const RtHostScene* s) { return s ? static_cast<const RtGPUScene*>(s->scene) : nullptr; }
RtAccel* rt_host_scene_accel(const RtHostScene* s) { return s ? s->scene->GetAccel() : nullptr; }

int rt_host_tracer_create(void* surface, void* last, int width, int height, uint64_t pitch, const RtHostTracerOptions* o, RtHostTracer** out)
{
    if (!out || width <= 0 || height <= 0) return rt_internal::Fail(RT_E_INVALID, "rt_host_tracer_create: bad argument");
    return Guard([&] {
        CUDARayTracer::Options opt;
        opt.stanford_bunny = o && o->bunny_obj; opt.bunny_obj = opt.stanford_bunny ? o->bunny_obj : "";
        opt.env_dds = (o && o->env_dds) ? o->env_dds : "";
        opt.cornell_box = o ? o->cornell_box != 0 : true;
        if (o && o->samples_per_launch > 0) opt.samples_per_launch = o->samples_per_launch;
        if (o && o->max_bounces > 0) opt.max_bounces = o->max_bounces;
        if (o) { opt.seed = o->seed; opt.kernel_variant = o->kernel_variant; }
        CUDARayTracer::SurfaceData sd; sd.surface = surface; sd.last_frame_surface = last; sd.width = width; sd.height = height; sd.pitch = pitch;
        auto* t = new RtHostTracer{ new CUDARayTracer(sd, opt), { nullptr, false } };
        t->scene_handle.scene = &t->tracer->GetScene();
        *out = t;
        return RT_OK;
    });
}

void rt_host_tracer_destroy(RtHostTracer* t) { if (t) { delete t->tracer; delete t; } }

int rt_host_tracer_process(RtHostTracer* t)
{
    if (!t) return rt_internal::Fail(RT_E_INVALID, "null tracer");
    if (rt_device_count() <= 0) return rt_internal::Fail(RT_E_CUDA, "rt_host_tracer_process: no CUDA device (there is no CPU fallback)");
    return Guard([&] { t->tracer->Process(); return RT_OK; });
}

int rt_host_tracer_on_resize(RtHostTracer* t, void* surface, void* last, int width, int height, uint64_t pitch)
{
    if (!t) return rt_internal::Fail(RT_E_INVALID, "null tracer");
    CUDARayTracer::SurfaceData sd; sd.surface = surface; sd.last_frame_surface = last; sd.width = width; sd.height = height; sd.pitch = pitch;
    return Guard([&] { t->tracer->OnResize(sd); return RT_OK; });
}

int rt_host_tracer_render_to_host(RtHostTracer* t, int frames, void* host_pixels, uint64_t host_pitch, uint64_t* h2d_bytes)
{
    if (!t) return rt_internal::Fail(RT_E_INVALID, "null tracer");
    if (rt_device_count() <= 0) return rt_internal::Fail(RT_E_CUDA, "rt_host_tracer_render_to_host: no CUDA device (there is no CPU fallback)");
    return Guard([&] { const size_t n = t->tracer->RenderToHost(frames, host_pixels, host_pitch); if (h2d_bytes) *h2d_bytes = n; return RT_OK; });
}

int rt_host_tracer_save_checkpoint(RtHostTracer* t, const char* path)
{
    if (!t || !path) return rt_internal::Fail(RT_E_INVALID, "rt_host_tracer_save_checkpoint: bad argument");
    return Guard([&] { t->tracer->SaveCheckpoint(path); return RT_OK; });
}

int rt_host_tracer_load_checkpoint(RtHostTracer* t, const char* path)
{
    if (!t || !path) return rt_internal::Fail(RT_E_INVALID, "rt_host_tracer_load_checkpoint: bad argument");
    if (rt_device_count() <= 0) return rt_internal::Fail(RT_E_CUDA, "rt_host_tracer_load_checkpoint: no CUDA device (there is no CPU fallback)");
    return Guard([&] { t->tracer->LoadCheckpoint(path); return RT_OK; });
}

RtHostScene* rt_host_tracer_scene(RtHostTracer* t) { return t ? &t->scene_handle : nullptr; }
int rt_host_tracer_frame_index(const RtHostTracer* t) { return t ? t->tracer->GetFrameIndex() : -1; }

int rt_host_tracer_set_sampling(RtHostTracer* t, int spp, int bounces, int variant)
{
    if (!t || spp <= 0 || bounces <= 0) return rt_internal::Fail(RT_E_INVALID, "rt_host_tracer_set_sampling: bad argument");
    auto& o = t->tracer->GetOptions();
    o.samples_per_launch = spp; o.max_bounces = bounces; o.kernel_variant = variant;
    return RT_OK;
}

unsigned long long rt_host_tracer_ray_count(const RtHostTracer* t) { return t ? t->tracer->GetRayCount() : 0; }

int rt_host_multi_tracer_create(int width, int height, const RtHostTracerOptions* o, const int* devices, int device_count, RtHostMultiTracer** out)
{
    if (!out || width <= 0 || height <= 0 || device_count < 0 || (device_count && !devices)) return rt_internal::Fail(RT_E_INVALID, "rt_host_multi_tracer_create: bad argument");
    if (rt_device_count() <= 0) return rt_internal::Fail(RT_E_CUDA, "rt_host_multi_tracer_create: no CUDA device (there is no CPU fallback)");
    return Guard([&] {
        MultiGPUTracer::Options m;
        m.width = width; m.height = height;
        m.devices.assign(devices, devices + device_count);
        if (o)
        {
            m.tracer.stanford_bunny = o->bunny_obj != nullptr;
            if (o->bunny_obj) m.tracer.bunny_obj = o->bunny_obj;
            m.tracer.env_dds = o->env_dds ? o->env_dds : "";
            m.tracer.cornell_box = o->cornell_box != 0;
            if (o->max_bounces > 0) m.tracer.max_bounces = o->max_bounces;
            m.tracer.seed = o->seed;
        }
        *out = new RtHostMultiTracer{ new MultiGPUTracer(m) };
        return RT_OK;
    });
}

void rt_host_multi_tracer_destroy(RtHostMultiTracer* t) { if (t) { delete t->tracer; delete t; } }
int rt_host_multi_tracer_device_count(const RtHostMultiTracer* t) { return t ? t->tracer->DeviceCount() : 0; }

int rt_host_multi_tracer_set_camera(RtHostMultiTracer* t, const float* position3, float angle_x_deg, float angle_y_deg)
{
    if (!t) return rt_internal::Fail(RT_E_INVALID, "null tracer");
    return Guard([&] { t->tracer->SetCamera(position3, angle_x_deg, angle_y_deg); return RT_OK; });
}

int rt_host_multi_tracer_reseed(RtHostMultiTracer* t, uint32_t seed)
{
    if (!t) return rt_internal::Fail(RT_E_INVALID, "null tracer");
    return Guard([&] { t->tracer->Reseed(seed); return RT_OK; });
}

int rt_host_multi_tracer_render(RtHostMultiTracer* t, int total_spp, int spp_per_launch, void* host_pixels, uint64_t host_pitch, double* timing_ms3)
{
    if (!t) return rt_internal::Fail(RT_E_INVALID, "null tracer");
    return Guard([&] {
        t->tracer->Render(total_spp, spp_per_launch, host_pixels, host_pitch);
        if (timing_ms3) { const auto& tm = t->tracer->LastTiming(); timing_ms3[0] = tm.render_ms; timing_ms3[1] = tm.reduce_ms; timing_ms3[2] = tm.total_ms; }
        return RT_OK;
    });
}

unsigned long long rt_host_multi_tracer_ray_count(const RtHostMultiTracer* t) { return t ? t->tracer->RayCount() : 0; }

int rt_host_write_ppm(const char* path, const uint8_t* rgba8, int width, int height)
{
    if (!path || !rgba8 || width <= 0 || height <= 0) return rt_internal::Fail(RT_E_INVALID, "rt_host_write_ppm: bad argument");
    FILE* f = fopen(path, "wb");
    if (!f) return rt_internal::Fail(RT_E_IO, std::string("cannot write ") + path);
    fprintf(f, "P6\n%d %d\n255\n", width, height);
    std::vector<uint8_t> row((size_t)width * 3);
    for (int y = 0; y < height; y++)
    {
        for (int x = 0; x < width; x++) for (int c = 0; c < 3; c++) row[3 * x + c] = rgba8[((size_t)y * width + x) * 4 + c];
        fwrite(row.data(), 1, row.size(), f);
    }
    fclose(f);
    return RT_OK;
}

}  // extern "C"
