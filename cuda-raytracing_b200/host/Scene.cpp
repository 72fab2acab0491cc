#include <algorithm>
#include "Scene.h"
#include "BVH.h"

namespace RayTracing
{
    void Camera::Update()  // RayTracing/Scene.cpp:15-36
    {
        aspect = viewport_size.x / viewport_size.y;
        transform = Math::ComposeMatrix(vec3(origin), quat_from_euler(vec3(radians(GetXAngle()), radians(GetYAngle()), 0)), vec3(1.0f));
        projection = perspectiveRH_NO(radians(fov_y), aspect, 1.0f, 1000.0f);
        const mat4 inv_proj = inverse(projection);
        const vec4 lower_left_corner4 = inv_proj * vec4(-1, -1, -1, 1);
        const vec4 upper_right_corner4 = inv_proj * vec4(1, 1, -1, 1);
        // lower left corner of the front clip rect in camera space
        const vec3 llc = (lower_left_corner4 / lower_left_corner4.w).xyz();
        const vec3 upper_right_corner = (upper_right_corner4 / upper_right_corner4.w).xyz();
        const vec3 size = upper_right_corner - llc;
        viewport_worldspace_size = RtVec2{ size.x, size.y };
        // to world space
        horizontal = (transform * vec4(viewport_worldspace_size.x, 0, 0, 0)).xyz();
        vertical = (transform * vec4(0, viewport_worldspace_size.y, 0, 0)).xyz();
        lower_left_corner = (transform * vec4(llc, 1)).xyz();
    }

    void Scene::AddTriangle(vec3 a, vec3 b, vec3 c, int material)  // RayTracing/Scene.cpp:46-67
    {
        const vec3 normal = normalize(cross(c - b, a - b));
        const uint32_t index0 = (uint32_t)vertices.size();
        const vec3 p[3] = { a, b, c };
        for (const vec3& q : p)
        {
            RtVertex v; v.position = q; v.normal = normal; v.uv = RtVec2{ 0, 0 };
            vertices.push_back(v);
        }
        // aggregate order of GPUFace is {v0, v2, v1, material}: the kernel sees (a, c, b)
        faces.push_back(RtFace{ index0, index0 + 1, index0 + 2, (uint32_t)material });
        AddDirtyFlag(DirtyFlagValue::SceneMemory);
        AddDirtyFlag(DirtyFlagValue::BVH);  // the reference forgets this (harmless there: flags start all-dirty)
    }

    void Scene::AddSphere(vec3 position, float radius, int material)  // Scene.cpp:69-73
    {
        RtSphere s{}; s.position = position; s.radius = radius; s.material = material;
        spheres.push_back(s);
        AddDirtyFlag(DirtyFlagValue::SceneMemory);
    }

    void Scene::AddLoadedScene(const Loader::LoadedMesh& mesh, const mat4& transform, int default_material)  // Scene.cpp:75-132
    {
        AddDirtyFlag(DirtyFlagValue::BVH);
        AddDirtyFlag(DirtyFlagValue::SceneMemory);
        if (mesh.normals.size() != mesh.positions.size()) throw std::runtime_error("no normals");
        const uint32_t index_offset = (uint32_t)vertices.size();
        const mat4 mesh_transform = transform * mesh.transform;
        const uint32_t nv = mesh.VertexCount(), nf = mesh.FaceCount();
        // geometric growth: an exact reserve per call would re-copy the whole scene for every instance of a mesh grid
        const size_t need_v = vertices.size() + nv + size_t(nf) * 3, need_f = faces.size() + size_t(nf) * 2;
        if (vertices.capacity() < need_v) vertices.reserve(std::max(need_v, vertices.capacity() * 2));
        if (faces.capacity() < need_f) faces.reserve(std::max(need_f, faces.capacity() * 2));
        auto P = [&](uint32_t i) { return vec4(mesh.positions[3 * i], mesh.positions[3 * i + 1], mesh.positions[3 * i + 2], 1.0f); };
        for (uint32_t v = 0; v < nv; v++)
        {
            RtVertex vertex;
            vertex.position = (mesh_transform * P(v)).xyz();
            vertex.normal = (mesh_transform * vec4(mesh.normals[3 * v], mesh.normals[3 * v + 1], mesh.normals[3 * v + 2], 0.0f)).xyz();
            vertex.uv = RtVec2{ 0, 0 };
            vertices.push_back(vertex);
        }
        for (uint32_t f = 0; f < nf; f++)
        {
            const uint32_t i0 = mesh.indices[3 * f], i1 = mesh.indices[3 * f + 1], i2 = mesh.indices[3 * f + 2];
            RtFace gpu_face;  // smooth, indexed copy: fields set by NAME (Scene.cpp:110-115)
            gpu_face.v0 = i0 + index_offset; gpu_face.v1 = i1 + index_offset; gpu_face.v2 = i2 + index_offset;
            gpu_face.material = (uint32_t)default_material;
            faces.push_back(gpu_face);
            // ... and the same triangle again as a flat-shaded one (Scene.cpp:117-127)
            AddTriangle((mesh_transform * P(i0)).xyz(), (mesh_transform * P(i1)).xyz(), (mesh_transform * P(i2)).xyz(), default_material);
        }
    }

    uint32_t Scene::AddMaterial(Material material)  // Scene.cpp:134-139
    {
        AddDirtyFlag(DirtyFlagValue::SceneMemory);
        materials.push_back(material);
        return (uint32_t)materials.size() - 1;
    }

    void Scene::Update(float)  // Scene.cpp:141-180 minus GetAsyncKeyState polling (Win32)
    {
        camera.Update();
    }

    void Scene::BuildBVH()
    {
        if (bvh_current && bvh) return;  // nothing changed since the last build (an explicit BuildBVH() followed by Upload())
        if (!bvh) bvh = std::make_unique<BVH>();
        bvh->Calculate(vertices, faces);
        bvh_current = true;
    }
}
