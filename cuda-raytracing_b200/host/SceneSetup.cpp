// The fixed scene the reference's CUDARayTracer constructor assembles (host only, no device work).
#include "RayTracing.h"

namespace RayTracing
{
    void CUDARayTracer::SetupCornellBox(Scene& scene)  // RayTracing/RayTracing.cpp:79-203
    {
        struct Wall { vec3 A, B, C, D, albedo, emissive; };
        const Wall walls[6] = {
            // back wall, floor, ceiling, left wall, right wall, light
            { vec3(-12.6f, -12.6f, 25.0f), vec3(12.6f, -12.6f, 25.0f), vec3(12.6f, 12.6f, 25.0f), vec3(-12.6f, 12.6f, 25.0f), vec3(0.7f, 0.7f, 0.7f), vec3(0.0f) },
            { vec3(-12.6f, -12.45f, 25.0f), vec3(12.6f, -12.45f, 25.0f), vec3(12.6f, -12.45f, 15.0f), vec3(-12.6f, -12.45f, 15.0f), vec3(0.7f, 0.7f, 0.7f), vec3(0.0f) },
            { vec3(-12.6f, 12.5f, 25.0f), vec3(12.6f, 12.5f, 25.0f), vec3(12.6f, 12.5f, 15.0f), vec3(-12.6f, 12.5f, 15.0f), vec3(0.7f, 0.7f, 0.7f), vec3(0.0f) },
            { vec3(-12.5f, -12.6f, 25.0f), vec3(-12.5f, -12.6f, 15.0f), vec3(-12.5f, 12.6f, 15.0f), vec3(-12.5f, 12.6f, 25.0f), vec3(0.1f, 0.7f, 0.1f), vec3(0.0f) },
            { vec3(12.5f, -12.6f, 25.0f), vec3(12.5f, -12.6f, 15.0f), vec3(12.5f, 12.6f, 15.0f), vec3(12.5f, 12.6f, 25.0f), vec3(0.7f, 0.1f, 0.1f), vec3(0.0f) },
            { vec3(-5.0f, 12.4f, 22.5f), vec3(5.0f, 12.4f, 22.5f), vec3(5.0f, 12.4f, 17.5f), vec3(-5.0f, 12.4f, 17.5f), vec3(0.0f), vec3(1.0f, 0.9f, 0.7f) * 20.0f },
        };
        for (const Wall& w : walls)
            scene.AddQuad(w.A, w.B, w.C, w.D, (int)scene.AddMaterial(Material(w.albedo, w.emissive)));

        auto ball = [&scene](vec3 albedo, float specular_percent, vec4 specular, float roughness, vec3 position, float radius)
        {
            Material m(albedo);
            m.specular_percent = specular_percent; m.specular = specular; m.roughness = roughness;
            scene.AddSphere(position, radius, (int)scene.AddMaterial(m));
        };
        ball(vec3(0.9f, 0.9f, 0.50f), 0.5f, vec4(0.9f, 0.9f, 0.9f, 0), 0.2f, vec3(-9.0f, -9.5f, 20.0f), 3);
        ball(vec3(0.9f, 0.5f, 0.90f), 0.3f, vec4(0.9f, 0.9f, 0.9f, 0), 0.2f, vec3(0.0f, -9.5f, 20.0f), 3);
        ball(vec3(0.0f, 0.0f, 1.0f), 0.5f, vec4(1.0f, 0.0f, 0.0f, 0), 0.4f, vec3(9.0f, -9.5f, 20.0f), 3);
        scene.GetCamera().SetYAndle(180);  // RayTracing.cpp:159
        // shiny green balls of varying roughness
        const float roughness[5] = { 0.0f, 0.25f, 0.5f, 0.75f, 0.97f };
        for (int i = 0; i < 5; i++)
            ball(vec3(1.0f), 1.0f, vec4(0.3f, 1.0f, 0.3f, 0.0f), roughness[i], vec3(-10.0f + 5.0f * (float)i, 0.0f, 23.0f), 1.75f);
    }

    void CUDARayTracer::SetupStanfordBunny(Scene& scene, const Loader::LoadedMesh& mesh)  // RayTracing.cpp:33-69
    {
        mat4 m = translate(mat4(1.0f), vec3(30, -18, 20));
        m = rotate(m, -(float)M_PI, vec3(0, 1, 0));
        m = rotate(m, (float)M_PI / 2, vec3(1, 0, 0));
        m = scale(m, vec3(150.0f));
        Material material(vec3(1.0f));
        material.specular_percent = 0.5f;
        material.roughness = 0.8f;
        material.specular = vec4(0.3f, 1.0f, 0.3f, 0.0f);
        scene.AddLoadedScene(mesh, m, (int)scene.AddMaterial(material));
        {   // 100 x 100 floor
            const vec3 offset(20, 0, 0), scl(50, 1, 50);
            const vec3 A = scl * vec3(-1.0f, -12.45f, 1.0f) + offset, B = scl * vec3(1.0f, -12.45f, 1.0f) + offset;
            const vec3 C = scl * vec3(1.0f, -12.45f, -1.0f) + offset, D = scl * vec3(-1.0f, -12.45f, -1.0f) + offset;
            scene.AddQuad(A, B, C, D, (int)scene.AddMaterial(Material(vec3(0.7f, 0.7f, 0.7f), vec3(0.0f))));
        }
        // emissive sphere
        scene.AddSphere(vec3(30, 10, 40), 8, (int)scene.AddMaterial(Material(vec3(0.0f), vec3(0.3f, 0.9f, 0.7f) * 10.0f)));
    }

    void CUDARayTracer::SetupStanfordBunny(Scene& scene, const std::string& obj_path)
    {
        SetupStanfordBunny(scene, Loader::ImportScene(obj_path));
    }
}
