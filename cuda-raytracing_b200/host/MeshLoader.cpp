#include "MeshLoader.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <unordered_map>

namespace Loader
{
    using namespace Math;

    mat4 ImportRootTransform()
    {
        // Assimp matrix4x4.inl RotationX: b2 = c3 = cos(a); c2 = sin(a); b3 = -sin(a) (row-major rows a..d);
        // CopyNodes reinterprets it as column-major and transposes, i.e. column j of the result holds (a_j, b_j, c_j, d_j).
        const float a = -(float)M_PI / 2;
        const float c = std::cos(a), s = std::sin(a);
        mat4 m(1.0f);
        m[1][1] = c;   // b2
        m[2][2] = c;   // c3
        m[1][2] = s;   // c2  (row c, column 2)
        m[2][1] = -s;  // b3  (row b, column 3)
        return m;
    }

    void ReadObj(const std::string& path, std::vector<float>& positions, std::vector<uint32_t>& indices)
    {
        FILE* f = fopen(path.c_str(), "rb");
        if (!f) throw std::runtime_error("Failed loading scene: " + path);
        char line[1024];
        std::vector<int64_t> corner;
        while (fgets(line, sizeof(line), f))
        {
            const char* p = line;
            while (*p == ' ' || *p == '\t') p++;
            if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t'))
            {
                char* e = nullptr; p += 2;
                for (int k = 0; k < 3; k++) { positions.push_back(strtof(p, &e)); p = e; }
            }
            else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t'))
            {
                corner.clear(); p += 2;
                for (;;)
                {
                    while (*p == ' ' || *p == '\t') p++;
                    if (*p == 0 || *p == '\n' || *p == '\r') break;
                    char* e = nullptr;
                    long v = strtol(p, &e, 10);
                    if (e == p) break;
                    corner.push_back(v);
                    p = e;
                    while (*p && *p != ' ' && *p != '\t' && *p != '\n' && *p != '\r') p++;  // skip /vt/vn
                }
                const int64_t nv = (int64_t)(positions.size() / 3);
                for (auto& c : corner) c = c < 0 ? nv + c : c - 1;
                for (size_t k = 1; k + 1 < corner.size(); k++)  // fan triangulation
                {
                    indices.push_back((uint32_t)corner[0]); indices.push_back((uint32_t)corner[k]); indices.push_back((uint32_t)corner[k + 1]);
                }
            }
        }
        fclose(f);
        const uint32_t nv = (uint32_t)(positions.size() / 3);
        for (uint32_t i : indices) if (i >= nv) throw std::runtime_error("OBJ face index out of range: " + path);
    }

    LoadedMesh BuildSmoothMesh(const std::vector<float>& positions, const std::vector<uint32_t>& indices, float max_angle_deg)
    {
        const uint32_t nv = (uint32_t)(positions.size() / 3), nf = (uint32_t)(indices.size() / 3);
        auto P = [&](uint32_t i) { return vec3(positions[3 * i], positions[3 * i + 1], positions[3 * i + 2]); };
        std::vector<vec3> face_normal(nf);
        for (uint32_t f = 0; f < nf; f++)
        {
            const vec3 p1 = P(indices[3 * f]), p2 = P(indices[3 * f + 1]), p3 = P(indices[3 * f + 2]);
            const vec3 n = cross(p2 - p1, p3 - p1);
            const float len = std::sqrt(dot(n, n));
            face_normal[f] = len > 0 ? n / len : n;
        }
        // position -> incident faces (CSR)
        std::vector<uint32_t> start(nv + 1, 0);
        for (uint32_t i : indices) start[i + 1]++;
        for (uint32_t v = 0; v < nv; v++) start[v + 1] += start[v];
        std::vector<uint32_t> incident(indices.size()), fill(start.begin(), start.end() - 1);
        for (uint32_t f = 0; f < nf; f++) for (int k = 0; k < 3; k++) incident[fill[indices[3 * f + k]]++] = f;

        const float limit = std::cos(radians(max_angle_deg));
        LoadedMesh mesh; mesh.transform = ImportRootTransform();
        mesh.indices.resize(indices.size());
        struct Key { uint32_t v; uint32_t n[3]; bool operator==(const Key& o) const { return memcmp(this, &o, sizeof(Key)) == 0; } };
        struct KeyHash { size_t operator()(const Key& k) const { uint64_t h = 1469598103934665603ull; const uint32_t w[4] = { k.v, k.n[0], k.n[1], k.n[2] }; for (uint32_t x : w) { h ^= x; h *= 1099511628211ull; } return (size_t)h; } };
        std::unordered_map<Key, uint32_t, KeyHash> joined; joined.reserve(nv * 2);
        for (uint32_t f = 0; f < nf; f++)
            for (int k = 0; k < 3; k++)
            {
                const uint32_t v = indices[3 * f + k];
                const vec3 own = face_normal[f];
                vec3 sum(0.0f);
                for (uint32_t s = start[v]; s < start[v + 1]; s++)
                {
                    const vec3 other = face_normal[incident[s]];
                    if (dot(other, own) >= limit * std::sqrt(dot(other, other)) * std::sqrt(dot(own, own))) sum = sum + other;
                }
                const float len = std::sqrt(dot(sum, sum));
                const vec3 n = len > 0 ? sum / len : sum;
                Key key; key.v = v; memcpy(key.n, &n, 12);
                auto it = joined.find(key);
                if (it == joined.end())
                {
                    it = joined.emplace(key, mesh.VertexCount()).first;
                    const vec3 p = P(v);
                    mesh.positions.insert(mesh.positions.end(), { p.x, p.y, p.z });
                    mesh.normals.insert(mesh.normals.end(), { n.x, n.y, n.z });
                }
                mesh.indices[3 * f + k] = it->second;
            }
        return mesh;
    }

    LoadedMesh ImportScene(const std::string& path)
    {
        std::vector<float> positions; std::vector<uint32_t> indices;
        ReadObj(path, positions, indices);
        return BuildSmoothMesh(positions, indices, 100.0f);
    }
}
