// TEST INFRASTRUCTURE (oracle). Not part of the product path; only tests/, smoke() and the
// bench's cpu_baseline / --impl reference legs may execute anything under oracle/.
//
// CPU restatement of what `texCubemapLod<float4>(tex, x, y, z, 0)` returns for the texture
// object the reference creates in utils/CUDATexture.cpp:156-171 (cubemap, RGBA32F, linear
// filter, normalised coordinates, seamless edges, level 0 only), used by the reference's miss
// shader at RayTracing/main_raytracing.cu:151-153.
//
// Parity: UNPINNED against hardware (no GPU in the build container, no graphics stack); the
// hardware filter uses 1.8 fixed-point weights (CUDA C Programming Guide, "Linear Filtering"),
// which this restatement emulates. Images are compared by RMSE, never bitwise.
#pragma once
#include <cmath>
#include <cstdint>

struct OracleCubemap
{
    const float* texels = nullptr;  // level 0, face-major: [6][size][size][4], faces +x -x +y -y +z -z
    int size = 0;
};

struct OracleFloat4 { float x, y, z, w; };

namespace oracle_cubemap_detail
{
    // Cube face selection and (s,t) in [-1,1] (CUDA C Programming Guide, "Cubemap Textures").
    inline int select_face(float x, float y, float z, float& sc, float& tc, float& ma)
    {
        const float ax = std::fabs(x), ay = std::fabs(y), az = std::fabs(z);
        if (ax >= ay && ax >= az) { ma = ax; if (x >= 0) { sc = -z; tc = -y; return 0; } sc = z; tc = -y; return 1; }
        if (ay >= az)             { ma = ay; if (y >= 0) { sc = x; tc = z; return 2; }  sc = x; tc = -z; return 3; }
        ma = az; if (z >= 0) { sc = x; tc = -y; return 4; } sc = -x; tc = -y; return 5;
    }

    // Inverse of select_face: direction through face-plane point (sc,tc), |ma| = 1.
    inline void face_to_dir(int face, float sc, float tc, float& x, float& y, float& z)
    {
        switch (face)
        {
        case 0: x = 1;  z = -sc; y = -tc; break;
        case 1: x = -1; z = sc;  y = -tc; break;
        case 2: y = 1;  x = sc;  z = tc;  break;
        case 3: y = -1; x = sc;  z = -tc; break;
        case 4: z = 1;  x = sc;  y = -tc; break;
        default: z = -1; x = -sc; y = -tc; break;
        }
    }

    inline const float* texel(const OracleCubemap& cm, int face, int i, int j)
    {
        return cm.texels + ((size_t(face) * cm.size + j) * cm.size + i) * 4;
    }

    // Fetch tap (i,j) of `face`; taps outside the face are taken from the adjacent face
    // (seamless filtering). Returns false for a corner tap (outside in both axes).
    inline bool tap(const OracleCubemap& cm, int face, int i, int j, const float*& out)
    {
        const int n = cm.size;
        const bool oi = i < 0 || i >= n, oj = j < 0 || j >= n;
        if (!oi && !oj) { out = texel(cm, face, i, j); return true; }
        if (oi && oj) return false;
        const float sc = (2.0f * (i + 0.5f)) / n - 1.0f, tc = (2.0f * (j + 0.5f)) / n - 1.0f;
        float x, y, z; face_to_dir(face, sc, tc, x, y, z);
        float s2, t2, ma; const int f2 = select_face(x, y, z, s2, t2, ma);
        int i2 = (int)std::floor((s2 / ma * 0.5f + 0.5f) * n), j2 = (int)std::floor((t2 / ma * 0.5f + 0.5f) * n);
        i2 = i2 < 0 ? 0 : (i2 >= n ? n - 1 : i2); j2 = j2 < 0 ? 0 : (j2 >= n ? n - 1 : j2);
        out = texel(cm, f2, i2, j2);
        return true;
    }
}

inline OracleFloat4 oracle_tex_cubemap_lod0(const OracleCubemap& cm, float x, float y, float z)
{
    using namespace oracle_cubemap_detail;
    float sc, tc, ma;
    const int face = select_face(x, y, z, sc, tc, ma);
    const int n = cm.size;
    const float s = (sc / ma + 1.0f) * 0.5f, t = (tc / ma + 1.0f) * 0.5f;
    const float xb = s * n - 0.5f, yb = t * n - 0.5f;
    const float fi = std::floor(xb), fj = std::floor(yb);
    // 1.8 fixed-point filter weights
    const float a = std::floor((xb - fi) * 256.0f + 0.5f) / 256.0f;
    const float b = std::floor((yb - fj) * 256.0f + 0.5f) / 256.0f;
    const int i = (int)fi, j = (int)fj;
    const float w[4] = { (1 - a) * (1 - b), a * (1 - b), (1 - a) * b, a * b };
    const int di[4] = { 0, 1, 0, 1 }, dj[4] = { 0, 0, 1, 1 };
    float acc[4] = { 0, 0, 0, 0 }, wsum = 0;
    for (int k = 0; k < 4; k++)
    {
        const float* p;
        if (!tap(cm, face, i + di[k], j + dj[k], p)) continue;
        wsum += w[k];
        for (int c = 0; c < 4; c++) acc[c] += w[k] * p[c];
    }
    if (wsum > 0 && wsum != 1.0f) for (int c = 0; c < 4; c++) acc[c] /= wsum;
    return { acc[0], acc[1], acc[2], acc[3] };
}
