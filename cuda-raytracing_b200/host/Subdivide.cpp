// Loop subdivision of an indexed triangle mesh (BASELINE.json config 4a: "loop-subdivided bunny").
// Deterministic: new vertices are numbered in first-seen edge order; face k splits into faces
// 4k..4k+3 = (a,ab,ca), (b,bc,ab), (c,ca,bc), (ab,bc,ca).
#include <cmath>
#include <unordered_map>
#include "MeshLoader.h"

namespace Loader
{
    void SubdivideLoop(std::vector<float>& positions, std::vector<uint32_t>& indices)
    {
        const uint32_t nv = (uint32_t)(positions.size() / 3), nf = (uint32_t)(indices.size() / 3);
        struct Edge { uint32_t opposite[2]; uint32_t count; uint32_t vertex; };
        std::unordered_map<uint64_t, Edge> edges; edges.reserve((size_t)nf * 2);
        auto key = [](uint32_t a, uint32_t b) { return a < b ? ((uint64_t)a << 32) | b : ((uint64_t)b << 32) | a; };
        for (uint32_t f = 0; f < nf; f++)
            for (int k = 0; k < 3; k++)
            {
                const uint32_t a = indices[3 * f + k], b = indices[3 * f + (k + 1) % 3], c = indices[3 * f + (k + 2) % 3];
                Edge& e = edges.try_emplace(key(a, b), Edge{ { 0, 0 }, 0, 0 }).first->second;
                if (e.count < 2) e.opposite[e.count] = c;
                e.count++;
            }
        // neighbours for the vertex rule
        std::vector<std::vector<uint32_t>> ring(nv);
        std::vector<uint8_t> boundary(nv, 0);
        for (const auto& kv : edges)
        {
            const uint32_t a = (uint32_t)(kv.first >> 32), b = (uint32_t)kv.first;
            ring[a].push_back(b); ring[b].push_back(a);
            if (kv.second.count != 2) { boundary[a] = 1; boundary[b] = 1; }
        }
        auto P = [&](uint32_t i, int c) { return (double)positions[3 * i + c]; };
        std::vector<float> out((size_t)nv * 3);
        for (uint32_t v = 0; v < nv; v++)
        {
            if (boundary[v])
            {   // crease rule: 3/4 v + 1/8 (two boundary neighbours)
                double acc[3] = { 0, 0, 0 }; int nb = 0;
                for (uint32_t n : ring[v])
                {
                    const auto it = edges.find(key(v, n));
                    if (it->second.count != 2) { for (int c = 0; c < 3; c++) acc[c] += P(n, c); nb++; }
                }
                for (int c = 0; c < 3; c++) out[3 * v + c] = nb == 2 ? (float)(0.75 * P(v, c) + 0.125 * acc[c]) : positions[3 * v + c];
            }
            else
            {
                const size_t n = ring[v].size();
                const double t = 0.375 + 0.25 * std::cos(2.0 * M_PI / (double)n);
                const double beta = (0.625 - t * t) / (double)n;
                double acc[3] = { 0, 0, 0 };
                for (uint32_t r : ring[v]) for (int c = 0; c < 3; c++) acc[c] += P(r, c);
                for (int c = 0; c < 3; c++) out[3 * v + c] = (float)((1.0 - n * beta) * P(v, c) + beta * acc[c]);
            }
        }
        // edge vertices, numbered in face/edge order (deterministic)
        std::vector<uint32_t> new_indices; new_indices.reserve((size_t)nf * 12);
        uint32_t next = nv;
        for (uint32_t f = 0; f < nf; f++)
        {
            uint32_t mid[3];
            for (int k = 0; k < 3; k++)
            {
                const uint32_t a = indices[3 * f + k], b = indices[3 * f + (k + 1) % 3];
                Edge& e = edges.find(key(a, b))->second;
                if (e.vertex == 0)
                {
                    e.vertex = next++;
                    for (int c = 0; c < 3; c++)
                    {
                        const double v = e.count == 2 ? 0.375 * (P(a, c) + P(b, c)) + 0.125 * (P(e.opposite[0], c) + P(e.opposite[1], c))
                                                      : 0.5 * (P(a, c) + P(b, c));
                        out.push_back((float)v);
                    }
                }
                mid[k] = e.vertex;
            }
            const uint32_t a = indices[3 * f], b = indices[3 * f + 1], c = indices[3 * f + 2];
            const uint32_t ab = mid[0], bc = mid[1], ca = mid[2];
            const uint32_t t[12] = { a, ab, ca, b, bc, ab, c, ca, bc, ab, bc, ca };
            new_indices.insert(new_indices.end(), t, t + 12);
        }
        positions.swap(out);
        indices.swap(new_indices);
    }
}
