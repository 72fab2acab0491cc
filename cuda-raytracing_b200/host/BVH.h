// CPU BVH builder producing the reference's topology: same node numbering, same leaf contents,
// same face-index permutation as RayTracing/BVH.cpp:8-124 (spatial-midpoint split on the node
// bounds, longest axis first with the two-axis fallback, Hoare-style partition whose permutation
// survives a failed axis, children numbered consecutively, left subtree numbered before right,
// no leaf-size bound). Built in parallel (OpenMP): level-synchronous top tree, depth-first
// subtrees one thread each, then the reference's node numbering from subtree sizes - see BVH.cpp.
#pragma once
#include <cstdint>
#include <span>
#include <vector>
#include "rt_types.h"

namespace RayTracing
{
    class BVH
    {
    public:
        // RayTracing/BVH.h:38
        void Calculate(std::span<const RtVertex> vertices, std::span<const RtFace> faces);
        std::span<const RtBVHNode> GetGPUBVHNodes() const { return std::span<const RtBVHNode>(nodes.data(), nodes_used); }
        std::span<const uint32_t> GetFaceIndices() const { return face_indices; }

    private:
        struct Centroid { float c[3]; uint32_t index; float lo[3], hi[3]; };  // centroid, face id, the triangle's own box

        std::vector<RtBVHNode> nodes;
        std::vector<Centroid> triangles;
        std::vector<uint32_t> face_indices;
        std::span<const RtVertex> vertices;
        std::span<const RtFace> faces;
        uint32_t nodes_used = 0;
    };
}
