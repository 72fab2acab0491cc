// Mesh ingest for the path-trace hot path: a small Wavefront-OBJ reader plus a smooth-normal pass.
//
// Stands in for utils/AssimpLoader.cpp:29-51 (Assimp 5.2.4, vcpkg.json:6 - source not in the
// reference tree). What is reproduced from the call site's flags (Triangulate | JoinIdenticalVertices
// | GenSmoothNormals | SortByPType, AI_CONFIG_PP_GSN_MAX_SMOOTHING_ANGLE = 100):
//   * face k of the result is the k-th `f` line (fan-triangulated) - hit IDs hang on this;
//   * positions are the file's values; * per-corner normals = normalised sum of the unit face
//     normals around that position lying within 100 degrees of the corner's own face normal;
//   * corners with equal (position, normal) are joined.
// PARITY UNPINNED for normals/vertex numbering (no Assimp here; the reference has no test at this
// boundary) - confined to shading normals, never to hit IDs or t. See DESIGN.md.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "Math.h"

namespace Loader
{
    struct LoadedMesh
    {
        std::vector<float> positions;   // 3 per vertex
        std::vector<float> normals;     // 3 per vertex
        std::vector<uint32_t> indices;  // 3 per face
        Math::mat4 transform;           // node transform (root pre-rotation, utils/AssimpLoader.cpp:46-48)
        uint32_t VertexCount() const { return (uint32_t)(positions.size() / 3); }
        uint32_t FaceCount() const { return (uint32_t)(indices.size() / 3); }
    };

    // utils/AssimpLoader.cpp:46-48,10-13 : RotationX(-(float)M_PI/2) as a column-major matrix
    Math::mat4 ImportRootTransform();

    // Raw OBJ: positions + triangulated 0-based position indices. Throws std::runtime_error.
    void ReadObj(const std::string& path, std::vector<float>& positions, std::vector<uint32_t>& indices);

    // GenSmoothNormals + JoinIdenticalVertices equivalent.
    LoadedMesh BuildSmoothMesh(const std::vector<float>& positions, const std::vector<uint32_t>& indices, float max_angle_deg = 100.0f);

    // One round of Loop subdivision on raw OBJ data (scaled-mesh config generator; Subdivide.cpp)
    void SubdivideLoop(std::vector<float>& positions, std::vector<uint32_t>& indices);

    // utils/AssimpLoader.cpp:29 ImportScene
    LoadedMesh ImportScene(const std::string& path);
}
