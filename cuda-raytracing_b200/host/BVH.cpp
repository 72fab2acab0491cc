#include "BVH.h"
#include <algorithm>

namespace RayTracing
{
    void BVH::UpdateBounds(uint32_t node_index)  // RayTracing/BVH.cpp:45-57
    {
        RtBVHNode& node = nodes[node_index];
        float lo[3] = { 1e30f, 1e30f, 1e30f }, hi[3] = { -1e30f, -1e30f, -1e30f };
        for (uint32_t i = 0; i < node.prim_count; i++)
        {
            const RtFace& f = faces[triangles[node.first_index + i].index];
            const RtVec3* p[3] = { &vertices[f.v0].position, &vertices[f.v1].position, &vertices[f.v2].position };
            for (int k = 0; k < 3; k++)
            {
                lo[0] = std::min(lo[0], p[k]->x); lo[1] = std::min(lo[1], p[k]->y); lo[2] = std::min(lo[2], p[k]->z);
                hi[0] = std::max(hi[0], p[k]->x); hi[1] = std::max(hi[1], p[k]->y); hi[2] = std::max(hi[2], p[k]->z);
            }
        }
        node.bounds_min = RtVec3{ lo[0], lo[1], lo[2] };
        node.bounds_max = RtVec3{ hi[0], hi[1], hi[2] };
    }

    void BVH::Calculate(std::span<const RtVertex> vertices_, std::span<const RtFace> faces_)
    {
        vertices = vertices_; faces = faces_;
        nodes.assign(faces.empty() ? 1 : faces.size() * 2 - 1, RtBVHNode{});
        triangles.clear(); triangles.reserve(faces.size());
        for (uint32_t i = 0; i < faces.size(); i++)
        {   // centroid = ((v0 + v1) + v2) / 3, RayTracing/BVH.cpp:21
            const RtVec3& a = vertices[faces[i].v0].position; const RtVec3& b = vertices[faces[i].v1].position; const RtVec3& c = vertices[faces[i].v2].position;
            triangles.push_back(Centroid{ { ((a.x + b.x) + c.x) / 3.0f, ((a.y + b.y) + c.y) / 3.0f, ((a.z + b.z) + c.z) / 3.0f }, i });
        }
        nodes_used = 1;
        nodes[0].first_index = 0;
        nodes[0].prim_count = (uint32_t)faces.size();
        UpdateBounds(0);

        // Pre-order work stack. The reference numbers a node's two children when the node is
        // subdivided and recurses left first, so popping left before right reproduces its numbering.
        std::vector<uint32_t> work; work.push_back(0);
        while (!work.empty())
        {
            const uint32_t node_index = work.back(); work.pop_back();
            RtBVHNode& node = nodes[node_index];
            const float lo[3] = { node.bounds_min.x, node.bounds_min.y, node.bounds_min.z };
            const float extent[3] = { node.bounds_max.x - lo[0], node.bounds_max.y - lo[1], node.bounds_max.z - lo[2] };
            int axis1 = 0;                                  // RayTracing/BVH.cpp:64-72
            if (extent[1] > extent[0]) axis1 = 1;
            if (extent[2] > extent[axis1]) axis1 = 2;
            int axis2 = (axis1 + 1) % 3, axis3 = (axis2 + 1) % 3;
            if (extent[axis3] > extent[axis2]) std::swap(axis2, axis3);
            const int try_axis[3] = { axis1, axis2, axis3 };

            bool found = false; int i = 0, left_count = 0;
            for (int axis : try_axis)
            {
                const float split_pos = lo[axis] + extent[axis] * 0.5f;
                i = (int)node.first_index;
                int j = i + (int)node.prim_count - 1;
                while (i <= j)                              // RayTracing/BVH.cpp:85-91
                {
                    if (triangles[i].c[axis] < split_pos) i++;
                    else std::swap(triangles[i], triangles[j--]);
                }
                left_count = i - (int)node.first_index;
                if (left_count != 0 && left_count != (int)node.prim_count) { found = true; break; }
            }
            if (!found) continue;                           // leaf (possibly very large, BVH.cpp:104-107)

            const uint32_t left = nodes_used++, right = nodes_used++;
            nodes[left].first_index = node.first_index; nodes[left].prim_count = (uint32_t)left_count;
            nodes[right].first_index = (uint32_t)i;      nodes[right].prim_count = node.prim_count - (uint32_t)left_count;
            node.first_index = left; node.prim_count = 0;
            UpdateBounds(left); UpdateBounds(right);
            work.push_back(right); work.push_back(left);
        }
        face_indices.resize(faces.size());
        for (size_t k = 0; k < faces.size(); k++) face_indices[k] = triangles[k].index;
    }
}
