#include "BVH.h"
#include <algorithm>
#include <memory>
#include <numeric>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#ifdef _OPENMP
#include <omp.h>
#endif

// Parallel, topology-identical restatement of the reference builder.
//
// The reference (RayTracing/BVH.cpp:59-124) is a depth-first recursion whose node numbering depends on the sizes of the
// subtrees numbered earlier: Subdivide(n) takes the next two free indices for n's children (L = k, R = k + 1), then numbers
// ALL of L's descendants before any of R's. So the descendants of n occupy one contiguous block
//       block(n) = [L, R] ++ block(L) ++ block(R)
// and every index follows from subtree sizes alone. The build therefore runs in two phases:
//   1. level-synchronous construction of a temporary TOP tree (ids in breadth-first order) down to nodes of at most
//      kSubtree triangles. The splits of one level touch disjoint triangle ranges, so they run in parallel; each single split
//      is the reference's sequential Hoare partition (its permutation - including the one a failed axis leaves behind - must
//      be reproduced exactly), only the bounds of very large nodes are reduced in parallel (min/max is exact in any order);
//   2. every node the top tree stopped at is built depth-first by ONE thread (cache-resident, the sequential pre-order
//      work-stack builder) into a private node array in block-relative reference numbering; subtrees run in parallel;
//   3. subtree sizes bottom-up, then the block rule top-down, writing top nodes and subtree blocks at their reference indices.
namespace RayTracing
{
    namespace
    {
        constexpr uint32_t kBigNode = 1u << 19;  // nodes above this size get an inner-parallel bounds pass
        constexpr uint32_t kSubtree = 1u << 14;  // nodes at or below this size are finished depth-first by one thread
    }

    void BVH::Calculate(std::span<const RtVertex> vertices_, std::span<const RtFace> faces_)
    {
        vertices = vertices_; faces = faces_;
        const size_t n = faces.size();
        const bool timing = std::getenv("B200RT_BVH_TIMING") != nullptr;
        auto clock0 = std::chrono::steady_clock::now();
        auto lap = [&](const char* what) { if (!timing) return; auto now = std::chrono::steady_clock::now(); std::fprintf(stderr, "[bvh] %-22s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - clock0).count()); clock0 = now; };
        triangles.resize(n);
#pragma omp parallel for schedule(static) if (n > 100000)
        for (int64_t i = 0; i < (int64_t)n; i++)
        {   // centroid = ((v0 + v1) + v2) / 3, RayTracing/BVH.cpp:21
            const RtVec3& a = vertices[faces[i].v0].position; const RtVec3& b = vertices[faces[i].v1].position; const RtVec3& c = vertices[faces[i].v2].position;
            triangles[i] = Centroid{ { ((a.x + b.x) + c.x) / 3.0f, ((a.y + b.y) + c.y) / 3.0f, ((a.z + b.z) + c.z) / 3.0f }, (uint32_t)i,
                                     { std::min(std::min(a.x, b.x), c.x), std::min(std::min(a.y, b.y), c.y), std::min(std::min(a.z, b.z), c.z) },
                                     { std::max(std::max(a.x, b.x), c.x), std::max(std::max(a.y, b.y), c.y), std::max(std::max(a.z, b.z), c.z) } };
        }

        lap("centroids");
        struct Temp { float lo[3], hi[3]; uint32_t first, count, left; };  // left = temp id of the first child (right = left + 1), 0 = leaf
        std::vector<Temp> temp; temp.reserve(1024);

        // RayTracing/BVH.cpp:45-57 (min/max over the node's vertices; exact in any grouping). The per-triangle box travels
        // with the centroid: no vertex gathers here. Two forms: an OpenMP region costs about a microsecond even when it runs
        // on one thread, and there are as many small nodes as triangles.
        auto bounds_serial = [&](uint32_t first, uint32_t count, float* lo, float* hi)
        {
            float l0 = 1e30f, l1 = 1e30f, l2 = 1e30f, h0 = -1e30f, h1 = -1e30f, h2 = -1e30f;
            for (const Centroid* t = triangles.data() + first, *e = t + count; t != e; ++t)
            {
                l0 = std::min(l0, t->lo[0]); l1 = std::min(l1, t->lo[1]); l2 = std::min(l2, t->lo[2]);
                h0 = std::max(h0, t->hi[0]); h1 = std::max(h1, t->hi[1]); h2 = std::max(h2, t->hi[2]);
            }
            lo[0] = l0; lo[1] = l1; lo[2] = l2; hi[0] = h0; hi[1] = h1; hi[2] = h2;
        };
        auto bounds_parallel = [&](uint32_t first, uint32_t count, float* lo, float* hi)
        {
            float l0 = 1e30f, l1 = 1e30f, l2 = 1e30f, h0 = -1e30f, h1 = -1e30f, h2 = -1e30f;
#pragma omp parallel for schedule(static) reduction(min : l0, l1, l2) reduction(max : h0, h1, h2)
            for (int64_t i = 0; i < (int64_t)count; i++)
            {
                const Centroid& t = triangles[first + i];
                l0 = std::min(l0, t.lo[0]); l1 = std::min(l1, t.lo[1]); l2 = std::min(l2, t.lo[2]);
                h0 = std::max(h0, t.hi[0]); h1 = std::max(h1, t.hi[1]); h2 = std::max(h2, t.hi[2]);
            }
            lo[0] = l0; lo[1] = l1; lo[2] = l2; hi[0] = h0; hi[1] = h1; hi[2] = h2;
        };

        // the reference's split of one node (BVH.cpp:59-107); returns the left count, 0 when the node stays a leaf
        auto split = [&](const Temp& node) -> uint32_t
        {
            const float extent[3] = { node.hi[0] - node.lo[0], node.hi[1] - node.lo[1], node.hi[2] - node.lo[2] };
            int axis1 = 0;                                  // RayTracing/BVH.cpp:64-72
            if (extent[1] > extent[0]) axis1 = 1;
            if (extent[2] > extent[axis1]) axis1 = 2;
            int axis2 = (axis1 + 1) % 3, axis3 = (axis2 + 1) % 3;
            if (extent[axis3] > extent[axis2]) std::swap(axis2, axis3);
            const int try_axis[3] = { axis1, axis2, axis3 };
            for (int axis : try_axis)
            {
                const float split_pos = node.lo[axis] + extent[axis] * 0.5f;
                int64_t i = node.first, j = (int64_t)node.first + node.count - 1;
                while (i <= j)                              // RayTracing/BVH.cpp:85-91
                {
                    if (triangles[i].c[axis] < split_pos) i++;
                    else std::swap(triangles[i], triangles[j--]);
                }
                const uint32_t left_count = (uint32_t)(i - node.first);
                if (left_count != 0 && left_count != node.count) return left_count;
            }
            return 0;                                       // leaf (possibly very large, BVH.cpp:104-107)
        };

        Temp root{}; root.first = 0; root.count = (uint32_t)n; root.left = 0;
        if (n > kBigNode) bounds_parallel(0, (uint32_t)n, root.lo, root.hi); else bounds_serial(0, (uint32_t)n, root.lo, root.hi);
        temp.push_back(root);
        std::vector<uint32_t> frontier{ 0 }, next, left_counts;
        std::vector<uint32_t> subtree_roots;
        while (!frontier.empty() && n)
        {
            // nodes small enough leave the level-synchronous phase
            next.clear();
            for (uint32_t id : frontier) { if (temp[id].count <= kSubtree) subtree_roots.push_back(id); else next.push_back(id); }
            frontier.swap(next);
            const size_t m = frontier.size();
            left_counts.assign(m, 0);
            // (1) splits of this level: disjoint ranges, each one sequential
#pragma omp parallel for schedule(dynamic, 1) if (m > 1)
            for (int64_t k = 0; k < (int64_t)m; k++) left_counts[k] = split(temp[frontier[k]]);
            // (2) temp ids of the children (deterministic: frontier order)
            next.clear();
            for (size_t k = 0; k < m; k++)
            {
                if (!left_counts[k]) continue;
                Temp& node = temp[frontier[k]];
                const uint32_t id = (uint32_t)temp.size();
                Temp l{}, r{};
                l.first = node.first; l.count = left_counts[k];
                r.first = node.first + left_counts[k]; r.count = node.count - left_counts[k];
                node.left = id;
                temp.push_back(l); temp.push_back(r);
                next.push_back(id); next.push_back(id + 1);
            }
            // (3) bounds of the new nodes: big ones one at a time with an inner-parallel reduction, the rest in parallel
            for (uint32_t id : next)
                if (temp[id].count > kBigNode) bounds_parallel(temp[id].first, temp[id].count, temp[id].lo, temp[id].hi);
#pragma omp parallel for schedule(dynamic, 64) if (next.size() > 64)
            for (int64_t k = 0; k < (int64_t)next.size(); k++)
            {
                Temp& t = temp[next[k]];
                if (t.count <= kBigNode) bounds_serial(t.first, t.count, t.lo, t.hi);
            }
            frontier.swap(next);
        }

        lap("top tree");
        // phase 2: depth-first subtrees, one thread each. local[0] is the subtree root, local[1..] its descendants in the
        // reference's order relative to the root's block; child links are local indices.
        // Blocks live in per-thread chunked arenas (one allocation per ~32 MB, not per subtree: concurrent large malloc/free
        // pairs serialise on the process's address-space lock).
        struct Block { const RtBVHNode* nodes = nullptr; uint32_t size = 0; };
        std::vector<Block> local(subtree_roots.size());
        int arena_count = 1;
#ifdef _OPENMP
        arena_count = omp_get_max_threads();
#endif
        std::vector<std::vector<std::unique_ptr<RtBVHNode[]>>> arenas(arena_count);
        constexpr size_t kChunkNodes = 1u << 20;
#pragma omp parallel if (subtree_roots.size() > 1)
        {
            int me = 0;
#ifdef _OPENMP
            me = omp_get_thread_num();
#endif
            auto& chunks = arenas[me];
            size_t chunk_used = kChunkNodes;
            std::vector<RtBVHNode> scratch; std::vector<uint32_t> work;
#pragma omp for schedule(dynamic, 1)
            for (int64_t s = 0; s < (int64_t)subtree_roots.size(); s++)
            {
                const Temp& r = temp[subtree_roots[s]];
                scratch.assign(r.count ? 2 * (size_t)r.count - 1 : 1, RtBVHNode{});
                scratch[0].bounds_min = RtVec3{ r.lo[0], r.lo[1], r.lo[2] }; scratch[0].bounds_max = RtVec3{ r.hi[0], r.hi[1], r.hi[2] };
                scratch[0].first_index = r.first; scratch[0].prim_count = r.count;
                uint32_t used = 1;
                work.clear(); work.push_back(0);
                while (!work.empty())
                {   // pre-order: the reference numbers both children at subdivision and recurses left first
                    const uint32_t node_index = work.back(); work.pop_back();
                    RtBVHNode& node = scratch[node_index];
                    Temp t{ { node.bounds_min.x, node.bounds_min.y, node.bounds_min.z }, { node.bounds_max.x, node.bounds_max.y, node.bounds_max.z }, node.first_index, node.prim_count, 0 };
                    const uint32_t left_count = split(t);
                    if (!left_count) continue;
                    const uint32_t left = used++, right = used++;
                    float lo[3], hi[3];
                    bounds_serial(t.first, left_count, lo, hi);
                    scratch[left].bounds_min = RtVec3{ lo[0], lo[1], lo[2] }; scratch[left].bounds_max = RtVec3{ hi[0], hi[1], hi[2] };
                    scratch[left].first_index = t.first; scratch[left].prim_count = left_count;
                    bounds_serial(t.first + left_count, t.count - left_count, lo, hi);
                    scratch[right].bounds_min = RtVec3{ lo[0], lo[1], lo[2] }; scratch[right].bounds_max = RtVec3{ hi[0], hi[1], hi[2] };
                    scratch[right].first_index = t.first + left_count; scratch[right].prim_count = t.count - left_count;
                    node.first_index = left; node.prim_count = 0;
                    work.push_back(right); work.push_back(left);
                }
                RtBVHNode* dst;
                if (used > kChunkNodes) { chunks.emplace_back(new RtBVHNode[used]); dst = chunks.back().get(); chunk_used = kChunkNodes; }
                else
                {
                    if (chunk_used + used > kChunkNodes) { chunks.emplace_back(new RtBVHNode[kChunkNodes]); chunk_used = 0; }
                    dst = chunks.back().get() + chunk_used;
                    chunk_used += used;
                }
                std::copy(scratch.begin(), scratch.begin() + used, dst);
                local[s] = Block{ dst, used };
            }
        }

        lap("subtrees");
        // phase 3: reference numbering from subtree sizes
        const size_t total = temp.size();
        std::vector<uint32_t> desc(total, 0), final_index(total, 0), block_base(total, 0);
        for (size_t s = 0; s < subtree_roots.size(); s++) desc[subtree_roots[s]] = local[s].size - 1;
        for (size_t t = total; t-- > 0;)
            if (temp[t].left) desc[t] = 2 + desc[temp[t].left] + desc[temp[t].left + 1];
        final_index[0] = 0; block_base[0] = 1;
        for (size_t t = 0; t < total; t++)
        {
            if (!temp[t].left) continue;
            const uint32_t l = temp[t].left, r = l + 1, base = block_base[t];
            final_index[l] = base; final_index[r] = base + 1;
            block_base[l] = base + 2; block_base[r] = base + 2 + desc[l];
        }
        nodes_used = 1 + desc[0];
        nodes.assign(nodes_used, RtBVHNode{});
        for (size_t t = 0; t < total; t++)
        {   // top-tree nodes: inner ones, and leaves too large for a subtree that no axis could split (subtree roots are
            // overwritten with their blocks below)
            RtBVHNode& out = nodes[final_index[t]];
            out.bounds_min = RtVec3{ temp[t].lo[0], temp[t].lo[1], temp[t].lo[2] };
            out.bounds_max = RtVec3{ temp[t].hi[0], temp[t].hi[1], temp[t].hi[2] };
            if (temp[t].left) { out.first_index = final_index[temp[t].left]; out.prim_count = 0; }
            else { out.first_index = temp[t].first; out.prim_count = temp[t].count; }
        }
#pragma omp parallel for schedule(dynamic, 1) if (subtree_roots.size() > 1)
        for (int64_t s = 0; s < (int64_t)subtree_roots.size(); s++)
        {
            const uint32_t t = subtree_roots[s], base = block_base[t];  // local i >= 1 -> base + i - 1
            const Block& src = local[s];
            for (size_t i = 0; i < src.size; i++)
            {
                RtBVHNode out = src.nodes[i];
                if (out.prim_count == 0) out.first_index = base + out.first_index - 1;
                nodes[i ? base + (uint32_t)i - 1 : final_index[t]] = out;
            }
        }
        lap("numbering + copy");
        face_indices.resize(n);
#pragma omp parallel for schedule(static) if (n > 100000)
        for (int64_t k = 0; k < (int64_t)n; k++) face_indices[k] = triangles[k].index;
        lap("face indices");
    }
}
