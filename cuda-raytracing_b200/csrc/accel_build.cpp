// Host re-layout of the reference BVH into WideNode / TriRecord / ShadeRecord arrays (accel.h).
// Compiled with -ffp-contract=off: e1/e2 must equal the fp32 differences the reference computes.
// Builder switches read from the environment at every build (A/B and diagnosis; the defaults are the product):
//   B200RT_TOPOLOGY=ref      keep BVH.cpp's midpoint-split topology instead of the SAH tree over its leaves
//   B200RT_REINSERT=<n>      insertion-based optimisation passes over the SAH tree (default 1; 0 = none)
//   B200RT_REINSERT_MAX=<n>  at most the n largest nodes per pass (default: all up to 2^18 leaves, else 4 096)
//   B200RT_BVH_TIMING=1      per-phase build times and the tree's surface-area cost on stderr
#include "accel.h"
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include "b200rt.h"

namespace
{
    struct Box { float lo[3], hi[3]; };

    inline Box EmptyBox() { return Box{ { 1e30f, 1e30f, 1e30f }, { -1e30f, -1e30f, -1e30f } }; }
    inline void Grow(Box& b, const float* p) { for (int k = 0; k < 3; k++) { b.lo[k] = std::min(b.lo[k], p[k]); b.hi[k] = std::max(b.hi[k], p[k]); } }
    inline void Grow(Box& b, const Box& o) { for (int k = 0; k < 3; k++) { b.lo[k] = std::min(b.lo[k], o.lo[k]); b.hi[k] = std::max(b.hi[k], o.hi[k]); } }
    inline float HalfArea(const Box& b)
    {
        const float x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
        return x * y + y * z + z * x;
    }
    inline Box FromNode(const RtBVHNode& n) { return Box{ { n.bounds_min.x, n.bounds_min.y, n.bounds_min.z }, { n.bounds_max.x, n.bounds_max.y, n.bounds_max.z } }; }

    // Sub-tree boxes have no reference counterpart: they only have to be supersets. Inflate generously so
    // that any ray the (rounding-afflicted) triangle test accepts also passes the box (DESIGN.md "fat leaves").
    inline Box Inflate(const Box& b)
    {
        float ext = 0, mag = 0;
        for (int k = 0; k < 3; k++) { ext = std::max(ext, b.hi[k] - b.lo[k]); mag = std::max(mag, std::max(std::fabs(b.lo[k]), std::fabs(b.hi[k]))); }
        const float pad = 0.01f * ext + 1e-4f * mag + 1e-6f;
        Box r = b;
        for (int k = 0; k < 3; k++) { r.lo[k] -= pad; r.hi[k] += pad; }
        return r;
    }

    inline uint32_t LeafRef(uint32_t first, uint32_t count) { return RT_LEAF_BIT | ((count - 1u) << 27) | first; }

    // centre and half extent of child `which` exactly as the kernels read them (traverse.cuh center_slabs)
    inline void ChildCentreHalf(const WideNode& w, int which, float ctr[3], float half[3])
    {
        const RtFloat4& q = which == 0 ? w.a : w.b;
        ctr[0] = q.x; ctr[1] = q.y; ctr[2] = which == 0 ? w.c.x : w.c.y;
        half[0] = q.z; half[1] = q.w; half[2] = which == 0 ? w.c.z : w.c.w;
    }

    inline void SetChildBox(WideNode& w, int which, const Box& b)
    {
        float c[3], h[3];
        for (int k = 0; k < 3; k++)
        {   // [c - h, c + h] in exact arithmetic holds [lo, hi]; h carries the 2^-21 the kernel's rounding of h * |1/d| needs
            const double lo = b.lo[k], hi = b.hi[k];
            if (!(lo <= hi)) { c[k] = 0.0f; h[k] = -1e30f; continue; }  // empty box: exit far before entry, never hit
            // a bound at infinity (a vertex at infinity) or beyond 1e30: the axis does not constrain - centre 0, half extent 1e30, so that
            // mid = -o / d stays finite and entry / exit go to -+ huge (a huge CENTRE would overflow mid and turn the test into NaN = miss)
            if (!(lo > -1e30 && hi < 1e30)) { c[k] = 0.0f; h[k] = 1e30f; continue; }
            c[k] = (float)(0.5 * (lo + hi));
            const double need = std::max(hi - (double)c[k], (double)c[k] - lo) * (1.0 + 4.8e-7) + 1e-37;
            float hf = (float)need;
            if ((double)hf < need) hf = std::nextafter(hf, INFINITY);
            h[k] = hf;
        }
        if (which == 0) { w.a = { c[0], c[1], h[0], h[1] }; w.c.x = c[2]; w.c.z = h[2]; }
        else { w.b = { c[0], c[1], h[0], h[1] }; w.c.y = c[2]; w.c.w = h[2]; }
    }

    struct Builder
    {
        const RtBVHNode* nodes; uint32_t node_count; uint32_t threshold;
        AccelHost& out;
        std::vector<Box> slot_box;      // per slot, for sub-trees
        std::vector<float> slot_centroid;
        int error = 0;

        Box SlotBox(uint32_t s) const
        {
            const TriRecord& t = out.tris[s];
            Box b = EmptyBox();
            float v1[3], v2[3];
            for (int k = 0; k < 3; k++) { v1[k] = t.v0[k] + t.e1[k]; v2[k] = t.v0[k] + t.e2[k]; }
            Grow(b, t.v0); Grow(b, v1); Grow(b, v2);
            // v0+e1 may differ from v1 by an ulp; the sub-tree boxes are inflated far beyond that
            return b;
        }

        // SAH sweep sub-tree over slots [first, first+count); permutes tris/shade inside the range.
        uint32_t BuildSub(uint32_t first, uint32_t count, uint32_t depth, Box& bounds_out)
        {
            Box bounds = EmptyBox();
            for (uint32_t s = first; s < first + count; s++) Grow(bounds, SlotBox(s));
            bounds_out = bounds;
            if (count <= 4) return LeafRef(first, count);

            std::vector<uint32_t> order(count);
            std::vector<Box> boxes(count);
            std::vector<float> right_area(count);
            float best_cost = 3.4e38f; int best_axis = -1; uint32_t best_split = 0;
            for (int axis = 0; axis < 3; axis++)
            {
                for (uint32_t i = 0; i < count; i++) { order[i] = first + i; }
                std::vector<float> cen(count);
                for (uint32_t i = 0; i < count; i++) { const Box b = SlotBox(first + i); cen[i] = b.lo[axis] + b.hi[axis]; }
                std::vector<uint32_t> idx(count);
                for (uint32_t i = 0; i < count; i++) idx[i] = i;
                std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return cen[a] < cen[b]; });
                Box acc = EmptyBox();
                for (uint32_t i = count; i-- > 0;) { Grow(acc, SlotBox(first + idx[i])); right_area[i] = HalfArea(acc); }
                acc = EmptyBox();
                for (uint32_t i = 0; i + 1 < count; i++)
                {
                    Grow(acc, SlotBox(first + idx[i]));
                    const float cost = HalfArea(acc) * float(i + 1) + right_area[i + 1] * float(count - i - 1);
                    if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = i + 1; }
                }
            }
            // apply the best permutation (no finite cost on any axis - NaN / infinite vertices: halve the range as it stands)
            if (best_axis < 0 || best_split == 0 || best_split >= count) best_split = count / 2;
            else
            {
                std::vector<float> cen(count);
                for (uint32_t i = 0; i < count; i++) { const Box b = SlotBox(first + i); cen[i] = b.lo[best_axis] + b.hi[best_axis]; }
                std::vector<uint32_t> idx(count);
                for (uint32_t i = 0; i < count; i++) idx[i] = i;
                std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return cen[a] < cen[b]; });
                std::vector<TriRecord> t(count); std::vector<ShadeRecord> sh(count);
                for (uint32_t i = 0; i < count; i++) { t[i] = out.tris[first + idx[i]]; sh[i] = out.shade[first + idx[i]]; }
                std::copy(t.begin(), t.end(), out.tris.begin() + first);
                std::copy(sh.begin(), sh.end(), out.shade.begin() + first);
            }
            const uint32_t w = (uint32_t)out.wide.size();
            out.wide.push_back(WideNode{});
            out.wide_depth = std::max(out.wide_depth, depth + 1);
            Box b0, b1;
            const uint32_t c0 = BuildSub(first, best_split, depth + 1, b0);
            const uint32_t c1 = BuildSub(first + best_split, count - best_split, depth + 1, b1);
            WideNode& node = out.wide[w];
            SetChildBox(node, 0, Inflate(b0)); SetChildBox(node, 1, Inflate(b1));
            node.child[0] = c0; node.child[1] = c1;
            return w;
        }

        // Child reference for reference node `n` (whose exact box the PARENT stores).
        uint32_t Emit(uint32_t n, uint32_t depth)
        {
            if (n >= node_count) { error = RT_E_INVALID; return RT_SENTINEL; }
            const RtBVHNode& node = nodes[n];
            out.ref_depth = std::max(out.ref_depth, depth);
            if (node.prim_count > 0)
            {
                out.ref_leaf_count++;
                if ((uint64_t)node.first_index + node.prim_count > out.tris.size()) { error = RT_E_INVALID; return RT_SENTINEL; }
                for (uint32_t s = node.first_index; s < node.first_index + node.prim_count; s++) out.tris[s].ref_leaf = n;
                if (node.prim_count <= threshold) return LeafRef(node.first_index, node.prim_count);
                out.fat_leaf_count++;
                Box unused;
                return BuildSub(node.first_index, node.prim_count, depth, unused);
            }
            if (depth > 4096) { error = RT_E_DEPTH; return RT_SENTINEL; }
            const uint32_t w = (uint32_t)out.wide.size();
            out.wide.push_back(WideNode{});
            out.wide_depth = std::max(out.wide_depth, depth + 1);
            const uint32_t l = node.first_index, r = node.first_index + 1;
            if (r >= node_count) { error = RT_E_INVALID; return RT_SENTINEL; }
            const uint32_t c0 = Emit(l, depth + 1);
            const uint32_t c1 = Emit(r, depth + 1);
            WideNode& wn = out.wide[w];
            SetChildBox(wn, 0, FromNode(nodes[l])); SetChildBox(wn, 1, FromNode(nodes[r]));  // exact reference boxes
            wn.child[0] = c0; wn.child[1] = c1;
            return w;
        }

        // ---- own topology over the reference's leaves -------------------------------------------------------------------
        // The hit rule (smallest t, ties to the lowest reference visiting order) and the deferred exact reference-leaf box
        // check (traverse.cuh) make the result a function of WHICH triangles a ray is tested against, not of the tree above
        // them: any binary tree whose boxes hold the exact reference leaf boxes below them returns the reference's hits.
        // BVH.cpp:61-107 splits at the spatial midpoint of the longest axis; this builds a surface-area-heuristic tree over the
        // SAME leaves (slots, leaf contents and TriRecords untouched: the reference-order kernels keep reading them as they are).
        struct Item { Box box; float cen[3]; uint32_t ref; float weight; };
        std::vector<Item> items;
        uint32_t fat_depth = 0;

        // reference leaves in visiting order; fat ones get their private sub-tree (appended to out.wide) first
        // One walk of the reference tree: its leaves in visiting order (validated), its depth.
        std::vector<uint32_t> leaf_nodes;
        void WalkReference()
        {
            std::vector<std::pair<uint32_t, uint32_t>> stack; stack.push_back({ 0u, 0u });
            uint64_t visited = 0;
            while (!stack.empty())
            {
                const auto [n, depth] = stack.back(); stack.pop_back();
                if (n >= node_count || ++visited > node_count) { error = RT_E_INVALID; return; }
                const RtBVHNode& node = nodes[n];
                out.ref_depth = std::max(out.ref_depth, depth);
                if (node.prim_count == 0)
                {
                    if (depth > 4096) { error = RT_E_DEPTH; return; }
                    stack.push_back({ node.first_index + 1, depth + 1 }); stack.push_back({ node.first_index, depth + 1 });
                    continue;
                }
                if ((uint64_t)node.first_index + node.prim_count > out.tris.size()) { error = RT_E_INVALID; return; }
                leaf_nodes.push_back(n);
            }
        }

        // items[] from leaf_nodes[]; fat leaves get their private sub-tree (appended to out.wide, which already holds the own tree's nodes)
        void CollectLeaves()
        {
            const int64_t count = (int64_t)leaf_nodes.size();
            items.resize(leaf_nodes.size());
            out.ref_leaf_count = (uint32_t)leaf_nodes.size();
#pragma omp parallel for schedule(static) if (count > 100000)
            for (int64_t i = 0; i < count; i++)
            {
                const uint32_t n = leaf_nodes[i];
                const RtBVHNode& node = nodes[n];
                for (uint32_t s = node.first_index; s < node.first_index + node.prim_count; s++) out.tris[s].ref_leaf = n;
                Item& it = items[i];
                it.box = FromNode(node);
                for (int k = 0; k < 3; k++) { it.cen[k] = 0.5f * it.box.lo[k] + 0.5f * it.box.hi[k]; if (!std::isfinite(it.cen[k])) it.cen[k] = 0.0f; }  // (sort keys must be ordered)
                it.ref = LeafRef(node.first_index, std::min(node.prim_count, threshold)); it.weight = (float)node.prim_count;
            }
            for (int64_t i = 0; i < count; i++)
            {
                const RtBVHNode& node = nodes[leaf_nodes[i]];
                if (node.prim_count <= threshold) continue;
                out.fat_leaf_count++;
                const uint32_t before = out.wide_depth;
                out.wide_depth = 0;
                Box unused;
                items[i].ref = BuildSub(node.first_index, node.prim_count, 0, unused);
                fat_depth = std::max(fat_depth, out.wide_depth);
                out.wide_depth = before;
                items[i].weight = 2.0f * std::log2((float)node.prim_count) + 4.0f;  // expected steps inside the sub-tree
            }
        }

        static uint32_t CeilLog2(uint32_t n) { uint32_t d = 0; while ((1ull << d) < n) d++; return d; }

        // Partitions items[lo, hi) (hi - lo >= 2) and returns the split position m (lo < m < hi).
        uint32_t Split(uint32_t lo, uint32_t hi, bool balanced)
        {
            const uint32_t n = hi - lo;
            Item* it = items.data() + lo;
            Box cb = EmptyBox();
            for (uint32_t i = 0; i < n; i++) Grow(cb, it[i].cen);
            int widest = 0;
            for (int k = 1; k < 3; k++) if (cb.hi[k] - cb.lo[k] > cb.hi[widest] - cb.lo[widest]) widest = k;
            auto median = [&]() -> uint32_t
            {
                if (cb.hi[widest] > cb.lo[widest])
                    std::nth_element(it, it + n / 2, it + n, [&](const Item& a, const Item& b) { return a.cen[widest] < b.cen[widest]; });
                return lo + n / 2;
            };
            if (balanced || n == 2) return median();
            if (n <= 1024)   // (binned from 8, 32 or 128 items on gives the same cost to three digits and the same build time)
            {   // full sweep on the three axes
                float best_cost = 3.4e38f; int best_axis = -1, sorted_axis = -1; uint32_t best_split = 0;
                std::vector<float> right_area(n), right_weight(n);
                for (int axis = 0; axis < 3; axis++)
                {
                    if (!(cb.hi[axis] > cb.lo[axis])) continue;
                    std::stable_sort(it, it + n, [&](const Item& a, const Item& b) { return a.cen[axis] < b.cen[axis]; });
                    sorted_axis = axis;
                    Box acc = EmptyBox(); float w = 0;
                    for (uint32_t i = n; i-- > 0;) { Grow(acc, it[i].box); w += it[i].weight; right_area[i] = HalfArea(acc); right_weight[i] = w; }
                    acc = EmptyBox(); w = 0;
                    for (uint32_t i = 0; i + 1 < n; i++)
                    {
                        Grow(acc, it[i].box); w += it[i].weight;
                        const float cost = HalfArea(acc) * w + right_area[i + 1] * right_weight[i + 1];
                        if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = i + 1; }
                    }
                }
                if (best_axis < 0) return median();
                if (best_axis != sorted_axis) std::stable_sort(it, it + n, [&](const Item& a, const Item& b) { return a.cen[best_axis] < b.cen[best_axis]; });
                return lo + best_split;
            }
            // binned: one pass over the items fills the bins of all three axes; large ranges in chunks that the team's idle threads take
            // (the top levels of a big tree would otherwise run on one thread while the others wait for tasks)
            constexpr int kBins = 64;
            struct Bins { Box box[3][kBins]; float weight[3][kBins]; uint32_t count[3][kBins]; };
            auto clear = [](Bins& b) { for (int a = 0; a < 3; a++) for (int k = 0; k < kBins; k++) { b.box[a][k] = EmptyBox(); b.weight[a][k] = 0; b.count[a][k] = 0; } };
            float scale3[3];
            for (int a = 0; a < 3; a++) { const float extent = cb.hi[a] - cb.lo[a]; scale3[a] = extent > 0 ? (float)kBins / extent : 0.0f; }
            auto fill = [&](Bins& b, uint32_t from, uint32_t to)
            {
                for (uint32_t i = from; i < to; i++)
                    for (int a = 0; a < 3; a++)
                    {
                        const int k = std::min(kBins - 1, std::max(0, (int)((it[i].cen[a] - cb.lo[a]) * scale3[a])));
                        Grow(b.box[a][k], it[i].box); b.weight[a][k] += it[i].weight; b.count[a][k]++;
                    }
            };
            std::unique_ptr<Bins> bins(new Bins); clear(*bins);
            if (n < (1u << 18)) fill(*bins, 0, n);
            else
            {
                const int chunks = 64;
                std::vector<Bins> part(chunks);
#pragma omp taskloop default(shared) grainsize(1)
                for (int c = 0; c < chunks; c++) { clear(part[c]); fill(part[c], (uint32_t)((uint64_t)n * c / chunks), (uint32_t)((uint64_t)n * (c + 1) / chunks)); }
                for (int c = 0; c < chunks; c++)
                    for (int a = 0; a < 3; a++) for (int k = 0; k < kBins; k++)
                        if (part[c].count[a][k]) { Grow(bins->box[a][k], part[c].box[a][k]); bins->weight[a][k] += part[c].weight[a][k]; bins->count[a][k] += part[c].count[a][k]; }
            }
            float best_cost = 3.4e38f; int best_axis = -1, best_plane = 0;
            for (int axis = 0; axis < 3; axis++)
            {
                if (!(cb.hi[axis] - cb.lo[axis] > 0)) continue;
                const Box* bin_box = bins->box[axis]; const float* bin_weight = bins->weight[axis]; const uint32_t* bin_count = bins->count[axis];
                float right_area[kBins], right_weight[kBins]; uint32_t right_count[kBins];
                Box acc = EmptyBox(); float w = 0; uint32_t c = 0;
                for (int b = kBins; b-- > 0;) { if (bin_count[b]) Grow(acc, bin_box[b]); w += bin_weight[b]; c += bin_count[b]; right_area[b] = HalfArea(acc); right_weight[b] = w; right_count[b] = c; }
                acc = EmptyBox(); w = 0; c = 0;
                for (int b = 0; b + 1 < kBins; b++)
                {
                    if (bin_count[b]) Grow(acc, bin_box[b]);
                    w += bin_weight[b]; c += bin_count[b];
                    if (c == 0 || right_count[b + 1] == 0) continue;
                    const float cost = HalfArea(acc) * w + right_area[b + 1] * right_weight[b + 1];
                    if (cost < best_cost) { best_cost = cost; best_axis = axis; best_plane = b; }
                }
            }
            if (best_axis < 0) return median();
            const float scale = (float)kBins / (cb.hi[best_axis] - cb.lo[best_axis]);
            const float origin = cb.lo[best_axis];
            Item* mid = std::partition(it, it + n, [&](const Item& a)
            { return std::min(kBins - 1, std::max(0, (int)((a.cen[best_axis] - origin) * scale))) <= best_plane; });
            const uint32_t m = lo + (uint32_t)(mid - it);
            return (m == lo || m == hi) ? median() : m;
        }

        // Working tree: tree[0 .. leaves-1) inner nodes (children: inner index, or kItem | item index), boxes = exact unions.
        static constexpr uint32_t kItem = 0x80000000u, kNone = 0xFFFFFFFFu;
        struct TNode { Box box; uint32_t child[2]; uint32_t parent; };
        std::vector<TNode> tree;
        std::vector<uint32_t> item_parent;
        uint32_t tree_root = 0;

        const Box& BoxOf(uint32_t ref) const { return (ref & kItem) ? items[ref & ~kItem].box : tree[ref].box; }
        uint32_t& ParentOf(uint32_t ref) { return (ref & kItem) ? item_parent[ref & ~kItem] : tree[ref].parent; }
        static Box Union(const Box& a, const Box& b) { Box r = a; Grow(r, b); return r; }

        // Sub-tree over items[lo, hi) (>= 2 items) rooted at tree[node]; it occupies tree[node, node + (hi - lo) - 1) in pre-order, so
        // the index of every node follows from the split positions alone (tasks need no shared counter).
        void BuildOwn(uint32_t lo, uint32_t hi, uint32_t node, uint32_t depth, uint32_t depth_limit)
        {
            const uint32_t n = hi - lo;
            const bool balanced = depth + 1 + CeilLog2(n) >= depth_limit;  // SAH may chain: from here on halve by count
            const uint32_t m = Split(lo, hi, balanced);
            const uint32_t left = node + 1, right = node + (m - lo);
            TNode& t = tree[node];
            t.child[0] = (m - lo == 1) ? (kItem | lo) : left;
            t.child[1] = (hi - m == 1) ? (kItem | m) : right;
            if (m - lo >= 2)
            {
                tree[left].parent = node;
#pragma omp task default(shared) firstprivate(lo, m, left, depth, depth_limit) if (m - lo > 20000)
                BuildOwn(lo, m, left, depth + 1, depth_limit);
            }
            else item_parent[lo] = node;
            if (hi - m >= 2)
            {
                tree[right].parent = node;
#pragma omp task default(shared) firstprivate(m, hi, right, depth, depth_limit) if (hi - m > 20000)
                BuildOwn(m, hi, right, depth + 1, depth_limit);
            }
            else item_parent[m] = node;
        }

        void Refit(uint32_t node)
        {
            while (node != kNone)
            {
                TNode& t = tree[node];
                const Box b = Union(BoxOf(t.child[0]), BoxOf(t.child[1]));
                if (!std::memcmp(&b, &t.box, sizeof(Box))) break;
                t.box = b;
                node = t.parent;
            }
        }

        uint32_t TreeDepth() const
        {
            uint32_t deepest = 0;
            std::vector<std::pair<uint32_t, uint32_t>> stack; stack.push_back({ tree_root, 1u });
            while (!stack.empty())
            {
                const auto [ref, depth] = stack.back(); stack.pop_back();
                deepest = std::max(deepest, depth);
                for (int c = 0; c < 2; c++) if (!(tree[ref].child[c] & kItem)) stack.push_back({ tree[ref].child[c], depth + 1 });
            }
            return deepest;
        }

        double TreeCost() const
        {
            double cost = 0;
            for (size_t i = 0; i < tree.size(); i++) cost += HalfArea(tree[i].box);
            return cost;
        }

        // Insertion-based optimisation (after Bittner, Hapala, Havran 2013): take a sub-tree out (its parent is replaced by its
        // sibling), find by branch and bound the position where putting it back adds the least surface area to the tree, put it there.
        // Topology only: leaves (= reference leaves) and their boxes stay what they are.
        void Optimize(int passes, uint32_t depth_limit)
        {
            struct Entry { float induced; uint32_t ref; bool operator<(const Entry& o) const { return induced > o.induced; } };
            std::vector<Entry> heap;
            std::vector<std::pair<float, uint32_t>> order;
            for (int pass = 0; pass < passes; pass++)
            {
                uint64_t visits = 0;
                order.clear();
                for (uint32_t i = 0; i < tree.size(); i++) if (i != tree_root && tree[i].parent != tree_root) order.push_back({ HalfArea(tree[i].box), i });
                for (uint32_t i = 0; i < items.size(); i++) if (item_parent[i] != tree_root) order.push_back({ HalfArea(items[i].box), kItem | i });
                auto larger = [](const std::pair<float, uint32_t>& a, const std::pair<float, uint32_t>& b) { return a.first > b.first || (a.first == b.first && a.second < b.second); };
                // every node of a tree of up to 2^18 leaves (0.4 s on the bunny scene's 69 376); beyond that the 4 096 largest - nearly all of
                // the gain sits there (walls, floors, the boxes of whole objects: bunny scene 6.43 -> 5.63 expected node steps from the 100
                // largest alone, 5.61 from all), and the search per node does not grow with the scene
                const size_t most = std::getenv("B200RT_REINSERT_MAX") ? (size_t)std::atoll(std::getenv("B200RT_REINSERT_MAX")) : (items.size() <= (1u << 18) ? (size_t)-1 : (size_t)4096);
                if (order.size() > most) { std::nth_element(order.begin(), order.begin() + most, order.end(), larger); order.resize(most); }
                std::sort(order.begin(), order.end(), larger);
                // the branch-and-bound search prunes well in ordinary scenes (bunny scene: 36 nodes looked at per reinsertion) and badly where
                // everything overlaps everything; largest nodes first, so running out of budget drops the least valuable ones
                const uint64_t budget = 256ull * order.size() + (1ull << 20);
                for (const auto& entry : order)
                {
                    if (visits > budget) break;
                    const uint32_t n = entry.second;
                    const uint32_t p = ParentOf(n);
                    if (p == tree_root) continue;                     // (the tree changed since the list was made)
                    const uint32_t g = tree[p].parent;
                    const int side = tree[p].child[0] == n ? 0 : 1;
                    const uint32_t sibling = tree[p].child[side ^ 1];
                    // take n (and its parent p) out
                    const int pside = tree[g].child[0] == p ? 0 : 1;
                    tree[g].child[pside] = sibling; ParentOf(sibling) = g;
                    Refit(g);
                    // where to put it back
                    const Box nb = BoxOf(n);
                    const float na = HalfArea(nb);
                    float best_cost = 3.4e38f; uint32_t best = sibling;
                    heap.clear(); heap.push_back({ 0.0f, tree_root });
                    while (!heap.empty())
                    {
                        std::pop_heap(heap.begin(), heap.end()); const Entry e = heap.back(); heap.pop_back();
                        visits++;
                        if (e.induced + na >= best_cost) break;
                        const Box& xb = BoxOf(e.ref);
                        const float direct = HalfArea(Union(xb, nb));
                        const float total = e.induced + direct;
                        if (total < best_cost) { best_cost = total; best = e.ref; }
                        const float child_induced = total - HalfArea(xb);
                        if (!(e.ref & kItem) && child_induced + na < best_cost)
                            for (int c = 0; c < 2; c++) { heap.push_back({ child_induced, tree[e.ref].child[c] }); std::push_heap(heap.begin(), heap.end()); }
                    }
                    // p takes best's place, with children (best, n)
                    const uint32_t bp = ParentOf(best);
                    tree[p].child[0] = best; tree[p].child[1] = n; tree[p].parent = bp;
                    if (bp == kNone) tree_root = p; else tree[bp].child[tree[bp].child[0] == best ? 0 : 1] = p;
                    ParentOf(best) = p; ParentOf(n) = p;
                    tree[p].box = Union(BoxOf(best), nb);
                    Refit(bp);
                }
                (void)depth_limit;
                if (std::getenv("B200RT_BVH_TIMING")) std::fprintf(stderr, "[accel]   reinsertion pass %d: %zu candidates, %llu nodes looked at (budget %llu)\n", pass, order.size(), (unsigned long long)visits, (unsigned long long)budget);
            }
        }

        // tree -> out.wide[0 ..), pre-order from the root
        void EmitTree()
        {
            std::vector<uint32_t> index(tree.size(), kNone);
            std::vector<uint32_t> stack; stack.push_back(tree_root);
            uint32_t next = 0;
            std::vector<uint32_t> sequence; sequence.reserve(tree.size());
            while (!stack.empty())
            {
                const uint32_t t = stack.back(); stack.pop_back();
                index[t] = next++; sequence.push_back(t);
                if (!(tree[t].child[1] & kItem)) stack.push_back(tree[t].child[1]);
                if (!(tree[t].child[0] & kItem)) stack.push_back(tree[t].child[0]);
            }
            const int64_t count = (int64_t)sequence.size();
#pragma omp parallel for schedule(static) if (count > 100000)
            for (int64_t i = 0; i < count; i++)
            {
                const uint32_t t = sequence[i];
                WideNode& w = out.wide[index[t]];
                for (int c = 0; c < 2; c++)
                {
                    const uint32_t ref = tree[t].child[c];
                    SetChildBox(w, c, BoxOf(ref));   // unions of exact reference leaf boxes; SetChildBox rounds outwards
                    w.child[c] = (ref & kItem) ? items[ref & ~kItem].ref : index[ref];
                }
                w.pad[0] = w.pad[1] = 0;
            }
        }

        // Root reference of the own-topology tree (RT_SENTINEL + error on failure).
        uint32_t BuildOwnTopology()
        {
            const bool timing = std::getenv("B200RT_BVH_TIMING") != nullptr;
            auto clock0 = std::chrono::steady_clock::now();
            auto lap = [&](const char* what) { if (!timing) return; auto now = std::chrono::steady_clock::now(); std::fprintf(stderr, "[accel]   %-18s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - clock0).count()); clock0 = now; };
            WalkReference();
            if (error) return RT_SENTINEL;
            const uint32_t leaves = (uint32_t)leaf_nodes.size();
            // the own tree's nodes at the front (root = wide[0]); fat-leaf sub-trees are appended behind them
            out.wide.assign(leaves > 1 ? leaves - 1 : 0, WideNode{});
            lap("walk reference tree");
            CollectLeaves();
            lap("collect leaves");
            if (error) return RT_SENTINEL;
            if (leaves == 1) return items[0].ref;
            if (CeilLog2(leaves) + fat_depth + 4 > RT_STACK_SIZE) { error = RT_E_DEPTH; return RT_SENTINEL; }
            const uint32_t depth_limit = RT_STACK_SIZE - 3 - fat_depth;
            tree.assign(leaves - 1, TNode{});
            item_parent.assign(leaves, kNone);
            tree[0].parent = kNone; tree_root = 0;
#pragma omp parallel
#pragma omp single
            BuildOwn(0, leaves, 0, 0, depth_limit);
            lap("SAH splits");
            for (uint32_t i = leaves - 1; i-- > 0;) tree[i].box = Union(BoxOf(tree[i].child[0]), BoxOf(tree[i].child[1]));  // children follow parents
            lap("inner boxes");
            const int passes = std::getenv("B200RT_REINSERT") ? std::atoi(std::getenv("B200RT_REINSERT")) : 1;
            if (passes > 0 && leaves > 3)
            {
                const std::vector<TNode> before = tree; const std::vector<uint32_t> before_parent = item_parent;
                const double cost0 = TreeCost();
                Optimize(passes, depth_limit);
                if (std::getenv("B200RT_BVH_TIMING")) std::fprintf(stderr, "[accel] reinsertion: area sum %.4g -> %.4g, depth %u\n", cost0, TreeCost(), TreeDepth());
                if (TreeDepth() + fat_depth + 3 > RT_STACK_SIZE) { tree = before; item_parent = before_parent; tree_root = 0; }
                lap("reinsertion");
            }
            EmitTree();
            lap("emit wide nodes");
            return 0;
        }

        // depth of the finished tree (internal index references), by walking it
        uint32_t MeasureDepth(uint32_t root) const
        {
            uint32_t deepest = 0;
            std::vector<std::pair<uint32_t, uint32_t>> stack; stack.push_back({ root, 0u });
            while (!stack.empty())
            {
                const auto [ref, depth] = stack.back(); stack.pop_back();
                if (ref == RT_SENTINEL || (ref & RT_LEAF_BIT)) continue;
                deepest = std::max(deepest, depth + 1);
                stack.push_back({ out.wide[ref].child[0], depth + 1 }); stack.push_back({ out.wide[ref].child[1], depth + 1 });
            }
            return deepest;
        }
    };
}

int BuildAccelHost(const RtBVHNode* nodes, uint32_t node_count, const uint32_t* face_indices,
                   const RtFace* faces, uint32_t face_count, const RtVertex* vertices, uint32_t vertex_count,
                   uint32_t fat_leaf_threshold, AccelHost& out)
{
    if (!nodes || !node_count || !face_indices || !faces || !vertices || !face_count) return RT_E_INVALID;
    if (face_count >= (1u << 27)) return RT_E_INVALID;
    if (fat_leaf_threshold == 0) fat_leaf_threshold = 8;
    if (fat_leaf_threshold > RT_MAX_LEAF_PRIMS) fat_leaf_threshold = RT_MAX_LEAF_PRIMS;

    const bool timing = std::getenv("B200RT_BVH_TIMING") != nullptr;
    auto clock0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) { if (!timing) return; auto now = std::chrono::steady_clock::now(); std::fprintf(stderr, "[accel] %-20s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - clock0).count()); clock0 = now; };
    out = AccelHost();
    out.tris.resize(face_count); out.shade.resize(face_count);  // (0.4 s of page faults at 14 M slots: first touch of 1.3 GB)
    lap("allocate");
    int invalid = 0;
#pragma omp parallel for schedule(static) reduction(| : invalid) if (face_count > 100000)
    for (int64_t s = 0; s < (int64_t)face_count; s++)
    {   // gather: one TriRecord + one ShadeRecord per BVH slot
        const uint32_t id = face_indices[s];
        if (id >= face_count) { invalid |= 1; continue; }
        const RtFace& f = faces[id];
        if (f.v0 >= vertex_count || f.v1 >= vertex_count || f.v2 >= vertex_count) { invalid |= 1; continue; }
        const RtVertex& a = vertices[f.v0]; const RtVertex& b = vertices[f.v1]; const RtVertex& c = vertices[f.v2];
        TriRecord& t = out.tris[s];
        t.v0[0] = a.position.x; t.v0[1] = a.position.y; t.v0[2] = a.position.z; t.face_id = id;
        t.e1[0] = b.position.x - a.position.x; t.e1[1] = b.position.y - a.position.y; t.e1[2] = b.position.z - a.position.z; t.ref_leaf = 0;
        t.e2[0] = c.position.x - a.position.x; t.e2[1] = c.position.y - a.position.y; t.e2[2] = c.position.z - a.position.z; t.orig_slot = (uint32_t)s;
        ShadeRecord& h = out.shade[s];
        h.n0[0] = a.normal.x; h.n0[1] = a.normal.y; h.n0[2] = a.normal.z; h.material = f.material;
        h.n1[0] = b.normal.x; h.n1[1] = b.normal.y; h.n1[2] = b.normal.z; h.pad0 = 0;
        h.n2[0] = c.normal.x; h.n2[1] = c.normal.y; h.n2[2] = c.normal.z; h.pad1 = 0;
    }
    if (invalid) return RT_E_INVALID;
    lap("gather records");
    Builder builder{ nodes, node_count, fat_leaf_threshold, out, {}, {}, 0 };
    const char* topology = std::getenv("B200RT_TOPOLOGY");
    // The own tree rests on one property of the reference's: every inner box holds its children's boxes, so a ray that passes a leaf's
    // box passes all of its ancestors' (the slab test is monotone in the box). BVH.cpp's UpdateBounds gives that for finite vertices;
    // NaN or infinite vertices can break it (a NaN bound fails every comparison: the reference never enters such a node, whatever
    // lies below). Then the reference's own topology is kept, boxes as they are.
    bool nested = true;
    for (uint32_t n = 0; n < node_count && nested; n++)
    {
        const RtBVHNode& p = nodes[n];
        const float b[6] = { p.bounds_min.x, p.bounds_min.y, p.bounds_min.z, p.bounds_max.x, p.bounds_max.y, p.bounds_max.z };
        for (float v : b) if (!std::isfinite(v)) nested = false;
        if (p.prim_count || !nested) continue;
        for (uint32_t c = p.first_index; c < p.first_index + 2 && c < node_count; c++)
        {
            const RtBVHNode& q = nodes[c];
            if (!(p.bounds_min.x <= q.bounds_min.x && p.bounds_min.y <= q.bounds_min.y && p.bounds_min.z <= q.bounds_min.z &&
                  p.bounds_max.x >= q.bounds_max.x && p.bounds_max.y >= q.bounds_max.y && p.bounds_max.z >= q.bounds_max.z)) nested = false;
        }
    }
    if (!nested || (topology && !std::strcmp(topology, "ref")))
    {   // the reference's own midpoint-split topology (A/B switch; scenes whose boxes do not nest)
        out.wide.reserve(node_count / 2 + 16);
        out.root_ref = builder.Emit(0, 0);
    }
    else
    {
        out.root_ref = builder.BuildOwnTopology();
        if (!builder.error) out.wide_depth = builder.MeasureDepth(out.root_ref);
    }
    lap("wide nodes");
    if (builder.error) return builder.error;
    if (out.wide_depth + 2 > RT_STACK_SIZE || out.ref_depth + 2 > RT_STACK_SIZE) return RT_E_DEPTH;
    if (out.wide.empty()) out.wide.push_back(WideNode{});  // keep the device array non-empty
    // internal (count-1) << 27 | index references -> byte offsets into the (wide nodes | triangle records) arena (accel.h)
    const uint64_t tri_byte0 = (uint64_t)out.wide.size() * sizeof(WideNode);
    if (tri_byte0 + (uint64_t)out.tris.size() * sizeof(TriRecord) + 64 > RT_REF_BYTE_MASK) return RT_E_INVALID;  // > 2 GiB of nodes + triangles
    out.tri_byte0 = (uint32_t)tri_byte0;
    auto to_offset = [&](uint32_t r) -> uint32_t
    {
        if (r == RT_SENTINEL) return r;
        if (r & RT_LEAF_BIT) return RT_LEAF_BIT | (uint32_t)(tri_byte0 + (uint64_t)(r & 0x07FFFFFFu) * sizeof(TriRecord)) | ((r >> 27) & 0xFu);
        return r * (uint32_t)sizeof(WideNode);
    };
    for (WideNode& w : out.wide) { w.child[0] = to_offset(w.child[0]); w.child[1] = to_offset(w.child[1]); }
    out.root_ref = to_offset(out.root_ref);
    lap("byte-offset refs");
    if (timing && !(out.root_ref & RT_LEAF_BIT))
    {   // expected steps of a random ray that hits the root box (surface-area measure): one node step per entered inner child, one
        // triangle step per triangle of an entered leaf
        auto area = [&](const WideNode& w, int c) { float ctr[3], h[3]; ChildCentreHalf(w, c, ctr, h); return h[0] < 0 ? 0.0 : 4.0 * ((double)h[0] * h[1] + (double)h[1] * h[2] + (double)h[2] * h[0]); };
        Box rb = EmptyBox();
        for (int c = 0; c < 2; c++)
        {
            float ctr[3], h[3]; ChildCentreHalf(out.wide[0], c, ctr, h);
            if (h[0] < 0) continue;
            float lo[3], hi[3];
            for (int k = 0; k < 3; k++) { lo[k] = ctr[k] - h[k]; hi[k] = ctr[k] + h[k]; }
            Grow(rb, lo); Grow(rb, hi);
        }
        const double root_area = HalfArea(rb);
        double node_steps = 1.0, tri_steps = 0.0, sub_node_steps = 0.0, sub_tri_steps = 0.0;
        const size_t own_nodes = out.ref_leaf_count > 1 ? out.ref_leaf_count - 1 : 0;   // the fat leaves' private sub-trees follow them
        for (size_t i = 0; i < out.wide.size(); i++)
            for (int c = 0; c < 2; c++)
            {
                const WideNode& w = out.wide[i];
                if (w.child[c] == RT_SENTINEL) continue;
                const double a = area(w, c) / root_area;
                if (w.child[c] & RT_LEAF_BIT) { tri_steps += a * ((w.child[c] & RT_REF_COUNT_MASK) + 1u); if (i >= own_nodes) sub_tri_steps += a * ((w.child[c] & RT_REF_COUNT_MASK) + 1u); }
                else { node_steps += a; if (i >= own_nodes) sub_node_steps += a; }
            }
        if (builder.items.size()) std::fprintf(stderr, "[accel] of which inside fat-leaf sub-trees: %.2f node steps + %.2f triangle steps (%u fat leaves)\n", sub_node_steps, sub_tri_steps, out.fat_leaf_count);
        std::fprintf(stderr, "[accel] surface-area cost: %.2f node steps + %.2f triangle steps = %.2f, depth %u\n", node_steps, tri_steps, node_steps + tri_steps, out.wide_depth);
    }
    return 0;
}

// Host-only structural check of a re-layout (no device): walks the wide tree from the root through the byte-offset references and
// verifies what the traversal kernels rely on. Returns 0 or a negative RT_E_*; `why` receives the first violation.
int CheckAccelHost(const AccelHost& a, const RtBVHNode* nodes, uint32_t node_count, uint64_t* info4, const char** why)
{
    auto fail = [&](const char* w) { if (why) *why = w; return (int)RT_E_INVALID; };
    const uint64_t arena = (uint64_t)a.wide.size() * sizeof(WideNode) + (uint64_t)a.tris.size() * sizeof(TriRecord);
    if (a.tri_byte0 != a.wide.size() * sizeof(WideNode)) return fail("tri_byte0 is not the end of the wide nodes");
    std::vector<uint8_t> seen(a.tris.size(), 0);
    std::vector<std::pair<uint32_t, uint32_t>> stack;  // (reference, depth)
    stack.push_back({ a.root_ref, 0 });
    uint64_t leaves = 0, inner = 0, max_depth = 0;
    while (!stack.empty())
    {
        const auto [ref, depth] = stack.back(); stack.pop_back();
        if (ref == RT_SENTINEL) return fail("sentinel used as a child reference");
        max_depth = std::max<uint64_t>(max_depth, depth);
        const uint32_t off = ref & RT_REF_BYTE_MASK;
        if (ref & RT_LEAF_BIT)
        {
            const uint32_t count = (ref & RT_REF_COUNT_MASK) + 1u;
            if (off < a.tri_byte0 || (off - a.tri_byte0) % sizeof(TriRecord)) return fail("leaf reference does not address a triangle record");
            const uint64_t first = (off - a.tri_byte0) / sizeof(TriRecord);
            if (((off - a.tri_byte0) >> 4) * 0xAAAAAAABu != (uint32_t)first) return fail("offset -> slot inverse-of-3 division is not exact");
            if (first + count > a.tris.size() || count > RT_MAX_LEAF_PRIMS) return fail("leaf range outside the triangle records");
            uint32_t leaf_node = a.tris[first].ref_leaf & RT_REF_LEAF_MASK;
            for (uint64_t s = first; s < first + count; s++)
            {
                if (seen[s]++) return fail("slot reachable twice");
                const TriRecord& t = a.tris[s];
                const uint32_t rl = t.ref_leaf & RT_REF_LEAF_MASK;
                if (rl >= node_count || nodes[rl].prim_count == 0) return fail("ref_leaf is not a reference leaf");
                if (t.orig_slot < nodes[rl].first_index || t.orig_slot >= nodes[rl].first_index + nodes[rl].prim_count) return fail("orig_slot outside its reference leaf");
                if (rl != leaf_node && nodes[rl].prim_count <= RT_MAX_LEAF_PRIMS && nodes[leaf_node].prim_count <= RT_MAX_LEAF_PRIMS) return fail("product leaf mixes reference leaves");
            }
            leaves++;
        }
        else
        {
            if (off % sizeof(WideNode) || off >= a.tri_byte0 || (ref & RT_REF_COUNT_MASK)) return fail("inner reference does not address a wide node");
            if (depth + 2 > RT_STACK_SIZE) return fail("deeper than the traversal stack");
            const WideNode& w = a.wide[off / sizeof(WideNode)];
            inner++;
            stack.push_back({ w.child[0], depth + 1 }); stack.push_back({ w.child[1], depth + 1 });
        }
    }
    for (uint8_t v : seen) if (v != 1) return fail("slot not reachable from the root");
    if (arena + 64 > RT_REF_BYTE_MASK) return fail("arena larger than the reference range");
    // Box containment, bottom-up: the stored box of every wide-node child (centre + half extent) must
    // hold, in exact arithmetic, the reference leaf boxes of everything below it (a fat leaf's private sub-tree: the triangles below).
    // This is what makes the conservative node test safe to cull with; exactness is the deferred reference-leaf check's business.
    struct DBox { double lo[3], hi[3]; };
    bool contained = true;
    auto bounds = [&](auto&& self, uint32_t ref) -> DBox
    {
        DBox b{ { 1e300, 1e300, 1e300 }, { -1e300, -1e300, -1e300 } };
        auto grow = [&](const double* lo, const double* hi) { for (int k = 0; k < 3; k++) { b.lo[k] = std::min(b.lo[k], std::max(lo[k], -1e30)); b.hi[k] = std::max(b.hi[k], std::min(hi[k], 1e30)); } };  // (SetChildBox: beyond 1e30 = unconstrained)
        const uint32_t off = ref & RT_REF_BYTE_MASK;
        if (ref & RT_LEAF_BIT)
        {
            const uint32_t count = (ref & RT_REF_COUNT_MASK) + 1u;
            const uint64_t first = (off - a.tri_byte0) / sizeof(TriRecord);
            const RtBVHNode& rn = nodes[a.tris[first].ref_leaf & RT_REF_LEAF_MASK];
            if (rn.prim_count == count)
            {
                const double lo[3] = { rn.bounds_min.x, rn.bounds_min.y, rn.bounds_min.z }, hi[3] = { rn.bounds_max.x, rn.bounds_max.y, rn.bounds_max.z };
                grow(lo, hi);
            }
            else
                for (uint64_t s = first; s < first + count; s++)
                {
                    const TriRecord& t = a.tris[s];
                    for (int v = 0; v < 3; v++)
                    {
                        double p[3];
                        for (int k = 0; k < 3; k++) p[k] = (double)t.v0[k] + (v == 1 ? (double)t.e1[k] : v == 2 ? (double)t.e2[k] : 0.0);
                        grow(p, p);
                    }
                }
            return b;
        }
        const WideNode& w = a.wide[off / sizeof(WideNode)];
        for (int c = 0; c < 2; c++)
        {
            const DBox cb = self(self, w.child[c]);
            double lo[3], hi[3];
            float ctr[3], half[3];
            ChildCentreHalf(w, c, ctr, half);
            for (int k = 0; k < 3; k++) { lo[k] = (double)ctr[k] - (double)half[k]; hi[k] = (double)ctr[k] + (double)half[k]; }
            for (int k = 0; k < 3; k++) if (!(lo[k] <= cb.lo[k] && hi[k] >= cb.hi[k])) contained = false;
            grow(cb.lo, cb.hi);
        }
        return b;
    };
    bounds(bounds, a.root_ref);
    if (!contained) return fail("a wide-node child box does not contain the reference boxes below it");
    // The node test itself, replayed on the host in the kernel's own fp32 arithmetic (traverse.cuh center_slabs + the relative widening:
    // float subtract, float multiply, fmaf - this file is compiled with -ffp-contract=off): rays that hit what lies below a child, by an
    // exact double-precision slab test of the contained bounds, must pass it. Deterministic sample of children, rays from near, far, and
    // from points on the box faces, towards points inside the contained bounds.
    {
        uint64_t rng = 0x9E3779B97F4A7C15ull;
        auto uni = [&]() { rng = rng * 6364136223846793005ull + 1442695040888963407ull; return (double)(rng >> 11) * (1.0 / 9007199254740992.0); };
        const size_t stride = std::max<size_t>(1, a.wide.size() / 4096);
        const float kWiden = 1.0e-6f;
        for (size_t wi = 0; wi < a.wide.size() && a.root_ref != RT_SENTINEL && !(a.root_ref & RT_LEAF_BIT); wi += stride)
        {
            const WideNode& w = a.wide[wi];
            for (int c = 0; c < 2; c++)
            {
                const DBox cb = bounds(bounds, w.child[c]);
                float ctr[3], half[3];
                ChildCentreHalf(w, c, ctr, half);
                double ext = 1e-6;
                for (int k = 0; k < 3; k++) ext = std::max(ext, cb.hi[k] - cb.lo[k]);
                if (!(ext < 1e30)) continue;   // clamped infinite bounds: the sampled origins would overflow fp32
                for (int r = 0; r < 12; r++)
                {
                    double target[3], od[3];
                    for (int k = 0; k < 3; k++) target[k] = cb.lo[k] + (cb.hi[k] - cb.lo[k]) * (0.05 + 0.9 * uni());
                    const double reach = r < 4 ? 0.6 * ext : r < 8 ? 40.0 * ext : 3000.0 * ext;
                    for (int k = 0; k < 3; k++) od[k] = target[k] + (2.0 * uni() - 1.0) * reach;
                    if (r % 4 == 3) { const int k = (int)(uni() * 3.0) % 3; od[k] = uni() < 0.5 ? cb.lo[k] : cb.hi[k]; }  // origin on a face plane
                    float o[3], d[3];
                    double len = 0;
                    for (int k = 0; k < 3; k++) { o[k] = (float)od[k]; len += (target[k] - o[k]) * (target[k] - o[k]); }
                    if (!(len > 0)) continue;
                    for (int k = 0; k < 3; k++) d[k] = (float)((target[k] - o[k]) / std::sqrt(len));
                    if (r % 6 == 5) d[(int)(uni() * 3.0) % 3] = 0.0f;  // axis-parallel component
                    // ground truth: exact slabs of the contained bounds for the ray (o, d) as the kernel sees it
                    double t0 = 0.0, t1 = 1e300;
                    bool truth = true;
                    for (int k = 0; k < 3 && truth; k++)
                    {
                        // a ray that runs INSIDE the plane of a zero-thickness box (an axis-aligned wall: lo == hi, d == 0, o on the plane) is
                        // not a witness: everything below such a box is triangles in that plane, which the ray sees edge-on (det == 0: no
                        // hit in the reference either), and the centre form gives it the slab [-h * 1e30, +h * 1e30] around t = 0 with h ~ 1e-37
                        if (d[k] == 0.0f && cb.lo[k] == cb.hi[k]) { truth = false; continue; }
                        if (d[k] == 0.0f) { truth = (double)o[k] >= cb.lo[k] && (double)o[k] <= cb.hi[k]; continue; }
                        const double ta = (cb.lo[k] - o[k]) / d[k], tb = (cb.hi[k] - o[k]) / d[k];
                        t0 = std::max(t0, std::min(ta, tb)); t1 = std::min(t1, std::max(ta, tb));
                    }
                    if (!truth || !(t0 <= t1)) continue;
                    // the kernel's test, with the reciprocal direction as IEEE gives it and 2 ulp to either side (room for an
                    // approximate reciprocal, RT_TW_RCP_APPROX)
                    for (int variant = 0; variant < 3; variant++)
                    {
                        const float scale = variant == 0 ? 1.0f : variant == 1 ? 1.0f + 2.4e-7f : 1.0f - 2.4e-7f;
                        float mn = -INFINITY, mx = INFINITY;
                        for (int k = 0; k < 3; k++)
                        {
                            const float inv = d[k] != 0.0f ? (1.0f / d[k]) * scale : 1e30f;
                            const float diff = ctr[k] - o[k];
                            const float mid = diff * inv;
                            mn = std::fmax(mn, std::fmaf(half[k], -std::fabs(inv), mid));
                            mx = std::fmin(mx, std::fmaf(half[k], std::fabs(inv), mid));
                        }
                        const float lo = std::fmaf(-kWiden, std::fabs(mn), mn), hi = std::fmaf(kWiden, std::fabs(mx), mx);
                        if (!(hi >= lo && lo <= 1e30f && hi > 0.0f)) return fail("the centre + half-extent node test rejects a ray that hits the bounds below the child");
                    }
                }
            }
        }
    }
    if (info4) { info4[0] = inner; info4[1] = leaves; info4[2] = max_depth; info4[3] = arena; }
    return 0;
}
