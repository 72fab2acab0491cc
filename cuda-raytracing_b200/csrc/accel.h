// Device layout of the acceleration structure ("GPUScene upload re-layout", north-star item 1).
//
// The reference keeps 32-B nodes that name ONE box each, a u32 face-index permutation, 16-B faces
// and 32-B vertices: a triangle test chases index -> face -> 3 vertices = 116 B through three
// dependent loads (RayTracing/main_raytracing.cu:53-56). Here the SAME leaves (same leaf contents, same
// order of slots as BVH.cpp's face_indices, every leaf with its exact reference box) are stored as
//
//   WideNode (64 B, 4 x 128-bit loads): the boxes of BOTH children of an inner node plus two child
//            references, so one pop tests two boxes and orders them near-first. The tree ABOVE the
//            leaves is the product's own: a surface-area-heuristic tree (binned / swept SAH, then one
//            insertion-based optimisation pass) over the reference's leaves, not BVH.cpp's spatial-midpoint
//            tree - the hit rule and the deferred exact leaf-box check (traverse.cuh) make the result a
//            function of which triangles a ray meets, not of the tree above them (accel_build.cpp;
//            19.2 -> 14.5 node steps per ray on the bench workload, 44.8 -> 37.6 ms, frames bit-identical);
//   TriRecord (48 B, 3 x 128-bit loads) per BVH slot: v0, e1 = v1-v0, e2 = v2-v0 pre-gathered
//            (bit-identical to what glm::intersectRayTriangle computes first, intersect.inl:38-39),
//            with the face id and the owning reference leaf in the spare lanes;
//   ShadeRecord (48 B) per slot: the three vertex normals and the material index, read once per hit.
//
// Child reference (u32). WideNodes and TriRecords live in ONE device arena (< 2 GiB) and a reference IS the byte offset of its
// record, so the traversal loop forms the address of either kind with one mask and one add, no select and no scaling:
//   bit31 = 0 -> WideNode at byte offset bits 4..30 (bits 0..3 zero);
//   bit31 = 1 -> leaf: bits 4..30 = byte offset of its first TriRecord (records of a leaf are consecutive, 48 B each),
//                bits 0..3 = prim_count-1 (<= 16 prims; the offset's own low four bits are zero: records are 16-B aligned).
//   "Next triangle of this leaf" = ref + 48 - 1. 0xFFFFFFFF = stack sentinel.
//   Slot (index into tris / shade) = (offset - tri_byte0) / 48. 2 GiB of nodes + triangles = ~27 M triangles.
// Reference leaves with more than `fat_leaf_threshold` prims (BVH.cpp never bounds leaf size; the
// bunny scene has one of 345) keep their exact box in the parent and get a private sub-tree of
// WideNodes over their own slots with conservatively inflated boxes (leaf contents unchanged).
#pragma once
#include <cstdint>
#include <vector>
#include "rt_types.h"

#define RT_LEAF_BIT 0x80000000u
#define RT_SENTINEL 0xFFFFFFFFu
#define RT_REF_BYTE_MASK 0x7FFFFFF0u
#define RT_REF_COUNT_MASK 0x0000000Fu
#define RT_MAX_LEAF_PRIMS 16u
// Wide-node boxes are stored as centre + half extent: the entry / exit distance per axis is mid -/+ half * |1/d| with
// mid = (centre - o) / d, all on the fma pipe - no min/max per axis on the ALU pipe (49.8 -> 46.0 ms against the lo/hi form, bit-identical
// results). Half extents are rounded up and inflated by 2^-21 (+ the centre's own rounding) at build time, so the stored box is a
// superset; exactness is the reference-leaf check's business (traverse.cuh).
#define RT_REF_LEAF_MASK 0x7FFFFFFFu
#ifndef RT_STACK_SIZE
#define RT_STACK_SIZE 64   /* traversal stack entries per thread (local memory); the builder refuses deeper trees */
#endif

struct RtFloat4 { float x, y, z, w; };

struct alignas(64) WideNode
{
    // register-pair friendly order for Blackwell's packed fp32 ops (add/mul/fma .f32x2): every aligned pair of a 128-bit load
    // is combined with one per-ray constant pair, (x, y) or (z, z)
    RtFloat4 a;   // centre0.x centre0.y | half0.x half0.y
    RtFloat4 b;   // centre1.x centre1.y | half1.x half1.y
    RtFloat4 c;   // centre0.z centre1.z | half0.z half1.z
    uint32_t child[2];
    uint32_t pad[2];
};

struct alignas(16) TriRecord
{
    float v0[3]; uint32_t face_id;
    float e1[3]; uint32_t ref_leaf;   // index of the reference leaf node (for the exact box re-check)
    float e2[3]; uint32_t orig_slot; // slot before any fat-leaf permutation (reference visiting order)
};

struct alignas(16) ShadeRecord
{
    float n0[3]; uint32_t material;
    float n1[3]; uint32_t pad0;
    float n2[3]; uint32_t pad1;
};

// Host-side result of the re-layout (accel_build.cpp); uploaded verbatim.
struct AccelHost
{
    std::vector<WideNode> wide;        // wide[0] is the root (unless root_ref is a leaf)
    std::vector<TriRecord> tris;       // one per BVH slot
    std::vector<ShadeRecord> shade;    // one per BVH slot
    uint32_t root_ref = RT_SENTINEL;
    uint32_t tri_byte0 = 0;            // byte offset of tris[0] in the arena = 64 * wide.size()
    uint32_t ref_leaf_count = 0, fat_leaf_count = 0, wide_depth = 0, ref_depth = 0;
};

// Returns 0, or a negative RT_E_* (include/b200rt.h).
int BuildAccelHost(const RtBVHNode* nodes, uint32_t node_count, const uint32_t* face_indices,
                   const RtFace* faces, uint32_t face_count, const RtVertex* vertices, uint32_t vertex_count,
                   uint32_t fat_leaf_threshold, AccelHost& out);

// Host-only structural check (every slot reachable exactly once through byte-offset references, ranges, depths, leaf bookkeeping).
int CheckAccelHost(const AccelHost& a, const RtBVHNode* nodes, uint32_t node_count, uint64_t* info4, const char** why);

// Device-side view passed to kernels by value.
struct AccelView
{
    const char* arena;           // wide nodes, then triangle records (what references address, as byte offsets)
    const WideNode* wide;        // = arena
    const TriRecord* tris;       // = arena + tri_byte0
    const ShadeRecord* shade;
    const RtBVHNode* ref_nodes;  // device copy of the reference nodes (exact leaf boxes, reference-order kernel)
    uint32_t root_ref;
    uint32_t node_count;
    uint32_t tri_byte0;
};

struct RtAccel
{
    AccelView view{};
    uint64_t device_bytes = 0;
    uint32_t wide_count = 0, slot_count = 0, ref_leaf_count = 0, fat_leaf_count = 0, wide_depth = 0, ref_depth = 0;
};
