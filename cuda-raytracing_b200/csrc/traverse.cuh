// Closest-hit traversal for sm_100a. This TU is compiled with -fmad=false: every `a*b+c` below is
// two IEEE roundings, exactly like the host-compiled reference the parity tests compare against
// (DESIGN.md "contraction policy"); FMAs are written explicitly (fmaf / __ffma2_rn) where parity does
// not hang on them (the conservative node test).
//
// Result rule (BASELINE.json): min t over the triangles the reference would test, exact ties ->
// lowest face id. Triangle arithmetic restates glm::intersectRayTriangle
// (include/glm/gtx/intersect.inl:29-94) on the pre-gathered v0/e1/e2; the exact leaf-box test
// restates Math::IntersectAABB (RayTracing/Math.h:50-61) as BVHRayHit calls it
// (RayTracing/main_raytracing.cu:46: UNNORMALISED direction, normalised-units length).
//
// This header holds the shipped path only. Every variant that was measured and lost (loop shapes, stack forms, node formats,
// cull forms, warp-cooperative schemes) lives in profiles/experiments/ as a patch with its A/B line; DESIGN.md section 7 lists them.
#pragma once
#include <cfloat>
#include <cstdint>
#include "accel.h"

#ifndef RT_TW_RCP_APPROX
#define RT_TW_RCP_APPROX 1   /* reciprocal direction of the CONSERVATIVE node test by MUFU.RCP: 45.05 -> 44.92 ms (3 interleaved rounds, +-0.02), bit-identical */
#endif
// RT_TW_FORCE_REDO=<mask> (test build, tests/test_gpu_redo_build.py): rays whose winning slot has (slot & mask) == 0 take the redo
// path (per-candidate exact leaf-box check) although their deferred check passed - the path a grazing ray takes about never.

struct Vec3 { float x, y, z; };

__device__ __forceinline__ Vec3 mk3(float x, float y, float z) { Vec3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ Vec3 operator+(Vec3 a, Vec3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ Vec3 operator-(Vec3 a, Vec3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ Vec3 operator-(Vec3 a) { return mk3(-a.x, -a.y, -a.z); }
// The MULTIPLIES of the vec3 algebra two at a time (Blackwell FMUL2 on the x and y lanes; IEEE round-to-nearest per lane, factors
// exchanged at most - exact), every ADD scalar and in GLM's order. Adds are deliberately NOT packed: ptxas fuses a packed multiply
// that feeds a packed add into FFMA2 even under -fmad=false (one rounding instead of two; tests/test_layout_and_scene.py guards
// the shipped SASS against any FFMA2 outside the node test).
__device__ __forceinline__ float2 xy_(Vec3 a) { return make_float2(a.x, a.y); }
__device__ __forceinline__ Vec3 operator*(Vec3 a, Vec3 b) { const float2 m = __fmul2_rn(xy_(a), xy_(b)); return mk3(m.x, m.y, a.z * b.z); }
__device__ __forceinline__ Vec3 operator*(Vec3 a, float s) { const float2 m = __fmul2_rn(xy_(a), make_float2(s, s)); return mk3(m.x, m.y, a.z * s); }
__device__ __forceinline__ Vec3 operator*(float s, Vec3 a) { const float2 m = __fmul2_rn(make_float2(s, s), xy_(a)); return mk3(m.x, m.y, s * a.z); }
__device__ __forceinline__ Vec3 operator/(Vec3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
// glm/detail/func_geometric.inl:48-54,68-78,82-89 ; func_exponential.inl:136-139
__device__ __forceinline__ float dot3(Vec3 a, Vec3 b) { const float2 m = __fmul2_rn(xy_(a), xy_(b)); return m.x + m.y + a.z * b.z; }
__device__ __forceinline__ Vec3 cross3(Vec3 x, Vec3 y)
{
    const float2 x_yx = make_float2(x.y, x.x);
    const float2 a = __fmul2_rn(xy_(y), make_float2(x.z, x.z));  // y.x x.z | y.y x.z
    const float2 b = __fmul2_rn(xy_(y), x_yx);                   // y.x x.y | y.y x.x
    const float2 c = __fmul2_rn(x_yx, make_float2(y.z, y.z));    // x.y y.z | x.x y.z
    return mk3(c.x - a.y, a.x - c.y, b.y - b.x);
}
__device__ __forceinline__ Vec3 normalize3(Vec3 v) { return v * (1.0f / sqrtf(dot3(v, v))); }

struct HitRecord
{
    float t;        // normalised-direction units; 1e30f = miss
    uint32_t id;    // face id, RT_SPHERE_ID_BIT | sphere, or RT_MISS_ID
    uint32_t slot;  // BVH slot of the winning triangle (ShadeRecord index)
    uint32_t key;   // exact-tie key of the winner (face id, or original slot = reference leaf order)
    float bu, bv;   // glm baryPosition.x / .y of the winner
};

__device__ __forceinline__ float4 ldg4(const void* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

// Node / leaf references are byte offsets into the accel arena (accel.h): one address form for both kinds of record.
// byte offset -> slot: exact division by 48 = shift by 4, then multiplication by the inverse of 3 modulo 2^32 (the quotient is exact)
__device__ __forceinline__ uint32_t offset_slot(const AccelView& A, uint32_t offset) { return ((offset - A.tri_byte0) >> 4) * 0xAAAAAAABu; }
__device__ __forceinline__ uint32_t ref_slot(const AccelView& A, uint32_t ref) { return offset_slot(A, ref & RT_REF_BYTE_MASK); }
__device__ __forceinline__ uint32_t ref_count(uint32_t ref) { return (ref & RT_REF_COUNT_MASK) + 1u; }
__device__ __forceinline__ bool ref_has_more(uint32_t ref) { return (ref & RT_REF_COUNT_MASK) != 0u; }
__device__ __forceinline__ uint32_t ref_next_tri(uint32_t ref) { return ref + 47u; }  // +48 bytes, -1 prim; caller checked the count

// Math::IntersectAABB with the reference's operand roles. Order-independent use: ray_length = the
// distance the BVH search STARTED with (sphere hit or 1e30).
// kInclusive: `tmin <= ray_length` instead of the reference's `<` - used only by the lowest-id tie rule, where a node whose
// entry distance EQUALS the current best can still hold an exact-tie candidate with a lower id.
template <bool kInclusive = false>
__device__ __forceinline__ bool exact_box(Vec3 o, Vec3 d, const RtBVHNode* node, float ray_length)
{
    const float4 q0 = ldg4(node);                                    // min.xyz max.x
    const float4 q1 = ldg4(reinterpret_cast<const char*>(node) + 16);  // max.yz first count
    const float tx1 = (q0.x - o.x) / d.x, tx2 = (q0.w - o.x) / d.x;
    float tmin = fminf(tx1, tx2), tmax = fmaxf(tx1, tx2);
    const float ty1 = (q0.y - o.y) / d.y, ty2 = (q1.x - o.y) / d.y;
    tmin = fmaxf(tmin, fminf(ty1, ty2)); tmax = fminf(tmax, fmaxf(ty1, ty2));
    const float tz1 = (q0.z - o.z) / d.z, tz2 = (q1.y - o.z) / d.z;
    tmin = fmaxf(tmin, fminf(tz1, tz2)); tmax = fminf(tmax, fmaxf(tz1, tz2));
    return tmax >= tmin && (kInclusive ? tmin <= ray_length : tmin < ray_length) && tmax > 0;
}

// Rays shorter than unit length (|d|^2 below this) make the reference's box cull NON-conservative (Math.h:50-61 compares an
// unnormalised-units entry distance with a normalised-units best, Appendix A.4): its result then depends on its own visiting
// order, so such rays take the reference-order traversal in every variant. Camera rays have |d| >= 1 and bounce rays are
// normalised, so the path tracer never takes this branch; it exists for rt_trace_rays' "direction as given" contract.
#define RT_SHORT_DIR2 0.99999f

// glm::intersectRayTriangle on a TriRecord (r0 = v0 | face id, r1 = e1 | ref leaf, r2 = e2 | original slot). Returns true with
// (t, bu, bv) when GLM would. The same products and sums, two multiplies / adds per instruction where operands pair up (Blackwell
// FMUL2 / FADD2, IEEE round-to-nearest per lane): every product below is one of GLM's with its factors possibly exchanged (exact),
// every sum has GLM's operands in GLM's order. Pairs: (x, y) of v0 / e1 / e2 straight from the 128-bit loads, (x, y) and (y, x) of
// the ray. No packed multiply feeds a packed add (ptxas would fuse the pair into FFMA2: one rounding instead of two).
__device__ __forceinline__ bool tri_test(Vec3 o, Vec3 nd, float4 r0, float4 r1, float4 r2, float& t, float& bu, float& bv)
{
    const float2 e1xy = make_float2(r1.x, r1.y), e2xy = make_float2(r2.x, r2.y);
    const float2 nd_xy = make_float2(nd.x, nd.y), nd_yx = make_float2(nd.y, nd.x);
    const float2 pa = __fmul2_rn(e2xy, make_float2(nd.z, nd.z));   // e2.x nd.z | e2.y nd.z
    const float2 pb = __fmul2_rn(e2xy, nd_yx);                     // e2.x nd.y | e2.y nd.x
    const float2 pc = __fmul2_rn(nd_yx, make_float2(r2.z, r2.z));  // nd.y e2.z | nd.x e2.z
    const float2 pxy = make_float2(pc.x - pa.y, pa.x - pc.y);      // cross(nd, e2).xy
    const float pz = pb.y - pb.x;
    const float2 dd = __fmul2_rn(e1xy, pxy);
    const float det = (dd.x + dd.y) + r1.z * pz;
    const float2 dxy = __fadd2_rn(make_float2(o.x, o.y), make_float2(-r0.x, -r0.y));  // o - v0
    const float dz = o.z - r0.z;
    const float2 uu = __fmul2_rn(dxy, pxy);
    const float u = (uu.x + uu.y) + dz * pz;
    const float2 d_yx = make_float2(dxy.y, dxy.x);
    const float2 qa = __fmul2_rn(e1xy, make_float2(dz, dz));       // e1.x d.z | e1.y d.z
    const float2 qb = __fmul2_rn(e1xy, d_yx);                      // e1.x d.y | e1.y d.x
    const float2 qc = __fmul2_rn(d_yx, make_float2(r1.z, r1.z));   // d.y e1.z | d.x e1.z
    const float2 qxy = make_float2(qc.x - qa.y, qa.x - qc.y);      // cross(dist, e1).xy
    const float qz = qb.y - qb.x;
    const float2 vv = __fmul2_rn(nd_xy, qxy);
    const float v = (vv.x + vv.y) + nd.z * qz;
    {   // GLM's two sign cases (intersect.inl:44-63) as one: for det < 0 its tests are the det > 0 tests of (-u, -v, -(u + v))
        // against -det, and flipping a sign bit is exact - same decisions (NaNs included; tests/test_signfree_miss_equivalence.py),
        // no data-dependent branch on the sign of det (lanes of a warp see front and back faces mixed)
        const uint32_t flip = __float_as_uint(det) & 0x80000000u;
        const float ad = fabsf(det);
        const float us = __uint_as_float(__float_as_uint(u) ^ flip), vs = __uint_as_float(__float_as_uint(v) ^ flip);
        const float uvs = __uint_as_float(__float_as_uint(u + v) ^ flip);
        if (!(ad > FLT_EPSILON) || (us < 0.0f || us > ad) || (vs < 0.0f || uvs > ad)) return false;
    }
    const float inv_det = 1.0f / det;
    const float2 tt = __fmul2_rn(e2xy, qxy);
    t = ((tt.x + tt.y) + r2.z * qz) * inv_det;
    bu = u * inv_det; bv = v * inv_det;
    return true;
}

// Candidate acceptance: main_raytracing.cu:62 (`distance >= best || distance < 0` rejects). The
// reference keeps the FIRST-VISITED candidate on exact ties, which depends on traversal order; here
// ties are resolved by an order-independent key: the face id (kTieById, BASELINE.json's rule for the
// hit-id parity config) or the original BVH slot (= the reference's visiting order inside a leaf,
// used by the path tracer so shading picks the same duplicate the reference does).
__device__ __forceinline__ bool accept(float t, uint32_t key, const HitRecord& h)
{
    // (a NaN t - a vertex at infinity: e2 . q = inf - inf - is no hit. The reference's `>=` / `<` let a NaN distance through, after which
    // every comparison with it is false and what it returns is whatever it visits last; no order-independent rule can mirror that.)
    if (!(t <= h.t) || t < 0.0f) return false;
    return t < h.t || (key < h.key && !(h.id & RT_SPHERE_ID_BIT));  // a sphere at the same t keeps the hit (:62, `>=`)
}

// One triangle record (already loaded) against the ray: glm test, tie rule, optional per-candidate exact leaf-box check.
template <bool kExactBoxPerCandidate, bool kTieById>
__device__ __forceinline__ void leaf_test_one(const AccelView& A, float4 r0, float4 r1, float4 r2, uint32_t slot, Vec3 o, Vec3 d, Vec3 nd,
                                              float start_length, HitRecord& h)
{
    float t, bu, bv;
    if (!tri_test(o, nd, r0, r1, r2, t, bu, bv)) return;
    const uint32_t id = __float_as_uint(r0.w);
    const uint32_t key = kTieById ? id : __float_as_uint(r2.w);
    if (!accept(t, key, h)) return;
    // the reference only tests this triangle if its leaf's box passes (main_raytracing.cu:46)
    if (kExactBoxPerCandidate && !exact_box(o, d, A.ref_nodes + (__float_as_uint(r1.w) & RT_REF_LEAF_MASK), start_length)) return;
    h.t = t; h.id = id; h.slot = slot; h.key = key; h.bu = bu; h.bv = bv;
}

template <bool kExactBoxPerCandidate, bool kTieById>
__device__ __forceinline__ void leaf_tests(const AccelView& A, uint32_t first, uint32_t count, Vec3 o, Vec3 d, Vec3 nd,
                                           float start_length, HitRecord& h)
{
    const TriRecord* tr = A.tris + first;
    for (uint32_t i = 0; i < count; i++)
    {
        const float4 r0 = ldg4(&tr[i]), r1 = ldg4(reinterpret_cast<const char*>(&tr[i]) + 16), r2 = ldg4(reinterpret_cast<const char*>(&tr[i]) + 32);
        leaf_test_one<kExactBoxPerCandidate, kTieById>(A, r0, r1, r2, first + i, o, d, nd, start_length, h);
    }
}

// Entry / exit distances of both child boxes of a centre + half-extent wide node (accel.h): per axis mid = (centre - o) * inv,
// entry = mid - half * |inv|, exit = mid + half * |inv| - twelve packed fma-pipe instructions and four 3-input min/max, no per-axis
// min/max pair. Conservative, not exact: mid carries ~1 ulp of |mid| <= |entry| + half |inv|; the half |inv| share is covered by
// the 2^-21 the builder adds to every half extent, the rest by the callers' relative widening (kWiden).
__device__ __forceinline__ void center_slabs(float4 a, float4 b, float4 c, float2 no_xy, float2 no_zz, float2 i_xy, float2 i_zz,
                                             float2 ai_xy, float2 ai_zz, float& mn0, float& mx0, float& mn1, float& mx1)
{
    const float2 m0 = __fmul2_rn(__fadd2_rn(make_float2(a.x, a.y), no_xy), i_xy);
    const float2 m1 = __fmul2_rn(__fadd2_rn(make_float2(b.x, b.y), no_xy), i_xy);
    const float2 mz = __fmul2_rn(__fadd2_rn(make_float2(c.x, c.y), no_zz), i_zz);
    const float2 h0 = make_float2(a.z, a.w), h1 = make_float2(b.z, b.w), hz = make_float2(c.z, c.w);
    const float2 nai_xy = make_float2(-ai_xy.x, -ai_xy.y), nai_zz = make_float2(-ai_zz.x, -ai_zz.y);
    const float2 n0 = __ffma2_rn(h0, nai_xy, m0), f0 = __ffma2_rn(h0, ai_xy, m0);
    const float2 n1 = __ffma2_rn(h1, nai_xy, m1), f1 = __ffma2_rn(h1, ai_xy, m1);
    const float2 nzz = __ffma2_rn(hz, nai_zz, mz), fzz = __ffma2_rn(hz, ai_zz, mz);
    mn0 = fmaxf(fmaxf(n0.x, n0.y), nzz.x); mx0 = fminf(fminf(f0.x, f0.y), fzz.x);
    mn1 = fmaxf(fmaxf(n1.x, n1.y), nzz.y); mx1 = fminf(fminf(f1.x, f1.y), fzz.y);
}

// Per-ray constants of the conservative node test: reciprocal of the normalised direction (zero components -> huge finite),
// negated origin, |reciprocal|, in the register pairs center_slabs() wants.
struct SlabRay
{
    float2 i_xy, i_zz, no_xy, no_zz, ai_xy, ai_zz;
};
__device__ __forceinline__ SlabRay make_slab_ray(Vec3 o, Vec3 nd)
{
    const float big = 1e30f;
#if RT_TW_RCP_APPROX
    // the reciprocal only feeds the conservative node test, whose 8-ulp relative widening has room for MUFU.RCP's ~1 ulp (the host
    // replay of the test in accel_build.cpp CheckAccelHost also runs with the reciprocal 2 ulp to either side)
    auto rcp = [](float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; };
    const float ix = nd.x != 0.0f ? rcp(nd.x) : big, iy = nd.y != 0.0f ? rcp(nd.y) : big, iz = nd.z != 0.0f ? rcp(nd.z) : big;
#else
    const float ix = nd.x != 0.0f ? 1.0f / nd.x : big, iy = nd.y != 0.0f ? 1.0f / nd.y : big, iz = nd.z != 0.0f ? 1.0f / nd.z : big;
#endif
    SlabRay r;
    r.i_xy = make_float2(ix, iy); r.i_zz = make_float2(iz, iz);
    r.no_xy = make_float2(-o.x, -o.y); r.no_zz = make_float2(-o.z, -o.z);
    r.ai_xy = make_float2(fabsf(ix), fabsf(iy)); r.ai_zz = make_float2(fabsf(iz), fabsf(iz));
    return r;
}

// Product traversal: wide nodes, near child first, conservative node test in normalised units, ONE loop whose
// iteration is either one wide node (two boxes) or one triangle of a leaf ("if-if"; the leaf reference is re-encoded with its
// remaining range, no extra state), ONE fetch for both kinds of step (56 bytes from the lane's own record - a 64-byte wide node or
// a 48-byte triangle record - before the kinds diverge, so their latencies overlap), traversal stack in local memory addressed
// through a pointer.
// `h` enters holding the sphere result (or a miss with t = 1e30).
//   kDeferred: the exact reference-leaf box re-check (6 IEEE divisions, ~110 instructions) is not done per accepted candidate but
//     ONCE, on the winner, after the loop. Unverified candidates only ever shrink the search interval, so any candidate that
//     beats them is still found; if the WINNER's leaf box fails (a ray grazing a leaf box within rounding - rare) the function
//     returns false and the caller redoes the ray with the per-candidate check. Exact by construction.
//   work (optional, measurement instantiations): [0] += wide-node steps, [1] += triangle tests of this ray.
template <bool kTieById, bool kDeferred>
__device__ __forceinline__ bool trace_wide_impl(const AccelView& A, Vec3 o, Vec3 d, Vec3 nd, HitRecord& h, uint32_t* work = nullptr)
{
    const float start_length = h.t;
    const float kWiden = 1.0e-6f;
    const SlabRay R = make_slab_ray(o, nd);
    uint32_t stack[RT_STACK_SIZE];
    uint32_t* sp = stack;   // push = store + add, pop = load + add (no index -> address LEA per access)
    *sp++ = RT_SENTINEL;
    uint32_t node = A.root_ref;
    while (node != RT_SENTINEL)
    {
        const bool is_leaf = (node & RT_LEAF_BIT) != 0u;
        const uint32_t slot = node & RT_REF_BYTE_MASK;  // arena byte offset; a winner's h.slot is converted to its slot after the loop
        const char* rec = A.arena + slot;                // one address form for wide nodes and triangle records
        const float4 q0 = ldg4(rec), q1 = ldg4(rec + 16), q2 = ldg4(rec + 32);
        const uint2 q3 = __ldg(reinterpret_cast<const uint2*>(rec + 48));
        if (!is_leaf)
        {
            if (work) work[0]++;
            float mn0, mx0, mn1, mx1;
            center_slabs(q0, q1, q2, R.no_xy, R.no_zz, R.i_xy, R.i_zz, R.ai_xy, R.ai_zz, mn0, mx0, mn1, mx1);
            // widen (superset of the reference's divisions), cull against the current best
            const float lo0 = fmaf(-kWiden, fabsf(mn0), mn0), hi0 = fmaf(kWiden, fabsf(mx0), mx0);
            const float lo1 = fmaf(-kWiden, fabsf(mn1), mn1), hi1 = fmaf(kWiden, fabsf(mx1), mx1);
            const bool hit0 = hi0 >= lo0 && lo0 <= h.t && hi0 > 0.0f;
            const bool hit1 = hi1 >= lo1 && lo1 <= h.t && hi1 > 0.0f;
            if (hit0 && hit1)
            {
                const bool swap = mn1 < mn0;
                *sp++ = swap ? q3.x : q3.y;
                node = swap ? q3.y : q3.x;
            }
            else if (hit0) node = q3.x;
            else if (hit1) node = q3.y;
            else node = *--sp;
        }
        else
        {
            if (work) work[1]++;
            leaf_test_one<!kDeferred, kTieById>(A, q0, q1, q2, slot, o, d, nd, start_length, h);
            if (ref_has_more(node)) node = ref_next_tri(node);
            else node = *--sp;
        }
    }
    // h entered as a sphere hit or a miss, so a triangle id means "found in this loop": its h.slot is still an arena byte offset
    if (h.id != RT_MISS_ID && !(h.id & RT_SPHERE_ID_BIT))
    {
        h.slot = offset_slot(A, h.slot);
        if (kDeferred)
        {   // verify the winner's reference leaf (main_raytracing.cu:46)
            const uint32_t leaf = __ldg(&A.tris[h.slot].ref_leaf) & RT_REF_LEAF_MASK;
#ifdef RT_TW_FORCE_REDO
            if ((h.slot & RT_TW_FORCE_REDO) == 0u) return false;
#endif
            return exact_box(o, d, A.ref_nodes + leaf, start_length);
        }
    }
    return true;
}

template <bool kTieById>
__device__ __noinline__ void trace_wide_checked(const AccelView& A, Vec3 o, Vec3 d, Vec3 nd, HitRecord& h)
{
    trace_wide_impl<kTieById, false>(A, o, d, nd, h);
}

// `h` enters holding what reinit(h) produces (a miss, or the sphere result); reinit is only called again on the rare redo.
template <bool kTieById, class Reinit>
__device__ __forceinline__ void trace_wide(const AccelView& A, Vec3 o, Vec3 d, Vec3 nd, HitRecord& h, Reinit reinit, uint32_t* work = nullptr)
{
    if (trace_wide_impl<kTieById, true>(A, o, d, nd, h, work)) return;
    reinit(h);
    trace_wide_checked<kTieById>(A, o, d, nd, h);
}

// Reference-order traversal over the reference-layout nodes: right child first, exact box test at
// every node against the CURRENT best (main_raytracing.cu:43-78). By-construction parity kernel.
template <bool kTieById>
__device__ __forceinline__ void trace_reference_order(const AccelView& A, Vec3 o, Vec3 d, Vec3 nd, HitRecord& h,
                                                      uint32_t* nodes_popped = nullptr, uint32_t* tris_tested = nullptr)
{
    uint32_t stack[RT_STACK_SIZE];
    int sp = 0;
    stack[sp++] = 0;
    const bool inclusive = kTieById && dot3(d, d) >= RT_SHORT_DIR2;
    while (sp)
    {
        const RtBVHNode* n = A.ref_nodes + stack[--sp];
        if (nodes_popped) (*nodes_popped)++;
        if (!(inclusive ? exact_box<true>(o, d, n, h.t) : exact_box<false>(o, d, n, h.t))) continue;
        const float4 q1 = ldg4(reinterpret_cast<const char*>(n) + 16);
        const uint32_t first = __float_as_uint(q1.z), count = __float_as_uint(q1.w);
        if (count)
        {
            if (tris_tested) (*tris_tested) += count;
            leaf_tests<false, kTieById>(A, first, count, o, d, nd, 0.0f, h);
        }
        else { stack[sp++] = first; stack[sp++] = first + 1; }
    }
}

// Spheres first, as GetRayHit does (main_raytracing.cu:89-103, glm intersect.inl:135-153).
__device__ __forceinline__ void sphere_tests(const RtSphere* spheres, int count, Vec3 o, Vec3 nd, HitRecord& h)
{
    for (int i = 0; i < count; i++)
    {
        const float4 s = ldg4(&spheres[i]);  // position, radius
        const Vec3 diff = mk3(s.x, s.y, s.z) - o;
        const float t0 = dot3(diff, nd);
        const float r2 = s.w * s.w;
        const float d2 = dot3(diff, diff) - t0 * t0;
        if (d2 > r2) continue;
        const float t1 = sqrtf(r2 - d2);
        const float t = t0 > t1 + FLT_EPSILON ? t0 - t1 : t0 + t1;
        if (!(t > FLT_EPSILON)) continue;
        if (t >= h.t) continue;
        h.t = t; h.id = RT_SPHERE_ID_BIT | (uint32_t)i; h.slot = (uint32_t)i; h.key = 0;
    }
}
