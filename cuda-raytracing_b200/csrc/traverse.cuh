// Closest-hit traversal for sm_100a. This TU is compiled with -fmad=false: every `a*b+c` below is
// two IEEE roundings, exactly like the host-compiled reference the parity tests compare against
// (DESIGN.md "contraction policy"); FMAs are written explicitly (fmaf) where parity does not hang
// on them (conservative box tests).
//
// Result rule (BASELINE.json): min t over the triangles the reference would test, exact ties ->
// lowest face id. Triangle arithmetic restates glm::intersectRayTriangle
// (include/glm/gtx/intersect.inl:29-94) on the pre-gathered v0/e1/e2; the exact leaf-box test
// restates Math::IntersectAABB (RayTracing/Math.h:50-61) as BVHRayHit calls it
// (RayTracing/main_raytracing.cu:46: UNNORMALISED direction, normalised-units length).
#pragma once
#include <cfloat>
#include <cstdint>
#include "accel.h"

#ifndef RT_TW_BRANCHLESS
#define RT_TW_BRANCHLESS 0
#endif
#ifndef RT_TW_SPPTR
#define RT_TW_SPPTR 1   /* traversal stack pointer as a pointer (push = store + add, pop = load + add): 50.8 -> 50.3 ms */
#endif
#ifndef RT_VEC3_PACKED_MUL
#define RT_VEC3_PACKED_MUL 1   /* vec3 multiplies as FMUL2 (adds stay scalar), bit-identical: 51.3 -> 50.8 ms */
#endif
#ifndef RT_TW_TRI_PACKED
#define RT_TW_TRI_PACKED 1   /* tri_test in packed fp32 (FMUL2 / FADD2), bit-identical: 52.9 -> 51.3 ms */
#endif
#ifndef RT_TW_BOX_PREFILTER
#define RT_TW_BOX_PREFILTER 0
#endif
#ifndef RT_TW_RCP_APPROX
#define RT_TW_RCP_APPROX 1   /* reciprocal direction of the CONSERVATIVE node test by MUFU.RCP: 45.05 -> 44.92 ms (3 interleaved rounds, +-0.02), bit-identical */
#endif
#ifndef RT_TW_STOPREF
#define RT_TW_STOPREF 0   /* mode 7, NOT MEASURED yet (round-2 candidate): the end-of-ray marker is a valid leaf-flagged reference (offset 0), so the
                             loop needs no sentinel compare per iteration - only triangle-side iterations check for it; costs one extra fetch per ray */
#endif
#define RT_STOP_REF RT_LEAF_BIT
#ifndef RT_TW_TRI_SIGNFLIP
#define RT_TW_TRI_SIGNFLIP 1   /* one sign-free miss test instead of two divergent sign cases, bit-identical: 45.95 -> 45.02 ms */
#endif
#ifndef RT_TW_TRI_STAGED
#define RT_TW_TRI_STAGED 0   /* tri_test: u range test before the second triple product is evaluated: one more divergent branch, 54.9 vs 52.9 ms, off */
#endif

struct Vec3 { float x, y, z; };

__device__ __forceinline__ Vec3 mk3(float x, float y, float z) { Vec3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ Vec3 operator+(Vec3 a, Vec3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ Vec3 operator-(Vec3 a, Vec3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ Vec3 operator-(Vec3 a) { return mk3(-a.x, -a.y, -a.z); }
#if RT_VEC3_PACKED_MUL
// The MULTIPLIES of the vec3 algebra two at a time (Blackwell FMUL2 on the x and y lanes; IEEE round-to-nearest per lane, factors
// exchanged at most - exact), every ADD scalar and in GLM's order. Adds are deliberately NOT packed: ptxas fuses a packed multiply
// that feeds a packed add into FFMA2 even under -fmad=false (one rounding instead of two; tests/test_layout_and_scene.py guards
// the shipped SASS against any FFMA2).
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
#else
__device__ __forceinline__ Vec3 operator*(Vec3 a, Vec3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ Vec3 operator*(Vec3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ Vec3 operator*(float s, Vec3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ Vec3 operator/(Vec3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
// glm/detail/func_geometric.inl:48-54,68-78,82-89 ; func_exponential.inl:136-139
__device__ __forceinline__ float dot3(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ Vec3 cross3(Vec3 x, Vec3 y) { return mk3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
#endif
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
__device__ __forceinline__ const char* ref_record(const AccelView& A, uint32_t ref) { return A.arena + (ref & RT_REF_BYTE_MASK); }
__device__ __forceinline__ const WideNode* ref_wide(const AccelView& A, uint32_t ref) { return reinterpret_cast<const WideNode*>(ref_record(A, ref)); }
// byte offset -> slot: exact division by 48 = shift by 4, then multiplication by the inverse of 3 modulo 2^32 (the quotient is exact)
__device__ __forceinline__ uint32_t offset_slot(const AccelView& A, uint32_t offset) { return ((offset - A.tri_byte0) >> 4) * 0xAAAAAAABu; }
__device__ __forceinline__ uint32_t ref_slot(const AccelView& A, uint32_t ref) { return offset_slot(A, ref & RT_REF_BYTE_MASK); }
__device__ __forceinline__ uint32_t ref_count(uint32_t ref) { return (ref & RT_REF_COUNT_MASK) + 1u; }
__device__ __forceinline__ bool ref_has_more(uint32_t ref) { return (ref & RT_REF_COUNT_MASK) != 0u; }
__device__ __forceinline__ uint32_t ref_next_tri(uint32_t ref, uint32_t step = 1u) { return ref + 47u * step; }  // +48 bytes, -1 prim each; caller checked the count

// Math::IntersectAABB with the reference's operand roles. Order-independent use: ray_length = the
// distance the BVH search STARTED with (sphere hit or 1e30).
// kInclusive: `tmin <= ray_length` instead of the reference's `<` - used only by the lowest-id tie rule, where a node whose
// entry distance EQUALS the current best can still hold an exact-tie candidate with a lower id.
template <bool kInclusive = false>
__device__ __forceinline__ bool exact_box(Vec3 o, Vec3 d, const RtBVHNode* node, float ray_length)
{
    const float4 q0 = ldg4(node);                                    // min.xyz max.x
    const float4 q1 = ldg4(reinterpret_cast<const char*>(node) + 16);  // max.yz first count
#if RT_TW_BOX_PREFILTER
    {   // measured once (45.39 vs 45.02 ms, bit-identical), off: the six IEEE divisions below decide almost always far from their thresholds. The same slabs
        // with MUFU.RCP instead: each quotient fl(s * rcp(d)) is within 2^-22 of the exact test's fl(s / d) (same s = fl(plane - o);
        // rcp.approx: 2^-23, two roundings: 2^-24 each), and so are the entry / exit distances (a min/max picks another element only
        // where the two are that close). If the approximate test passes with 1e-6 to spare on every comparison the exact one passes.
        // Anything else - including every inf / NaN an axis-parallel or denormal component produces - takes the exact path.
        auto rcp_approx = [](float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; };
        const float rx = rcp_approx(d.x), ry = rcp_approx(d.y), rz = rcp_approx(d.z);
        const float ax1 = (q0.x - o.x) * rx, ax2 = (q0.w - o.x) * rx, ay1 = (q0.y - o.y) * ry, ay2 = (q1.x - o.y) * ry;
        const float az1 = (q0.z - o.z) * rz, az2 = (q1.y - o.z) * rz;
        const float lo = fmaxf(fmaxf(fminf(ax1, ax2), fminf(ay1, ay2)), fminf(az1, az2));
        const float hi = fminf(fminf(fmaxf(ax1, ax2), fmaxf(ay1, ay2)), fmaxf(az1, az2));
        const float lo_up = fmaf(1.0e-6f, fabsf(lo), lo) + 1e-30f, hi_dn = fmaf(-1.0e-6f, fabsf(hi), hi) - 1e-30f;
        if (hi_dn >= lo_up && lo_up < ray_length && hi_dn > 0.0f) return true;
    }
#endif
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

// glm::intersectRayTriangle on a TriRecord. Returns true with (t, bu, bv) when GLM would.
// kNegO: `o` holds the NEGATED ray origin (the packed slab test keeps -o in registers; using it here as well takes the origin
// itself out of the traversal loop's live set). Every intermediate is then the exact negation of GLM's (IEEE negation is exact
// and round-to-nearest is symmetric: fl(-a + b) = -fl(a - b), fl((-a) * b) = -fl(a * b)), the comparisons are mirrored, and the
// results are negated back: bit-identical t, u, v.
//
// twin_clear_miss (optional): every imported mesh face exists twice in the reference's scene (Scene.cpp:110-127: the smooth
// indexed face, then AddTriangle's flat copy with v1/v2 exchanged), and both copies always fall into the same leaf (same centroid).
// The copy's test is the SAME three scalar triple products with e1/e2 exchanged (det' = -det, u' = -v, v' = -u in exact
// arithmetic, on bit-identical inputs dist/nd/e1/e2), so when this record misses by more than the rounding error of both
// evaluations the copy misses too and need not be fetched. Rounding bound: a computed triple product c.(a x b) of fp32 vectors is
// within 30 eps |a|inf |b|inf |c|inf of the exact one (two roundings per cross component, three products + two sums in the dot),
// i.e. E <= 30 eps M (Tm + M) with M = max(|e1|inf, |e2|inf), Tm = |dist|inf, |nd|inf <= 1, eps = 2^-24. The copy is proven to
// miss when a range test fails by delta > 7 E (3 E for this evaluation's u, v, det, 3 E for the copy's, 1 E for the two u + v
// roundings) and |det| > delta + FLT_EPSILON (so the copy takes the opposite-sign branch); delta = 2^-14 M (Tm + M) = 1024 eps M (Tm + M)
// is five times that. Only rays within ~1 % of an edge (in barycentric units at the bunny's scale) are "not clear"; those fetch and
// test the copy as before, so results stay bit-identical (tests/test_gpu_parity.py).
template <bool kNegO = false>
__device__ __forceinline__ bool tri_test(Vec3 o, Vec3 nd, float4 r0, float4 r1, float4 r2, float& t, float& bu, float& bv,
                                         bool* twin_clear_miss = nullptr)
{
#if RT_TW_TRI_PACKED
    if (kNegO && !twin_clear_miss)
    {   // packed form on the NEGATED origin (the slab test's own -o pair): every intermediate that involves dist is the exact negation
        // of GLM's (round-to-nearest is sign-symmetric), the range tests are mirrored, and the division by -det hands back GLM's t, u, v
        const float2 e1xy = make_float2(r1.x, r1.y), e2xy = make_float2(r2.x, r2.y);
        const float2 nd_xy = make_float2(nd.x, nd.y), nd_yx = make_float2(nd.y, nd.x);
        const float2 pa = __fmul2_rn(e2xy, make_float2(nd.z, nd.z));
        const float2 pb = __fmul2_rn(e2xy, nd_yx);
        const float2 pc = __fmul2_rn(nd_yx, make_float2(r2.z, r2.z));
        const float2 pxy = make_float2(pc.x - pa.y, pa.x - pc.y);
        const float pz = pb.y - pb.x;
        const float2 dd = __fmul2_rn(e1xy, pxy);
        const float det = (dd.x + dd.y) + r1.z * pz;
        const float2 dxy = __fadd2_rn(make_float2(o.x, o.y), make_float2(r0.x, r0.y));  // (-origin) + v0 = -(origin - v0)
        const float dz = o.z + r0.z;
        const float2 uu = __fmul2_rn(dxy, pxy);
        const float un = (uu.x + uu.y) + dz * pz;                                        // -u
        const float2 d_yx = make_float2(dxy.y, dxy.x);
        const float2 qa = __fmul2_rn(e1xy, make_float2(dz, dz));
        const float2 qb = __fmul2_rn(e1xy, d_yx);
        const float2 qc = __fmul2_rn(d_yx, make_float2(r1.z, r1.z));
        const float2 qxy = make_float2(qc.x - qa.y, qa.x - qc.y);                         // -cross(dist, e1).xy
        const float qz = qb.y - qb.x;
        const float2 vv = __fmul2_rn(nd_xy, qxy);
        const float vn = (vv.x + vv.y) + nd.z * qz;                                       // -v
        const float ndet = -det;
        if (det > FLT_EPSILON)
        {
            if (un > 0.0f || un < ndet) return false;
            if (vn > 0.0f || (un + vn) < ndet) return false;
        }
        else if (det < -FLT_EPSILON)
        {
            if (un < 0.0f || un > ndet) return false;
            if (vn < 0.0f || (un + vn) > ndet) return false;
        }
        else return false;
        const float ninv = 1.0f / ndet;                                                   // = -(1 / det) exactly
        const float2 tt = __fmul2_rn(e2xy, qxy);
        t = ((tt.x + tt.y) + r2.z * qz) * ninv;
        bu = un * ninv; bv = vn * ninv;
        return true;
    }
    if (!kNegO && !twin_clear_miss)
    {   // the same products and sums, two multiplies / adds per instruction where operands pair up (Blackwell FMUL2 / FADD2, IEEE
        // round-to-nearest per lane): every product below is one of GLM's with its factors possibly exchanged (exact), every sum has
        // GLM's operands in GLM's order. Pairs: (x, y) of v0 / e1 / e2 straight from the 128-bit loads, (x, y) and (y, x) of the ray.
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
#if RT_TW_TRI_SIGNFLIP
        {   // GLM's two sign cases (intersect.inl:44-63) as one: for det < 0 its tests are the det > 0 tests of (-u, -v, -(u + v))
            // against -det, and flipping a sign bit is exact - same decisions (NaNs included), no data-dependent branch on the sign
            // of det (lanes of a warp see front and back faces mixed: the two cases used to run one after the other)
            const uint32_t flip = __float_as_uint(det) & 0x80000000u;
            const float ad = fabsf(det);
            const float us = __uint_as_float(__float_as_uint(u) ^ flip), vs = __uint_as_float(__float_as_uint(v) ^ flip);
            const float uvs = __uint_as_float(__float_as_uint(u + v) ^ flip);
            if (!(ad > FLT_EPSILON) || (us < 0.0f || us > ad) || (vs < 0.0f || uvs > ad)) return false;
        }
#else
        if (det > FLT_EPSILON)
        {
            if (u < 0.0f || u > det) return false;
            if (v < 0.0f || (u + v) > det) return false;
        }
        else if (det < -FLT_EPSILON)
        {
            if (u > 0.0f || u < det) return false;
            if (v > 0.0f || (u + v) < det) return false;
        }
        else return false;
#endif
        const float inv_det = 1.0f / det;
        const float2 tt = __fmul2_rn(e2xy, qxy);
        t = ((tt.x + tt.y) + r2.z * qz) * inv_det;
        bu = u * inv_det; bv = v * inv_det;
        return true;
    }
#endif
    const Vec3 v0 = mk3(r0.x, r0.y, r0.z), e1 = mk3(r1.x, r1.y, r1.z), e2 = mk3(r2.x, r2.y, r2.z);
    const Vec3 p = cross3(nd, e2);
    const float det = dot3(e1, p);
    if (!kNegO)
    {
        const Vec3 dist = o - v0;
        const float u = dot3(dist, p);
        const bool want_twin = twin_clear_miss && *twin_clear_miss;  // in: this record's copy sits in the next slot
        if (twin_clear_miss) *twin_clear_miss = false;               // out: the copy provably misses as well
#if RT_TW_TRI_STAGED
        // the u range test first, cross(dist, e1) and v only for the lanes that pass it (same operations, same roundings: GLM's own
        // order of the TESTS, intersect.inl:55-63; without this the compiler evaluates both triple products before the first branch)
        if (!want_twin)
        {
            if (det > FLT_EPSILON) { if (u < 0.0f || u > det) return false; }
            else if (det < -FLT_EPSILON) { if (u > 0.0f || u < det) return false; }
            else return false;
        }
#endif
        const Vec3 perp = cross3(dist, e1);
        const float v = dot3(nd, perp);
        bool miss;
        if (det > FLT_EPSILON) miss = (u < 0.0f || u > det) || (v < 0.0f || (u + v) > det);
        else if (det < -FLT_EPSILON) miss = (u > 0.0f || u < det) || (v > 0.0f || (u + v) < det);
        else return false;
        if (miss)
        {
            if (want_twin)
            {
                const float M = fmaxf(fmaxf(fmaxf(fabsf(e1.x), fabsf(e1.y)), fmaxf(fabsf(e1.z), fabsf(e2.x))), fmaxf(fabsf(e2.y), fabsf(e2.z)));
                const float Tm = fmaxf(fmaxf(fabsf(dist.x), fabsf(dist.y)), fabsf(dist.z));
                const float delta = (M * 6.103515625e-5f) * (Tm + M);
                const float ad = fabsf(det), us = det > 0.0f ? u : -u, vs = det > 0.0f ? v : -v;
                *twin_clear_miss = ad > delta + FLT_EPSILON && (fminf(us, vs) < -delta || fmaxf(us, us + vs) > ad + delta);
            }
            return false;
        }
        const float inv_det = 1.0f / det;
        t = dot3(e2, perp) * inv_det;
        bu = u * inv_det; bv = v * inv_det;
        return true;
    }
    else
    {
        const Vec3 ndist = o + v0;                 // = -(origin - v0)
        const float un = dot3(ndist, p);           // = -u
        const Vec3 nperp = cross3(ndist, e1);      // = -perp
        const float vn = dot3(nd, nperp);          // = -v
        const float ndet = -det;
        if (det > FLT_EPSILON)
        {
            if (un > 0.0f || un < ndet) return false;
            if (vn > 0.0f || (un + vn) < ndet) return false;
        }
        else if (det < -FLT_EPSILON)
        {
            if (un < 0.0f || un > ndet) return false;
            if (vn < 0.0f || (un + vn) > ndet) return false;
        }
        else return false;
        const float inv_det = 1.0f / det;
        t = -(dot3(e2, nperp) * inv_det);
        bu = -(un * inv_det); bv = -(vn * inv_det);
        return true;
    }
}

// Candidate acceptance: main_raytracing.cu:62 (`distance >= best || distance < 0` rejects). The
// reference keeps the FIRST-VISITED candidate on exact ties, which depends on traversal order; here
// ties are resolved by an order-independent key: the face id (kTieById, BASELINE.json's rule for the
// hit-id parity config) or the original BVH slot (= the reference's visiting order inside a leaf,
// used by the path tracer so shading picks the same duplicate the reference does).
__device__ __forceinline__ bool accept(float t, uint32_t key, const HitRecord& h)
{
    if (t > h.t || t < 0.0f) return false;
    return t < h.t || (key < h.key && !(h.id & RT_SPHERE_ID_BIT));  // a sphere at the same t keeps the hit (:62, `>=`)
}

// One triangle record (already loaded) against the ray: glm test, tie rule, optional per-candidate exact leaf-box check.
#if defined(RT_TW_MODE) && RT_TW_MODE == 8
static __device__ unsigned long long g_tw_hist[66];  // [n] iterations with n lanes alive; [33 + n] pending stack entries summed
#endif

template <bool kLazyExactBox, bool kTieById, bool kNegO = false>
__device__ __forceinline__ void leaf_test_one(const AccelView& A, float4 r0, float4 r1, float4 r2, uint32_t slot, Vec3 o, Vec3 d, Vec3 nd,
                                              float start_length, HitRecord& h, uint32_t* verified_leaf = nullptr, bool* twin_clear_miss = nullptr)
{
    static_assert(!(kLazyExactBox && kNegO), "the per-candidate leaf-box check needs the origin itself");
    float t, bu, bv;
    if (twin_clear_miss) *twin_clear_miss = (__float_as_uint(r1.w) & RT_TWIN_BIT) != 0u;
    if (!tri_test<kNegO>(o, nd, r0, r1, r2, t, bu, bv, twin_clear_miss)) return;
    const uint32_t id = __float_as_uint(r0.w);
    const uint32_t key = kTieById ? id : __float_as_uint(r2.w);
    if (!accept(t, key, h)) return;
    if (kLazyExactBox)
    {   // the reference only tests this triangle if its leaf's box passes (main_raytracing.cu:46); optional one-entry cache of
        // the last leaf that passed (the smooth/flat duplicates of a face share a leaf and are accepted back to back)
        const uint32_t leaf = __float_as_uint(r1.w) & RT_REF_LEAF_MASK;
        if (!verified_leaf || *verified_leaf != leaf)
        {
            if (!exact_box(o, d, A.ref_nodes + leaf, start_length)) return;
            if (verified_leaf) *verified_leaf = leaf;
        }
    }
    h.t = t; h.id = id; h.slot = slot; h.key = key; h.bu = bu; h.bv = bv;
}

// Same, for the deferred-check traversal with the hit details (id, slot, key, u, v) parked in shared memory: only the best t stays
// in a register across the traversal loop (the kernel runs at a 48-register cap; the details are written ~once per ray).
// sh points at this thread's column of a [5][blockDim] array; stride = blockDim.x.
template <bool kTieById>
__device__ __forceinline__ void leaf_test_one_parked(float4 r0, float4 r1, float4 r2, uint32_t slot, Vec3 o, Vec3 nd, float& best_t,
                                                     uint32_t* sh, uint32_t stride)
{
    float t, bu, bv;
    if (!tri_test(o, nd, r0, r1, r2, t, bu, bv)) return;
    if (t > best_t || t < 0.0f) return;
    const uint32_t id = __float_as_uint(r0.w);
    const uint32_t key = kTieById ? id : __float_as_uint(r2.w);
    if (!(t < best_t) && !(key < sh[2 * stride] && !(sh[0] & RT_SPHERE_ID_BIT))) return;  // exact tie: accept()'s key rule
    best_t = t;
    sh[0] = id; sh[stride] = slot; sh[2 * stride] = key; sh[3 * stride] = __float_as_uint(bu); sh[4 * stride] = __float_as_uint(bv);
}

template <bool kLazyExactBox, bool kTieById>
__device__ __forceinline__ void leaf_tests(const AccelView& A, uint32_t first, uint32_t count, Vec3 o, Vec3 d, Vec3 nd,
                                           float start_length, HitRecord& h, uint32_t* verified_leaf = nullptr)
{
    const TriRecord* tr = A.tris + first;
    for (uint32_t i = 0; i < count; i++)
    {
        const float4 r0 = ldg4(&tr[i]), r1 = ldg4(reinterpret_cast<const char*>(&tr[i]) + 16), r2 = ldg4(reinterpret_cast<const char*>(&tr[i]) + 32);
        leaf_test_one<kLazyExactBox, kTieById>(A, r0, r1, r2, first + i, o, d, nd, start_length, h, verified_leaf);
    }
}

#if RT_TW_CENTER
// Entry / exit distances of both child boxes of a centre + half-extent wide node (accel.h): per axis mid = (centre - o) * inv,
// entry = mid - half * |inv|, exit = mid + half * |inv| - twelve packed fma-pipe instructions and four 3-input min/max, no per-axis
// min/max pair. Conservative, not exact: mid carries ~1 ulp of |mid| <= |entry| + half |inv|; the half |inv| share is covered by
// the 2^-21 the builder adds to every half extent, the rest by the callers' relative widening (kWiden).
__device__ __forceinline__ void center_slabs(float4 a, float4 b, float4 c, float2 no_xy, float2 no_zz, float2 i_xy, float2 i_zz,
                                             float2 ai_xy, float2 ai_zz, float& mn0, float& mx0, float& mn1, float& mx1)
{
#if RT_TW_NODE48
    // 48-byte node (accel.h): a = centre0.xy centre1.xy, b = centre0.z centre1.z | two half-extent words, c.x = the z half-extent word
    // (c.y, c.z = child references). Of each word the high bf16 is used as the word stands, the low one as word << 16.
    const float2 m0 = __fmul2_rn(__fadd2_rn(make_float2(a.x, a.y), no_xy), i_xy);
    const float2 m1 = __fmul2_rn(__fadd2_rn(make_float2(a.z, a.w), no_xy), i_xy);
    const float2 mz = __fmul2_rn(__fadd2_rn(make_float2(b.x, b.y), no_zz), i_zz);
    const float2 h0 = make_float2(__uint_as_float(__float_as_uint(b.z) << 16), b.z);
    const float2 h1 = make_float2(__uint_as_float(__float_as_uint(b.w) << 16), b.w);
    const float2 hz = make_float2(__uint_as_float(__float_as_uint(c.x) << 16), c.x);
#else
    const float2 m0 = __fmul2_rn(__fadd2_rn(make_float2(a.x, a.y), no_xy), i_xy);
    const float2 m1 = __fmul2_rn(__fadd2_rn(make_float2(b.x, b.y), no_xy), i_xy);
    const float2 mz = __fmul2_rn(__fadd2_rn(make_float2(c.x, c.y), no_zz), i_zz);
    const float2 h0 = make_float2(a.z, a.w), h1 = make_float2(b.z, b.w), hz = make_float2(c.z, c.w);
#endif
#if RT_TW_NODE48 == 2
    // scalar fmas for the half-extent products: the decoded halves do not sit in aligned register pairs, and assembling the pairs costs
    // six moves per node step (same fma-pipe time either way)
    const float2 n0 = make_float2(fmaf(h0.x, -ai_xy.x, m0.x), fmaf(h0.y, -ai_xy.y, m0.y)), f0 = make_float2(fmaf(h0.x, ai_xy.x, m0.x), fmaf(h0.y, ai_xy.y, m0.y));
    const float2 n1 = make_float2(fmaf(h1.x, -ai_xy.x, m1.x), fmaf(h1.y, -ai_xy.y, m1.y)), f1 = make_float2(fmaf(h1.x, ai_xy.x, m1.x), fmaf(h1.y, ai_xy.y, m1.y));
    const float2 nzz = make_float2(fmaf(hz.x, -ai_zz.x, mz.x), fmaf(hz.y, -ai_zz.y, mz.y)), fzz = make_float2(fmaf(hz.x, ai_zz.x, mz.x), fmaf(hz.y, ai_zz.y, mz.y));
#else
    const float2 nai_xy = make_float2(-ai_xy.x, -ai_xy.y), nai_zz = make_float2(-ai_zz.x, -ai_zz.y);
    const float2 n0 = __ffma2_rn(h0, nai_xy, m0), f0 = __ffma2_rn(h0, ai_xy, m0);
    const float2 n1 = __ffma2_rn(h1, nai_xy, m1), f1 = __ffma2_rn(h1, ai_xy, m1);
    const float2 nzz = __ffma2_rn(hz, nai_zz, mz), fzz = __ffma2_rn(hz, ai_zz, mz);
#endif
    mn0 = fmaxf(fmaxf(n0.x, n0.y), nzz.x); mx0 = fminf(fminf(f0.x, f0.y), fzz.x);
    mn1 = fmaxf(fmaxf(n1.x, n1.y), nzz.y); mx1 = fminf(fminf(f1.x, f1.y), fzz.y);
}
#endif

// Product traversal: wide nodes, near-child-first, while-while, conservative box culls in
// normalised units, exact leaf box re-check only on candidate hits.
// `h` enters holding the sphere result (or a miss with t = 1e30).
// Tuning switches of trace_wide, A/B-measured on B200 (profiles/r01c_traversal_ab.txt); results are bit-identical either way.
//   kFma (template): slab test as fma(plane, inv, -(o*inv)) + an absolute slack instead of (plane - o) * inv.
//     Path-trace megakernel: 77.2 -> 74.9 ms/frame (72.3 with 5 resident CTAs/SM); bare trace_rays kernel: 5.13 -> 5.64 ms
//     (the four extra live registers cost more than the 12 saved FADDs) => on for the path tracer, off for the query kernels.
//   RT_TW_CACHE: one-entry cache of the last reference leaf whose exact box passed: slower (+2..3 %), off.
//   RT_TW_TOS: top of the traversal stack in a register: neutral to slower, off.
#ifndef RT_TW_CACHE
#define RT_TW_CACHE 0
#endif
#ifndef RT_TW_TOS
#define RT_TW_TOS 0
#endif
#ifndef RT_TW_FOLD
#define RT_TW_FOLD 1
#endif
#ifndef RT_TW_PREFETCH_CHILD
#define RT_TW_PREFETCH_CHILD 0
#endif
#ifndef RT_TW_SHSTACK
#define RT_TW_SHSTACK 0
#endif
#ifndef RT_TW_NEGO
#define RT_TW_NEGO 0
#endif
#ifndef RT_TW_PARK
#define RT_TW_PARK 0
#endif
#ifndef RT_TW_CULL
#define RT_TW_CULL 0   /* only wired into RT_TW_MODE 7 (tcut refresh) */
#endif
#ifndef RT_TW_PACKED
#define RT_TW_PACKED 1
#endif
#ifndef RT_TW_AXIS_SLACK
#define RT_TW_AXIS_SLACK 0
#endif
#ifndef RT_TW_PREFETCH
#define RT_TW_PREFETCH 0
#endif
#ifndef RT_TW_NODES
#define RT_TW_NODES 2
#endif
#ifndef RT_TW_BATCH
#define RT_TW_BATCH 8
#endif
#ifndef RT_TW_DEFER
#define RT_TW_DEFER 1
#endif
#ifndef RT_TW_MODE
#define RT_TW_MODE 7
#endif
#ifndef RT_TW_CARRY_PERIOD
#define RT_TW_CARRY_PERIOD 1
#endif
#ifndef RT_TW_LEAN
#define RT_TW_LEAN 0   /* mode 7, deferred: keep only (t, slot, key) of the best candidate in the loop; id, u, v are recomputed from the winner's record after it */
#endif
// RT_TW_TWIN (accel.h), mode 7: skip the duplicate of a mesh face when the first copy misses by a proven margin (tri_test): bit-identical,
// measured slower (57.9 vs 56.3 ms), off

//   kDeferred (template): the exact reference-leaf box re-check (6 IEEE divisions, ~110 instructions, executed by the ~2 lanes
//     that just accepted a candidate: 12 % of the megakernel's issue slots) is not done per accepted candidate but ONCE, on the
//     winner, after the loop. Unverified candidates only ever shrink the search interval, so any candidate that beats them is
//     still found; if the WINNER's leaf box fails (a ray grazing a leaf box within rounding - rare) the function returns false
//     and the caller redoes the ray with the per-candidate check (trace_wide below). Exact by construction.
// Suspend / resume of one ray's traversal (mode 7): when only `threshold` lanes of the warp are still inside the loop (and more
// than that entered it together), the stragglers leave with their state (current node reference, stack depth; the stack itself
// lives in the caller's array, the best hit so far in `h`) and finish the ray in the caller's NEXT round, side by side with the
// other lanes' new rays, instead of running the tail of this round alone. Per lane nothing changes: same steps, same order.
struct TraceCarry
{
    uint32_t* stack;     // caller-owned, RT_STACK_SIZE entries, untouched by the caller between suspend and resume
    uint32_t node;
    int sp;
    bool resume;         // in: continue the suspended ray (h, node, sp, stack as left)
    int threshold;       // suspend when at most this many lanes are left
};
enum : int { RT_TRACE_REDO = 0, RT_TRACE_DONE = 1, RT_TRACE_SUSPENDED = 2 };

template <bool kTieById, bool kFma, bool kDeferred>
__device__ __forceinline__ int trace_wide_impl(const AccelView& A, Vec3 o, Vec3 d, Vec3 nd, HitRecord& h, uint32_t* steps = nullptr,
                                               TraceCarry* carry = nullptr, float carry_start_length = 0.0f)
{
    const float start_length = (carry && carry->resume) ? carry_start_length : h.t;
    // reciprocal of the normalised direction; zero components -> huge finite (conservative)
    const float big = 1e30f;
#if RT_TW_RCP_APPROX
    // round-2 candidate, measured once (44.87 vs 45.02 ms, bit-identical: inside the noise): the reciprocal direction only feeds the conservative node test, whose relative widening (8 ulp)
    // has room for MUFU.RCP's ~1 ulp - three IEEE divisions (~25 instructions with their slow-path checks) per ray become three MUFUs
    auto rcp_approx = [](float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; };
    const float ix = nd.x != 0.0f ? rcp_approx(nd.x) : big, iy = nd.y != 0.0f ? rcp_approx(nd.y) : big, iz = nd.z != 0.0f ? rcp_approx(nd.z) : big;
#else
    const float ix = nd.x != 0.0f ? 1.0f / nd.x : big, iy = nd.y != 0.0f ? 1.0f / nd.y : big, iz = nd.z != 0.0f ? 1.0f / nd.z : big;
#endif
    const float kWiden = 1.0e-6f;
    // kFma: abs error of fma(plane, inv, c) is bounded by ~3 ulp of |o * inv|: `slack` covers it, kWiden the relative part
    const float ax = o.x * ix, ay = o.y * iy, az = o.z * iz;
    const float cx = -ax, cy = -ay, cz = -az;
    const float slack = kFma ? 4.0e-7f * fmaxf(fabsf(ax), fmaxf(fabsf(ay), fabsf(az))) + 1e-30f : 0.0f;
#define RT_SLAB(p, inv, c, oc) (kFma ? fmaf(p, inv, c) : (((p) - (oc)) * (inv)))
    uint32_t local_stack[RT_STACK_SIZE];
    uint32_t* const stack = carry ? carry->stack : local_stack;
    int sp = 0;
#if !RT_TW_SHSTACK
    stack[sp++] = (RT_TW_STOPREF && RT_TW_MODE == 7) ? RT_STOP_REF : RT_SENTINEL;
#endif
#if RT_TW_TOS || RT_TW_SHSTACK || !RT_TW_SPPTR
    uint32_t* sp_ptr = nullptr; (void)sp_ptr;  // only the pointer-stack form uses it; the (compiled-out) branchless block of RT_NODE_MATH names it
#endif
#if RT_TW_TOS
    uint32_t tos = RT_SENTINEL;
#define RT_PUSH(x) do { stack[sp++] = tos; tos = (x); } while (0)
#define RT_POP() (node = tos, tos = stack[--sp])
#elif RT_TW_SHSTACK
    // experiment (north-star "shared-memory short stack"): the first RT_TW_SHSTACK entries per thread live in shared memory
    // ([entry][thread], conflict-free), deeper ones in the local-memory array. Sized for 128-thread CTAs (the megakernel).
    __shared__ uint32_t sh_stack[RT_TW_SHSTACK * 128];
    uint32_t* const shs = sh_stack + (threadIdx.x & 127);
#define RT_PUSH(x) do { if (sp < RT_TW_SHSTACK) shs[sp * 128] = (x); else stack[sp - RT_TW_SHSTACK] = (x); sp++; } while (0)
#define RT_POP() do { --sp; node = sp < RT_TW_SHSTACK ? shs[sp * 128] : stack[sp - RT_TW_SHSTACK]; } while (0)
#elif RT_TW_SPPTR
    // stack pointer as a POINTER into the local array: push = store + add, pop = load + add (no index -> address LEA per access)
    uint32_t* sp_ptr = stack + sp;
#define RT_PUSH(x) do { *sp_ptr++ = (x); } while (0)
#define RT_POP() (node = *--sp_ptr)
#else
#define RT_PUSH(x) do { stack[sp++] = (x); } while (0)
#define RT_POP() (node = stack[--sp])
#endif
#if RT_TW_SHSTACK
    RT_PUSH(RT_SENTINEL);
#endif
    uint32_t node = A.root_ref;
    if (carry && carry->resume)
    {   // (stack[0] is the sentinel either way)
        node = carry->node; sp = carry->sp;
#if RT_TW_SPPTR && !RT_TW_TOS && !RT_TW_SHSTACK
        sp_ptr = stack + sp;
#endif
    }
    // suspend limit of this round: -1 (never) unless more lanes entered together than the threshold - a round that starts with
    // few lanes runs to completion, so every round finishes at least one ray
    const int carry_limit = (carry && __popc(__activemask()) > carry->threshold) ? carry->threshold : -1;
#if RT_TW_CACHE
    uint32_t verified = RT_SENTINEL;
    uint32_t* const verified_ptr = &verified;
#else
    uint32_t* const verified_ptr = nullptr;
#endif

    // RT_TW_FOLD: the three culls (exit >= entry, entry <= best, exit > 0) as ONE compare max(entry, 0) <= widen(min(exit, best)):
    // k1 = (1 + kWiden) / (1 - kWiden) and k2 = 2 slack / (1 - kWiden), both rounded up, move the whole widening onto the exit side,
    // so the entry side needs none (a superset of the two-sided form term by term; -6 instructions per wide node).
    const float k1 = 1.0f + 2.5e-6f, k2 = 2.1f * slack;
    const float kExitWiden = 1.0f + 3.0e-6f, kBestWiden = 1.0f + 2.0e-6f;
    float tcut = h.t * kBestWiden;  // RT_TW_CULL: refreshed whenever a triangle step may have changed h.t
    // RT_TW_AXIS_SLACK (kFma only): the absolute error of fma(plane, inv, -(o*inv)) is ~ulp(|o*inv|) of THAT axis, and |o*inv| is
    // huge exactly on the axes the ray is nearly perpendicular to. One slack = the max over the axes (above) makes every entry/exit
    // that wrong: on the 13.9 M-triangle grid (coordinates ~300, many near-axis-parallel rays) it is larger than the leaves and the
    // cull against the current best stops working (path trace 23.6 ms vs 14.5 ms with the exact (plane - o) * inv form). Here each
    // axis carries its own slack inside its fma constant: entry planes use c - eps, exit planes c + eps. Which plane is the entry
    // one follows from the sign of inv (fma is monotonic in the plane), so the min/max pairs become selects: same instruction count.
    const float2 i_xy = make_float2(ix, iy), i_zz = make_float2(iz, iz), no_xy = make_float2(-o.x, -o.y), no_zz = make_float2(-o.z, -o.z);
    const float2 c_xy = make_float2(cx, cy), c_zz = make_float2(cz, cz);
#if RT_TW_CENTER
    const float2 ai_xy = make_float2(fabsf(ix), fabsf(iy)), ai_zz = make_float2(fabsf(iz), fabsf(iz));
#endif
    const bool neg_x = ix < 0.0f, neg_y = iy < 0.0f, neg_z = iz < 0.0f;
    const float ex = 4.0e-7f * fabsf(ax) + 1e-30f, ey = 4.0e-7f * fabsf(ay) + 1e-30f, ez = 4.0e-7f * fabsf(az) + 1e-30f;
    const float cnx = cx - ex, cny = cy - ey, cnz = cz - ez, cfx = cx + ex, cfy = cy + ey, cfz = cz + ez;
#define RT_NODE_STEP()                                                                                                     \
    do {                                                                                                                    \
        const WideNode* w = ref_wide(A, node);                                                                              \
        const float4 a = ldg4(w), b = ldg4(reinterpret_cast<const char*>(w) + 16), c = ldg4(reinterpret_cast<const char*>(w) + 32); \
        const uint2 ch = RT_TW_NODE48 ? make_uint2(__float_as_uint(c.y), __float_as_uint(c.z))                              \
                                      : __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const char*>(w) + 48));       \
        RT_NODE_MATH(a, b, c, ch);                                                                                          \
    } while (0)
#if RT_TW_CENTER
#define RT_CENTER_SLABS(a, b, c) if (true) center_slabs(a, b, c, no_xy, no_zz, i_xy, i_zz, ai_xy, ai_zz, mn0, mx0, mn1, mx1); else
#else
#define RT_CENTER_SLABS(a, b, c)
#endif
#define RT_NODE_MATH(a, b, c, ch)                                                                                          \
    do {                                                                                                                    \
        if (RT_TW_PREFETCH_CHILD)                                                                                           \
        {   /* experiment: start the L2->L1 fetch of both children's records while this node's boxes are tested */        \
            const char* c0 = ref_record(A, ch.x); \
            const char* c1 = ref_record(A, ch.y); \
            asm volatile("prefetch.global.L1 [%0];" ::"l"(c0));                                                             \
            asm volatile("prefetch.global.L1 [%0];" ::"l"(c1));                                                             \
        }                                                                                                                   \
        if (kFma && RT_TW_AXIS_SLACK)                                                                                       \
        {                                                                                                                   \
            const float en0 = fmaxf(fmaxf(fmaf(neg_x ? a.z : a.x, ix, cnx), fmaf(neg_y ? a.w : a.y, iy, cny)), fmaf(neg_z ? c.y : c.x, iz, cnz)); \
            const float ex0 = fminf(fminf(fmaf(neg_x ? a.x : a.z, ix, cfx), fmaf(neg_y ? a.y : a.w, iy, cfy)), fmaf(neg_z ? c.x : c.y, iz, cfz)); \
            const float en1 = fmaxf(fmaxf(fmaf(neg_x ? b.z : b.x, ix, cnx), fmaf(neg_y ? b.w : b.y, iy, cny)), fmaf(neg_z ? c.w : c.z, iz, cnz)); \
            const float ex1 = fminf(fminf(fmaf(neg_x ? b.x : b.z, ix, cfx), fmaf(neg_y ? b.y : b.w, iy, cfy)), fmaf(neg_z ? c.z : c.w, iz, cfz)); \
            const bool h0 = fmaxf(en0, 0.0f) <= fminf(ex0, h.t) * k1, h1 = fmaxf(en1, 0.0f) <= fminf(ex1, h.t) * k1;        \
            if (h0 && h1)                                                                                                   \
            {                                                                                                               \
                const bool swap = en1 < en0;                                                                                \
                RT_PUSH(swap ? ch.x : ch.y);                                                                                \
                node = swap ? ch.y : ch.x;                                                                                  \
            }                                                                                                               \
            else if (h0) node = ch.x;                                                                                       \
            else if (h1) node = ch.y;                                                                                       \
            else RT_POP();                                                                                                  \
            break;                                                                                                          \
        }                                                                                                                   \
        float mn0, mx0, mn1, mx1;                                                                                           \
        RT_CENTER_SLABS(a, b, c)                                                                                            \
        if (RT_TW_PACKED)                                                                                                   \
        {   /* Blackwell packed fp32: six 2-wide ops per slab stage instead of twelve scalar ones; same IEEE results */     \
            float2 d0, d1, d2, d3, d4, d5;                                                                                  \
            if (kFma)                                                                                                       \
            {                                                                                                               \
                d0 = __ffma2_rn(make_float2(a.x, a.y), i_xy, c_xy); d1 = __ffma2_rn(make_float2(a.z, a.w), i_xy, c_xy);     \
                d2 = __ffma2_rn(make_float2(b.x, b.y), i_xy, c_xy); d3 = __ffma2_rn(make_float2(b.z, b.w), i_xy, c_xy);     \
                d4 = __ffma2_rn(make_float2(c.x, c.y), i_zz, c_zz); d5 = __ffma2_rn(make_float2(c.z, c.w), i_zz, c_zz);     \
            }                                                                                                               \
            else                                                                                                            \
            {                                                                                                               \
                d0 = __fmul2_rn(__fadd2_rn(make_float2(a.x, a.y), no_xy), i_xy); d1 = __fmul2_rn(__fadd2_rn(make_float2(a.z, a.w), no_xy), i_xy);\
                d2 = __fmul2_rn(__fadd2_rn(make_float2(b.x, b.y), no_xy), i_xy); d3 = __fmul2_rn(__fadd2_rn(make_float2(b.z, b.w), no_xy), i_xy);\
                d4 = __fmul2_rn(__fadd2_rn(make_float2(c.x, c.y), no_zz), i_zz); d5 = __fmul2_rn(__fadd2_rn(make_float2(c.z, c.w), no_zz), i_zz);\
            }                                                                                                               \
            mn0 = fmaxf(fmaxf(fminf(d0.x, d1.x), fminf(d0.y, d1.y)), fminf(d4.x, d4.y));                                    \
            mx0 = fminf(fminf(fmaxf(d0.x, d1.x), fmaxf(d0.y, d1.y)), fmaxf(d4.x, d4.y));                                    \
            mn1 = fmaxf(fmaxf(fminf(d2.x, d3.x), fminf(d2.y, d3.y)), fminf(d5.x, d5.y));                                    \
            mx1 = fminf(fminf(fmaxf(d2.x, d3.x), fmaxf(d2.y, d3.y)), fmaxf(d5.x, d5.y));                                    \
        }                                                                                                                   \
        else                                                                                                                \
        {                                                                                                                   \
            float t1 = RT_SLAB(a.x, ix, cx, o.x), t2 = RT_SLAB(a.z, ix, cx, o.x);                                               \
            mn0 = fminf(t1, t2); mx0 = fmaxf(t1, t2);                                                                         \
            t1 = RT_SLAB(a.y, iy, cy, o.y); t2 = RT_SLAB(a.w, iy, cy, o.y);                                                     \
            mn0 = fmaxf(mn0, fminf(t1, t2)); mx0 = fminf(mx0, fmaxf(t1, t2));                                                   \
            t1 = RT_SLAB(c.x, iz, cz, o.z); t2 = RT_SLAB(c.y, iz, cz, o.z);                                                     \
            mn0 = fmaxf(mn0, fminf(t1, t2)); mx0 = fminf(mx0, fmaxf(t1, t2));                                                   \
            t1 = RT_SLAB(b.x, ix, cx, o.x); t2 = RT_SLAB(b.z, ix, cx, o.x);                                                     \
            mn1 = fminf(t1, t2); mx1 = fmaxf(t1, t2);                                                                         \
            t1 = RT_SLAB(b.y, iy, cy, o.y); t2 = RT_SLAB(b.w, iy, cy, o.y);                                                     \
            mn1 = fmaxf(mn1, fminf(t1, t2)); mx1 = fminf(mx1, fmaxf(t1, t2));                                                   \
            t1 = RT_SLAB(c.z, iz, cz, o.z); t2 = RT_SLAB(c.w, iz, cz, o.z);                                                     \
            mn1 = fmaxf(mn1, fminf(t1, t2)); mx1 = fminf(mx1, fmaxf(t1, t2));                                                   \
        }                                                                                                                   \
        bool hit0, hit1;                                                                                                    \
        if (RT_TW_FOLD && kFma)                                                                                             \
        {                                                                                                                   \
            hit0 = fmaxf(mn0, 0.0f) <= fmaf(fminf(mx0, h.t), k1, k2);                                                       \
            hit1 = fmaxf(mn1, 0.0f) <= fmaf(fminf(mx1, h.t), k1, k2);                                                       \
        }                                                                                                                   \
        else if (RT_TW_CULL && !kFma)                                                                                      \
        {   /* exact slab form: only relative errors. exit * (1+3k) >= entry, entry <= best * (1+2k) (= tcut, kept per ray),    \
               exit > 0: term by term a superset of the two-sided widening below, 4 instead of 7 instructions per child */  \
            hit0 = mx0 * kExitWiden >= mn0 && mn0 <= tcut && mx0 > 0.0f;                                                    \
            hit1 = mx1 * kExitWiden >= mn1 && mn1 <= tcut && mx1 > 0.0f;                                                    \
        }                                                                                                                   \
        else                                                                                                                \
        {   /* widen (superset of the reference's divisions), cull against the current best */                             \
            const float lo0 = fmaf(-kWiden, fabsf(mn0), mn0) - slack, hi0 = fmaf(kWiden, fabsf(mx0), mx0) + slack;          \
            const float lo1 = fmaf(-kWiden, fabsf(mn1), mn1) - slack, hi1 = fmaf(kWiden, fabsf(mx1), mx1) + slack;          \
            hit0 = hi0 >= lo0 && lo0 <= h.t && hi0 > 0.0f;                                                                  \
            hit1 = hi1 >= lo1 && lo1 <= h.t && hi1 > 0.0f;                                                                  \
        }                                                                                                                   \
        if (RT_TW_BRANCHLESS && RT_TW_SPPTR && !RT_TW_TOS && !RT_TW_SHSTACK)                                               \
        {   /* one select chain instead of three short divergent branches: the far child is stored unconditionally and kept   \
               (pointer advanced) only when both boxes were hit; the slot above the top is always inside the array */        \
            const bool take1 = hit1 && (!hit0 || mn1 < mn0);                                                                \
            const uint32_t nxt = take1 ? ch.y : ch.x;                                                                       \
            *sp_ptr = take1 ? ch.x : ch.y;                                                                                  \
            sp_ptr += (hit0 && hit1) ? 1 : 0;                                                                               \
            if (hit0 || hit1) node = nxt;                                                                                   \
            else RT_POP();                                                                                                  \
            break;                                                                                                          \
        }                                                                                                                   \
        if (hit0 && hit1)                                                                                                   \
        {                                                                                                                   \
            const bool swap = mn1 < mn0;                                                                                    \
            RT_PUSH(swap ? ch.x : ch.y);                                                                                    \
            node = swap ? ch.y : ch.x;                                                                                      \
        }                                                                                                                   \
        else if (hit0) node = ch.x;                                                                                         \
        else if (hit1) node = ch.y;                                                                                         \
        else RT_POP();                                                                                                      \
    } while (0)
#define RT_LEAF_STEP(ref) leaf_tests<!kDeferred, kTieById>(A, ref_slot(A, (ref)), ref_count(ref), o, d, nd, start_length, h, verified_ptr)

#if RT_TW_PARK && RT_TW_MODE == 7
    __shared__ uint32_t park_store[5 * 256];  // blocks of this library's traversal kernels have at most 256 threads
    uint32_t* const park = park_store + threadIdx.x;
    const uint32_t park_stride = blockDim.x;
    if (kDeferred)
    {
        park[0] = h.id; park[park_stride] = h.slot; park[2 * park_stride] = h.key;
        park[3 * park_stride] = __float_as_uint(h.bu); park[4 * park_stride] = __float_as_uint(h.bv);
    }
#else
    uint32_t* const park = nullptr; const uint32_t park_stride = 0;
#endif
#if RT_TW_MODE == 1
    // if-if: one node of either kind per iteration, no inner loop to wait in
    while (node != RT_SENTINEL)
    {
        if (!(node & RT_LEAF_BIT)) RT_NODE_STEP();
        else { RT_LEAF_STEP(node); RT_POP(); }
    }
#elif RT_TW_MODE == 3
    // if-if, one TRIANGLE per iteration: the leaf reference is re-encoded with the remaining range, no extra state
    while (node != RT_SENTINEL)
    {
        if (!(node & RT_LEAF_BIT)) RT_NODE_STEP();
        else
        {
            leaf_tests<!kDeferred, kTieById>(A, ref_slot(A, node), 1u, o, d, nd, start_length, h, verified_ptr);
            if (ref_has_more(node)) node = ref_next_tri(node);
            else RT_POP();
        }
    }
#elif RT_TW_MODE == 4
    // mode 3 + warp vote: lanes that sit on a leaf wait until RT_TW_BATCH of them do (or nobody has an inner node left), so the
    // triangle code runs with more lanes. All lanes that entered together stay in the loop until the last one is done.
    {
        const unsigned entry = __activemask();
        for (;;)
        {
            const bool inner = !(node & RT_LEAF_BIT);                       // (the sentinel has the leaf bit set)
            const bool leaf = (node & RT_LEAF_BIT) && node != RT_SENTINEL;
            const unsigned m_inner = __ballot_sync(entry, inner), m_leaf = __ballot_sync(entry, leaf);
            if ((m_inner | m_leaf) == 0u) break;
            if (inner) RT_NODE_STEP();
            if (leaf && (__popc(m_leaf) >= RT_TW_BATCH || m_inner == 0u))
            {
                leaf_tests<!kDeferred, kTieById>(A, ref_slot(A, node), 1u, o, d, nd, start_length, h, verified_ptr);
                if (ref_has_more(node)) node = ref_next_tri(node);
                else RT_POP();
            }
        }
    }
#elif RT_TW_MODE == 7
    // mode 3 with ONE fetch for both kinds of step: every lane loads 56 bytes from its own record (a 64-byte wide node or a
    // 48-byte triangle record; the arrays are padded so the last triangle's +48 read stays inside) before the kinds diverge, so
    // the memory latency of the node lanes and of the triangle lanes overlaps instead of adding up within an iteration.
#if RT_TW_LEAN
    uint32_t win_slot = RT_SENTINEL, win_key = 0u;  // kDeferred only: the best TRIANGLE so far (h.t follows; id/slot/key/u/v of h stay the entry's until the end)
#endif
#if RT_TW_STOPREF
#if RT_TW_SHSTACK || RT_TW_CARRY_PERIOD != 1
#error "RT_TW_STOPREF: only with the local-memory stack and without the unrolled suspend vote"
#endif
    if (node == RT_SENTINEL) node = RT_STOP_REF;  // empty scene
    for (;;)
#else
    while (node != RT_SENTINEL)
#endif
    {
        if (carry && __popc(__activemask()) <= carry_limit)
        {
#if RT_TW_SPPTR && !RT_TW_TOS && !RT_TW_SHSTACK
            sp = (int)(sp_ptr - stack);
#endif
            carry->node = node; carry->sp = sp;
            return RT_TRACE_SUSPENDED;
        }
#if RT_TW_CARRY_PERIOD > 1
      // the suspend vote costs ~7 instructions: take it once per RT_TW_CARRY_PERIOD steps (unrolled copies of the step)
#pragma unroll
      for (int carry_u = 0; carry_u < (carry ? RT_TW_CARRY_PERIOD : 1); carry_u++)
      {
        if (carry_u && node == RT_SENTINEL) break;
#endif
        if (steps) (*steps)++;  // per-lane work measure (node + triangle steps) for cost-binned pixel scheduling
        const bool is_leaf = (node & RT_LEAF_BIT) != 0u;
        const uint32_t slot = node & RT_REF_BYTE_MASK;  // arena byte offset; a winner's h.slot is converted to its slot after the loop
        const char* rec = A.arena + slot;                // one address form for wide nodes and triangle records
        const float4 q0 = ldg4(rec), q1 = ldg4(rec + 16), q2 = ldg4(rec + 32);
#if RT_TW_NODE48
        const uint2 q3 = make_uint2(__float_as_uint(q2.y), __float_as_uint(q2.z));  // 48 bytes for either kind: no fourth load
#else
        const uint2 q3 = __ldg(reinterpret_cast<const uint2*>(rec + 48));
#endif
#if RT_TW_PREFETCH
        if (is_leaf && ref_has_more(node))
        {   // the next record of this leaf is the next iteration's fetch: start it now (matters when the scene is not L2-resident)
            asm volatile("prefetch.global.L1 [%0];" ::"l"(rec + 64));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(rec + 95));
        }
#endif
        if (!is_leaf) RT_NODE_MATH(q0, q1, q2, q3);
        else
        {
#if RT_TW_STOPREF
            if (node == RT_STOP_REF) break;  // popped the stack bottom (the fetch above read the first wide node: valid memory, unused)
#endif
#if RT_TW_LEAN
            if (kDeferred)
            {
                float t, bu, bv;
                const Vec3 oo = RT_TW_NEGO ? mk3(no_xy.x, no_xy.y, no_zz.x) : o;
                if (tri_test<RT_TW_NEGO != 0>(oo, nd, q0, q1, q2, t, bu, bv) && !(t > h.t || t < 0.0f))
                {
                    const uint32_t key = kTieById ? __float_as_uint(q0.w) : __float_as_uint(q2.w);
                    if (t < h.t || (win_slot != RT_SENTINEL && key < win_key)) { h.t = t; win_slot = slot; win_key = key; }
                }
                if (ref_has_more(node)) node = ref_next_tri(node);
                else RT_POP();
                continue;
            }
#endif
            if (RT_TW_PARK && kDeferred) leaf_test_one_parked<kTieById>(q0, q1, q2, slot, o, nd, h.t, park, park_stride);
            else if (RT_TW_NEGO && kDeferred) leaf_test_one<false, kTieById, true>(A, q0, q1, q2, slot, mk3(no_xy.x, no_xy.y, no_zz.x), d, nd, start_length, h, verified_ptr);
#if RT_TW_TWIN
            else
            {   // a record whose copy (Scene.cpp:110-127) is the next slot of this leaf and provably misses too: skip the copy
                bool skip_twin;
                leaf_test_one<!kDeferred, kTieById>(A, q0, q1, q2, slot, o, d, nd, start_length, h, verified_ptr, &skip_twin);
                tcut = h.t * kBestWiden;
                const uint32_t step = skip_twin ? 2u : 1u;
                if ((node & RT_REF_COUNT_MASK) >= step) node = ref_next_tri(node, step);
                else RT_POP();
                continue;
            }
#else
            else leaf_test_one<!kDeferred, kTieById>(A, q0, q1, q2, slot, o, d, nd, start_length, h, verified_ptr);
#endif
            tcut = h.t * kBestWiden;
            if (ref_has_more(node)) node = ref_next_tri(node);
            else RT_POP();
        }
#if RT_TW_CARRY_PERIOD > 1
      }
#endif
    }
#if RT_TW_LEAN
    if (kDeferred && win_slot != RT_SENTINEL)
    {   // the winner's details, once per ray: the same arithmetic on the same record gives the same t, u, v bit for bit
        const char* rec = A.arena + win_slot;
        const float4 q0 = ldg4(rec), q1 = ldg4(rec + 16), q2 = ldg4(rec + 32);
        float t, bu, bv;
        tri_test(o, nd, q0, q1, q2, t, bu, bv);
        h.id = __float_as_uint(q0.w); h.slot = win_slot; h.key = win_key; h.bu = bu; h.bv = bv;
    }
#endif
    if (RT_TW_PARK && kDeferred)
    {
        h.id = park[0]; h.slot = park[park_stride]; h.key = park[2 * park_stride];
        h.bu = __uint_as_float(park[3 * park_stride]); h.bv = __uint_as_float(park[4 * park_stride]);
    }
#elif RT_TW_MODE == 8
    // measurement build (benchmarks/lane_histogram.py): mode 7 stepped warp-synchronously, with a histogram of how many lanes
    // still have a ray per iteration and of their pending stack entries. Not a product mode.
    {
        const unsigned entry = __activemask();
        for (;;)
        {
            const bool busy = node != RT_SENTINEL;
            const unsigned alive = __ballot_sync(entry, busy);
            if (alive == 0u) break;
            const unsigned pending = __reduce_add_sync(entry, busy ? (unsigned)(sp - 1) : 0u);
            if ((threadIdx.x & 31) == (__ffs(entry) - 1))
            {
                atomicAdd(&g_tw_hist[__popc(alive)], 1ull);
                atomicAdd(&g_tw_hist[33 + __popc(alive)], (unsigned long long)pending);
            }
            if (busy)
            {
                const bool is_leaf = (node & RT_LEAF_BIT) != 0u;
                const uint32_t slot = node & RT_REF_BYTE_MASK;
                const char* rec = A.arena + slot;
                const float4 q0 = ldg4(rec), q1 = ldg4(rec + 16), q2 = ldg4(rec + 32);
                const uint2 q3 = RT_TW_NODE48 ? make_uint2(__float_as_uint(q2.y), __float_as_uint(q2.z)) : __ldg(reinterpret_cast<const uint2*>(rec + 48));
                if (!is_leaf) RT_NODE_MATH(q0, q1, q2, q3);
                else
                {
                    leaf_test_one<!kDeferred, kTieById>(A, q0, q1, q2, slot, o, d, nd, start_length, h, verified_ptr);
                    if (ref_has_more(node)) node = ref_next_tri(node);
                    else RT_POP();
                }
            }
        }
    }
#elif RT_TW_MODE == 6
    // if-if, up to RT_TW_NODES wide nodes then one triangle per iteration (amortises the serialised triangle path)
    while (node != RT_SENTINEL)
    {
        if (!(node & RT_LEAF_BIT))
        {
            RT_NODE_STEP();
#pragma unroll
            for (int k = 1; k < RT_TW_NODES; k++) if (!(node & RT_LEAF_BIT)) RT_NODE_STEP();
        }
        else
        {
            leaf_tests<!kDeferred, kTieById>(A, ref_slot(A, node), 1u, o, d, nd, start_length, h, verified_ptr);
            if (ref_has_more(node)) node = ref_next_tri(node);
            else RT_POP();
        }
    }
#elif RT_TW_MODE == 5
    // if-if with up to two wide nodes per iteration (a 2-triangle leaf costs about two node steps)
    while (node != RT_SENTINEL)
    {
        if (!(node & RT_LEAF_BIT))
        {
            RT_NODE_STEP();
            if (node != RT_SENTINEL && !(node & RT_LEAF_BIT)) RT_NODE_STEP();
        }
        else { RT_LEAF_STEP(node); RT_POP(); }
    }
#elif RT_TW_MODE == 2
    // while-while with one postponed leaf: a lane that reaches a leaf keeps descending until it holds two (or runs dry)
    while (node != RT_SENTINEL)
    {
        uint32_t pending = RT_SENTINEL;
        while (true)
        {
            if (node == RT_SENTINEL) break;
            if (node & RT_LEAF_BIT)
            {
                if (pending != RT_SENTINEL) break;
                pending = node; RT_POP();
                continue;
            }
            RT_NODE_STEP();
        }
        if (pending != RT_SENTINEL) RT_LEAF_STEP(pending);
        if (node != RT_SENTINEL) { RT_LEAF_STEP(node); RT_POP(); }
    }
#else
    while (node != RT_SENTINEL)
    {
        while (!(node & RT_LEAF_BIT)) RT_NODE_STEP();
        if (node == RT_SENTINEL) break;
        RT_LEAF_STEP(node);
        RT_POP();
    }
#endif
#if RT_TW_MODE == 7 || RT_TW_MODE == 8
    // h entered as a sphere hit or a miss, so a triangle id means "found in this loop": its h.slot is still an arena byte offset
    if (h.id != RT_MISS_ID && !(h.id & RT_SPHERE_ID_BIT)) h.slot = offset_slot(A, h.slot);
#endif
#undef RT_NODE_STEP
#undef RT_NODE_MATH
#undef RT_LEAF_STEP
#undef RT_SLAB
#undef RT_PUSH
#undef RT_POP
    if (kDeferred && h.id != RT_MISS_ID && !(h.id & RT_SPHERE_ID_BIT))
    {   // h entered as a sphere hit or a miss, so a triangle id means "found here": verify its reference leaf (main_raytracing.cu:46)
        const uint32_t leaf = __ldg(&A.tris[h.slot].ref_leaf) & RT_REF_LEAF_MASK;
#ifdef RT_TW_FORCE_REDO
        if ((h.slot & RT_TW_FORCE_REDO) == 0u) return RT_TRACE_REDO;  // test hook: exercise the redo path on a fraction of the rays
#endif
        return exact_box(o, d, A.ref_nodes + leaf, start_length) ? RT_TRACE_DONE : RT_TRACE_REDO;
    }
    return RT_TRACE_DONE;
}

template <bool kTieById, bool kFma>
__device__ __noinline__ void trace_wide_checked(const AccelView& A, Vec3 o, Vec3 d, Vec3 nd, HitRecord& h)
{
    trace_wide_impl<kTieById, kFma, false>(A, o, d, nd, h);
}

// `h` enters holding what reinit(h) produces (a miss, or the sphere result); reinit is only called again on the rare redo.
template <bool kTieById, bool kFma = false, class Reinit>
__device__ __forceinline__ void trace_wide(const AccelView& A, Vec3 o, Vec3 d, Vec3 nd, HitRecord& h, Reinit reinit, uint32_t* steps = nullptr)
{
#if RT_TW_DEFER
    if (trace_wide_impl<kTieById, kFma, true>(A, o, d, nd, h, steps) != RT_TRACE_REDO) return;
    reinit(h);
    trace_wide_checked<kTieById, kFma>(A, o, d, nd, h);
#else
    trace_wide_impl<kTieById, kFma, false>(A, o, d, nd, h);
#endif
}

// trace_wide with suspend / resume (TraceCarry). Returns false when the ray was suspended: the caller keeps `h`, `carry` and
// `start_length` untouched and calls again with carry.resume = true (and the same o, d, nd). Deferred leaf-box check as above.
template <bool kTieById, bool kFma = false, class Reinit>
__device__ __forceinline__ bool trace_wide_resumable(const AccelView& A, Vec3 o, Vec3 d, Vec3 nd, HitRecord& h, Reinit reinit,
                                                     TraceCarry& carry, float& start_length)
{
    if (!carry.resume) start_length = h.t;
    const int r = trace_wide_impl<kTieById, kFma, true>(A, o, d, nd, h, nullptr, &carry, start_length);
    if (r == RT_TRACE_SUSPENDED) return false;
    if (r == RT_TRACE_REDO) { reinit(h); trace_wide_checked<kTieById, kFma>(A, o, d, nd, h); }
    return true;
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
