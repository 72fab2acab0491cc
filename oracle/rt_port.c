/*
 * TEST INFRASTRUCTURE - plain-C restatement ("port") of the reference's path-trace hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this; the product
 * (cuda-raytracing_b200/) never does.
 *
 * Parity status: PINNED. Every function here is checked bit for bit against the reference's own code
 * compiled host-side (oracle/_ref: unmodified BVHRayHit/GetRayHit/raytracing_kernel_main + the toolkit's
 * cuRAND header) by tests/test_oracle_port.py, and against the committed golden vectors those produced.
 * Exception, stated in DESIGN.md: the cubemap fetch (hardware filter) is restated from the CUDA
 * programming guide and is "parity unpinned" against hardware - images are compared by RMSE.
 *
 * Build: oracle/build_port.sh -> oracle/_build/librt_port.so, gcc -O2 -ffp-contract=off (plain IEEE
 * mul/add, no FMA: the same contraction policy as oracle/_ref and the -fmad=false device code).
 *
 * What follows what (reference file:line):
 *   port_intersect_aabb      RayTracing/Math.h:50-61 (host build: std::min/std::max, utils/CUDAHelper.h:39-40)
 *   port_intersect_triangle  include/glm/gtx/intersect.inl:29-94
 *   port_intersect_sphere    include/glm/gtx/intersect.inl:135-153
 *   port_bvh_hit             RayTracing/main_raytracing.cu:33-81   (+ a face id, + work counters)
 *   port_get_ray_hit         RayTracing/main_raytracing.cu:83-109
 *   port_ray_color           RayTracing/main_raytracing.cu:111-160
 *   port_render_pixel        RayTracing/main_raytracing.cu:162-200
 *   xorwow_*                 curand_kernel.h:863-874 (curand), :742-790 (curand_init), curand_uniform.h:69-72
 *   port_point_on_sphere     RayTracing/Random.h:23-31
 *   port_ambient_occlusion   BASELINE.json config 3: primary hit at the pixel centre, then ao_rays rays in the
 *                            reference's own diffuse lobe normalize(n + S) from p + 0.01 n (main_raytracing.cu:129,135)
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define PORT_API __attribute__((visibility("default")))

typedef struct { float x, y, z; } v3;
typedef struct { float x, y, z, w; } v4;

/* reference-layout PODs (RayTracing/GPUScene.h:25-74), byte-compatible */
typedef struct { v3 position; v3 normal; float uv[2]; } PVertex;                       /* 32 B */
typedef struct { uint32_t v0, v2, v1, material; } PFace;                               /* 16 B, order v0,v2,v1 (GPUScene.h:32-38) */
typedef struct { v3 bmin; v3 bmax; uint32_t first_index, prim_count; } PNode;          /* 32 B */
typedef struct { v3 position; float radius; int32_t material; int32_t pad[3]; } PSphere; /* 32 B */
typedef struct { v4 albedo, emissive, specular; float roughness, specular_percent, IOR, pad; } PMaterial; /* 64 B */
typedef struct { v3 origin; float vp[2]; float aspect; v3 horizontal, vertical, lower_left; } PCamera;    /* 60 B */
typedef struct { uint32_t d, v[5]; int32_t bm_flag, bm_flag_d; float bm_extra; double bm_extra_d; } PRng; /* 48 B */

typedef struct
{
    const PVertex* vertices; const PFace* faces; const PNode* nodes; const uint32_t* face_indices;
    const PSphere* spheres; int sphere_count; const PMaterial* materials;
    const float* cubemap; int cubemap_size;   /* level 0, [6][n][n][4] */
    PCamera camera;
} PScene;

typedef struct
{
    float distance; v3 position, normal; uint32_t material;
    uint32_t id;                      /* face id, 0x80000000|sphere, 0xFFFFFFFF */
} PHit;

static inline v3 V(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static inline v3 add(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 mul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 scale(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
static inline v3 neg(v3 a) { return V(-a.x, -a.y, -a.z); }
/* glm/detail/func_geometric.inl:48-54 (dot = (x + y) + z of the products), :68-78 (cross), :82-89 + inversesqrt */
static inline float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 cross(v3 x, v3 y) { return V(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
static inline v3 normalize(v3 a) { return scale(a, 1.0f / sqrtf(dot(a, a))); }
static inline float mixf(float x, float y, float a) { return x * (1.0f - a) + y * a; }   /* glm mix */
static inline v3 mix3(v3 x, v3 y, float a) { return V(mixf(x.x, y.x, a), mixf(x.y, y.y, a), mixf(x.z, y.z, a)); }
static inline float fmaxg(float a, float b) { return (a < b) ? b : a; }  /* glm::max / std::max */
static inline float fming(float a, float b) { return (b < a) ? b : a; }  /* glm::min / std::min */

/* Math.h:50-61, called with the UNNORMALISED direction and a normalised-units length (main_raytracing.cu:46) */
static int port_intersect_aabb(v3 o, v3 d, v3 bmin, v3 bmax, float ray_length)
{
    float tx1 = (bmin.x - o.x) / d.x, tx2 = (bmax.x - o.x) / d.x;
    float tmin = fming(tx1, tx2), tmax = fmaxg(tx1, tx2);
    float ty1 = (bmin.y - o.y) / d.y, ty2 = (bmax.y - o.y) / d.y;
    tmin = fmaxg(tmin, fming(ty1, ty2)); tmax = fming(tmax, fmaxg(ty1, ty2));
    float tz1 = (bmin.z - o.z) / d.z, tz2 = (bmax.z - o.z) / d.z;
    tmin = fmaxg(tmin, fming(tz1, tz2)); tmax = fming(tmax, fmaxg(tz1, tz2));
    return tmax >= tmin && tmin < ray_length && tmax > 0;
}

/* glm::intersectRayTriangle, intersect.inl:29-94 */
static int port_intersect_triangle(v3 orig, v3 dir, v3 vert0, v3 vert1, v3 vert2, float* bx, float* by, float* distance)
{
    const v3 edge1 = sub(vert1, vert0), edge2 = sub(vert2, vert0);
    const v3 p = cross(dir, edge2);
    const float det = dot(edge1, p);
    v3 perp;
    if (det > FLT_EPSILON)
    {
        const v3 dist = sub(orig, vert0);
        *bx = dot(dist, p);
        if (*bx < 0.0f || *bx > det) return 0;
        perp = cross(dist, edge1);
        *by = dot(dir, perp);
        if (*by < 0.0f || (*bx + *by) > det) return 0;
    }
    else if (det < -FLT_EPSILON)
    {
        const v3 dist = sub(orig, vert0);
        *bx = dot(dist, p);
        if (*bx > 0.0f || *bx < det) return 0;
        perp = cross(dist, edge1);
        *by = dot(dir, perp);
        if (*by > 0.0f || (*bx + *by) < det) return 0;
    }
    else return 0;
    const float inv_det = 1.0f / det;
    *distance = dot(edge2, perp) * inv_det;
    *bx *= inv_det; *by *= inv_det;
    return 1;
}

/* glm::intersectRaySphere (distance form), intersect.inl:135-153 */
static int port_intersect_sphere(v3 start, v3 dir, v3 center, float radius2, float* distance)
{
    const v3 diff = sub(center, start);
    const float t0 = dot(diff, dir);
    const float d2 = dot(diff, diff) - t0 * t0;
    if (d2 > radius2) return 0;
    const float t1 = sqrtf(radius2 - d2);
    *distance = t0 > t1 + FLT_EPSILON ? t0 - t1 : t0 + t1;
    return *distance > FLT_EPSILON;
}

/*
 * main_raytracing.cu:33-81. tie_mode 0 = the reference's rule (strictly closer wins => first visited keeps exact ties);
 * tie_mode 1 = BASELINE.json's order-independent rule: minimum t over every triangle whose leaf passes the box test
 * against the distance the search STARTED with, exact ties -> lowest face id (what the B200 kernels implement).
 * work[0] += nodes popped, work[1] += triangle tests (may be NULL).
 */
static int port_bvh_hit(const PScene* s, v3 o, v3 d, PHit* result, int tie_mode, uint64_t* work)
{
    uint32_t stack[64]; uint32_t sp = 0;
    stack[sp++] = 0;
    const v3 nd = normalize(d);
    const float original = result->distance;
    /* |d| < 1: the reference's cull is not conservative (Appendix A.4) and its result depends on its own visiting order;
     * the lowest-id rule then only re-resolves exact ties among the candidates the reference order visits. */
    const int order_independent = tie_mode && dot(d, d) >= 0.99999f;
    while (sp)
    {
        const PNode* node = &s->nodes[stack[--sp]];
        if (work) work[0]++;
        if (!port_intersect_aabb(o, d, node->bmin, node->bmax, order_independent ? original : result->distance)) continue;
        if (node->prim_count > 0)
        {
            if (work) work[1] += node->prim_count;
            for (uint32_t i = 0; i < node->prim_count; i++)
            {
                const uint32_t id = s->face_indices[node->first_index + i];
                const PFace* face = &s->faces[id];
                const PVertex* v0 = &s->vertices[face->v0]; const PVertex* v1 = &s->vertices[face->v1]; const PVertex* v2 = &s->vertices[face->v2];
                float distance, bx, by;
                if (!port_intersect_triangle(o, nd, v0->position, v1->position, v2->position, &bx, &by, &distance)) continue;
                if (tie_mode == 0) { if (distance >= result->distance || distance < 0.0f) continue; }
                else
                {
                    if (distance > result->distance || distance < 0.0f) continue;
                    if (distance == result->distance && ((result->id & 0x80000000u) || id > result->id)) continue;  /* sphere keeps a tie (:62) */
                }
                const float bz = 1.0f - bx - by;
                result->distance = distance;
                result->position = add(o, scale(nd, distance));
                result->normal = normalize(add(add(scale(v0->normal, bx), scale(v1->normal, by)), scale(v2->normal, bz)));  /* :66, weights as the reference assigns them */
                result->material = face->material;
                result->id = id;
                if (dot(nd, result->normal) >= 0.0f) result->normal = neg(result->normal);
            }
        }
        else { stack[sp++] = node->first_index; stack[sp++] = node->first_index + 1; }
    }
    return result->distance < original;
}

/* main_raytracing.cu:83-109 */
static int port_get_ray_hit(const PScene* s, v3 o, v3 d, PHit* result, int with_spheres, int tie_mode, uint64_t* work)
{
    const float max_distance = 1e30f;
    const v3 nd = normalize(d);
    result->distance = max_distance; result->id = 0xFFFFFFFFu;
    for (int i = 0; with_spheres && i < s->sphere_count; i++)
    {
        const PSphere* sp = &s->spheres[i];
        float distance;
        if (port_intersect_sphere(o, nd, sp->position, sp->radius * sp->radius, &distance))
        {
            if (distance >= result->distance) continue;
            result->distance = distance;
            result->position = add(o, scale(nd, distance));
            result->normal = V((result->position.x - sp->position.x) / sp->radius, (result->position.y - sp->position.y) / sp->radius,
                               (result->position.z - sp->position.z) / sp->radius);
            result->material = (uint32_t)sp->material;
            result->id = 0x80000000u | (uint32_t)i;
        }
    }
    if (s->nodes) port_bvh_hit(s, o, d, result, tie_mode, work);
    return result->distance < max_distance;
}

/* ---------------------------------------------------------------------------------------------
 * RNG: cuRAND XORWOW restated (curand_kernel.h). Skip-ahead matrices are derived from the
 * generator's definition (T = one step on v[5]; one subsequence = T^(2^67)), not copied tables.
 * ------------------------------------------------------------------------------------------- */
typedef struct { uint32_t row[160][5]; } XMat;
static XMat g_jump[41];  /* g_jump[k] = T^(2^(67+k)) */
static int g_jump_ready = 0;

static void xorwow_step(uint32_t v[5])  /* curand_kernel.h:866-871 without the Weyl term */
{
    const uint32_t t = v[0] ^ (v[0] >> 2);
    v[0] = v[1]; v[1] = v[2]; v[2] = v[3]; v[3] = v[4];
    v[4] = (v[4] ^ (v[4] << 4)) ^ (t ^ (t << 1));
}
static void xmat_vec(const uint32_t* vec, const XMat* m, uint32_t* out)  /* __curand_matvec, curand_kernel.h:336-352 */
{
    uint32_t r[5] = { 0, 0, 0, 0, 0 };
    for (int i = 0; i < 5; i++)
        for (int j = 0; j < 32; j++)
            if (vec[i] & (1u << j))
                for (int k = 0; k < 5; k++) r[k] ^= m->row[i * 32 + j][k];
    memcpy(out, r, sizeof(r));
}
static void xmat_square(XMat* m)
{
    static XMat tmp;
    for (int r = 0; r < 160; r++) xmat_vec(m->row[r], m, tmp.row[r]);
    *m = tmp;
}
static void xorwow_prepare(void)
{
    if (g_jump_ready) return;
    static XMat m;
    for (int r = 0; r < 160; r++)
    {
        uint32_t v[5] = { 0, 0, 0, 0, 0 };
        v[r / 32] = 1u << (r % 32);
        xorwow_step(v);
        memcpy(m.row[r], v, sizeof(v));
    }
    for (int i = 0; i < 67; i++) xmat_square(&m);
    for (int k = 0; k <= 40; k++) { g_jump[k] = m; xmat_square(&m); }
    g_jump_ready = 1;
}

/* curand_init(seed, subsequence, 0, state), curand_kernel.h:772-790 + _skipahead_sequence_inplace */
PORT_API void port_rng_init(PRng* s, uint64_t seed, uint64_t subsequence)
{
    xorwow_prepare();
    const uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
    const uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
    const uint32_t t0 = 1099087573u * s0;
    const uint32_t t1 = 2591861531u * s1;
    memset(s, 0, sizeof(*s));
    s->d = 6615241u + t1 + t0;
    s->v[0] = 123456789u + t0; s->v[1] = 362436069u ^ t0; s->v[2] = 521288629u + t1; s->v[3] = 88675123u ^ t1; s->v[4] = 5783321u + t0;
    for (int k = 0; k <= 40 && (subsequence >> k); k++)
        if ((subsequence >> k) & 1u) xmat_vec(s->v, &g_jump[k], s->v);
    /* d advances by 362437 * 2^67 * n = 0 mod 2^32 */
}
static inline uint32_t port_curand(PRng* s)  /* curand_kernel.h:863-874 */
{
    const uint32_t t = s->v[0] ^ (s->v[0] >> 2);
    s->v[0] = s->v[1]; s->v[1] = s->v[2]; s->v[2] = s->v[3]; s->v[3] = s->v[4];
    s->v[4] = (s->v[4] ^ (s->v[4] << 4)) ^ (t ^ (t << 1));
    s->d += 362437u;
    return s->v[4] + s->d;
}
/* curand_uniform.h:69-72: x * 2^-32 + 2^-33; the product is exact so FMA or not rounds once, identically */
static inline float port_uniform(PRng* s) { return (float)port_curand(s) * 2.3283064e-10f + (2.3283064e-10f / 2.0f); }
static v3 port_point_on_sphere(PRng* s)  /* Random.h:23-31 (host libm sin/cos here; device sinf/cosf differ by <= 2 ulp) */
{
    const float z = port_uniform(s) * 2.0f - 1.0f;
    const float a = port_uniform(s) * 3.14159265358979323846f * 2.0f;
    const float r = sqrtf(1.0f - z * z);
    return V(r * cosf(a), r * sinf(a), z);
}

/* ---------------------------------------------------------------------------------------------
 * Cubemap fetch of the miss shader (main_raytracing.cu:152; sampler state utils/CUDATexture.cpp:156-171):
 * level 0, linear, normalised, seamless. Restated from the CUDA programming guide (face selection,
 * 1.8 fixed-point weights). PARITY UNPINNED against hardware - see the header note.
 * ------------------------------------------------------------------------------------------- */
static int cm_select(float x, float y, float z, float* sc, float* tc, float* ma)
{
    const float ax = fabsf(x), ay = fabsf(y), az = fabsf(z);
    if (ax >= ay && ax >= az) { *ma = ax; if (x >= 0) { *sc = -z; *tc = -y; return 0; } *sc = z; *tc = -y; return 1; }
    if (ay >= az) { *ma = ay; if (y >= 0) { *sc = x; *tc = z; return 2; } *sc = x; *tc = -z; return 3; }
    *ma = az; if (z >= 0) { *sc = x; *tc = -y; return 4; } *sc = -x; *tc = -y; return 5;
}
static void cm_face_dir(int face, float sc, float tc, float* x, float* y, float* z)
{
    switch (face)
    {
    case 0: *x = 1; *z = -sc; *y = -tc; break;
    case 1: *x = -1; *z = sc; *y = -tc; break;
    case 2: *y = 1; *x = sc; *z = tc; break;
    case 3: *y = -1; *x = sc; *z = -tc; break;
    case 4: *z = 1; *x = sc; *y = -tc; break;
    default: *z = -1; *x = -sc; *y = -tc; break;
    }
}
static const float* cm_tap(const PScene* s, int face, int i, int j)
{
    const int n = s->cubemap_size;
    const int oi = i < 0 || i >= n, oj = j < 0 || j >= n;
    if (oi && oj) return NULL;
    if (oi || oj)
    {
        const float sc = (2.0f * (i + 0.5f)) / n - 1.0f, tc = (2.0f * (j + 0.5f)) / n - 1.0f;
        float x, y, z, s2, t2, ma;
        cm_face_dir(face, sc, tc, &x, &y, &z);
        face = cm_select(x, y, z, &s2, &t2, &ma);
        i = (int)floorf((s2 / ma * 0.5f + 0.5f) * n); j = (int)floorf((t2 / ma * 0.5f + 0.5f) * n);
        i = i < 0 ? 0 : (i >= n ? n - 1 : i); j = j < 0 ? 0 : (j >= n ? n - 1 : j);
    }
    return s->cubemap + (((size_t)face * n + j) * n + i) * 4;
}
static v4 port_tex_cubemap(const PScene* s, float x, float y, float z)
{
    float sc, tc, ma;
    const int face = cm_select(x, y, z, &sc, &tc, &ma);
    const int n = s->cubemap_size;
    const float u = (sc / ma + 1.0f) * 0.5f, v = (tc / ma + 1.0f) * 0.5f;
    const float xb = u * n - 0.5f, yb = v * n - 0.5f;
    const float fi = floorf(xb), fj = floorf(yb);
    const float a = floorf((xb - fi) * 256.0f + 0.5f) / 256.0f, b = floorf((yb - fj) * 256.0f + 0.5f) / 256.0f;
    const int i = (int)fi, j = (int)fj;
    const float w[4] = { (1 - a) * (1 - b), a * (1 - b), (1 - a) * b, a * b };
    const int di[4] = { 0, 1, 0, 1 }, dj[4] = { 0, 0, 1, 1 };
    float acc[4] = { 0, 0, 0, 0 }, wsum = 0;
    for (int k = 0; k < 4; k++)
    {
        const float* p = cm_tap(s, face, i + di[k], j + dj[k]);
        if (!p) continue;
        wsum += w[k];
        for (int c = 0; c < 4; c++) acc[c] += w[k] * p[c];
    }
    if (wsum > 0 && wsum != 1.0f) for (int c = 0; c < 4; c++) acc[c] /= wsum;
    v4 r = { acc[0], acc[1], acc[2], acc[3] };
    return r;
}

/* ---------------------------------------------------------------------------------------------
 * Shading + frame: main_raytracing.cu:111-200
 * ------------------------------------------------------------------------------------------- */
typedef struct { float qx, qy, qz, qw; } PQuat;
static PQuat port_env_quat(void)  /* glm quat(vec3(0, PI, 0)), type_quat.inl:204-213, PI = 3.1415926536f (:7) */
{
    const float PI = 3.1415926536f;
    const float cx = cosf(0.0f * 0.5f), cy = cosf(PI * 0.5f), cz = cosf(0.0f * 0.5f);
    const float sx = sinf(0.0f * 0.5f), sy = sinf(PI * 0.5f), sz = sinf(0.0f * 0.5f);
    PQuat q;
    q.qw = cx * cy * cz + sx * sy * sz;
    q.qx = sx * cy * cz - cx * sy * sz;
    q.qy = cx * sy * cz + sx * cy * sz;
    q.qz = cx * cy * sz - sx * sy * cz;
    return q;
}

static v3 port_ray_color(const PScene* s, v3 ro, v3 rd, PRng* rng, int num_bounces, int tie_mode, uint64_t* rays, uint64_t* work)
{
    v3 result = V(0, 0, 0), throughput = V(1, 1, 1);
    v3 o = ro, d = rd;
    for (int i = 0; i < num_bounces; i++)
    {
        PHit hit;
        if (rays) (*rays)++;
        if (port_get_ray_hit(s, o, d, &hit, 1, tie_mode, work))
        {
            const PMaterial* m = &s->materials[hit.material];
            const float do_specular = (port_uniform(rng) < m->specular_percent) ? 1.0f : 0.0f;                          /* :124 */
            result = add(result, mul(throughput, V(m->emissive.x, m->emissive.y, m->emissive.z)));                       /* :126 */
            throughput = mul(throughput, mix3(V(m->albedo.x, m->albedo.y, m->albedo.z), V(m->specular.x, m->specular.y, m->specular.z), do_specular)); /* :127 */
            const v3 diffuse_dir = normalize(add(hit.normal, port_point_on_sphere(rng)));                                /* :129 */
            const v3 refl = sub(d, scale(scale(hit.normal, dot(hit.normal, d)), 2.0f));                                  /* glm::reflect: I - N*dot(N,I)*2 */
            v3 specular_dir = normalize(refl);                                                                           /* :130 */
            specular_dir = normalize(mix3(specular_dir, diffuse_dir, m->roughness * m->roughness));                      /* :132 */
            const v3 new_dir = normalize(mix3(diffuse_dir, specular_dir, do_specular));                                  /* :133 */
            o = add(hit.position, scale(hit.normal, 0.01f));                                                             /* :135 */
            d = new_dir;
            const float p = fmaxg(throughput.x, fmaxg(throughput.y, throughput.z));                                      /* :141 */
            if (port_uniform(rng) > p) break;                                                                            /* :142 */
            throughput = scale(throughput, 1.0f / p);                                                                    /* :146 */
        }
        else
        {
            if (s->cubemap)
            {   /* :151-153; glm quat * vec3: v + ((cross(q,v) * w) + cross(q, cross(q,v))) * 2 (type_quat.inl:343-350) */
                const PQuat q = port_env_quat();
                const v3 qv = V(q.qx, q.qy, q.qz);
                const v3 uv = cross(qv, d), uuv = cross(qv, uv);
                const v3 r = add(d, scale(add(scale(uv, q.qw), uuv), 2.0f));
                const v4 c = port_tex_cubemap(s, r.x, r.y, r.z);
                const v3 e = V(fming(fmaxg(c.x, 0.0f), 50.0f), fming(fmaxg(c.y, 0.0f), 50.0f), fming(fmaxg(c.z, 0.0f), 50.0f));
                result = add(result, mul(throughput, e));
            }
            break;
        }
    }
    return result;
}

/* GPUCamera::GetRay, GPUScene.h:13 (direction not normalised) */
static void port_camera_ray(const PCamera* c, float u, float v, v3* o, v3* d)
{
    *o = c->origin;
    *d = sub(add(add(c->lower_left, scale(c->horizontal, u)), scale(c->vertical, v)), c->origin);
}

/* ---------------------------------------------------------------------------------------------
 * Exported API (ctypes: oracle/port_oracle.py)
 * ------------------------------------------------------------------------------------------- */
PORT_API void port_scene_fill(PScene* s, const void* vertices, const void* faces, const void* nodes, const uint32_t* face_indices,
                              const void* spheres, int sphere_count, const void* materials, const float* cubemap, int cubemap_size,
                              const void* camera60)
{
    s->vertices = vertices; s->faces = faces; s->nodes = nodes; s->face_indices = face_indices;
    s->spheres = spheres; s->sphere_count = sphere_count; s->materials = materials;
    s->cubemap = cubemap; s->cubemap_size = cubemap_size;
    if (camera60) memcpy(&s->camera, camera60, sizeof(PCamera)); else memset(&s->camera, 0, sizeof(PCamera));
}
PORT_API uint32_t port_sizeof_scene(void) { return (uint32_t)sizeof(PScene); }

/* Pixel-centre primary rays. work2 (optional) = per-pixel {nodes popped, triangle tests}. */
PORT_API void port_primary_hits(const PScene* s, int width, int height, int with_spheres, int tie_mode,
                                uint32_t* out_id, float* out_t, uint32_t* work2)
{
#pragma omp parallel for schedule(dynamic, 8)
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++)
        {
            const float u = ((float)x + 0.5f) / (float)width, v = ((float)y + 0.5f) / (float)height;
            v3 o, d; port_camera_ray(&s->camera, u, v, &o, &d);
            PHit h; uint64_t w[2] = { 0, 0 };
            port_get_ray_hit(s, o, d, &h, with_spheres, tie_mode, w);
            const size_t i = (size_t)y * width + x;
            out_id[i] = h.id; out_t[i] = h.distance;
            if (work2) { work2[2 * i] = (uint32_t)w[0]; work2[2 * i + 1] = (uint32_t)w[1]; }
        }
}

PORT_API void port_trace_rays(const PScene* s, const float* rays, uint64_t n, int with_spheres, int tie_mode, uint32_t* out_id, float* out_t)
{
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < (int64_t)n; i++)
    {
        PHit h;
        port_get_ray_hit(s, V(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]), V(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]), &h, with_spheres, tie_mode, NULL);
        out_id[i] = h.id; out_t[i] = h.distance;
    }
}

PORT_API void port_rng_init_many(PRng* states, uint64_t count, uint64_t seed, uint64_t first_subsequence)
{
    xorwow_prepare();
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)count; i++) port_rng_init(&states[i], seed, first_subsequence + (uint64_t)i);
}
PORT_API void port_rng_draw(PRng* s, uint32_t n, uint32_t* out) { for (uint32_t i = 0; i < n; i++) out[i] = port_curand(s); }
PORT_API void port_rng_uniform(PRng* s, uint32_t n, float* out) { for (uint32_t i = 0; i < n; i++) out[i] = port_uniform(s); }
PORT_API void port_rng_sphere(PRng* s, float* out3) { const v3 p = port_point_on_sphere(s); out3[0] = p.x; out3[1] = p.y; out3[2] = p.z; }

/*
 * BASELINE.json config 3, the incoherent-ray stress: per pixel one pixel-centre primary ray through GetRayHit
 * (main_raytracing.cu:83-109); on a hit, `ao_rays` closest-hit queries along normalize(hit.normal + GetRandomPointOnSphere())
 * (the diffuse lobe of main_raytracing.cu:129, two draws per ray, Random.h:23-31) from hit.position + hit.normal * 0.01
 * (:135). out_ao = unoccluded / ao_rays (1 where the primary ray misses: no draws); out_id (optional) = primary hit id.
 * The reference has no AO kernel: this is its lobe, offset, RNG stream and closest-hit query in the sequence
 * csrc/pathtrace.cu ambient_occlusion_kernel issues them.
 */
PORT_API void port_ambient_occlusion(const PScene* s, PRng* rng_states, int width, int height, int ao_rays, int tie_mode,
                                     float* out_ao, uint32_t* out_id, uint64_t* ray_count)
{
    uint64_t c_rays = 0;
#pragma omp parallel for schedule(dynamic, 2) reduction(+ : c_rays)
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++)
        {
            const size_t i = (size_t)y * width + x;
            PRng* rng = &rng_states[i];
            const float u = ((float)x + 0.5f) / (float)width, v = ((float)y + 0.5f) / (float)height;
            v3 o, d; port_camera_ray(&s->camera, u, v, &o, &d);
            PHit h;
            c_rays++;
            const int hit = port_get_ray_hit(s, o, d, &h, 1, tie_mode, NULL);
            if (out_id) out_id[i] = h.id;
            if (!hit) { out_ao[i] = 1.0f; continue; }
            const v3 base = add(h.position, scale(h.normal, 0.01f));
            int unoccluded = 0;
            for (int k = 0; k < ao_rays; k++)
            {
                const v3 dir = normalize(add(h.normal, port_point_on_sphere(rng)));
                PHit h2;
                c_rays++;
                if (!port_get_ray_hit(s, base, dir, &h2, 1, tie_mode, NULL)) unoccluded++;
            }
            out_ao[i] = ao_rays ? (float)unoccluded / (float)ao_rays : 1.0f;
        }
    if (ray_count) *ray_count += c_rays;
}

/*
 * raytracing_kernel_main (main_raytracing.cu:162-200) for rows [row_begin, row_end) stepping row_stride.
 * blend_mode 0: *pixel = lerp(last, mean, 1/(frame_index+1)), alpha = 1.  blend_mode 1: surface += sum, alpha += spp.
 * counters (optional): [0] += rays (GetRayHit calls), [1] += nodes popped, [2] += triangle tests.
 */
PORT_API void port_render(const PScene* s, PRng* rng_states, float* surface, const float* surface_last, int width, int height,
                          uint64_t pitch_bytes, int frame_index, int spp, int num_bounces, int blend_mode, int tie_mode,
                          int row_begin, int row_end, int row_stride, uint64_t* counters)
{
    if (row_end <= 0 || row_end > height) row_end = height;
    if (row_stride <= 0) row_stride = 1;
    uint64_t c_rays = 0, c_nodes = 0, c_tris = 0;
#pragma omp parallel for schedule(dynamic, 2) reduction(+ : c_rays, c_nodes, c_tris)
    for (int y = row_begin; y < row_end; y += row_stride)
    {
        uint64_t rays = 0, work[2] = { 0, 0 };
        for (int x = 0; x < width; x++)
        {
            PRng* rng = &rng_states[(size_t)y * width + x];                                   /* GPUScene.h:95 */
            float* px = (float*)((char*)surface + (size_t)y * pitch_bytes) + 4 * x;
            v3 sum = V(0, 0, 0);
            for (int i = 0; i < spp; i++)
            {
                const float jx = port_uniform(rng), jy = port_uniform(rng);                   /* :190, x first = nvcc device order */
                const float u = ((float)x + jx) / (float)width, v = ((float)y + jy) / (float)height;
                v3 o, d; port_camera_ray(&s->camera, u, v, &o, &d);
                sum = add(sum, port_ray_color(s, o, d, rng, num_bounces, tie_mode, &rays, counters ? work : NULL));
            }
            if (blend_mode == 0)
            {
                const float n = (float)spp;
                const v3 mean = V(sum.x / n, sum.y / n, sum.z / n);                           /* :195 */
                const float a = frame_index > 0 ? 1.0f / (float)(frame_index + 1) : 1.0f;     /* :196 */
                const float* last = (const float*)((const char*)surface_last + (size_t)y * pitch_bytes) + 4 * x;
                px[0] = mixf(last[0], mean.x, a); px[1] = mixf(last[1], mean.y, a); px[2] = mixf(last[2], mean.z, a); px[3] = 1.0f;
            }
            else { px[0] += sum.x; px[1] += sum.y; px[2] += sum.z; px[3] += (float)spp; }
        }
        c_rays += rays; c_nodes += work[0]; c_tris += work[1];
    }
    if (counters) { counters[0] += c_rays; counters[1] += c_nodes; counters[2] += c_tris; }
}
