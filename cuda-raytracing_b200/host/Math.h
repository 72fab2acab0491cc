// Host-side fp32 vector/matrix math for scene assembly and the camera.
//
// The reference does this with GLM 0.9.9.8 (RayTracing/Math.h:12-23). Triangle IDs and the BVH
// topology hang on the exact fp32 vertex positions, so every routine here evaluates the SAME
// expression tree, in the same order, as the GLM routine it stands in for (cited per function);
// tests/test_scene_parity.py checks the resulting arrays bit-for-bit against a GLM build.
// Compile with -ffp-contract=off (no FMA), as the oracle is.
#pragma once
#include <cmath>
#include <cstdint>
#include "rt_types.h"

namespace Math
{
    struct vec2 { float x = 0, y = 0; vec2() = default; vec2(float x_, float y_) : x(x_), y(y_) {} explicit vec2(float s) : x(s), y(s) {} };

    struct vec3
    {
        float x = 0, y = 0, z = 0;
        vec3() = default;
        vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
        explicit vec3(float s) : x(s), y(s), z(s) {}
        vec3(const RtVec3& v) : x(v.x), y(v.y), z(v.z) {}
        operator RtVec3() const { return RtVec3{ x, y, z }; }
        float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
        float& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    };

    struct vec4
    {
        float x = 0, y = 0, z = 0, w = 0;
        vec4() = default;
        vec4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
        explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
        vec4(const vec3& v, float w_) : x(v.x), y(v.y), z(v.z), w(w_) {}
        operator RtVec4() const { return RtVec4{ x, y, z, w }; }
        vec3 xyz() const { return vec3(x, y, z); }
        float operator[](int i) const { return (&x)[i]; }
        float& operator[](int i) { return (&x)[i]; }
    };

    inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
    inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
    inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
    inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
    inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
    inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
    inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
    inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
    inline vec4 operator-(vec4 a, vec4 b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
    inline vec4 operator*(vec4 a, vec4 b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
    inline vec4 operator*(vec4 a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
    inline vec4 operator/(vec4 a, float s) { return vec4(a.x / s, a.y / s, a.z / s, a.w / s); }

    // glm/detail/func_geometric.inl:48-54 (dot: products then x+y+z), :68-78 (cross), :82-89 (normalize = v * inversesqrt)
    inline float dot(vec3 a, vec3 b) { const vec3 t = a * b; return t.x + t.y + t.z; }
    inline vec3 cross(vec3 x, vec3 y) { return vec3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
    inline float inversesqrt(float x) { return 1.0f / std::sqrt(x); }  // glm/detail/func_exponential.inl:136-139
    inline vec3 normalize(vec3 v) { return v * inversesqrt(dot(v, v)); }
    inline float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }  // glm/detail/func_trigonometric.inl

    struct quat { float x = 0, y = 0, z = 0, w = 1; };

    // glm/detail/type_quat.inl:204-213 - quaternion from Euler angles (pitch, yaw, roll)
    inline quat quat_from_euler(vec3 e)
    {
        const vec3 c(std::cos(e.x * 0.5f), std::cos(e.y * 0.5f), std::cos(e.z * 0.5f));
        const vec3 s(std::sin(e.x * 0.5f), std::sin(e.y * 0.5f), std::sin(e.z * 0.5f));
        quat q;
        q.w = c.x * c.y * c.z + s.x * s.y * s.z;
        q.x = s.x * c.y * c.z - c.x * s.y * s.z;
        q.y = c.x * s.y * c.z + s.x * c.y * s.z;
        q.z = c.x * c.y * s.z - s.x * s.y * c.z;
        return q;
    }

    struct mat4
    {
        vec4 c[4];  // column-major, as GLM
        mat4() : mat4(1.0f) {}
        explicit mat4(float d) { c[0] = vec4(d, 0, 0, 0); c[1] = vec4(0, d, 0, 0); c[2] = vec4(0, 0, d, 0); c[3] = vec4(0, 0, 0, d); }
        const vec4& operator[](int i) const { return c[i]; }
        vec4& operator[](int i) { return c[i]; }
    };

    // glm/detail/type_mat4x4.inl:561-572 : (m0*x + m1*y) + (m2*z + m3*w)
    inline vec4 operator*(const mat4& m, vec4 v)
    {
        const vec4 Mul0 = m[0] * v.x, Mul1 = m[1] * v.y; const vec4 Add0 = Mul0 + Mul1;
        const vec4 Mul2 = m[2] * v.z, Mul3 = m[3] * v.w; const vec4 Add1 = Mul2 + Mul3;
        return Add0 + Add1;
    }

    // glm/detail/type_mat4x4.inl:630-648 : left-to-right sum of the four scaled columns
    inline mat4 operator*(const mat4& a, const mat4& b)
    {
        mat4 r;
        for (int i = 0; i < 4; i++) r[i] = a[0] * b[i][0] + a[1] * b[i][1] + a[2] * b[i][2] + a[3] * b[i][3];
        return r;
    }
    inline mat4 operator*(const mat4& a, float s) { mat4 r; for (int i = 0; i < 4; i++) r[i] = a[i] * s; return r; }

    // glm/ext/matrix_transform.inl:10-15
    inline mat4 translate(const mat4& m, vec3 v) { mat4 r(m); r[3] = m[0] * v.x + m[1] * v.y + m[2] * v.z + m[3]; return r; }
    // glm/ext/matrix_transform.inl:78-86
    inline mat4 scale(const mat4& m, vec3 v) { mat4 r; r[0] = m[0] * v.x; r[1] = m[1] * v.y; r[2] = m[2] * v.z; r[3] = m[3]; return r; }
    // glm/ext/matrix_transform.inl:18-46
    inline mat4 rotate(const mat4& m, float angle, vec3 v)
    {
        const float a = angle, c = std::cos(a), s = std::sin(a);
        const vec3 axis(normalize(v));
        const vec3 temp((1.0f - c) * axis);
        float R[3][3];
        R[0][0] = c + temp[0] * axis[0]; R[0][1] = temp[0] * axis[1] + s * axis[2]; R[0][2] = temp[0] * axis[2] - s * axis[1];
        R[1][0] = temp[1] * axis[0] - s * axis[2]; R[1][1] = c + temp[1] * axis[1]; R[1][2] = temp[1] * axis[2] + s * axis[0];
        R[2][0] = temp[2] * axis[0] + s * axis[1]; R[2][1] = temp[2] * axis[1] - s * axis[0]; R[2][2] = c + temp[2] * axis[2];
        mat4 r;
        r[0] = m[0] * R[0][0] + m[1] * R[0][1] + m[2] * R[0][2];
        r[1] = m[0] * R[1][0] + m[1] * R[1][1] + m[2] * R[1][2];
        r[2] = m[0] * R[2][0] + m[1] * R[2][1] + m[2] * R[2][2];
        r[3] = m[3];
        return r;
    }

    // glm/gtc/quaternion.inl:41-72 (mat3_cast widened to mat4)
    inline mat4 mat4_cast(quat q)
    {
        mat4 r(1.0f);
        const float qxx(q.x * q.x), qyy(q.y * q.y), qzz(q.z * q.z), qxz(q.x * q.z), qxy(q.x * q.y), qyz(q.y * q.z);
        const float qwx(q.w * q.x), qwy(q.w * q.y), qwz(q.w * q.z);
        r[0][0] = 1.0f - 2.0f * (qyy + qzz); r[0][1] = 2.0f * (qxy + qwz); r[0][2] = 2.0f * (qxz - qwy);
        r[1][0] = 2.0f * (qxy - qwz); r[1][1] = 1.0f - 2.0f * (qxx + qzz); r[1][2] = 2.0f * (qyz + qwx);
        r[2][0] = 2.0f * (qxz + qwy); r[2][1] = 2.0f * (qyz - qwx); r[2][2] = 1.0f - 2.0f * (qxx + qyy);
        return r;
    }

    // glm/ext/matrix_clip_space.inl:249-262
    inline mat4 perspectiveRH_NO(float fovy, float aspect, float zNear, float zFar)
    {
        const float tanHalfFovy = std::tan(fovy / 2.0f);
        mat4 r(0.0f);
        r[0][0] = 1.0f / (aspect * tanHalfFovy);
        r[1][1] = 1.0f / (tanHalfFovy);
        r[2][2] = -(zFar + zNear) / (zFar - zNear);
        r[2][3] = -1.0f;
        r[3][2] = -(2.0f * zFar * zNear) / (zFar - zNear);
        return r;
    }

    mat4 inverse(const mat4& m);  // glm/detail/func_matrix.inl:294-352, in Math.cpp

    // RayTracing/Math.h:63-70
    inline mat4 ComposeMatrix(vec3 position, quat rotation, vec3 scl)
    {
        mat4 result(1.0f);
        result = translate(result, position);
        result = result * mat4_cast(rotation);
        result = scale(result, scl);
        return result;
    }
}
