#include "Math.h"

namespace Math
{
    // Cofactor inverse in the exact evaluation order of glm/detail/func_matrix.inl:294-352
    // (used only on the projection matrix, RayTracing/Scene.cpp:22).
    mat4 inverse(const mat4& m)
    {
        const float Coef00 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
        const float Coef02 = m[1][2] * m[3][3] - m[3][2] * m[1][3];
        const float Coef03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
        const float Coef04 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
        const float Coef06 = m[1][1] * m[3][3] - m[3][1] * m[1][3];
        const float Coef07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
        const float Coef08 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
        const float Coef10 = m[1][1] * m[3][2] - m[3][1] * m[1][2];
        const float Coef11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
        const float Coef12 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
        const float Coef14 = m[1][0] * m[3][3] - m[3][0] * m[1][3];
        const float Coef15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
        const float Coef16 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
        const float Coef18 = m[1][0] * m[3][2] - m[3][0] * m[1][2];
        const float Coef19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
        const float Coef20 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
        const float Coef22 = m[1][0] * m[3][1] - m[3][0] * m[1][1];
        const float Coef23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];

        const vec4 Fac0(Coef00, Coef00, Coef02, Coef03);
        const vec4 Fac1(Coef04, Coef04, Coef06, Coef07);
        const vec4 Fac2(Coef08, Coef08, Coef10, Coef11);
        const vec4 Fac3(Coef12, Coef12, Coef14, Coef15);
        const vec4 Fac4(Coef16, Coef16, Coef18, Coef19);
        const vec4 Fac5(Coef20, Coef20, Coef22, Coef23);

        const vec4 Vec0(m[1][0], m[0][0], m[0][0], m[0][0]);
        const vec4 Vec1(m[1][1], m[0][1], m[0][1], m[0][1]);
        const vec4 Vec2(m[1][2], m[0][2], m[0][2], m[0][2]);
        const vec4 Vec3(m[1][3], m[0][3], m[0][3], m[0][3]);

        const vec4 Inv0(Vec1 * Fac0 - Vec2 * Fac1 + Vec3 * Fac2);
        const vec4 Inv1(Vec0 * Fac0 - Vec2 * Fac3 + Vec3 * Fac4);
        const vec4 Inv2(Vec0 * Fac1 - Vec1 * Fac3 + Vec3 * Fac5);
        const vec4 Inv3(Vec0 * Fac2 - Vec1 * Fac4 + Vec2 * Fac5);

        const vec4 SignA(+1, -1, +1, -1);
        const vec4 SignB(-1, +1, -1, +1);
        mat4 Inverse;
        Inverse[0] = Inv0 * SignA; Inverse[1] = Inv1 * SignB; Inverse[2] = Inv2 * SignA; Inverse[3] = Inv3 * SignB;

        const vec4 Row0(Inverse[0][0], Inverse[1][0], Inverse[2][0], Inverse[3][0]);
        const vec4 Dot0(m[0] * Row0);
        const float Dot1 = (Dot0.x + Dot0.y) + (Dot0.z + Dot0.w);
        const float OneOverDeterminant = 1.0f / Dot1;
        return Inverse * OneOverDeterminant;
    }
}
