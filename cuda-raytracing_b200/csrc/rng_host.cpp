// Host side of the XORWOW skip-ahead: the GF(2) jump matrices for `curand_init(seed, subsequence, 0)`.
// cuRAND ships them as tables (curand_precalc.h, powers of 4 of T^(2^67)); here they are derived from
// the generator's definition: T = one XORWOW step on the 160-bit v[] (curand_kernel.h:866-871),
// J = T^(2^67) (one subsequence, curand_kernel.h:742-746), jump[k] = J^(2^k).
// Matrix convention follows __curand_matvec (curand_kernel.h:336-352): row r is the image of bit r.
#include "rng_host.h"
#include <cstring>

namespace
{
    constexpr int kWords = 5, kBits = 160;
    struct Mat { uint32_t row[kBits][kWords]; };

    void Step(uint32_t v[kWords])
    {
        const uint32_t t = v[0] ^ (v[0] >> 2);
        v[0] = v[1]; v[1] = v[2]; v[2] = v[3]; v[3] = v[4];
        v[4] = (v[4] ^ (v[4] << 4)) ^ (t ^ (t << 1));
    }
    void MatVec(const uint32_t* vec, const Mat& m, uint32_t* out)
    {
        uint32_t r[kWords] = { 0, 0, 0, 0, 0 };
        for (int i = 0; i < kWords; i++)
            for (int j = 0; j < 32; j++)
                if (vec[i] & (1u << j))
                    for (int k = 0; k < kWords; k++) r[k] ^= m.row[i * 32 + j][k];
        memcpy(out, r, sizeof(r));
    }
    void Square(Mat& m)
    {
        static Mat tmp;
        for (int r = 0; r < kBits; r++) MatVec(m.row[r], m, tmp.row[r]);
        m = tmp;
    }
}

void XorwowSequenceJumpMatrices(uint32_t* out /* [levels][160][5] */, int levels)
{
    static Mat m;
    for (int r = 0; r < kBits; r++)
    {
        uint32_t v[kWords] = { 0, 0, 0, 0, 0 };
        v[r / 32] = 1u << (r % 32);
        Step(v);
        memcpy(m.row[r], v, sizeof(v));
    }
    for (int i = 0; i < 67; i++) Square(m);  // T^(2^67)
    for (int k = 0; k < levels; k++)
    {
        memcpy(out + size_t(k) * kBits * kWords, m.row, sizeof(m.row));
        Square(m);
    }
}

void XorwowSeedState(uint64_t seed, uint32_t* d, uint32_t v[5])  // curand_kernel.h:772-790
{
    const uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
    const uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
    const uint32_t t0 = 1099087573u * s0;
    const uint32_t t1 = 2591861531u * s1;
    *d = 6615241u + t1 + t0;
    v[0] = 123456789u + t0;
    v[1] = 362436069u ^ t0;
    v[2] = 521288629u + t1;
    v[3] = 88675123u ^ t1;
    v[4] = 5783321u + t0;
}
