#pragma once
#include <cstddef>
#include <cstdint>
// jump[k] = (T^(2^67))^(2^k), k < levels; each matrix is 160 rows x 5 words (row r = image of state bit r)
void XorwowSequenceJumpMatrices(uint32_t* out, int levels);
void XorwowSeedState(uint64_t seed, uint32_t* d, uint32_t v[5]);
