// Legacy-DDS reader for the one layout the reference's asset uses (data/sunset_uncompressed.dds):
// FourCC 116 (D3DFMT_A32B32G32R32F -> DXGI_FORMAT_R32G32B32A32_FLOAT, utils/image/DDSLoaderHelpers.h:745)
// or a DX10 header with format 2, cubemap with all six faces, face-major mip chains
// (utils/image/DDSLoader.cpp:37-133). Other formats are out of scope (never exercised).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

struct DdsCubemap
{
    uint32_t size = 0;       // edge length of level 0
    uint32_t mip_count = 0;
    // mip-major re-pack (utils/CUDATexture.cpp:195-217): levels[m] = 6 faces x (size>>m)^2 x RGBA32F
    std::vector<std::vector<float>> levels;
};

// returns 0 or a negative RT_E_* code; `error` receives a one-line reason
int ParseDdsCubemap(const void* bytes, size_t size, DdsCubemap& out, std::string& error);
int LoadFileBytes(const char* path, std::vector<uint8_t>& out);
