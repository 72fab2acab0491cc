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

// ---- general DDS container (utils/image/DDSLoader.cpp:135-371, DDSLoaderHelpers.h:24-189,350-510,512-755) ----
// Everything GetDDSSubresources accepts: legacy and DX10 headers, 1D/2D/3D, arrays, cubemaps, mip chains, block-compressed
// and packed formats. Table-driven; the subresource walk (array element major, mips inside) is FillInitData's.
struct DxgiFormatInfo
{
    uint32_t dxgi = 0;
    uint32_t bits_per_pixel = 0;   // DDSLoaderHelpers.h BitsPerPixel
    uint32_t block_bytes = 0;      // 8 / 16 for BC1..BC7 (4x4 texel blocks), 4 for the 2x1-packed R8G8_B8G8 family, else 0
    int cx = 0, cy = 0, cz = 0, cw = 0;  // CUDA channel bits (utils/CUDATexture.cpp:9-75)
    int cuda_kind = -1;            // cudaChannelFormatKind, -1 = no CUDA array format (parse only)
    bool normalized = false;       // cudaReadModeNormalizedFloat
    const char* name = "";
};
bool LookupDxgiFormat(uint32_t dxgi, DxgiFormatInfo& out);

struct DdsSubresource { size_t offset = 0, row_pitch = 0, slice_pitch = 0; uint32_t width = 0, height = 0, depth = 0; };
struct DdsImage
{
    uint32_t width = 0, height = 0, depth = 0, mip_count = 0, array_size = 0;  // array_size already x6 for cubemaps
    bool is_cubemap = false;
    int num_dimensions = 0;
    DxgiFormatInfo format;
    std::vector<DdsSubresource> subresources;  // index = element * mip_count + mip, offsets into the file bytes
};
int ParseDds(const void* bytes, size_t size, DdsImage& out, std::string& error);
