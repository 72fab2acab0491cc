#include "dds.h"
#include <cstdio>
#include <cstring>
#include "b200rt.h"

namespace
{
    struct DdsPixelFormat { uint32_t size, flags, fourCC, RGBBitCount, RBitMask, GBitMask, BBitMask, ABitMask; };
    struct DdsHeader
    {
        uint32_t size, flags, height, width, pitchOrLinearSize, depth, mipMapCount, reserved1[11];
        DdsPixelFormat ddspf;
        uint32_t caps, caps2, caps3, caps4, reserved2;
    };
    static_assert(sizeof(DdsHeader) == 124, "DDS_HEADER is 124 bytes");
    constexpr uint32_t kMagic = 0x20534444u;         // "DDS "
    constexpr uint32_t kFourCC = 0x4u;               // DDPF_FOURCC
    constexpr uint32_t kCubemapAllFaces = 0xFE00u;   // DDSCAPS2_CUBEMAP | all six face bits
    constexpr uint32_t kFourCC_DX10 = 0x30315844u;   // "DX10"
    constexpr uint32_t kD3DFMT_A32B32G32R32F = 116u;
    constexpr uint32_t kDXGI_R32G32B32A32_FLOAT = 2u;
}

int LoadFileBytes(const char* path, std::vector<uint8_t>& out)
{
    FILE* f = fopen(path, "rb");
    if (!f) return RT_E_IO;
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (n < 0) { fclose(f); return RT_E_IO; }
    out.resize((size_t)n);
    const size_t got = n ? fread(out.data(), 1, (size_t)n, f) : 0;
    fclose(f);
    return got == (size_t)n ? RT_OK : RT_E_IO;
}

int ParseDdsCubemap(const void* bytes, size_t size, DdsCubemap& out, std::string& error)
{
    const uint8_t* p = static_cast<const uint8_t*>(bytes);
    if (size < 4 + sizeof(DdsHeader)) { error = "DDS: file too small"; return RT_E_IO; }
    uint32_t magic; memcpy(&magic, p, 4);
    DdsHeader h; memcpy(&h, p + 4, sizeof(h));
    if (magic != kMagic || h.size != 124 || h.ddspf.size != 32) { error = "DDS: bad magic/header"; return RT_E_IO; }
    size_t offset = 4 + sizeof(DdsHeader);
    if (!(h.ddspf.flags & kFourCC)) { error = "DDS: only FourCC float formats are supported"; return RT_E_IO; }
    if (h.ddspf.fourCC == kFourCC_DX10)
    {
        if (size < offset + 20) { error = "DDS: truncated DX10 header"; return RT_E_IO; }
        uint32_t dx10[5]; memcpy(dx10, p + offset, 20); offset += 20;
        if (dx10[0] != kDXGI_R32G32B32A32_FLOAT || !(dx10[2] & 0x4u)) { error = "DDS: DX10 header is not an RGBA32F cubemap"; return RT_E_IO; }
    }
    else if (h.ddspf.fourCC != kD3DFMT_A32B32G32R32F) { error = "DDS: unsupported FourCC (only 116 = RGBA32F)"; return RT_E_IO; }
    else if ((h.caps2 & kCubemapAllFaces) != kCubemapAllFaces) { error = "DDS: not a six-face cubemap"; return RT_E_IO; }
    if (h.width != h.height || h.width == 0) { error = "DDS: cubemap faces must be square"; return RT_E_IO; }

    out.size = h.width;
    out.mip_count = h.mipMapCount ? h.mipMapCount : 1;
    // nothing is allocated on the header's word alone: a full chain has floor(log2(size)) + 1 levels (and at most 16 here, sizes up
    // to 16384 like ParseDds), and the pixel data the header promises must actually be in the file
    if (h.width > 16384) { error = "DDS: cubemap face larger than 16384"; return RT_E_IO; }
    uint32_t full_chain = 1;
    while ((h.width >> full_chain) != 0 && full_chain < 16) full_chain++;
    if (out.mip_count > full_chain) { error = "DDS: more mip levels than the face size allows"; return RT_E_IO; }
    {
        size_t need = 0;
        for (uint32_t m = 0; m < out.mip_count; m++) { const size_t w = (out.size >> m) ? (out.size >> m) : 1; need += 6 * w * w * 16; }
        if (need > size - offset) { error = "DDS: truncated pixel data"; return RT_E_IO; }
    }
    out.levels.assign(out.mip_count, {});
    for (uint32_t m = 0; m < out.mip_count; m++)
    {
        const uint32_t w = out.size >> m ? out.size >> m : 1;
        out.levels[m].resize((size_t)6 * w * w * 4);
    }
    // file order: face-major (face0 mip0..mipN, face1 ...), DDSLoader.cpp:60-128; re-packed mip-major
    for (uint32_t face = 0; face < 6; face++)
        for (uint32_t m = 0; m < out.mip_count; m++)
        {
            const uint32_t w = out.size >> m ? out.size >> m : 1;
            const size_t bytes_here = (size_t)w * w * 16;
            if (offset + bytes_here > size) { error = "DDS: truncated pixel data"; return RT_E_IO; }
            memcpy(reinterpret_cast<uint8_t*>(out.levels[m].data()) + face * bytes_here, p + offset, bytes_here);
            offset += bytes_here;
        }
    return RT_OK;
}

// ------------------------------------------------------------------------------------------------
// General container
// ------------------------------------------------------------------------------------------------
namespace
{
    // cudaChannelFormatKind values (driver_types.h) without pulling the CUDA headers into this host TU
    enum { K_S = 0, K_U = 1, K_F = 2, K_BC1 = 17, K_BC2 = 19, K_BC3 = 21, K_BC4U = 23, K_BC4S = 24, K_BC5U = 25, K_BC5S = 26,
           K_BC6U = 27, K_BC6S = 28, K_BC7 = 29, K_NONE = -1 };

    struct Row { uint32_t dxgi, bpp, block; int x, y, z, w, kind; bool norm; const char* name; };
    // dxgi value, bits/pixel, block bytes, CUDA channel bits, kind, normalized read
    const Row kFormats[] = {
        { 1, 128, 0, 0, 0, 0, 0, K_NONE, false, "R32G32B32A32_TYPELESS" }, { 2, 128, 0, 32, 32, 32, 32, K_F, false, "R32G32B32A32_FLOAT" },
        { 3, 128, 0, 32, 32, 32, 32, K_U, false, "R32G32B32A32_UINT" }, { 4, 128, 0, 32, 32, 32, 32, K_S, false, "R32G32B32A32_SINT" },
        { 5, 96, 0, 0, 0, 0, 0, K_NONE, false, "R32G32B32_TYPELESS" }, { 6, 96, 0, 0, 0, 0, 0, K_NONE, false, "R32G32B32_FLOAT" },
        { 7, 96, 0, 0, 0, 0, 0, K_NONE, false, "R32G32B32_UINT" }, { 8, 96, 0, 0, 0, 0, 0, K_NONE, false, "R32G32B32_SINT" },
        { 9, 64, 0, 0, 0, 0, 0, K_NONE, false, "R16G16B16A16_TYPELESS" }, { 10, 64, 0, 16, 16, 16, 16, K_F, false, "R16G16B16A16_FLOAT" },
        { 11, 64, 0, 16, 16, 16, 16, K_U, true, "R16G16B16A16_UNORM" }, { 12, 64, 0, 16, 16, 16, 16, K_U, false, "R16G16B16A16_UINT" },
        { 13, 64, 0, 16, 16, 16, 16, K_S, true, "R16G16B16A16_SNORM" }, { 14, 64, 0, 16, 16, 16, 16, K_S, false, "R16G16B16A16_SINT" },
        { 15, 64, 0, 0, 0, 0, 0, K_NONE, false, "R32G32_TYPELESS" }, { 16, 64, 0, 32, 32, 0, 0, K_F, false, "R32G32_FLOAT" },
        { 17, 64, 0, 32, 32, 0, 0, K_U, false, "R32G32_UINT" }, { 18, 64, 0, 32, 32, 0, 0, K_S, false, "R32G32_SINT" },
        { 19, 64, 0, 0, 0, 0, 0, K_NONE, false, "R32G8X24_TYPELESS" }, { 20, 64, 0, 0, 0, 0, 0, K_NONE, false, "D32_FLOAT_S8X24_UINT" },
        { 21, 64, 0, 0, 0, 0, 0, K_NONE, false, "R32_FLOAT_X8X24_TYPELESS" }, { 22, 64, 0, 0, 0, 0, 0, K_NONE, false, "X32_TYPELESS_G8X24_UINT" },
        { 23, 32, 0, 0, 0, 0, 0, K_NONE, false, "R10G10B10A2_TYPELESS" }, { 24, 32, 0, 0, 0, 0, 0, K_NONE, false, "R10G10B10A2_UNORM" },
        { 25, 32, 0, 0, 0, 0, 0, K_NONE, false, "R10G10B10A2_UINT" }, { 26, 32, 0, 0, 0, 0, 0, K_NONE, false, "R11G11B10_FLOAT" },
        { 27, 32, 0, 0, 0, 0, 0, K_NONE, false, "R8G8B8A8_TYPELESS" }, { 28, 32, 0, 8, 8, 8, 8, K_U, true, "R8G8B8A8_UNORM" },
        { 29, 32, 0, 8, 8, 8, 8, K_U, true, "R8G8B8A8_UNORM_SRGB" }, { 30, 32, 0, 8, 8, 8, 8, K_U, false, "R8G8B8A8_UINT" },
        { 31, 32, 0, 8, 8, 8, 8, K_S, true, "R8G8B8A8_SNORM" }, { 32, 32, 0, 8, 8, 8, 8, K_S, false, "R8G8B8A8_SINT" },
        { 33, 32, 0, 0, 0, 0, 0, K_NONE, false, "R16G16_TYPELESS" }, { 34, 32, 0, 16, 16, 0, 0, K_F, false, "R16G16_FLOAT" },
        { 35, 32, 0, 16, 16, 0, 0, K_U, true, "R16G16_UNORM" }, { 36, 32, 0, 16, 16, 0, 0, K_U, false, "R16G16_UINT" },
        { 37, 32, 0, 16, 16, 0, 0, K_S, true, "R16G16_SNORM" }, { 38, 32, 0, 16, 16, 0, 0, K_S, false, "R16G16_SINT" },
        { 39, 32, 0, 0, 0, 0, 0, K_NONE, false, "R32_TYPELESS" }, { 40, 32, 0, 0, 0, 0, 0, K_NONE, false, "D32_FLOAT" },
        { 41, 32, 0, 32, 0, 0, 0, K_F, false, "R32_FLOAT" }, { 42, 32, 0, 32, 0, 0, 0, K_U, false, "R32_UINT" }, { 43, 32, 0, 32, 0, 0, 0, K_S, false, "R32_SINT" },
        { 44, 32, 0, 0, 0, 0, 0, K_NONE, false, "R24G8_TYPELESS" }, { 45, 32, 0, 0, 0, 0, 0, K_NONE, false, "D24_UNORM_S8_UINT" },
        { 46, 32, 0, 0, 0, 0, 0, K_NONE, false, "R24_UNORM_X8_TYPELESS" }, { 47, 32, 0, 0, 0, 0, 0, K_NONE, false, "X24_TYPELESS_G8_UINT" },
        { 48, 16, 0, 0, 0, 0, 0, K_NONE, false, "R8G8_TYPELESS" }, { 49, 16, 0, 8, 8, 0, 0, K_U, true, "R8G8_UNORM" }, { 50, 16, 0, 8, 8, 0, 0, K_U, false, "R8G8_UINT" },
        { 51, 16, 0, 8, 8, 0, 0, K_S, true, "R8G8_SNORM" }, { 52, 16, 0, 8, 8, 0, 0, K_S, false, "R8G8_SINT" },
        { 53, 16, 0, 0, 0, 0, 0, K_NONE, false, "R16_TYPELESS" }, { 54, 16, 0, 16, 0, 0, 0, K_F, false, "R16_FLOAT" }, { 55, 16, 0, 0, 0, 0, 0, K_NONE, false, "D16_UNORM" },
        { 56, 16, 0, 16, 0, 0, 0, K_U, true, "R16_UNORM" }, { 57, 16, 0, 16, 0, 0, 0, K_U, false, "R16_UINT" },
        { 58, 16, 0, 16, 0, 0, 0, K_S, true, "R16_SNORM" }, { 59, 16, 0, 16, 0, 0, 0, K_S, false, "R16_SINT" },
        { 60, 8, 0, 0, 0, 0, 0, K_NONE, false, "R8_TYPELESS" }, { 61, 8, 0, 8, 0, 0, 0, K_U, true, "R8_UNORM" }, { 62, 8, 0, 8, 0, 0, 0, K_U, false, "R8_UINT" },
        { 63, 8, 0, 8, 0, 0, 0, K_S, true, "R8_SNORM" }, { 64, 8, 0, 8, 0, 0, 0, K_S, false, "R8_SINT" }, { 65, 8, 0, 8, 0, 0, 0, K_U, true, "A8_UNORM" },
        { 66, 1, 0, 0, 0, 0, 0, K_NONE, false, "R1_UNORM" }, { 67, 32, 0, 0, 0, 0, 0, K_NONE, false, "R9G9B9E5_SHAREDEXP" },
        { 68, 32, 4, 0, 0, 0, 0, K_NONE, false, "R8G8_B8G8_UNORM" }, { 69, 32, 4, 0, 0, 0, 0, K_NONE, false, "G8R8_G8B8_UNORM" },
        { 70, 4, 8, 0, 0, 0, 0, K_NONE, false, "BC1_TYPELESS" }, { 71, 4, 8, 8, 8, 8, 8, K_BC1, true, "BC1_UNORM" }, { 72, 4, 8, 8, 8, 8, 8, K_BC1, true, "BC1_UNORM_SRGB" },
        { 73, 8, 16, 0, 0, 0, 0, K_NONE, false, "BC2_TYPELESS" }, { 74, 8, 16, 8, 8, 8, 8, K_BC2, true, "BC2_UNORM" }, { 75, 8, 16, 8, 8, 8, 8, K_BC2, true, "BC2_UNORM_SRGB" },
        { 76, 8, 16, 0, 0, 0, 0, K_NONE, false, "BC3_TYPELESS" }, { 77, 8, 16, 8, 8, 8, 8, K_BC3, true, "BC3_UNORM" }, { 78, 8, 16, 8, 8, 8, 8, K_BC3, true, "BC3_UNORM_SRGB" },
        { 79, 4, 8, 0, 0, 0, 0, K_NONE, false, "BC4_TYPELESS" }, { 80, 4, 8, 8, 0, 0, 0, K_BC4U, true, "BC4_UNORM" }, { 81, 4, 8, 8, 0, 0, 0, K_BC4S, true, "BC4_SNORM" },
        { 82, 8, 16, 0, 0, 0, 0, K_NONE, false, "BC5_TYPELESS" }, { 83, 8, 16, 8, 8, 0, 0, K_BC5U, true, "BC5_UNORM" }, { 84, 8, 16, 8, 8, 0, 0, K_BC5S, true, "BC5_SNORM" },
        { 85, 16, 0, 0, 0, 0, 0, K_NONE, false, "B5G6R5_UNORM" }, { 86, 16, 0, 0, 0, 0, 0, K_NONE, false, "B5G5R5A1_UNORM" },
        { 87, 32, 0, 0, 0, 0, 0, K_NONE, false, "B8G8R8A8_UNORM" }, { 88, 32, 0, 0, 0, 0, 0, K_NONE, false, "B8G8R8X8_UNORM" },
        { 89, 32, 0, 0, 0, 0, 0, K_NONE, false, "R10G10B10_XR_BIAS_A2_UNORM" }, { 90, 32, 0, 0, 0, 0, 0, K_NONE, false, "B8G8R8A8_TYPELESS" },
        { 91, 32, 0, 0, 0, 0, 0, K_NONE, false, "B8G8R8A8_UNORM_SRGB" }, { 92, 32, 0, 0, 0, 0, 0, K_NONE, false, "B8G8R8X8_TYPELESS" },
        { 93, 32, 0, 0, 0, 0, 0, K_NONE, false, "B8G8R8X8_UNORM_SRGB" },
        { 94, 8, 16, 0, 0, 0, 0, K_NONE, false, "BC6H_TYPELESS" }, { 95, 8, 16, 16, 16, 16, 0, K_BC6U, false, "BC6H_UF16" }, { 96, 8, 16, 16, 16, 16, 0, K_BC6S, false, "BC6H_SF16" },
        { 97, 8, 16, 0, 0, 0, 0, K_NONE, false, "BC7_TYPELESS" }, { 98, 8, 16, 8, 8, 8, 8, K_BC7, true, "BC7_UNORM" }, { 99, 8, 16, 8, 8, 8, 8, K_BC7, true, "BC7_UNORM_SRGB" },
        { 115, 16, 0, 0, 0, 0, 0, K_NONE, false, "B4G4R4A4_UNORM" },
    };

    constexpr uint32_t DDPF_ALPHA = 0x2u, DDPF_RGB = 0x40u, DDPF_LUMINANCE = 0x20000u, DDPF_BUMPDUDV = 0x80000u;
    constexpr uint32_t DDSD_HEIGHT = 0x2u, DDSD_DEPTH = 0x800000u;
    constexpr uint32_t DDSCAPS2_CUBEMAP = 0x200u;
    constexpr uint32_t FourCC(char a, char b, char c, char d) { return (uint32_t)(uint8_t)a | ((uint32_t)(uint8_t)b << 8) | ((uint32_t)(uint8_t)c << 16) | ((uint32_t)(uint8_t)d << 24); }

    // legacy DDS_PIXELFORMAT -> DXGI (DDSLoaderHelpers.h:512-755), as data: {flag, bit count (0 = any), masks R G B A, dxgi}
    struct LegacyMask { uint32_t flag, bits, r, g, b, a, dxgi; };
    const LegacyMask kLegacyMasks[] = {
        { DDPF_RGB, 32, 0x000000ffu, 0x0000ff00u, 0x00ff0000u, 0xff000000u, 28 }, { DDPF_RGB, 32, 0x00ff0000u, 0x0000ff00u, 0x000000ffu, 0xff000000u, 87 },
        { DDPF_RGB, 32, 0x00ff0000u, 0x0000ff00u, 0x000000ffu, 0, 88 }, { DDPF_RGB, 32, 0x3ff00000u, 0x000ffc00u, 0x000003ffu, 0xc0000000u, 24 },
        { DDPF_RGB, 32, 0x0000ffffu, 0xffff0000u, 0, 0, 35 }, { DDPF_RGB, 32, 0xffffffffu, 0, 0, 0, 41 },
        { DDPF_RGB, 16, 0x7c00u, 0x03e0u, 0x001fu, 0x8000u, 86 }, { DDPF_RGB, 16, 0xf800u, 0x07e0u, 0x001fu, 0, 85 }, { DDPF_RGB, 16, 0x0f00u, 0x00f0u, 0x000fu, 0xf000u, 115 },
        { DDPF_LUMINANCE, 8, 0xffu, 0, 0, 0, 61 }, { DDPF_LUMINANCE, 8, 0x00ffu, 0, 0, 0xff00u, 49 },
        { DDPF_LUMINANCE, 16, 0xffffu, 0, 0, 0, 56 }, { DDPF_LUMINANCE, 16, 0x00ffu, 0, 0, 0xff00u, 49 },
        { DDPF_BUMPDUDV, 16, 0x00ffu, 0xff00u, 0, 0, 51 }, { DDPF_BUMPDUDV, 32, 0x000000ffu, 0x0000ff00u, 0x00ff0000u, 0xff000000u, 31 },
        { DDPF_BUMPDUDV, 32, 0x0000ffffu, 0xffff0000u, 0, 0, 37 },
    };
    struct LegacyFourCC { uint32_t code, dxgi; };
    const LegacyFourCC kLegacyFourCC[] = {
        { FourCC('D', 'X', 'T', '1'), 71 }, { FourCC('D', 'X', 'T', '2'), 74 }, { FourCC('D', 'X', 'T', '3'), 74 }, { FourCC('D', 'X', 'T', '4'), 77 },
        { FourCC('D', 'X', 'T', '5'), 77 }, { FourCC('A', 'T', 'I', '1'), 80 }, { FourCC('B', 'C', '4', 'U'), 80 }, { FourCC('B', 'C', '4', 'S'), 81 },
        { FourCC('A', 'T', 'I', '2'), 83 }, { FourCC('B', 'C', '5', 'U'), 83 }, { FourCC('B', 'C', '5', 'S'), 84 },
        { FourCC('R', 'G', 'B', 'G'), 68 }, { FourCC('G', 'R', 'G', 'B'), 69 },
        { 36, 11 }, { 110, 13 }, { 111, 54 }, { 112, 34 }, { 113, 10 }, { 114, 41 }, { 115, 16 }, { 116, 2 },
    };

    uint32_t LegacyToDxgi(const DdsPixelFormat& pf)
    {
        if (pf.flags & DDPF_RGB) { for (const auto& m : kLegacyMasks) if (m.flag == DDPF_RGB && m.bits == pf.RGBBitCount && m.r == pf.RBitMask && m.g == pf.GBitMask && m.b == pf.BBitMask && m.a == pf.ABitMask) return m.dxgi; return 0; }
        if (pf.flags & DDPF_LUMINANCE) { for (const auto& m : kLegacyMasks) if (m.flag == DDPF_LUMINANCE && m.bits == pf.RGBBitCount && m.r == pf.RBitMask && m.g == pf.GBitMask && m.b == pf.BBitMask && m.a == pf.ABitMask) return m.dxgi; return 0; }
        if (pf.flags & DDPF_ALPHA) return pf.RGBBitCount == 8 ? 65u : 0u;
        if (pf.flags & DDPF_BUMPDUDV) { for (const auto& m : kLegacyMasks) if (m.flag == DDPF_BUMPDUDV && m.bits == pf.RGBBitCount && m.r == pf.RBitMask && m.g == pf.GBitMask && m.b == pf.BBitMask && m.a == pf.ABitMask) return m.dxgi; return 0; }
        if (pf.flags & kFourCC) { for (const auto& f : kLegacyFourCC) if (f.code == pf.fourCC) return f.dxgi; }
        return 0;
    }

    // bytes of one 2D slice and of one row (of blocks) - DDSLoaderHelpers.h:350-510 GetSurfaceInfo, for the formats of the table
    void SurfaceBytes(const DxgiFormatInfo& f, uint64_t w, uint64_t h, uint64_t& row, uint64_t& slice)
    {
        if (f.block_bytes >= 8) { const uint64_t bw = w ? (w + 3) / 4 : 0, bh = h ? (h + 3) / 4 : 0; row = (bw ? bw : (w ? 1 : 0)) * f.block_bytes; slice = row * bh; }
        else if (f.block_bytes == 4) { row = ((w + 1) >> 1) * 4; slice = row * h; }   // 2x1-packed
        else { row = (w * f.bits_per_pixel + 7) / 8; slice = row * h; }
    }
}

bool LookupDxgiFormat(uint32_t dxgi, DxgiFormatInfo& out)
{
    for (const Row& r : kFormats)
        if (r.dxgi == dxgi)
        {
            out.dxgi = r.dxgi; out.bits_per_pixel = r.bpp; out.block_bytes = r.block; out.cx = r.x; out.cy = r.y; out.cz = r.z; out.cw = r.w;
            out.cuda_kind = r.kind; out.normalized = r.norm; out.name = r.name;
            return true;
        }
    return false;
}

int ParseDds(const void* bytes, size_t size, DdsImage& out, std::string& error)
{
    out = DdsImage();
    const uint8_t* p = static_cast<const uint8_t*>(bytes);
    if (!p || size < 4 + sizeof(DdsHeader)) { error = "DDS: file too small"; return RT_E_IO; }
    uint32_t magic; memcpy(&magic, p, 4);
    DdsHeader h; memcpy(&h, p + 4, sizeof(h));
    if (magic != kMagic || h.size != 124 || h.ddspf.size != 32) { error = "DDS: bad magic/header"; return RT_E_IO; }
    size_t offset = 4 + sizeof(DdsHeader);
    uint32_t width = h.width, height = h.height, depth = h.depth, array = 1, dxgi = 0;
    int dims = 0;
    bool cube = false;
    const uint32_t mips = h.mipMapCount ? h.mipMapCount : 1;
    if ((h.ddspf.flags & kFourCC) && h.ddspf.fourCC == kFourCC_DX10)
    {
        if (size < offset + 20) { error = "DDS: truncated DX10 header"; return RT_E_IO; }
        uint32_t dx10[5]; memcpy(dx10, p + offset, 20); offset += 20;  // dxgiFormat, resourceDimension, miscFlag, arraySize, miscFlags2
        array = dx10[3];
        if (array == 0) { error = "DDS: DX10 arraySize is 0"; return RT_E_IO; }
        dxgi = dx10[0];
        if (dxgi >= 111 && dxgi <= 114) { error = "DDS: video/palette formats are not supported"; return RT_E_IO; }
        switch (dx10[1])
        {
        case 2: if ((h.flags & DDSD_HEIGHT) && height != 1) { error = "DDS: 1D texture with height != 1"; return RT_E_IO; } height = depth = 1; dims = 1; break;
        case 3: if (dx10[2] & 0x4u) { array *= 6; cube = true; } depth = 1; dims = 2; break;
        case 4: if (!(h.flags & DDSD_DEPTH)) { error = "DDS: 3D texture without the depth flag"; return RT_E_IO; }
                if (array > 1) { error = "DDS: volume textures are not texture arrays"; return RT_E_IO; } dims = 3; break;
        default: error = "DDS: unsupported resource dimension"; return RT_E_IO;
        }
    }
    else
    {
        dxgi = LegacyToDxgi(h.ddspf);
        if (!dxgi) { error = "DDS: legacy pixel format has no DXGI equivalent"; return RT_E_IO; }
        if (h.flags & DDSD_DEPTH) dims = 3;
        else
        {
            if (h.caps2 & DDSCAPS2_CUBEMAP)
            {
                if ((h.caps2 & kCubemapAllFaces) != kCubemapAllFaces) { error = "DDS: partial cubemaps are not supported"; return RT_E_IO; }
                array = 6; cube = true;
            }
            depth = 1; dims = 2;
        }
    }
    DxgiFormatInfo f;
    if (!LookupDxgiFormat(dxgi, f) || f.bits_per_pixel == 0) { error = "DDS: unknown DXGI format " + std::to_string(dxgi); return RT_E_IO; }
    if (mips > 16) { error = "DDS: too many mip levels"; return RT_E_IO; }
    const bool too_big = dims == 3 ? (array > 1 || width > 2048 || height > 2048 || depth > 2048)
                                   : (array > 2048 || width > 16384 || (dims == 2 && height > 16384));
    if (too_big || width == 0 || height == 0 || depth == 0) { error = "DDS: resource dimensions out of range"; return RT_E_IO; }

    out.width = width; out.height = height; out.depth = depth; out.mip_count = mips; out.array_size = array;
    out.is_cubemap = cube; out.num_dimensions = dims; out.format = f;
    out.subresources.reserve((size_t)mips * array);
    for (uint32_t e = 0; e < array; e++)
    {
        uint64_t w = width, hh = height, d = depth;
        for (uint32_t m = 0; m < mips; m++)
        {
            uint64_t row, slice; SurfaceBytes(f, w, hh, row, slice);
            if (row > 0xFFFFFFFFull || slice > 0xFFFFFFFFull) { error = "DDS: surface too large"; return RT_E_IO; }
            if (offset + slice * d > size) { error = "DDS: truncated pixel data"; return RT_E_IO; }
            DdsSubresource s; s.offset = offset; s.row_pitch = (size_t)row; s.slice_pitch = (size_t)slice; s.width = (uint32_t)w; s.height = (uint32_t)hh; s.depth = (uint32_t)d;
            out.subresources.push_back(s);
            offset += (size_t)(slice * d);
            w = w > 1 ? w >> 1 : 1; hh = hh > 1 ? hh >> 1 : 1; d = d > 1 ? d >> 1 : 1;
        }
    }
    return RT_OK;
}
