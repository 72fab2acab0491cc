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
