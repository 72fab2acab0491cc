// TEST INFRASTRUCTURE. C entry over the reference's own DDS container parser (utils/image/DDSLoader.cpp GetDDSSubresources +
// DDSLoaderHelpers.h LoadTextureDataFromMemory / BitsPerPixel / IsCompressed), compiled from /root/reference with the
// Windows.h stand-in of oracle/ref_shim/winshim. Same info layout as rt_dds_info (include/b200rt.h) so tests compare arrays.
#include "utils/image/DDSLoader.h"

extern "C" int ref_dds_info(const void* bytes, size_t size, uint64_t* info, int capacity)
{
    using namespace DirectX;
    const DDS_HEADER* header = nullptr; const uint8_t* bits = nullptr; size_t bit_size = 0;
    if (FAILED(LoadTextureDataFromMemory(static_cast<const uint8_t*>(bytes), size, &header, &bits, &bit_size))) return -1;
    auto sub = GetDDSSubresources(header, bits, bit_size);
    if (!sub) return -2;
    if (capacity < 12) return -3;
    info[0] = sub->width; info[1] = sub->height; info[2] = sub->depth; info[3] = sub->mip_count; info[4] = sub->array_size;
    info[5] = sub->is_cubemap; info[6] = sub->num_dimensions; info[7] = (uint64_t)sub->format; info[8] = BitsPerPixel(sub->format);
    info[9] = 0; info[10] = 0; info[11] = sub->subresources.size();
    const int n = (int)sub->subresources.size();
    for (int i = 0; i < n && 12 + 3 * i + 2 < capacity; i++)
    {
        info[12 + 3 * i] = (uint64_t)(static_cast<const uint8_t*>(sub->subresources[i].pSysMem) - static_cast<const uint8_t*>(bytes));
        info[13 + 3 * i] = sub->subresources[i].SysMemPitch; info[14 + 3 * i] = sub->subresources[i].SysMemSlicePitch;
    }
    return n;
}
