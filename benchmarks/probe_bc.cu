// Probe (not product): which extents do CUDA's block-compressed array formats take? 8x8 BC1 texture, one solid colour per 4x4 block.
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>
__global__ void sample(cudaTextureObject_t t, float4* out)
{
    const int i = threadIdx.x;  // 4 blocks
    const float u = ((i & 1) * 4 + 2.0f) / 8.0f, v = ((i >> 1) * 4 + 2.0f) / 8.0f;
    out[i] = tex2DLod<float4>(t, u, v, 0.0f);
}
static void block(uint8_t* b, uint16_t c565) { memcpy(b, &c565, 2); memcpy(b + 2, &c565, 2); memset(b + 4, 0, 4); }
int main()
{
    uint8_t data[4 * 8];
    block(data, 0xF800); block(data + 8, 0x07E0); block(data + 16, 0x001F); block(data + 24, 0xFFFF);  // R G / B W
    for (int mode = 0; mode < 2; mode++)
    {
        const cudaChannelFormatDesc d = cudaCreateChannelDesc<cudaChannelFormatKindUnsignedBlockCompressed1>();
        cudaMipmappedArray_t arr = nullptr;
        const int aw = mode == 0 ? 8 : 2, ah = mode == 0 ? 8 : 2;  // 0: texel extent, 1: block extent (what the reference passes)
        cudaError_t e = cudaMallocMipmappedArray(&arr, &d, make_cudaExtent(aw, ah, 0), 1, 0);
        printf("mode %d malloc(%d x %d): %s\n", mode, aw, ah, cudaGetErrorString(e));
        if (e != cudaSuccess) { cudaGetLastError(); continue; }
        cudaArray_t lvl; cudaGetMipmappedArrayLevel(&lvl, arr, 0);
        for (int cm = 0; cm < 3; cm++)
        {
            cudaMemcpy3DParms p = {};
            const int bw = 2, bh = 2;
            if (cm == 0) { p.srcPtr = make_cudaPitchedPtr(data, bw * 8, bw, bh); p.extent = make_cudaExtent(bw, bh, 1); }        // blocks
            if (cm == 1) { p.srcPtr = make_cudaPitchedPtr(data, bw * 8, 8, 8); p.extent = make_cudaExtent(8, 8, 1); }              // texels
            if (cm == 2) { e = cudaMemcpy2DToArray(lvl, 0, 0, data, bw * 8, bw * 8, bh, cudaMemcpyHostToDevice); }
            else { p.dstArray = lvl; p.kind = cudaMemcpyHostToDevice; e = cudaMemcpy3D(&p); }
            printf("  copy mode %d: %s\n", cm, cudaGetErrorString(e));
            cudaGetLastError();
            if (e != cudaSuccess) continue;
            cudaResourceDesc r; memset(&r, 0, sizeof(r)); r.resType = cudaResourceTypeMipmappedArray; r.res.mipmap.mipmap = arr;
            cudaTextureDesc t; memset(&t, 0, sizeof(t)); t.normalizedCoords = 1; t.filterMode = cudaFilterModePoint; t.readMode = cudaReadModeNormalizedFloat;
            cudaTextureObject_t tex = 0; e = cudaCreateTextureObject(&tex, &r, &t, nullptr);
            printf("  texobj: %s\n", cudaGetErrorString(e));
            if (e != cudaSuccess) { cudaGetLastError(); continue; }
            float4* o; cudaMalloc(&o, 64); sample<<<1, 4>>>(tex, o); float4 h[4]; e = cudaMemcpy(h, o, 64, cudaMemcpyDeviceToHost);
            printf("  sample: %s\n", cudaGetErrorString(e));
            for (int i = 0; i < 4; i++) printf("   block %d: %.3f %.3f %.3f %.3f\n", i, h[i].x, h[i].y, h[i].z, h[i].w);
            cudaDestroyTextureObject(tex); cudaFree(o);
        }
        cudaFreeMipmappedArray(arr);
    }
    return 0;
}
