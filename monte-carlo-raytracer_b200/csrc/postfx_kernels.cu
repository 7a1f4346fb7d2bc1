// postfx_kernels.cu -- the two post-processing stages that follow reconstruction in the reference's
// pipelines (PathTracingApp.cpp:241-255), on float4 device buffers instead of GL-interop images:
//   k_bilateral_denoise   BilateralDenoise     assets/kernels/Denoise.cl:6-47
//   k_reinhard_tonemap    ReinhardToneMapping  assets/kernels/ToneMapping.cl:41-63
// Same expressions and the same tap order (x outer, y inner) as the reference, so the sums round alike;
// expf is CUDA's (<= 2 ulp from glibc's), which is the only source of difference.
#include "pt_kernels.h"

#include "pt_math.cuh"

namespace mcrt {

// The (2r+1)^2 taps of a pixel are read from a shared-memory tile (block + halo) that the CTA loads once
// with coalesced float4 reads; clamp-to-edge addressing is applied when the tile is filled, so a tap at
// a clamped coordinate finds the edge pixel at the position the unclamped tap would have had.
#define DN_BX 32
#define DN_BY 8

__global__ void __launch_bounds__(DN_BX * DN_BY)
k_bilateral_denoise(int width, int height, int radius, float sigmaSpatial, float sigmaRange,
                    const float4* __restrict__ src, float4* __restrict__ dst) {
    extern __shared__ float4 tile[];
    const int tw = DN_BX + 2 * radius, th = DN_BY + 2 * radius;
    const int x0 = blockIdx.x * DN_BX - radius, y0 = blockIdx.y * DN_BY - radius;
    for (int i = threadIdx.y * DN_BX + threadIdx.x; i < tw * th; i += DN_BX * DN_BY) {
        const int tx = i % tw, ty = i / tw;
        const int sx = min(max(x0 + tx, 0), width - 1), sy = min(max(y0 + ty, 0), height - 1);
        tile[i] = src[(size_t)sy * width + sx];
    }
    __syncthreads();
    const int gx = blockIdx.x * DN_BX + threadIdx.x, gy = blockIdx.y * DN_BY + threadIdx.y;
    if (gx >= width || gy >= height) return;
    const float sdSq = sigmaSpatial * sigmaSpatial;
    const float srSq = sigmaRange * sigmaRange;
    const float4 original = tile[(threadIdx.y + radius) * tw + threadIdx.x + radius];
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    float weightSum = 0.0f;
    for (int rx = -radius; rx <= radius; ++rx) {
        const int x = min(max(rx + gx, 0), width - 1);
        for (int ry = -radius; ry <= radius; ++ry) {
            const int y = min(max(ry + gy, 0), height - 1);
            const float4 k = tile[(y - y0) * tw + (x - x0)];
            const float dx = original.x - k.x, dy = original.y - k.y, dz = original.z - k.z, dw = original.w - k.w;
            const float d2 = dx * dx + dy * dy + dz * dz + dw * dw;     // dot(colorDiff, colorDiff), left to right
            const int sp = (gx - x) * (gx - x) + (gy - y) * (gy - y);
            const float w = expf(-sp / (2.0f * sdSq) - d2 / (2.0f * srSq));
            weightSum += w;
            acc.x += w * k.x;
            acc.y += w * k.y;
            acc.z += w * k.z;
            acc.w += w * k.w;
        }
    }
    dst[(size_t)gy * width + gx] = make_float4(acc.x / weightSum, acc.y / weightSum, acc.z / weightSum, acc.w / weightSum);
}

__global__ void __launch_bounds__(256)
k_reinhard_tonemap(int n, float Lwhite, const float4* __restrict__ src, float4* __restrict__ dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 rgba = src[i];
    const float L = 0.212671f * rgba.x + 0.715160f * rgba.y + 0.072169f * rgba.z;      // colors.cl:19-22
    const float tL = L * (1.0f + L / (Lwhite * Lwhite)) / (1.0f + L);                    // toneMapControlled
    const float s = tL / L;
    rgba.x *= s;
    rgba.y *= s;
    rgba.z *= s;
    dst[i] = rgba;
}

cudaError_t launch_bilateral_denoise(int width, int height, int radius, float sigmaSpatial, float sigmaRange,
                                     const float4* src, float4* dst, cudaStream_t stream) {
    // Denoise.cl:17-18: nothing is written for a degenerate filter
    if (radius <= 0 || sigmaSpatial <= 0.0f || sigmaRange <= 0.0f) return cudaSuccess;
    const dim3 block(DN_BX, DN_BY), grid((width + DN_BX - 1) / DN_BX, (height + DN_BY - 1) / DN_BY);
    const size_t smem = (size_t)(DN_BX + 2 * radius) * (DN_BY + 2 * radius) * sizeof(float4);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_bilateral_denoise, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    k_bilateral_denoise<<<grid, block, smem, stream>>>(width, height, radius, sigmaSpatial, sigmaRange, src, dst);
    return cudaGetLastError();
}

cudaError_t launch_reinhard_tonemap(int width, int height, float Lwhite, const float4* src, float4* dst,
                                    cudaStream_t stream) {
    const int n = width * height;
    k_reinhard_tonemap<<<(n + 255) / 256, 256, 0, stream>>>(n, Lwhite, src, dst);
    return cudaGetLastError();
}

}  // namespace mcrt
