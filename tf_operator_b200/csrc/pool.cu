// MaxPool 3x3 / stride 2 / pad 1 for channel-last bf16 (the ResNet stem pool).
//
// Forward writes the pooled tensor plus a 1-byte argmax code (0..8 = kh*3+kw)
// per element; backward is a pure gather (every input pixel looks at the <= 4
// windows that cover it), so there are no atomics and dx is written exactly
// once. Both kernels move 16-byte vectors of 8 channels.
//
// No counterpart in the reference (it ships no kernels); part of
// the compute side of the PyTorchJob ResNet-50 config.
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256) maxpool3x3s2_fwd_kernel(
    const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
    uint8_t* __restrict__ idx, int N, int H, int W, int C, int Ho, int Wo) {
    // 32-bit index arithmetic (the host checks the element count fits): 64-bit
    // divisions dominated the first version of this kernel
    const unsigned int CV = C >> 3;
    const unsigned int total = (unsigned int)N * Ho * Wo * CV;
    const unsigned int stride = gridDim.x * blockDim.x;
    for (unsigned int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int cv = (int)(t % CV);
        unsigned int p = t / CV;
        const int wo = (int)(p % Wo);
        p /= Wo;
        const int ho = (int)(p % Ho);
        const int n = (int)(p / Ho);
        float best[8];
        unsigned int code[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            best[i] = -INFINITY;
            code[i] = 0;
        }
        bool first = true;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int h = 2 * ho - 1 + kh;
            if (h < 0 || h >= H) continue;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int w = 2 * wo - 1 + kw;
                if (w < 0 || w >= W) continue;
                const uint4 raw =
                    ld_nc_v4(x + (((size_t)n * H + h) * W + w) * C + (size_t)cv * 8);
                float f[8];
                bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), f);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    // first valid tap initialises (so an all -inf window still gets a
                    // valid code); NaN wins like in torch
                    if (first || f[i] > best[i] || f[i] != f[i]) {
                        best[i] = f[i];
                        code[i] = kh * 3 + kw;
                    }
                }
                first = false;
            }
        }
        bf16x8 pk = f32_to_bf16x8(best);
        *reinterpret_cast<uint4*>(y + (size_t)t * 8) = *reinterpret_cast<uint4*>(&pk);
        uint2 cpk;
        cpk.x = code[0] | (code[1] << 8) | (code[2] << 16) | (code[3] << 24);
        cpk.y = code[4] | (code[5] << 8) | (code[6] << 16) | (code[7] << 24);
        *reinterpret_cast<uint2*>(idx + (size_t)t * 8) = cpk;
    }
}

__global__ void __launch_bounds__(256) maxpool3x3s2_bwd_kernel(
    const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx,
    __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int Ho, int Wo) {
    const unsigned int CV = C >> 3;
    const unsigned int total = (unsigned int)N * H * W * CV;
    const unsigned int stride = gridDim.x * blockDim.x;
    for (unsigned int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int cv = (int)(t % CV);
        unsigned int p = t / CV;
        const int w = (int)(p % W);
        p /= W;
        const int h = (int)(p % H);
        const int n = (int)(p / H);
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
        // windows covering row h: ho with 2*ho-1 <= h <= 2*ho+1
        const int ho_lo = h >> 1;              // ceil((h-1)/2) for h >= 0
        const int ho_hi = (h + 1) >> 1;        // floor((h+1)/2)
        const int wo_lo = w >> 1, wo_hi = (w + 1) >> 1;
        for (int ho = ho_lo; ho <= ho_hi; ++ho) {
            if (ho >= Ho) continue;
            const int kh = h - (2 * ho - 1);
            for (int wo = wo_lo; wo <= wo_hi; ++wo) {
                if (wo >= Wo) continue;
                const unsigned int me = (unsigned int)(kh * 3 + (w - (2 * wo - 1)));
                const size_t o = ((((size_t)n * Ho + ho) * Wo + wo) * CV + cv) * 8;
                const uint2 cpk = __ldg(reinterpret_cast<const uint2*>(idx + o));
                const uint4 raw = ld_nc_v4(dy + o);
                float g[8];
                bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), g);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const unsigned int cd = ((i < 4 ? cpk.x : cpk.y) >> (8 * (i & 3))) & 0xffu;
                    if (cd == me) acc[i] += g[i];
                }
            }
        }
        bf16x8 pk = f32_to_bf16x8(acc);
        *reinterpret_cast<uint4*>(dx + (size_t)t * 8) = *reinterpret_cast<uint4*>(&pk);
    }
}

int grid_for(size_t total) {
    size_t b = (total + 255) / 256;
    const size_t cap = 148 * 16;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace

// x: [N,H,W,C] bf16 (C % 8 == 0) -> y: [N,Ho,Wo,C], idx: uint8 [N,Ho,Wo,C];
// Ho = (H + 1) / 2, Wo = (W + 1) / 2  (3x3 window, stride 2, padding 1).
TFOP_API int tfop_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C,
                                   cudaStream_t stream) {
    TFOP_REQUIRE(C % 8 == 0 && C > 0, "channels must be a multiple of 8");
    TFOP_REQUIRE(N > 0 && H > 0 && W > 0, "empty tensor");
    TFOP_REQUIRE((size_t)N * H * W * (C / 8) < 0xffffffffull, "tensor too large for 32-bit indexing");
    const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
    const size_t total = (size_t)N * Ho * Wo * (C / 8);
    maxpool3x3s2_fwd_kernel<<<grid_for(total), 256, 0, stream>>>(
        (const __nv_bfloat16*)x, (__nv_bfloat16*)y, (uint8_t*)idx, N, H, W, C, Ho, Wo);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

TFOP_API int tfop_maxpool3x3s2_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W,
                                   int C, cudaStream_t stream) {
    TFOP_REQUIRE(C % 8 == 0 && C > 0, "channels must be a multiple of 8");
    TFOP_REQUIRE((size_t)N * H * W * (C / 8) < 0xffffffffull, "tensor too large for 32-bit indexing");
    const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
    const size_t total = (size_t)N * H * W * (C / 8);
    maxpool3x3s2_bwd_kernel<<<grid_for(total), 256, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const uint8_t*)idx, (__nv_bfloat16*)dx, N, H, W, C, Ho, Wo);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
