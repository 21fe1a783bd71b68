// Input pipeline tail on the device: uint8 NHWC image batch (as it arrives from
// pinned host memory) -> normalised bf16 NHWC, one pass.
//
// Reference parity: the per-step feed of the example trainers
// (examples/v1/dist-mnist/dist_mnist.py:275-281 feeds fp32 from host every
// step); here the host sends 1 byte per value and the cast/normalise is fused
// on the device.
#include "common.cuh"

namespace {

struct Norm {
    float scale[4];  // 1 / (255 * std[c])
    float bias[4];   // -mean[c] / std[c]
};

__global__ void __launch_bounds__(256) u8_to_bf16_norm_kernel(const uint8_t* __restrict__ in,
                                                              __nv_bfloat16* __restrict__ out,
                                                              size_t n, int channels, Norm nm) {
    const size_t nv = n / 16;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += stride) {
        const uint4 raw = ld_stream_v4(in + v * 16);
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
        int c = (int)((v * 16) % channels);
        float f[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float x = (float)((w[i >> 2] >> (8 * (i & 3))) & 0xffu);
            f[i] = x * nm.scale[c] + nm.bias[c];
            c = (c + 1 == channels) ? 0 : c + 1;
        }
        bf16x8 lo = f32_to_bf16x8(f);
        bf16x8 hi = f32_to_bf16x8(f + 8);
        st_stream_v4(out + v * 16, *reinterpret_cast<uint4*>(&lo));
        st_stream_v4(out + v * 16 + 8, *reinterpret_cast<uint4*>(&hi));
    }
    // tail (n not a multiple of 16)
    if (blockIdx.x == 0) {
        for (size_t e = nv * 16 + threadIdx.x; e < n; e += blockDim.x) {
            const int c = (int)(e % channels);
            out[e] = __float2bfloat16_rn((float)in[e] * nm.scale[c] + nm.bias[c]);
        }
    }
}

}  // namespace

// in: uint8 [.., C] channel-last, n total values; mean/std: C floats (C <= 4) in [0,1] units.
TFOP_API int tfop_u8_to_bf16_normalize(const void* in_u8, void* out_bf16, size_t n, int channels,
                                       const float* mean, const float* stdv,
                                       cudaStream_t stream) {
    TFOP_REQUIRE(channels >= 1 && channels <= 4, "channels must be in [1, 4]");
    TFOP_REQUIRE((((uintptr_t)in_u8) & 15) == 0 && (((uintptr_t)out_bf16) & 15) == 0,
                 "buffers must be 16-byte aligned");
    if (n == 0) return 0;
    Norm nm;
    for (int c = 0; c < 4; ++c) {
        const float s = c < channels ? stdv[c] : 1.f;
        const float m = c < channels ? mean[c] : 0.f;
        nm.scale[c] = 1.f / (255.f * s);
        nm.bias[c] = -m / s;
    }
    size_t blocks = (n / 16 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    u8_to_bf16_norm_kernel<<<(int)blocks, 256, 0, stream>>>(
        (const uint8_t*)in_u8, (__nv_bfloat16*)out_bf16, n, channels, nm);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// Stem space-to-depth: uint8 [N,H,W,3] -> bf16 [N, H/2+3, W/2+3, 16], normalised.
// A 7x7 / stride-2 / pad-3 convolution over 3 channels equals a 4x4 / stride-1 / pad-0
// convolution over this tensor (2x2 pixel blocks folded into 12 channels, padded to 16;
// two zero block-rows/cols in front and one behind supply the padding), which lets the
// library pick tensor-core kernels instead of the 3-channel special case.
// out[n, j+2, l+2, (bh*2+bw)*3 + ch] = norm(in[n, 2j+bh, 2l+bw, ch])
// ---------------------------------------------------------------------------
namespace {

__global__ void __launch_bounds__(256) stem_s2d_kernel(const uint8_t* __restrict__ in,
                                                       __nv_bfloat16* __restrict__ out, int N,
                                                       int H, int W, Norm nm) {
    const int Hb = H / 2, Wb = W / 2;       // blocks
    const int Ho = Hb + 3, Wo = Wb + 3;     // padded block grid
    const unsigned int total = (unsigned int)N * Ho * Wo;
    const unsigned int stride = gridDim.x * blockDim.x;
    for (unsigned int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int l = (int)(t % Wo) - 2;
        unsigned int p = t / Wo;
        const int j = (int)(p % Ho) - 2;
        const int n = (int)(p / Ho);
        float f[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) f[i] = 0.f;
        if (j >= 0 && j < Hb && l >= 0 && l < Wb) {
#pragma unroll
            for (int bh = 0; bh < 2; ++bh) {
                // the two horizontally adjacent pixels are 6 consecutive bytes
                const uint8_t* src = in + (((size_t)n * H + (2 * j + bh)) * W + 2 * l) * 3;
#pragma unroll
                for (int k = 0; k < 6; ++k) {
                    const int ch = k % 3;
                    f[bh * 6 + k] = (float)src[k] * nm.scale[ch] + nm.bias[ch];
                }
            }
        }
        bf16x8 lo = f32_to_bf16x8(f);
        bf16x8 hi = f32_to_bf16x8(f + 8);
        uint4* dst = reinterpret_cast<uint4*>(out + (size_t)t * 16);
        dst[0] = *reinterpret_cast<uint4*>(&lo);
        dst[1] = *reinterpret_cast<uint4*>(&hi);
    }
}

}  // namespace

// in: uint8 [N,H,W,3] (H, W even); out: bf16 [N, H/2+3, W/2+3, 16].
TFOP_API int tfop_stem_space_to_depth(const void* in_u8, void* out_bf16, int N, int H, int W,
                                      const float* mean, const float* stdv,
                                      cudaStream_t stream) {
    TFOP_REQUIRE(N > 0 && H > 0 && W > 0 && H % 2 == 0 && W % 2 == 0, "H and W must be even");
    TFOP_REQUIRE((((uintptr_t)out_bf16) & 15) == 0, "output must be 16-byte aligned");
    Norm nm;
    for (int c = 0; c < 4; ++c) {
        const float s = c < 3 ? stdv[c] : 1.f;
        const float m = c < 3 ? mean[c] : 0.f;
        nm.scale[c] = 1.f / (255.f * s);
        nm.bias[c] = -m / s;
    }
    const size_t total = (size_t)N * (H / 2 + 3) * (W / 2 + 3);
    TFOP_REQUIRE(total < 0xffffffffull, "tensor too large for 32-bit indexing");
    size_t blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    stem_s2d_kernel<<<(int)blocks, 256, 0, stream>>>((const uint8_t*)in_u8,
                                                     (__nv_bfloat16*)out_bf16, N, H, W, nm);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
