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
