// Multi-tensor pack: gather up to 128 gradient tensors (fp32 or bf16) into
// their slots of a flat bf16 bucket in one launch, with the cast and an
// optional scale fused. Segment descriptors travel as kernel parameters, so a
// captured CUDA graph needs no device-side table.
//
// Reference parity: the "pack the gradients into few tensors" step TF performs
// before the collective in examples/v1/distribution_strategy/keras-API/
// multi_worker_strategy-with-keras.py:76-77 (first row).
#include "common.cuh"

#define TFOP_PACK_MAX_SEGS 128

struct tfop_pack_seg {
    const void* src;
    unsigned long long dst_off;  // element offset in the flat destination (multiple of 8)
    unsigned long long n;        // elements
};

namespace {

struct PackArgs {
    tfop_pack_seg seg[TFOP_PACK_MAX_SEGS];
};

template <bool SRC_BF16>
__global__ void __launch_bounds__(256) multi_pack_kernel(const __grid_constant__ PackArgs a,
                                                         __nv_bfloat16* __restrict__ dst,
                                                         float scale) {
    const tfop_pack_seg s = a.seg[blockIdx.y];
    __nv_bfloat16* d = dst + s.dst_off;
    const size_t n = s.n;
    const size_t nv = n / 8;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const bool aligned = (((uintptr_t)s.src) & 15) == 0;
    if (aligned) {
        for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += stride) {
            float f[8];
            if (SRC_BF16) {
                const uint4 raw = ld_stream_v4((const __nv_bfloat16*)s.src + v * 8);
                bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), f);
            } else {
                const uint4 r0 = ld_stream_v4((const float*)s.src + v * 8);
                const uint4 r1 = ld_stream_v4((const float*)s.src + v * 8 + 4);
                f[0] = __uint_as_float(r0.x); f[1] = __uint_as_float(r0.y);
                f[2] = __uint_as_float(r0.z); f[3] = __uint_as_float(r0.w);
                f[4] = __uint_as_float(r1.x); f[5] = __uint_as_float(r1.y);
                f[6] = __uint_as_float(r1.z); f[7] = __uint_as_float(r1.w);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] *= scale;
            bf16x8 pk = f32_to_bf16x8(f);
            *reinterpret_cast<uint4*>(d + v * 8) = *reinterpret_cast<uint4*>(&pk);
        }
    }
    // scalar tail (or the whole segment when the source is not 16-byte aligned)
    const size_t start = aligned ? nv * 8 : 0;
    for (size_t e = start + (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        const float v = SRC_BF16 ? __bfloat162float(((const __nv_bfloat16*)s.src)[e])
                                 : ((const float*)s.src)[e];
        d[e] = __float2bfloat16_rn(v * scale);
    }
}

}  // namespace

// segs: host array of nseg descriptors (nseg <= 128). dst: flat bf16 buffer,
// 16-byte aligned; every dst_off a multiple of 8.
TFOP_API int tfop_multi_pack_bf16(const tfop_pack_seg* segs, int nseg, int src_dtype, void* dst,
                                  float scale, cudaStream_t stream) {
    TFOP_REQUIRE(nseg >= 0 && nseg <= TFOP_PACK_MAX_SEGS, "at most 128 segments per launch");
    TFOP_REQUIRE(src_dtype == TFOP_F32 || src_dtype == TFOP_BF16, "bad src dtype");
    if (nseg == 0) return 0;
    PackArgs a;
    memset(&a, 0, sizeof(a));
    size_t maxn = 0;
    for (int i = 0; i < nseg; ++i) {
        TFOP_REQUIRE(segs[i].dst_off % 8 == 0, "dst_off must be a multiple of 8 elements");
        a.seg[i] = segs[i];
        if (segs[i].n > maxn) maxn = segs[i].n;
    }
    size_t bx = (maxn / 8 + 255) / 256;
    if (bx < 1) bx = 1;
    if (bx > 64) bx = 64;
    dim3 grid((unsigned)bx, (unsigned)nseg);
    if (src_dtype == TFOP_BF16)
        multi_pack_kernel<true><<<grid, 256, 0, stream>>>(a, (__nv_bfloat16*)dst, scale);
    else
        multi_pack_kernel<false><<<grid, 256, 0, stream>>>(a, (__nv_bfloat16*)dst, scale);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
