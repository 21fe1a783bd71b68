// Vectorised typed element loads/stores shared by the collective kernels: VEC consecutive
// elements of an fp32 or bf16 array <-> VEC floats in registers, with a scalar tail path.
#pragma once
#include "common.cuh"

template <int VEC>
__device__ __forceinline__ void load_elems(const void* in, int dtype, size_t e0, size_t n,
                                           bool aligned, float* f) {
    if (aligned && e0 + VEC <= n) {
        if (dtype == TFOP_F32) {
            const float4* p = reinterpret_cast<const float4*>((const float*)in + e0);
#pragma unroll
            for (int i = 0; i < VEC / 4; ++i) {
                float4 t = p[i];
                f[4 * i] = t.x;
                f[4 * i + 1] = t.y;
                f[4 * i + 2] = t.z;
                f[4 * i + 3] = t.w;
            }
        } else {
            const __nv_bfloat162* p =
                reinterpret_cast<const __nv_bfloat162*>((const __nv_bfloat16*)in + e0);
            if (VEC == 8) {
                uint4 raw = *reinterpret_cast<const uint4*>(p);
                bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), f);
            } else {
                uint2 raw = *reinterpret_cast<const uint2*>(p);
                float2 a = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&raw.x));
                float2 b = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&raw.y));
                f[0] = a.x;
                f[1] = a.y;
                f[2] = b.x;
                f[3] = b.y;
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            size_t e = e0 + i;
            float v = 0.f;
            if (e < n)
                v = dtype == TFOP_F32 ? ((const float*)in)[e]
                                      : __bfloat162float(((const __nv_bfloat16*)in)[e]);
            f[i] = v;
        }
    }
}

template <int VEC>
__device__ __forceinline__ void store_elems(void* out, int dtype, size_t e0, size_t n,
                                            bool aligned, const float* f) {
    if (aligned && e0 + VEC <= n) {
        if (dtype == TFOP_F32) {
            float4* p = reinterpret_cast<float4*>((float*)out + e0);
#pragma unroll
            for (int i = 0; i < VEC / 4; ++i)
                p[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
        } else {
            if (VEC == 8) {
                bf16x8 pk = f32_to_bf16x8(f);
                *reinterpret_cast<uint4*>((__nv_bfloat16*)out + e0) =
                    *reinterpret_cast<uint4*>(&pk);
            } else {
                __nv_bfloat162 a = __floats2bfloat162_rn(f[0], f[1]);
                __nv_bfloat162 b = __floats2bfloat162_rn(f[2], f[3]);
                uint2 raw;
                raw.x = *reinterpret_cast<uint32_t*>(&a);
                raw.y = *reinterpret_cast<uint32_t*>(&b);
                *reinterpret_cast<uint2*>((__nv_bfloat16*)out + e0) = raw;
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            size_t e = e0 + i;
            if (e < n) {
                if (dtype == TFOP_F32)
                    ((float*)out)[e] = f[i];
                else
                    ((__nv_bfloat16*)out)[e] = __float2bfloat16_rn(f[i]);
            }
        }
    }
}

