// Shared device/host helpers for libtfop_b200 (sm_100a only).
//
// Everything in this library is exposed through a plain C ABI (see
// include/tfop_b200.h) so the Python side binds it with ctypes and no
// torch C++ headers are needed at build time.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define TFOP_API extern "C" __attribute__((visibility("default")))

// Thread-local last-error string, readable through tfop_last_error().
void tfop_set_error(const char* file, int line, const char* msg);

#define TFOP_CUDA(x)                                                      \
    do {                                                                  \
        cudaError_t e_ = (x);                                             \
        if (e_ != cudaSuccess) {                                          \
            tfop_set_error(__FILE__, __LINE__, cudaGetErrorString(e_));   \
            return (int)e_;                                               \
        }                                                                 \
    } while (0)

#define TFOP_REQUIRE(cond, msg)                                           \
    do {                                                                  \
        if (!(cond)) {                                                    \
            tfop_set_error(__FILE__, __LINE__, msg);                      \
            return -1;                                                    \
        }                                                                 \
    } while (0)

// dtype tags used across the ABI
enum { TFOP_F32 = 0, TFOP_BF16 = 1, TFOP_F16 = 2 };

#define TFOP_MAX_WORLD 8

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
#ifdef __CUDACC__

struct __align__(16) bf16x8 {
    __nv_bfloat162 v[4];
};

__device__ __forceinline__ void bf16x8_to_f32(const bf16x8& p, float* f) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(p.v[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}

__device__ __forceinline__ bf16x8 f32_to_bf16x8(const float* f) {
    bf16x8 p;
#pragma unroll
    for (int i = 0; i < 4; ++i) p.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return p;
}

// 16-byte global loads/stores. The *_peer variants are used on pointers that
// may map a peer GPU over NVLink: they are plain coherent accesses (peer
// memory bypasses the local L2, see B300_MICROARCH "NVLink").
// Read-only streaming loads. These are compiler intrinsics rather than asm
// volatile so ptxas is free to hoist a batch of them ahead of their uses (an
// asm volatile load was observed to be scheduled load-use-load-use, i.e. one
// request in flight per thread).
__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {   // evict-first: data read once
    return __ldcs(reinterpret_cast<const uint4*>(p));
}
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {       // default policy: re-read soon
    return __ldg(reinterpret_cast<const uint4*>(p));
}
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
    uint4 r;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ void st_peer_v4(void* p, const uint4& v) {
    asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
                 "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void st_stream_v4(void* p, const uint4& v) {
    __stcs(reinterpret_cast<uint4*>(p), v);
}

// system-scope flag signalling (K8 primitives)
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() {
    asm volatile("fence.acq_rel.sys;" ::: "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

#endif  // __CUDACC__
