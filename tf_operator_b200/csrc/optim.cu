// Fused flat-buffer optimizers (one launch for the whole model).
//
// The replica runtime keeps all parameters in flat buffers: fp32 master
// weights, fp32 optimizer state, bf16 model weights (what the convs / GEMMs
// read) and a bf16 gradient bucket that lives in symmetric peer memory (what
// the all-reduce reduces in place). One step = one pass over those buffers.
// Hyper-parameters are read from device memory so a captured CUDA graph keeps
// following the LR schedule.
//
// Reference parity: the optimizer the example workloads apply after the
// gradient exchange (Adam in examples/v1/dist-mnist/dist_mnist.py:194-208 and
// keras-API/multi_worker_strategy-with-keras.py:66-70); new code, no
// counterpart kernel in the reference.
#include "common.cuh"

namespace {

// hp layout (device float[8]): lr, momentum, weight_decay, grad_scale, beta1, beta2, eps, step
enum { HP_LR = 0, HP_MOM, HP_WD, HP_GSCALE, HP_B1, HP_B2, HP_EPS, HP_STEP };

template <bool GRAD_BF16>
__global__ void __launch_bounds__(256) sgd_momentum_kernel(
    float* __restrict__ master, float* __restrict__ mom, const void* __restrict__ grad,
    __nv_bfloat16* __restrict__ weights, size_t n, size_t n_decay, const float* __restrict__ hp,
    int nesterov) {
    const float lr = hp[HP_LR], mu = hp[HP_MOM], wd = hp[HP_WD], gs = hp[HP_GSCALE];
    const size_t nv = n / 8;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += stride) {
        const size_t e0 = v * 8;
        float g[8], w[8], m[8];
        if (GRAD_BF16) {
            uint4 raw = ld_stream_v4((const __nv_bfloat16*)grad + e0);
            bf16x8_to_f32(*reinterpret_cast<bf16x8*>(&raw), g);
        } else {
            uint4 r0 = ld_stream_v4((const float*)grad + e0);
            uint4 r1 = ld_stream_v4((const float*)grad + e0 + 4);
            g[0] = __uint_as_float(r0.x); g[1] = __uint_as_float(r0.y);
            g[2] = __uint_as_float(r0.z); g[3] = __uint_as_float(r0.w);
            g[4] = __uint_as_float(r1.x); g[5] = __uint_as_float(r1.y);
            g[6] = __uint_as_float(r1.z); g[7] = __uint_as_float(r1.w);
        }
        float4 w0 = *reinterpret_cast<float4*>(master + e0);
        float4 w1 = *reinterpret_cast<float4*>(master + e0 + 4);
        float4 m0 = *reinterpret_cast<float4*>(mom + e0);
        float4 m1 = *reinterpret_cast<float4*>(mom + e0 + 4);
        w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w;
        w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
        m[0] = m0.x; m[1] = m0.y; m[2] = m0.z; m[3] = m0.w;
        m[4] = m1.x; m[5] = m1.y; m[6] = m1.z; m[7] = m1.w;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float decay = (e0 + i) < n_decay ? wd : 0.f;
            float gi = g[i] * gs + decay * w[i];
            m[i] = mu * m[i] + gi;
            const float upd = nesterov ? gi + mu * m[i] : m[i];
            w[i] -= lr * upd;
        }
        *reinterpret_cast<float4*>(master + e0) = make_float4(w[0], w[1], w[2], w[3]);
        *reinterpret_cast<float4*>(master + e0 + 4) = make_float4(w[4], w[5], w[6], w[7]);
        *reinterpret_cast<float4*>(mom + e0) = make_float4(m[0], m[1], m[2], m[3]);
        *reinterpret_cast<float4*>(mom + e0 + 4) = make_float4(m[4], m[5], m[6], m[7]);
        if (weights) {
            bf16x8 pk = f32_to_bf16x8(w);
            *reinterpret_cast<uint4*>(weights + e0) = *reinterpret_cast<uint4*>(&pk);
        }
    }
}

// Each thread owns 4 consecutive elements per trip: the fp32 streams move as one float4 per
// thread (a warp covers 512 contiguous bytes per instruction), the bf16 streams as 8 bytes.
// Two independent trips are in flight per iteration.
template <bool GRAD_BF16>
__global__ void __launch_bounds__(256) adamw_kernel(
    float* __restrict__ master, float* __restrict__ m1, float* __restrict__ m2,
    const void* __restrict__ grad, __nv_bfloat16* __restrict__ weights, size_t n, size_t n_decay,
    const float* __restrict__ hp) {
    const float lr = hp[HP_LR], wd = hp[HP_WD], gs = hp[HP_GSCALE];
    const float b1 = hp[HP_B1], b2 = hp[HP_B2], eps = hp[HP_EPS], step = hp[HP_STEP];
    // -expm1(step * log(beta)) = 1 - beta^step without cancellation (beta2 = 0.999: 1e-3 at step 1)
    const float c1 = -1.f / expm1f(step * logf(b1));
    const float c2 = -1.f / expm1f(step * logf(b2));
    const float omb1 = 1.f - b1, omb2 = 1.f - b2;
    const size_t nv = n / 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    constexpr int U = 2;
    for (size_t v0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v0 < nv; v0 += U * stride) {
        float4 w[U], a[U], b[U], g[U];
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t v = v0 + u * stride;
            ok[u] = v < nv;
            if (ok[u]) {
                w[u] = __ldcs(reinterpret_cast<const float4*>(master) + v);
                a[u] = __ldcs(reinterpret_cast<const float4*>(m1) + v);
                b[u] = __ldcs(reinterpret_cast<const float4*>(m2) + v);
                if (GRAD_BF16) {
                    const uint2 raw = __ldcs(reinterpret_cast<const uint2*>(grad) + v);
                    const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
                    const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
                    g[u] = make_float4(lo.x, lo.y, hi.x, hi.y);
                } else {
                    g[u] = __ldcs(reinterpret_cast<const float4*>(grad) + v);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (!ok[u]) continue;
            const size_t v = v0 + u * stride;
            const size_t e0 = v * 4;
            float* wf = reinterpret_cast<float*>(&w[u]);
            float* af = reinterpret_cast<float*>(&a[u]);
            float* bf = reinterpret_cast<float*>(&b[u]);
            const float* gf = reinterpret_cast<const float*>(&g[u]);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float gi = gf[i] * gs;
                af[i] = b1 * af[i] + omb1 * gi;
                bf[i] = b2 * bf[i] + omb2 * gi * gi;
                const float decay = (e0 + i) < n_decay ? wd : 0.f;
                wf[i] -= lr * ((af[i] * c1) / (sqrtf(bf[i] * c2) + eps) + decay * wf[i]);
            }
            __stcs(reinterpret_cast<float4*>(master) + v, w[u]);
            __stcs(reinterpret_cast<float4*>(m1) + v, a[u]);
            __stcs(reinterpret_cast<float4*>(m2) + v, b[u]);
            if (weights) {
                __nv_bfloat162 lo = __floats2bfloat162_rn(wf[0], wf[1]);
                __nv_bfloat162 hi = __floats2bfloat162_rn(wf[2], wf[3]);
                uint2 pk;
                pk.x = *reinterpret_cast<uint32_t*>(&lo);
                pk.y = *reinterpret_cast<uint32_t*>(&hi);
                // the next forward reads these: keep them cacheable
                *(reinterpret_cast<uint2*>(weights) + v) = pk;
            }
        }
    }
}

__global__ void __launch_bounds__(256) zero_kernel(uint4* __restrict__ p, size_t nvec) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride)
        p[v] = make_uint4(0, 0, 0, 0);
}

int grid_for(size_t nvec) {
    size_t b = (nvec + 255) / 256;
    const size_t cap = 148 * 8;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace

// n must be a multiple of 8 (flat buffers are padded by the runtime). The first
// n_decay elements get weight decay, the rest (BN / bias region) do not.
TFOP_API int tfop_sgd_momentum(float* master, float* mom, const void* grad, int grad_dtype,
                               void* weights_bf16, size_t n, size_t n_decay, const float* hp_dev,
                               int nesterov, cudaStream_t stream) {
    TFOP_REQUIRE(n % 8 == 0, "flat optimizer buffers must be padded to a multiple of 8 elements");
    TFOP_REQUIRE(grad_dtype == TFOP_BF16 || grad_dtype == TFOP_F32, "bad grad dtype");
    if (n == 0) return 0;
    const int grid = grid_for(n / 8);
    if (grad_dtype == TFOP_BF16)
        sgd_momentum_kernel<true><<<grid, 256, 0, stream>>>(
            master, mom, grad, (__nv_bfloat16*)weights_bf16, n, n_decay, hp_dev, nesterov);
    else
        sgd_momentum_kernel<false><<<grid, 256, 0, stream>>>(
            master, mom, grad, (__nv_bfloat16*)weights_bf16, n, n_decay, hp_dev, nesterov);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

TFOP_API int tfop_adamw(float* master, float* m1, float* m2, const void* grad, int grad_dtype,
                        void* weights_bf16, size_t n, size_t n_decay, const float* hp_dev,
                        cudaStream_t stream) {
    TFOP_REQUIRE(n % 8 == 0, "flat optimizer buffers must be padded to a multiple of 8 elements");
    TFOP_REQUIRE(grad_dtype == TFOP_BF16 || grad_dtype == TFOP_F32, "bad grad dtype");
    if (n == 0) return 0;
    const int grid = grid_for(n / 8);  // two 4-element trips per thread and iteration
    if (grad_dtype == TFOP_BF16)
        adamw_kernel<true><<<grid, 256, 0, stream>>>(master, m1, m2, grad,
                                                     (__nv_bfloat16*)weights_bf16, n, n_decay,
                                                     hp_dev);
    else
        adamw_kernel<false><<<grid, 256, 0, stream>>>(master, m1, m2, grad,
                                                      (__nv_bfloat16*)weights_bf16, n, n_decay,
                                                      hp_dev);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// Zero a 16-byte-aligned region (used for the gradient bucket between steps).
TFOP_API int tfop_zero(void* p, size_t bytes, cudaStream_t stream) {
    TFOP_REQUIRE((((uintptr_t)p) & 15) == 0 && bytes % 16 == 0, "tfop_zero needs 16-byte alignment");
    if (bytes == 0) return 0;
    zero_kernel<<<grid_for(bytes / 16), 256, 0, stream>>>((uint4*)p, bytes / 16);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
