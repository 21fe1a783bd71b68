// K10: softmax cross-entropy over bf16 logits with ignore_index, mean over the labelled rows -
// the masked-LM loss of BASELINE.json config 3 (BERT), where 85 % of the rows carry no label.
//
// The torch composition (logits.float() -> log_softmax -> nll_loss, and the same again in
// backward, then a cast back to bf16) moves ~6 GB for an 8192 x 30528 logits matrix. Here the
// forward reads only the labelled rows once (online max / sum-exp, fp32) and keeps one fp32
// log-sum-exp per row; the backward re-reads those rows and writes bf16 dlogits directly,
// storing zeros for unlabelled rows without reading them.
//
// One 256-thread block per row; 16-byte vector accesses when the row pitch allows it.
// The reference ships no model code; the op is defined by BASELINE.json config 3.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ float block_max(float v, float* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    v = l < kThreads / 32 ? red[l] : -INFINITY;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return __shfl_sync(0xffffffffu, v, 0);
}
__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    v = l < kThreads / 32 ? red[l] : 0.f;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return __shfl_sync(0xffffffffu, v, 0);
}

// row_loss[r] = (lse - logit[label]) * inv_n for labelled rows, 0 otherwise; lse[r] saved.
__global__ void __launch_bounds__(kThreads)
xent_fwd_kernel(const __nv_bfloat16* __restrict__ logits, long ld, const long long* __restrict__ labels,
                int V, long long ignore_index, const float* __restrict__ inv_n,
                float* __restrict__ row_loss, float* __restrict__ lse_out) {
    __shared__ float red[kThreads / 32];
    const int row = blockIdx.x;
    const long long label = labels[row];
    if (label == ignore_index || label < 0 || label >= V) {
        if (threadIdx.x == 0) {
            row_loss[row] = 0.f;
            lse_out[row] = 0.f;
        }
        return;
    }
    const __nv_bfloat16* x = logits + (size_t)row * ld;
    const bool vec_ok = (ld % 8 == 0) && ((((uintptr_t)logits) & 15) == 0);
    const int nvec = vec_ok ? V / 8 : 0;
    // online max / sum-exp per thread, merged across the block afterwards
    float m = -INFINITY, s = 0.f;
    for (int v = threadIdx.x; v < nvec; v += kThreads) {
        const uint4 raw = __ldg(reinterpret_cast<const uint4*>(x) + v);
        float f[8];
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), f);
        float mx = f[0];
#pragma unroll
        for (int j = 1; j < 8; ++j) mx = fmaxf(mx, f[j]);
        if (mx > m) {
            s *= __expf(m - mx);
            m = mx;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) s += __expf(f[j] - m);
    }
    for (int i = nvec * 8 + threadIdx.x; i < V; i += kThreads) {
        const float f = __bfloat162float(x[i]);
        if (f > m) {
            s *= __expf(m - f);
            m = f;
        }
        s += __expf(f - m);
    }
    const float gm = block_max(m, red);
    const float gs = block_sum(m == -INFINITY ? 0.f : s * __expf(m - gm), red);
    if (threadIdx.x == 0) {
        const float lse = gm + __logf(gs);
        lse_out[row] = lse;
        row_loss[row] = (lse - __bfloat162float(x[label])) * inv_n[0];
    }
}

// dlogits[r, j] = (exp(x - lse) - [j == label]) * inv_n * grad_out for labelled rows, 0 otherwise
__global__ void __launch_bounds__(kThreads)
xent_bwd_kernel(const __nv_bfloat16* __restrict__ logits, long ld, const long long* __restrict__ labels,
                int V, long long ignore_index, const float* __restrict__ inv_n,
                const float* __restrict__ lse_in, const float* __restrict__ grad_out,
                __nv_bfloat16* __restrict__ dlogits, long ld_out) {
    const int row = blockIdx.x;
    const long long label = labels[row];
    __nv_bfloat16* d = dlogits + (size_t)row * ld_out;
    const bool out_vec = (ld_out % 8 == 0) && ((((uintptr_t)dlogits) & 15) == 0);
    if (label == ignore_index || label < 0 || label >= V) {
        const int nv = out_vec ? V / 8 : 0;
        for (int v = threadIdx.x; v < nv; v += kThreads)
            reinterpret_cast<uint4*>(d)[v] = make_uint4(0, 0, 0, 0);
        for (int i = nv * 8 + threadIdx.x; i < V; i += kThreads) d[i] = __float2bfloat16_rn(0.f);
        return;
    }
    const __nv_bfloat16* x = logits + (size_t)row * ld;
    const bool in_vec = (ld % 8 == 0) && ((((uintptr_t)logits) & 15) == 0);
    const float lse = lse_in[row];
    const float scale = inv_n[0] * grad_out[0];
    const int nvec = (in_vec && out_vec) ? V / 8 : 0;
    const int lab = (int)label;
    for (int v = threadIdx.x; v < nvec; v += kThreads) {
        const uint4 raw = __ldcs(reinterpret_cast<const uint4*>(x) + v);
        float f[8];
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float p = __expf(f[j] - lse);
            f[j] = (p - ((v * 8 + j) == lab ? 1.f : 0.f)) * scale;
        }
        bf16x8 pk = f32_to_bf16x8(f);
        reinterpret_cast<uint4*>(d)[v] = *reinterpret_cast<uint4*>(&pk);
    }
    for (int i = nvec * 8 + threadIdx.x; i < V; i += kThreads) {
        const float p = __expf(__bfloat162float(x[i]) - lse);
        d[i] = __float2bfloat16_rn((p - (i == lab ? 1.f : 0.f)) * scale);
    }
}

}  // namespace

// Forward: row_loss / lse fp32 [rows]; inv_n is a device scalar = 1 / max(#labelled rows, 1).
// Labels are int64; rows whose label is ignore_index (or out of range) contribute nothing.
TFOP_API int tfop_xent_fwd(const void* logits, long ld, const void* labels, int rows, int V,
                           long long ignore_index, const float* inv_n, float* row_loss,
                           float* lse, cudaStream_t stream) {
    TFOP_REQUIRE(rows > 0 && V > 0 && ld >= V, "xent: bad shape");
    xent_fwd_kernel<<<rows, kThreads, 0, stream>>>((const __nv_bfloat16*)logits, ld,
                                                   (const long long*)labels, V, ignore_index,
                                                   inv_n, row_loss, lse);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// Backward: dlogits bf16 [rows, V] (pitch ld_out); grad_out is the device scalar dL/dloss.
TFOP_API int tfop_xent_bwd(const void* logits, long ld, const void* labels, int rows, int V,
                           long long ignore_index, const float* inv_n, const float* lse,
                           const float* grad_out, void* dlogits, long ld_out,
                           cudaStream_t stream) {
    TFOP_REQUIRE(rows > 0 && V > 0 && ld >= V && ld_out >= V, "xent: bad shape");
    xent_bwd_kernel<<<rows, kThreads, 0, stream>>>((const __nv_bfloat16*)logits, ld,
                                                   (const long long*)labels, V, ignore_index,
                                                   inv_n, lse, grad_out, (__nv_bfloat16*)dlogits,
                                                   ld_out);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
