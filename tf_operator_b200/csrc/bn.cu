// Fused BatchNorm (+ residual add) (+ ReLU) for channel-last bf16 activations.
//
// ResNet-50's step is half convolution (cuDNN, a library call) and half
// memory-bound BN / ReLU / add traffic. These kernels do the latter in the
// minimum number of HBM passes:
//   forward : stats pass (x read once; last block to finish folds the
//             per-block partials into mean/invstd, updates the running stats
//             and emits per-channel scale/shift) + apply pass (x [+ residual]
//             -> y, plus a 1-bit/element ReLU mask)
//   backward: reduce pass (dgamma, dbeta) + apply pass (dx [, dresidual])
// Layout: x is [M, C] with C contiguous (NHWC), C a multiple of 64 up to 4096.
// A block column owns 64 channels and a thread owns 8 consecutive channels for
// the whole kernel, so no per-element channel arithmetic is needed.
//
// Scratch: the cross-block accumulators and the "last block" ticket counters
// live in a caller-provided workspace (tfop_bn_workspace_bytes()).
//
// Reference parity: no counterpart in the reference (it ships no kernels,
// SURVEY.md 2.12); this is the compute side of BASELINE config 2 (ResNet-50).
#include "common.cuh"

namespace {

constexpr int kT = 256;          // 8 channel-vectors (64 channels) x 32 row lanes
constexpr int kGroupC = 64;      // channels per block column
constexpr int kMaxC = 4096;
constexpr int kMaxGroups = kMaxC / kGroupC;
constexpr int kTargetBlocks = 148 * 8;
constexpr int kU = 4;            // rows in flight per thread (memory-level parallelism)

__device__ __forceinline__ float ld_param(const void* p, int dtype, int i) {
    return dtype == TFOP_F32 ? ((const float*)p)[i]
                             : __bfloat162float(((const __nv_bfloat16*)p)[i]);
}
__device__ __forceinline__ void st_param(void* p, int dtype, int i, float v) {
    if (dtype == TFOP_F32)
        ((float*)p)[i] = v;
    else
        ((__nv_bfloat16*)p)[i] = __float2bfloat16_rn(v);
}

// Fold the 16 per-thread accumulators over the 32 row lanes of the block and
// add the block's 128 sums into the global accumulators (fp32 RED; 128
// distinct addresses per channel group, so contention is per-address only).
// accum layout: [2][C]
__device__ __forceinline__ void block_fold_atomic(const float* a, const float* b, int C,
                                                  int c_base, float* accum,
                                                  float* smem /*16*kT*/) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        smem[k * kT + threadIdx.x] = a[k];
        smem[(8 + k) * kT + threadIdx.x] = b[k];
    }
    __syncthreads();
    if (threadIdx.x < 128) {
        const int k = threadIdx.x >> 3;  // 0..15 : which accumulator
        const int tx = threadIdx.x & 7;  // channel vector inside the group
        float s = 0.f;
#pragma unroll 8
        for (int ty = 0; ty < 32; ++ty) s += smem[k * kT + ty * 8 + tx];
        const int which = k >> 3, kk = k & 7;
        atomicAdd(accum + which * C + c_base + tx * 8 + kk, s);
    }
}

// Returns true in exactly one block per channel group (the last to arrive),
// after every block's atomics are visible to it.
__device__ __forceinline__ bool last_block_ticket(unsigned int* counter, int* s_flag) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(counter, 1u);
        *s_flag = (t == gridDim.x - 1);
    }
    __syncthreads();
    const bool last = *s_flag != 0;
    if (last) __threadfence();
    return last;
}

__global__ void __launch_bounds__(kT, 4) bn_fwd_stats_kernel(
    const __nv_bfloat16* __restrict__ x, size_t M, int C, float* __restrict__ accum,
    unsigned int* __restrict__ counters, const void* __restrict__ gamma,
    const void* __restrict__ beta, int param_dtype, float* __restrict__ running_mean,
    float* __restrict__ running_var, float momentum, float eps, float* __restrict__ scale_shift,
    float* __restrict__ save_mean, float* __restrict__ save_invstd) {
    __shared__ float smem[16 * kT];
    __shared__ int s_flag;
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    const int c_base = blockIdx.y * kGroupC;
    const __nv_bfloat16* xp = x + c_base + tx * 8;
    float s[8], q[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = q[i] = 0.f;
    const size_t rstride = (size_t)gridDim.x * 32;
    for (size_t r = (size_t)blockIdx.x * 32 + ty; r < M; r += rstride * kU) {
        uint4 raw[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const size_t rr = r + (size_t)u * rstride;
            raw[u] = rr < M ? ld_nc_v4(xp + rr * C) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            float f[8];
            bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw[u]), f);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                s[i] += f[i];
                q[i] = fmaf(f[i], f[i], q[i]);
            }
        }
    }
    block_fold_atomic(s, q, C, c_base, accum, smem);
    if (!last_block_ticket(counters + blockIdx.y, &s_flag)) return;
    if (threadIdx.x < kGroupC) {
        const int c = c_base + threadIdx.x;
        const float S = atomicExch(accum + c, 0.f);
        const float Q = atomicExch(accum + C + c, 0.f);
        const float invM = 1.f / (float)M;
        const float mean = S * invM;
        float var = fmaf(-mean, mean, Q * invM);
        var = fmaxf(var, 0.f);
        const float invstd = rsqrtf(var + eps);
        const float g = ld_param(gamma, param_dtype, c);
        const float bt = ld_param(beta, param_dtype, c);
        const float a = g * invstd;
        scale_shift[c] = a;
        scale_shift[C + c] = fmaf(-mean, a, bt);
        save_mean[c] = mean;
        save_invstd[c] = invstd;
        if (running_mean) {
            const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
            running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
        }
    }
    if (threadIdx.x == 0) counters[blockIdx.y] = 0u;
}

// Statistics from sums the producer already reduced: the convolution epilogue (conv_tc.cu) adds
// the per-channel sum and sum of squares of its output to sums[2][C]; this turns them into
// mean / invstd / scale_shift / running statistics and clears the sums for the next step, so the
// read of the convolution output that bn_fwd_stats_kernel makes never happens.
__global__ void __launch_bounds__(256) bn_finalize_kernel(
    float* __restrict__ sums, size_t M, int C, const void* __restrict__ gamma,
    const void* __restrict__ beta, int param_dtype, float* __restrict__ running_mean,
    float* __restrict__ running_var, float momentum, float eps, float* __restrict__ scale_shift,
    float* __restrict__ save_mean, float* __restrict__ save_invstd) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float S = sums[c], Q = sums[C + c];
    sums[c] = 0.f;
    sums[C + c] = 0.f;
    const float invM = 1.f / (float)M;
    const float mean = S * invM;
    float var = fmaf(-mean, mean, Q * invM);
    var = fmaxf(var, 0.f);
    const float invstd = rsqrtf(var + eps);
    const float g = ld_param(gamma, param_dtype, c);
    const float bt = ld_param(beta, param_dtype, c);
    const float a = g * invstd;
    scale_shift[c] = a;
    scale_shift[C + c] = fmaf(-mean, a, bt);
    save_mean[c] = mean;
    save_invstd[c] = invstd;
    if (running_mean) {
        const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
    }
}

template <bool RELU, bool RES>
__global__ void __launch_bounds__(kT, 4) bn_fwd_apply_kernel(
    const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
    __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ mask,
    const float* __restrict__ scale_shift, size_t M, int C) {
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    const int c0 = blockIdx.y * kGroupC + tx * 8;
    const int CV = C >> 3;
    float a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        a[i] = scale_shift[c0 + i];
        b[i] = scale_shift[C + c0 + i];
    }
    const size_t rstride = (size_t)gridDim.x * 32;
    constexpr int U = RES ? 2 : 4;
    for (size_t r0 = (size_t)blockIdx.x * 32 + ty; r0 < M; r0 += rstride * U) {
        uint4 raw[U], rres[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t r = r0 + (size_t)u * rstride;
            if (r < M) {
                raw[u] = ld_stream_v4(x + r * C + c0);
                if (RES) rres[u] = ld_stream_v4(res + r * C + c0);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t r = r0 + (size_t)u * rstride;
            if (r >= M) break;
            const size_t e = r * C + c0;
            float f[8];
            bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw[u]), f);
            if (RES) {
                float rv[8];
                bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&rres[u]), rv);
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = fmaf(f[i], a[i], b[i]) + rv[i];
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = fmaf(f[i], a[i], b[i]);
            }
            if (RELU) {
                unsigned int m = 0;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    if (f[i] > 0.f)
                        m |= 1u << i;
                    else
                        f[i] = 0.f;
                }
                if (mask) mask[r * CV + (c0 >> 3)] = (uint8_t)m;
            }
            bf16x8 pk = f32_to_bf16x8(f);
            *reinterpret_cast<uint4*>(y + e) = *reinterpret_cast<uint4*>(&pk);
        }
    }
}

template <bool RELU>
__global__ void __launch_bounds__(kT) bn_bwd_reduce_kernel(
    const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
    const uint8_t* __restrict__ mask, const float* __restrict__ save_mean,
    const float* __restrict__ save_invstd, size_t M, int C, float* __restrict__ accum,
    unsigned int* __restrict__ counters, void* __restrict__ dgamma, void* __restrict__ dbeta,
    int param_dtype, float* __restrict__ coef) {
    __shared__ float smem[16 * kT];
    __shared__ int s_flag;
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    const int c_base = blockIdx.y * kGroupC;
    const int c0 = c_base + tx * 8;
    const int CV = C >> 3;
    float mu[8], is[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        mu[i] = save_mean[c0 + i];
        is[i] = save_invstd[c0 + i];
    }
    float sb[8], sg[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) sb[i] = sg[i] = 0.f;
    const size_t rstride = (size_t)gridDim.x * 32;
    for (size_t r = (size_t)blockIdx.x * 32 + ty; r < M; r += rstride) {
        const size_t e = r * C + c0;
        const uint4 rd = ld_nc_v4(dy + e);
        const uint4 rx = ld_nc_v4(x + e);
        float d[8], f[8];
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&rd), d);
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&rx), f);
        unsigned int m = 0xffu;
        if (RELU) m = mask[r * CV + (c0 >> 3)];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float dz = (m >> i) & 1u ? d[i] : 0.f;
            sb[i] += dz;
            sg[i] = fmaf(dz, (f[i] - mu[i]) * is[i], sg[i]);
        }
    }
    block_fold_atomic(sb, sg, C, c_base, accum, smem);
    if (!last_block_ticket(counters + blockIdx.y, &s_flag)) return;
    if (threadIdx.x < kGroupC) {
        const int c = c_base + threadIdx.x;
        const float B = atomicExch(accum + c, 0.f);
        const float G = atomicExch(accum + C + c, 0.f);
        const float invM = 1.f / (float)M;
        st_param(dbeta, param_dtype, c, B);
        st_param(dgamma, param_dtype, c, G);
        coef[c] = B * invM;
        coef[C + c] = G * invM;
    }
    if (threadIdx.x == 0) counters[blockIdx.y] = 0u;
}

template <bool RELU, bool RES>
__global__ void __launch_bounds__(kT) bn_bwd_apply_kernel(
    const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
    const uint8_t* __restrict__ mask, const float* __restrict__ save_mean,
    const float* __restrict__ save_invstd, const void* __restrict__ gamma, int param_dtype,
    const float* __restrict__ coef, __nv_bfloat16* __restrict__ dx,
    __nv_bfloat16* __restrict__ dres, size_t M, int C) {
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    const int c0 = blockIdx.y * kGroupC + tx * 8;
    const int CV = C >> 3;
    float mu[8], is[8], gi[8], c1[8], c2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = c0 + i;
        mu[i] = save_mean[c];
        is[i] = save_invstd[c];
        gi[i] = ld_param(gamma, param_dtype, c) * is[i];
        c1[i] = coef[c];
        c2[i] = coef[C + c];
    }
    const size_t rstride = (size_t)gridDim.x * 32;
    for (size_t r = (size_t)blockIdx.x * 32 + ty; r < M; r += rstride) {
        const size_t e = r * C + c0;
        const uint4 rd = ld_stream_v4(dy + e);
        const uint4 rx = ld_stream_v4(x + e);
        float d[8], f[8], o[8];
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&rd), d);
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&rx), f);
        unsigned int m = 0xffu;
        if (RELU) m = mask[r * CV + (c0 >> 3)];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float dz = (m >> i) & 1u ? d[i] : 0.f;
            d[i] = dz;
            const float xh = (f[i] - mu[i]) * is[i];
            o[i] = gi[i] * (dz - c1[i] - xh * c2[i]);
        }
        bf16x8 pk = f32_to_bf16x8(o);
        *reinterpret_cast<uint4*>(dx + e) = *reinterpret_cast<uint4*>(&pk);
        if (RES) {
            bf16x8 pr = f32_to_bf16x8(d);
            *reinterpret_cast<uint4*>(dres + e) = *reinterpret_cast<uint4*>(&pr);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Stem tail fused: y = maxpool3x3/2/pad1(relu(bn(x))) without materialising the BN output.
// Forward applies the affine + ReLU per tap while pooling (one read of x, one write of the 4x
// smaller pooled tensor + a 1-byte argmax code; code 15 = every tap <= 0, i.e. no gradient).
// Backward never builds the pooled-gradient scatter either: both BN backward passes gather
// dz(n,h,w,c) = sum of dpool over the <= 4 windows whose argmax is this tap, on the fly.
// Replaces bn_fwd_apply + maxpool_fwd (341 us -> ~100 us at batch 256) and maxpool_bwd +
// bn_bwd_reduce + bn_bwd_apply (736 us -> ~260 us) for the 112x112x64 stem activation.
constexpr unsigned int kDeadCode = 15u;

__global__ void __launch_bounds__(kT) stem_bn_relu_pool_fwd_kernel(
    const __nv_bfloat16* __restrict__ x, const float* __restrict__ scale_shift,
    __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ idx, int N, int H, int W, int C, int Ho,
    int Wo) {
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    const int c0 = blockIdx.y * kGroupC + tx * 8;
    float a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        a[i] = scale_shift[c0 + i];
        b[i] = scale_shift[C + c0 + i];
    }
    const unsigned int Mo = (unsigned int)N * Ho * Wo;
    const unsigned int rstride = gridDim.x * 32u;
    for (unsigned int r = blockIdx.x * 32u + ty; r < Mo; r += rstride) {
        const int wo = (int)(r % Wo);
        const unsigned int q = r / Wo;
        const int ho = (int)(q % Ho);
        const int n = (int)(q / Ho);
        float best[8];
        unsigned int code[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            best[i] = 0.f;          // the ReLU floor: only a positive tap can win
            code[i] = kDeadCode;
        }
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int h = 2 * ho - 1 + kh;
            if (h < 0 || h >= H) continue;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int w = 2 * wo - 1 + kw;
                if (w < 0 || w >= W) continue;
                const uint4 raw = ld_nc_v4(x + (((size_t)n * H + h) * W + w) * C + c0);
                float f[8];
                bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), f);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float v = fmaf(f[i], a[i], b[i]);
                    if (v > best[i]) {
                        best[i] = v;
                        code[i] = kh * 3 + kw;
                    }
                }
            }
        }
        const size_t o = (size_t)r * C + c0;
        bf16x8 pk = f32_to_bf16x8(best);
        *reinterpret_cast<uint4*>(y + o) = *reinterpret_cast<uint4*>(&pk);
        uint2 cpk;
        cpk.x = code[0] | (code[1] << 8) | (code[2] << 16) | (code[3] << 24);
        cpk.y = code[4] | (code[5] << 8) | (code[6] << 16) | (code[7] << 24);
        *reinterpret_cast<uint2*>(idx + o) = cpk;
    }
}

// Gradient wrt the (never materialised) BN output at input position (n, h, w), 8 channels.
__device__ __forceinline__ void stem_gather_dz(const __nv_bfloat16* __restrict__ dpool,
                                               const uint8_t* __restrict__ idx, int n, int h,
                                               int w, int C, int Ho, int Wo, int c0, float* dz) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dz[i] = 0.f;
    const int ho_lo = h >> 1, ho_hi = (h + 1) >> 1;
    const int wo_lo = w >> 1, wo_hi = (w + 1) >> 1;
    for (int ho = ho_lo; ho <= ho_hi; ++ho) {
        if (ho >= Ho) continue;
        const int kh = h - (2 * ho - 1);
        for (int wo = wo_lo; wo <= wo_hi; ++wo) {
            if (wo >= Wo) continue;
            const unsigned int me = (unsigned int)(kh * 3 + (w - (2 * wo - 1)));
            const size_t o = (((size_t)n * Ho + ho) * Wo + wo) * C + c0;
            const uint2 cpk = __ldg(reinterpret_cast<const uint2*>(idx + o));
            // byte-wise compare of 4 codes at a time; most windows select some other tap, so
            // skip the gradient load unless one of the 8 channels points here
            const unsigned int me4 = me * 0x01010101u;
            const unsigned int hit_lo = __vcmpeq4(cpk.x, me4), hit_hi = __vcmpeq4(cpk.y, me4);
            if ((hit_lo | hit_hi) == 0u) continue;
            const uint4 raw = ld_nc_v4(dpool + o);
            float g[8];
            bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), g);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const unsigned int hit = ((i < 4 ? hit_lo : hit_hi) >> (8 * (i & 3))) & 1u;
                dz[i] += hit ? g[i] : 0.f;
            }
        }
    }
}

__global__ void __launch_bounds__(kT) stem_bn_pool_bwd_reduce_kernel(
    const __nv_bfloat16* __restrict__ dpool, const uint8_t* __restrict__ idx,
    const __nv_bfloat16* __restrict__ x, const float* __restrict__ save_mean,
    const float* __restrict__ save_invstd, int N, int H, int W, int C, int Ho, int Wo,
    float* __restrict__ accum, unsigned int* __restrict__ counters, void* __restrict__ dgamma,
    void* __restrict__ dbeta, int param_dtype, float* __restrict__ coef) {
    __shared__ float smem[16 * kT];
    __shared__ int s_flag;
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    const int c_base = blockIdx.y * kGroupC;
    const int c0 = c_base + tx * 8;
    float mu[8], is[8], sb[8], sg[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        mu[i] = save_mean[c0 + i];
        is[i] = save_invstd[c0 + i];
        sb[i] = sg[i] = 0.f;
    }
    const unsigned int M = (unsigned int)N * H * W;
    const unsigned int rstride = gridDim.x * 32u;
    for (unsigned int r = blockIdx.x * 32u + ty; r < M; r += rstride) {
        const int w = (int)(r % W);
        const unsigned int q = r / W;
        const int h = (int)(q % H);
        const int n = (int)(q / H);
        const uint4 rx = ld_nc_v4(x + (size_t)r * C + c0);
        float dz[8], f[8];
        stem_gather_dz(dpool, idx, n, h, w, C, Ho, Wo, c0, dz);
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&rx), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            sb[i] += dz[i];
            sg[i] = fmaf(dz[i], (f[i] - mu[i]) * is[i], sg[i]);
        }
    }
    block_fold_atomic(sb, sg, C, c_base, accum, smem);
    if (!last_block_ticket(counters + blockIdx.y, &s_flag)) return;
    if (threadIdx.x < kGroupC) {
        const int c = c_base + threadIdx.x;
        const float B = atomicExch(accum + c, 0.f);
        const float G = atomicExch(accum + C + c, 0.f);
        const float invM = 1.f / (float)M;
        st_param(dbeta, param_dtype, c, B);
        st_param(dgamma, param_dtype, c, G);
        coef[c] = B * invM;
        coef[C + c] = G * invM;
    }
    if (threadIdx.x == 0) counters[blockIdx.y] = 0u;
}

__global__ void __launch_bounds__(kT) stem_bn_pool_bwd_apply_kernel(
    const __nv_bfloat16* __restrict__ dpool, const uint8_t* __restrict__ idx,
    const __nv_bfloat16* __restrict__ x, const float* __restrict__ save_mean,
    const float* __restrict__ save_invstd, const void* __restrict__ gamma, int param_dtype,
    const float* __restrict__ coef, __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C,
    int Ho, int Wo) {
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    const int c0 = blockIdx.y * kGroupC + tx * 8;
    float mu[8], is[8], gi[8], c1[8], c2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = c0 + i;
        mu[i] = save_mean[c];
        is[i] = save_invstd[c];
        gi[i] = ld_param(gamma, param_dtype, c) * is[i];
        c1[i] = coef[c];
        c2[i] = coef[C + c];
    }
    const unsigned int M = (unsigned int)N * H * W;
    const unsigned int rstride = gridDim.x * 32u;
    for (unsigned int r = blockIdx.x * 32u + ty; r < M; r += rstride) {
        const int w = (int)(r % W);
        const unsigned int q = r / W;
        const int h = (int)(q % H);
        const int n = (int)(q / H);
        const size_t e = (size_t)r * C + c0;
        const uint4 rx = ld_stream_v4(x + e);
        float dz[8], f[8], o[8];
        stem_gather_dz(dpool, idx, n, h, w, C, Ho, Wo, c0, dz);
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&rx), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float xh = (f[i] - mu[i]) * is[i];
            o[i] = gi[i] * (dz[i] - c1[i] - xh * c2[i]);
        }
        bf16x8 pk = f32_to_bf16x8(o);
        *reinterpret_cast<uint4*>(dx + e) = *reinterpret_cast<uint4*>(&pk);
    }
}

dim3 grid_for(size_t M, int C) {
    const int ny = C / kGroupC;
    size_t nx = (M + 127) / 128;  // >= 4 rows per thread when there is enough work
    size_t cap = kTargetBlocks / ny;
    if (cap < 1) cap = 1;
    if (nx > cap) nx = cap;
    if (nx < 1) nx = 1;
    return dim3((unsigned)nx, (unsigned)ny);
}

int check_shape(size_t M, int C) {
    TFOP_REQUIRE(M >= 1, "M must be >= 1");
    TFOP_REQUIRE(C >= kGroupC && C <= kMaxC && C % kGroupC == 0,
                 "channel count must be a multiple of 64 in [64, 4096]");
    return 0;
}

}  // namespace

// Workspace: [2][kMaxC] fp32 accumulators, then kMaxGroups u32 ticket counters.
// It must start zeroed; every kernel leaves it zeroed again, so one workspace
// serves every BN layer as long as the calls are ordered on one stream.
TFOP_API size_t tfop_bn_workspace_bytes(void) {
    return (size_t)2 * kMaxC * sizeof(float) + kMaxGroups * sizeof(unsigned int);
}

static inline unsigned int* ws_counter(void* ws) {
    return (unsigned int*)((char*)ws + (size_t)2 * kMaxC * sizeof(float));
}

// Training-mode statistics. Outputs: scale_shift[2*C], save_mean[C], save_invstd[C];
// running_mean/var (fp32, may be NULL) are updated with `momentum`.
TFOP_API int tfop_bn_fwd_stats(const void* x, size_t M, int C, const void* gamma,
                               const void* beta, int param_dtype, float* running_mean,
                               float* running_var, float momentum, float eps,
                               float* scale_shift, float* save_mean, float* save_invstd,
                               void* workspace, cudaStream_t stream) {
    if (check_shape(M, C) != 0) return -1;
    const dim3 nb = grid_for(M, C);
    bn_fwd_stats_kernel<<<nb, kT, 0, stream>>>(
        (const __nv_bfloat16*)x, M, C, (float*)workspace, ws_counter(workspace), gamma, beta,
        param_dtype, running_mean, running_var, momentum, eps, scale_shift, save_mean,
        save_invstd);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// Same outputs as tfop_bn_fwd_stats, from sums[2][C] (fp32: per-channel sum and sum of squares
// over the M rows) that the producing convolution accumulated (tfop_conv_fwd `stats`). Clears
// `sums` so the array can be handed to the next step's convolution as it is.
TFOP_API int tfop_bn_finalize(float* sums, size_t M, int C, const void* gamma, const void* beta,
                              int param_dtype, float* running_mean, float* running_var,
                              float momentum, float eps, float* scale_shift, float* save_mean,
                              float* save_invstd, cudaStream_t stream) {
    TFOP_REQUIRE(sums != nullptr && M > 0 && C > 0, "bn_finalize: bad arguments");
    bn_finalize_kernel<<<(C + 255) / 256, 256, 0, stream>>>(sums, M, C, gamma, beta, param_dtype,
                                                            running_mean, running_var, momentum,
                                                            eps, scale_shift, save_mean,
                                                            save_invstd);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// y = [relu]( x * scale + shift [+ res] ); mask (M*C/8 bytes) may be NULL.
TFOP_API int tfop_bn_fwd_apply(const void* x, const void* res, void* y, void* mask,
                               const float* scale_shift, size_t M, int C, int relu,
                               cudaStream_t stream) {
    if (check_shape(M, C) != 0) return -1;
    const dim3 nb = grid_for(M, C);
    const __nv_bfloat16* xp = (const __nv_bfloat16*)x;
    const __nv_bfloat16* rp = (const __nv_bfloat16*)res;
    __nv_bfloat16* yp = (__nv_bfloat16*)y;
    uint8_t* mp = (uint8_t*)mask;
    if (relu && res)
        bn_fwd_apply_kernel<true, true><<<nb, kT, 0, stream>>>(xp, rp, yp, mp, scale_shift, M, C);
    else if (relu)
        bn_fwd_apply_kernel<true, false><<<nb, kT, 0, stream>>>(xp, rp, yp, mp, scale_shift, M, C);
    else if (res)
        bn_fwd_apply_kernel<false, true><<<nb, kT, 0, stream>>>(xp, rp, yp, mp, scale_shift, M, C);
    else
        bn_fwd_apply_kernel<false, false><<<nb, kT, 0, stream>>>(xp, rp, yp, mp, scale_shift, M, C);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// dgamma/dbeta (param_dtype) and coef[2*C] = {sum(dz)/M, sum(dz*xhat)/M}.
TFOP_API int tfop_bn_bwd_reduce(const void* dy, const void* x, const void* mask,
                                const float* save_mean, const float* save_invstd, size_t M,
                                int C, void* dgamma, void* dbeta, int param_dtype, float* coef,
                                void* workspace, cudaStream_t stream) {
    if (check_shape(M, C) != 0) return -1;
    const dim3 nb = grid_for(M, C);
    if (mask)
        bn_bwd_reduce_kernel<true><<<nb, kT, 0, stream>>>(
            (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const uint8_t*)mask, save_mean,
            save_invstd, M, C, (float*)workspace, ws_counter(workspace), dgamma, dbeta,
            param_dtype, coef);
    else
        bn_bwd_reduce_kernel<false><<<nb, kT, 0, stream>>>(
            (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, nullptr, save_mean, save_invstd, M,
            C, (float*)workspace, ws_counter(workspace), dgamma, dbeta, param_dtype, coef);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// dx = gamma*invstd*(dz - c1 - xhat*c2); dres = dz when requested.
TFOP_API int tfop_bn_bwd_apply(const void* dy, const void* x, const void* mask,
                               const float* save_mean, const float* save_invstd,
                               const void* gamma, int param_dtype, const float* coef, void* dx,
                               void* dres, size_t M, int C, cudaStream_t stream) {
    if (check_shape(M, C) != 0) return -1;
    const dim3 nb = grid_for(M, C);
    const __nv_bfloat16* dyp = (const __nv_bfloat16*)dy;
    const __nv_bfloat16* xp = (const __nv_bfloat16*)x;
    const uint8_t* mp = (const uint8_t*)mask;
    __nv_bfloat16* dxp = (__nv_bfloat16*)dx;
    __nv_bfloat16* drp = (__nv_bfloat16*)dres;
    if (mask && dres)
        bn_bwd_apply_kernel<true, true><<<nb, kT, 0, stream>>>(
            dyp, xp, mp, save_mean, save_invstd, gamma, param_dtype, coef, dxp, drp, M, C);
    else if (mask)
        bn_bwd_apply_kernel<true, false><<<nb, kT, 0, stream>>>(
            dyp, xp, mp, save_mean, save_invstd, gamma, param_dtype, coef, dxp, drp, M, C);
    else if (dres)
        bn_bwd_apply_kernel<false, true><<<nb, kT, 0, stream>>>(
            dyp, xp, mp, save_mean, save_invstd, gamma, param_dtype, coef, dxp, drp, M, C);
    else
        bn_bwd_apply_kernel<false, false><<<nb, kT, 0, stream>>>(
            dyp, xp, mp, save_mean, save_invstd, gamma, param_dtype, coef, dxp, drp, M, C);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// Allocation-free variant of tfop_bn_fwd_stats.
//
// Who needs it: a caller that must not allocate device memory when it runs BatchNorm
// statistics. Two cases we have met: (a) a host application that reserves the whole GPU
// for its own pool allocator at start-up, so a later cudaMalloc from a plug-in either
// fails or eats into memory the pool believes it owns; (b) a latency-critical loop where
// cudaMalloc / cudaFree are off limits because they synchronise the device. Such a caller
// cannot create the workspace tfop_bn_fwd_stats() asks for, so this entry point brings
// its own: the cross-block accumulators and ticket counters live in the file-scope
// __device__ array below, which every GPU that loads this library carries until exit.
//
// Limit: kNoAllocMaxC = 4096 channels (2*4096 floats + 64 counters = 33,024 bytes). A
// larger C is rejected with an error; nothing is truncated.
//
// Sharing: there is ONE scratch array per device and every call uses it, keeping whatever
// the previous call left. The accumulators and counters must start at zero, so each call
// clears them on `stream` before the kernel. Two calls that overlap on the GPU (different
// streams, or different host threads on different streams) race on the array: serialise
// them on one stream.
//
// Set-up: in host code the array's name is only a host-side stand-in, not a device
// address, so the first call on each device looks the real address up with
// cudaGetSymbolAddress (that device being current) and caches it. Make that first call
// before starting any stream capture.
// ---------------------------------------------------------------------------
namespace {

constexpr int kNoAllocMaxC = kMaxC;
constexpr size_t kNoAllocScratchBytes =
    (size_t)2 * kNoAllocMaxC * sizeof(float) + kMaxGroups * sizeof(unsigned int);

static __device__ __align__(256) unsigned char g_bn_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
void* g_noalloc_ptr[kMaxDevices] = {nullptr};

// Device address of the scratch array on the current device (looked up once per device).
int noalloc_scratch(void** out) {
    int dev = -1;
    TFOP_CUDA(cudaGetDevice(&dev));
    TFOP_REQUIRE(dev >= 0 && dev < kMaxDevices, "device index out of range for the scratch table");
    if (g_noalloc_ptr[dev] == nullptr) {
        void* p = nullptr;
        TFOP_CUDA(cudaGetSymbolAddress(&p, g_bn_noalloc_scratch));
        TFOP_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}

}  // namespace

// Same contract and outputs as tfop_bn_fwd_stats, minus the workspace argument; performs
// no cudaMalloc / cudaFree and calls no allocating library routine. C must be a multiple
// of 64 and at most 4096. See the block comment above for sharing and first-call rules.
TFOP_API int tfop_bn_fwd_stats_noalloc(const void* x, size_t M, int C, const void* gamma,
                                       const void* beta, int param_dtype, float* running_mean,
                                       float* running_var, float momentum, float eps,
                                       float* scale_shift, float* save_mean, float* save_invstd,
                                       cudaStream_t stream) {
    TFOP_REQUIRE(C <= kNoAllocMaxC,
                 "tfop_bn_fwd_stats_noalloc: C exceeds the built-in scratch (max 4096 channels)");
    if (check_shape(M, C) != 0) return -1;
    void* scratch = nullptr;
    if (noalloc_scratch(&scratch) != 0) return -1;
    // the array is shared and keeps the last call's contents: start from zero every time
    TFOP_CUDA(cudaMemsetAsync(scratch, 0, kNoAllocScratchBytes, stream));
    const dim3 nb = grid_for(M, C);
    bn_fwd_stats_kernel<<<nb, kT, 0, stream>>>(
        (const __nv_bfloat16*)x, M, C, (float*)scratch, ws_counter(scratch), gamma, beta,
        param_dtype, running_mean, running_var, momentum, eps, scale_shift, save_mean,
        save_invstd);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// ---- stem tail: BN + ReLU + 3x3/2 max-pool fused (kernels above) ---------------------------
// x: [N,H,W,C] bf16 NHWC conv output; scale_shift from tfop_bn_fwd_stats on the same x.
// y: [N,Ho,Wo,C] bf16, idx: uint8 [N,Ho,Wo,C] argmax codes; Ho = (H+1)/2, Wo = (W+1)/2.
TFOP_API int tfop_stem_bn_relu_pool_fwd(const void* x, const float* scale_shift, void* y,
                                        void* idx, int N, int H, int W, int C,
                                        cudaStream_t stream) {
    TFOP_REQUIRE(N > 0 && H > 0 && W > 0, "empty tensor");
    TFOP_REQUIRE((size_t)N * H * W < 0x7fffffffull, "tensor too large for 32-bit row indexing");
    const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
    const size_t Mo = (size_t)N * Ho * Wo;
    if (check_shape(Mo, C) != 0) return -1;
    stem_bn_relu_pool_fwd_kernel<<<grid_for(Mo, C), kT, 0, stream>>>(
        (const __nv_bfloat16*)x, scale_shift, (__nv_bfloat16*)y, (uint8_t*)idx, N, H, W, C, Ho, Wo);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// dgamma / dbeta / coef of the stem BN from the pooled gradient dpool [N,Ho,Wo,C] (gathered on
// the fly through idx); workspace as for tfop_bn_bwd_reduce.
TFOP_API int tfop_stem_bn_pool_bwd_reduce(const void* dpool, const void* idx, const void* x,
                                          const float* save_mean, const float* save_invstd, int N,
                                          int H, int W, int C, void* dgamma, void* dbeta,
                                          int param_dtype, float* coef, void* workspace,
                                          cudaStream_t stream) {
    TFOP_REQUIRE((size_t)N * H * W < 0x7fffffffull, "tensor too large for 32-bit row indexing");
    const size_t M = (size_t)N * H * W;
    if (check_shape(M, C) != 0) return -1;
    stem_bn_pool_bwd_reduce_kernel<<<grid_for(M, C), kT, 0, stream>>>(
        (const __nv_bfloat16*)dpool, (const uint8_t*)idx, (const __nv_bfloat16*)x, save_mean,
        save_invstd, N, H, W, C, (H + 1) / 2, (W + 1) / 2, (float*)workspace,
        ws_counter(workspace), dgamma, dbeta, param_dtype, coef);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// dx [N,H,W,C] = gradient wrt the conv output x.
TFOP_API int tfop_stem_bn_pool_bwd_apply(const void* dpool, const void* idx, const void* x,
                                         const float* save_mean, const float* save_invstd,
                                         const void* gamma, int param_dtype, const float* coef,
                                         void* dx, int N, int H, int W, int C,
                                         cudaStream_t stream) {
    TFOP_REQUIRE((size_t)N * H * W < 0x7fffffffull, "tensor too large for 32-bit row indexing");
    const size_t M = (size_t)N * H * W;
    if (check_shape(M, C) != 0) return -1;
    stem_bn_pool_bwd_apply_kernel<<<grid_for(M, C), kT, 0, stream>>>(
        (const __nv_bfloat16*)dpool, (const uint8_t*)idx, (const __nv_bfloat16*)x, save_mean,
        save_invstd, gamma, param_dtype, coef, (__nv_bfloat16*)dx, N, H, W, C, (H + 1) / 2,
        (W + 1) / 2);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
