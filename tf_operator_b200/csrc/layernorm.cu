// K9: fused (bias + residual + LayerNorm) forward / backward and (bias + GeLU) forward /
// backward for the transformer block (BASELINE.json config 3, BERT-large data-parallel).
//
// What is fused, and why: the GEMM output h of the attention projection / second MLP layer
// goes through  s = h + bias + residual;  y = LN(s).  As separate torch ops that is an add
// kernel, a LayerNorm kernel, and in backward a grad-input kernel, a gamma/beta kernel that ran
// at 0.8 TB/s, a residual-grad add and a bias-grad column reduction - six passes over the
// [tokens, hidden] tensor. Here forward is one pass and backward is one pass that also
// produces the bias gradient (the column sum of dx) and dgamma / dbeta.
//
// Layout: one warp per row, the row lives in registers (hidden <= 1536, multiple of 8), each
// lane owns 16-byte vectors lane, lane+32, ...; statistics are fp32 and computed on the
// bf16-rounded sum so that forward and backward see the same s.
//
// Column reductions (dgamma, dbeta, dbias): each warp walks rows with a grid stride and keeps
// per-column partials in registers, the block folds its warps through shared memory, then one
// fp32 RED per column and block into pre-zeroed outputs.
//
// The reference has no model code; the op list is defined by BASELINE.json config 3.
#include "common.cuh"

namespace {

constexpr int kWarps = 8;  // per block
constexpr float kInvSqrt2 = 0.70710678118654752f;
constexpr float kInvSqrt2Pi = 0.3989422804014327f;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float* f) {
    const uint4 raw = __ldg(reinterpret_cast<const uint4*>(p));
    bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), f);
}
__device__ __forceinline__ void load8_stream(const __nv_bfloat16* p, float* f) {
    const uint4 raw = __ldcs(reinterpret_cast<const uint4*>(p));
    bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), f);
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float* f) {
    bf16x8 pk = f32_to_bf16x8(f);
    *reinterpret_cast<uint4*>(p) = *reinterpret_cast<uint4*>(&pk);
}
__device__ __forceinline__ float round_bf16(float v) {
    return __bfloat162float(__float2bfloat16_rn(v));
}

// GeLU(t) = t * Phi(t) and its derivative Phi(t) + t * phi(t), with erf from Abramowitz-Stegun
// 7.1.26 (|error| <= 1.5e-7, far below bf16 resolution): one rcp, one ex2 and a degree-5 Horner
// instead of erff's ~25 instructions - the GeLU passes were instruction-bound (ncu: 33 / 47
// warp instructions per element, SM throughput 65-69 %). exp(-u^2) with u = t / sqrt(2) is
// exp(-t^2 / 2), i.e. the Gaussian pdf up to its constant, so backward gets phi(t) for free.
__device__ __forceinline__ void gelu_parts(float t, float& cdf, float& gauss) {
    const float u = fabsf(t) * kInvSqrt2;
    const float k = __fdividef(1.f, fmaf(0.3275911f, u, 1.f));
    float p = fmaf(k, 1.061405429f, -1.453152027f);
    p = fmaf(k, p, 1.421413741f);
    p = fmaf(k, p, -0.284496736f);
    p = fmaf(k, p, 0.254829592f);
    gauss = __expf(-u * u);
    const float erf_abs = fmaf(-p * k, gauss, 1.f);
    cdf = 0.5f * (1.f + copysignf(erf_abs, t));
}

// y = LN(h + bias + res) * gamma + beta; s (the bf16 LN input), mean, rstd saved for backward.
template <int NV>
__global__ void __launch_bounds__(kWarps * 32)
ln_fwd_kernel(const __nv_bfloat16* __restrict__ h, const __nv_bfloat16* __restrict__ bias,
              const __nv_bfloat16* __restrict__ res, const __nv_bfloat16* __restrict__ gamma,
              const __nv_bfloat16* __restrict__ beta, __nv_bfloat16* __restrict__ y,
              __nv_bfloat16* __restrict__ s_out, float* __restrict__ mean_out,
              float* __restrict__ rstd_out, int rows, int H, float eps) {
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int nvec = H >> 3;
    const size_t base = (size_t)row * H;
    float v[NV][8];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const int vec = lane + 32 * i;
        if (vec < nvec) {
            load8_stream(h + base + vec * 8, v[i]);
            float t[8];
            if (bias) {
                load8(bias + vec * 8, t);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[i][j] += t[j];
            }
            if (res) {
                load8_stream(res + base + vec * 8, t);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[i][j] += t[j];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                v[i][j] = round_bf16(v[i][j]);
                sum += v[i][j];
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[i][j] = 0.f;
        }
    }
    const float mean = warp_sum(sum) / (float)H;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        if (lane + 32 * i < nvec) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float d = v[i][j] - mean;
                sq += d * d;
            }
        }
    }
    const float rstd = rsqrtf(warp_sum(sq) / (float)H + eps);
    if (lane == 0) {
        mean_out[row] = mean;
        rstd_out[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const int vec = lane + 32 * i;
        if (vec < nvec) {
            if (s_out) store8(s_out + base + vec * 8, v[i]);
            float g[8], b[8], o[8];
            load8(gamma + vec * 8, g);
            load8(beta + vec * 8, b);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * g[j] + b[j];
            store8(y + base + vec * 8, o);
        }
    }
}

// dx = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * gamma,  xhat = (s - mean) * rstd
// dgamma += dy * xhat, dbeta += dy, dbias += dx   (column sums, fp32 RED into zeroed buffers)
template <int NV>
__global__ void __launch_bounds__(kWarps * 32)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ s,
              const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
              const __nv_bfloat16* __restrict__ gamma, __nv_bfloat16* __restrict__ dx,
              float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dbias,
              int rows, int H) {
    __shared__ float red[3][kWarps][32 * 8 + 8];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nvec = H >> 3;
    const float inv_h = 1.f / (float)H;
    float gam[NV][8], ag[NV][8], ab[NV][8], ax[NV][8];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const int vec = lane + 32 * i;
        if (vec < nvec) load8(gamma + vec * 8, gam[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) ag[i][j] = ab[i][j] = ax[i][j] = 0.f;
    }
    for (int row = blockIdx.x * kWarps + warp; row < rows; row += gridDim.x * kWarps) {
        const size_t base = (size_t)row * H;
        const float mean = mean_in[row], rstd = rstd_in[row];
        float d[NV][8], xh[NV][8];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int vec = lane + 32 * i;
            if (vec < nvec) {
                load8_stream(dy + base + vec * 8, d[i]);
                load8_stream(s + base + vec * 8, xh[i]);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    xh[i][j] = (xh[i][j] - mean) * rstd;
                    ag[i][j] += d[i][j] * xh[i][j];
                    ab[i][j] += d[i][j];
                    d[i][j] *= gam[i][j];
                    c1 += d[i][j];
                    c2 += d[i][j] * xh[i][j];
                }
            }
        }
        c1 = warp_sum(c1) * inv_h;
        c2 = warp_sum(c2) * inv_h;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int vec = lane + 32 * i;
            if (vec < nvec) {
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    o[j] = round_bf16(rstd * (d[i][j] - c1 - xh[i][j] * c2));
                    ax[i][j] += o[j];
                }
                store8(dx + base + vec * 8, o);
            }
        }
    }
    // fold the block's warps, one 256-column slab (i) at a time, all three quantities per round
    // (the fold costs as much as a few rows: one resident wave of blocks, few rounds)
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            red[0][warp][lane * 8 + j] = ag[i][j];
            red[1][warp][lane * 8 + j] = ab[i][j];
            red[2][warp][lane * 8 + j] = ax[i][j];
        }
        __syncthreads();
        const int col = i * 256 + threadIdx.x;  // 256 threads <-> 256 columns
        if (col < H) {
            float t0 = 0.f, t1 = 0.f, t2 = 0.f;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                t0 += red[0][w][threadIdx.x];
                t1 += red[1][w][threadIdx.x];
                t2 += red[2][w][threadIdx.x];
            }
            atomicAdd(dgamma + col, t0);
            atomicAdd(dbeta + col, t1);
            if (dbias) atomicAdd(dbias + col, t2);
        }
    }
}

// Tiled LayerNorm backward for H % 256 == 0 (the transformer sizes). A row is spread over
// H/8 threads (each owns one 16-byte column vector), a 256-thread block holds 256/(H/8) row
// lanes, and every thread processes R rows per trip - so 2R 16-byte loads are in flight per
// thread and only 24 column partials live in registers (the warp-per-row kernel above keeps the
// whole row plus 3 row-length accumulators: 224 registers, 8 warps per SM, 1.1 TB/s). Row sums
// go warp shuffle -> shared memory -> every thread of the row lane; two barriers per trip.
template <int R>
__global__ void __launch_bounds__(256, R <= 2 ? 3 : 2)
ln_bwd_tiled_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ s,
                    const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                    const __nv_bfloat16* __restrict__ gamma, __nv_bfloat16* __restrict__ dx,
                    float* __restrict__ dgamma, float* __restrict__ dbeta,
                    float* __restrict__ dbias, int rows, int H) {
    __shared__ float part[2][8][2 * R];   // [parity][warp][c1 / c2 of each of the R rows]
    __shared__ float fold[3][256 * 8];
    const int vpr = H >> 3;               // threads per row, multiple of 32
    const int lanes = 256 / vpr;          // row lanes per block
    const int tid = threadIdx.x;
    const int cv = tid % vpr;
    const int rl = tid / vpr;
    const bool active = rl < lanes;
    const int warp = tid >> 5, lane = tid & 31;
    const int wpr = vpr >> 5;             // warps per row lane
    const int w0 = rl * wpr;              // first warp of this row lane
    const float inv_h = 1.f / (float)H;
    float gam[8], ag[8], ab[8], ax[8];
    load8(gamma + cv * 8, gam);
#pragma unroll
    for (int j = 0; j < 8; ++j) ag[j] = ab[j] = ax[j] = 0.f;
    const int rows_per_trip = lanes * R;
    int parity = 0;
    for (int row0 = blockIdx.x * rows_per_trip; row0 < rows; row0 += gridDim.x * rows_per_trip) {
        float d[R][8], xh[R][8], rs[R];
        float p[2 * R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int row = row0 + rl * R + r;
            const bool ok = active && row < rows;
            if (ok) {
                const size_t off = (size_t)row * H + cv * 8;
                load8_stream(dy + off, d[r]);
                load8_stream(s + off, xh[r]);
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) d[r][j] = xh[r][j] = 0.f;
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int row = row0 + rl * R + r;
            const bool ok = active && row < rows;
            const float mean = ok ? mean_in[row] : 0.f;
            rs[r] = ok ? rstd_in[row] : 0.f;
            float c1 = 0.f, c2 = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                xh[r][j] = (xh[r][j] - mean) * rs[r];
                ag[j] += d[r][j] * xh[r][j];
                ab[j] += d[r][j];
                d[r][j] *= gam[j];
                c1 += d[r][j];
                c2 += d[r][j] * xh[r][j];
            }
            p[2 * r] = c1;
            p[2 * r + 1] = c2;
        }
#pragma unroll
        for (int i = 0; i < 2 * R; ++i) p[i] = warp_sum(p[i]);
        if (lane == 0) {
#pragma unroll
            for (int i = 0; i < 2 * R; ++i) part[parity][warp][i] = p[i];
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 2 * R; ++i) {
            float t = 0.f;
            if (active)
                for (int w = 0; w < wpr; ++w) t += part[parity][w0 + w][i];
            p[i] = t * inv_h;
        }
        parity ^= 1;   // the next trip writes the other buffer: one barrier per trip is enough
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int row = row0 + rl * R + r;
            if (active && row < rows) {
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    o[j] = round_bf16(rs[r] * (d[r][j] - p[2 * r] - xh[r][j] * p[2 * r + 1]));
                    ax[j] += o[j];
                }
                store8(dx + (size_t)row * H + cv * 8, o);
            }
        }
    }
    // fold the row lanes of this block, then one RED per column and quantity
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        fold[0][tid * 8 + j] = active ? ag[j] : 0.f;
        fold[1][tid * 8 + j] = active ? ab[j] : 0.f;
        fold[2][tid * 8 + j] = active ? ax[j] : 0.f;
    }
    __syncthreads();
    for (int col = tid; col < H; col += 256) {
        float t0 = 0.f, t1 = 0.f, t2 = 0.f;
        for (int l = 0; l < lanes; ++l) {
            t0 += fold[0][l * H + col];
            t1 += fold[1][l * H + col];
            t2 += fold[2][l * H + col];
        }
        atomicAdd(dgamma + col, t0);
        atomicAdd(dbeta + col, t1);
        if (dbias) atomicAdd(dbias + col, t2);
    }
}

// y = gelu(h + bias) (erf form); nothing saved - backward recomputes from h and bias.
// Block = 32 column-vectors x 8 row lanes walking a row band: the bias slice is loaded once
// per thread and no per-element index arithmetic is left (a 64-bit modulo per vector made the
// first version of this kernel instruction-bound at 2.9 TB/s).
__global__ void __launch_bounds__(256)
bias_gelu_fwd_kernel(const __nv_bfloat16* __restrict__ h, const __nv_bfloat16* __restrict__ bias,
                     __nv_bfloat16* __restrict__ y, int rows, int H, int rows_per_block) {
    const int lx = threadIdx.x & 31;
    const int ly = threadIdx.x >> 5;
    const int vec = blockIdx.x * 32 + lx;
    if (vec >= (H >> 3)) return;
    const int r0 = blockIdx.y * rows_per_block;
    const int r1 = min(rows, r0 + rows_per_block);
    float b[8];
    load8(bias + vec * 8, b);
#pragma unroll 2
    for (int r = r0 + ly; r < r1; r += 8) {
        const size_t off = (size_t)r * H + vec * 8;
        float x[8], o[8];
        load8_stream(h + off, x);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float t = round_bf16(x[j] + b[j]);
            float cdf, gauss;
            gelu_parts(t, cdf, gauss);
            o[j] = t * cdf;
        }
        store8(y + off, o);
    }
}

// dh = dy * gelu'(h + bias), dbias += colsum(dh). act = 0: dh = dy (pure column sum, dh may be
// null). Block = 32 column-vectors x 8 row lanes; each block walks a row band.
__global__ void __launch_bounds__(256)
bias_act_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ h,
                    const __nv_bfloat16* __restrict__ bias, __nv_bfloat16* __restrict__ dh,
                    float* __restrict__ dbias, int rows, int H, int act, int rows_per_block) {
    __shared__ float red[8][32 * 8 + 8];
    const int lx = threadIdx.x & 31;   // column vector within the 256-column slab
    const int ly = threadIdx.x >> 5;   // row lane
    const int vec = blockIdx.x * 32 + lx;
    const int nvec = H >> 3;
    const int r0 = blockIdx.y * rows_per_block;
    const int r1 = min(rows, r0 + rows_per_block);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    if (vec < nvec) {
        float b[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) b[j] = 0.f;
        if (act && bias) load8(bias + vec * 8, b);
        if (act) {
#pragma unroll 2
            for (int r = r0 + ly; r < r1; r += 8) {
                const size_t off = (size_t)r * H + vec * 8;
                float d[8], x[8], o[8];
                load8_stream(dy + off, d);
                load8_stream(h + off, x);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float t = round_bf16(x[j] + b[j]);
                    float cdf, gauss;
                    gelu_parts(t, cdf, gauss);
                    o[j] = round_bf16(d[j] * fmaf(t * kInvSqrt2Pi, gauss, cdf));
                    acc[j] += o[j];
                }
                store8(dh + off, o);
            }
        } else {
            // pure column sum: four independent 16-byte loads in flight per thread
            int r = r0 + ly;
            for (; r + 24 < r1; r += 32) {
                float d[4][8];
#pragma unroll
                for (int u = 0; u < 4; ++u) load8_stream(dy + (size_t)(r + 8 * u) * H + vec * 8, d[u]);
#pragma unroll
                for (int u = 0; u < 4; ++u)
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] += d[u][j];
            }
            for (; r < r1; r += 8) {
                float d[8];
                load8_stream(dy + (size_t)r * H + vec * 8, d);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += d[j];
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[ly][lx * 8 + j] = acc[j];
    __syncthreads();
    const int col = blockIdx.x * 256 + threadIdx.x;
    if (col < H) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
        atomicAdd(dbias + col, t);
    }
}

template <int NV>
int launch_fwd(const void* h, const void* bias, const void* res, const void* gamma,
               const void* beta, void* y, void* s_out, float* mean, float* rstd, int rows, int H,
               float eps, cudaStream_t stream) {
    const int grid = (rows + kWarps - 1) / kWarps;
    ln_fwd_kernel<NV><<<grid, kWarps * 32, 0, stream>>>(
        (const __nv_bfloat16*)h, (const __nv_bfloat16*)bias, (const __nv_bfloat16*)res,
        (const __nv_bfloat16*)gamma, (const __nv_bfloat16*)beta, (__nv_bfloat16*)y,
        (__nv_bfloat16*)s_out, mean, rstd, rows, H, eps);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

template <int NV>
int launch_bwd(const void* dy, const void* s, const float* mean, const float* rstd,
               const void* gamma, void* dx, float* dgamma, float* dbeta, float* dbias, int rows,
               int H, cudaStream_t stream) {
    int grid = (rows + kWarps - 1) / kWarps;
    if (grid > 148) grid = 148;  // 227 registers: one block per SM, keep it to one wave
    ln_bwd_kernel<NV><<<grid, kWarps * 32, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)s, mean, rstd, (const __nv_bfloat16*)gamma,
        (__nv_bfloat16*)dx, dgamma, dbeta, dbias, rows, H);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace

static int g_ln_bwd_rows = 2;
// Benchmark hook: rows per thread and trip of the tiled LayerNorm backward (2 or 4).
TFOP_API int tfop_ln_set_bwd_rows(int r) {
    TFOP_REQUIRE(r == 2 || r == 4, "rows per trip: 2 or 4");
    g_ln_bwd_rows = r;
    return 0;
}
static int g_ln_force_warp_rows = 0;
// Test hook: 1 routes every LayerNorm backward through the warp-per-row kernel.
TFOP_API int tfop_ln_force_warp_rows(int on) {
    g_ln_force_warp_rows = on;
    return 0;
}

#define TFOP_LN_DISPATCH(FN, ...)                        \
    switch ((H + 255) / 256) {                           \
        case 1: return FN<1>(__VA_ARGS__);               \
        case 2: return FN<2>(__VA_ARGS__);               \
        case 3: return FN<3>(__VA_ARGS__);               \
        case 4: return FN<4>(__VA_ARGS__);               \
        case 5: return FN<5>(__VA_ARGS__);               \
        default: return FN<6>(__VA_ARGS__);              \
    }

// y[rows,H] = LayerNorm(h + bias + res) * gamma + beta, all bf16; bias / res / s_out may be
// null. mean / rstd: fp32 [rows]. H: multiple of 8, <= 1536 (the row and its three column
// partials stay in registers: 227 at H = 1024, no spills); tensors 16-byte aligned.
TFOP_API int tfop_ln_fwd(const void* h, const void* bias, const void* res, const void* gamma,
                         const void* beta, void* y, void* s_out, float* mean, float* rstd,
                         int rows, int H, float eps, cudaStream_t stream) {
    TFOP_REQUIRE(rows > 0 && H > 0 && H % 8 == 0 && H <= 1536,
                 "layernorm: hidden must be a multiple of 8 and <= 1536");
    TFOP_LN_DISPATCH(launch_fwd, h, bias, res, gamma, beta, y, s_out, mean, rstd, rows, H, eps, stream);
}

// Backward of tfop_ln_fwd. dgamma / dbeta / dbias are fp32 [H] that the caller zeroed (they are
// accumulated with RED); dbias may be null.
TFOP_API int tfop_ln_bwd(const void* dy, const void* s, const float* mean, const float* rstd,
                         const void* gamma, void* dx, float* dgamma, float* dbeta, float* dbias,
                         int rows, int H, cudaStream_t stream) {
    TFOP_REQUIRE(rows > 0 && H > 0 && H % 8 == 0 && H <= 1536,
                 "layernorm: hidden must be a multiple of 8 and <= 1536");
    TFOP_REQUIRE(dgamma != nullptr && dbeta != nullptr, "layernorm bwd needs dgamma and dbeta");
    if (H % 256 == 0 && !g_ln_force_warp_rows) {
        // R rows per thread and trip. ncu on R = 4 (128 registers, 2 blocks = 16 warps per SM):
        // 1.2 TB/s, long-scoreboard stalls dominate, issue slots 38 % busy - latency-bound for
        // lack of warps. R = 2 fits three blocks per SM (tfop_ln_set_bwd_rows switches, for A/B).
        const int R = g_ln_bwd_rows;
        const int rows_per_trip = (256 / (H / 8)) * R;
        int grid = (rows + rows_per_trip - 1) / rows_per_trip;
        const int per_sm = R <= 2 ? 3 : 2;
        if (grid > 148 * per_sm) grid = 148 * per_sm;
        if (R <= 2)
            ln_bwd_tiled_kernel<2><<<grid, 256, 0, stream>>>(
                (const __nv_bfloat16*)dy, (const __nv_bfloat16*)s, mean, rstd,
                (const __nv_bfloat16*)gamma, (__nv_bfloat16*)dx, dgamma, dbeta, dbias, rows, H);
        else
            ln_bwd_tiled_kernel<4><<<grid, 256, 0, stream>>>(
                (const __nv_bfloat16*)dy, (const __nv_bfloat16*)s, mean, rstd,
                (const __nv_bfloat16*)gamma, (__nv_bfloat16*)dx, dgamma, dbeta, dbias, rows, H);
        TFOP_CUDA(cudaGetLastError());
        return 0;
    }
    TFOP_LN_DISPATCH(launch_bwd, dy, s, mean, rstd, gamma, dx, dgamma, dbeta, dbias, rows, H, stream);
}

// y = gelu(h + bias), bf16 [rows, H], H % 8 == 0.
TFOP_API int tfop_bias_gelu_fwd(const void* h, const void* bias, void* y, int rows, int H,
                                cudaStream_t stream) {
    TFOP_REQUIRE(rows > 0 && H > 0 && H % 8 == 0, "bias_gelu: H must be a multiple of 8");
    const int gx = (H / 8 + 31) / 32;
    int gy = (148 * 16 + gx - 1) / gx;
    int rpb = (rows + gy - 1) / gy;
    if (rpb < 32) rpb = 32;
    gy = (rows + rpb - 1) / rpb;
    bias_gelu_fwd_kernel<<<dim3(gx, gy), 256, 0, stream>>>(
        (const __nv_bfloat16*)h, (const __nv_bfloat16*)bias, (__nv_bfloat16*)y, rows, H, rpb);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// act = 1: dh = dy * gelu'(h + bias) and dbias += colsum(dh); act = 0: dbias += colsum(dy)
// (h, bias, dh unused). dbias: fp32 [H], zeroed by the caller.
TFOP_API int tfop_bias_act_bwd(const void* dy, const void* h, const void* bias, void* dh,
                               float* dbias, int rows, int H, int act, cudaStream_t stream) {
    TFOP_REQUIRE(rows > 0 && H > 0 && H % 8 == 0, "bias_act_bwd: H must be a multiple of 8");
    TFOP_REQUIRE(dbias != nullptr, "bias_act_bwd needs dbias");
    TFOP_REQUIRE(act == 0 || (h != nullptr && dh != nullptr), "bias_act_bwd: act needs h and dh");
    const int gx = (H / 8 + 31) / 32;
    // enough row bands to fill the machine a few times over, at least 64 rows each
    int gy = (148 * 8 + gx - 1) / gx;
    int rpb = (rows + gy - 1) / gy;
    if (rpb < 64) rpb = 64;
    gy = (rows + rpb - 1) / rpb;
    bias_act_bwd_kernel<<<dim3(gx, gy), 256, 0, stream>>>(
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)h, (const __nv_bfloat16*)bias,
        (__nv_bfloat16*)dh, dbias, rows, H, act, rpb);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
