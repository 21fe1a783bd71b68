// Gradient-histogram construction for histogram-based gradient boosting (the per-worker
// compute step in front of the K6 histogram all-reduce).
//
//   hist[f][bin] += (grad[r], hess[r])   for every row r of the node, every feature f
//
// bins: uint8 [rows, F] row-major (256 bins). A block owns 8 consecutive features (one 8-byte
// load per row) and a slice of the node's rows; it accumulates in a 16 KB shared-memory
// histogram (shared atomics, spread over 2048 addresses) and flushes once with global REDs.
//
// Reference parity: the histogram build inside XGBoost's `hist` tree method that the
// reference's XGBoostJob example runs out of tree (examples/xgboost/xgboostjob.yaml:19-25);
// shapes from BASELINE.json config 5 (256 bins x 500 features x {grad, hess}).
#include "common.cuh"

namespace {

constexpr int kFT = 8;        // features per block
constexpr int kBins = 256;

__global__ void __launch_bounds__(256) hist_build_kernel(
    const uint8_t* __restrict__ bins, const float* __restrict__ grad,
    const float* __restrict__ hess, const int* __restrict__ row_idx, size_t n_rows, int F,
    float* __restrict__ out /* [F][256][2] */) {
    __shared__ float sh[kFT * kBins * 2];
    for (int i = threadIdx.x; i < kFT * kBins * 2; i += blockDim.x) sh[i] = 0.f;
    __syncthreads();
    const int f0 = blockIdx.y * kFT;
    const int nf = min(kFT, F - f0);
    const bool vec_ok = (nf == kFT) && (F % 8 == 0);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_rows; i += stride) {
        const size_t r = row_idx ? (size_t)row_idx[i] : i;
        const float g = grad[r], h = hess[r];
        const uint8_t* p = bins + r * (size_t)F + f0;
        if (vec_ok) {
            const uint2 raw = *reinterpret_cast<const uint2*>(p);
#pragma unroll
            for (int f = 0; f < kFT; ++f) {
                const unsigned b = ((f < 4 ? raw.x : raw.y) >> (8 * (f & 3))) & 0xffu;
                atomicAdd(&sh[(f * kBins + b) * 2], g);
                atomicAdd(&sh[(f * kBins + b) * 2 + 1], h);
            }
        } else {
            for (int f = 0; f < nf; ++f) {
                const unsigned b = p[f];
                atomicAdd(&sh[(f * kBins + b) * 2], g);
                atomicAdd(&sh[(f * kBins + b) * 2 + 1], h);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nf * kBins * 2; i += blockDim.x) {
        const float v = sh[i];
        if (v != 0.f) atomicAdd(out + (size_t)f0 * kBins * 2 + i, v);
    }
}

}  // namespace

// out must be zeroed by the caller (tfop_zero); row_idx may be NULL (= all rows 0..n_rows).
TFOP_API int tfop_hist_build(const void* bins_u8, const float* grad, const float* hess,
                             const int* row_idx, size_t n_rows, int F, float* out,
                             cudaStream_t stream) {
    TFOP_REQUIRE(F >= 1, "need at least one feature");
    if (n_rows == 0) return 0;
    const int fy = (F + kFT - 1) / kFT;
    size_t bx = (n_rows + 256 * 16 - 1) / (256 * 16);
    const size_t cap = (size_t)(148 * 8 + fy - 1) / fy;
    if (bx > cap) bx = cap;
    if (bx < 1) bx = 1;
    dim3 grid((unsigned)bx, (unsigned)fy);
    hist_build_kernel<<<grid, 256, 0, stream>>>((const uint8_t*)bins_u8, grad, hess, row_idx,
                                                n_rows, F, out);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
