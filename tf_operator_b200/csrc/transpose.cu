// bf16 matrix transpose, out[C, R] = in[R, C]^T, tiled through shared memory with 16-byte
// global accesses on both sides. The TN tcgen05 GEMM needs K-major operands; the weight-gradient
// GEMM dW = dY^T X therefore takes dY^T and X^T, and torch's generic strided copy (measured:
// 32 % of a BERT-large step) is replaced by this kernel.
//
// No counterpart in the reference (it ships no kernels).
#include "common.cuh"

namespace {

constexpr int TILE = 64;
constexpr int PITCH = TILE + 2;   // elements; 33 words per row -> conflict-free column reads

__global__ void __launch_bounds__(256) transpose_bf16_kernel(const __nv_bfloat16* __restrict__ in,
                                                             __nv_bfloat16* __restrict__ out,
                                                             int R, int C, int ld_in, int ld_out) {
    __shared__ __nv_bfloat16 tile[TILE][PITCH];
    const int r0 = blockIdx.y * TILE, c0 = blockIdx.x * TILE;
    const int t = threadIdx.x;
    const bool vec_in = (ld_in % 8 == 0) && ((((uintptr_t)in) & 15) == 0);
    const bool vec_out = (ld_out % 8 == 0) && ((((uintptr_t)out) & 15) == 0);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int r = (t >> 3) + 32 * i, c8 = (t & 7) * 8;
        const int gr = r0 + r, gc = c0 + c8;
        __nv_bfloat16 v[8];
        if (gr < R && gc + 8 <= C && vec_in) {
            *reinterpret_cast<uint4*>(v) = ld_nc_v4(in + (size_t)gr * ld_in + gc);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                v[j] = (gr < R && gc + j < C) ? in[(size_t)gr * ld_in + gc + j] : __float2bfloat16(0.f);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) tile[r][c8 + j] = v[j];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int orow = (t >> 3) + 32 * i, o8 = (t & 7) * 8;   // out row = input col, out cols = input rows
        const int gr = c0 + orow, gc = r0 + o8;
        __nv_bfloat16 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = tile[o8 + j][orow];
        if (gr < C) {
            if (gc + 8 <= R && vec_out) {
                *reinterpret_cast<uint4*>(out + (size_t)gr * ld_out + gc) = *reinterpret_cast<uint4*>(v);
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (gc + j < R) out[(size_t)gr * ld_out + gc + j] = v[j];
            }
        }
    }
}

}  // namespace

// out[C, R] (row pitch ld_out) = transpose of in[R, C] (row pitch ld_in); bf16.
TFOP_API int tfop_transpose_bf16(const void* in, void* out, int R, int C, int ld_in, int ld_out,
                                 cudaStream_t stream) {
    TFOP_REQUIRE(R > 0 && C > 0 && ld_in >= C && ld_out >= R, "bad transpose shape");
    dim3 grid((C + TILE - 1) / TILE, (R + TILE - 1) / TILE);
    TFOP_REQUIRE(grid.y <= 65535, "too many rows for one launch");
    transpose_bf16_kernel<<<grid, 256, 0, stream>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, R,
                                                    C, ld_in, ld_out);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
