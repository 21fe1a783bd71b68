// K11: the ResNet stem convolution as an implicit GEMM on the tcgen05 tensor cores.
//
// After the space-to-depth preprocessing (preproc.cu) the 7x7/2 stem is a 4x4/1 convolution
// over a padded NHWC input with 16 channels: y[n,oh,ow,:] = sum_{kh,kw,c} x[n,oh+kh,ow+kw,c] *
// w[:,kh,kw,c]. For a fixed kh the 4 kw x 16 c taps of one output pixel are 64 CONTIGUOUS bf16
// (128 bytes) of the input row, so the im2col matrix never has to exist: a 4-D TMA tensor map
// over x with dims {64 elements, ow, ih, n} whose second stride is ONE PIXEL (32 bytes) - i.e.
// rows that overlap their neighbours by 96 bytes - delivers a [128 pixels x 64] K-major operand
// tile, already 128B-swizzled, per kh. (Variant 1 keeps the windows disjoint per load: a 5-D map
// {c, kw, ow, ih, n}, one box per kw, 32-byte rows with 32B swizzle. A 5-D box with a 32-byte
// inner dimension under 128B swizzle does NOT land as dense 128-byte rows - measured, wrong
// numerics - which is why the inner box dimension always equals the swizzle span here.)
//
//   GEMM view: M = N*OH*OW output pixels (tiles of 8 oh x 16 ow), N = 64 output channels,
//   K = 256 = 4 k-blocks (one per kh). The 32 KB weight matrix is loaded once per CTA and stays
//   in shared memory; only x streams. fp32 accumulators in TMEM (double buffered), bf16 output
//   through swizzled boxes and 4-D TMA tensor stores.
//
// cuDNN runs this layer with a 256x64 sm80-style kernel at 1.48 ms (batch 256, 224x224) and its
// weight gradient with an sm80 xmma kernel at 0.75 ms; the layer moves ~520 MB, i.e. ~80 us at
// HBM speed. Reference parity: the stem of the ResNet-50 trainer of BASELINE.json config 2 (the
// reference ships no model code).
#include "common.cuh"
#include "gemm_common.cuh"

int tfop_gemm_num_sms(int* out);  // gemm_tcgen05.cu (library-internal)

namespace {

using namespace tfop_gemm;

constexpr int kCout = 64;
constexpr int kTileOW = 16, kTileOH = 8;          // 128 output pixels per tile
constexpr int kKH = 4;                            // k-blocks
constexpr int kABytes = BLOCK_M * BLOCK_K * 2;    // 16 KB per (tile, kh)
constexpr int kWBytes = kCout * BLOCK_K * 2;      // 8 KB per kh
constexpr int kStagesA = 8;
constexpr int kStoreBox = 32 * 128;               // per epilogue warp: 32 pixels x 64 ch bf16
constexpr int kSmem = kStagesA * kABytes + kKH * kWBytes + 4 * 2 * kStoreBox + 256 + 1024;
constexpr int kTmemCols = 128;                    // 2 accumulator stages x 64 columns

__device__ __forceinline__ void tma_load_5d(const CUtensorMap* map, uint64_t* bar, void* dst,
                                            int c0, int c1, int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* dst,
                                            int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
// K-major operand in SWIZZLE_32B layout: rows of 32 bytes (one UMMA_K step of bf16), 8-row
// groups 256 bytes apart.
__device__ __forceinline__ uint64_t make_smem_desc_sw32(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(256 >> 4) << 32;   // SBO
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)6 << 61;            // SWIZZLE_32B
    return d;
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, const void* src, int c0,
                                             int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
            map),
        "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

__global__ void __launch_bounds__(kNumThreads, 1)
conv_stem_fwd_kernel(const __grid_constant__ CUtensorMap map_x,
                     const __grid_constant__ CUtensorMap map_w,
                     const __grid_constant__ CUtensorMap map_y, int n_img, int oh_blocks,
                     int ow_blocks, int mode) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_a = smem;
    uint8_t* smem_w = smem + kStagesA * kABytes;
    uint8_t* smem_store = smem_w + kKH * kWBytes;
    uint64_t* bars = (uint64_t*)(smem_store + 4 * 2 * kStoreBox);
    uint64_t* full = bars;
    uint64_t* empty = bars + kStagesA;
    uint64_t* tmem_full = bars + 2 * kStagesA;
    uint64_t* tmem_empty = tmem_full + 2;
    uint64_t* w_full = tmem_empty + 2;
    uint32_t* tmem_slot = (uint32_t*)(w_full + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int tiles_per_img = oh_blocks * ow_blocks;
    const int num_tiles = n_img * tiles_per_img;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_y) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStagesA; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], 4);
        }
        mbar_init(w_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            // the whole weight matrix [64, 256], one 64x64 box per kh, resident for the kernel
            mbar_expect_tx(w_full, kKH * kWBytes);
            for (int kh = 0; kh < kKH; ++kh)
                tma_load_2d(&map_w, w_full, smem_w + kh * kWBytes, kh * BLOCK_K, 0);
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int n = t / tiles_per_img;
                const int r = t - n * tiles_per_img;
                const int oh0 = (r / ow_blocks) * kTileOH;
                const int ow0 = (r % ow_blocks) * kTileOW;
                for (int kh = 0; kh < kKH; ++kh) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    mbar_expect_tx(&full[stage], kABytes);
                    if (mode == 0) {
                        // 4-D map, rows = overlapping 64-element windows one pixel apart:
                        // box {64, 16 ow, 8 ih, 1 n} = 128 rows of 128 bytes, 128B swizzle
                        tma_load_4d(&map_x, &full[stage], smem_a + stage * kABytes, 0, ow0,
                                    oh0 + kh, n);
                    } else {
                        // 5-D map {c, kw, ow, ih, n}, one box {16, 1, 16, 8, 1} per kw:
                        // 128 rows of 32 bytes, 32B swizzle (one UMMA_K step each)
#pragma unroll
                        for (int kw = 0; kw < 4; ++kw)
                            tma_load_5d(&map_x, &full[stage], smem_a + stage * kABytes + kw * 4096,
                                        0, kw, ow0, oh0 + kh, n);
                    }
                    if (++stage == kStagesA) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            const uint32_t idesc = make_idesc(BLOCK_M, kCout);
            mbar_wait(w_full, 0);
            tcgen05_fence_after();
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(acc * kCout);
                for (int kh = 0; kh < kKH; ++kh) {
                    mbar_wait(&full[stage], phase);
                    tcgen05_fence_after();
                    const uint32_t a_addr = smem_u32(smem_a + stage * kABytes);
                    const uint32_t b_addr = smem_u32(smem_w + kh * kWBytes);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        const uint64_t adesc = mode == 0 ? make_smem_desc(a_addr + k * UMMA_K * 2)
                                                         : make_smem_desc_sw32(a_addr + k * 4096);
                        umma_bf16(tmem_d, adesc, make_smem_desc(b_addr + k * UMMA_K * 2), idesc,
                                  (kh | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty[stage]);
                    if (++stage == kStagesA) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(&tmem_full[acc]);
                if (++acc == 2) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    } else if (warp >= kEpiWarp0) {
        // ===================== epilogue: TMEM -> bf16 -> swizzled box -> 4-D TMA store ======
        const int ew = warp - kEpiWarp0;
        uint8_t* box0 = smem_store + ew * 2 * kStoreBox;
        const uint32_t row_off = (uint32_t)lane * 128u;
        const uint32_t sw = (uint32_t)(lane & 7);
        int acc = 0, buf = 0;
        uint32_t acc_phase = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            const int n = t / tiles_per_img;
            const int r = t - n * tiles_per_img;
            const int oh0 = (r / ow_blocks) * kTileOH;
            const int ow0 = (r % ow_blocks) * kTileOW;
            mbar_wait(&tmem_full[acc], acc_phase);
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + (uint32_t)(acc * kCout) + ((uint32_t)(ew * 32) << 16);
            uint32_t ra[32], rb[32];
            tmem_ld_32x32b_x32(taddr, ra);
            tmem_ld_32x32b_x32(taddr + 32u, rb);
            tmem_ld_wait();
            // accumulator stage is free again as soon as it is in registers
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[acc]);
            uint8_t* box = box0 + buf * kStoreBox;
            if (lane == 0) bulk_wait_read<1>();
            __syncwarp();
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(ra[j]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                bf16x8 pk = f32_to_bf16x8(f + 8 * j);
                *reinterpret_cast<uint4*>(box + row_off + (((uint32_t)j ^ sw) << 4)) =
                    *reinterpret_cast<uint4*>(&pk);
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(rb[j]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                bf16x8 pk = f32_to_bf16x8(f + 8 * j);
                *reinterpret_cast<uint4*>(box + row_off + (((uint32_t)(j + 4) ^ sw) << 4)) =
                    *reinterpret_cast<uint4*>(&pk);
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
                // this warp's 32 pixels = 2 output rows x 16 columns
                tma_store_4d(&map_y, box, 0, ow0, oh0 + 2 * ew, n);
                bulk_commit();
            }
            buf ^= 1;
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1;
            }
        }
        if (lane == 0) bulk_wait_read<0>();
        __syncwarp();
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 2) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
    }
}

// ---------------------------------------------------------------------------------------------
// Weight gradient: dW[o, kh, kw, c] = sum over pixels of dy[p, o] * x_window[p, (kh, kw, c)].
// As a GEMM: D[m = (kh,kw,c) 0..255, n = o 0..63] = sum_p A[p, m] * dy[p, n] - the contraction
// runs over pixels, so BOTH operands are MN-major exactly as they sit in memory (the same trick
// as the linear-layer wgrad in gemm_tcgen05.cu): a k-block is a patch of 4 oh x 16 ow = 64
// pixels; the A operand is four 64-pixel x 64-element boxes (one per kh, from the overlapping-
// window map), the B operand one 64-pixel x 64-channel box of dy. M = 256 runs as two 128-row
// MMAs into two TMEM accumulators. Split-K over all CTAs (the output is only 64 KB): every CTA
// accumulates its share of the 50k k-blocks in TMEM and adds it to the fp32 result with RED.
constexpr int kWgA = 4 * kMnBoxBytes;              // 32 KB: 4 kh boxes
constexpr int kWgB = kMnBoxBytes;                  // 8 KB: dy
constexpr int kWgStage = kWgA + kWgB;
constexpr int kWgStages = 4;
constexpr int kWgSmem = kWgStages * kWgStage + 256 + 1024;
constexpr int kWgPatchOH = 4;

__global__ void __launch_bounds__(kNumThreads, 1)
conv_stem_wgrad_kernel(const __grid_constant__ CUtensorMap map_x,
                       const __grid_constant__ CUtensorMap map_dy, float* __restrict__ dw,
                       int n_img, int oh_blocks, int ow_blocks) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = (uint64_t*)(smem + kWgStages * kWgStage);
    uint64_t* full = bars;
    uint64_t* empty = bars + kWgStages;
    uint64_t* done = bars + 2 * kWgStages;
    uint32_t* tmem_slot = (uint32_t*)(done + 1);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int per_img = oh_blocks * ow_blocks;
    const int num_kb = n_img * per_img;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_dy) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kWgStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        mbar_init(done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = blockIdx.x; kb < num_kb; kb += gridDim.x) {
                const int n = kb / per_img;
                const int r = kb - n * per_img;
                const int oh0 = (r / ow_blocks) * kWgPatchOH;
                const int ow0 = (r % ow_blocks) * kTileOW;
                mbar_wait(&empty[stage], phase ^ 1);
                mbar_expect_tx(&full[stage], kWgStage);
                uint8_t* sa = smem + stage * kWgStage;
#pragma unroll
                for (int kh = 0; kh < kKH; ++kh)
                    tma_load_4d(&map_x, &full[stage], sa + kh * kMnBoxBytes, 0, ow0, oh0 + kh, n);
                tma_load_4d(&map_dy, &full[stage], sa + kWgA, 0, ow0, oh0, n);
                if (++stage == kWgStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            // both operands MN-major (bits 15 and 16)
            const uint32_t idesc = make_idesc_major(BLOCK_M, kCout, 1, 1);
            int stage = 0;
            uint32_t phase = 0;
            bool first = true;
            for (int kb = blockIdx.x; kb < num_kb; kb += gridDim.x) {
                mbar_wait(&full[stage], phase);
                tcgen05_fence_after();
                const uint32_t a_addr = smem_u32(smem + stage * kWgStage);
                const uint32_t b_addr = a_addr + kWgA;
#pragma unroll
                for (int k = 0; k < 64 / UMMA_K; ++k) {
                    const uint64_t bdesc = make_smem_desc_mn(b_addr + k * 2048);
#pragma unroll
                    for (int h = 0; h < 2; ++h)   // m-halves: kh {0,1} and kh {2,3}
                        umma_bf16(tmem_base + (uint32_t)(h * kCout),
                                  make_smem_desc_mn(a_addr + h * 2 * kMnBoxBytes + k * 2048), bdesc,
                                  idesc, (first && k == 0) ? 0u : 1u);
                }
                first = false;
                umma_commit(&empty[stage]);
                if (++stage == kWgStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit(done);
        }
    } else if (warp >= kEpiWarp0 && (int)blockIdx.x < num_kb) {
        const int ew = warp - kEpiWarp0;
        mbar_wait(done, 0);
        tcgen05_fence_after();
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            const int m = h * 128 + ew * 32 + lane;       // (kh, kw, c) index of this lane's row
            const uint32_t taddr = tmem_base + (uint32_t)(h * kCout) + ((uint32_t)(ew * 32) << 16);
#pragma unroll 1
            for (int c = 0; c < kCout; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(taddr + (uint32_t)c, r);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    atomicAdd(dw + (size_t)(c + j) * 256 + m, __uint_as_float(r[j]));
            }
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 2) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
    }
}

PFN_cuTensorMapEncodeTiled_v12000 g_enc = nullptr;

int g_mode = 0;  // 0: 4-D overlapping-window map + 128B swizzle, 1: 5-D map + 32B swizzle

int encode(CUtensorMap* map, void* base, int rank, const cuuint64_t* dims,
           const cuuint64_t* strides, const cuuint32_t* box, const char* what,
           CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = g_enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, base, dims, strides, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char msg[128];
        snprintf(msg, sizeof(msg), "cuTensorMapEncodeTiled(%s) failed with CUresult %d", what, (int)r);
        tfop_set_error(__FILE__, __LINE__, msg);
        return -1;
    }
    return 0;
}

}  // namespace

// Test hook: operand-fetch variant (0 = 4-D overlapping windows, 128B swizzle; 1 = 5-D, one load
// per kw, 32B swizzle).
TFOP_API int tfop_conv_stem_set_mode(int mode) {
    TFOP_REQUIRE(mode == 0 || mode == 1, "conv_stem mode must be 0 or 1");
    g_mode = mode;
    return 0;
}

// y[N, OH, OW, 64] = conv4x4/1(x[N, OH+3, OW+3, 16], w[64, 4, 4, 16]); all bf16, NHWC / KRSC,
// dense. OH % 8 == 0, OW % 16 == 0. x is the output of tfop_stem_space_to_depth.
TFOP_API int tfop_conv_stem_s2d_fwd(const void* x, const void* w, void* y, int n, int oh, int ow,
                                    cudaStream_t stream) {
    TFOP_REQUIRE(n > 0 && oh > 0 && ow > 0, "conv_stem: empty problem");
    TFOP_REQUIRE(oh % kTileOH == 0 && ow % kTileOW == 0,
                 "conv_stem: OH must be a multiple of 8 and OW a multiple of 16");
    TFOP_REQUIRE(((((uintptr_t)x) | ((uintptr_t)w) | ((uintptr_t)y)) & 15) == 0,
                 "conv_stem: tensors must be 16-byte aligned");
    {   // bind the primary context on this thread (see gemm_tcgen05.cu: bind_context)
        int dev_ = 0;
        TFOP_CUDA(cudaGetDevice(&dev_));
        TFOP_CUDA(cudaSetDevice(dev_));
    }
    if (!g_enc) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        TFOP_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        TFOP_REQUIRE(qres == cudaDriverEntryPointSuccess && fn != nullptr,
                     "cuTensorMapEncodeTiled not available from the driver");
        g_enc = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    }
    const int ih = oh + 3, iw = ow + 3, C = 16;
    CUtensorMap mx, mw, my;
    int mode = g_mode;
    if (mode == 0) {
        // rows of 64 contiguous elements (4 kw x 16 c) starting one pixel (32 bytes) apart
        cuuint64_t dims[4] = {64, (cuuint64_t)ow, (cuuint64_t)ih, (cuuint64_t)n};
        cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)iw * C * 2,
                                 (cuuint64_t)ih * iw * C * 2};
        cuuint32_t box[4] = {64, kTileOW, kTileOH, 1};
        if (encode(&mx, const_cast<void*>(x), 4, dims, strides, box, "x (4-D windows)")) mode = 1;
    }
    if (mode == 1) {
        // {c, kw, ow, ih, n}; kw and ow both advance one pixel; one kw per load
        cuuint64_t dims[5] = {(cuuint64_t)C, 4, (cuuint64_t)ow, (cuuint64_t)ih, (cuuint64_t)n};
        cuuint64_t strides[4] = {(cuuint64_t)C * 2, (cuuint64_t)C * 2, (cuuint64_t)iw * C * 2,
                                 (cuuint64_t)ih * iw * C * 2};
        cuuint32_t box[5] = {(cuuint32_t)C, 1, kTileOW, kTileOH, 1};
        if (encode(&mx, const_cast<void*>(x), 5, dims, strides, box, "x (5-D)",
                   CU_TENSOR_MAP_SWIZZLE_32B))
            return -1;
    }
    {   // w [64, 256] K-major, one 64 x 64 box per kh
        cuuint64_t dims[2] = {256, kCout};
        cuuint64_t strides[1] = {256 * 2};
        cuuint32_t box[2] = {BLOCK_K, kCout};
        if (encode(&mw, const_cast<void*>(w), 2, dims, strides, box, "w")) return -1;
    }
    {   // y {c, ow, oh, n}; one store = 32 pixels (2 rows x 16 columns) x 64 channels
        cuuint64_t dims[4] = {kCout, (cuuint64_t)ow, (cuuint64_t)oh, (cuuint64_t)n};
        cuuint64_t strides[3] = {kCout * 2, (cuuint64_t)ow * kCout * 2,
                                 (cuuint64_t)oh * ow * kCout * 2};
        cuuint32_t box[4] = {kCout, kTileOW, 2, 1};
        if (encode(&my, y, 4, dims, strides, box, "y")) return -1;
    }
    static bool attr_set[64] = {};   // per device
    int dev = 0;
    TFOP_CUDA(cudaGetDevice(&dev));
    TFOP_REQUIRE(dev >= 0 && dev < 64, "device index out of range");
    if (!attr_set[dev]) {
        TFOP_CUDA(cudaFuncSetAttribute(conv_stem_fwd_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        attr_set[dev] = true;
    }
    int sms = 0;
    if (tfop_gemm_num_sms(&sms) != 0) return -1;
    const int oh_blocks = oh / kTileOH, ow_blocks = ow / kTileOW;
    const long tiles = (long)n * oh_blocks * ow_blocks;
    const int grid = (int)(tiles < sms ? tiles : sms);
    conv_stem_fwd_kernel<<<grid, kNumThreads, kSmem, stream>>>(mx, mw, my, n, oh_blocks, ow_blocks,
                                                               mode);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// dw[64, 4, 4, 16] (fp32, ZEROED by the caller; accumulated with RED) = weight gradient of
// tfop_conv_stem_s2d_fwd for dy[N, OH, OW, 64] and x[N, OH+3, OW+3, 16] (bf16, NHWC).
TFOP_API int tfop_conv_stem_s2d_wgrad(const void* x, const void* dy, float* dw, int n, int oh,
                                      int ow, cudaStream_t stream) {
    TFOP_REQUIRE(n > 0 && oh > 0 && ow > 0, "conv_stem: empty problem");
    TFOP_REQUIRE(oh % kWgPatchOH == 0 && ow % kTileOW == 0,
                 "conv_stem wgrad: OH must be a multiple of 4 and OW a multiple of 16");
    TFOP_REQUIRE(((((uintptr_t)x) | ((uintptr_t)dy) | ((uintptr_t)dw)) & 15) == 0,
                 "conv_stem: tensors must be 16-byte aligned");
    {   // bind the primary context on this thread (see gemm_tcgen05.cu: bind_context)
        int dev_ = 0;
        TFOP_CUDA(cudaGetDevice(&dev_));
        TFOP_CUDA(cudaSetDevice(dev_));
    }
    if (!g_enc) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        TFOP_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        TFOP_REQUIRE(qres == cudaDriverEntryPointSuccess && fn != nullptr,
                     "cuTensorMapEncodeTiled not available from the driver");
        g_enc = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    }
    const int ih = oh + 3, iw = ow + 3, C = 16;
    CUtensorMap mx, mdy;
    {
        cuuint64_t dims[4] = {64, (cuuint64_t)ow, (cuuint64_t)ih, (cuuint64_t)n};
        cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)iw * C * 2,
                                 (cuuint64_t)ih * iw * C * 2};
        cuuint32_t box[4] = {64, kTileOW, kWgPatchOH, 1};
        if (encode(&mx, const_cast<void*>(x), 4, dims, strides, box, "x (wgrad windows)")) return -1;
    }
    {
        cuuint64_t dims[4] = {kCout, (cuuint64_t)ow, (cuuint64_t)oh, (cuuint64_t)n};
        cuuint64_t strides[3] = {kCout * 2, (cuuint64_t)ow * kCout * 2,
                                 (cuuint64_t)oh * ow * kCout * 2};
        cuuint32_t box[4] = {kCout, kTileOW, kWgPatchOH, 1};
        if (encode(&mdy, const_cast<void*>(dy), 4, dims, strides, box, "dy")) return -1;
    }
    static bool attr_set[64] = {};   // per device
    int dev = 0;
    TFOP_CUDA(cudaGetDevice(&dev));
    TFOP_REQUIRE(dev >= 0 && dev < 64, "device index out of range");
    if (!attr_set[dev]) {
        TFOP_CUDA(cudaFuncSetAttribute(conv_stem_wgrad_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, kWgSmem));
        attr_set[dev] = true;
    }
    int sms = 0;
    if (tfop_gemm_num_sms(&sms) != 0) return -1;
    const int oh_blocks = oh / kWgPatchOH, ow_blocks = ow / kTileOW;
    const long kbs = (long)n * oh_blocks * ow_blocks;
    const int grid = (int)(kbs < sms ? kbs : sms);
    conv_stem_wgrad_kernel<<<grid, kNumThreads, kWgSmem, stream>>>(mx, mdy, dw, n, oh_blocks,
                                                                    ow_blocks);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
