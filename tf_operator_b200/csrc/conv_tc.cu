// K12: NHWC bf16 convolutions (1x1 and 3x3, stride 1 and 2) as implicit GEMMs on the tcgen05
// tensor cores - forward, data gradient and weight gradient of every convolution of the ResNet
// family except the stem (conv_stem.cu). No im2col matrix, no padded copy, no transposed weight
// ever exists in memory:
//
//   forward   y[p, co]  = sum_{tap, ci} x[p + d(tap), ci] * w[co, tap, ci]
//             GEMM view: M = output pixels in tiles of 128 (a tw x th x nb box of the NHW grid,
//             powers of two picked per layer so the boxes tile the grid without waste),
//             N = Cout, K = taps x Cin. A k-block is (tap, 64 channels): ONE 4-D TMA box of x
//             shifted by the tap offset - the hardware zero-fills what falls outside the image,
//             which is the padding - and one 2-D box of the KRSC weight matrix.
//             Stride 2 reads x through four "parity" tensor maps (even/odd rows x even/odd
//             columns, strides doubled), so every tap is again a dense shifted box.
//   dgrad     the same kernel on dy with the taps mirrored; the weight matrix is read as it is
//             stored ([Cout, tap, Cin] = MN-major for this product). Stride 2 runs once per
//             parity class of dx (1, 2, 2 and 4 taps) and stores through a strided tensor map.
//             `accumulate` turns the TMA stores into TMA reduce-adds: the residual branch's
//             gradient is added in the epilogue instead of by a separate elementwise kernel.
//   epilogue  TMEM -> registers -> bf16 -> swizzled shared-memory boxes -> TMA tensor stores.
//             With `stats` the per-channel sum and sum of squares of the (rounded) outputs are
//             reduced from the staged boxes and added to a [2, Cout] fp32 array: BatchNorm's
//             statistics pass over the convolution output disappears (bn.cu: tfop_bn_finalize).
//   wgrad     dW[co, tap, ci] = sum_p dy[p, co] * x[p + d(tap), ci]: contraction over pixels, both
//             operands MN-major as they sit in memory. M = two (tap, 64-channel) slices of x,
//             N = Cout (<= 256 per tile), K = pixels in blocks of 64, split over all SMs; fp32
//             partial tiles are added to dW by TMA reduce-add stores.
//
// Mainloop, barriers and CTA pairing are those of gemm_tcgen05.cu (warp 0 TMA producer, warp 1
// one-thread tcgen05.mma issuer with cta_group::2 pairs on adjacent pixel tiles, warps 4-7
// epilogue, two TMEM accumulator stages).
//
// Reference parity: the convolutions of the CNN trained by
// examples/v1/distribution_strategy/keras-API/multi_worker_strategy-with-keras.py:42-60 (Keras
// Conv2D -> cuDNN in the reference stack); here they are the ResNet-50 layers of BASELINE.json
// config 2.
#include "common.cuh"
#include "gemm_common.cuh"

int tfop_gemm_num_sms(int* out);  // gemm_tcgen05.cu (library-internal)

namespace {

using namespace tfop_gemm;

constexpr int kMaxTaps = 9;

// n / d and n % d for 0 <= n < 2^24 by one 64-bit multiply (d < 2^16): the tile loops decode
// (w, h, n) block indices for every tile and k-block, and a runtime-divisor IDIV is ~25 instructions.
struct FastDiv {
    unsigned long long magic;  // ceil(2^40 / d)
    int d;
    __host__ void set(int div) {
        d = div;
        magic = ((1ull << 40) + (unsigned long long)div - 1) / (unsigned long long)div;
    }
    __device__ __forceinline__ void divmod(int n, int& q, int& r) const {
        q = (int)(((unsigned long long)(unsigned)n * magic) >> 40);
        r = n - q * d;
    }
};

struct ConvMaps {
    CUtensorMap a[4];  // activation operand: dense map, or the four parity maps of a stride-2 read
    CUtensorMap b;     // weights [Cout, taps * Cin]
    CUtensorMap c;     // output (store / reduce-add target)
};

struct ConvArgs {
    int n_taps, cblocks;           // k-blocks = n_taps * cblocks (cblocks = contraction channels / 64)
    int tap_map[kMaxTaps], tap_dh[kMaxTaps], tap_dw[kMaxTaps], tap_bcol[kMaxTaps];
    int tw, th, nb;                // pixel tile: tw * th * nb == 128
    int wblk, hblk, nblk;          // tiles per dimension
    FastDiv div_wblk, div_hblk, div_nblocks;
    int sw, sh;                    // 32-row store box: sw x sh x (32 / (sw * sh))
    int W, H, N;                   // output pixel grid (for store guards)
    int ncols;                     // GEMM N = output channels of this product
    int b_mn;                      // weights read MN-major (dgrad)
    int accumulate;                // TMA reduce-add instead of store
    float* stats;                  // [2, ncols] fp32 sums, or null
};

__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* dst,
                                            int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d_pair(const CUtensorMap* map, uint64_t* bar, void* dst,
                                                 int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, const void* src, int c0,
                                             int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
            map),
        "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma_reduce_add_4d(const CUtensorMap* map, const void* src, int c0,
                                                  int c1, int c2, int c3) {
    asm volatile(
        "cp.reduce.async.bulk.tensor.4d.global.shared::cta.add.tile.bulk_group"
        " [%0, {%2, %3, %4, %5}], [%1];" ::"l"(map),
        "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

constexpr int kConvThreads = 384;  // warps 0-3 as in the GEMM, 4-7 and 8-11 two epilogue groups

// barrier among the 128 threads of one epilogue group (ids 1 and 2; 0 is __syncthreads)
__device__ __forceinline__ void epi_group_sync(int grp) {
    if (grp == 0)
        asm volatile("bar.sync 1, 128;" ::: "memory");
    else
        asm volatile("bar.sync 2, 128;" ::: "memory");
}

template <int BN, int kPair>
struct CCfg {
    // one stage fewer than the plain GEMM: the statistics partials take its place
    static constexpr int kStages = kPair == 2 ? (BN == 256 ? 5 : (BN == 128 ? 7 : 8))
                                              : (BN == 256 ? 3 : (BN == 128 ? 5 : 6));
    static_assert(BN <= 256, "two accumulator stages of BN columns must fit TMEM");
    static constexpr int kABytes = BLOCK_M * BLOCK_K * 2;
    static constexpr int kBBytes = (BN / kPair) * BLOCK_K * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kBarBytes = 256;
    static constexpr int kStoreBoxBytes = 32 * 128;
    static constexpr int kStoreBytes = 8 * kStoreBoxBytes;   // one staging box per epilogue warp
    static constexpr int kStatBytes = 8 * 2 * BN * 4;  // per epilogue warp: sum and sum of squares
    static constexpr int kSmemBytes =
        kStages * kStageBytes + kStoreBytes + kBarBytes + kStatBytes + 1024;
    static_assert(kSmemBytes <= 227 * 1024, "shared memory budget of one CTA");
};

// Packed fp32 pairs (FADD2 / FFMA2): one instruction for the two channels a lane owns.
__device__ __forceinline__ unsigned long long pack_f32x2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ unsigned long long add_f32x2(unsigned long long a, unsigned long long b) {
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ unsigned long long fma_f32x2(unsigned long long a, unsigned long long b,
                                                        unsigned long long c) {
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// Per-channel sum / sum of squares over the 32 rows of a staged (swizzled) 32 x 64 bf16 box:
// lane l owns channels 2l and 2l+1. All lanes read the same row at once: conflict-free. Rows
// whose bit in `rows` is clear lie outside the image (partial tiles: a 3x3 tap still reaches
// valid input from there, so they are not zero) and are left out. The swizzle term depends on
// r & 7 only: eight base pointers, then every row is one LDS with an immediate offset, a shift,
// a mask and two packed fp32 operations (was ~13 instructions per row: ncu, source page).
template <bool kMasked>
__device__ __forceinline__ void box_col_sums(const uint8_t* box, int lane, uint32_t rows, float& s0,
                                             float& s1, float& q0, float& q1) {
    const uint32_t j = (uint32_t)lane >> 2, o = ((uint32_t)lane & 3u) << 2;
    const uint8_t* base[8];
#pragma unroll
    for (uint32_t k = 0; k < 8; ++k) base[k] = box + ((j ^ k) << 4) + o;
    unsigned long long s[2] = {0ull, 0ull}, q[2] = {0ull, 0ull};
#pragma unroll
    for (uint32_t r = 0; r < 32; ++r) {
        uint32_t v = *reinterpret_cast<const uint32_t*>(base[r & 7u] + r * 128u);
        if (kMasked && !((rows >> r) & 1u)) v = 0u;
        const unsigned long long a =
            pack_f32x2(__uint_as_float(v << 16), __uint_as_float(v & 0xffff0000u));
        s[r & 1u] = add_f32x2(s[r & 1u], a);
        q[r & 1u] = fma_f32x2(a, a, q[r & 1u]);
    }
    const unsigned long long st = add_f32x2(s[0], s[1]), qt = add_f32x2(q[0], q[1]);
    asm("mov.b64 {%0, %1}, %2;" : "=f"(s0), "=f"(s1) : "l"(st));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(q0), "=f"(q1) : "l"(qt));
}

template <int BN, int kCluster>
__global__ void __maxnreg__(128)
conv_igemm_kernel(const __grid_constant__ ConvMaps maps, const __grid_constant__ ConvArgs args) {
    using cfg = CCfg<BN, kCluster>;
    constexpr bool kPair = kCluster == 2;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + cfg::kStages * cfg::kABytes;
    uint8_t* smem_store = smem + cfg::kStages * cfg::kStageBytes;
    uint64_t* bars = (uint64_t*)(smem_store + cfg::kStoreBytes);
    float* sstat = (float*)(smem_store + cfg::kStoreBytes + cfg::kBarBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + cfg::kStages;
    uint64_t* tmem_full = bars + 2 * cfg::kStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_slot = (uint32_t*)(tmem_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m_blocks = args.wblk * args.hblk * args.nblk;
    const int n_blocks = (args.ncols + BN - 1) / BN;
    const int k_blocks = args.n_taps * args.cblocks;
    const int crank = kCluster == 2 ? (int)cluster_ctarank() : 0;
    const int m_units = (m_blocks + kCluster - 1) / kCluster;
    const int num_tiles = m_units * n_blocks;
    const int t_first = blockIdx.x / kCluster;
    const int t_step = gridDim.x / kCluster;

    if (warp == 0 && lane == 0) {
        for (int i = 0; i < 4; ++i)
            asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.a[i]) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.b) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.c) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < cfg::kStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], 4 * kCluster);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        if (kPair) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             smem_u32(tmem_slot)),
                         "r"((uint32_t)cfg::kTmemCols)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             smem_u32(tmem_slot)),
                         "r"((uint32_t)cfg::kTmemCols)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (kCluster == 2) cluster_sync_all();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // pixel tile index -> origin in the (w, h, n) grid
    auto tile_origin = [&](int mb, int& w0, int& h0, int& n0) {
        int q, wb, hb, nbk;
        args.div_wblk.divmod(mb, q, wb);
        args.div_hblk.divmod(q, nbk, hb);
        w0 = wb * args.tw;
        h0 = hb * args.th;
        n0 = nbk * args.nb;
    };
    // Tile order: the n-blocks of one pixel tile are consecutive, so the CTAs that need the same
    // activation boxes run at the same time (the weights are small and always in L2).
    auto tile_decode = [&](int t, int& mu, int& nb) { args.div_nblocks.divmod(t, mu, nb); };

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = t_first; t < num_tiles; t += t_step) {
                int mu, nb;
                tile_decode(t, mu, nb);
                int w0, h0, n0;
                tile_origin(mu * kCluster + crank, w0, h0, n0);
                const int col0 = nb * BN + (kPair ? crank * (BN / 2) : 0);
                for (int tap = 0; tap < args.n_taps; ++tap) {
                    const CUtensorMap* ma = &maps.a[args.tap_map[tap]];
                    const int ww = w0 + args.tap_dw[tap], hh = h0 + args.tap_dh[tap];
                    const int bcol = args.tap_bcol[tap];
                    for (int cb = 0; cb < args.cblocks; ++cb) {
                        mbar_wait(&empty[stage], phase ^ 1);
                        uint8_t* sa = smem_a + stage * cfg::kABytes;
                        uint8_t* sb = smem_b + stage * cfg::kBBytes;
                        if (kPair) {
                            if (crank == 0) mbar_expect_tx(&full[stage], 2 * cfg::kStageBytes);
                            tma_load_4d_pair(ma, &full[stage], sa, cb * BLOCK_K, ww, hh, n0);
                            if (!args.b_mn) {
                                tma_load_2d_pair(&maps.b, &full[stage], sb, bcol + cb * BLOCK_K, col0);
                            } else {
#pragma unroll
                                for (int j = 0; j < BN / 128; ++j)
                                    tma_load_2d_pair(&maps.b, &full[stage], sb + j * kMnBoxBytes,
                                                     bcol + col0 + 64 * j, cb * BLOCK_K);
                            }
                        } else {
                            mbar_expect_tx(&full[stage], cfg::kStageBytes);
                            tma_load_4d(ma, &full[stage], sa, cb * BLOCK_K, ww, hh, n0);
                            if (!args.b_mn) {
                                tma_load_2d(&maps.b, &full[stage], sb, bcol + cb * BLOCK_K, col0);
                            } else {
#pragma unroll
                                for (int j = 0; j < BN / 64; ++j)
                                    tma_load_2d(&maps.b, &full[stage], sb + j * kMnBoxBytes,
                                                bcol + col0 + 64 * j, cb * BLOCK_K);
                            }
                        }
                        if (++stage == cfg::kStages) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (the even CTA of a pair) =====================
        if ((!kPair || crank == 0) && elect_one()) {
            const uint32_t idesc = make_idesc_major(BLOCK_M * kCluster, BN, 0, args.b_mn);
            const uint32_t b_step = args.b_mn ? 2048u : (uint32_t)(UMMA_K * 2);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int t = t_first; t < num_tiles; t += t_step) {
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tcgen05_fence_after();
                    const uint32_t a_addr = smem_u32(smem_a + stage * cfg::kABytes);
                    const uint32_t b_addr = smem_u32(smem_b + stage * cfg::kBBytes);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        const uint64_t adesc = make_smem_desc(a_addr + k * (UMMA_K * 2));
                        const uint64_t bdesc = args.b_mn ? make_smem_desc_mn(b_addr + k * b_step)
                                                         : make_smem_desc(b_addr + k * b_step);
                        if (kPair)
                            umma_bf16_pair(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                        else
                            umma_bf16(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    if (kPair)
                        umma_commit_pair(&empty[stage]);
                    else
                        umma_commit(&empty[stage]);
                    if (++stage == cfg::kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                if (kPair)
                    umma_commit_pair(&tmem_full[acc]);
                else
                    umma_commit(&tmem_full[acc]);
                if (++acc == 2) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    } else if (warp >= kEpiWarp0) {
        // ===================== epilogue: two groups of four warps =====================
        // Group g drains accumulator stage g, i.e. every other tile of this CTA, so two tiles'
        // epilogues (TMEM loads, conversion, statistics) run side by side on the four SM
        // sub-partitions: with short K (1x1 layers off 64 channels: one k-block per tile) the
        // epilogue, not the tensor pipe, is what a tile costs.
        const int grp = (warp - kEpiWarp0) >> 2;
        const int ew = warp & 3;  // TMEM lane quadrant of this warp
        const int acc = grp;
        uint32_t acc_phase = 0;
        uint8_t* box = smem_store + (grp * 4 + ew) * cfg::kStoreBoxBytes;
        const uint32_t row_off = (uint32_t)lane * 128u;
        const uint32_t sw = (uint32_t)(lane & 7);
        float* gstat = sstat + grp * 8 * BN;
        float* mystat = gstat + ew * 2 * BN;
        const bool want_stats = args.stats != nullptr;
        // this warp's 32 rows of the tile as a sub-box of the pixel tile
        const int r0 = ew * 32;
        const int tw0 = r0 % args.tw;
        const int th0 = (r0 / args.tw) % args.th;
        const int nb0 = r0 / (args.tw * args.th);
        // this lane's row inside the warp's 32-row sub-box
        const int lw = lane % args.sw, lh = (lane / args.sw) % args.sh, ln = lane / (args.sw * args.sh);
        for (int t = t_first + grp * t_step; t < num_tiles; t += 2 * t_step) {
            int mu, nb;
            tile_decode(t, mu, nb);
            int w0, h0, n0;
            tile_origin(mu * kCluster + crank, w0, h0, n0);
            w0 += tw0;
            h0 += th0;
            n0 += nb0;
            const bool in_grid = w0 < args.W && h0 < args.H && n0 < args.N;
            // which of this warp's 32 rows are pixels of the image (lane = row)
            uint32_t rows = 0xffffffffu;
            if (want_stats)
                rows = __ballot_sync(0xffffffffu,
                                     w0 + lw < args.W && h0 + lh < args.H && n0 + ln < args.N);
            const int col0 = nb * BN;
            mbar_wait(&tmem_full[acc], acc_phase);
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + (uint32_t)(acc * BN) + ((uint32_t)(ew * 32) << 16);
            uint32_t ra[32], rb[32];
            tmem_ld_32x32b_x32(taddr, ra);
#pragma unroll 1
            for (int c = 0; c < BN; c += 64) {
                tmem_ld_wait();
                tmem_ld_32x32b_x32(taddr + (uint32_t)(c + 32), rb);
                // one staging box per warp: the previous chunk's store must have read it
                if (lane == 0) bulk_wait_read<0>();
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
                tmem_ld_wait();
                if (c + 64 < BN) tmem_ld_32x32b_x32(taddr + (uint32_t)(c + 64), ra);
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
                    if (in_grid && col0 + c < args.ncols) {
                        if (args.accumulate)
                            tma_reduce_add_4d(&maps.c, box, col0 + c, w0, h0, n0);
                        else
                            tma_store_4d(&maps.c, box, col0 + c, w0, h0, n0);
                    }
                    bulk_commit();
                }
                if (want_stats) {
                    float s0, s1, q0, q1;
                    if (rows == 0xffffffffu)
                        box_col_sums<false>(box, lane, rows, s0, s1, q0, q1);
                    else
                        box_col_sums<true>(box, lane, rows, s0, s1, q0, q1);
                    *reinterpret_cast<float2*>(mystat + c + 2 * lane) = make_float2(s0, s1);
                    *reinterpret_cast<float2*>(mystat + BN + c + 2 * lane) = make_float2(q0, q1);
                }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (kPair)
                    mbar_arrive_leader(&tmem_empty[acc]);
                else
                    mbar_arrive(&tmem_empty[acc]);
            }
            acc_phase ^= 1;
            if (want_stats) {
                // fold the four warps' partials, one RED per channel and moment per tile
                epi_group_sync(grp);
                for (int i = ew * 32 + lane; i < 2 * BN; i += 128) {
                    const int ch = i < BN ? i : i - BN;
                    if (col0 + ch < args.ncols) {
                        const float v = gstat[i] + gstat[2 * BN + i] + gstat[4 * BN + i] +
                                        gstat[6 * BN + i];
                        atomicAdd(args.stats + (i < BN ? 0 : args.ncols) + col0 + ch, v);
                    }
                }
                epi_group_sync(grp);
            }
        }
        if (lane == 0) bulk_wait_read<0>();
        __syncwarp();
    }

    tcgen05_fence_before();
    __syncthreads();
    if (kCluster == 2) cluster_sync_all();
    if (warp == 2) {
        tcgen05_fence_after();
        if (kPair)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                         "r"((uint32_t)cfg::kTmemCols)
                         : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                         "r"((uint32_t)cfg::kTmemCols)
                         : "memory");
    }
}

// ---------------------------------------------------------------------------------------------
// Weight gradient. One scheduling unit = one (M tile, N tile, pixel range): the M tile is a run of
// (tap, 64 input channels) slices of x, the N tile up to 256 output channels of dy, K the unit's
// share of the 64-pixel blocks. Both operands MN-major. Units with the same pixel range and
// different tiles are adjacent in the grid, so the boxes they share come out of L2.
//   BN >= 128: units run on CTA pairs (cta_group::2): 256 x BN MMAs, each CTA stages its own two
//              slices of x and HALF of the dy tile - 32 KB per k-block and CTA instead of 48 KB.
//              The kernel is bound by how fast TMA can fill shared memory (ncu: tensor pipe 39 %
//              busy, both the producer and the MMA thread waiting), so bytes per MMA is the lever.
//   BN == 64:  a CTA holds four slices (two 128-row MMAs into two accumulators per k-step) so the
//              8 KB dy box is shared by twice the work.
struct WgradMaps {
    CUtensorMap x[4];
    CUtensorMap dy;
    CUtensorMap dw;  // fp32 [Cout rows, n_taps * Cin], boxes of 32 x 32, reduce-add target
};
struct WgradArgs {
    int n_taps, cblocks;   // M slices = n_taps * cblocks (cblocks = Cin / 64)
    int tap_map[kMaxTaps], tap_dh[kMaxTaps], tap_dw[kMaxTaps];
    int pw, ph, pn;        // pixel block: pw * ph * pn == 64
    int wblk, hblk, nblk;  // pixel blocks per dimension
    FastDiv div_wblk, div_hblk;
    int m_units, n_tiles, splits;
    int Cin, Cout;
};

template <int BN, int CL>
struct WCfg {
    static constexpr int kSlices = BN == 64 ? 4 : 2;      // slices of x per CTA
    static constexpr int kABytes = kSlices * kMnBoxBytes;
    static constexpr int kBBytes = (BN / CL / 64) * kMnBoxBytes;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = kStageBytes <= 32768 ? 6 : 4;
    static constexpr int kBoxBytes = 32 * 32 * 4;  // per epilogue warp: a transposed 32 x 32 fp32 box
    static constexpr int kTmemCols = (kSlices / 2) * BN;
    static constexpr int kSmemBytes = kStages * kStageBytes + 4 * kBoxBytes + 256 + 1024;
    static_assert(kSmemBytes <= 227 * 1024, "shared memory budget of one CTA");
    static_assert(BN / CL >= 64, "a CTA stages whole 64-column boxes of dy");
};

template <int BN, int CL>
__global__ void __launch_bounds__(kNumThreads, 1)
conv_wgrad_kernel(const __grid_constant__ WgradMaps maps, const __grid_constant__ WgradArgs args) {
    using cfg = WCfg<BN, CL>;
    constexpr bool kPair = CL == 2;
    constexpr int MS = cfg::kSlices;
    constexpr int kStages = cfg::kStages;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_box = smem + kStages * cfg::kStageBytes;
    uint64_t* bars = (uint64_t*)(smem_box + 4 * cfg::kBoxBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + kStages;
    uint64_t* done = bars + 2 * kStages;
    uint32_t* tmem_slot = (uint32_t*)(done + 1);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int crank = kPair ? (int)cluster_ctarank() : 0;

    // unit -> (tile, split): tiles of one pixel range are neighbours
    const int unit = blockIdx.x / CL;
    const int tiles = args.m_units * args.n_tiles;
    const int tile = unit % tiles;
    const int split = unit / tiles;
    const int mt = (tile % args.m_units) * CL + crank;   // this CTA's run of MS slices
    const int nt = tile / args.m_units;
    const int pblocks = args.wblk * args.hblk * args.nblk;
    const int pb0 = (int)((long)pblocks * split / args.splits);
    const int pb1 = (int)((long)pblocks * (split + 1) / args.splits);
    const int n_slices = args.n_taps * args.cblocks;

    if (warp == 0 && lane == 0) {
        for (int i = 0; i < 4; ++i)
            asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.x[i]) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.dy) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.dw) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        mbar_init(done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        if (kPair) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             smem_u32(tmem_slot)),
                         "r"((uint32_t)cfg::kTmemCols)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             smem_u32(tmem_slot)),
                         "r"((uint32_t)cfg::kTmemCols)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (kPair) cluster_sync_all();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (elect_one()) {
            // this CTA's slices; a slice past the end is fetched from outside the batch (all
            // zeros) and skipped in the epilogue
            int s_map[MS], s_dh[MS], s_dw[MS], s_c0[MS];
            bool s_ok[MS];
#pragma unroll
            for (int j = 0; j < MS; ++j) {
                const int sl = mt * MS + j;
                s_ok[j] = sl < n_slices;
                const int tap = s_ok[j] ? sl / args.cblocks : 0;
                s_map[j] = args.tap_map[tap];
                s_dh[j] = args.tap_dh[tap];
                s_dw[j] = args.tap_dw[tap];
                s_c0[j] = s_ok[j] ? (sl - tap * args.cblocks) * 64 : 0;
            }
            const int n_out = args.nblk * args.pn + 64;
            const int dy_col = nt * BN + crank * (BN / CL);
            int stage = 0;
            uint32_t phase = 0;
            for (int pb = pb0; pb < pb1; ++pb) {
                int q, wb, hb, nbk;
                args.div_wblk.divmod(pb, q, wb);
                args.div_hblk.divmod(q, nbk, hb);
                const int w0 = wb * args.pw, h0 = hb * args.ph, n0 = nbk * args.pn;
                mbar_wait(&empty[stage], phase ^ 1);
                uint8_t* sa = smem + stage * cfg::kStageBytes;
                if (kPair) {
                    if (crank == 0) mbar_expect_tx(&full[stage], 2 * cfg::kStageBytes);
#pragma unroll
                    for (int j = 0; j < MS; ++j)
                        tma_load_4d_pair(&maps.x[s_map[j]], &full[stage], sa + j * kMnBoxBytes,
                                         s_c0[j], w0 + s_dw[j], h0 + s_dh[j], s_ok[j] ? n0 : n_out);
#pragma unroll
                    for (int j = 0; j < BN / CL / 64; ++j)
                        tma_load_4d_pair(&maps.dy, &full[stage],
                                         sa + cfg::kABytes + j * kMnBoxBytes, dy_col + 64 * j, w0,
                                         h0, n0);
                } else {
                    mbar_expect_tx(&full[stage], cfg::kStageBytes);
#pragma unroll
                    for (int j = 0; j < MS; ++j)
                        tma_load_4d(&maps.x[s_map[j]], &full[stage], sa + j * kMnBoxBytes, s_c0[j],
                                    w0 + s_dw[j], h0 + s_dh[j], s_ok[j] ? n0 : n_out);
#pragma unroll
                    for (int j = 0; j < BN / 64; ++j)
                        tma_load_4d(&maps.dy, &full[stage], sa + cfg::kABytes + j * kMnBoxBytes,
                                    dy_col + 64 * j, w0, h0, n0);
                }
                if (++stage == kStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if ((!kPair || crank == 0) && elect_one()) {
            const uint32_t idesc = make_idesc_major(BLOCK_M * CL, BN, 1, 1);
            int stage = 0;
            uint32_t phase = 0;
            for (int pb = pb0; pb < pb1; ++pb) {
                mbar_wait(&full[stage], phase);
                tcgen05_fence_after();
                const uint32_t a_addr = smem_u32(smem + stage * cfg::kStageBytes);
                const uint32_t b_addr = a_addr + cfg::kABytes;
#pragma unroll
                for (int k = 0; k < 64 / UMMA_K; ++k) {
                    const uint64_t bdesc = make_smem_desc_mn(b_addr + k * 2048);
#pragma unroll
                    for (int h = 0; h < MS / 2; ++h) {
                        const uint64_t adesc =
                            make_smem_desc_mn(a_addr + h * 2 * kMnBoxBytes + k * 2048);
                        const uint32_t accf = (pb != pb0 || k != 0) ? 1u : 0u;
                        if (kPair)
                            umma_bf16_pair(tmem_base + (uint32_t)(h * BN), adesc, bdesc, idesc, accf);
                        else
                            umma_bf16(tmem_base + (uint32_t)(h * BN), adesc, bdesc, idesc, accf);
                    }
                }
                if (kPair)
                    umma_commit_pair(&empty[stage]);
                else
                    umma_commit(&empty[stage]);
                if (++stage == kStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            if (kPair)
                umma_commit_pair(done);
            else
                umma_commit(done);
        }
    } else if (warp >= kEpiWarp0 && pb1 > pb0) {
        // TMEM -> registers -> transposed into a [32 co][32 m] fp32 box -> one TMA reduce-add per
        // box: row m of the accumulator (lane) is a run of consecutive addresses of dW[co, :],
        // so the lanes of one STS write one 128-byte row of the box. (Per-element RED from the
        // registers cost ~40 issue cycles per warp instruction: 22 us per CTA at 128 x 256.)
        const int ew = warp - kEpiWarp0;
        mbar_wait(done, 0);
        tcgen05_fence_after();
        float* box = (float*)(smem_box + ew * cfg::kBoxBytes);
#pragma unroll 1
        for (int h = 0; h < MS / 2; ++h) {
            const int m0 = mt * (64 * MS) + h * 128 + ew * 32;   // column of dW = (tap, ci) flattened
            const bool ok = m0 < n_slices * 64;
            const uint32_t taddr = tmem_base + (uint32_t)(h * BN) + ((uint32_t)(ew * 32) << 16);
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(taddr + (uint32_t)c, r);
                tmem_ld_wait();
                if (lane == 0) bulk_wait_read<0>();
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 32; ++j) box[j * 32 + lane] = __uint_as_float(r[j]);
                fence_async_smem();
                __syncwarp();
                if (lane == 0) {
                    if (ok && nt * BN + c < args.Cout)
                        tma_reduce_add_2d(&maps.dw, box, m0, nt * BN + c);
                    bulk_commit();
                }
            }
        }
        if (lane == 0) bulk_wait_read<0>();
        __syncwarp();
    }

    tcgen05_fence_before();
    __syncthreads();
    if (kPair) cluster_sync_all();
    if (warp == 2) {
        tcgen05_fence_after();
        if (kPair)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                         "r"((uint32_t)cfg::kTmemCols)
                         : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                         "r"((uint32_t)cfg::kTmemCols)
                         : "memory");
    }
}

// ---------------------------------------------------------------------------------------------
// host side
PFN_cuTensorMapEncodeTiled_v12000 g_enc = nullptr;

// A driver-API call (cuTensorMapEncodeTiled) needs a current context on the calling thread. An
// autograd worker thread whose first GPU work is one of these entry points may have none yet:
// cudaSetDevice on the thread's current device binds the primary context (legal under capture).
static int bind_context() {
    int dev = 0;
    TFOP_CUDA(cudaGetDevice(&dev));
    TFOP_CUDA(cudaSetDevice(dev));
    return 0;
}

int get_enc() {
    if (bind_context() != 0) return -1;
    if (g_enc) return 0;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    TFOP_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    TFOP_REQUIRE(qres == cudaDriverEntryPointSuccess && fn != nullptr,
                 "cuTensorMapEncodeTiled not available from the driver");
    g_enc = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    return 0;
}

int encode(CUtensorMap* map, const void* base, int rank, const cuuint64_t* dims,
           const cuuint64_t* strides, const cuuint32_t* box, const char* what) {
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = g_enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), dims,
                       strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char msg[160];
        snprintf(msg, sizeof(msg), "cuTensorMapEncodeTiled(%s) failed with CUresult %d", what, (int)r);
        tfop_set_error(__FILE__, __LINE__, msg);
        return -1;
    }
    return 0;
}

// NHWC activation tensor [N, H, W, C] read with pixel stride `s` starting at pixel (pr, pc):
// dims {C, ceil((W - pc) / s), ceil((H - pr) / s), N}; box {64, bw, bh, bn}.
int act_map(CUtensorMap* map, const void* base, int N, int H, int W, int C, int s, int pr, int pc,
            int bw, int bh, int bn, const char* what) {
    const int w = (W - pc + s - 1) / s, h = (H - pr + s - 1) / s;
    if (w <= 0 || h <= 0) {  // an empty parity plane (1-pixel images): any valid map, never hit
        return act_map(map, base, N, H, W, C, 1, 0, 0, bw, bh, bn, what);
    }
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)s * C * 2, (cuuint64_t)s * W * C * 2,
                             (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)bw, (cuuint32_t)bh, (cuuint32_t)bn};
    const uint8_t* b = (const uint8_t*)base + ((size_t)pr * W + pc) * C * 2;
    return encode(map, b, 4, dims, strides, box, what);
}

// weights [rows = Cout, cols = taps * Cin] row-major, box {64 cols, box_rows}
int weight_map(CUtensorMap* map, const void* w, int rows, int cols, int box_rows) {
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    return encode(map, w, 2, dims, strides, box, "weights");
}

struct Geom {
    int tw, th, nb;
};

// Split `total` (128 or 64, a power of two) pixels into a tw x th x nb box (powers of two) that
// tiles the W x H x N grid with the least padding; on near-ties prefer long runs along W.
Geom pick_geom(int W, int H, int N, int total) {
    Geom best{1, 1, total};
    double best_score = -1.0;
    for (int tw = 1; tw <= total; tw *= 2)
        for (int th = 1; tw * th <= total; th *= 2) {
            const int nb = total / (tw * th);
            if (tw > 256 || th > 256 || nb > 256) continue;
            const double covered = (double)((W + tw - 1) / tw * tw) * ((H + th - 1) / th * th) *
                                   ((N + nb - 1) / nb * nb);
            double score = (double)W * H * N / covered;
            // two images' worth of penalty per halving of the run length below 16 pixels
            int run = tw < 16 ? tw : 16, want = W < 16 ? W : 16;
            while (run < want) {
                score -= 0.03;
                run *= 2;
            }
            if (score > best_score + 1e-9) {
                best_score = score;
                best = Geom{tw, th, nb};
            }
        }
    return best;
}

struct DevInfo {
    bool init = false;
    int sms = 0;
    int max_ctas[3][2] = {};   // [BN index][cluster - 1]
    int wg_slots[3][2] = {};   // wgrad kernel: resident CTAs / clusters, 0 = not set up yet
};
DevInfo g_dev[16];

template <int BN, int CL>
int launch_igemm(const ConvMaps& maps, ConvArgs& args, DevInfo& di, int bn_idx, cudaStream_t stream) {
    auto kern = conv_igemm_kernel<BN, CL>;
    int& max_ctas = di.max_ctas[bn_idx][CL - 1];
    if (max_ctas == 0) {
        TFOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       CCfg<BN, CL>::kSmemBytes));
        max_ctas = di.sms;
        if (CL == 2) {
            cudaLaunchConfig_t q = {};
            q.gridDim = dim3(di.sms / 2 * 2);
            q.blockDim = dim3(kConvThreads);
            q.dynamicSmemBytes = CCfg<BN, CL>::kSmemBytes;
            cudaLaunchAttribute qa[1];
            qa[0].id = cudaLaunchAttributeClusterDimension;
            qa[0].val.clusterDim.x = 2;
            qa[0].val.clusterDim.y = 1;
            qa[0].val.clusterDim.z = 1;
            q.attrs = qa;
            q.numAttrs = 1;
            int n = 0;
            TFOP_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &q));
            TFOP_REQUIRE(n > 0, "no 2-CTA cluster of the convolution kernel fits on this device");
            max_ctas = 2 * n;
        }
    }
    const int m_blocks = args.wblk * args.hblk * args.nblk;
    const int m_units = (m_blocks + CL - 1) / CL;
    const int tiles = m_units * ((args.ncols + BN - 1) / BN);
    args.div_nblocks.set((args.ncols + BN - 1) / BN);
    int grid = tiles * CL;
    if (grid > max_ctas) grid = max_ctas;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kConvThreads);
    cfg.dynamicSmemBytes = CCfg<BN, CL>::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    TFOP_CUDA(cudaLaunchKernelEx(&cfg, kern, maps, args));
    return 0;
}

int g_conv_cluster_mode = -1;  // -1 auto, 0 single CTAs, 1 pairs wherever the tile allows

int get_dev(DevInfo** out) {
    int dev = 0;
    TFOP_CUDA(cudaGetDevice(&dev));
    TFOP_REQUIRE(dev >= 0 && dev < 16, "device index out of range");
    DevInfo& di = g_dev[dev];
    if (!di.init) {
        TFOP_CUDA(cudaDeviceGetAttribute(&di.sms, cudaDevAttrMultiProcessorCount, dev));
        di.init = true;
    }
    *out = &di;
    return 0;
}

// Shared by forward and dgrad: output pixel grid (N, OHt, OWt), `ncols` output channels,
// activation maps / taps already in `maps` / `args`.
int run_igemm(ConvMaps& maps, ConvArgs& args, const void* w, int w_rows, int w_cols,
              cudaStream_t stream) {
    DevInfo* di;
    if (get_dev(&di) != 0) return -1;
    const int ncols = args.ncols;
    const int bn = ncols >= 256 ? 256 : (ncols >= 128 ? 128 : 64);
    const int m_blocks = args.wblk * args.hblk * args.nblk;
    // pairs need two pixel tiles and at least a 64-row (MN-major: 64-column) half of B per CTA
    bool pair = m_blocks >= 2 && bn >= 128;
    if (g_conv_cluster_mode == 0) pair = false;
    const int box_rows = args.b_mn ? 64 : (pair ? bn / 2 : bn);
    if (weight_map(&maps.b, w, w_rows, w_cols, box_rows) != 0) return -1;
    if (bn == 256)
        return pair ? launch_igemm<256, 2>(maps, args, *di, 2, stream)
                    : launch_igemm<256, 1>(maps, args, *di, 2, stream);
    if (bn == 128)
        return pair ? launch_igemm<128, 2>(maps, args, *di, 1, stream)
                    : launch_igemm<128, 1>(maps, args, *di, 1, stream);
    return launch_igemm<64, 1>(maps, args, *di, 0, stream);
}

void set_geom(ConvArgs& a, int W, int H, int N) {
    const Geom g = pick_geom(W, H, N, 128);
    a.tw = g.tw;
    a.th = g.th;
    a.nb = g.nb;
    a.wblk = (W + g.tw - 1) / g.tw;
    a.hblk = (H + g.th - 1) / g.th;
    a.nblk = (N + g.nb - 1) / g.nb;
    a.div_wblk.set(a.wblk);
    a.div_hblk.set(a.hblk);
    a.sw = g.tw < 32 ? g.tw : 32;
    a.sh = g.th < 32 / a.sw ? g.th : 32 / a.sw;
    a.W = W;
    a.H = H;
    a.N = N;
}

int check_conv_args(const void* a, const void* b, const void* c, int N, int H, int W, int Cin,
                    int Cout, int ksize, int stride) {
    TFOP_REQUIRE(N > 0 && H > 0 && W > 0, "conv: empty problem");
    TFOP_REQUIRE(ksize == 1 || ksize == 3, "conv: kernel size must be 1 or 3");
    TFOP_REQUIRE(stride == 1 || stride == 2, "conv: stride must be 1 or 2");
    TFOP_REQUIRE(Cin % 64 == 0 && Cout % 64 == 0 && Cin > 0 && Cout > 0,
                 "conv: channel counts must be multiples of 64");
    TFOP_REQUIRE(((((uintptr_t)a) | ((uintptr_t)b) | ((uintptr_t)c)) & 15) == 0,
                 "conv: tensors must be 16-byte aligned");
    return 0;
}

}  // namespace

// Test / benchmark hook: -1 auto, 0 never pair CTAs, 1 pair wherever possible.
TFOP_API int tfop_conv_set_cluster_mode(int mode) {
    TFOP_REQUIRE(mode >= -1 && mode <= 1, "cluster mode must be -1, 0 or 1");
    g_conv_cluster_mode = mode;
    return 0;
}

// y[N, OH, OW, Cout] = conv(x[N, H, W, Cin], w[Cout, k, k, Cin]), padding k / 2, bf16 NHWC / KRSC.
// stats: null, or [2, Cout] fp32 to which the per-channel sum and sum of squares of y are ADDED.
TFOP_API int tfop_conv_fwd(const void* x, const void* w, void* y, float* stats, int N, int H,
                           int W, int Cin, int Cout, int ksize, int stride, cudaStream_t stream) {
    if (check_conv_args(x, w, y, N, H, W, Cin, Cout, ksize, stride) != 0) return -1;
    if (get_enc() != 0) return -1;
    const int pad = ksize / 2;
    const int OH = (H + 2 * pad - ksize) / stride + 1, OW = (W + 2 * pad - ksize) / stride + 1;
    ConvMaps maps;
    ConvArgs a = {};
    set_geom(a, OW, OH, N);
    a.cblocks = Cin / 64;
    a.ncols = Cout;
    a.stats = stats;
    bool used[4] = {false, false, false, false};
    for (int kh = 0; kh < ksize; ++kh)
        for (int kw = 0; kw < ksize; ++kw) {
            const int t = a.n_taps++;
            const int eh = kh - pad, ew = kw - pad;
            if (stride == 1) {
                a.tap_map[t] = 0;
                a.tap_dh[t] = eh;
                a.tap_dw[t] = ew;
            } else {
                const int pr = eh & 1, pc = ew & 1;
                a.tap_map[t] = pr * 2 + pc;
                a.tap_dh[t] = (eh - pr) / 2;
                a.tap_dw[t] = (ew - pc) / 2;
            }
            used[a.tap_map[t]] = true;
            a.tap_bcol[t] = (kh * ksize + kw) * Cin;
        }
    for (int i = 0; i < 4; ++i) {
        const int m = used[i] ? i : 0;  // unused slots alias a valid map (they are only prefetched)
        if (stride == 1 ? act_map(&maps.a[i], x, N, H, W, Cin, 1, 0, 0, a.tw, a.th, a.nb, "x")
                        : act_map(&maps.a[i], x, N, H, W, Cin, 2, m >> 1, m & 1, a.tw, a.th, a.nb,
                                  "x (parity)"))
            return -1;
    }
    if (act_map(&maps.c, y, N, OH, OW, Cout, 1, 0, 0, a.sw, a.sh, 32 / (a.sw * a.sh), "y")) return -1;
    return run_igemm(maps, a, w, Cout, ksize * ksize * Cin, stream);
}

// dx[N, H, W, Cin] (+)= conv_transpose(dy[N, OH, OW, Cout], w[Cout, k, k, Cin]): the data gradient
// of tfop_conv_fwd. accumulate != 0 adds into dx (bf16 add at L2) instead of overwriting it. A
// 1x1 stride-2 gradient only touches the even pixels: without `accumulate` dx is zeroed first.
TFOP_API int tfop_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin,
                             int Cout, int ksize, int stride, int accumulate, cudaStream_t stream) {
    if (check_conv_args(dy, w, dx, N, H, W, Cin, Cout, ksize, stride) != 0) return -1;
    if (get_enc() != 0) return -1;
    const int pad = ksize / 2;
    const int OH = (H + 2 * pad - ksize) / stride + 1, OW = (W + 2 * pad - ksize) / stride + 1;
    if (stride == 2 && ksize == 1 && !accumulate)
        TFOP_CUDA(cudaMemsetAsync(dx, 0, (size_t)N * H * W * Cin * 2, stream));
    for (int p = 0; p < stride; ++p)
        for (int q = 0; q < stride; ++q) {
            const int Hp = (H - p + stride - 1) / stride, Wq = (W - q + stride - 1) / stride;
            if (Hp <= 0 || Wq <= 0) continue;
            ConvMaps maps;
            ConvArgs a = {};
            set_geom(a, Wq, Hp, N);
            a.cblocks = Cout / 64;
            a.ncols = Cin;
            a.b_mn = 1;
            a.accumulate = accumulate;
            for (int kh = 0; kh < ksize; ++kh) {
                if ((p + pad - kh) % stride != 0) continue;
                for (int kw = 0; kw < ksize; ++kw) {
                    if ((q + pad - kw) % stride != 0) continue;
                    const int t = a.n_taps++;
                    a.tap_map[t] = 0;
                    a.tap_dh[t] = (p + pad - kh) / stride;
                    a.tap_dw[t] = (q + pad - kw) / stride;
                    a.tap_bcol[t] = (kh * ksize + kw) * Cin;
                }
            }
            if (a.n_taps == 0) continue;  // 1x1 stride 2: odd pixels get no gradient
            for (int i = 0; i < 4; ++i)
                if (act_map(&maps.a[i], dy, N, OH, OW, Cout, 1, 0, 0, a.tw, a.th, a.nb, "dy"))
                    return -1;
            if (act_map(&maps.c, dx, N, H, W, Cin, stride, p, q, a.sw, a.sh, 32 / (a.sw * a.sh),
                        "dx"))
                return -1;
            if (run_igemm(maps, a, w, Cout, ksize * ksize * Cin, stream) != 0) return -1;
        }
    return 0;
}

// dw[Cout, k, k, Cin] (fp32) += weight gradient of tfop_conv_fwd for x[N, H, W, Cin] and
// dy[N, OH, OW, Cout]. The caller zeroes dw (partial sums of all CTAs are added with RED).
TFOP_API int tfop_conv_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W,
                             int Cin, int Cout, int ksize, int stride, cudaStream_t stream) {
    if (check_conv_args(x, dy, dw, N, H, W, Cin, Cout, ksize, stride) != 0) return -1;
    if (get_enc() != 0) return -1;
    DevInfo* di;
    if (get_dev(&di) != 0) return -1;
    const int pad = ksize / 2;
    const int OH = (H + 2 * pad - ksize) / stride + 1, OW = (W + 2 * pad - ksize) / stride + 1;
    WgradMaps maps;
    WgradArgs a = {};
    const Geom g = pick_geom(OW, OH, N, 64);
    a.pw = g.tw;
    a.ph = g.th;
    a.pn = g.nb;
    a.wblk = (OW + g.tw - 1) / g.tw;
    a.hblk = (OH + g.th - 1) / g.th;
    a.nblk = (N + g.nb - 1) / g.nb;
    a.cblocks = Cin / 64;
    a.Cin = Cin;
    a.Cout = Cout;
    bool used[4] = {false, false, false, false};
    for (int kh = 0; kh < ksize; ++kh)
        for (int kw = 0; kw < ksize; ++kw) {
            const int t = a.n_taps++;
            const int eh = kh - pad, ew = kw - pad;
            if (stride == 1) {
                a.tap_map[t] = 0;
                a.tap_dh[t] = eh;
                a.tap_dw[t] = ew;
            } else {
                const int pr = eh & 1, pc = ew & 1;
                a.tap_map[t] = pr * 2 + pc;
                a.tap_dh[t] = (eh - pr) / 2;
                a.tap_dw[t] = (ew - pc) / 2;
            }
            used[a.tap_map[t]] = true;
        }
    for (int i = 0; i < 4; ++i) {
        const int m = used[i] ? i : 0;
        if (stride == 1 ? act_map(&maps.x[i], x, N, H, W, Cin, 1, 0, 0, a.pw, a.ph, a.pn, "x")
                        : act_map(&maps.x[i], x, N, H, W, Cin, 2, m >> 1, m & 1, a.pw, a.ph, a.pn,
                                  "x (parity)"))
            return -1;
    }
    if (act_map(&maps.dy, dy, N, OH, OW, Cout, 1, 0, 0, a.pw, a.ph, a.pn, "dy")) return -1;
    {   // dW as a 2-D fp32 matrix [Cout, taps * Cin]; unswizzled 32 x 32 boxes
        const int cols = ksize * ksize * Cin;
        cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)Cout};
        cuuint64_t strides[1] = {(cuuint64_t)cols * 4};
        cuuint32_t box[2] = {32, 32};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = g_enc(&maps.dw, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, dw, dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                           CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            tfop_set_error(__FILE__, __LINE__, "cuTensorMapEncodeTiled(dw) failed");
            return -1;
        }
    }
    const int bn = Cout >= 256 ? 256 : (Cout >= 128 ? 128 : 64);
    const int slices = a.n_taps * a.cblocks;
    const int ms = bn == 64 ? 4 : 2;                 // slices per CTA
    const int m_tiles = (slices + ms - 1) / ms;
    const bool pair = bn >= 128 && m_tiles >= 2 && g_conv_cluster_mode != 0;
    const int cl = pair ? 2 : 1;
    a.m_units = (m_tiles + cl - 1) / cl;
    a.n_tiles = (Cout + bn - 1) / bn;
    a.div_wblk.set(a.wblk);
    a.div_hblk.set(a.hblk);
    const int tiles = a.m_units * a.n_tiles;
    const int pblocks = a.wblk * a.hblk * a.nblk;
    const int bi = bn == 256 ? 2 : (bn == 128 ? 1 : 0);
    int& slots = di->wg_slots[bi][cl - 1];
#define TFOP_WG_PREP(BN_, CL_)                                                                    \
    do {                                                                                          \
        TFOP_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<BN_, CL_>,                               \
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,               \
                                       WCfg<BN_, CL_>::kSmemBytes));                              \
        slots = di->sms;                                                                          \
        if (CL_ == 2) {                                                                           \
            cudaLaunchConfig_t q = {};                                                            \
            q.gridDim = dim3(di->sms / 2 * 2);                                                    \
            q.blockDim = dim3(kNumThreads);                                                       \
            q.dynamicSmemBytes = WCfg<BN_, CL_>::kSmemBytes;                                      \
            cudaLaunchAttribute qa[1];                                                            \
            qa[0].id = cudaLaunchAttributeClusterDimension;                                       \
            qa[0].val.clusterDim.x = 2;                                                           \
            qa[0].val.clusterDim.y = 1;                                                           \
            qa[0].val.clusterDim.z = 1;                                                           \
            q.attrs = qa;                                                                         \
            q.numAttrs = 1;                                                                       \
            int n = 0;                                                                            \
            TFOP_CUDA(cudaOccupancyMaxActiveClusters(&n, conv_wgrad_kernel<BN_, CL_>, &q));       \
            TFOP_REQUIRE(n > 0, "no 2-CTA cluster of the wgrad kernel fits on this device");      \
            slots = n;                                                                            \
        }                                                                                         \
    } while (0)
    if (slots == 0) {
        if (bn == 256 && pair) TFOP_WG_PREP(256, 2);
        else if (bn == 256) TFOP_WG_PREP(256, 1);
        else if (bn == 128 && pair) TFOP_WG_PREP(128, 2);
        else if (bn == 128) TFOP_WG_PREP(128, 1);
        else TFOP_WG_PREP(64, 1);
    }
#undef TFOP_WG_PREP
    // Pixel ranges per tile: every unit pays ~8 k-blocks' worth of set-up, pipeline fill and
    // epilogue, and the units run in waves of `slots`: take the split count with the least
    // (waves x (k-blocks per unit + 8)).
    int splits = 1;
    {
        long best = -1;
        const int smax = pblocks < 4 * slots ? pblocks : 4 * slots;
        for (int sp = 1; sp <= smax; ++sp) {
            const long waves = ((long)tiles * sp + slots - 1) / slots;
            const long cost = waves * ((pblocks + sp - 1) / sp + 8);
            if (best < 0 || cost < best) {
                best = cost;
                splits = sp;
            }
        }
    }
    a.splits = splits;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(tiles * splits * cl);
    cfg.blockDim = dim3(kNumThreads);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cl;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
#define TFOP_WG_LAUNCH(BN_, CL_)                                                                  \
    do {                                                                                          \
        cfg.dynamicSmemBytes = WCfg<BN_, CL_>::kSmemBytes;                                        \
        TFOP_CUDA(cudaLaunchKernelEx(&cfg, conv_wgrad_kernel<BN_, CL_>, maps, a));                \
    } while (0)
    if (bn == 256 && pair) TFOP_WG_LAUNCH(256, 2);
    else if (bn == 256) TFOP_WG_LAUNCH(256, 1);
    else if (bn == 128 && pair) TFOP_WG_LAUNCH(128, 2);
    else if (bn == 128) TFOP_WG_LAUNCH(128, 1);
    else TFOP_WG_LAUNCH(64, 1);
#undef TFOP_WG_LAUNCH
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
