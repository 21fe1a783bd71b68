// K7: bf16 GEMM on the 5th-gen tensor cores, written directly against the sm_100a ISA.
//
//   C[M,N] = act(op(A) * op(B)^T + bias[N]),  fp32 accumulation in TMEM
//   op(X) = X stored K-major ([rows, K]) or MN-major ([K, rows]) - so F.linear, its dgrad and its
//   wgrad all read the tensors as they sit in memory (no transposed copies).
//
// Structure (persistent, warp-specialised, one CTA per SM, 256 threads):
//   warp 0    TMA producer: cp.async.bulk.tensor into 128B-swizzled shared-memory stages
//   warp 1    MMA issuer: one elected thread, tcgen05.mma.cta_group::2 - CTAs run as pairs
//             (2-CTA clusters) on vertically adjacent tiles, one MMA covers 256 x BN, each SM
//             stages only half of the B tile (6 stages of 32 KB at BN = 256)
//   warp 2    TMEM allocator (2 accumulator stages: epilogue of tile i overlaps MMAs of tile i+1)
//   warps 4-7 epilogue: tcgen05.ld (double buffered) -> bias / GeLU / ReLU from a shared-memory
//             bias slice -> bf16 / fp32 -> swizzled 32 x 128 B boxes -> TMA tensor stores
//             (reduce-add stores under split-K)
// Tile order: bands of m-blocks sized so the band's A panel (~32 MB) stays in L2.
// Split-K: when the tiles alone fill < 75 % of a wave (weight-gradient shapes).
// Measured: 1.63 PFLOP/s at 8192^3 (cuBLAS 1.70), tensor pipe 94 % busy at the sustained clock;
// history and ncu numbers in profiles/r1_gemm_2cta_analysis.md.
//
// Reference parity: the dense layers of the example trainers
// (examples/v1/dist-mnist/dist_mnist.py:188-192 matmuls -> cuBLAS/Eigen in the reference stack);
// BASELINE.json requires the GEMMs on the named paths to be tcgen05/TMEM/TMA kernels rather than
// library calls.
#include "common.cuh"
#include "gemm_common.cuh"

namespace {

using namespace tfop_gemm;

enum { EPI_NONE = 0, EPI_BIAS = 1, EPI_BIAS_GELU = 2, EPI_BIAS_RELU = 3, EPI_BIAS_GELU_TANH = 4 };

// bias + activation on 32 consecutive accumulator columns; sb = the fp32 bias of those columns
// in shared memory (every lane reads the same address: one broadcast LDS.128 per 4 columns).
// Loading bias from global memory per element put ~256 dependent L1 round trips into every
// tile's epilogue and made the K=1024 forward GEMMs 3-4x slower than the bias-free ones.
__device__ __forceinline__ void apply_epilogue(float* f, const float* sb, int epi) {
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
        const float4 b4 = *reinterpret_cast<const float4*>(sb + j);
        f[j] += b4.x;
        f[j + 1] += b4.y;
        f[j + 2] += b4.z;
        f[j + 3] += b4.w;
    }
    if (epi == EPI_BIAS_GELU) {
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = gelu_erf(f[j]);
    } else if (epi == EPI_BIAS_GELU_TANH) {
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = gelu_tanh(f[j]);
    } else if (epi == EPI_BIAS_RELU) {
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
    }
}

// The 128 epilogue threads copy the tile's BN bias values (zero past N) into shared memory.
template <int BN>
__device__ __forceinline__ void stage_bias(float* sbias, const void* bias, int bias_dtype, int n0,
                                           int N, int epi_tid) {
    epi_bar_sync();  // the previous tile's readers are done
    for (int i = epi_tid; i < BN; i += 128) {
        const int col = n0 + i;
        float v = 0.f;
        if (col < N)
            v = bias_dtype == TFOP_F32 ? __ldg((const float*)bias + col)
                                       : __bfloat162float(((const __nv_bfloat16*)bias)[col]);
        sbias[i] = v;
    }
    epi_bar_sync();
}

// kTmaStore: the epilogue converts into 128B-swizzled shared-memory boxes (32 rows x 128 bytes
// per warp, double buffered) and one lane issues a TMA tensor store per box, so C is written in
// whole 128-byte lines and the next tcgen05.ld is in flight while the current chunk is
// converted. The direct-store path stays for C layouts TMA cannot describe (ldc*elt % 16 != 0).
//
// kCluster = 2: CTAs run as cta_group::2 pairs on vertically adjacent tiles (same n-block). One
// tcgen05.mma covers 256 x BN: each SM multiplies its own 128 rows of A with the whole B tile,
// of which it stages only its BN/2 rows - the tensor cores fetch the other half from the peer.
// Per k-block a CTA writes and reads 32 KB of shared memory instead of 48 KB (BN = 256), which
// is what the single-CTA kernel is bound by (multicasting B alone changed nothing). Only the
// even CTA issues MMAs; TMA loads of both CTAs count on its `full` barrier, its commits are
// multicast to both CTAs' `empty` / `tmem_full` barriers, and both epilogues arrive on its
// `tmem_empty`.
// Register cap 136 (not the 255 that __launch_bounds__(256, 1) would allow): the CTA then holds
// 34 K of the SM's 64 K registers, which leaves room for a 512-thread block of the all-reduce /
// pack kernels to be resident NEXT TO a GEMM CTA - with 241 registers the gradient all-reduce of
// a data-parallel step could only run between GEMMs (+1.1 ms per BERT-large step at 2 and at 8
// GPUs alike).
//
// kFuse (gemm_allreduce.cu, NVLS flavour): C is this rank's copy of a multicast-bound region. An
// epilogue warp that has stored its 32 rows of a tile waits for the stores to land, then bumps
// the tile's arrival counter on the OWNER rank (tile id % world) with a release RED over NVLink.
// Warps 8-11 take this rank's owned tiles as they become complete on every rank: the NVSwitch
// sums the ranks' copies (multimem.ld_reduce), the result goes to every rank (multimem.st), and
// each rank's `delivered` counter is bumped. The kernel ends when this rank's whole C is in.
template <int BN, bool kTmaStore, int kCluster, bool kFuse = false>
__global__ void __maxnreg__(136)
gemm_tn_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const __grid_constant__ CUtensorMap map_c, void* __restrict__ C, const void* __restrict__ bias, int bias_dtype, int M, int N,
               int K, int ldc, int out_dtype, int epi, int a_mn, int b_mn, int group_m, int splits,
               const GemmFuse fz) {
    using cfg = Cfg<BN, kCluster>;
    constexpr bool kPair = kCluster == 2;
    static_assert(!kFuse || kTmaStore, "the fused mode stores its tiles with TMA");
    extern __shared__ uint8_t smem_raw[];
    // 128B swizzle atoms need 1024-byte aligned tiles
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + cfg::kStages * cfg::kABytes;
    uint8_t* smem_store = smem + cfg::kStages * cfg::kStageBytes;
    uint64_t* bars = (uint64_t*)(smem_store + cfg::kStoreBytes);
    float* sbias = (float*)(smem_store + cfg::kStoreBytes + cfg::kBarBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + cfg::kStages;
    uint64_t* tmem_full = bars + 2 * cfg::kStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_slot = (uint32_t*)(tmem_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m_blocks = (M + BLOCK_M - 1) / BLOCK_M;
    const int n_blocks = (N + BN - 1) / BN;
    const int k_blocks = (K + BLOCK_K - 1) / BLOCK_K;
    // scheduling unit: one m-block, or a vertical pair of m-blocks per cluster
    const int crank = kCluster == 2 ? (int)cluster_ctarank() : 0;
    const int m_units = (m_blocks + kCluster - 1) / kCluster;
    // split-K: a scheduling unit is (tile, K range); the `splits` units of a tile are adjacent
    const int num_tiles = m_units * n_blocks * splits;
    const int t_first = blockIdx.x / kCluster;
    const int t_step = gridDim.x / kCluster;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        if (kTmaStore) asm volatile("prefetch.tensormap [%0];" ::"l"(&map_c) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < cfg::kStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], 4 * kCluster);  // one arrive per epilogue warp (of the pair)
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
    // the peer's barriers must be initialised before anything of ours can arrive on them
    if (kCluster == 2) cluster_sync_all();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // fused mode: this call's generation (CTA 0 writes it back when the output is complete)
    uint32_t fz_gen = 0;
    if (kFuse) fz_gen = *(fz.arrived[fz.rank] + fz.tiles_per_owner + 1) + 1;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = t_first; t < num_tiles; t += t_step) {
                int mb, nb;
                tile_coords_rt(t / splits, m_units, n_blocks, group_m, mb, nb);
                const int m0 = (mb * kCluster + crank) * BLOCK_M;
                const int n0 = nb * BN;
                const int sp = t % splits;
                const int kb0 = (int)((long)k_blocks * sp / splits);
                const int kb1 = (int)((long)k_blocks * (sp + 1) / splits);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t* sa = smem_a + stage * cfg::kABytes;
                    uint8_t* sb = smem_b + stage * cfg::kBBytes;
                    if (kPair) {
                        // the leader's barrier counts the bytes of both CTAs
                        if (crank == 0) mbar_expect_tx(&full[stage], 2 * cfg::kStageBytes);
                        if (!a_mn) {
                            tma_load_2d_pair(&map_a, &full[stage], sa, kb * BLOCK_K, m0);
                        } else {
#pragma unroll
                            for (int j = 0; j < BLOCK_M / 64; ++j)
                                tma_load_2d_pair(&map_a, &full[stage], sa + j * kMnBoxBytes,
                                                 m0 + 64 * j, kb * BLOCK_K);
                        }
                        // this CTA's half of the B tile
                        const int nh = n0 + crank * (BN / 2);
                        if (!b_mn) {
                            tma_load_2d_pair(&map_b, &full[stage], sb, kb * BLOCK_K, nh);
                        } else {
#pragma unroll
                            for (int j = 0; j < BN / 128; ++j)
                                tma_load_2d_pair(&map_b, &full[stage], sb + j * kMnBoxBytes,
                                                 nh + 64 * j, kb * BLOCK_K);
                        }
                    } else {
                        mbar_expect_tx(&full[stage], cfg::kStageBytes);
                        if (!a_mn) {
                            tma_load_2d(&map_a, &full[stage], sa, kb * BLOCK_K, m0);
                        } else {
                            // A stored [K, M]: boxes of 64 k-rows x 64 m-elements
#pragma unroll
                            for (int j = 0; j < BLOCK_M / 64; ++j)
                                tma_load_2d(&map_a, &full[stage], sa + j * kMnBoxBytes,
                                            m0 + 64 * j, kb * BLOCK_K);
                        }
                        if (!b_mn) {
                            tma_load_2d(&map_b, &full[stage], sb, kb * BLOCK_K, n0);
                        } else {
#pragma unroll
                            for (int j = 0; j < BN / 64; ++j)
                                tma_load_2d(&map_b, &full[stage], sb + j * kMnBoxBytes,
                                            n0 + 64 * j, kb * BLOCK_K);
                        }
                    }
                    if (++stage == cfg::kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (the even CTA of a pair) =====================
        if ((!kPair || crank == 0) && elect_one()) {
            const uint32_t idesc = make_idesc_major(BLOCK_M * kCluster, BN, a_mn, b_mn);
            // per-UMMA_K start-address step: 32 bytes along a K-major swizzle row, 16 k-rows
            // (2048 bytes) of an MN-major box
            const uint32_t a_step = a_mn ? 2048u : (uint32_t)(UMMA_K * 2);
            const uint32_t b_step = b_mn ? 2048u : (uint32_t)(UMMA_K * 2);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int t = t_first; t < num_tiles; t += t_step) {
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
                const int sp = t % splits;
                const int kb0 = (int)((long)k_blocks * sp / splits);
                const int kb1 = (int)((long)k_blocks * (sp + 1) / splits);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tcgen05_fence_after();
                    const uint32_t a_addr = smem_u32(smem_a + stage * cfg::kABytes);
                    const uint32_t b_addr = smem_u32(smem_b + stage * cfg::kBBytes);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        const uint64_t adesc = a_mn ? make_smem_desc_mn(a_addr + k * a_step)
                                                    : make_smem_desc(a_addr + k * a_step);
                        const uint64_t bdesc = b_mn ? make_smem_desc_mn(b_addr + k * b_step)
                                                    : make_smem_desc(b_addr + k * b_step);
                        if (kPair)
                            umma_bf16_pair(tmem_d, adesc, bdesc, idesc, (kb != kb0 || k != 0) ? 1u : 0u);
                        else
                            umma_bf16(tmem_d, adesc, bdesc, idesc, (kb != kb0 || k != 0) ? 1u : 0u);
                    }
                    // frees the smem slot when the MMAs retire (in both CTAs of a pair)
                    if (kPair)
                        umma_commit_pair(&empty[stage]);
                    else
                        umma_commit(&empty[stage]);
                    if (++stage == cfg::kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                // accumulator ready for the epilogue (of both CTAs of a pair)
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
    } else if (warp >= kEpiWarp0 && warp < kEpiWarp0 + 4) {
        // ===================== epilogue =====================
        const int ew = warp - kEpiWarp0;  // == warp % 4: TMEM lane quadrant
        int acc = 0;
        uint32_t acc_phase = 0;
        if (kTmaStore) {
            // chunk = 128 bytes of one output row: 64 bf16 or 32 fp32 columns
            const int chunk_cols = out_dtype == TFOP_BF16 ? 64 : 32;
            uint8_t* box0 = smem_store + ew * 2 * cfg::kStoreBoxBytes;
            const uint32_t row_off = (uint32_t)lane * 128u;
            const uint32_t sw = (uint32_t)(lane & 7);
            int buf = 0;
            for (int t = t_first; t < num_tiles; t += t_step) {
                int mb, nb;
                tile_coords_rt(t / splits, m_units, n_blocks, group_m, mb, nb);
                const int m0 = (mb * kCluster + crank) * BLOCK_M;
                const int n0 = nb * BN;
                if (epi != EPI_NONE)
                    stage_bias<BN>(sbias, bias, bias_dtype, n0, N, ew * 32 + lane);
                mbar_wait(&tmem_full[acc], acc_phase);
                tcgen05_fence_after();
                const uint32_t taddr =
                    tmem_base + (uint32_t)(acc * BN) + ((uint32_t)(ew * 32) << 16);
                uint32_t ra[32], rb[32];
                tmem_ld_32x32b_x32(taddr, ra);
#pragma unroll 1
                for (int c = 0; c < BN; c += 64) {
                    // ---- columns c .. c+31 (in ra); prefetch c+32 .. c+63 into rb
                    tmem_ld_wait();
                    tmem_ld_32x32b_x32(taddr + (uint32_t)(c + 32), rb);
                    uint8_t* box = box0 + buf * cfg::kStoreBoxBytes;
                    // the store that last read this box must have drained it
                    if (lane == 0) bulk_wait_read<1>();
                    __syncwarp();
                    float f[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(ra[j]);
                    if (epi != EPI_NONE) apply_epilogue(f, sbias + c, epi);
                    if (out_dtype == TFOP_BF16) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            bf16x8 pk = f32_to_bf16x8(f + 8 * j);
                            *reinterpret_cast<uint4*>(box + row_off + (((uint32_t)j ^ sw) << 4)) =
                                *reinterpret_cast<uint4*>(&pk);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            *reinterpret_cast<float4*>(box + row_off + (((uint32_t)j ^ sw) << 4)) =
                                make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
                        fence_async_smem();
                        __syncwarp();
                        if (lane == 0) {
                            if (n0 + c < N && m0 + ew * 32 < M) {
                                if (splits > 1)
                                    tma_reduce_add_2d(&map_c, box, n0 + c, m0 + ew * 32);
                                else
                                    tma_store_2d(&map_c, box, n0 + c, m0 + ew * 32);
                            }
                            bulk_commit();
                        }
                        buf ^= 1;
                        box = box0 + buf * cfg::kStoreBoxBytes;
                        if (lane == 0) bulk_wait_read<1>();
                        __syncwarp();
                    }
                    // ---- columns c+32 .. c+63 (in rb); prefetch the next chunk into ra
                    tmem_ld_wait();
                    if (c + 64 < BN) tmem_ld_32x32b_x32(taddr + (uint32_t)(c + 64), ra);
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(rb[j]);
                    if (epi != EPI_NONE) apply_epilogue(f, sbias + c + 32, epi);
                    if (out_dtype == TFOP_BF16) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            bf16x8 pk = f32_to_bf16x8(f + 8 * j);
                            *reinterpret_cast<uint4*>(box + row_off +
                                                      (((uint32_t)(j + 4) ^ sw) << 4)) =
                                *reinterpret_cast<uint4*>(&pk);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            *reinterpret_cast<float4*>(box + row_off + (((uint32_t)j ^ sw) << 4)) =
                                make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
                    }
                    fence_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        const int cc = out_dtype == TFOP_BF16 ? n0 + c : n0 + c + 32;
                        if (cc < N && m0 + ew * 32 < M) {
                            if (splits > 1)
                                tma_reduce_add_2d(&map_c, box, cc, m0 + ew * 32);
                            else
                                tma_store_2d(&map_c, box, cc, m0 + ew * 32);
                        }
                        bulk_commit();
                    }
                    buf ^= 1;
                }
                // every tcgen05.ld of this accumulator has completed: hand it back to the MMA warp
                tcgen05_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (kPair)
                        mbar_arrive_leader(&tmem_empty[acc]);
                    else
                        mbar_arrive(&tmem_empty[acc]);
                }
                if (kFuse) {
                    // my 32 rows of the tile are in this GPU's copy: tell the tile's owner
                    if (lane == 0 && m0 < M) {
                        bulk_wait_all();
                        asm volatile("fence.proxy.async;" ::: "memory");
                        fence_acq_rel_sys();
                        // owner and slot follow the PRODUCTION order (the same on every rank):
                        // the reducers then meet their tiles in the order they complete
                        const int q = t * kCluster + crank;
                        fuse_red_release_sys(fz.arrived[q % fz.world] + q / fz.world, 1u);
                    }
                    __syncwarp();
                }
                if (++acc == 2) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
            (void)chunk_cols;
            if (lane == 0) bulk_wait_read<0>();  // smem must outlive the last store's read
            __syncwarp();
        } else {
        for (int t = t_first; t < num_tiles; t += t_step) {
            int mb, nb;
            tile_coords_rt(t / splits, m_units, n_blocks, group_m, mb, nb);
            const int m0 = (mb * kCluster + crank) * BLOCK_M;
            const int n0 = nb * BN;
            if (epi != EPI_NONE) stage_bias<BN>(sbias, bias, bias_dtype, n0, N, ew * 32 + lane);
            mbar_wait(&tmem_full[acc], acc_phase);
            tcgen05_fence_after();
            const int row = m0 + ew * 32 + lane;
            const uint32_t taddr = tmem_base + (uint32_t)(acc * BN) + ((uint32_t)(ew * 32) << 16);
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(taddr + (uint32_t)c, r);
                tmem_ld_wait();
                const int col0 = n0 + c;
                if (row < M && col0 < N) {
                    float f[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(r[j]);
                    if (epi != EPI_NONE) apply_epilogue(f, sbias + c, epi);
                    if (out_dtype == TFOP_BF16) {
                        __nv_bfloat16* crow = (__nv_bfloat16*)C + (size_t)row * ldc + col0;
                        if (col0 + 32 <= N && (((uintptr_t)crow) & 15) == 0) {
#pragma unroll
                            for (int j = 0; j < 32; j += 8) {
                                bf16x8 pk = f32_to_bf16x8(f + j);
                                *reinterpret_cast<uint4*>(crow + j) = *reinterpret_cast<uint4*>(&pk);
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < N) crow[j] = __float2bfloat16_rn(f[j]);
                        }
                    } else {
                        float* crow = (float*)C + (size_t)row * ldc + col0;
                        if (col0 + 32 <= N && (((uintptr_t)crow) & 15) == 0) {
#pragma unroll
                            for (int j = 0; j < 32; j += 4)
                                *reinterpret_cast<float4*>(crow + j) =
                                    make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < N) crow[j] = f[j];
                        }
                    }
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
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1;
            }
        }
        }
    }

    if (kFuse && warp >= kFuseRedWarp0) {
        // ===================== reducers (fused mode), concurrent with the GEMM warps ==========
        const int rt = threadIdx.x - kFuseRedWarp0 * 32;
        const int W = fz.world;
        const int real_tiles = m_blocks * n_blocks;
        uint32_t* my_arrived = fz.arrived[fz.rank];
        // Tiles in production order q = t * kCluster + crank; this rank owns q = j * W + rank.
        // A unit of reducer work is one 32-row slab of an owned tile, units dealt round-robin
        // over the CTAs: a few owned tiles still keep many SMs' worth of requests in flight.
        const int seq = num_tiles * kCluster;
        const int owned = (seq - fz.rank + W - 1) / W;
        bool alive = true;
        int done = 0;
        for (int unit = blockIdx.x; unit < owned * 4 && alive; unit += gridDim.x) {
            const int j = unit >> 2, slab = unit & 3;
            const int q = j * W + fz.rank;
            int mb, nb;
            tile_coords_rt(q / kCluster, m_units, n_blocks, group_m, mb, nb);
            const int r0 = (mb * kCluster + q % kCluster) * BLOCK_M + slab * 32;
            const int c0 = nb * BN;
            if ((mb * kCluster + q % kCluster) * BLOCK_M >= M) continue;   // no such tile
            if (rt == 0) {
                const uint32_t target = fz_gen * 4u * (uint32_t)W;   // 4 epilogue warps per rank
                const long long t0 = clock64();
                while ((int32_t)(ld_acquire_sys_u32(my_arrived + j) - target) < 0) {
                    if (clock64() - t0 > fz.timeout) {
                        atomicExch(fz.err, 1u);
                        break;
                    }
                }
            }
            fuse_reducer_bar();
            if (*(volatile uint32_t*)fz.err != 0u) alive = false;     // a peer never arrived
            // the whole slab in one round: 16 bytes x UN per thread in flight through the switch
            constexpr int kVecs = 32 * BN / 8;
            constexpr int UN = kVecs / 128;
            for (int v0 = rt; alive && v0 < kVecs; v0 += 128 * UN) {
                uint4 r[UN];
                size_t off[UN];
                bool ok[UN];
#pragma unroll
                for (int u = 0; u < UN; ++u) {
                    const int v = v0 + u * 128;
                    const int gr = r0 + v / (BN / 8), gc = c0 + (v % (BN / 8)) * 8;
                    ok[u] = gr < M && gc < N;
                    off[u] = ((size_t)gr * fz.ldo + gc) * 2;
                    if (ok[u]) r[u] = fuse_mm_ld_reduce_bf16x8(fz.mc_out + off[u]);
                }
#pragma unroll
                for (int u = 0; u < UN; ++u) {
                    if (!ok[u]) continue;
                    if (fz.scale != 1.f) {
                        float f[8];
                        bf16x8_to_f32(*reinterpret_cast<bf16x8*>(&r[u]), f);
#pragma unroll
                        for (int i = 0; i < 8; ++i) f[i] *= fz.scale;
                        bf16x8 pk = f32_to_bf16x8(f);
                        r[u] = *reinterpret_cast<uint4*>(&pk);
                    }
                    fuse_mm_st_16(fz.mc_out + off[u], r[u]);
                }
            }
            if (alive) ++done;
        }
        // one release per CTA, not per slab: every rank learns how many slabs this CTA delivered
        fuse_reducer_bar();
        if (done > 0 && rt < W) {
            fence_acq_rel_sys();
            fuse_red_release_sys(fz.arrived[rt] + fz.tiles_per_owner, (uint32_t)done);
        }
        if (blockIdx.x == 0 && rt == 0) {
            uint32_t* my_delivered = my_arrived + fz.tiles_per_owner;
            const uint32_t target = fz_gen * (uint32_t)real_tiles * 4u;   // slabs
            const long long t0 = clock64();
            while ((int32_t)(ld_acquire_sys_u32(my_delivered) - target) < 0) {
                if (clock64() - t0 > fz.timeout || *(volatile uint32_t*)fz.err != 0u) {
                    atomicExch(fz.err, 1u);
                    break;
                }
            }
            my_delivered[1] = fz_gen;
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    // the leader's commits still land on this CTA's barriers and its MMAs read this CTA's
    // shared memory: leave together
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

PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;

// A driver-API call (cuTensorMapEncodeTiled) needs a current context on the calling thread. An
// autograd worker thread whose first GPU work is one of these entry points may have none yet:
// cudaSetDevice on the thread's current device binds the primary context (legal under capture).
static int bind_context() {
    int dev = 0;
    TFOP_CUDA(cudaGetDevice(&dev));
    TFOP_CUDA(cudaSetDevice(dev));
    return 0;
}

int get_encode() {
    if (bind_context() != 0) return -1;
    if (g_encode) return 0;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    TFOP_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    TFOP_REQUIRE(qres == cudaDriverEntryPointSuccess && fn != nullptr,
                 "cuTensorMapEncodeTiled not available from the driver");
    g_encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    return 0;
}

// 2D bf16 row-major [rows, K] (ld elements between rows), box = [box_rows, 64] with 128B swizzle
int make_map(CUtensorMap* map, const void* base, int rows, int K, int ld, int box_rows) {
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)BLOCK_K, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                          strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char msg[96];
        snprintf(msg, sizeof(msg), "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
        tfop_set_error(__FILE__, __LINE__, msg);
        return -1;
    }
    return 0;
}

// C [M, N] row-major (ldc elements between rows) as a store target: boxes of 32 rows x 128
// bytes (64 bf16 / 32 fp32 columns), 128B swizzle to match what the epilogue writes.
int make_map_c(CUtensorMap* map, void* base, int M, int N, int ldc, int out_dtype) {
    const int elt = out_dtype == TFOP_BF16 ? 2 : 4;
    cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)M};
    cuuint64_t strides[1] = {(cuuint64_t)ldc * elt};
    cuuint32_t box[2] = {(cuuint32_t)(128 / elt), 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(map,
                          out_dtype == TFOP_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                                 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                          2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char msg[96];
        snprintf(msg, sizeof(msg), "cuTensorMapEncodeTiled (C) failed with CUresult %d", (int)r);
        tfop_set_error(__FILE__, __LINE__, msg);
        return -1;
    }
    return 0;
}

// Per-DEVICE launch state: SM count, whether the kernel instantiation's shared-memory attribute
// has been set on that device, and how many CTAs / 2-CTA clusters fit there. A process that
// drives several GPUs (one replica per GPU is the norm, but nothing forbids more) must not run
// the second device with the first one's numbers or with an unset attribute.
constexpr int kMaxDevices = 64;
int g_num_sms_dev[kMaxDevices] = {};
int g_force_direct_store = 0;

int current_device(int* dev) {
    TFOP_CUDA(cudaGetDevice(dev));
    TFOP_REQUIRE(*dev >= 0 && *dev < kMaxDevices, "device index out of range");
    return 0;
}
int num_sms(int* out) {
    int dev = 0;
    if (current_device(&dev) != 0) return -1;
    if (g_num_sms_dev[dev] == 0)
        TFOP_CUDA(cudaDeviceGetAttribute(&g_num_sms_dev[dev], cudaDevAttrMultiProcessorCount, dev));
    *out = g_num_sms_dev[dev];
    return 0;
}

int g_cluster_mode = -1;  // -1 auto, 0 off, 1 force clusters of two
int g_group_m = 0;        // 0: default band height; > 0: override (benchmark hook)
int g_splits = 0;         // 0: split-K heuristic, > 0: forced split count, -1: never split

template <int BN, bool kTmaStore, int kCluster>
int launch(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, void* C,
           const void* bias, int bias_dtype, int M, int N, int K, int ldc, int out_dtype, int epi,
           int a_mn, int b_mn, int splits_req, bool can_split, cudaStream_t stream) {
    static int max_ctas_dev[kMaxDevices] = {};   // per instantiation and device; 0 = not set up
    auto kern = gemm_tn_kernel<BN, kTmaStore, kCluster>;
    int dev = 0, g_num_sms = 0;
    if (current_device(&dev) != 0 || num_sms(&g_num_sms) != 0) return -1;
    int& max_ctas = max_ctas_dev[dev];
    if (max_ctas == 0) {
        TFOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg<BN, kCluster>::kSmemBytes));
        max_ctas = g_num_sms;
        if (kCluster == 2) {
            // how many 2-CTA clusters can be resident at once (GPCs with an odd SM count lose one)
            cudaLaunchConfig_t q = {};
            q.gridDim = dim3(g_num_sms / 2 * 2);
            q.blockDim = dim3(kNumThreads);
            q.dynamicSmemBytes = Cfg<BN, kCluster>::kSmemBytes;
            cudaLaunchAttribute qa[1];
            qa[0].id = cudaLaunchAttributeClusterDimension;
            qa[0].val.clusterDim.x = 2;
            qa[0].val.clusterDim.y = 1;
            qa[0].val.clusterDim.z = 1;
            q.attrs = qa;
            q.numAttrs = 1;
            int n = 0;
            TFOP_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &q));
            TFOP_REQUIRE(n > 0, "no 2-CTA cluster of the GEMM kernel fits on this device");
            max_ctas = 2 * n;
        }
    }
    const int m_units = ((M + BLOCK_M - 1) / BLOCK_M + kCluster - 1) / kCluster;
    const int tiles = m_units * ((N + BN - 1) / BN);
    // Split-K when the tiles alone leave much of the machine idle (small M x N, long K - the
    // weight-gradient shapes): pick the split count with the best wave efficiency, 3 % penalty
    // per extra split for the reduction traffic, at least 16 k-blocks per split. The partial
    // tiles are added into a zeroed C by TMA reduce-add stores.
    int splits = 1;
    if (can_split) {
        const int slots = max_ctas / kCluster;
        const int k_blocks = (K + BLOCK_K - 1) / BLOCK_K;
        if (splits_req > 0) {
            splits = splits_req;
        } else {
            auto eff = [&](int s) {
                const int u = tiles * s;
                return (double)u / (double)(((u + slots - 1) / slots) * slots) - 0.03 * (s - 1);
            };
            if (eff(1) < 0.75)
                for (int s = 2; s <= 4; ++s)
                    if (k_blocks / s >= 16 && eff(s) > eff(splits) + 1e-9) splits = s;
        }
        if (splits > k_blocks) splits = k_blocks;
        if (splits > 1) {
            const size_t elt = out_dtype == TFOP_BF16 ? 2 : 4;
            TFOP_CUDA(cudaMemset2DAsync(C, (size_t)ldc * elt, 0, (size_t)N * elt, (size_t)M, stream));
        }
    }
    const int units = tiles * splits;
    int grid = units * kCluster;
    if (grid > max_ctas) grid = max_ctas;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = Cfg<BN, kCluster>::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kCluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // Band height in scheduling units (m-blocks, or pairs of m-blocks): as tall as keeps the
    // band's A panel (rows x K) around 32 MB, i.e. resident in L2 while the wave sweeps N, so B
    // is re-read once per band and A once. Measured (benchmarks/gemm_group_sweep.py): 8192^3
    // best at 8 pairs (1.56 vs 1.45 PFLOP/s at 32), 8192x30528x1024 best at 32 (1.50 vs 1.36).
    int group_m = g_group_m;
    if (group_m <= 0) {
        const long panel_row_bytes = (long)BLOCK_M * kCluster * K * 2;
        group_m = (int)((32L << 20) / panel_row_bytes);
        if (group_m < 2) group_m = 2;
        if (group_m > m_units) group_m = m_units;
    }
    TFOP_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mb, mc, C, bias, bias_dtype, M, N, K, ldc,
                                 out_dtype, epi, a_mn, b_mn, group_m, splits, GemmFuse{}));
    return 0;
}

// Fused GEMM -> in-switch all-reduce: the 2-CTA kernel with the reducer warps. Every CTA of every
// rank must be resident at once (the reducers wait for the other ranks' tiles).
template <int BN>
int launch_fused(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, void* C,
                 int M, int N, int K, int ldc, int a_mn, int b_mn, const GemmFuse& fz, int max_ctas_req,
                 cudaStream_t stream) {
    static int max_ctas_dev[kMaxDevices] = {};
    auto kern = gemm_tn_kernel<BN, true, 2, true>;
    int dev = 0, sms = 0;
    if (current_device(&dev) != 0 || num_sms(&sms) != 0) return -1;
    int& max_ctas = max_ctas_dev[dev];
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    if (max_ctas == 0) {
        TFOP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg<BN, 2>::kSmemBytes));
        cudaLaunchConfig_t q = {};
        q.gridDim = dim3(sms / 2 * 2);
        q.blockDim = dim3(kFuseThreads);
        q.dynamicSmemBytes = Cfg<BN, 2>::kSmemBytes;
        q.attrs = attr;
        q.numAttrs = 1;
        int n = 0;
        TFOP_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &q));
        TFOP_REQUIRE(n > 0, "no 2-CTA cluster of the fused GEMM kernel fits on this device");
        max_ctas = 2 * n;
    }
    const int m_units = ((M + BLOCK_M - 1) / BLOCK_M + 1) / 2;
    const int units = m_units * ((N + BN - 1) / BN);
    int grid = units * 2;
    if (grid > max_ctas) grid = max_ctas;
    if (max_ctas_req > 0 && grid > max_ctas_req) grid = max_ctas_req / 2 * 2;
    if (grid < 2) grid = 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kFuseThreads);
    cfg.dynamicSmemBytes = Cfg<BN, 2>::kSmemBytes;
    cfg.stream = stream;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const long panel_row_bytes = (long)BLOCK_M * 2 * K * 2;
    int group_m = (int)((32L << 20) / panel_row_bytes);
    if (group_m < 2) group_m = 2;
    if (group_m > m_units) group_m = m_units;
    TFOP_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mb, mc, C, (const void*)nullptr, 0, M, N, K, ldc,
                                 (int)TFOP_BF16, (int)EPI_NONE, a_mn, b_mn, group_m, 1, fz));
    return 0;
}

}  // namespace

// Host helpers shared with gemm_allreduce.cu (library-internal, not part of the C ABI).
int tfop_gemm_make_tensor_map(CUtensorMap* map, const void* base, int rows, int K, int ld,
                              int box_rows) {
    if (get_encode() != 0) return -1;
    return make_map(map, base, rows, K, ld, box_rows);
}
int tfop_gemm_num_sms(int* out) { return num_sms(out); }

// The fused GEMM -> in-switch all-reduce launch (gemm_allreduce.cu owns the protocol state).
// C_local: this rank's copy of the multicast-bound output region, [M, N] bf16, row pitch ldo.
// tile_n 128 or 256; a_mn / b_mn as in tfop_gemm_bf16.
int tfop_gemm_launch_fused(const void* A, const void* B, void* C_local, int M, int N, int K,
                           int lda, int ldb, int ldo, int a_mn, int b_mn, int tile_n,
                           const tfop_gemm::GemmFuse* fz, int max_ctas, cudaStream_t stream) {
    if (get_encode() != 0) return -1;
    CUtensorMap ma, mb, mc;
    if (a_mn ? make_map(&ma, A, K, M, lda, BLOCK_K) : make_map(&ma, A, M, K, lda, BLOCK_M)) return -1;
    if (b_mn ? make_map(&mb, B, K, N, ldb, BLOCK_K) : make_map(&mb, B, N, K, ldb, tile_n / 2))
        return -1;
    if (make_map_c(&mc, C_local, M, N, ldo, TFOP_BF16) != 0) return -1;
    if (tile_n == 256)
        return launch_fused<256>(ma, mb, mc, C_local, M, N, K, ldo, a_mn, b_mn, *fz, max_ctas, stream);
    return launch_fused<128>(ma, mb, mc, C_local, M, N, K, ldo, a_mn, b_mn, *fz, max_ctas, stream);
}

// C[M,N] = epi(op(A) * op(B)^T + bias[N]), bf16 operands, fp32 accumulation in TMEM.
//   a_mn = 0: A is [M, K] row-major (K-major), lda elements between rows
//   a_mn = 1: A is [K, M] row-major (MN-major: the stored matrix is A^T), lda between K rows
//   b_mn = 0: B is [N, K] row-major; b_mn = 1: B is [K, N] row-major
// so (0,0) is F.linear(x, W); (0,1) is x @ W (dgrad straight from the stored weight); (1,1) is
// A_stored^T @ B_stored (wgrad straight from dY and X) - no transposed copies anywhere.
// Leading dimensions are multiples of 8 elements, base pointers 16-byte aligned.
// out_dtype: TFOP_BF16 or TFOP_F32. epilogue: 0 none, 1 +bias, 2 +bias,GeLU(erf),
// 3 +bias,ReLU, 4 +bias,GeLU(tanh). tile_n: 0 auto, 128 or 256.
TFOP_API int tfop_gemm_bf16(const void* A, const void* B, void* C, const void* bias,
                            int bias_dtype, int M, int N, int K, int lda, int ldb, int ldc,
                            int out_dtype, int epilogue, int tile_n, int a_mn, int b_mn,
                            cudaStream_t stream) {
    TFOP_REQUIRE(M > 0 && N > 0 && K > 0, "empty GEMM");
    TFOP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "lda/ldb must be multiples of 8 elements (TMA)");
    TFOP_REQUIRE((((uintptr_t)A) & 15) == 0 && (((uintptr_t)B) & 15) == 0,
                 "A and B must be 16-byte aligned");
    TFOP_REQUIRE(out_dtype == TFOP_BF16 || out_dtype == TFOP_F32, "bad out dtype");
    TFOP_REQUIRE(epilogue >= 0 && epilogue <= 4, "bad epilogue");
    TFOP_REQUIRE(epilogue == 0 || bias != nullptr, "epilogue needs a bias pointer");
    if (get_encode() != 0) return -1;
    int g_num_sms = 0;
    if (num_sms(&g_num_sms) != 0) return -1;
    if (tile_n == 0) {
        // wide tiles when they still fill the machine
        const long tiles256 = (long)((M + BLOCK_M - 1) / BLOCK_M) * ((N + 255) / 256);
        tile_n = (N >= 256 && tiles256 >= g_num_sms) ? 256 : 128;
    }
    TFOP_REQUIRE(tile_n == 128 || tile_n == 256, "tile_n must be 0, 128 or 256");
    CUtensorMap ma, mb;
    // K-major: rows = MN extent, inner = K, box [BLOCK_M | tile_n rows, 64]
    // MN-major: rows = K extent, inner = MN, box [64 k-rows, 64]
    if (a_mn ? make_map(&ma, A, K, M, lda, BLOCK_K) : make_map(&ma, A, M, K, lda, BLOCK_M)) return -1;
    // CTA pairs (cta_group::2) take two vertically adjacent m-blocks: worth it from two m-blocks on
    const int m_blocks = (M + BLOCK_M - 1) / BLOCK_M;
    const bool cluster = g_cluster_mode == 1 || (g_cluster_mode == -1 && m_blocks >= 2);
    // (a K-major B box is the part one CTA fetches: the whole tile, or half of it per CTA)
    if (b_mn ? make_map(&mb, B, K, N, ldb, BLOCK_K)
             : make_map(&mb, B, N, K, ldb, cluster ? tile_n / 2 : tile_n))
        return -1;
    // TMA can describe C when its row pitch and base are 16-byte multiples
    const int elt = out_dtype == TFOP_BF16 ? 2 : 4;
    const bool tma_store = !g_force_direct_store && ((size_t)ldc * elt) % 16 == 0 &&
                           (((uintptr_t)C) & 15) == 0;
    CUtensorMap mc = ma;
    if (tma_store && make_map_c(&mc, C, M, N, ldc, out_dtype) != 0) return -1;
#define TFOP_GEMM_LAUNCH(BN_, TMA_, CL_)                                                       \
    launch<BN_, TMA_, CL_>(ma, mb, mc, C, bias, bias_dtype, M, N, K, ldc, out_dtype, epilogue, \
                           a_mn, b_mn, g_splits, tma_store && epilogue == 0 && g_splits >= 0,  \
                           stream)
#define TFOP_GEMM_LAUNCH2(BN_, TMA_) \
    (cluster ? TFOP_GEMM_LAUNCH(BN_, TMA_, 2) : TFOP_GEMM_LAUNCH(BN_, TMA_, 1))
    if (tile_n == 256)
        return tma_store ? TFOP_GEMM_LAUNCH2(256, true) : TFOP_GEMM_LAUNCH2(256, false);
    return tma_store ? TFOP_GEMM_LAUNCH2(128, true) : TFOP_GEMM_LAUNCH2(128, false);
#undef TFOP_GEMM_LAUNCH2
#undef TFOP_GEMM_LAUNCH
}

// The "TN" special case kept as its own entry point: F.linear(x, W, b).
TFOP_API int tfop_gemm_bf16_tn(const void* A, const void* B, void* C, const void* bias,
                               int bias_dtype, int M, int N, int K, int lda, int ldb, int ldc,
                               int out_dtype, int epilogue, int tile_n, cudaStream_t stream) {
    return tfop_gemm_bf16(A, B, C, bias, bias_dtype, M, N, K, lda, ldb, ldc, out_dtype, epilogue,
                          tile_n, 0, 0, stream);
}

// Benchmark / test hook: -1 auto (default), 0 never, 1 always run the CTAs as cta_group::2
// pairs (256-row MMAs, each SM staging half of the B tile).
TFOP_API int tfop_gemm_set_cluster_mode(int mode) {
    TFOP_REQUIRE(mode >= -1 && mode <= 1, "cluster mode must be -1, 0 or 1");
    g_cluster_mode = mode;
    return 0;
}

// Benchmark / test hook: split-K factor (0 = heuristic, -1 = never, n > 0 = force n splits).
TFOP_API int tfop_gemm_set_splits(int s) {
    TFOP_REQUIRE(s >= -1 && s <= 64, "splits out of range");
    g_splits = s;
    return 0;
}

// Benchmark hook: band height of the tile order in scheduling units (0 = default).
TFOP_API int tfop_gemm_set_group_m(int g) {
    TFOP_REQUIRE(g >= 0 && g <= 1024, "group_m out of range");
    g_group_m = g;
    return 0;
}

// Test hook: 1 forces the direct (per-thread 16-byte) C stores even where TMA stores apply.
TFOP_API int tfop_gemm_force_direct_store(int on) {
    g_force_direct_store = on;
    return 0;
}
