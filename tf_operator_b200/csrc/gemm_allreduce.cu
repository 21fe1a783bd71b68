// Fused GEMM -> reduce-scatter -> all-gather over NVLink peer memory, ONE kernel.
//
// Data-parallel weight gradient: every rank computes its partial  P_r = A_r * B_r^T  (the
// tcgen05/TMEM/TMA pipeline of gemm_tcgen05.cu) and the ranks need  C = scale * sum_r P_r.
// Instead of GEMM-then-all-reduce (two kernels, the whole C written and re-read, no overlap),
// the GEMM epilogue pushes each finished 128x256 tile as bf16 straight into the OWNER rank's
// HBM (tile t belongs to rank t % W) and bumps that tile's arrival counter with a
// release-scope RED over NVLink. As soon as a tile has arrived from all W ranks its owner sums
// the W partials in fp32, scales, and stores the result tile into every rank's output with
// peer stores, then bumps each rank's "delivered" counter. Transfers of early tiles overlap
// the math of later tiles, tile by tile; C never makes a round trip through local HBM.
//
//   warps 0-7  (persistent GEMM)  : tiles -> owner slots   [UTC*MMA + peer st + RED]
//   warps 8-11 (reducers, CONCURRENT): owned tiles as they arrive: wait, reduce, broadcast
//   exit only when this rank's own output has been fully delivered
//
// Symmetric-buffer layout at `ws_off` (see tfop_gemm_allreduce_workspace_bytes):
//   [W][tiles_per_owner][128*256] bf16 partial slots | arrived[tiles_per_owner] u32 |
//   delivered u32 | generation u32
// and the result region at `out_off`: [M, N] bf16 (e.g. the parameter's slot of the DDP
// gradient bucket). Counters are monotone (target = generation * k), never reset; a workspace
// is tied to one (M, N, world): zero it before using it for another shape.
//
// Reference parity: gradient GEMM + NCCL all-reduce of the reference stack
// (examples/v1/distribution_strategy/keras-API/multi_worker_strategy-with-keras.py:76-77,
// 117-120); BASELINE.json: "write ONE kernel that does both over peer memory".
#include "common.cuh"
#include "comm.h"
#include "gemm_common.cuh"

#include <cooperative_groups.h>

int tfop_gemm_make_tensor_map(CUtensorMap* map, const void* base, int rows, int K, int ld,
                              int box_rows);
int tfop_gemm_num_sms(int* out);
int tfop_gemm_launch_fused(const void* A, const void* B, void* C_local, int M, int N, int K,
                           int lda, int ldb, int ldo, int a_mn, int b_mn, int tile_n,
                           const tfop_gemm::GemmFuse* fz, int max_ctas, cudaStream_t stream);

namespace {

using namespace tfop_gemm;

constexpr int BN = 256;
constexpr int kTileElems = BLOCK_M * BN;
constexpr int kGarThreads = 384;   // 4 control + 4 epilogue + 4 reducer warps
constexpr int kRedWarp0 = 8;
constexpr int kRedThreads = 128;
constexpr int kGarStages = 3;          // one pipeline stage traded for the epilogue staging buffer
constexpr int kStgPitch = BN * 2 + 16;  // bytes per staged row (+16: spread rows over smem banks)
constexpr int kStgBytes = 4 * 32 * kStgPitch;
constexpr int kGarSmemBytes = kGarStages * Cfg<BN>::kStageBytes + kStgBytes + 256 + 1024;
long long g_spin_timeout = 30LL * 1000 * 1000 * 1000;   // cycles (~15 s); see tfop_gemm_allreduce_set_timeout

struct GarArgs {
    char* data[TFOP_MAX_WORLD];     // symmetric buffers (peer-mapped)
    size_t ws_off, out_off;
    uint32_t* err;
    int M, N, K, ldo;               // ldo: row pitch (elements) of the output region
    int rank, world;
    int tiles_per_owner;
    float scale;
    long long timeout;
    char* nv_uc;                    // NVLS mode: local / multicast view of the bound buffer
    char* nv_mc;
};

struct GarEmu {
    GarArgs r[TFOP_MAX_WORLD];
};

__device__ __forceinline__ void red_release_sys_add(uint32_t* p, uint32_t v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void reducer_bar() {   // named barrier 1: the 128 reducer threads
    asm volatile("bar.sync 1, 128;" ::: "memory");
}
__device__ __forceinline__ size_t slots_bytes(int world, int tpo) {
    return (size_t)world * tpo * kTileElems * 2;
}
__device__ __forceinline__ uint32_t* arrived_ptr(char* base, const GarArgs& a) {
    return (uint32_t*)(base + a.ws_off + (a.nv_mc ? 0 : slots_bytes(a.world, a.tiles_per_owner)));
}
// 16 bytes of bf16 summed over every rank's copy inside the NVSwitch / stored to every copy
__device__ __forceinline__ uint4 mm_ld_reduce_bf16x8(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
__device__ __forceinline__ void mm_st_16(void* mc, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc),
                 "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                 "f"(__uint_as_float(v.w))
                 : "memory");
}
__device__ __forceinline__ bool spin_until(const uint32_t* p, uint32_t target, uint32_t* err,
                                           long long timeout) {
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys_u32(p) - target) < 0) {
        if (clock64() - t0 > timeout) {
            atomicExch(err, 1u);
            return false;
        }
    }
    return true;
}

template <bool NVLS>
__device__ void gemm_allreduce_body(const CUtensorMap& map_a, const CUtensorMap& map_b,
                                    const GarArgs& a) {
    using cfg = Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + kGarStages * cfg::kABytes;
    uint8_t* smem_stg = smem + kGarStages * cfg::kStageBytes;
    uint64_t* bars = (uint64_t*)(smem_stg + kStgBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + kGarStages;
    uint64_t* tmem_full = bars + 2 * kGarStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_slot = (uint32_t*)(tmem_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m_blocks = (a.M + BLOCK_M - 1) / BLOCK_M;
    const int n_blocks = (a.N + BN - 1) / BN;
    const int num_tiles = m_blocks * n_blocks;
    const int k_blocks = (a.K + BLOCK_K - 1) / BLOCK_K;
    const int W = a.world;

    char* self = a.data[a.rank];
    uint32_t* my_arrived = arrived_ptr(self, a);
    uint32_t* my_delivered = my_arrived + a.tiles_per_owner;
    uint32_t* my_gen = my_delivered + 1;
    const uint32_t gen = *my_gen + 1;     // every thread reads; CTA 0 writes it back at the end

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kGarStages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)),
                     "r"((uint32_t)cfg::kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // ------------------------------ phase 1: GEMM tiles -> owner slots ----------------------
    if (warp == 0) {
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m0 = (t / n_blocks) * BLOCK_M;
                const int n0 = (t % n_blocks) * BN;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    mbar_expect_tx(&full[stage], cfg::kStageBytes);
                    tma_load_2d(&map_a, &full[stage], smem_a + stage * cfg::kABytes, kb * BLOCK_K, m0);
                    tma_load_2d(&map_b, &full[stage], smem_b + stage * cfg::kBBytes, kb * BLOCK_K, n0);
                    if (++stage == kGarStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            const uint32_t idesc = make_idesc(BLOCK_M, BN);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tcgen05_fence_after();
                    const uint32_t a_addr = smem_u32(smem_a + stage * cfg::kABytes);
                    const uint32_t b_addr = smem_u32(smem_b + stage * cfg::kBBytes);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                        umma_bf16(tmem_d, make_smem_desc(a_addr + k * UMMA_K * 2),
                                  make_smem_desc(b_addr + k * UMMA_K * 2), idesc,
                                  (kb | k) != 0 ? 1u : 0u);
                    umma_commit(&empty[stage]);
                    if (++stage == kGarStages) {
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
    } else if (warp >= kEpiWarp0 && warp < kRedWarp0) {
        const int ew = warp - kEpiWarp0;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            const int owner = t % W;
            const int slot = t / W;
            // partial tile of THIS rank inside the owner's HBM: [src rank][slot][128][256] bf16
            __nv_bfloat16* dst;
            uint32_t row_bytes = BN * 2;
            if (NVLS) {
                // NVLS mode: the partial stays LOCAL, written at its final place in the
                // multicast-bound output region; the switch sums the ranks' copies later
                const int gr = (t / n_blocks) * BLOCK_M + ew * 32 + lane;
                const int gc = (t % n_blocks) * BN;
                dst = (__nv_bfloat16*)(a.nv_uc + a.out_off) + (size_t)gr * a.ldo + gc;
                const int cols = min(BN, a.N - gc);
                row_bytes = gr < a.M ? (uint32_t)cols * 2 : 0;
            } else {
                dst = (__nv_bfloat16*)(a.data[owner] + a.ws_off) +
                      ((size_t)a.rank * a.tiles_per_owner + slot) * kTileElems +
                      (size_t)(ew * 32 + lane) * BN;
            }
            uint8_t* my_row = smem_stg + (size_t)(ew * 32 + lane) * kStgPitch;
            mbar_wait(&tmem_full[acc], acc_phase);
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + (uint32_t)(acc * BN) + ((uint32_t)(ew * 32) << 16);
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(taddr + (uint32_t)c, r);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                    float f[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(r[j + i]);
                    bf16x8 pk = f32_to_bf16x8(f);
                    *reinterpret_cast<uint4*>(my_row + (c + j) * 2) = *reinterpret_cast<uint4*>(&pk);
                }
            }
            // the accumulator is drained: hand TMEM back to the MMA warp before the transfer
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[acc]);
            // one 512-byte bulk store per row, smem -> owner's HBM over NVLink (async proxy)
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            if (row_bytes)
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst),
                             "r"(smem_u32(my_row)), "r"(row_bytes)
                             : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            asm volatile("fence.proxy.async;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                fence_acq_rel_sys();
                red_release_sys_add(arrived_ptr(a.data[owner], a) + slot, 1u);
            }
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1;
            }
        }
    }

    // ------------------ phase 2, CONCURRENT with phase 1: the reducer warps take this rank's
    // owned tiles as they arrive from all ranks (wait, reduce in fp32, scale, broadcast) while
    // the TMA / MMA / epilogue warps keep computing later tiles.
    if (warp >= kRedWarp0) {
        const int rt = threadIdx.x - kRedWarp0 * 32;
        const int owned = (num_tiles - a.rank + W - 1) / W;   // tiles t = j*W + rank < num_tiles
        for (int j = blockIdx.x; j < owned; j += gridDim.x) {
            const int t = j * W + a.rank;
            const int m0 = (t / n_blocks) * BLOCK_M;
            const int n0 = (t % n_blocks) * BN;
            if (rt == 0) spin_until(my_arrived + j, gen * 4u * (uint32_t)W, a.err, a.timeout);
            reducer_bar();
            if (NVLS) {
                constexpr int UN = 4;
                for (int v0 = rt; v0 < kTileElems / 8; v0 += kRedThreads * UN) {
                    uint4 r[UN];
                    size_t off[UN];
                    bool ok[UN];
#pragma unroll
                    for (int u = 0; u < UN; ++u) {
                        const int v = v0 + u * kRedThreads;
                        const int gr = m0 + v / (BN / 8), gc = n0 + (v % (BN / 8)) * 8;
                        ok[u] = gr < a.M && gc < a.N;
                        off[u] = a.out_off + ((size_t)gr * a.ldo + gc) * 2;
                        if (ok[u]) r[u] = mm_ld_reduce_bf16x8(a.nv_mc + off[u]);   // summed in the switch
                    }
#pragma unroll
                    for (int u = 0; u < UN; ++u) {
                        if (!ok[u]) continue;
                        if (a.scale != 1.f) {
                            float f[8];
                            bf16x8_to_f32(*reinterpret_cast<bf16x8*>(&r[u]), f);
#pragma unroll
                            for (int i = 0; i < 8; ++i) f[i] *= a.scale;
                            bf16x8 pk = f32_to_bf16x8(f);
                            r[u] = *reinterpret_cast<uint4*>(&pk);
                        }
                        mm_st_16(a.nv_mc + off[u], r[u]);                          // to every rank
                    }
                }
            }
            const __nv_bfloat16* slots = (const __nv_bfloat16*)(self + a.ws_off);
            constexpr int U = 4;   // vectors in flight per thread and source
            for (int v0 = rt; !NVLS && v0 < kTileElems / 8; v0 += kRedThreads * U) {
                float acc8[U][8];
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int i = 0; i < 8; ++i) acc8[u][i] = 0.f;
                for (int src = 0; src < W; ++src) {
                    const __nv_bfloat16* base =
                        slots + ((size_t)src * a.tiles_per_owner + j) * kTileElems;
                    uint4 raw[U];
#pragma unroll
                    for (int u = 0; u < U; ++u)
                        raw[u] = ld_peer_v4(base + (size_t)(v0 + u * kRedThreads) * 8);
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        float f[8];
                        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw[u]), f);
#pragma unroll
                        for (int i = 0; i < 8; ++i) acc8[u][i] += f[i];
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int v = v0 + u * kRedThreads;
                    const int row = v / (BN / 8), cv = v % (BN / 8);
#pragma unroll
                    for (int i = 0; i < 8; ++i) acc8[u][i] *= a.scale;
                    const int gr = m0 + row, gc = n0 + cv * 8;
                    if (gr < a.M && gc < a.N) {
                        const bf16x8 pk = f32_to_bf16x8(acc8[u]);
                        const size_t off = a.out_off + ((size_t)gr * a.ldo + gc) * 2;
                        if (gc + 8 <= a.N && (off & 15) == 0) {
                            for (int p = 0; p < W; ++p)
                                st_peer_v4(a.data[p] + off, *reinterpret_cast<const uint4*>(&pk));
                        } else {
                            const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&pk);
                            for (int p = 0; p < W; ++p)
                                for (int i = 0; i < 8 && gc + i < a.N; ++i)
                                    ((__nv_bfloat16*)(a.data[p] + off))[i] = e[i];
                        }
                    }
                }
            }
            reducer_bar();
            if (rt < W) {
                fence_acq_rel_sys();
                uint32_t* peer_delivered = arrived_ptr(a.data[rt], a) + a.tiles_per_owner;
                red_release_sys_add(peer_delivered, 1u);
            }
        }
        // this rank's output is complete once every tile has been delivered by its owner
        if (blockIdx.x == 0 && rt == 0) {
            spin_until(my_delivered, gen * (uint32_t)num_tiles, a.err, a.timeout);
            *my_gen = gen;
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 2) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                     "r"((uint32_t)cfg::kTmemCols)
                     : "memory");
    }
}

__global__ void __launch_bounds__(kGarThreads, 1)
gemm_allreduce_kernel(const __grid_constant__ CUtensorMap map_a,
                      const __grid_constant__ CUtensorMap map_b, const GarArgs a) {
    gemm_allreduce_body<false>(map_a, map_b, a);
}

// NVLS flavour: partial tiles stay local, the NVSwitch reduces and broadcasts them
__global__ void __launch_bounds__(kGarThreads, 1)
gemm_allreduce_nvls_kernel(const __grid_constant__ CUtensorMap map_a,
                           const __grid_constant__ CUtensorMap map_b, const GarArgs a) {
    gemm_allreduce_body<true>(map_a, map_b, a);
}

struct EmuMaps {
    CUtensorMap a[4];
    CUtensorMap b[4];
};

// all ranks on one GPU, one cooperative launch: blockIdx.y is the rank (world <= 4)
__global__ void __launch_bounds__(kGarThreads, 1)
gemm_allreduce_kernel_emu(const __grid_constant__ EmuMaps maps, const GarEmu e) {
    gemm_allreduce_body<false>(maps.a[blockIdx.y], maps.b[blockIdx.y], e.r[blockIdx.y]);
}

int tiles_per_owner(int M, int N, int W) {
    const int tiles = ((M + BLOCK_M - 1) / BLOCK_M) * ((N + BN - 1) / BN);
    return (tiles + W - 1) / W;
}

size_t ws_bytes(int M, int N, int W) {
    const int tpo = tiles_per_owner(M, N, W);
    size_t b = (size_t)W * tpo * kTileElems * 2 + ((size_t)tpo + 2) * 4;
    return (b + 255) & ~(size_t)255;
}

int fill(tfop_comm* c, GarArgs* a, int M, int N, int K, int ldo, float scale, size_t ws_off,
         size_t out_off) {
    TFOP_REQUIRE(M > 0 && N > 0 && K > 0, "empty GEMM");
    TFOP_REQUIRE((ws_off & 255) == 0 && (out_off & 1) == 0, "ws_off must be 256-byte aligned");
    TFOP_REQUIRE(ws_off + ws_bytes(M, N, c->world) <= c->data_bytes,
                 "workspace does not fit the symmetric buffer");
    TFOP_REQUIRE(out_off + ((size_t)(M - 1) * ldo + N) * 2 <= c->data_bytes,
                 "output region does not fit the symmetric buffer");
    memset(a, 0, sizeof(*a));
    for (int p = 0; p < c->world; ++p) a->data[p] = (char*)c->data[p];
    a->ws_off = ws_off;
    a->out_off = out_off;
    a->err = c->sig[c->rank] + TFOP_SIG_WORDS;
    a->M = M;
    a->N = N;
    a->K = K;
    a->ldo = ldo;
    a->rank = c->rank;
    a->world = c->world;
    a->tiles_per_owner = tiles_per_owner(M, N, c->world);
    a->scale = scale;
    a->timeout = g_spin_timeout;
    return 0;
}

}  // namespace

// Spin time-out of the in-kernel waits, in SM clock cycles (default 30e9, about 15 s).
TFOP_API int tfop_gemm_allreduce_set_timeout(long long cycles) {
    g_spin_timeout = cycles;
    return 0;
}

// Bytes of symmetric workspace needed at ws_off (must start zeroed; reusable across calls).
TFOP_API size_t tfop_gemm_allreduce_workspace_bytes(int M, int N, int world) {
    return ws_bytes(M, N, world);
}

// out[M,N] (bf16, in EVERY rank's symmetric buffer at out_off, row pitch ldo) =
//     scale * sum over ranks of A_r[M,K] * B_r[N,K]^T.   A, B: local bf16, K-major.
// All ranks call with the same shapes/offsets. max_ctas: 0 = one CTA per SM.
TFOP_API int tfop_gemm_allreduce_bf16_tn(tfop_comm* c, const void* A, const void* B, int M, int N,
                                         int K, int lda, int ldb, int ldo, float scale,
                                         size_t ws_off, size_t out_off, int max_ctas,
                                         cudaStream_t stream) {
    TFOP_REQUIRE(c != nullptr && !c->emulated, "needs a real communicator");
    TFOP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "lda/ldb must be multiples of 8 elements (TMA)");
    GarArgs a;
    if (fill(c, &a, M, N, K, ldo, scale, ws_off, out_off) != 0) return -1;
    CUtensorMap ma, mb;
    if (tfop_gemm_make_tensor_map(&ma, A, M, K, lda, BLOCK_M) != 0) return -1;
    if (tfop_gemm_make_tensor_map(&mb, B, N, K, ldb, BN) != 0) return -1;
    {   // per call: cheap, and right for whichever device is current
        TFOP_CUDA(cudaFuncSetAttribute(gemm_allreduce_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kGarSmemBytes));
    }
    int sms = 0;
    if (tfop_gemm_num_sms(&sms) != 0) return -1;
    const int tiles = ((M + BLOCK_M - 1) / BLOCK_M) * ((N + BN - 1) / BN);
    int grid = tiles < sms ? tiles : sms;
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
    gemm_allreduce_kernel<<<grid, kGarThreads, kGarSmemBytes, stream>>>(ma, mb, a);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// Test / benchmark hooks of the NVLS flavour: tile width (0 = heuristic, 128, 256) and 1 to run
// the first-generation kernel (one CTA per tile, 3-stage pipeline) for comparison.
static int g_nvls_tile_n = 0;
static int g_nvls_legacy = 0;
TFOP_API int tfop_gemm_allreduce_set_tile_n(int tile_n) {
    TFOP_REQUIRE(tile_n == 0 || tile_n == 128 || tile_n == 256, "tile_n must be 0, 128 or 256");
    g_nvls_tile_n = tile_n;
    return 0;
}
TFOP_API int tfop_gemm_allreduce_set_legacy(int on) {
    g_nvls_legacy = on;
    return 0;
}

// NVLS flavour: out[M,N] (bf16, at out_off of the communicator's multicast-bound buffer
// (tfop_nvls_bind), row pitch ldo, in EVERY rank's copy) = scale * sum over ranks of
// op(A_r) * op(B_r)^T with a_mn / b_mn as in tfop_gemm_bf16 ((1,1) = the weight gradient
// dY^T X straight from the stored activations). One kernel: the cta_group::2 GEMM pipeline of
// gemm_tcgen05.cu stores each partial tile into the local copy, tells the tile's owner, and
// reducer warps in the same CTAs let the NVSwitch sum (multimem.ld_reduce) and broadcast
// (multimem.st) the tiles as they complete. Partial tiles never cross NVLink as such.
// ws_off: (tiles + 2) * 4 bytes of zeroed counters in the IPC symmetric buffer, tiles counted
// at 128 x 128 (tfop_gemm_allreduce_nvls_workspace_bytes); tied to one (M, N, world).
// N and ldo must be multiples of 8. All ranks call with the same arguments.
TFOP_API size_t tfop_gemm_allreduce_nvls_workspace_bytes(int M, int N, int world) {
    const int tiles = (((M + BLOCK_M - 1) / BLOCK_M) + 1) / 2 * 2 * ((N + 127) / 128);
    return (((size_t)(tiles + world - 1) / world + 2) * 4 + 255) & ~(size_t)255;
}

TFOP_API int tfop_gemm_allreduce_nvls_bf16(tfop_comm* c, const void* A, const void* B, int M,
                                           int N, int K, int lda, int ldb, int ldo, int a_mn,
                                           int b_mn, float scale, size_t ws_off, size_t out_off,
                                           int max_ctas, cudaStream_t stream) {
    TFOP_REQUIRE(c != nullptr && !c->emulated, "needs a real communicator");
    TFOP_REQUIRE(c->mc_ptr != nullptr, "NVLS is not set up on this communicator");
    TFOP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "lda/ldb must be multiples of 8 elements (TMA)");
    TFOP_REQUIRE(N % 8 == 0 && ldo % 8 == 0 && (out_off & 15) == 0,
                 "NVLS mode needs N, ldo multiples of 8 and a 16-byte aligned out_off");
    TFOP_REQUIRE(M > 0 && N > 0 && K > 0, "empty GEMM");
    TFOP_REQUIRE(out_off + ((size_t)(M - 1) * ldo + N) * 2 <= c->mc_bytes,
                 "output region does not fit the NVLS buffer");
    const int m_blocks = (M + BLOCK_M - 1) / BLOCK_M;
    const bool legacy = g_nvls_legacy || (m_blocks < 2 && !a_mn && !b_mn);
    int sms = 0;
    if (tfop_gemm_num_sms(&sms) != 0) return -1;
    if (!legacy) {
        // wide tiles unless they would leave most of the machine idle (measured on 8 GPUs:
        // 256 wins from 48 tile pairs up, 128 at 16 - r2_gemm_allreduce_8gpu.json)
        int tile_n = g_nvls_tile_n;
        if (tile_n == 0) {
            const long pairs256 = (long)((m_blocks + 1) / 2) * ((N + 255) / 256);
            tile_n = (N >= 256 && pairs256 >= 32) ? 256 : 128;
        }
        // tiles are owned by production sequence number: 2 per scheduling unit (CTA pair)
        const int tiles = (m_blocks + 1) / 2 * 2 * ((N + tile_n - 1) / tile_n);
        const int tpo = (tiles + c->world - 1) / c->world;
        TFOP_REQUIRE((ws_off & 15) == 0 && ws_off + ((size_t)tpo + 2) * 4 <= c->data_bytes,
                     "counter workspace does not fit the symmetric buffer");
        tfop_gemm::GemmFuse fz;
        memset(&fz, 0, sizeof(fz));
        for (int p = 0; p < c->world; ++p) fz.arrived[p] = (uint32_t*)((char*)c->data[p] + ws_off);
        fz.mc_out = (char*)c->mc_ptr + out_off;
        fz.err = c->sig[c->rank] + TFOP_SIG_WORDS;
        fz.timeout = g_spin_timeout;
        fz.rank = c->rank;
        fz.world = c->world;
        fz.tiles_per_owner = tpo;
        fz.ldo = ldo;
        fz.scale = scale;
        return tfop_gemm_launch_fused(A, B, (char*)c->mc_uc_ptr + out_off, M, N, K, lda, ldb, ldo,
                                      a_mn, b_mn, tile_n, &fz, max_ctas, stream);
    }
    TFOP_REQUIRE(!a_mn && !b_mn, "the first-generation kernel takes K-major operands only");
    const int tpo = tiles_per_owner(M, N, c->world);
    TFOP_REQUIRE((ws_off & 15) == 0 && ws_off + ((size_t)tpo + 2) * 4 <= c->data_bytes,
                 "counter workspace does not fit the symmetric buffer");
    GarArgs a;
    memset(&a, 0, sizeof(a));
    for (int p = 0; p < c->world; ++p) a.data[p] = (char*)c->data[p];
    a.ws_off = ws_off;
    a.out_off = out_off;
    a.err = c->sig[c->rank] + TFOP_SIG_WORDS;
    a.M = M; a.N = N; a.K = K; a.ldo = ldo;
    a.rank = c->rank; a.world = c->world;
    a.tiles_per_owner = tpo;
    a.scale = scale;
    a.timeout = g_spin_timeout;
    a.nv_uc = (char*)c->mc_uc_ptr;
    a.nv_mc = (char*)c->mc_ptr;
    CUtensorMap ma, mb;
    if (tfop_gemm_make_tensor_map(&ma, A, M, K, lda, BLOCK_M) != 0) return -1;
    if (tfop_gemm_make_tensor_map(&mb, B, N, K, ldb, BN) != 0) return -1;
    {   // per call: cheap, and right for whichever device is current
        TFOP_CUDA(cudaFuncSetAttribute(gemm_allreduce_nvls_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, kGarSmemBytes));
    }
    const int tiles = ((M + BLOCK_M - 1) / BLOCK_M) * ((N + BN - 1) / BN);
    int grid = tiles < sms ? tiles : sms;
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
    gemm_allreduce_nvls_kernel<<<grid, kGarThreads, kGarSmemBytes, stream>>>(ma, mb, a);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// K-major operands (F.linear layout): the entry point of round 1, kept.
TFOP_API int tfop_gemm_allreduce_nvls_bf16_tn(tfop_comm* c, const void* A, const void* B, int M,
                                              int N, int K, int lda, int ldb, int ldo, float scale,
                                              size_t ws_off, size_t out_off, int max_ctas,
                                              cudaStream_t stream) {
    return tfop_gemm_allreduce_nvls_bf16(c, A, B, M, N, K, lda, ldb, ldo, 0, 0, scale, ws_off,
                                         out_off, max_ctas, stream);
}

// Single-GPU emulation for tests: all `world` (<= 4) ranks of an emulated communicator group
// in one cooperative launch. As/Bs: per-rank operand pointers.
TFOP_API int tfop_gemm_allreduce_emulated(tfop_comm** comms, int world, const void* const* As,
                                          const void* const* Bs, int M, int N, int K, int lda,
                                          int ldb, int ldo, float scale, size_t ws_off,
                                          size_t out_off, cudaStream_t stream) {
    TFOP_REQUIRE(world >= 1 && world <= 4, "emulation supports up to 4 ranks");
    GarEmu e;
    EmuMaps maps;
    memset(&e, 0, sizeof(e));
    memset(&maps, 0, sizeof(maps));
    for (int r = 0; r < world; ++r) {
        TFOP_REQUIRE(comms[r]->emulated, "comm is not emulated");
        if (fill(comms[r], &e.r[r], M, N, K, ldo, scale, ws_off, out_off) != 0) return -1;
        if (tfop_gemm_make_tensor_map(&maps.a[r], As[r], M, K, lda, BLOCK_M) != 0) return -1;
        if (tfop_gemm_make_tensor_map(&maps.b[r], Bs[r], N, K, ldb, BN) != 0) return -1;
    }
    {   // per call: cheap, and right for whichever device is current
        TFOP_CUDA(cudaFuncSetAttribute(gemm_allreduce_kernel_emu,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kGarSmemBytes));
    }
    int sms = 0;
    if (tfop_gemm_num_sms(&sms) != 0) return -1;
    const int tiles = ((M + BLOCK_M - 1) / BLOCK_M) * ((N + BN - 1) / BN);
    int gx = sms / world;                 // every CTA of every rank must be co-resident
    if (gx > tiles) gx = tiles;
    if (gx < 1) gx = 1;
    dim3 grid(gx, world), block(kGarThreads);
    void* params[] = {(void*)&maps, (void*)&e};
    TFOP_CUDA(cudaLaunchCooperativeKernel((const void*)gemm_allreduce_kernel_emu, grid, block,
                                          params, kGarSmemBytes, stream));
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
