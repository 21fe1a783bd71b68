// Device-side building blocks of the tcgen05 / TMEM / TMA GEMM kernels (sm_100a): tile
// constants, mbarrier / TMA / tcgen05 PTX wrappers, shared-memory and instruction descriptors,
// the banded tile order. Shared by gemm_tcgen05.cu and gemm_allreduce.cu.
#pragma once
#include "common.cuh"

#include <cuda.h>
#include <cudaTypedefs.h>

namespace tfop_gemm {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;   // 64 bf16 = 128 bytes = one swizzle-128B row
constexpr int UMMA_K = 16;
constexpr int kNumThreads = 256;
constexpr int kEpiWarp0 = 4;

// kPair = 2: the CTA is one half of a cta_group::2 pair and stages only BN/2 rows of B per
// k-block (the tensor cores read the other half from the peer), which buys two more stages.
template <int BN, int kPair = 1>
struct Cfg {
    static constexpr int kStages = kPair == 2 ? (BN == 256 ? 6 : 8) : (BN == 256 ? 4 : 6);
    static constexpr int kABytes = BLOCK_M * BLOCK_K * 2;
    static constexpr int kBBytes = (BN / kPair) * BLOCK_K * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kTmemCols = 2 * BN;  // two accumulator stages
    static constexpr int kBarBytes = 256;
    // epilogue staging for TMA stores: per epilogue warp two 32-row x 128-byte swizzled boxes
    static constexpr int kStoreBoxBytes = 32 * 128;
    static constexpr int kStoreBytes = 4 * 2 * kStoreBoxBytes;
    static constexpr int kBiasBytes = BN * 4;  // the tile's bias slice as fp32
    static constexpr int kSmemBytes =
        kStages * kStageBytes + kStoreBytes + kBarBytes + kBiasBytes + 1024;  // + align slack
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst,
                                            int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// TMA store of one swizzled shared-memory box into a 2D tensor; rows/columns outside the
// tensor are clipped by the hardware, so M/N tails need no predicates.
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* src, int c0,
                                             int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map),
        "r"(smem_u32(src)), "r"(c0), "r"(c1)
        : "memory");
}
// Same, but global += shared (element-wise add in the tensor's data type, done by the TMA unit
// at L2): the split-K epilogue, where several CTAs own different K ranges of one output tile.
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* map, const void* src, int c0,
                                                  int c1) {
    asm volatile(
        "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::
            "l"(map),
        "r"(smem_u32(src)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void bulk_commit() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// barrier among the 128 epilogue threads only (id 0 is __syncthreads)
__device__ __forceinline__ void epi_bar_sync() {
    asm volatile("bar.sync 1, 128;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile(
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
            smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                          uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// K-major operand tile in SWIZZLE_128B layout: rows of 128 bytes, 8-row groups
// 1024 bytes apart (SBO), LBO unused for swizzled K-major.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;             // LBO (ignored for 128B swizzle, K-major)
    d |= (uint64_t)(1024 >> 4) << 32;   // SBO
    d |= (uint64_t)1 << 46;             // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;             // SWIZZLE_128B
    return d;
}
// MN-major operand tile (the operand is stored [K, MN] with MN contiguous) in SWIZZLE_128B
// layout, as TMA delivers it in boxes of 64 K-rows x 64 MN-elements (128 bytes): k-rows 128
// bytes apart, 8-row groups 1024 bytes apart (SBO), the next 64 MN-elements in the next box
// 8192 bytes on (LBO). One UMMA_K=16 step advances the start address by 16 rows = 2048 bytes.
constexpr int kMnBoxBytes = 64 * 128;
__device__ __forceinline__ uint64_t make_smem_desc_mn(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(kMnBoxBytes >> 4) << 16;  // LBO
    d |= (uint64_t)(1024 >> 4) << 32;         // SBO
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// kind::f16 instruction descriptor: D=f32, A=B=bf16, both K-major.
__device__ __forceinline__ uint32_t make_idesc(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}
// same with per-operand major-ness (bit 15: A is MN-major, bit 16: B is MN-major)
__device__ __forceinline__ uint32_t make_idesc_major(int M, int N, int a_mn, int b_mn) {
    return make_idesc(M, N) | ((uint32_t)(a_mn != 0) << 15) | ((uint32_t)(b_mn != 0) << 16);
}
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]),
          "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]),
          "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ float gelu_tanh(float x) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    const float u = k0 * x * fmaf(k1 * x, x, 1.f);
    return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float gelu_erf(float x) {
    return 0.5f * x * (1.f + erff(x * 0.70710678118654752f));
}

// Tile order: walk down a band of kGroupM m-blocks before moving to the next n-block, so the
// ~148 tiles in flight at any time form a near-square patch of C. Measured on 8192^3 with the
// plain n-fastest order: L2 hit rate 74 %, every B tile re-read from HBM by each m-row; the
// banded order shares each A tile ~9x and each B tile ~16x within a wave.
constexpr int kGroupM = 16;
template <int GROUP>
__device__ __forceinline__ void tile_coords_g(int t, int m_blocks, int n_blocks, int& mb, int& nb) {
    const int per_group = GROUP * n_blocks;
    const int group = t / per_group;
    const int first_m = group * GROUP;
    const int gm = min(GROUP, m_blocks - first_m);
    const int r = t - group * per_group;
    mb = first_m + r % gm;
    nb = r / gm;
}
__device__ __forceinline__ void tile_coords_rt(int t, int m_blocks, int n_blocks, int group,
                                               int& mb, int& nb) {
    const int per_group = group * n_blocks;
    const int g = t / per_group;
    const int first_m = g * group;
    const int gm = min(group, m_blocks - first_m);
    const int r = t - g * per_group;
    mb = first_m + r % gm;
    nb = r / gm;
}
__device__ __forceinline__ void tile_coords(int t, int m_blocks, int n_blocks, int& mb, int& nb) {
    tile_coords_g<kGroupM>(t, m_blocks, n_blocks, mb, nb);
}

// ---- thread-block cluster helpers (2-CTA clusters sharing the B tile by TMA multicast) ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// ---- cta_group::2 (CTA pair) variants -----------------------------------------------------
// Shared-window addresses carry the CTA's rank within the pair in bit 24: clearing it names
// the same offset in the even ("leader") CTA.
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;

// TMA load issued by either CTA of a pair into its own shared memory; the bytes are counted on
// the LEADER's barrier, which the (leader-only) MMA thread waits on.
__device__ __forceinline__ void tma_load_2d_pair(const CUtensorMap* map, uint64_t* bar, void* dst,
                                                 int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
        : "memory");
}
// One 256 x N x 16 MMA executed by both SMs of the pair (each on its 128 rows of A / D and its
// N/2 rows of B); issued by the leader CTA only.
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                               uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Completion of the pair's MMAs, delivered to the barrier at this offset in both CTAs.
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
        " [%0], %1;" ::"r"(smem_u32(bar)),
        "h"((uint16_t)3)
        : "memory");
}
// Arrive on the leader CTA's barrier from either CTA of the pair.
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(
                     smem_u32(bar) & kPeerBitMask)
                 : "memory");
}

// ---- GEMM -> all-reduce in one kernel (NVLS flavour, see gemm_allreduce.cu) -----------------
// The GEMM kernel run with kFuse writes its bf16 partial tiles with TMA stores into the LOCAL
// copy of the multicast-bound output region, tells the tile's owner rank, and four extra warps
// per CTA reduce this rank's owned tiles inside the NVSwitch as they become complete on all
// ranks. arrived[r]: rank r's counter block in its symmetric buffer as mapped here:
//   arrived[tiles_per_owner] | delivered | generation      (monotone, never reset)
constexpr int kFuseThreads = 384;   // 4 control + 4 epilogue + 4 reducer warps
constexpr int kFuseRedWarp0 = 8;
struct GemmFuse {
    uint32_t* arrived[8];
    char* mc_out;            // multicast view of the output region, row pitch ldo elements
    uint32_t* err;
    long long timeout;
    int rank, world, tiles_per_owner, ldo;
    float scale;
};

__device__ __forceinline__ void bulk_wait_all() {      // writes complete, not only reads
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void fuse_red_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 fuse_mm_ld_reduce_bf16x8(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
__device__ __forceinline__ void fuse_mm_st_16(void* mc, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc),
                 "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                 "f"(__uint_as_float(v.w))
                 : "memory");
}
__device__ __forceinline__ void fuse_reducer_bar() {   // named barrier 2: the 128 reducer threads
    asm volatile("bar.sync 2, 128;" ::: "memory");
}

}  // namespace tfop_gemm
