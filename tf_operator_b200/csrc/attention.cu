// K13: multi-head self-attention (head dimension 64, no mask) on the tcgen05 tensor cores -
// forward and backward, reading the PACKED projection output qkv[B*S, 3*H] exactly as the QKV
// GEMM wrote it and writing the context as [B*S, H], so no transposed / unbound copy of Q, K or V
// (nor their gradients) ever exists.
//
// Forward: persistent CTAs (one per SM) walk the (batch, head, 128-query) tiles. Two passes over
// the keys of a tile (128 at a time, TMA ring), software-pipelined over THREE S accumulators in
// TMEM and TWO P tiles in shared memory so the MMA thread runs ahead of the softmax warps (the
// first version - one S, one P, everything a serial MMA -> softmax -> MMA round trip - left the
// tensor pipe 18 % and the MUFU pipe 27 % busy: ncu, profiles/r2_ncu_attention.md):
//   pass 1  S = Q K^T (tcgen05.mma, fp32 in TMEM) -> exact row maxima
//   pass 2  S again -> P = exp2((S - max) * scale * log2 e) -> bf16 into a swizzled K-major
//           shared-memory tile -> O += P V (V read as stored, MN-major) accumulated in TMEM
//   end     O / rowsum -> bf16 -> TMA store; log-sum-exp (base 2) kept for the backward.
// The second QK^T costs a third more tensor work (the kernel is bound by the exponentials, not
// by the MMAs) and buys an O accumulator that never has to be rescaled.
//
// Reference parity: the attention of the BERT-large TFJob of BASELINE.json config 3 (the worker
// matmuls of examples/v1/dist-mnist/dist_mnist.py:188-192 in the reference stack); replaces the
// cuDNN scaled_dot_product_attention call.
#include "common.cuh"
#include "gemm_common.cuh"

namespace {

using namespace tfop_gemm;

constexpr int kD = 64;          // head dimension
constexpr int kBQ = 128;        // queries per CTA
constexpr int kBK = 128;        // keys per block
constexpr int kTile = 128 * 64 * 2;   // one [128 x 64] bf16 tile = 16 KB
constexpr int kRing = 4;        // K / V ring slots
constexpr int kSBuf = 3;        // S accumulators in TMEM (128 columns each)
constexpr int kFwdThreads = 256;
constexpr int kFwdSmem = 2 * kTile /*Q x2*/ + kRing * kTile + 2 * 2 * kTile /*P x2*/ +
                         4 * 4096 /*O staging*/ + 256 + 1024;
constexpr int kFwdTmemCols = 512;  // S_b: [128 b, 128 b + 128), b < 3; O: [384, 448)
static_assert(kFwdSmem <= 227 * 1024, "shared memory budget of one CTA");

struct AttnMaps {
    CUtensorMap qkv;   // [B*S, 3H] bf16, box {64, 128}
    CUtensorMap out;   // [B*S, H] bf16, box {64, 32} (store)
};

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__global__ void __launch_bounds__(kFwdThreads, 1)
attn_fwd_kernel(const __grid_constant__ AttnMaps maps, float* __restrict__ lse, int S, int H,
                int heads, int n_items, float scale_log2e) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sq = smem;                          // 2 x Q tile
    uint8_t* sring = sq + 2 * kTile;             // K / V ring
    uint8_t* sp = sring + kRing * kTile;         // 2 x P tile (2 boxes of 64 keys each)
    uint8_t* sstage = sp + 4 * kTile;            // 4 x 4 KB O staging
    uint64_t* bars = (uint64_t*)(sstage + 4 * 4096);
    uint64_t* q_full = bars;                     // [2]
    uint64_t* q_empty = q_full + 2;              // [2]
    uint64_t* r_full = q_empty + 2;              // [kRing]
    uint64_t* r_empty = r_full + kRing;          // [kRing]
    uint64_t* s_full = r_empty + kRing;          // [kSBuf] MMA -> softmax
    uint64_t* s_free = s_full + kSBuf;           // [kSBuf] softmax -> MMA (S read)
    uint64_t* p_full = s_free + kSBuf;           // [2] softmax -> MMA (P written)
    uint64_t* p_free = p_full + 2;               // [2] MMA -> softmax (P V has read P)
    uint64_t* o_done = p_free + 2;               // MMA -> softmax: O of this tile complete
    uint64_t* o_free = o_done + 1;               // softmax -> MMA: O read out
    uint32_t* tmem_slot = (uint32_t*)(o_free + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nkb = S / kBK, nqb = S / kBQ;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.qkv) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.out) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&q_full[i], 1);
            mbar_init(&q_empty[i], 1);
            mbar_init(&p_full[i], 4);
            mbar_init(&p_free[i], 1);
        }
        for (int i = 0; i < kRing; ++i) {
            mbar_init(&r_full[i], 1);
            mbar_init(&r_empty[i], 1);
        }
        for (int i = 0; i < kSBuf; ++i) {
            mbar_init(&s_full[i], 1);
            mbar_init(&s_free[i], 4);
        }
        mbar_init(o_done, 1);
        mbar_init(o_free, 4);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)),
                     "r"((uint32_t)kFwdTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_o = tmem_base + 384u;

    // tile -> (batch, head, query block); query blocks of one (batch, head) are neighbours, so
    // the CTAs working on them at the same time share its K / V through L2
    auto decode = [&](int w, int& b, int& head, int& qb) {
        qb = w % nqb;
        const int t = w / nqb;
        head = t % heads;
        b = t / heads;
    };

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            uint32_t L = 0;
            auto load_kv = [&](int which, int b, int head, int j) {
                const int slot = (int)(L % kRing);
                mbar_wait(&r_empty[slot], ((L / kRing) & 1u) ^ 1u);
                mbar_expect_tx(&r_full[slot], kTile);
                tma_load_2d(&maps.qkv, &r_full[slot], sring + slot * kTile, which * H + head * kD,
                            b * S + j * kBK);
                ++L;
            };
            int it = 0;
            for (int w = blockIdx.x; w < n_items; w += gridDim.x, ++it) {
                int b, head, qb;
                decode(w, b, head, qb);
                const int qs = it & 1;
                mbar_wait(&q_empty[qs], (((uint32_t)it >> 1) & 1u) ^ 1u);
                mbar_expect_tx(&q_full[qs], kTile);
                tma_load_2d(&maps.qkv, &q_full[qs], sq + qs * kTile, head * kD, b * S + qb * kBQ);
                for (int j = 0; j < nkb; ++j) load_kv(1, b, head, j);          // pass 1
                for (int j = 0; j < 2 && j < nkb; ++j) load_kv(1, b, head, j); // pass 2 runs 2 ahead
                for (int j = 0; j < nkb; ++j) {
                    if (j + 2 < nkb) load_kv(1, b, head, j + 2);
                    load_kv(2, b, head, j);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            const uint32_t idesc_s = make_idesc(kBQ, kBK);                   // S = Q K^T
            const uint32_t idesc_o = make_idesc_major(kBQ, kD, 0, 1);        // O += P V (V MN-major)
            uint32_t L = 0, T = 0, J = 0;    // ring loads / S tiles / P V products consumed so far
            int it = 0;
            for (int w = blockIdx.x; w < n_items; w += gridDim.x, ++it) {
                const int qs = it & 1;
                mbar_wait(&q_full[qs], ((uint32_t)it >> 1) & 1u);
                tcgen05_fence_after();
                const uint32_t q_addr = smem_u32(sq + qs * kTile);
                auto issue_s = [&]() {
                    const int slot = (int)(L % kRing);
                    mbar_wait(&r_full[slot], (L / kRing) & 1u);
                    const int sb = (int)(T % kSBuf);
                    if (T >= kSBuf) mbar_wait(&s_free[sb], ((T / kSBuf) - 1u) & 1u);
                    tcgen05_fence_after();
                    const uint32_t k_addr = smem_u32(sring + slot * kTile);
#pragma unroll
                    for (int k = 0; k < kD / UMMA_K; ++k)
                        umma_bf16(tmem_base + (uint32_t)(sb * 128), make_smem_desc(q_addr + k * 32),
                                  make_smem_desc(k_addr + k * 32), idesc_s, k != 0 ? 1u : 0u);
                    umma_commit(&r_empty[slot]);
                    umma_commit(&s_full[sb]);
                    ++L;
                    ++T;
                };
                for (int j = 0; j < nkb; ++j) issue_s();
                for (int j = 0; j < 2 && j < nkb; ++j) issue_s();
                for (int j = 0; j < nkb; ++j) {
                    if (j + 2 < nkb) issue_s();
                    const int slot = (int)(L % kRing);
                    mbar_wait(&r_full[slot], (L / kRing) & 1u);                  // V_j
                    const int pb = (int)(J & 1u);
                    mbar_wait(&p_full[pb], (J >> 1) & 1u);                       // P_j written
                    if (j == 0 && it > 0) mbar_wait(o_free, (uint32_t)(it - 1) & 1u);
                    tcgen05_fence_after();
                    const uint32_t v_addr = smem_u32(sring + slot * kTile);
                    const uint32_t p_addr = smem_u32(sp + pb * 2 * kTile);
#pragma unroll
                    for (int k = 0; k < kBK / UMMA_K; ++k) {
                        // P: two K-major boxes of 64 keys; V: 128 key rows of 128 bytes
                        const uint32_t pa = p_addr + (k >> 2) * kTile + (k & 3) * 32;
                        umma_bf16(tmem_o, make_smem_desc(pa), make_smem_desc_mn(v_addr + k * 2048),
                                  idesc_o, (j | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&r_empty[slot]);
                    umma_commit(&p_free[pb]);
                    ++L;
                    ++J;
                }
                umma_commit(o_done);
                umma_commit(&q_empty[qs]);
            }
        }
    } else if (warp >= kEpiWarp0) {
        // ===================== softmax: one thread per query row =====================
        const int ew = warp - kEpiWarp0;
        const int row = ew * 32 + lane;
        const uint32_t lane_sel = (uint32_t)(ew * 32) << 16;
        const uint32_t sw = (uint32_t)(row & 7);
        uint8_t* obox = sstage + ew * 4096;
        uint32_t T = 0, J = 0;
        int it = 0;
        for (int w = blockIdx.x; w < n_items; w += gridDim.x, ++it) {
            int b, head, qb;
            decode(w, b, head, qb);
            // Both passes keep ONE tcgen05.ld in flight while the previous chunk is being worked on
            // (tcgen05.wait::ld waits for every outstanding load, so the next load is issued right
            // after the wait and before the arithmetic): the first version waited out each load's
            // full latency 24 times per tile.
            float m = -INFINITY;
            {   // pass 1: exact row maximum; half tiles of 64 keys ping-pong between A and B
                uint32_t A0[32], A1[32], B0[32], B1[32];
                {
                    const int sb = (int)(T % kSBuf);
                    mbar_wait(&s_full[sb], (T / kSBuf) & 1u);
                    tcgen05_fence_after();
                    const uint32_t t_s = tmem_base + (uint32_t)(sb * 128) + lane_sel;
                    tmem_ld_32x32b_x32(t_s, A0);
                    tmem_ld_32x32b_x32(t_s + 32u, A1);
                }
                for (int j = 0; j < nkb; ++j, ++T) {
                    const int sb = (int)(T % kSBuf);
                    const uint32_t t_s = tmem_base + (uint32_t)(sb * 128) + lane_sel;
                    tmem_ld_wait();                                  // A: keys 0-63 of tile j
                    tmem_ld_32x32b_x32(t_s + 64u, B0);
                    tmem_ld_32x32b_x32(t_s + 96u, B1);
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        m = fmaxf(m, fmaxf(__uint_as_float(A0[i]), __uint_as_float(A1[i])));
                    tmem_ld_wait();                                  // B: every load of tile j is in
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&s_free[sb]);
                    if (j + 1 < nkb) {
                        const int sn = (int)((T + 1) % kSBuf);
                        mbar_wait(&s_full[sn], ((T + 1) / kSBuf) & 1u);
                        tcgen05_fence_after();
                        const uint32_t t_n = tmem_base + (uint32_t)(sn * 128) + lane_sel;
                        tmem_ld_32x32b_x32(t_n, A0);
                        tmem_ld_32x32b_x32(t_n + 32u, A1);
                    }
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        m = fmaxf(m, fmaxf(__uint_as_float(B0[i]), __uint_as_float(B1[i])));
                }
            }
            const float mc = m * scale_log2e;
            float l = 0.f;
            {   // pass 2: probabilities; chunks of 32 keys ping-pong between RA and RB
                uint32_t RA[32], RB[32];
                // exp2, row sum, bf16 pack and the swizzled store of one 32-key chunk
                auto emit = [&](const uint32_t* r, uint8_t* prow, int c) {
                    uint32_t pk[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float p0 = ex2_approx(fmaf(__uint_as_float(r[2 * i]), scale_log2e, -mc));
                        const float p1 = ex2_approx(fmaf(__uint_as_float(r[2 * i + 1]), scale_log2e, -mc));
                        l += p0 + p1;
                        const __nv_bfloat162 h2 = __floats2bfloat162_rn(p0, p1);
                        pk[i] = *reinterpret_cast<const uint32_t*>(&h2);
                    }
                    uint8_t* box = prow + (c >> 6) * kTile;      // keys 0-63 | 64-127
                    const uint32_t ch0 = (uint32_t)(c & 63) >> 3;  // first 16-byte chunk: 0 or 4
#pragma unroll
                    for (uint32_t q = 0; q < 4; ++q)
                        *reinterpret_cast<uint4*>(box + (((ch0 + q) ^ sw) << 4)) =
                            make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
                };
                {
                    const int sb = (int)(T % kSBuf);
                    mbar_wait(&s_full[sb], (T / kSBuf) & 1u);
                    tcgen05_fence_after();
                    tmem_ld_32x32b_x32(tmem_base + (uint32_t)(sb * 128) + lane_sel, RA);
                }
                for (int j = 0; j < nkb; ++j, ++T, ++J) {
                    const int sb = (int)(T % kSBuf);
                    const int pb = (int)(J & 1u);
                    const uint32_t t_s = tmem_base + (uint32_t)(sb * 128) + lane_sel;
                    uint8_t* prow = sp + pb * 2 * kTile + (uint32_t)row * 128u;
                    tmem_ld_wait();                                  // RA: keys 0-31
                    tmem_ld_32x32b_x32(t_s + 32u, RB);
                    // the P V product that last read this P buffer must have finished
                    if (J >= 2) mbar_wait(&p_free[pb], ((J >> 1) - 1u) & 1u);
                    emit(RA, prow, 0);
                    tmem_ld_wait();                                  // RB: keys 32-63
                    tmem_ld_32x32b_x32(t_s + 64u, RA);
                    emit(RB, prow, 32);
                    tmem_ld_wait();                                  // RA: keys 64-95
                    tmem_ld_32x32b_x32(t_s + 96u, RB);
                    emit(RA, prow, 64);
                    tmem_ld_wait();                                  // RB: every load of tile j is in
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&s_free[sb]);
                    if (j + 1 < nkb) {
                        const int sn = (int)((T + 1) % kSBuf);
                        mbar_wait(&s_full[sn], ((T + 1) / kSBuf) & 1u);
                        tcgen05_fence_after();
                        tmem_ld_32x32b_x32(tmem_base + (uint32_t)(sn * 128) + lane_sel, RA);
                    }
                    emit(RB, prow, 96);
                    fence_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&p_full[pb]);
                }
            }
            // ---- O / l -> bf16 -> swizzled box -> TMA store; lse ----
            mbar_wait(o_done, (uint32_t)it & 1u);
            tcgen05_fence_after();
            uint32_t ro[64];
            tmem_ld_32x32b_x32(tmem_o + lane_sel, ro);
            tmem_ld_32x32b_x32(tmem_o + lane_sel + 32u, ro + 32);
            tmem_ld_wait();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(o_free);
                bulk_wait_read<0>();          // the previous tile's store has drained the box
            }
            __syncwarp();
            const float inv = 1.f / l;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                float f[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(ro[8 * q + i]) * inv;
                bf16x8 v = f32_to_bf16x8(f);
                *reinterpret_cast<uint4*>(obox + (uint32_t)lane * 128u +
                                          (((uint32_t)q ^ (uint32_t)(lane & 7)) << 4)) =
                    *reinterpret_cast<uint4*>(&v);
            }
            lse[((size_t)b * heads + head) * S + qb * kBQ + row] = mc + log2f(l);
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
                tma_store_2d(&maps.out, obox, head * kD, b * S + qb * kBQ + ew * 32);
                bulk_commit();
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
                     "r"((uint32_t)kFwdTmemCols)
                     : "memory");
    }
}

// ---------------------------------------------------------------------------------------------
// Backward. One CTA per (128 keys, head, batch) - K_j and V_j stay in shared memory - looping over
// the query blocks i. Everything is computed TRANSPOSED (keys on the 128 TMEM lanes, queries on
// the columns), which makes every operand of the five products readable as it already sits in
// shared memory:
//   S^T  = K_j Q_i^T            A = K_j  (K-major)   B = Q_i  (K-major)
//   dP^T = V_j dO_i^T           A = V_j  (K-major)   B = dO_i (K-major)
//   threads: P^T = exp2(S^T c - lse_i), dS^T = P^T (dP^T - delta_i) / 8  -> bf16 tiles, K-major
//   dV_j += P^T  dO_i           A = P^T  (K-major)   B = dO_i (MN-major, as stored)
//   dK_j += dS^T Q_i            A = dS^T (K-major)   B = Q_i  (MN-major, as stored)
//   dQ_i  = dS   K_j            A = dS^T read MN-major (it IS dS stored [key][query])
//                                                    B = K_j  (MN-major, as stored)
// Each query block is processed as two halves of 64 queries that ping-pong between the tensor core
// and the threads (the first version, whole blocks in strict MMA -> threads -> MMA turns, kept
// the tensor pipe 14 % and the issue slots 21 % busy).
// dV_j / dK_j accumulate in TMEM over i and leave through TMA stores into the packed gradient
// dqkv[B*S, 3H]; the dQ_i tiles of the S/128 CTAs of one (batch, head) meet in an fp32 buffer by
// TMA reduce-add (attn_dq_convert_kernel rounds it into dqkv afterwards).
constexpr int kBwdThreads = 384;   // warps 0-3 as in the GEMM, 4-7 / 8-11: one thread group per half
constexpr int kBwdTmemCols = 512;   // S^T [0,128) dP^T [128,256) dV [256,320) dK [320,384) dQ [384,448)
constexpr int kBwdSmem = 2 * kTile /*K V*/ + 4 * kTile /*Q dO x2*/ + 2 * kTile /*P^T*/ +
                         2 * kTile /*dS^T*/ + 8 * 4096 /*staging*/ + 2 * 2 * 512 /*lse delta*/ +
                         256 + 1024;
static_assert(kBwdSmem <= 227 * 1024, "shared memory budget of one CTA");

__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]),
          "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}

struct AttnBwdMaps {
    CUtensorMap qkv;    // [B*S, 3H] bf16, box {64, 128}: Q, K, V loads
    CUtensorMap dout;   // [B*S, H] bf16, box {64, 128}: dO loads
    CUtensorMap dqkv;   // [B*S, 3H] bf16, box {64, 32}: dK, dV stores
    CUtensorMap dq32;   // [B*S, H] fp32, box {32, 32}: dQ reduce-add
};

// MN-major operand whose 64-element halves along MN are `lbo` bytes apart (tiles of 128 k-rows)
__device__ __forceinline__ uint64_t make_smem_desc_mn_lbo(uint32_t saddr, uint32_t lbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(lbo >> 4) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_kernel(const __grid_constant__ AttnBwdMaps maps, const float* __restrict__ lse,
                const float* __restrict__ delta, int S, int H, int heads, float scale_log2e,
                float scale) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sk = smem;
    uint8_t* sv = sk + kTile;
    uint8_t* sqd = sv + kTile;              // 2 stages x (Q_i, dO_i)
    uint8_t* spt = sqd + 4 * kTile;         // P^T   [128 keys][128 queries] as two 64-query boxes
    uint8_t* sdst = spt + 2 * kTile;        // dS^T  same layout
    uint8_t* sstage = sdst + 2 * kTile;     // 4 x 4 KB store staging
    float* svec = (float*)(sstage + 8 * 4096);   // [2 stages][lse | delta][128]
    uint64_t* bars = (uint64_t*)((uint8_t*)svec + 2 * 2 * 512);
    uint64_t* kv_full = bars;
    uint64_t* qd_full = bars + 1;           // [2]
    uint64_t* qd_empty = qd_full + 2;       // [2]
    uint64_t* st_full = qd_empty + 2;       // [2] MMA -> threads: S^T / dP^T of half h ready
    uint64_t* st_free = st_full + 2;        // [2] threads -> MMA: that half has been read
    uint64_t* pt_full = st_free + 2;        // [2] threads -> MMA: P^T / dS^T box h written
    uint64_t* g2_done = pt_full + 2;        // MMA -> threads: dV, dK, dQ products of block i done
    uint32_t* tmem_slot = (uint32_t*)(g2_done + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int jb = blockIdx.x, head = blockIdx.y, b = blockIdx.z;
    const int krow0 = b * S + jb * kBK;
    const int nqb = S / kBQ;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.qkv) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.dout) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.dqkv) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&maps.dq32) : "memory");
    }
    if (warp == 1 && lane == 0) {
        mbar_init(kv_full, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&qd_full[i], 1);
            mbar_init(&qd_empty[i], 1);
            mbar_init(&st_full[i], 1);
            mbar_init(&st_free[i], 4);
            mbar_init(&pt_full[i], 4);
        }
        mbar_init(g2_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)),
                     "r"((uint32_t)kBwdTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t t_st = tmem_base, t_dpt = tmem_base + 128u, t_dv = tmem_base + 256u,
                   t_dk = tmem_base + 320u, t_dq = tmem_base + 384u;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            mbar_expect_tx(kv_full, 2 * kTile);
            tma_load_2d(&maps.qkv, kv_full, sk, H + head * kD, krow0);
            tma_load_2d(&maps.qkv, kv_full, sv, 2 * H + head * kD, krow0);
            for (int i = 0; i < nqb; ++i) {
                const int st = i & 1;
                mbar_wait(&qd_empty[st], ((uint32_t)(i >> 1) & 1u) ^ 1u);
                mbar_expect_tx(&qd_full[st], 2 * kTile);
                tma_load_2d(&maps.qkv, &qd_full[st], sqd + st * 2 * kTile, head * kD, b * S + i * kBQ);
                tma_load_2d(&maps.dout, &qd_full[st], sqd + st * 2 * kTile + kTile, head * kD,
                            b * S + i * kBQ);
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            const uint32_t id_s = make_idesc(128, 64);                  // both K-major, 64 queries
            const uint32_t id_acc = make_idesc_major(128, kD, 0, 1);    // A K-major, B MN-major
            const uint32_t id_dq = make_idesc_major(128, kD, 1, 1);     // both MN-major
            const uint32_t k_addr = smem_u32(sk), v_addr = smem_u32(sv);
            const uint32_t pt_addr = smem_u32(spt), dst_addr = smem_u32(sdst);
            mbar_wait(kv_full, 0);
            // S^T / dP^T of one HALF (64 queries) of block i: the halves ping-pong, so the
            // threads work on one while the tensor core produces the other
            auto issue_s = [&](int i, int h) {
                const int st = i & 1;
                if (h == 0) mbar_wait(&qd_full[st], (uint32_t)(i >> 1) & 1u);
                if (i > 0) mbar_wait(&st_free[h], (uint32_t)(i - 1) & 1u);
                tcgen05_fence_after();
                const uint32_t q_addr = smem_u32(sqd + st * 2 * kTile) + h * 8192, do_addr = q_addr + kTile;
#pragma unroll
                for (int k = 0; k < kD / UMMA_K; ++k)
                    umma_bf16(t_st + (uint32_t)(h * 64), make_smem_desc(k_addr + k * 32),
                              make_smem_desc(q_addr + k * 32), id_s, k != 0 ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < kD / UMMA_K; ++k)
                    umma_bf16(t_dpt + (uint32_t)(h * 64), make_smem_desc(v_addr + k * 32),
                              make_smem_desc(do_addr + k * 32), id_s, k != 0 ? 1u : 0u);
                umma_commit(&st_full[h]);
            };
            issue_s(0, 0);
            issue_s(0, 1);
            for (int i = 0; i < nqb; ++i) {
                const int st = i & 1;
                const uint32_t q_addr = smem_u32(sqd + st * 2 * kTile), do_addr = q_addr + kTile;
                for (int h = 0; h < 2; ++h) {
                    mbar_wait(&pt_full[h], (uint32_t)i & 1u);
                    tcgen05_fence_after();
#pragma unroll
                    for (int k = 0; k < 64 / UMMA_K; ++k) {
                        // K-major A: box h holds queries h*64 .. h*64+63; MN-major B: those rows
                        const uint32_t aoff = h * kTile + k * 32, boff = h * 8192 + k * 2048;
                        umma_bf16(t_dv, make_smem_desc(pt_addr + aoff), make_smem_desc_mn(do_addr + boff),
                                  id_acc, (i | h | k) != 0 ? 1u : 0u);
                        umma_bf16(t_dk, make_smem_desc(dst_addr + aoff), make_smem_desc_mn(q_addr + boff),
                                  id_acc, (i | h | k) != 0 ? 1u : 0u);
                    }
                    if (h == 1) {
#pragma unroll
                        for (int k = 0; k < 128 / UMMA_K; ++k)
                            umma_bf16(t_dq, make_smem_desc_mn_lbo(dst_addr + k * 2048, kTile),
                                      make_smem_desc_mn(k_addr + k * 2048), id_dq, k != 0 ? 1u : 0u);
                        umma_commit(&qd_empty[st]);
                        umma_commit(g2_done);
                    }
                    // the next block's half h: its S^T / dP^T columns were read before pt_full[h]
                    if (i + 1 < nqb) issue_s(i + 1, h);
                }
            }
        }
    } else if (warp >= kEpiWarp0) {
        // ===================== one thread per key row =====================
        // Two groups of four warps: group h owns the 64-query half h of every block (its S^T /
        // dP^T columns, its P^T / dS^T box), so both halves are in flight at once and every SM
        // sub-partition has two warps to hide latencies with.
        const int grp = (warp - kEpiWarp0) >> 2;
        const int ew = warp & 3;                     // TMEM lane quadrant
        const int row = ew * 32 + lane;
        const uint32_t lane_sel = (uint32_t)(ew * 32) << 16;
        const uint32_t sw = (uint32_t)(row & 7);
        uint8_t* ptrow = spt + (uint32_t)row * 128u;
        uint8_t* dstrow = sdst + (uint32_t)row * 128u;
        uint8_t* stage = sstage + (grp * 4 + ew) * 4096;
        const size_t vec0 = ((size_t)b * heads + head) * S;
        // fp32 [32 rows x 32 cols] of a TMEM accumulator -> swizzled staging box -> TMA
        auto flush_f32 = [&](uint32_t taddr, int col, int grow) {
            {
                const int c = grp * 32;              // this group's 32 of the 64 columns
                uint32_t r[32];
                tmem_ld_32x32b_x32(taddr + lane_sel + (uint32_t)c, r);
                tmem_ld_wait();
                if (lane == 0) bulk_wait_read<0>();
                __syncwarp();
#pragma unroll
                for (uint32_t q = 0; q < 8; ++q)
                    *reinterpret_cast<uint4*>(stage + (uint32_t)lane * 128u + ((q ^ (uint32_t)(lane & 7)) << 4)) =
                        make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
                fence_async_smem();
                __syncwarp();
                if (lane == 0) {
                    tma_reduce_add_2d(&maps.dq32, stage, col + c, grow);
                    bulk_commit();
                }
            }
        };
        // lse (group 0) / delta (group 1) of the NEXT block are fetched while this block is worked
        // on; S^T / dP^T come out of TMEM in chunks of 16 queries, one chunk in flight while the
        // previous one is turned into P^T / dS^T (see the forward kernel).
        float vnext = grp == 0 ? lse[vec0 + row] : delta[vec0 + row];
        for (int i = 0; i < nqb; ++i) {
            float* vec = svec + (i & 1) * 256;
            vec[grp * 128 + row] = vnext;
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (i + 1 < nqb)
                vnext = grp == 0 ? lse[vec0 + (i + 1) * kBQ + row] : delta[vec0 + (i + 1) * kBQ + row];
            {
                const int h = grp;
                if (i > 0) {
                    // block i-1 is finished on the tensor core: its dQ tile can leave, and the
                    // P^T / dS^T boxes may be overwritten. (This wait comes BEFORE the one on
                    // st_full: S^T of half 1 is issued behind block i-1's dQ product, so the
                    // flush overlaps that MMA; the other order cost 15 us per call.)
                    mbar_wait(g2_done, (uint32_t)(i - 1) & 1u);
                    tcgen05_fence_after();
                    flush_f32(t_dq, head * kD, b * S + (i - 1) * kBQ + ew * 32);
                }
                mbar_wait(&st_full[h], (uint32_t)i & 1u);
                tcgen05_fence_after();
                const uint32_t boff = (uint32_t)h * kTile;
                uint32_t sa[16], da[16], sb2[16], db2[16];
                // P^T and dS^T of 16 queries: exp2, the dS product, bf16 pack, swizzled stores
                auto emit = [&](const uint32_t* rs, const uint32_t* rd, int c) {
                    const int qc = h * 64 + c;
                    uint32_t pk[8], dk[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const float2 l2 = *reinterpret_cast<const float2*>(vec + qc + 2 * k);
                        const float2 d2 = *reinterpret_cast<const float2*>(vec + 128 + qc + 2 * k);
                        const float p0 = ex2_approx(fmaf(__uint_as_float(rs[2 * k]), scale_log2e, -l2.x));
                        const float p1 = ex2_approx(fmaf(__uint_as_float(rs[2 * k + 1]), scale_log2e, -l2.y));
                        const float s0 = p0 * (__uint_as_float(rd[2 * k]) - d2.x) * scale;
                        const float s1 = p1 * (__uint_as_float(rd[2 * k + 1]) - d2.y) * scale;
                        const __nv_bfloat162 ph = __floats2bfloat162_rn(p0, p1);
                        const __nv_bfloat162 dh = __floats2bfloat162_rn(s0, s1);
                        pk[k] = *reinterpret_cast<const uint32_t*>(&ph);
                        dk[k] = *reinterpret_cast<const uint32_t*>(&dh);
                    }
                    const uint32_t ch0 = (uint32_t)c >> 3;
#pragma unroll
                    for (uint32_t q = 0; q < 2; ++q) {
                        const uint32_t off = boff + (((ch0 + q) ^ sw) << 4);
                        *reinterpret_cast<uint4*>(ptrow + off) =
                            make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
                        *reinterpret_cast<uint4*>(dstrow + off) =
                            make_uint4(dk[4 * q], dk[4 * q + 1], dk[4 * q + 2], dk[4 * q + 3]);
                    }
                };
                const uint32_t a_st = t_st + lane_sel + (uint32_t)(h * 64);
                const uint32_t a_dp = t_dpt + lane_sel + (uint32_t)(h * 64);
                tmem_ld_32x32b_x16(a_st, sa);
                tmem_ld_32x32b_x16(a_dp, da);
                tmem_ld_wait();
                tmem_ld_32x32b_x16(a_st + 16u, sb2);
                tmem_ld_32x32b_x16(a_dp + 16u, db2);
                emit(sa, da, 0);
                tmem_ld_wait();
                tmem_ld_32x32b_x16(a_st + 32u, sa);
                tmem_ld_32x32b_x16(a_dp + 32u, da);
                emit(sb2, db2, 16);
                tmem_ld_wait();
                tmem_ld_32x32b_x16(a_st + 48u, sb2);
                tmem_ld_32x32b_x16(a_dp + 48u, db2);
                emit(sa, da, 32);
                tmem_ld_wait();
                // every load of this half is in: the tensor core may overwrite S^T / dP^T
                tcgen05_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&st_free[h]);
                emit(sb2, db2, 48);
                fence_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&pt_full[h]);
            }
        }
        mbar_wait(g2_done, (uint32_t)(nqb - 1) & 1u);
        tcgen05_fence_after();
        flush_f32(t_dq, head * kD, b * S + (nqb - 1) * kBQ + ew * 32);
        // dV_j, dK_j: fp32 TMEM -> bf16 -> swizzled box -> TMA store into the packed gradient
        {
            const int which = grp;                   // group 0 stores dV_j, group 1 dK_j
            const uint32_t taddr = (which == 0 ? t_dv : t_dk) + lane_sel;
            if (lane == 0) bulk_wait_read<0>();
            __syncwarp();
#pragma unroll 1
            for (int c = 0; c < kD; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(taddr + (uint32_t)c, r);
                tmem_ld_wait();
                float f[32];
#pragma unroll
                for (int k = 0; k < 32; ++k) f[k] = __uint_as_float(r[k]);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    bf16x8 v = f32_to_bf16x8(f + 8 * q);
                    *reinterpret_cast<uint4*>(stage + (uint32_t)lane * 128u +
                                              ((((uint32_t)(c >> 3) + q) ^ (uint32_t)(lane & 7)) << 4)) =
                        *reinterpret_cast<uint4*>(&v);
                }
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
                tma_store_2d(&maps.dqkv, stage, (which == 0 ? 2 * H : H) + head * kD, krow0 + ew * 32);
                bulk_commit();
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
                     "r"((uint32_t)kBwdTmemCols)
                     : "memory");
    }
}

// delta[b, head, s] = sum_d dO[row, head*64 + d] * O[row, head*64 + d]  (8 lanes per (row, head))
__global__ void attn_delta_kernel(const __nv_bfloat16* __restrict__ dout,
                                  const __nv_bfloat16* __restrict__ out, float* __restrict__ delta,
                                  int B, int S, int heads) {
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long item = t >> 3;                       // (row, head)
    const int part = (int)(t & 7);
    const long n_items = (long)B * S * heads;
    float acc = 0.f;
    long row = 0;
    int head = 0;
    if (item < n_items) {
        row = item / heads;
        head = (int)(item - row * heads);
        const size_t off = (size_t)row * heads * kD + head * kD + part * 8;
        float a[8], o[8];
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(dout + off), a);
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(out + off), o);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc = fmaf(a[k], o[k], acc);
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    if (item < n_items && part == 0) {
        const long b = row / S, s = row - b * S;
        delta[((size_t)b * heads + head) * S + s] = acc;
    }
}

// dqkv[row, 0:H] = bf16(dq32[row, :])   (8 elements per thread)
__global__ void attn_dq_convert_kernel(const float* __restrict__ dq32, __nv_bfloat16* __restrict__ dqkv,
                                       long rows, int H) {
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int per_row = H / 8;
    if (t >= rows * per_row) return;
    const long row = t / per_row;
    const int c = (int)(t - row * per_row) * 8;
    const float4 a = *reinterpret_cast<const float4*>(dq32 + row * H + c);
    const float4 b = *reinterpret_cast<const float4*>(dq32 + row * H + c + 4);
    const float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    bf16x8 v = f32_to_bf16x8(f);
    *reinterpret_cast<uint4*>(dqkv + row * 3 * H + c) = *reinterpret_cast<uint4*>(&v);
}

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

// 2-D row-major [rows, cols] tensor, box {box_cols, box_rows}
int map2d(CUtensorMap* map, const void* base, long rows, long cols, int box_cols, int box_rows,
          CUtensorMapDataType dt, int elt, CUtensorMapSwizzle swz, const char* what) {
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * elt};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_enc(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char msg[128];
        snprintf(msg, sizeof(msg), "cuTensorMapEncodeTiled(%s) failed with CUresult %d", what, (int)r);
        tfop_set_error(__FILE__, __LINE__, msg);
        return -1;
    }
    return 0;
}

int check_attn(const void* a, const void* b, int B, int S, int heads, int d) {
    TFOP_REQUIRE(B > 0 && heads > 0, "attention: empty problem");
    TFOP_REQUIRE(d == kD, "attention: head dimension must be 64");
    TFOP_REQUIRE(S > 0 && S % 128 == 0, "attention: sequence length must be a multiple of 128");
    TFOP_REQUIRE(((((uintptr_t)a) | ((uintptr_t)b)) & 15) == 0, "attention: tensors must be 16-byte aligned");
    return 0;
}

}  // namespace

// out[B*S, H] = softmax(Q K^T / sqrt(64)) V per (batch, head), H = heads * 64, for the packed
// qkv[B*S, 3H] = [Q | K | V] (bf16, row-major). lse[B, heads, S] (fp32) receives the base-2
// log-sum-exp of the scaled scores, which tfop_attn_bwd needs.
TFOP_API int tfop_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int heads,
                           int head_dim, cudaStream_t stream) {
    if (check_attn(qkv, out, B, S, heads, head_dim) != 0) return -1;
    if (get_enc() != 0) return -1;
    const int H = heads * kD;
    AttnMaps maps;
    if (map2d(&maps.qkv, qkv, (long)B * S, 3L * H, 64, 128, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
              CU_TENSOR_MAP_SWIZZLE_128B, "qkv"))
        return -1;
    if (map2d(&maps.out, out, (long)B * S, H, 64, 32, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
              CU_TENSOR_MAP_SWIZZLE_128B, "out"))
        return -1;
    TFOP_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   kFwdSmem));
    const float scale_log2e = 0.125f * 1.4426950408889634f;
    int dev = 0, sms = 0;
    TFOP_CUDA(cudaGetDevice(&dev));
    TFOP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int n_items = B * heads * (S / kBQ);
    attn_fwd_kernel<<<n_items < sms ? n_items : sms, kFwdThreads, kFwdSmem, stream>>>(
        maps, lse, S, H, heads, n_items, scale_log2e);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// Gradient of tfop_attn_fwd. dqkv[B*S, 3H] (bf16) receives [dQ | dK | dV] in the layout of qkv,
// i.e. it IS the gradient of the packed projection output. `out` and `lse` are the forward's
// results. Scratch (caller-provided, so the call allocates nothing): delta[B, heads, S] fp32 and
// dq32[B*S, H] fp32; dq32 is zeroed here and holds the unrounded dQ afterwards.
// (Letting the rounding kernel hand dq32 back zeroed instead of clearing it here was tried: its
// zero stores made that kernel 27 us slower to save a 7 us memset.)
TFOP_API int tfop_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse,
                           void* dqkv, float* delta, float* dq32, int B, int S, int heads,
                           int head_dim, cudaStream_t stream) {
    if (check_attn(qkv, dqkv, B, S, heads, head_dim) != 0) return -1;
    TFOP_REQUIRE(((((uintptr_t)out) | ((uintptr_t)dout) | ((uintptr_t)dq32)) & 15) == 0,
                 "attention: tensors must be 16-byte aligned");
    if (get_enc() != 0) return -1;
    const int H = heads * kD;
    const long rows = (long)B * S;
    AttnBwdMaps maps;
    if (map2d(&maps.qkv, qkv, rows, 3L * H, 64, 128, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
              CU_TENSOR_MAP_SWIZZLE_128B, "qkv"))
        return -1;
    if (map2d(&maps.dout, dout, rows, H, 64, 128, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
              CU_TENSOR_MAP_SWIZZLE_128B, "dout"))
        return -1;
    if (map2d(&maps.dqkv, dqkv, rows, 3L * H, 64, 32, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
              CU_TENSOR_MAP_SWIZZLE_128B, "dqkv"))
        return -1;
    if (map2d(&maps.dq32, dq32, rows, H, 32, 32, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4,
              CU_TENSOR_MAP_SWIZZLE_128B, "dq32"))
        return -1;
    TFOP_CUDA(cudaMemsetAsync(dq32, 0, (size_t)rows * H * 4, stream));
    {
        const long threads = rows * heads * 8;
        attn_delta_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(
            (const __nv_bfloat16*)dout, (const __nv_bfloat16*)out, delta, B, S, heads);
    }
    TFOP_CUDA(cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   kBwdSmem));
    const float scale = 0.125f;
    attn_bwd_kernel<<<dim3(S / kBK, heads, B), kBwdThreads, kBwdSmem, stream>>>(
        maps, lse, delta, S, H, heads, scale * 1.4426950408889634f, scale);
    {
        const long threads = rows * (H / 8);
        attn_dq_convert_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(
            dq32, (__nv_bfloat16*)dqkv, rows, H);
    }
    TFOP_CUDA(cudaGetLastError());
    return 0;
}
