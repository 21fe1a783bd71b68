// K4 / K5: parameter-server embedding over NVLink peer memory.
//
// The embedding table is sharded over the PS replicas' HBM (row r lives on shard r % S at
// local row r / S) inside their symmetric buffers. Workers never talk to a PS *process*:
//
//   K4  ps_gather_rows   : a worker reads the rows it needs straight out of the owning
//                          shard's HBM with P2P loads (the "pull").
//   K5  ps_push_adagrad  : a worker reduces the gradients of duplicate ids (segment sum
//                          over the sorted id list) and applies Adagrad ON THE OWNING
//                          SHARD with vector fp32 atomics over NVLink:
//                              acc += g^2 ;  w -= lr * g / (sqrt(acc) + eps)
//                          one kernel, no staging copy, asynchronous across workers like the
//                          reference's async PS training.
//
// Reference parity: the variable pull / gradient push + optimizer-on-PS of
// examples/v1/dist-mnist/dist_mnist.py:161-208 (TF gRPC RecvTensor + CPU Adam in the
// reference stack); BASELINE.json config 4 (wide-and-deep, 10M-row sharded embedding,
// 2 PS + 6 workers).
#include "common.cuh"

#define TFOP_PS_MAX_SHARDS 8

struct tfop_ps_table {
    float* weight[TFOP_PS_MAX_SHARDS];   // shard base pointers as mapped in THIS process
    float* accum[TFOP_PS_MAX_SHARDS];    // Adagrad accumulators, same shape
    int num_shards;
    int dim;                             // multiple of 4
    long long rows_per_shard;
};

namespace {

__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
    float4 r;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
// fetch-and-add of 4 floats, system scope (returns the OLD values)
__device__ __forceinline__ float4 atom_add_f4_sys(float* p, float4 v) {
    float4 r;
    asm volatile("atom.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4], {%5,%6,%7,%8};"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
    return r;
}
__device__ __forceinline__ void red_add_f4_sys(float* p, float4 v) {
    asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
                 "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// One thread per (id, 4-float chunk): lanes of a warp cover consecutive chunks of the same
// row first, so a row of dim 64 is one 256-byte coalesced request.
template <bool OUT_BF16>
__global__ void __launch_bounds__(256) ps_gather_kernel(const tfop_ps_table t,
                                                        const long long* __restrict__ ids,
                                                        size_t n_ids, void* __restrict__ out) {
    const int chunks = t.dim >> 2;
    const size_t total = n_ids * (size_t)chunks;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const size_t k = i / chunks;
        const int c = (int)(i % chunks);
        const long long id = ids[k];
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (id >= 0) {
            const int shard = (int)(id % t.num_shards);
            const long long row = id / t.num_shards;
            if (row < t.rows_per_shard)
                v = ld_peer_f4(t.weight[shard] + (size_t)row * t.dim + (size_t)c * 4);
        }
        if (OUT_BF16) {
            __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
            __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&a);
            pk.y = *reinterpret_cast<uint32_t*>(&b);
            reinterpret_cast<uint2*>(out)[i] = pk;
        } else {
            reinterpret_cast<float4*>(out)[i] = v;
        }
    }
}

// One thread per (unique id, 4-float chunk). seg_off[u]..seg_off[u+1] are the positions in the
// SORTED id list that carry unique id u; perm maps a sorted position to the row of `grad`.
template <bool GRAD_BF16>
__global__ void __launch_bounds__(256) ps_push_adagrad_kernel(
    const tfop_ps_table t, const long long* __restrict__ uniq_ids,
    const long long* __restrict__ seg_off, const long long* __restrict__ perm, size_t n_uniq,
    const void* __restrict__ grad, float lr, float eps, float grad_scale) {
    const int chunks = t.dim >> 2;
    const size_t total = n_uniq * (size_t)chunks;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const size_t u = i / chunks;
        const int c = (int)(i % chunks);
        const long long id = uniq_ids[u];
        if (id < 0) continue;
        const int shard = (int)(id % t.num_shards);
        const long long row = id / t.num_shards;
        if (row >= t.rows_per_shard) continue;
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        for (long long s = seg_off[u]; s < seg_off[u + 1]; ++s) {      // segment sum
            const size_t src = (size_t)perm[s] * t.dim + (size_t)c * 4;
            if (GRAD_BF16) {
                const uint2 raw = *reinterpret_cast<const uint2*>((const __nv_bfloat16*)grad + src);
                const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
                const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
                g.x += a.x; g.y += a.y; g.z += b.x; g.w += b.y;
            } else {
                const float4 v = *reinterpret_cast<const float4*>((const float*)grad + src);
                g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
            }
        }
        g.x *= grad_scale; g.y *= grad_scale; g.z *= grad_scale; g.w *= grad_scale;
        const size_t off = (size_t)row * t.dim + (size_t)c * 4;
        const float4 sq = make_float4(g.x * g.x, g.y * g.y, g.z * g.z, g.w * g.w);
        const float4 old = atom_add_f4_sys(t.accum[shard] + off, sq);    // on the owning shard
        float4 d;
        d.x = -lr * g.x / (sqrtf(old.x + sq.x) + eps);
        d.y = -lr * g.y / (sqrtf(old.y + sq.y) + eps);
        d.z = -lr * g.z / (sqrtf(old.z + sq.z) + eps);
        d.w = -lr * g.w / (sqrtf(old.w + sq.w) + eps);
        red_add_f4_sys(t.weight[shard] + off, d);
    }
}

// ---- K5b: duplicate ids merged on the device, without a sort ---------------------------------
// A worker's batch names the same row many times (hot ids). Adagrad is not linear in the
// gradient, so the gradients of one id must be summed before the update. Four short kernels on
// one stream, no host round trip, no allocation, capturable in a CUDA graph:
//   1 insert  : every position puts its id into an open-addressing table (64-bit CAS, linear
//               probing) and remembers the slot it landed in;
//   2 compact : every occupied slot takes the next dense index, publishes (id -> index) and
//               empties itself again (the table is clean for the next call);
//   3 sum     : every position adds its gradient row into row `index` of an fp32 scratch
//               (vector red.add, duplicates meet here);
//   4 apply   : every dense index applies Adagrad on the owning shard over NVLink and zeroes
//               its scratch row again.
// The scratch (keys / slot_of / dense / uniq / count / sums) is the caller's, sized by
// tfop_ps_dedup_workspace, and must be zero-filled except keys (all ones) before first use.
constexpr unsigned long long kEmptyKey = ~0ull;

__device__ __forceinline__ unsigned hash_id(unsigned long long id) {
    id ^= id >> 33;
    id *= 0xff51afd7ed558ccdull;
    id ^= id >> 33;
    return (unsigned)id;
}

__global__ void __launch_bounds__(256) ps_dedup_insert_kernel(
    const long long* __restrict__ ids, size_t n_ids, unsigned long long* __restrict__ keys,
    unsigned cap_mask, int* __restrict__ slot_of, int* __restrict__ count) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    if (blockIdx.x == 0 && threadIdx.x == 0) *count = 0;          // kernel 2 counts from zero
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n_ids; k += stride) {
        const long long id = ids[k];
        int where = -1;
        if (id >= 0) {
            unsigned slot = hash_id((unsigned long long)id) & cap_mask;
            for (unsigned probe = 0; probe <= cap_mask; ++probe) {
                const unsigned long long seen =
                    atomicCAS(keys + slot, kEmptyKey, (unsigned long long)id);
                if (seen == kEmptyKey || seen == (unsigned long long)id) {
                    where = (int)slot;
                    break;
                }
                slot = (slot + 1) & cap_mask;
            }
        }
        slot_of[k] = where;
    }
}

__global__ void __launch_bounds__(256) ps_dedup_compact_kernel(
    unsigned long long* __restrict__ keys, unsigned cap, int* __restrict__ dense,
    long long* __restrict__ uniq, int* __restrict__ count) {
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned s = blockIdx.x * blockDim.x + threadIdx.x; s < cap; s += stride) {
        const unsigned long long k = keys[s];
        const bool used = k != kEmptyKey;
        // one atomic per warp: lanes with an occupied slot take consecutive indices
        const unsigned m = __ballot_sync(0xffffffffu, used);
        int base = 0;
        const int lane = threadIdx.x & 31;
        if (m != 0u) {
            const int leader = __ffs(m) - 1;
            if (lane == leader) base = atomicAdd(count, __popc(m));
            base = __shfl_sync(0xffffffffu, base, leader);
        }
        if (used) {
            const int idx = base + __popc(m & ((1u << lane) - 1u));
            dense[s] = idx;
            uniq[idx] = (long long)k;
            keys[s] = kEmptyKey;
        }
    }
}

template <bool GRAD_BF16>
__global__ void __launch_bounds__(256) ps_dedup_sum_kernel(
    const int* __restrict__ slot_of, const int* __restrict__ dense, size_t n_ids, int dim,
    const void* __restrict__ grad, float* __restrict__ sums) {
    const int chunks = dim >> 2;
    const size_t total = n_ids * (size_t)chunks;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const size_t k = i / chunks;
        const int c = (int)(i % chunks);
        const int slot = slot_of[k];
        if (slot < 0) continue;
        const size_t src = k * dim + (size_t)c * 4;
        float4 g;
        if (GRAD_BF16) {
            const uint2 raw = *reinterpret_cast<const uint2*>((const __nv_bfloat16*)grad + src);
            const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
            const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
            g = make_float4(a.x, a.y, b.x, b.y);
        } else {
            g = *reinterpret_cast<const float4*>((const float*)grad + src);
        }
        float* dst = sums + (size_t)dense[slot] * dim + (size_t)c * 4;
        asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst),
                     "f"(g.x), "f"(g.y), "f"(g.z), "f"(g.w)
                     : "memory");
    }
}

// kFetch: the accumulator update is a fetch-and-add (exact under concurrent pushes of the same
// row by several workers, one NVLink round trip per 16 bytes that nothing can overlap). !kFetch:
// the old accumulator is read with a plain peer load (pipelines like the gather), the two
// updates leave as fire-and-forget reductions; a concurrent push of the same row by another
// worker may then scale its step with an accumulator that misses this one's g^2 - the staleness
// asynchronous PS training has anyway (the reference's workers apply gradients computed on
// stale variables, examples/v1/dist-mnist/dist_mnist.py:161-208).
template <bool kFetch>
__global__ void __launch_bounds__(256) ps_dedup_apply_kernel(
    const tfop_ps_table t, const long long* __restrict__ uniq, const int* __restrict__ count,
    float* __restrict__ sums, float lr, float eps, float grad_scale) {
    const int chunks = t.dim >> 2;
    const size_t total = (size_t)(*count) * (size_t)chunks;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const size_t u = i / chunks;
        const int c = (int)(i % chunks);
        float4* sp = reinterpret_cast<float4*>(sums + u * t.dim + (size_t)c * 4);
        float4 g = *sp;
        *sp = make_float4(0.f, 0.f, 0.f, 0.f);                    // clean for the next call
        const long long id = uniq[u];
        const int shard = (int)(id % t.num_shards);
        const long long row = id / t.num_shards;
        if (row >= t.rows_per_shard) continue;
        g.x *= grad_scale; g.y *= grad_scale; g.z *= grad_scale; g.w *= grad_scale;
        const size_t off = (size_t)row * t.dim + (size_t)c * 4;
        const float4 sq = make_float4(g.x * g.x, g.y * g.y, g.z * g.z, g.w * g.w);
        float4 old;
        if (kFetch) {
            old = atom_add_f4_sys(t.accum[shard] + off, sq);
        } else {
            old = ld_peer_f4(t.accum[shard] + off);
            red_add_f4_sys(t.accum[shard] + off, sq);
        }
        float4 d;
        d.x = -lr * g.x / (sqrtf(old.x + sq.x) + eps);
        d.y = -lr * g.y / (sqrtf(old.y + sq.y) + eps);
        d.z = -lr * g.z / (sqrtf(old.z + sq.z) + eps);
        d.w = -lr * g.w / (sqrtf(old.w + sq.w) + eps);
        red_add_f4_sys(t.weight[shard] + off, d);
    }
}

int g_push_fetch = 0;   // tfop_ps_set_push_mode

size_t dedup_capacity(size_t n_ids) {
    size_t cap = 1024;
    while (cap < 2 * n_ids) cap <<= 1;
    return cap;
}

int grid_for(size_t total, int max_blocks) {
    size_t b = (total + 255) / 256;
    size_t cap = max_blocks > 0 ? (size_t)max_blocks : (size_t)148 * 8;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

int check_table(const tfop_ps_table* t) {
    TFOP_REQUIRE(t != nullptr, "null table");
    TFOP_REQUIRE(t->num_shards >= 1 && t->num_shards <= TFOP_PS_MAX_SHARDS, "1..8 shards");
    TFOP_REQUIRE(t->dim >= 4 && t->dim % 4 == 0, "embedding dim must be a multiple of 4");
    TFOP_REQUIRE(t->rows_per_shard > 0, "empty shard");
    return 0;
}

}  // namespace

// out[k, :] = table[ids[k], :] for k < n_ids (ids < 0 give zero rows); out is fp32 or bf16.
TFOP_API int tfop_ps_gather(const tfop_ps_table* table, const void* ids_i64, size_t n_ids,
                            void* out, int out_dtype, int max_blocks, cudaStream_t stream) {
    if (check_table(table) != 0) return -1;
    TFOP_REQUIRE(out_dtype == TFOP_F32 || out_dtype == TFOP_BF16, "bad out dtype");
    if (n_ids == 0) return 0;
    const size_t total = n_ids * (size_t)(table->dim / 4);
    const int grid = grid_for(total, max_blocks);
    if (out_dtype == TFOP_BF16)
        ps_gather_kernel<true><<<grid, 256, 0, stream>>>(*table, (const long long*)ids_i64, n_ids, out);
    else
        ps_gather_kernel<false><<<grid, 256, 0, stream>>>(*table, (const long long*)ids_i64, n_ids, out);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// Segment-sum the gradient rows of each unique id and apply Adagrad on the owning shard.
//   uniq_ids[n_uniq], seg_off[n_uniq+1], perm[n_sorted] : int64 (see kernel comment)
//   grad: [n_rows, dim] fp32 or bf16, the gradient w.r.t. the gathered rows
TFOP_API int tfop_ps_push_adagrad(const tfop_ps_table* table, const void* uniq_ids_i64,
                                  const void* seg_off_i64, const void* perm_i64, size_t n_uniq,
                                  const void* grad, int grad_dtype, float lr, float eps,
                                  float grad_scale, int max_blocks, cudaStream_t stream) {
    if (check_table(table) != 0) return -1;
    TFOP_REQUIRE(grad_dtype == TFOP_F32 || grad_dtype == TFOP_BF16, "bad grad dtype");
    if (n_uniq == 0) return 0;
    const size_t total = n_uniq * (size_t)(table->dim / 4);
    const int grid = grid_for(total, max_blocks);
    if (grad_dtype == TFOP_BF16)
        ps_push_adagrad_kernel<true><<<grid, 256, 0, stream>>>(
            *table, (const long long*)uniq_ids_i64, (const long long*)seg_off_i64,
            (const long long*)perm_i64, n_uniq, grad, lr, eps, grad_scale);
    else
        ps_push_adagrad_kernel<false><<<grid, 256, 0, stream>>>(
            *table, (const long long*)uniq_ids_i64, (const long long*)seg_off_i64,
            (const long long*)perm_i64, n_uniq, grad, lr, eps, grad_scale);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// Byte sizes of the six scratch arrays of tfop_ps_push_adagrad_dedup for up to n_ids positions
// of a table of width dim: keys (fill with 0xFF), slot_of, dense, uniq, count, sums (zero-fill).
TFOP_API int tfop_ps_dedup_workspace(size_t n_ids, int dim, size_t* sizes6) {
    TFOP_REQUIRE(sizes6 != nullptr && dim >= 4 && dim % 4 == 0, "bad arguments");
    const size_t cap = dedup_capacity(n_ids);
    sizes6[0] = cap * sizeof(unsigned long long);
    sizes6[1] = n_ids * sizeof(int);
    sizes6[2] = cap * sizeof(int);
    sizes6[3] = n_ids * sizeof(long long);
    sizes6[4] = 256;
    sizes6[5] = n_ids * (size_t)dim * sizeof(float);
    return 0;
}

// The push with the duplicate merge on the device (K5b): ids[n_ids] int64 as the batch names
// them (negative = padding), grad[n_ids, dim] fp32 or bf16. Allocates nothing, never blocks the
// host; the scratch comes back clean (same state as before the call).
TFOP_API int tfop_ps_push_adagrad_dedup(const tfop_ps_table* table, const void* ids_i64,
                                        size_t n_ids, const void* grad, int grad_dtype, float lr,
                                        float eps, float grad_scale, void* ws_keys,
                                        void* ws_slot_of, void* ws_dense, void* ws_uniq,
                                        void* ws_count, void* ws_sums, int max_blocks,
                                        cudaStream_t stream) {
    if (check_table(table) != 0) return -1;
    TFOP_REQUIRE(grad_dtype == TFOP_F32 || grad_dtype == TFOP_BF16, "bad grad dtype");
    TFOP_REQUIRE(ws_keys && ws_slot_of && ws_dense && ws_uniq && ws_count && ws_sums,
                 "null scratch");
    TFOP_REQUIRE(n_ids < ((size_t)1 << 30), "too many ids for one push");
    if (n_ids == 0) return 0;
    const size_t cap = dedup_capacity(n_ids);
    const int chunks = table->dim / 4;
    ps_dedup_insert_kernel<<<grid_for(n_ids, max_blocks), 256, 0, stream>>>(
        (const long long*)ids_i64, n_ids, (unsigned long long*)ws_keys, (unsigned)(cap - 1),
        (int*)ws_slot_of, (int*)ws_count);
    ps_dedup_compact_kernel<<<grid_for(cap, max_blocks), 256, 0, stream>>>(
        (unsigned long long*)ws_keys, (unsigned)cap, (int*)ws_dense, (long long*)ws_uniq,
        (int*)ws_count);
    const int grid = grid_for(n_ids * (size_t)chunks, max_blocks);
    if (grad_dtype == TFOP_BF16)
        ps_dedup_sum_kernel<true><<<grid, 256, 0, stream>>>(
            (const int*)ws_slot_of, (const int*)ws_dense, n_ids, table->dim, grad, (float*)ws_sums);
    else
        ps_dedup_sum_kernel<false><<<grid, 256, 0, stream>>>(
            (const int*)ws_slot_of, (const int*)ws_dense, n_ids, table->dim, grad, (float*)ws_sums);
    if (g_push_fetch)
        ps_dedup_apply_kernel<true><<<grid, 256, 0, stream>>>(
            *table, (const long long*)ws_uniq, (const int*)ws_count, (float*)ws_sums, lr, eps,
            grad_scale);
    else
        ps_dedup_apply_kernel<false><<<grid, 256, 0, stream>>>(
            *table, (const long long*)ws_uniq, (const int*)ws_count, (float*)ws_sums, lr, eps,
            grad_scale);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// 1: fetch-and-add on the accumulator (exact under concurrent pushes of one row); 0 (default):
// plain load + two fire-and-forget reductions (see ps_dedup_apply_kernel).
TFOP_API int tfop_ps_set_push_mode(int fetch) {
    g_push_fetch = fetch ? 1 : 0;
    return 0;
}
