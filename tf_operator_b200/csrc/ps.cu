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
