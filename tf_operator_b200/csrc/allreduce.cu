// K1 / K6 / K8: P2P all-reduce over NVLink peer memory, fused with the bucket
// cast + scale, plus the flag barrier it is built on.
//
//   one-shot : every rank reads all peers' staged data and reduces locally
//              (latency-optimal; 1 barrier; staging double-buffered by call parity)
//   two-shot : rank r reduces chunk r from all peers (reduce-scatter by pull)
//              and stores the result into every peer's buffer (all-gather by
//              push); 2 barriers; can run fully in place on symmetric memory
//
// The pack (in -> wire dtype) and unpack (wire -> out dtype, * scale) passes
// are part of the same kernel, so a DDP bucket of fp32 or bf16 gradients is
// cast, summed in fp32, scaled by 1/world and written back in ONE launch.
// Summation order is rank 0..W-1 on every rank, so all replicas get bitwise
// identical results (weights never drift apart).
//
// Reference parity: replaces the NCCL all-reduce the reference's example job
// calls through TF (examples/v1/distribution_strategy/keras-API/
// multi_worker_strategy-with-keras.py:76-77) and the rabit histogram
// all-reduce of examples/xgboost/xgboostjob.yaml:19-25 (wire dtype f32).
#include "common.cuh"
#include "comm.h"
#include "elems.cuh"

#include <cooperative_groups.h>

namespace {

constexpr int kThreads = 512;
// How long a block waits for a peer's flag before it gives up (SM clock cycles). Generous by
// default - a legitimate skew of many seconds exists (a checkpoint save on one rank, a first
// step's autotuning, an input stall) - and configurable: tfop_comm_set_timeout_ms().
long long g_spin_timeout_cycles = 240LL * 1000 * 1000 * 1000;  // ~2 minutes at 1.9 GHz

struct ARArgs {
    void* data[TFOP_MAX_WORLD];
    uint32_t* sig[TFOP_MAX_WORLD];
    uint32_t* counter;   // local per-block generation counters
    uint32_t* err;       // local error word
    const void* in;      // may be null: data already staged in symmetric memory
    void* out;           // may be null (two-shot): leave result in symmetric memory
    size_t n;            // elements
    size_t sym_off;      // byte offset of the staging region inside the symmetric buffer
    size_t region_bytes; // bytes of one staging region (16 B multiple)
    float scale;
    int rank;
    int in_dtype;
    int out_dtype;
    int algo;            // 1 one-shot, 2 two-shot
    long long timeout;   // spin limit of the flag barriers, in cycles
};

struct ARArgsEmu {
    ARArgs r[TFOP_MAX_WORLD];
};

template <typename T>
struct Wire;
template <>
struct Wire<__nv_bfloat16> {
    static constexpr int VEC = 8;
    __device__ static void unpack(const uint4& raw, float* f) {
        bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), f);
    }
    __device__ static uint4 pack(const float* f) {
        bf16x8 p = f32_to_bf16x8(f);
        return *reinterpret_cast<uint4*>(&p);
    }
};
template <>
struct Wire<float> {
    static constexpr int VEC = 4;
    __device__ static void unpack(const uint4& raw, float* f) {
        f[0] = __uint_as_float(raw.x);
        f[1] = __uint_as_float(raw.y);
        f[2] = __uint_as_float(raw.z);
        f[3] = __uint_as_float(raw.w);
    }
    __device__ static uint4 pack(const float* f) {
        return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                          __float_as_uint(f[3]));
    }
};

// K8: block-granular flag barrier across ranks. Block b of every rank
// rendezvous with block b of every other rank. `gen` increases by one per
// kernel launch, so flags never need resetting.
// Returns false when a peer never arrived (or an earlier barrier of this communicator already
// timed out): the caller then ABANDONS the collective - no reduce, no stores - so a dead or
// wedged peer can never turn into silently wrong gradients. The error word stays set; the
// engine polls it and fails the replica with a retryable exit code (parallel/engine.py).
template <int W>
__device__ __forceinline__ bool rank_barrier(const ARArgs& a, int phase, uint32_t gen) {
    __shared__ int s_ok;
    __syncthreads();
    if (threadIdx.x == 0) s_ok = 1;
    __syncthreads();
    if (threadIdx.x < W) {
        const int slot = (phase * TFOP_MAX_BLOCKS + blockIdx.x) * TFOP_MAX_WORLD;
        fence_acq_rel_sys();
        st_release_sys_u32(a.sig[threadIdx.x] + slot + a.rank, gen);
        const uint32_t* mine = a.sig[a.rank] + slot + threadIdx.x;
        const long long t0 = clock64();
        unsigned spins = 0;
        while ((int32_t)(ld_acquire_sys_u32(mine) - gen) < 0) {
            // a barrier that timed out anywhere on this rank poisons the rest quickly
            if (clock64() - t0 > a.timeout ||
                ((++spins & 0x3ffu) == 0 && *(volatile uint32_t*)a.err != 0u)) {
                atomicExch(a.err, 1u);
                s_ok = 0;
                break;
            }
        }
    }
    __syncthreads();
    return s_ok != 0;
}

template <typename T, int W>
__device__ __forceinline__ void reduce_vec(const ARArgs& a, size_t byte_off, float scale,
                                           float* acc) {
    constexpr int VEC = Wire<T>::VEC;
    uint4 raw[W];
    // issue all W loads back to back (all peers in flight at once), then sum in
    // fixed rank order for bitwise-identical results on all ranks
#pragma unroll
    for (int p = 0; p < W; ++p) raw[p] = ld_peer_v4((const char*)a.data[p] + byte_off);
#pragma unroll
    for (int i = 0; i < VEC; ++i) acc[i] = 0.f;
#pragma unroll
    for (int p = 0; p < W; ++p) {
        float f[VEC];
        Wire<T>::unpack(raw[p], f);
#pragma unroll
        for (int i = 0; i < VEC; ++i) acc[i] += f[i];
    }
#pragma unroll
    for (int i = 0; i < VEC; ++i) acc[i] *= scale;
}

template <typename T, int W>
__device__ void allreduce_body(const ARArgs& a) {
    constexpr int VEC = Wire<T>::VEC;
    const uint32_t gen = a.counter[blockIdx.x] + 1;
    const size_t total_vecs = (a.n + VEC - 1) / VEC;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nthreads = (size_t)gridDim.x * blockDim.x;
    const bool in_aligned = a.in && (((uintptr_t)a.in) & 15) == 0;
    const bool out_aligned = a.out && (((uintptr_t)a.out) & 15) == 0;

    if (a.algo == 1) {
        // ---------------- one-shot ----------------
        const size_t off = a.sym_off + (a.in ? (size_t)(gen & 1u) * a.region_bytes : 0);
        if (a.in) {
            char* stage = (char*)a.data[a.rank] + off;
            for (size_t v = tid; v < total_vecs; v += nthreads) {
                float f[VEC];
                load_elems<VEC>(a.in, a.in_dtype, v * VEC, a.n, in_aligned, f);
                *reinterpret_cast<uint4*>(stage + v * 16) = Wire<T>::pack(f);
            }
        }
        if (!rank_barrier<W>(a, 0, gen)) return;
        for (size_t v = tid; v < total_vecs; v += nthreads) {
            float acc[VEC];
            reduce_vec<T, W>(a, off + v * 16, a.scale, acc);
            store_elems<VEC>(a.out, a.out_dtype, v * VEC, a.n, out_aligned, acc);
        }
        // staged input is double-buffered by parity, so no trailing barrier is
        // needed; reading user-owned symmetric memory in place needs one.
        if (!a.in && !rank_barrier<W>(a, 1, gen)) return;
    } else {
        // ---------------- two-shot ----------------
        const size_t off = a.sym_off;
        const size_t chunk_vecs = (total_vecs + W - 1) / W;
        char* self = (char*)a.data[a.rank] + off;
        if (a.in) {
#pragma unroll 1
            for (int c = 0; c < W; ++c)
                for (size_t j = tid; j < chunk_vecs; j += nthreads) {
                    const size_t v = (size_t)c * chunk_vecs + j;
                    if (v >= total_vecs) break;
                    float f[VEC];
                    load_elems<VEC>(a.in, a.in_dtype, v * VEC, a.n, in_aligned, f);
                    *reinterpret_cast<uint4*>(self + v * 16) = Wire<T>::pack(f);
                }
        }
        if (!rank_barrier<W>(a, 0, gen)) return;
        {
            const size_t lo = (size_t)a.rank * chunk_vecs;
            for (size_t j = tid; j < chunk_vecs; j += nthreads) {
                const size_t v = lo + j;
                if (v >= total_vecs) break;
                float acc[VEC];
                reduce_vec<T, W>(a, off + v * 16, a.scale, acc);
                const uint4 pk = Wire<T>::pack(acc);
#pragma unroll
                for (int p = 0; p < W; ++p) st_peer_v4((char*)a.data[p] + off + v * 16, pk);
            }
        }
        if (!rank_barrier<W>(a, 1, gen)) return;
        if (a.out) {
#pragma unroll 1
            for (int c = 0; c < W; ++c)
                for (size_t j = tid; j < chunk_vecs; j += nthreads) {
                    const size_t v = (size_t)c * chunk_vecs + j;
                    if (v >= total_vecs) break;
                    float f[VEC];
                    Wire<T>::unpack(ld_peer_v4(self + v * 16), f);
                    store_elems<VEC>(a.out, a.out_dtype, v * VEC, a.n, out_aligned, f);
                }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) a.counter[blockIdx.x] = gen;
}

template <typename T, int W>
__global__ void __launch_bounds__(kThreads) allreduce_kernel(const ARArgs a) {
    allreduce_body<T, W>(a);
}

// Emulation: blockIdx.y selects the rank; all "ranks" live on this GPU and the
// launch is cooperative so every CTA is co-resident while they wait on flags.
template <typename T, int W>
__global__ void __launch_bounds__(kThreads) allreduce_kernel_emu(const ARArgsEmu e) {
    allreduce_body<T, W>(e.r[blockIdx.y]);
}

size_t g_oneshot_max_bytes = 512 * 1024;

int pick_blocks(size_t work_vecs, int max_blocks) {
    if (max_blocks <= 0 || max_blocks > TFOP_MAX_BLOCKS) max_blocks = TFOP_MAX_BLOCKS;
    size_t b = (work_vecs + (size_t)kThreads * 2 - 1) / ((size_t)kThreads * 2);
    if (b < 1) b = 1;
    if (b > (size_t)max_blocks) b = max_blocks;
    return (int)b;
}

int fill_args(tfop_comm* c, ARArgs* a, const void* in, int in_dtype, void* out, int out_dtype,
              int wire_dtype, size_t n, size_t sym_off, float scale, int algo) {
    const size_t wsz = wire_dtype == TFOP_F32 ? 4 : 2;
    const size_t region = ((n * wsz + 15) / 16) * 16;
    if (algo == 0) algo = (n * wsz <= g_oneshot_max_bytes && (in != nullptr || out != nullptr)) ? 1 : 2;
    TFOP_REQUIRE(algo == 1 || algo == 2, "algo must be 0 (auto), 1 (one-shot) or 2 (two-shot)");
    TFOP_REQUIRE(wire_dtype == TFOP_F32 || wire_dtype == TFOP_BF16, "wire dtype must be f32 or bf16");
    TFOP_REQUIRE(in == nullptr || in_dtype == TFOP_F32 || in_dtype == TFOP_BF16, "bad in dtype");
    TFOP_REQUIRE(out == nullptr || out_dtype == TFOP_F32 || out_dtype == TFOP_BF16, "bad out dtype");
    TFOP_REQUIRE((sym_off & 15) == 0, "sym_off must be 16-byte aligned");
    TFOP_REQUIRE(algo != 1 || out != nullptr, "one-shot needs an output pointer");
    const size_t need = sym_off + region * ((algo == 1 && in) ? 2 : 1);
    TFOP_REQUIRE(need <= c->data_bytes, "message does not fit the symmetric buffer region");
    for (int p = 0; p < TFOP_MAX_WORLD; ++p) {
        a->data[p] = p < c->world ? c->data[p] : nullptr;
        a->sig[p] = p < c->world ? c->sig[p] : nullptr;
    }
    a->counter = c->counters;
    a->err = c->sig[c->rank] + TFOP_SIG_WORDS;  // error word just past the flag slots
    a->in = in;
    a->out = out;
    a->n = n;
    a->sym_off = sym_off;
    a->region_bytes = region;
    a->scale = scale;
    a->rank = c->rank;
    a->in_dtype = in_dtype;
    a->out_dtype = out_dtype;
    a->algo = algo;
    a->timeout = g_spin_timeout_cycles;
    return 0;
}

template <typename T>
int launch_real(int world, const ARArgs& a, int blocks, cudaStream_t stream) {
    switch (world) {
        case 1: allreduce_kernel<T, 1><<<blocks, kThreads, 0, stream>>>(a); break;
        case 2: allreduce_kernel<T, 2><<<blocks, kThreads, 0, stream>>>(a); break;
        case 3: allreduce_kernel<T, 3><<<blocks, kThreads, 0, stream>>>(a); break;
        case 4: allreduce_kernel<T, 4><<<blocks, kThreads, 0, stream>>>(a); break;
        case 5: allreduce_kernel<T, 5><<<blocks, kThreads, 0, stream>>>(a); break;
        case 6: allreduce_kernel<T, 6><<<blocks, kThreads, 0, stream>>>(a); break;
        case 7: allreduce_kernel<T, 7><<<blocks, kThreads, 0, stream>>>(a); break;
        case 8: allreduce_kernel<T, 8><<<blocks, kThreads, 0, stream>>>(a); break;
        default: return -1;
    }
    return 0;
}

template <typename T>
int launch_emu(int world, ARArgsEmu& e, int blocks, cudaStream_t stream) {
    void* params[] = {(void*)&e};
    dim3 grid(blocks, world), block(kThreads);
    const void* fn = nullptr;
    switch (world) {
        case 1: fn = (const void*)allreduce_kernel_emu<T, 1>; break;
        case 2: fn = (const void*)allreduce_kernel_emu<T, 2>; break;
        case 3: fn = (const void*)allreduce_kernel_emu<T, 3>; break;
        case 4: fn = (const void*)allreduce_kernel_emu<T, 4>; break;
        case 5: fn = (const void*)allreduce_kernel_emu<T, 5>; break;
        case 6: fn = (const void*)allreduce_kernel_emu<T, 6>; break;
        case 7: fn = (const void*)allreduce_kernel_emu<T, 7>; break;
        case 8: fn = (const void*)allreduce_kernel_emu<T, 8>; break;
        default: return -1;
    }
    TFOP_CUDA(cudaLaunchCooperativeKernel(fn, grid, block, params, 0, stream));
    return 0;
}

}  // namespace

TFOP_API int tfop_allreduce_set_oneshot_max_bytes(size_t bytes) {
    g_oneshot_max_bytes = bytes;
    return 0;
}

// Fused all-reduce(sum) * scale with dtype casts; see file header.
//   in  == NULL : input already sits in the symmetric buffer at sym_off (wire dtype)
//   out == NULL : (two-shot only) leave the result in the symmetric buffer, in place
// A one-shot call with `in` needs 2 * region bytes at sym_off (parity double buffer).
// All ranks must issue the same sequence of calls with the same n/algo/max_blocks.
TFOP_API int tfop_allreduce(tfop_comm* c, const void* in, int in_dtype, void* out, int out_dtype,
                            int wire_dtype, size_t n, size_t sym_off, float scale, int algo,
                            int max_blocks, cudaStream_t stream) {
    TFOP_REQUIRE(c != nullptr, "null comm");
    TFOP_REQUIRE(!c->emulated, "emulated comms must use tfop_allreduce_emulated");
    if (n == 0) return 0;
    ARArgs a;
    int rc = fill_args(c, &a, in, in_dtype, out, out_dtype, wire_dtype, n, sym_off, scale, algo);
    if (rc != 0) return rc;
    const int vec = wire_dtype == TFOP_F32 ? 4 : 8;
    const size_t total_vecs = (n + vec - 1) / vec;
    const size_t work = a.algo == 1 ? total_vecs : (total_vecs + c->world - 1) / c->world;
    const int blocks = pick_blocks(work, max_blocks);
    rc = wire_dtype == TFOP_F32 ? launch_real<float>(c->world, a, blocks, stream)
                                : launch_real<__nv_bfloat16>(c->world, a, blocks, stream);
    TFOP_REQUIRE(rc == 0, "unsupported world size");
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// Test/emulation entry: every rank of a tfop_comm_create_emulated() group in
// one cooperative launch on the current GPU.
TFOP_API int tfop_allreduce_emulated(tfop_comm** comms, int world, const void* const* ins,
                                     int in_dtype, void* const* outs, int out_dtype,
                                     int wire_dtype, size_t n, size_t sym_off, float scale,
                                     int algo, int max_blocks, cudaStream_t stream) {
    TFOP_REQUIRE(world >= 1 && world <= TFOP_MAX_WORLD, "bad world");
    if (n == 0) return 0;
    ARArgsEmu e;
    memset(&e, 0, sizeof(e));
    for (int r = 0; r < world; ++r) {
        TFOP_REQUIRE(comms[r]->emulated, "comm is not emulated");
        int rc = fill_args(comms[r], &e.r[r], ins ? ins[r] : nullptr, in_dtype,
                           outs ? outs[r] : nullptr, out_dtype, wire_dtype, n, sym_off, scale,
                           algo);
        if (rc != 0) return rc;
    }
    const int vec = wire_dtype == TFOP_F32 ? 4 : 8;
    const size_t total_vecs = (n + vec - 1) / vec;
    const size_t work = e.r[0].algo == 1 ? total_vecs : (total_vecs + world - 1) / world;
    if (max_blocks <= 0 || max_blocks > 16) max_blocks = 16;  // world*blocks CTAs must be co-resident
    const int blocks = pick_blocks(work, max_blocks);
    int rc = wire_dtype == TFOP_F32 ? launch_emu<float>(world, e, blocks, stream)
                                    : launch_emu<__nv_bfloat16>(world, e, blocks, stream);
    if (rc != 0) return rc;
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

// Library-internal: the spin limit every flag barrier of this library uses (nvls.cu too).
long long tfop_spin_timeout_cycles() { return g_spin_timeout_cycles; }

// Spin limit of the cross-rank flag barriers, in milliseconds of a 2 GHz SM clock (default
// 120,000). When it expires the kernel abandons the collective and sets the communicator's
// error word. Applies to calls issued (or graphs captured) afterwards.
TFOP_API int tfop_comm_set_timeout_ms(long long ms) {
    TFOP_REQUIRE(ms > 0, "timeout must be positive");
    g_spin_timeout_cycles = ms * 2000000LL;
    return 0;
}

// Device address of the communicator's error word (uint32: 0 = healthy), for callers that copy
// it to the host asynchronously next to their own results instead of calling
// tfop_comm_check_error, which synchronises.
TFOP_API void* tfop_comm_error_ptr(tfop_comm* c) {
    return c ? (void*)(c->sig[c->rank] + TFOP_SIG_WORDS) : nullptr;
}

// Returns 1 if any barrier of this communicator ever timed out (a peer died).
TFOP_API int tfop_comm_check_error(tfop_comm* c, int* flag) {
    uint32_t v = 0;
    TFOP_CUDA(cudaMemcpy(&v, c->sig[c->rank] + TFOP_SIG_WORDS, 4, cudaMemcpyDeviceToHost));
    *flag = (int)v;
    return 0;
}
