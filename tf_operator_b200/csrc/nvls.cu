// K2: NVLS (NVLink SHARP) all-reduce — the NVSwitch does the reduction.
//
// Each rank owns a VMM allocation bound to one multicast object shared by all ranks. Through
// the multicast mapping, `multimem.ld_reduce` returns the SUM over every rank's copy of an
// address (reduced inside the switch) and `multimem.st` stores to every rank's copy at once.
// Rank r reduces chunk r and broadcasts it: each GPU moves ~S/W bytes in and S/W bytes out
// per all-reduce instead of 2*(W-1)/W*S, which is why NCCL's NVLS algorithm wins at large
// sizes — and what this kernel matches with the bucket cast/scale fused in.
//
// Set-up (host, once): rank 0 creates the multicast object and exports a POSIX fd; the fd
// travels to the other ranks (SCM_RIGHTS over a Unix socket, runtime/nvls.py); every rank
// adds its device, allocates its local memory, binds it and maps both views. The flag
// barrier reuses the communicator's IPC signal pad (phases 2 and 3).
//
// Availability is probed at run time (CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED); every NVLS
// path has the P2P twin in allreduce.cu.
//
// Reference parity: the all-reduce of examples/v1/distribution_strategy/keras-API/
// multi_worker_strategy-with-keras.py:76-77 (NCCL library call in the reference stack).
#include "common.cuh"
#include "comm.h"
#include "elems.cuh"

#include <cuda.h>
#include <unistd.h>

long long tfop_spin_timeout_cycles();  // allreduce.cu (library-internal)

namespace {

// ---- driver entry points (no link-time dependency on libcuda) --------------------------
#define DRV_LIST(X)                                                                         \
    X(cuDeviceGet) X(cuDeviceGetAttribute) X(cuMulticastCreate) X(cuMulticastAddDevice)     \
    X(cuMulticastBindMem) X(cuMulticastGetGranularity) X(cuMulticastUnbind) X(cuMemCreate)  \
    X(cuMemExportToShareableHandle) X(cuMemImportFromShareableHandle)                       \
    X(cuMemAddressReserve) X(cuMemMap) X(cuMemSetAccess) X(cuMemGetAllocationGranularity)   \
    X(cuMemUnmap) X(cuMemRelease) X(cuMemAddressFree) X(cuGetErrorString)

struct Drv {
#define X(name) decltype(&name) name##_ = nullptr;
    DRV_LIST(X)
#undef X
    bool loaded = false;
} g;

int load_driver() {
    if (g.loaded) return 0;
#define X(name)                                                                             \
    {                                                                                       \
        void* fn = nullptr;                                                                 \
        cudaDriverEntryPointQueryResult q;                                                  \
        TFOP_CUDA(cudaGetDriverEntryPoint(#name, &fn, cudaEnableDefault, &q));              \
        TFOP_REQUIRE(q == cudaDriverEntryPointSuccess && fn, "driver entry point missing: " #name); \
        g.name##_ = (decltype(&name))fn;                                                    \
    }
    DRV_LIST(X)
#undef X
    g.loaded = true;
    return 0;
}

#define TFOP_DRV(call)                                                          \
    do {                                                                        \
        CUresult r_ = (call);                                                   \
        if (r_ != CUDA_SUCCESS) {                                               \
            const char* s_ = nullptr;                                           \
            g.cuGetErrorString_(r_, &s_);                                       \
            char m_[256];                                                       \
            snprintf(m_, sizeof(m_), "%s -> %s", #call, s_ ? s_ : "CUDA error"); \
            tfop_set_error(__FILE__, __LINE__, m_);                             \
            return (int)r_;                                                     \
        }                                                                       \
    } while (0)

struct NvlsState {
    CUmemGenericAllocationHandle mc_handle = 0;
    CUmemGenericAllocationHandle mem_handle = 0;
    CUdeviceptr uc_va = 0, mc_va = 0;
    size_t bytes = 0;
    int device = 0;
    bool bound = false;
};

// one NVLS state per communicator, kept outside tfop_comm to leave that struct unchanged
constexpr int kMaxStates = 64;
struct Slot {
    tfop_comm* comm;
    NvlsState st;
} g_slots[kMaxStates];

NvlsState* state_of(tfop_comm* c, bool create) {
    for (auto& s : g_slots)
        if (s.comm == c) return &s.st;
    if (!create) return nullptr;
    for (auto& s : g_slots)
        if (s.comm == nullptr) {
            s.comm = c;
            s.st = NvlsState();
            return &s.st;
        }
    return nullptr;
}

CUmulticastObjectProp mc_prop(int world, size_t bytes) {
    CUmulticastObjectProp p;
    memset(&p, 0, sizeof(p));
    p.numDevices = (unsigned)world;
    p.size = bytes;
    p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}

// ---- device side --------------------------------------------------------------------------
constexpr int kThreads = 512;

struct NvlsArgs {
    uint32_t* sig[TFOP_MAX_WORLD];
    uint32_t* counter;     // per-block generation counters (second half of the comm's table)
    uint32_t* err;
    char* uc;              // local view of the bound memory
    char* mc;              // multicast view
    const void* in;
    void* out;
    size_t n;
    float scale;
    int rank, world;
    int in_dtype, out_dtype;
    long long timeout;     // spin limit of the flag barriers (tfop_comm_set_timeout_ms)
};

// false = a peer never arrived: the caller abandons the collective (see allreduce.cu).
__device__ __forceinline__ bool nvls_barrier(const NvlsArgs& a, int phase, uint32_t gen) {
    __shared__ int s_ok;
    __syncthreads();
    if (threadIdx.x == 0) s_ok = 1;
    __syncthreads();
    if ((int)threadIdx.x < a.world) {
        const int slot = (phase * TFOP_MAX_BLOCKS + blockIdx.x) * TFOP_MAX_WORLD;
        fence_acq_rel_sys();
        st_release_sys_u32(a.sig[threadIdx.x] + slot + a.rank, gen);
        const uint32_t* mine = a.sig[a.rank] + slot + threadIdx.x;
        const long long t0 = clock64();
        unsigned spins = 0;
        while ((int32_t)(ld_acquire_sys_u32(mine) - gen) < 0) {
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

// 16 bytes through the switch: sum over all ranks' copies (fp32 accumulation for bf16)
__device__ __forceinline__ uint4 mm_ld_reduce_bf16(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
__device__ __forceinline__ uint4 mm_ld_reduce_f32(const void* mc) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(mc)
                 : "memory");
    return make_uint4(__float_as_uint(r.x), __float_as_uint(r.y), __float_as_uint(r.z),
                      __float_as_uint(r.w));
}
__device__ __forceinline__ void mm_st_v4(void* mc, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc),
                 "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                 "f"(__uint_as_float(v.w))
                 : "memory");
}

template <bool BF16>
__global__ void __launch_bounds__(kThreads) nvls_allreduce_kernel(const NvlsArgs a) {
    constexpr int VEC = BF16 ? 8 : 4;
    const uint32_t gen = a.counter[blockIdx.x] + 1;
    const size_t total_vecs = (a.n + VEC - 1) / VEC;
    const size_t chunk_vecs = (total_vecs + a.world - 1) / a.world;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nthreads = (size_t)gridDim.x * blockDim.x;
    const bool in_al = a.in && (((uintptr_t)a.in) & 15) == 0;
    const bool out_al = a.out && (((uintptr_t)a.out) & 15) == 0;

    // pack: local input -> local copy of the bound buffer (wire dtype)
    if (a.in) {
        for (int c = 0; c < a.world; ++c)
            for (size_t j = tid; j < chunk_vecs; j += nthreads) {
                const size_t v = (size_t)c * chunk_vecs + j;
                if (v >= total_vecs) break;
                float f[VEC];
                load_elems<VEC>(a.in, a.in_dtype, v * VEC, a.n, in_al, f);
                uint4 pk;
                if (BF16) {
                    bf16x8 p = f32_to_bf16x8(f);
                    pk = *reinterpret_cast<uint4*>(&p);
                } else {
                    pk = make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]),
                                    __float_as_uint(f[2]), __float_as_uint(f[3]));
                }
                *reinterpret_cast<uint4*>(a.uc + v * 16) = pk;
            }
    }
    if (!nvls_barrier(a, 2, gen)) return;
    // in-switch reduce of my chunk, scale, broadcast to every rank's copy
    {
        const size_t lo = (size_t)a.rank * chunk_vecs;
        for (size_t j = tid; j < chunk_vecs; j += nthreads) {
            const size_t v = lo + j;
            if (v >= total_vecs) break;
            uint4 r = BF16 ? mm_ld_reduce_bf16(a.mc + v * 16) : mm_ld_reduce_f32(a.mc + v * 16);
            if (a.scale != 1.f) {
                if (BF16) {
                    float f[8];
                    bf16x8_to_f32(*reinterpret_cast<bf16x8*>(&r), f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) f[i] *= a.scale;
                    bf16x8 p = f32_to_bf16x8(f);
                    r = *reinterpret_cast<uint4*>(&p);
                } else {
                    r.x = __float_as_uint(__uint_as_float(r.x) * a.scale);
                    r.y = __float_as_uint(__uint_as_float(r.y) * a.scale);
                    r.z = __float_as_uint(__uint_as_float(r.z) * a.scale);
                    r.w = __float_as_uint(__uint_as_float(r.w) * a.scale);
                }
            }
            mm_st_v4(a.mc + v * 16, r);
        }
    }
    if (!nvls_barrier(a, 3, gen)) return;
    if (a.out) {
        for (int c = 0; c < a.world; ++c)
            for (size_t j = tid; j < chunk_vecs; j += nthreads) {
                const size_t v = (size_t)c * chunk_vecs + j;
                if (v >= total_vecs) break;
                const uint4 raw = ld_peer_v4(a.uc + v * 16);
                float f[VEC];
                if (BF16) {
                    bf16x8_to_f32(*reinterpret_cast<const bf16x8*>(&raw), f);
                } else {
                    f[0] = __uint_as_float(raw.x); f[1] = __uint_as_float(raw.y);
                    f[2] = __uint_as_float(raw.z); f[3] = __uint_as_float(raw.w);
                }
                store_elems<VEC>(a.out, a.out_dtype, v * VEC, a.n, out_al, f);
            }
    }
    __syncthreads();
    if (threadIdx.x == 0) a.counter[blockIdx.x] = gen;
}

uint32_t* g_nvls_counters[kMaxStates] = {nullptr};

}  // namespace

TFOP_API int tfop_nvls_supported(int* out) {
    *out = 0;
    if (load_driver() != 0) return -1;
    int dev = 0;
    TFOP_CUDA(cudaGetDevice(&dev));
    CUdevice cud;
    TFOP_DRV(g.cuDeviceGet_(&cud, dev));
    int v = 0;
    TFOP_DRV(g.cuDeviceGetAttribute_(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cud));
    *out = v;
    return 0;
}

// Size rounded up to the multicast granularity for `world` devices.
TFOP_API int tfop_nvls_round_size(int world, size_t bytes, size_t* out) {
    if (load_driver() != 0) return -1;
    CUmulticastObjectProp p = mc_prop(world, bytes);
    size_t gran = 0;
    TFOP_DRV(g.cuMulticastGetGranularity_(&gran, &p, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    *out = (bytes + gran - 1) / gran * gran;
    return 0;
}

// Rank 0: create the multicast object and export it as a POSIX fd (caller sends the fd to
// the other ranks and closes it).
TFOP_API int tfop_nvls_create(tfop_comm* c, size_t bytes, int* out_fd) {
    if (load_driver() != 0) return -1;
    NvlsState* st = state_of(c, true);
    TFOP_REQUIRE(st != nullptr, "too many NVLS communicators");
    TFOP_CUDA(cudaFree(0));
    size_t rounded = 0;
    if (tfop_nvls_round_size(c->world, bytes, &rounded) != 0) return -1;
    st->bytes = rounded;
    st->device = c->device;
    CUmulticastObjectProp p = mc_prop(c->world, rounded);
    TFOP_DRV(g.cuMulticastCreate_(&st->mc_handle, &p));
    int fd = -1;
    TFOP_DRV(g.cuMemExportToShareableHandle_(&fd, st->mc_handle,
                                             CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    *out_fd = fd;
    return 0;
}

// Other ranks: import the multicast object from the received fd.
TFOP_API int tfop_nvls_import(tfop_comm* c, size_t bytes, int fd) {
    if (load_driver() != 0) return -1;
    NvlsState* st = state_of(c, true);
    TFOP_REQUIRE(st != nullptr, "too many NVLS communicators");
    TFOP_CUDA(cudaFree(0));
    size_t rounded = 0;
    if (tfop_nvls_round_size(c->world, bytes, &rounded) != 0) return -1;
    st->bytes = rounded;
    st->device = c->device;
    TFOP_DRV(g.cuMemImportFromShareableHandle_(&st->mc_handle, (void*)(uintptr_t)fd,
                                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    return 0;
}

// All ranks: register this rank's device with the object. Barrier (host) afterwards: binding
// is only legal once every device has been added.
TFOP_API int tfop_nvls_add_device(tfop_comm* c) {
    NvlsState* st = state_of(c, false);
    TFOP_REQUIRE(st && st->mc_handle, "tfop_nvls_create/import first");
    CUdevice cud;
    TFOP_DRV(g.cuDeviceGet_(&cud, st->device));
    TFOP_DRV(g.cuMulticastAddDevice_(st->mc_handle, cud));
    return 0;
}

// All ranks (after the barrier): allocate local memory, bind it, map the unicast and the
// multicast view. Barrier (host) afterwards before the first collective.
TFOP_API int tfop_nvls_bind(tfop_comm* c) {
    NvlsState* st = state_of(c, false);
    TFOP_REQUIRE(st && st->mc_handle, "tfop_nvls_create/import first");
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = st->device;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    TFOP_DRV(g.cuMemCreate_(&st->mem_handle, st->bytes, &prop, 0));
    TFOP_DRV(g.cuMulticastBindMem_(st->mc_handle, 0, st->mem_handle, 0, st->bytes, 0));
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = st->device;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    TFOP_DRV(g.cuMemAddressReserve_(&st->uc_va, st->bytes, 0, 0, 0));
    TFOP_DRV(g.cuMemMap_(st->uc_va, st->bytes, 0, st->mem_handle, 0));
    TFOP_DRV(g.cuMemSetAccess_(st->uc_va, st->bytes, &acc, 1));
    TFOP_DRV(g.cuMemAddressReserve_(&st->mc_va, st->bytes, 0, 0, 0));
    TFOP_DRV(g.cuMemMap_(st->mc_va, st->bytes, 0, st->mc_handle, 0));
    TFOP_DRV(g.cuMemSetAccess_(st->mc_va, st->bytes, &acc, 1));
    TFOP_CUDA(cudaMemset((void*)st->uc_va, 0, st->bytes));
    TFOP_CUDA(cudaDeviceSynchronize());
    st->bound = true;
    c->mc_ptr = (void*)st->mc_va;
    c->mc_uc_ptr = (void*)st->uc_va;
    c->mc_bytes = st->bytes;
    for (int i = 0; i < kMaxStates; ++i)
        if (g_slots[i].comm == c && g_nvls_counters[i] == nullptr) {
            TFOP_CUDA(cudaMalloc((void**)&g_nvls_counters[i], TFOP_MAX_BLOCKS * sizeof(uint32_t)));
            TFOP_CUDA(cudaMemset(g_nvls_counters[i], 0, TFOP_MAX_BLOCKS * sizeof(uint32_t)));
        }
    return 0;
}

TFOP_API void* tfop_nvls_local_ptr(tfop_comm* c) { return c->mc_uc_ptr; }
TFOP_API size_t tfop_nvls_bytes(tfop_comm* c) { return c->mc_bytes; }

// Fused all-reduce(sum)*scale through the switch on the region at byte offset sym_off of the
// bound buffer. in == NULL: data already there (wire dtype); out == NULL: leave the result
// there. All ranks, same arguments.
TFOP_API int tfop_allreduce_nvls(tfop_comm* c, const void* in, int in_dtype, void* out,
                                 int out_dtype, int wire_dtype, size_t n, size_t sym_off,
                                 float scale, int max_blocks, cudaStream_t stream) {
    TFOP_REQUIRE(c && c->mc_ptr, "NVLS is not set up on this communicator");
    TFOP_REQUIRE(wire_dtype == TFOP_F32 || wire_dtype == TFOP_BF16, "bad wire dtype");
    if (n == 0) return 0;
    const size_t wsz = wire_dtype == TFOP_F32 ? 4 : 2;
    const int vec = wire_dtype == TFOP_F32 ? 4 : 8;
    TFOP_REQUIRE((sym_off & 15) == 0, "sym_off must be 16-byte aligned");
    TFOP_REQUIRE(sym_off + (n * wsz + 15) / 16 * 16 <= c->mc_bytes,
                 "message does not fit the NVLS buffer");
    NvlsArgs a;
    memset(&a, 0, sizeof(a));
    for (int p = 0; p < c->world; ++p) a.sig[p] = c->sig[p];
    a.counter = nullptr;
    for (int i = 0; i < kMaxStates; ++i)
        if (g_slots[i].comm == c) a.counter = g_nvls_counters[i];
    TFOP_REQUIRE(a.counter != nullptr, "NVLS counters missing");
    a.err = c->sig[c->rank] + TFOP_SIG_WORDS;
    a.uc = (char*)c->mc_uc_ptr + sym_off;
    a.mc = (char*)c->mc_ptr + sym_off;
    a.in = in;
    a.out = out;
    a.n = n;
    a.scale = scale;
    a.rank = c->rank;
    a.world = c->world;
    a.in_dtype = in_dtype;
    a.out_dtype = out_dtype;
    a.timeout = tfop_spin_timeout_cycles();
    const size_t total_vecs = (n + vec - 1) / vec;
    const size_t work = (total_vecs + c->world - 1) / c->world;
    if (max_blocks <= 0 || max_blocks > TFOP_MAX_BLOCKS) max_blocks = TFOP_MAX_BLOCKS;
    size_t b = (work + (size_t)kThreads * 2 - 1) / ((size_t)kThreads * 2);
    if (b < 1) b = 1;
    if (b > (size_t)max_blocks) b = max_blocks;
    if (wire_dtype == TFOP_BF16)
        nvls_allreduce_kernel<true><<<(int)b, kThreads, 0, stream>>>(a);
    else
        nvls_allreduce_kernel<false><<<(int)b, kThreads, 0, stream>>>(a);
    TFOP_CUDA(cudaGetLastError());
    return 0;
}

TFOP_API int tfop_nvls_destroy(tfop_comm* c) {
    NvlsState* st = state_of(c, false);
    if (!st) return 0;
    cudaDeviceSynchronize();
    if (st->mc_va) {
        g.cuMemUnmap_(st->mc_va, st->bytes);
        g.cuMemAddressFree_(st->mc_va, st->bytes);
    }
    if (st->uc_va) {
        g.cuMemUnmap_(st->uc_va, st->bytes);
        g.cuMemAddressFree_(st->uc_va, st->bytes);
    }
    if (st->bound) {
        CUdevice cud;
        if (g.cuDeviceGet_(&cud, st->device) == CUDA_SUCCESS)
            g.cuMulticastUnbind_(st->mc_handle, cud, 0, st->bytes);
    }
    if (st->mem_handle) g.cuMemRelease_(st->mem_handle);
    if (st->mc_handle) g.cuMemRelease_(st->mc_handle);
    for (int i = 0; i < kMaxStates; ++i)
        if (g_slots[i].comm == c) {
            if (g_nvls_counters[i]) cudaFree(g_nvls_counters[i]);
            g_nvls_counters[i] = nullptr;
            g_slots[i].comm = nullptr;
        }
    c->mc_ptr = c->mc_uc_ptr = nullptr;
    c->mc_bytes = 0;
    return 0;
}
