// Symmetric peer memory ("the Service DNS name becomes a peer-memory handle").
//
// One tfop_comm per replica process (one process per B200). Each rank owns a
// data buffer and a signal pad allocated with cudaMalloc; the 64-byte CUDA IPC
// handles of both are published through the launcher's rendezvous store (or
// any byte transport) and opened by every peer, after which kernels read and
// write peer HBM directly over NVLink 5 / NVSwitch.
//
// Reference parity: the reference operator only hands out DNS names + ports
// (pkg/controller.v1/pytorch/pytorch.go:13-68); the transport lived in user
// images. This file is the B200-native replacement for that transport's
// address exchange.
#include "common.cuh"
#include "comm.h"

#include <cstdlib>

#include <mutex>
#include <new>

static thread_local char g_err[512] = "";

void tfop_set_error(const char* file, int line, const char* msg) {
    snprintf(g_err, sizeof(g_err), "%s:%d: %s", file, line, msg);
}

TFOP_API const char* tfop_last_error(void) { return g_err; }

TFOP_API int tfop_version(void) { return 100; }

TFOP_API int tfop_comm_handle_bytes(void) { return (int)(2 * sizeof(cudaIpcMemHandle_t)); }

static int comm_alloc_local(tfop_comm* c) {
    TFOP_CUDA(cudaMalloc(&c->data[c->rank], c->data_bytes));
    TFOP_CUDA(cudaMemset(c->data[c->rank], 0, c->data_bytes));
    void* sig = nullptr;
    TFOP_CUDA(cudaMalloc(&sig, TFOP_SIG_BYTES));
    TFOP_CUDA(cudaMemset(sig, 0, TFOP_SIG_BYTES));
    c->sig[c->rank] = (uint32_t*)sig;
    TFOP_CUDA(cudaMalloc((void**)&c->counters, TFOP_MAX_BLOCKS * sizeof(uint32_t)));
    TFOP_CUDA(cudaMemset(c->counters, 0, TFOP_MAX_BLOCKS * sizeof(uint32_t)));
    TFOP_CUDA(cudaDeviceSynchronize());
    return 0;
}

TFOP_API int tfop_comm_create(int rank, int world, size_t data_bytes, tfop_comm** out) {
    TFOP_REQUIRE(world >= 1 && world <= TFOP_MAX_WORLD, "world must be in [1, 8]");
    TFOP_REQUIRE(rank >= 0 && rank < world, "rank out of range");
    tfop_comm* c = new (std::nothrow) tfop_comm();
    TFOP_REQUIRE(c != nullptr, "out of host memory");
    memset(c, 0, sizeof(*c));
    c->rank = rank;
    c->world = world;
    // Whole 2 MiB pages: a peer that maps this buffer gets large-page translations. With a size
    // that is only a multiple of 4 KiB the importing GPU walked small pages and random 256-byte
    // row reads out of a 5 GB table ran ten times slower. (TFOP_COMM_ALIGN overrides, for tests.)
    size_t align = size_t(2) << 20;
    if (const char* e = getenv("TFOP_COMM_ALIGN")) {
        const long v = atol(e);
        if (v >= 256 && (v & (v - 1)) == 0) align = (size_t)v;
    }
    c->data_bytes = (data_bytes + align - 1) & ~(align - 1);
    TFOP_CUDA(cudaGetDevice(&c->device));
    int rc = comm_alloc_local(c);
    if (rc != 0) {
        delete c;
        return rc;
    }
    *out = c;
    return 0;
}

TFOP_API int tfop_comm_get_handle(tfop_comm* c, void* out_bytes) {
    cudaIpcMemHandle_t h[2];
    TFOP_CUDA(cudaIpcGetMemHandle(&h[0], c->data[c->rank]));
    TFOP_CUDA(cudaIpcGetMemHandle(&h[1], (void*)c->sig[c->rank]));
    memcpy(out_bytes, h, sizeof(h));
    return 0;
}

// all_handles: world * tfop_comm_handle_bytes() bytes, rank-major.
TFOP_API int tfop_comm_open_peers(tfop_comm* c, const void* all_handles) {
    const cudaIpcMemHandle_t* h = (const cudaIpcMemHandle_t*)all_handles;
    for (int p = 0; p < c->world; ++p) {
        if (p == c->rank) continue;
        void* d = nullptr;
        void* s = nullptr;
        TFOP_CUDA(cudaIpcOpenMemHandle(&d, h[2 * p], cudaIpcMemLazyEnablePeerAccess));
        TFOP_CUDA(cudaIpcOpenMemHandle(&s, h[2 * p + 1], cudaIpcMemLazyEnablePeerAccess));
        c->data[p] = d;
        c->sig[p] = (uint32_t*)s;
        c->opened[p] = 1;
    }
    return 0;
}

// Single-process, single-GPU emulation: `world` comm objects whose peer
// tables point at each other's local allocations. Used by the GPU tests (a
// multi-rank kernel is then run as ONE cooperative launch, never as kernels
// that wait on each other across launches).
TFOP_API int tfop_comm_create_emulated(int world, size_t data_bytes, tfop_comm** out) {
    TFOP_REQUIRE(world >= 1 && world <= TFOP_MAX_WORLD, "world must be in [1, 8]");
    for (int r = 0; r < world; ++r) {
        int rc = tfop_comm_create(r, world, data_bytes, &out[r]);
        if (rc != 0) return rc;
        out[r]->emulated = 1;
    }
    for (int r = 0; r < world; ++r)
        for (int p = 0; p < world; ++p) {
            out[r]->data[p] = out[p]->data[p];
            out[r]->sig[p] = out[p]->sig[p];
        }
    return 0;
}

TFOP_API void* tfop_comm_buffer(tfop_comm* c, int peer) {
    if (peer < 0 || peer >= c->world) return nullptr;
    return c->data[peer];
}

TFOP_API size_t tfop_comm_buffer_bytes(tfop_comm* c) { return c->data_bytes; }
TFOP_API int tfop_comm_rank(tfop_comm* c) { return c->rank; }
TFOP_API int tfop_comm_world(tfop_comm* c) { return c->world; }

// The caller states that every peer has finished with this rank's buffers (it ran a cross-rank
// exit barrier after its last collective). Until then tfop_comm_destroy refuses to free.
TFOP_API int tfop_comm_quiesce(tfop_comm* c) {
    TFOP_REQUIRE(c != nullptr, "null comm");
    c->quiesced = 1;
    return 0;
}

// Frees the local buffers and unmaps the peers'. With peers (world > 1, real IPC mappings) the
// memory may still be read or written by their kernels over NVLink: freeing it under them is a
// fault on THEIR side. So destroy only proceeds after tfop_comm_quiesce(); otherwise it returns
// an error and leaves everything mapped (process exit reclaims it).
TFOP_API int tfop_comm_destroy(tfop_comm* c) {
    if (!c) return 0;
    TFOP_REQUIRE(c->world == 1 || c->emulated || c->quiesced,
                 "peers may still be using this rank's symmetric buffers: run an exit barrier "
                 "across the ranks, call tfop_comm_quiesce(), then destroy");
    cudaDeviceSynchronize();
    for (int p = 0; p < c->world; ++p) {
        if (p == c->rank) continue;
        if (c->opened[p]) {
            cudaIpcCloseMemHandle(c->data[p]);
            cudaIpcCloseMemHandle((void*)c->sig[p]);
        }
    }
    cudaFree(c->data[c->rank]);
    cudaFree((void*)c->sig[c->rank]);
    cudaFree((void*)c->counters);
    delete c;
    return 0;
}
