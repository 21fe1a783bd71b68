// Internal layout of the symmetric-memory communicator.
#pragma once
#include <stddef.h>
#include <stdint.h>

#define TFOP_MAX_WORLD 8
#define TFOP_MAX_BLOCKS 64
#define TFOP_SIG_PHASES 4
// signal pad: [phase][block][src_rank] u32 generation counters
#define TFOP_SIG_WORDS (TFOP_SIG_PHASES * TFOP_MAX_BLOCKS * TFOP_MAX_WORLD)
#define TFOP_SIG_BYTES (TFOP_SIG_WORDS * 4 + 256)  // + error word and padding

struct tfop_comm {
    int rank;
    int world;
    int device;
    int emulated;
    size_t data_bytes;
    void* data[TFOP_MAX_WORLD];      // data[rank] is local; others are IPC mappings
    uint32_t* sig[TFOP_MAX_WORLD];   // signal pads, same convention
    uint32_t* counters;              // local: per-block generation counters
    int opened[TFOP_MAX_WORLD];
    // NVLS (multicast) state, filled by nvls.cu when available
    void* mc_ptr;                    // multicast VA covering the data buffer (or null)
    void* mc_uc_ptr;                 // local unicast VA of the multicast-bound memory
    size_t mc_bytes;
    int quiesced;                    // set by tfop_comm_quiesce(): peers are done with our buffers
};
