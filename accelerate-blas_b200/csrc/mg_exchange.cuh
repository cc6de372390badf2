// mg_exchange.cuh -- the small peer-memory exchange protocol shared by mg.cu (gemv / dot collectives) and dot.cu
// (the dot kernel whose last CTA performs the scalar all-reduce itself).
//
// Every rank owns one exchange buffer, IPC-mapped into every peer: [parity 0 | parity 1], each parity =
// XFLAGS bytes of flags (one 64-byte line per source rank) + world * XSLOT bytes of deposit slots.  Collective
// number `seq` uses parity seq & 1: a rank deposits its piece in EVERY peer's slot for this rank, raises its flag
// there to `seq` (st.release.sys) and waits until all of its own flags read `seq` (ld.acquire.sys).
#pragma once
#include "common.cuh"

namespace accblas {

constexpr int MG_MAX_RANKS = 16;
constexpr size_t XSLOT = 1u << 20;                 // bytes one rank can deposit per collective (gemv T/H partial y)
constexpr size_t XFLAGS = 4096;                    // flag block per parity: one 64-byte line per source rank

struct PeerPtrs { char *p[MG_MAX_RANKS]; };

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void wait_flag(const uint32_t *p, uint32_t seq) {
    unsigned ns = 20;
    while (ld_acquire_sys(p) != seq) {             // a collective: waits for the peer like any other would
        __nanosleep(ns);
        if (ns < 1000) ns *= 2;
    }
}

// What the fused dot kernel needs to finish a sharded dot: the peers' exchange buffers, this collective's parity and
// sequence number, and the communicator's scratch (per-CTA partials + the self-resetting arrival counter).
struct DotExchange {
    PeerPtrs peers;
    int rank, world;
    size_t parity_off;
    uint32_t seq;
    void *partials;                                // >= 32 bytes x grid
    unsigned *counter;                             // zero between calls (atomicInc wraps it back)
};

// dot.cu: local dot of the slices + the scalar all-reduce in ONE launch (the last CTA to finish sums the partials
// in index order, deposits the rank's value in every peer, waits for theirs and adds them in rank order).
int dot_mg_dispatch(cudaStream_t s, int dtype, int conj_x, int64_t n, const void *x, const void *y, void *out,
                    const DotExchange &ex, int max_grid);

}  // namespace accblas
