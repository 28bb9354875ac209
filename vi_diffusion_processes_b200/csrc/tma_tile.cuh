// Blackwell-native tile movement for the parallel-in-time sweeps: 1-D bulk asynchronous copies (the TMA unit's
// cp.async.bulk, SASS UBLKCP) completed on an mbarrier (SYNCS), and bulk stores from shared memory.
//
// Every tile of a sweep is a CONTIGUOUS range of one sequence's array (9-35 KB), so one elected thread moves a whole
// array of a tile with one instruction instead of every thread issuing 17-34 eight-byte LDGSTS copies, and the
// copies of the NEXT tile are in flight while the current one is scanned and replayed (two-stage ring).
//
// Alignment.  cp.async.bulk wants 16-byte aligned addresses and sizes; rows of the reference layout are only 8-byte
// aligned (n = N + 1 doubles per sequence is odd for the usual grids; D = 3 blocks are 72 bytes).  A tile's array is
// therefore fetched from its start rounded DOWN to 16 bytes into a 16-byte aligned staging buffer -- the element
// wanted first lands at slot `off` in {0, 1} -- and a last odd element, which a bulk copy could only fetch by
// reading 8 bytes past the end of the array, is brought by one 8-byte LDGSTS.  Stores mirror this: the aligned
// middle goes out as one bulk store when shared and global parity agree, the odd head / tail elements as plain
// stores; arrays whose output position has the other parity (the shifted outputs of naturals_to_ssm) fall back to
// coalesced thread stores.
#pragma once
#include <stdint.h>

namespace vdp {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// make the initialised barriers visible to the asynchronous proxy (the TMA unit) before the first copy arrives
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "VDP_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra VDP_MBAR_DONE;\n"
        "bra VDP_MBAR_WAIT;\n"
        "VDP_MBAR_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned; completes `bytes` on `bar`
__device__ __forceinline__ void bulk_g2s(void* sdst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(sdst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// shared -> global bulk store (joins the thread's current bulk group)
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// the shared-memory source of every committed bulk store of this thread has been read (the buffer may be reused)
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// every committed bulk store of this thread is complete (before the CTA exits)
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy writes to shared memory (st.shared) become visible to the asynchronous proxy (bulk stores)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// start [g, g + count) on its way to L2 (the tile this CTA stages next: its bulk copy then finds the lines there);
// one thread, one instruction: the range is widened to 16-byte bounds inside the array's allocation
__device__ __forceinline__ void bulk_prefetch_l2(const double* g, int count) {
    if (count <= 1) return;
    const double* g0 = g - ((reinterpret_cast<uintptr_t>(g) >> 3) & 1u);
    const uint32_t bytes = (uint32_t)(((g + count) - g0) >> 1) << 4;
    if (bytes) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(g0), "r"(bytes) : "memory");
}

__device__ __forceinline__ void ldgsts8(double* s, const double* g) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(smem_u32(s)), "l"(g) : "memory");
}
__device__ __forceinline__ void ldgsts_wait_all() {
    asm volatile("cp.async.commit_group;\n" ::);
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

// slot of the first wanted element in a staging buffer filled by tile_fetch (0 or 1)
__device__ __forceinline__ int tile_off(const double* g) { return (int)((reinterpret_cast<uintptr_t>(g) >> 3) & 1u); }

// doubles a staging buffer must hold for `count` elements (room for the over-fetched head, rounded to 16 bytes)
__host__ __device__ constexpr int tile_slots(int count) { return (count + 2 + 1) & ~1; }

// ONE thread: bring [g, g + count) into `sraw` (16-byte aligned): element i lands at sraw[tile_off(g) + i].
// Returns the bytes the bulk copy will complete on `bar` (to be summed into one mbar_arrive_expect_tx, which may be
// issued after the copies of the same phase); the odd tail element, if any, travels by LDGSTS (wait with
// ldgsts_wait_all in the same thread, then a block barrier).
__device__ __forceinline__ uint32_t tile_fetch(double* sraw, const double* g, int count, uint64_t* bar) {
    if (count <= 0) return 0;
    const int off = tile_off(g);
    const int pairs = (off + count) >> 1;
    const uint32_t bytes = (uint32_t)pairs << 4;
    if (pairs > 0) bulk_g2s(sraw, g - off, bytes, bar);
    if ((off + count) & 1) ldgsts8(sraw + off + count - 1, g + count - 1);
    return bytes;
}

// ONE thread, after a block barrier that follows fence_async_smem() in the writing threads: store the `count` elements
// at s (= sraw + tile_off of the array the tile was fetched from) to [g, g + count).  Returns false when shared and
// global parity differ (nothing stored: the caller falls back to thread stores).
__device__ __forceinline__ bool tile_store_bulk(double* g, const double* s, int count) {
    if (count <= 0) return true;
    const int goff = tile_off(g);
    if (goff != (int)((smem_u32(s) >> 3) & 1u)) return false;
    int i0 = 0, i1 = count;
    if (goff) { g[0] = s[0]; i0 = 1; }                       // odd head
    if ((i1 - i0) & 1) { --i1; g[i1] = s[i1]; }               // odd tail
    if (i1 > i0) bulk_s2g(g + i0, s + i0, (uint32_t)(i1 - i0) << 3);
    return true;
}

// ALL T threads of the block (after fence_async_smem + barrier): store a tile's array, as one bulk store when the parity
// of shared and global position agree, else with coalesced thread stores.  Thread 0 commits the bulk group afterwards.
template <int T>
__device__ __forceinline__ void tile_store(double* __restrict__ g, const double* __restrict__ s, int count) {
    if (count <= 0) return;
    if (tile_off(g) == (int)((smem_u32(s) >> 3) & 1u)) {
        if (threadIdx.x == 0) tile_store_bulk(g, s, count);
    } else {
        for (int i = threadIdx.x; i < count; i += T) g[i] = s[i];
    }
}

}  // namespace vdp
