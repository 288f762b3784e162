// peer.cuh -- device side of the exchange of an evaluation's sums between pattern shards in different processes.
//
// The reference adds the per-packet lnL / df / ddf of its OpenMP threads with `reduction(+)` (tree/phylokernelnew.h:2377,
// 2792, 3273).  Across GPUs that is an all-reduce of 3 doubles; as a library call (ncclAllReduce) it costs a launch and
// ~40 us of latency per evaluation, as much as 1/8 of a cfg2 step split over 8 GPUs.  Here the kernel that FINISHES the
// evaluation does it itself over NVLink peer memory: one thread stores the rank's sums into its slot of every peer's
// exchange buffer, fences, raises a flag carrying the evaluation's sequence number in every peer, then waits until every
// peer's flag has arrived in its own buffer and adds the slots in RANK ORDER -- every rank computes bit-identical sums.
// Buffers alternate by the parity of the sequence number, so a rank that is one evaluation ahead never overwrites
// values a slower rank is still reading (a rank cannot get two ahead: it waits for everyone's flag each time).
#pragma once
#include "common.cuh"

// exchange `count` (<= 4) doubles at `vals` (this rank's sums, replaced by the totals).  Call from ONE thread.
__device__ __forceinline__ void peer_exchange(double *vals, int count, const PeerArgs &pa) {
    if (pa.world <= 1) return;
    const int par = (int)(pa.seq & 1ull);
    const size_t val_off = ((size_t)par * B200_PEER_MAX + pa.rank) * 4;
    const size_t flag_base = (size_t)2 * B200_PEER_MAX * 4;  // doubles before the flags
    for (int r = 0; r < pa.world; r++) {
        volatile double *dst = pa.bufs[r] + val_off;
        for (int k = 0; k < count; k++) dst[k] = vals[k];
    }
    __threadfence_system();
    for (int r = 0; r < pa.world; r++) {
        volatile unsigned long long *flag =
            reinterpret_cast<volatile unsigned long long *>(pa.bufs[r] + flag_base) + (size_t)par * B200_PEER_MAX + pa.rank;
        *flag = pa.seq;
    }
    volatile unsigned long long *mine =
        reinterpret_cast<volatile unsigned long long *>(pa.bufs[pa.rank] + flag_base) + (size_t)par * B200_PEER_MAX;
    for (int r = 0; r < pa.world; r++)
        while (mine[r] < pa.seq) __nanosleep(64);
    __threadfence_system();
    const volatile double *src = pa.bufs[pa.rank] + (size_t)par * B200_PEER_MAX * 4;
    for (int k = 0; k < count; k++) {
        double s = 0.0;
        for (int r = 0; r < pa.world; r++) s += src[(size_t)r * 4 + k];
        vals[k] = s;
    }
}
