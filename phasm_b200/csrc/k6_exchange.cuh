// k6_exchange.cuh -- the multi-GPU data path over peer-mapped HBM (NVLink 5 / NVSwitch).
//
// One process per GPU; every rank allocates an ARENA (ovl.cu) whose CUDA IPC handle the others
// open at connect time, so a rank holds a device pointer into every peer's arena.  All exchanges
// of a pass then happen on the device, without NCCL calls and without the host:
//
//   * the forward strands of the packed reads (the bulk: 0.25 B/base) are PUSHED into the arenas
//     of the peers that compare against them by the copy engines, in pieces over several streams,
//     each followed by an 8-byte copy of the step id into the peer's flag word (same stream:
//     ordered behind the data).  Which peers: verifier_rank() (ovl_common.cuh) places a candidate
//     on one of the two ranks that hold its reads so that rank r only needs the reads of ranks
//     r+1 .. r+floor(N/2) -- half of an all-gather -- and never a remote reverse complement;
//   * the small, urgent slices (prefix table, filter bitmap) are pushed by a few blocks of plain
//     stores (k_push): a copy-engine copy of them queued behind the bulk copy to the same peer;
//   * a consumer waits with k_wait (one warp spinning on its own arena's flag words, bounded by
//     a timeout) -- the only kernels that ever wait, and they wait on OTHER GPUs only;
//   * three item exchanges carry 8/16-byte records to the rank that needs them next: scan
//     survivors (k-mer hash, A entry, position) to the rank that owns the k-mer's slice of the
//     anchor table, candidate keys to the rank that compares them, verified hit keys to the rank
//     that owns their A entry.  They are a PULL: the producing kernel (k3_filter, k3_resolve,
//     k4a/k4b) sorts the 128-256 items of a warp's shared-memory staging buffer by destination,
//     reserves room with one atomic per destination on a counter in its OWN memory and stores the
//     runs into per-destination segments in its OWN arena; the per-destination counts travel with
//     the "phase done" signal (k_signal_counts), and the consuming kernel of the destination
//     (k3_resolve, the candidate grouping, k7_hist / k7_scatter) streams the segment out of the
//     producer's arena through the peer mapping, a whole grid of coalesced loads.  History:
//     (1) resolving survivors through peer-mapped table pointers (one 8-byte peer load per probe;
//     every such load moves a 32-byte sector plus packet headers): the scan of 2 GPUs took 2.2x the
//     time of one; (2) a routing kernel per exchange with one remote atomic per warp: ~150 000
//     atomics on one remote address, ~0.4 ms of serialised NVLink round trips; (3) the same with
//     one remote atomic per 1024 items: 0.10 / 0.06 / 0.04 ms per exchange at 2 GPUs; (4) PUSHED from
//     inside the producing kernels (remote stores in runs of 512 / N bytes): the producers slowed
//     down with the share of remote items -- filter pass 0.13 / 0.22 / 0.38 ms at 1 / 2 / 4 GPUs,
//     the stores competing with the bulk copies on the same links (profiles/r02l_*);
//   * the rows stay on the rank that emitted them (sharded by A range like the reads); rank 0
//     pulls them with copy-engine copies only when a consumer asks for all of them in one place.
//
// The reference has no counterpart (single process, single thread: SURVEY.md 2.1, 8(e)).
#pragma once
#include "ovl_common.cuh"

namespace ovl {

// phases a rank signals to its peers; flag[k][s] in MY arena header = the last step in which rank
// s completed phase k towards me
enum Flag { F_PRE = 0, F_FILT, F_SCAN, F_CAND, F_FWD, F_HITS, F_STATUS, F_COUNT };
// the three item lists of a rank
enum ListKind { L_SCAN = 0, L_CAND, L_HITS, L_COUNT };

constexpr int STATUS_WORDS = 8;
// status vector a rank publishes to everyone at the end of a pass.  ST_FILL = the fullest of this
// rank's buffers, in 1/1024 of its capacity (> 1024: it overflowed; the growth factor of the retry)
enum Status { ST_ERR = 0, ST_FILL, ST_NROWS, ST_FLAGS, ST_SPARE0, ST_SPARE1, ST_SPARE2, ST_SPARE3 };

struct alignas(256) ArenaHdr {
    unsigned long long flag[F_COUNT][MAX_RANKS];
    unsigned long long status[MAX_RANKS][STATUS_WORDS];
    unsigned long long seg_n[L_COUNT][MAX_RANKS];  // seg_n[k][s]: items rank s appended to ITS segment of my list k in this step
    unsigned long long step_word;  // current step id: the source of the 8-byte signal copies
    unsigned long long abort;      // != 0: a wait timed out, the pass is void
};

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Wait until every peer s in `mask` (s != me) has written flags[s] >= step (one lane per peer).
// Gives up after timeout_ns and raises the abort word, so that a dead peer costs a bounded time,
// not a hang.
__global__ void __launch_bounds__(32) k_wait(const unsigned long long* __restrict__ flags, uint32_t nr, uint32_t me,
                                             uint32_t mask, unsigned long long step, unsigned long long timeout_ns,
                                             unsigned long long* __restrict__ abort_word) {
    const uint32_t s = threadIdx.x;
    if (s < nr && s != me && ((mask >> s) & 1u)) {
        const unsigned long long t0 = global_ns();
        while (ld_acquire_sys(flags + s) < step) {
            if (global_ns() - t0 > timeout_ns) {
                atomicMax(abort_word, (unsigned long long)(s + 1));
                break;
            }
            __nanosleep(200);
        }
    }
    __syncwarp();
    __threadfence_system();
}

// Tell every peer that phase `kind` of step `step` is complete on this rank.  Launched behind
// the kernel whose remote writes it announces (kernel boundary = all of them performed).
__global__ void __launch_bounds__(32) k_signal(ArenaHdr* const* __restrict__ peer_hdr, uint32_t nr, uint32_t me,
                                               int kind, unsigned long long step) {
    const uint32_t s = threadIdx.x;
    __threadfence_system();
    if (s < nr && s != me) st_release_sys(&peer_hdr[s]->flag[kind][me], step);
}

// The same, with the item counts of list `list`: cnt[d] = items this rank appended to its segment
// in rank d's list (counted in this rank's memory by the producing kernel, which has ended).
__global__ void __launch_bounds__(32) k_signal_counts(ArenaHdr* const* __restrict__ peer_hdr, uint32_t nr, uint32_t me,
                                                      int kind, int list, const unsigned long long* __restrict__ cnt,
                                                      unsigned long long step) {
    const uint32_t d = threadIdx.x;
    if (d < nr) {
        peer_hdr[d]->seg_n[list][me] = cnt[d];
        __threadfence_system();
        if (d != me) st_release_sys(&peer_hdr[d]->flag[kind][me], step);
    }
}

// Small, urgent pushes (prefix slice, filter slice): plain 16-byte stores into every peer's arena
// from a few blocks.  (A copy-engine copy of the same bytes queued behind the bulk copy of the
// packed reads to the same peer and arrived 0.14 ms late at 2 GPUs.)  blockIdx.y picks the peer.
struct PushTargets {
    uint4* dst[MAX_RANKS];  // where this rank's slice lives in every peer's arena
    uint32_t nr, me;
};

__global__ void __launch_bounds__(256) k_push(const uint4* __restrict__ src, const PushTargets T, uint64_t n16) {
    const uint32_t r = (T.me + 1u + blockIdx.y) % T.nr;
    uint4* __restrict__ dst = T.dst[r];
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x)
        dst[i] = src[i];
    __threadfence_system();
}

// The item exchange proper: stream the segments the other ranks produced for this rank out of
// their arenas into this rank's inbox, with the whole grid and four independent 16-byte loads per
// thread in flight (measured on this box, tools/p2p_bw.cu, profiles/r02o_p2p_bandwidth.txt: a grid
// of >= 592 blocks pulls 740 GB/s over NVLink, a copy engine 760 GB/s -- but a copy engine needs
// the size on the host, and the counts only exist on the device).  The consumers then read local
// memory: a consumer that streamed the peer segments itself, one record ahead and with its own work
// between the loads, moved them at 150 GB/s (k3_resolve at 4 GPUs: 0.48 ms instead of 0.09).
struct PullParams {
    const uint4* src[MAX_RANKS];     // segment of source s for this rank, in s's arena
    uint4* dst[MAX_RANKS];           // slot s of this rank's inbox
    uint64_t cap[MAX_RANKS];         // items that fit (min of the two capacities)
    const unsigned long long* n;     // [nr] item counts (delivered with the phase signal)
    uint32_t item_bytes;             // 8 or 16
    uint32_t nr, me;
};

__global__ void __launch_bounds__(256) k_pull(const PullParams P) {
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (uint64_t)gridDim.x * blockDim.x;
    for (uint32_t j = 1; j < P.nr; ++j) {
        const uint32_t s = (P.me + j) % P.nr;  // ring order: at any moment the ranks read from different peers
        const unsigned long long n_all = P.n[s];
        const uint64_t n = n_all < P.cap[s] ? n_all : P.cap[s];
        const uint64_t n16 = (n * P.item_bytes + 15) / 16;
        const uint4* __restrict__ src = P.src[s];
        uint4* __restrict__ dst = P.dst[s];
        uint64_t i = tid;
        for (; i + 3 * nth < n16; i += 4 * nth) {
            const uint4 v0 = ldg_stream(src + i), v1 = ldg_stream(src + i + nth), v2 = ldg_stream(src + i + 2 * nth),
                        v3 = ldg_stream(src + i + 3 * nth);
            dst[i] = v0;
            dst[i + nth] = v1;
            dst[i + 2 * nth] = v2;
            dst[i + 3 * nth] = v3;
        }
        for (; i < n16; i += nth) dst[i] = ldg_stream(src + i);
    }
}

// The status vector of this rank, from the device counters (one thread).  Every buffer that
// somebody appends to without a size check on the host is compared with its capacity HERE, on the
// receiving side: a fill above 1024/1024 raises STF_OVERFLOW.  Every rank reads every rank's vector
// after the exchange, so all of them take the same "grow and repeat" decision without talking to
// each other on the host.
struct StatusParams {
    unsigned long long* st;               // [STATUS_WORDS] this rank's slot in its own arena
    const unsigned long long* err;        // first bad entry + 1
    const unsigned long long* flags;      // STF_* bits raised by kernels
    const unsigned long long* abort_word;
    const unsigned long long* seg_n;      // [L_COUNT][MAX_RANKS] of this rank's header
    unsigned long long* seg_snap;         // [L_COUNT][MAX_RANKS] copy of seg_n for the host (taken before F_STATUS goes out)
    unsigned long long seg_cap[L_COUNT][MAX_RANKS];  // capacity of the segment of every source in every list
    uint32_t nr;
    const unsigned long long* n_surv;     // verifier: candidates alive after the head
    unsigned long long cap_surv;
    const unsigned long long* n_rows;
    unsigned long long cap_rows;
};

__global__ void k9_status(const StatusParams P) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long fill = 0;
    auto see = [&](unsigned long long n, unsigned long long cap) {
        const unsigned long long f = cap ? (n * 1024ull + cap - 1) / cap : (n ? 1ull << 20 : 0ull);
        if (f > fill) fill = f;
    };
    for (int k = 0; k < L_COUNT; ++k)
        for (uint32_t s = 0; s < P.nr; ++s) {
            const unsigned long long n = P.seg_n[k * MAX_RANKS + s];
            P.seg_snap[k * MAX_RANKS + s] = n;
            see(n, P.seg_cap[k][s]);
        }
    see(*P.n_surv, P.cap_surv);
    see(*P.n_rows, P.cap_rows);
    unsigned long long fl = *P.flags | (*P.abort_word ? STF_TIMEOUT : 0ull);
    if (fill > 1024) fl |= STF_OVERFLOW;
    P.st[ST_ERR] = *P.err;
    P.st[ST_FILL] = fill;
    P.st[ST_NROWS] = *P.n_rows;
    P.st[ST_FLAGS] = fl;
    P.st[ST_SPARE0] = P.st[ST_SPARE1] = P.st[ST_SPARE2] = P.st[ST_SPARE3] = 0;
}

}  // namespace ovl
