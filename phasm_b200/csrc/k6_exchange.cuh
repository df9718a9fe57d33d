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
//   * three item exchanges route 8/16-byte records to the rank that needs them next: k3_route the
//     scan survivors (k-mer hash, A entry, position) to the rank that owns the k-mer's slice of
//     the anchor table, k6_route_cand the candidate keys to the rank that compares them,
//     k6_route the verified hit keys to the rank that owns their A entry.  A block ranks 1024
//     items per destination in shared memory, reserves room with ONE system-scope atomicAdd per
//     destination on the owner's counter and stores the items over NVLink (lanes with the same
//     destination write consecutive slots).  Looking the table up remotely instead (one 8-byte peer
//     load per probe) was measured first: every such load moves a 32-byte sector plus packet
//     headers, and the scan of 2 GPUs took 2.2x the time of one;
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

constexpr int STATUS_WORDS = 8;
// status vector a rank publishes to everyone at the end of a pass
enum Status { ST_ERR = 0, ST_NCAND, ST_NSCAN, ST_NHITS, ST_NROWS, ST_FLAGS, ST_NSTAGE, ST_SPARE };

struct alignas(256) ArenaHdr {
    unsigned long long flag[F_COUNT][MAX_RANKS];
    unsigned long long status[MAX_RANKS][STATUS_WORDS];
    unsigned long long inbox_n;    // hit keys appended to this rank's hit list (local and remote atomics)
    unsigned long long scan_n;     // scan survivors appended to this rank's probe list (idem)
    unsigned long long cand_n;     // candidate keys appended to this rank's verify list (idem)
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

// ---- item routing -----------------------------------------------------------------------
// Append every item to the list of the rank dest_of(item) says (scan survivors -> owner of the
// k-mer's table slice; hit keys -> owner of the A entry).  A block takes a tile of 256 * ITEMS
// items, ranks them per destination in shared memory (warp-aggregated), reserves the space with ONE
// system-scope atomicAdd per destination on the owner's counter and then stores the items -- lanes
// with the same destination write consecutive slots.  (The first form reserved per warp: ~150 000
// atomics on one remote address per pass, ~0.4 ms of serialised NVLink round trips at 2 GPUs.)
constexpr int ROUTE_ITEMS = 4;

template <class T>
struct RouteLists {
    T* inbox[MAX_RANKS];                     // list of every rank (peer-mapped)
    unsigned long long* inbox_n[MAX_RANKS];  // its append counter (in the owner's ArenaHdr)
    uint64_t inbox_cap[MAX_RANKS];
    uint32_t nr, me;
};

template <class T, class DestFn>
__device__ __forceinline__ void route_items(const T* __restrict__ items, uint64_t n, const RouteLists<T>& R, DestFn dest_of,
                                            unsigned long long* n_sent) {
    __shared__ unsigned int s_cnt[MAX_RANKS];
    __shared__ unsigned long long s_base[MAX_RANKS];
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t tile_items = (uint64_t)blockDim.x * ROUTE_ITEMS;
    const uint64_t n_tiles = (n + tile_items - 1) / tile_items;
    if (threadIdx.x < MAX_RANKS) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    unsigned long long sent = 0;
    for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        T it[ROUTE_ITEMS];
        uint32_t dest[ROUTE_ITEMS], off[ROUTE_ITEMS];
#pragma unroll
        for (int j = 0; j < ROUTE_ITEMS; ++j) {
            const uint64_t i = tile * tile_items + (uint64_t)j * blockDim.x + threadIdx.x;
            dest[j] = 0xffffffffu;
            if (i < n) {
                it[j] = items[i];
                dest[j] = dest_of(it[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < ROUTE_ITEMS; ++j) {
            const unsigned grp = __match_any_sync(FULL, dest[j]);
            const uint32_t leader = __ffs(grp) - 1;
            uint32_t base = 0;
            if (lane == leader && dest[j] != 0xffffffffu) base = atomicAdd(&s_cnt[dest[j]], (unsigned)__popc(grp));
            base = __shfl_sync(FULL, base, leader);
            off[j] = base + __popc(grp & ((1u << lane) - 1u));
        }
        __syncthreads();
        if (threadIdx.x < R.nr && s_cnt[threadIdx.x])
            s_base[threadIdx.x] = atomicAdd_system(R.inbox_n[threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]);
        __syncthreads();
#pragma unroll
        for (int j = 0; j < ROUTE_ITEMS; ++j)
            if (dest[j] != 0xffffffffu) {
                const unsigned long long pos = s_base[dest[j]] + off[j];
                if (pos < R.inbox_cap[dest[j]]) R.inbox[dest[j]][pos] = it[j];
                sent += dest[j] != R.me;
            }
        __syncthreads();
        if (threadIdx.x < MAX_RANKS) s_cnt[threadIdx.x] = 0;
        __syncthreads();
    }
    if (n_sent) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) sent += __shfl_xor_sync(FULL, sent, d);
        if (lane == 0 && sent) atomicAdd(n_sent, sent);
    }
    __threadfence_system();
}

// With strand-mirror halving a rank compares one candidate of each strand-mirror pair and
// produces BOTH hit keys, and on N > 1 GPUs a candidate is verified by the rank that owns its
// k-mer, not its A entry.  Rows are emitted by the rank that owns A (that keeps "smallest p per
// (a,b)" local).
struct RouteParams {
    const uint64_t* keys;              // verified hit keys of this rank's candidates (staging list)
    const unsigned long long* n_ptr;
    uint64_t cap;
    KeyFmt fmt;
    RankMap map;
    RouteLists<uint64_t> lists;        // hit list of every rank
    unsigned long long* n_sent;        // stat: keys that left this GPU
};

__global__ void __launch_bounds__(256) k6_route(const RouteParams P) {
    const unsigned long long n_all = *P.n_ptr;
    const uint64_t n = n_all < P.cap ? n_all : P.cap;
    route_items<uint64_t>(P.keys, n, P.lists, [&](uint64_t key) { return (uint32_t)owner_of(P.map, P.fmt.a(key)); }, P.n_sent);
}

// Candidate keys (flags clear) to the rank that compares them: verifier_rank() of the two entries'
// owners -- one of the two, so that one operand of every comparison is local, and chosen so that a
// rank needs the forward strands of only half of its peers (ovl_common.cuh).
__global__ void __launch_bounds__(256) k6_route_cand(const RouteParams P) {
    const unsigned long long n_all = *P.n_ptr;
    const uint64_t n = n_all < P.cap ? n_all : P.cap;
    route_items<uint64_t>(P.keys, n, P.lists, [&](uint64_t key) {
        const uint32_t a = P.fmt.a(key), b = P.fmt.b(key);
        return (uint32_t)verifier_rank(P.map.n, owner_of(P.map, a), owner_of(P.map, b), a, b);
    }, P.n_sent);
}

// The status vector of this rank, from the device counters (one thread).  A count above its
// capacity raises STF_OVERFLOW: every rank reads every rank's vector after the exchange, so all of
// them take the same "grow and repeat" decision without talking to each other on the host.
struct StatusParams {
    unsigned long long* st;              // [STATUS_WORDS] this rank's slot in its own arena
    const unsigned long long* ctr;       // device counter block
    int c_err, c_ncand, c_nscan, c_nstage, c_flags;
    const unsigned long long* n_probes;  // this rank's probe list (scan survivors routed to it)
    const unsigned long long* n_candin;  // this rank's verify list (candidate keys routed to it)
    const unsigned long long* n_hits;
    const unsigned long long* n_rows;
    const unsigned long long* abort_word;
    unsigned long long cap_cand, cap_candin, cap_scan, cap_stage, cap_hits, cap_rows;
};

__global__ void k9_status(const StatusParams P) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const unsigned long long nfound = P.ctr[P.c_ncand], ncandin = *P.n_candin, nstage = P.ctr[P.c_nstage];
    const unsigned long long ncand = nfound > ncandin ? nfound : ncandin;
    const unsigned long long nscan = P.ctr[P.c_nscan] > *P.n_probes ? P.ctr[P.c_nscan] : *P.n_probes;
    const unsigned long long nhits = *P.n_hits, nrows = *P.n_rows;
    unsigned long long fl = P.ctr[P.c_flags] | (*P.abort_word ? STF_TIMEOUT : 0ull);
    if (nfound > P.cap_cand || ncandin > P.cap_candin || nscan > P.cap_scan || nstage > P.cap_stage || nhits > P.cap_hits || nrows > P.cap_rows)
        fl |= STF_OVERFLOW;
    P.st[ST_ERR] = P.ctr[P.c_err];
    P.st[ST_NCAND] = ncand;
    P.st[ST_NSCAN] = nscan;
    P.st[ST_NSTAGE] = nstage;
    P.st[ST_NHITS] = nhits;
    P.st[ST_NROWS] = nrows;
    P.st[ST_FLAGS] = fl;
    P.st[ST_SPARE] = 0;
}

}  // namespace ovl
