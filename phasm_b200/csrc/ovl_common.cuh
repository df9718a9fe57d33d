// ovl_common.cuh -- device layout and small helpers shared by the overlapper kernels.
//
// HBM layout (persistent tables live outside, per-pass working data inside the ARENA, which on
// N > 1 GPUs is the peer-mapped allocation the other ranks push into; ovl.cu):
//
//   ascii     u8     staged sequence bytes, every record 16-byte aligned, 16 B slack at both
//                    ends
//   packed    uint4  one 16-byte GRANULE = 64 bases, 2 bits/base, base j of a granule in bits
//                    [2j, 2j+2) of the granule's 128-bit little-endian value; every entry
//                    starts on a granule boundary and is zero padded to it.
//                    code = (ascii >> 1) & 3  (A0 C1 T2 G3), so complement = code ^ 2.
//                    TWO REGIONS: [0, G_f) holds every entry that is not a reverse-complement
//                    twin, in entry order (the "forward region": on N GPUs a rank's slice of it
//                    is what travels over NVLink, straight into the peers' arrays), [G_f, G)
//                    holds the twins in entry order; two spare granules follow (the verifier
//                    reads one granule past an entry's end).
//   ent       uint2  [M]   {first granule, length in bases} of every entry
//   reg       uint4  [M+1] the entries in REGION order (= granule order): {first granule,
//                    entry, length, granules}; reg[M] = {G, 0, 0, 0} is the sentinel
//   len       u32    [M]   entry lengths in bases
//   src       u64    [M]   byte offset of the entry's source record in `ascii`
//   flags     u8     [M]   bit0: entry is the reverse complement of its source record (it then
//                          directly follows the forward entry of the same record)
//
// Candidate / hit key (64 bit): [a | b | p | sp | cont], most significant first.
//
// Reference semantics restated by these kernels: /root/reference/src/overlapper.cpp:28-150
// (see SURVEY.md 3.3 for the normative R1/R2 statement).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ovl {

constexpr int GRAN_BASES = 64;
constexpr unsigned FULL = 0xffffffffu;
constexpr int MAX_RANKS = 16;  // GPUs of one box (kernel parameter blocks carry per-rank pointer tables)

// Anchor hash-table slot: one 64-bit word {fingerprint (high 32) | list pointer (low 32)}.
// The table is PARTITIONED by k-mer hash over the ranks of a multi-GPU job (rank r owns the keys
// with owner_of_hash() == r and inserts only those, so table, filter slice and build time do not
// grow with the number of GPUs); the scan resolves a key in its owner's partition, through
// peer-mapped pointers when the owner is another GPU (TableView).  A list pointer is an anchor id
// (b * stride + delta) with bit 31 = "this node has a successor" (next[id] is only read, and
// only written, when it is set: most keys are singletons and resolve with ONE load).
// Two different k-mers that collide on slot AND fingerprint (~2^-32 per probe) share a list: the
// scan then emits a few false candidates, which the verifier rejects -- the table only has to
// return a superset.
typedef unsigned long long Slot;
constexpr unsigned long long EMPTY_SLOT = ~0ull;
constexpr uint32_t NIL = 0xffffffffu;
constexpr uint32_t HAS_NEXT = 0x80000000u;
constexpr uint32_t ID_MASK = 0x7fffffffu;

struct TableView {
    const Slot* tbl[MAX_RANKS];      // tbl[r]: partition of rank r (peer-mapped for r != me)
    const uint32_t* nxt[MAX_RANKS];  // nxt[r]: its successor array, indexed by anchor id
    uint32_t nr;                     // ranks (1: everything is local)
    int tbits;                       // log2 slots of every partition
};

struct Row {
    uint32_t a, b, astart, aend, bend;
};

// bits kernels raise in the pass's flag word (reported to every rank in the status vector)
constexpr unsigned long long STF_OVERFLOW = 1;  // a capacity was exceeded: grow and repeat the pass
constexpr unsigned long long STF_TABLE = 2;     // anchor partition full
constexpr unsigned long long STF_TIMEOUT = 4;   // a k_wait gave up
constexpr unsigned long long STF_GATHER = 8;    // rank 0's row array too small

// Bit layout of a 64-bit candidate key: [a | b | p | sp | cont], most significant first.
struct KeyFmt {
    int bits_e;  // bits of an entry index
    int bits_p;  // bits of a position
    __host__ __device__ int total() const { return 2 * bits_e + bits_p + 2; }
    __host__ __device__ uint64_t make(uint32_t a, uint32_t b, uint32_t p) const {
        return ((((uint64_t)a << bits_e) | b) << bits_p | p) << 2;
    }
    __host__ __device__ uint32_t p(uint64_t key) const {
        return (uint32_t)((key >> 2) & ((1ull << bits_p) - 1ull));
    }
    __host__ __device__ uint32_t b(uint64_t key) const {
        return (uint32_t)((key >> (2 + bits_p)) & ((1ull << bits_e) - 1ull));
    }
    __host__ __device__ uint32_t a(uint64_t key) const {
        return (uint32_t)(key >> (2 + bits_p + bits_e));
    }
};
// CONT sorts before SP so that, when the two items of one key arrive as separate entries (mirror
// mode, k4_verify.cuh), the SP entries of a pair stay contiguous at the end of the pair's run.
constexpr uint64_t HIT_CONT = 1;  // rem >= |b|: b wholly inside a (R2 row)
constexpr uint64_t HIT_SP = 2;    // rem <= |b|: suffix-prefix (R1 candidate)

// Entry ranges of the ranks of a multi-GPU job (rank r holds the reads of entries [e[r], e[r+1])).
struct RankMap {
    uint32_t e[MAX_RANKS + 1];
    int n;  // ranks; <= 1: single GPU
};
__host__ __device__ __forceinline__ int owner_of(const RankMap& R, uint32_t a) {
    int r = 0;
    while (r + 1 < R.n && a >= R.e[r + 1]) ++r;
    return r;
}
// The rank that compares a candidate whose entries live on ranks ra and rb.  It is one of the two
// (so one operand is always local), it is the same for a candidate and its strand-mirror twins
// (symmetric in the two READS), and it is chosen along the ring so that rank r only ever needs the
// reads of ranks r+1 .. r+floor(n/2): half of the all-gather a "verify where the candidate was
// found" rule needs.  d == n/2 (n even) is the tie: split by the parity of the two read numbers.
__host__ __device__ __forceinline__ int verifier_rank(int n, int ra, int rb, uint32_t a, uint32_t b) {
    if (ra == rb) return ra;
    int d = rb - ra;
    if (d < 0) d += n;
    if (2 * d < n) return ra;
    if (2 * d > n) return rb;
    const bool hi = (((a >> 1) + (b >> 1)) & 1u) != 0;
    return hi ? (ra > rb ? ra : rb) : (ra < rb ? ra : rb);
}
// does rank r compare candidates that involve reads of rank s (does s have to send them to r)?
__host__ __device__ __forceinline__ bool needs_reads_of(int n, int r, int s) {
    if (r == s) return false;
    int d = s - r;
    if (d < 0) d += n;
    return 2 * d <= n;
}

// Strand-mirror halving (k4_verify.cuh explains why it is exact).  When every read is present
// on both strands (entry e ^ 1 is the other strand of entry e) only one candidate of each
// strand-mirror pair is compared.  mirror_items() returns the HIT_* items candidate (a,p,b)
// has to verify and emit itself; 0 means its twins cover it.
struct MirrorRule {
    int on;
    RankMap map;
};
// SP twins (a,.,b) <-> (b^1,.,a^1): which of the two compares is decided by a rule that is
// symmetric under the swap and splits every rank's candidates about evenly (a plain "smaller a
// wins" would leave the last rank of a multi-GPU run with nothing to compare).
__host__ __device__ __forceinline__ bool sp_self(uint32_t a, uint32_t bt) {
    if (a == bt) return true;  // fixed point: the candidate is its own twin
    const bool lt = a < bt;
    return (((a ^ bt) >> 1) & 1u) ? !lt : lt;
}
// Candidates across two ranks: the twin maps (SP and CONT alike) turn the entry of the REMOTE read
// (the one the comparing rank does not hold) into its other strand, so "compare the one whose remote
// entry is the forward strand" picks exactly one of every twin pair -- and the reverse-complement
// strands of remote reads are never needed: only forward strands travel, none is rebuilt.
__host__ __device__ __forceinline__ uint32_t mirror_items(const MirrorRule& R, uint32_t a, uint32_t la, uint32_t b,
                                                 uint32_t lb, uint32_t p) {
    const uint32_t rem = la - p;
    uint32_t fl = (rem <= lb ? (uint32_t)HIT_SP : 0u) | (rem >= lb ? (uint32_t)HIT_CONT : 0u);
    if (R.on) {
        if (R.map.n > 1) {
            const int ra = owner_of(R.map, a), rb = owner_of(R.map, b);
            if (ra != rb) {
                const uint32_t remote = verifier_rank(R.map.n, ra, rb, a, b) == ra ? b : a;
                return (remote & 1u) ? 0u : fl;
            }
        }
        if ((fl & HIT_SP) && !sp_self(a, b ^ 1u)) fl &= ~(uint32_t)HIT_SP;
        if (a & 1u) fl &= ~(uint32_t)HIT_CONT;
    }
    return fl;
}

__host__ __device__ __forceinline__ uint64_t kmer_mask(int k) {
    return k >= 32 ? ~0ull : ((1ull << (2 * k)) - 1ull);
}

// ---- dense path (stride 1): shared-memory blocked bloom filter ---------------------------
// 32-bit hash of a packed k-mer: high half of a 64-bit multiplicative hash.
__host__ __device__ __forceinline__ uint32_t kmer_hash(uint64_t x) {
    return (uint32_t)((x * 0x9E3779B97F4A7C15ull) >> 32);
}
// Blocked bloom filter: one 32-bit word per key, two bits in it.
__host__ __device__ __forceinline__ uint32_t filter_word(uint32_t h, int fw_bits) {
    return h >> (32 - fw_bits);
}
__host__ __device__ __forceinline__ uint32_t filter_bits(uint32_t h) {
    return (1u << (h & 31)) | (1u << ((h >> 5) & 31));
}

// ---- table hash (both paths) ----------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x *= 0x9E3779B97F4A7C15ull;
    x ^= x >> 32;
    x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 29;
    return x;
}
// Owner rank of a key: the integer part of  (high 32 hash bits) * nr / 2^32;  the fractional part
// of the same product (uniform inside one owner) gives the slot.  nr = 1: owner 0, slot = top bits.
__host__ __device__ __forceinline__ uint32_t owner_of_hash(uint64_t h, uint32_t nr) {
    return nr <= 1 ? 0u : (uint32_t)((((uint64_t)(uint32_t)(h >> 32)) * nr) >> 32);
}
__host__ __device__ __forceinline__ uint32_t slot_of(uint64_t h, uint32_t nr, int tbits) {
    const uint32_t frac = (uint32_t)(h >> 32) * (nr ? nr : 1u);
    return frac >> (32 - tbits);
}
// Sparse-path filter: one bit per key hash in an L2-resident bitmap of nr slices of 2^slice_log2
// bits; slice r holds exactly the keys rank r owns (built by r, pushed to every peer).
__host__ __device__ __forceinline__ uint64_t bitmap_bit(uint64_t h, uint32_t nr, int slice_log2) {
    return ((uint64_t)owner_of_hash(h, nr) << slice_log2) | ((uint32_t)h & ((1u << slice_log2) - 1u));
}

// k-mer starting at base j (0..31) of word w0, continuing into w1.
__device__ __forceinline__ uint64_t kmer_at(uint64_t w0, uint64_t w1, int j) {
    const int s = 2 * j;
    return s ? ((w0 >> s) | (w1 << (64 - s))) : w0;
}

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// Lookup shared by the scan kernels: list pointer of the key with hash h in its owner's
// partition, NIL if absent.  (Measured and not kept, r02i: fetching a whole 32-byte sector = 4 slots
// per probe step -- 4.7 dependent loads per warp step became ~2, but k3_resolve stayed at 87 us:
// it is bound by its ~360 instructions per 32 survivors at 44 resident warps, not by this chain.)
__device__ __forceinline__ uint32_t table_find(const TableView& T, uint64_t h, uint32_t& owner) {
    owner = owner_of_hash(h, T.nr);
    const Slot* __restrict__ table = T.tbl[owner];
    const uint32_t tmask = (1u << T.tbits) - 1u;
    const uint32_t fp = (uint32_t)h;
    uint32_t s = slot_of(h, T.nr, T.tbits);
    for (uint32_t tries = 0; tries <= tmask; ++tries) {
        const unsigned long long v = __ldg(table + s);
        if (v == EMPTY_SLOT) return NIL;
        if ((uint32_t)(v >> 32) == fp) return (uint32_t)v;
        s = (s + 1) & tmask;
    }
    return NIL;
}

// ---- item lists between the stages (and, on N > 1 GPUs, between the ranks) ----------------------
// Three kinds of items flow through a pass: scan survivors (to the rank that owns their k-mer),
// candidate keys (to the rank that compares them), hit keys (to the rank that owns their A entry).
// The list of a rank has one SEGMENT per source rank.  A source keeps the segments it produces -- one
// per destination -- in ITS OWN arena (its outbox), reserves room with atomics on counters in its own
// memory and publishes the counts with its "phase done" signal (k6_exchange.cuh); the destination
// reads the segment in place over NVLink (SegList below).  No remote atomic, no remote store, no
// separate routing pass.  N = 1: one segment, one counter.
//
// Producer side.  Items are staged per warp in shared memory (WBUF items, appended
// warp-synchronously) and flushed with one atomic per (flush, destination), in runs of
// consecutive slots per destination.
constexpr int WBUF = 256;  // items per warp

template <class T>
struct Sink {
    T* seg[MAX_RANKS];          // the segment (in this rank's arena) of the items for rank r
    uint64_t cap[MAX_RANKS];    // its capacity
    unsigned long long* cnt;    // [nr] items appended per destination (may exceed cap: overflow is detected by the receiver)
    uint32_t nr;
};
typedef Sink<uint64_t> CandSink;

// per-warp scratch of a routed flush
struct RouteScratch {
    unsigned long long base[MAX_RANKS];
    uint32_t cnt[MAX_RANKS];
    uint8_t dest[WBUF];
};

template <class T, class DestFn>
__device__ __forceinline__ void wbuf_flush(T* buf, uint32_t& n, const Sink<T>& S, uint32_t lane, RouteScratch* rs,
                                           DestFn dest_of) {
    if (S.nr <= 1) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(S.cnt, (unsigned long long)n);
        base = __shfl_sync(FULL, base, 0);
        for (uint32_t i = lane; i < n; i += 32)
            if (base + i < S.cap[0]) S.seg[0][base + i] = buf[i];
    } else {
        // pass 1: destination of every item, items per destination
        if (lane < MAX_RANKS) rs->cnt[lane] = 0;
        __syncwarp();
        for (uint32_t i0 = 0; i0 < n; i0 += 32) {
            const uint32_t i = i0 + lane;
            uint32_t d = 0xffu;
            if (i < n) {
                d = dest_of(buf[i]);
                rs->dest[i] = (uint8_t)d;
            }
            const unsigned grp = __match_any_sync(FULL, d);
            if (d != 0xffu && lane == (uint32_t)(__ffs(grp) - 1)) rs->cnt[d] += __popc(grp);
            __syncwarp();
        }
        // one reservation per destination, on this rank's own counters
        if (lane < S.nr) {
            const uint32_t c = rs->cnt[lane];
            rs->base[lane] = c ? atomicAdd(S.cnt + lane, (unsigned long long)c) : 0ull;
            rs->cnt[lane] = 0;
        }
        __syncwarp();
        // pass 2: every item to its slot; the items of one destination fill consecutive slots
        for (uint32_t i0 = 0; i0 < n; i0 += 32) {
            const uint32_t i = i0 + lane;
            const uint32_t d = i < n ? rs->dest[i] : 0xffu;
            const unsigned grp = __match_any_sync(FULL, d);
            uint32_t off = 0;
            if (d != 0xffu) off = rs->cnt[d];
            __syncwarp();
            if (d != 0xffu) {
                if (lane == (uint32_t)(__ffs(grp) - 1)) rs->cnt[d] = off + __popc(grp);
                const unsigned long long pos = rs->base[d] + off + __popc(grp & ((1u << lane) - 1u));
                if (pos < S.cap[d]) S.seg[d][pos] = buf[i];
            }
            __syncwarp();
        }
    }
    n = 0;
    __syncwarp();
}

// Append `item` from every lane with have != 0.  All 32 lanes must call.
template <class T, class DestFn>
__device__ __forceinline__ void wbuf_push(T* buf, uint32_t& n, const Sink<T>& S, uint32_t lane, bool have, const T& item,
                                          RouteScratch* rs, DestFn dest_of) {
    const unsigned m = __ballot_sync(FULL, have);
    if (!m) return;
    if (n + 32 > WBUF) wbuf_flush(buf, n, S, lane, rs, dest_of);
    if (have) buf[n + __popc(m & ((1u << lane) - 1u))] = item;
    n += __popc(m);
    __syncwarp();
}

// Consumer side: the list as the concatenation of its segments.  Segment s holds the items source
// rank s produced for THIS rank and lives in s's arena (its outbox): the consumer reads it in place,
// through the peer mapping when s is another GPU -- a PULL.  (Round 2 first pushed the items from
// the producing kernels into the consumer's memory: with 3/4 of the items remote at 4 GPUs the filter
// pass took 0.38 ms instead of 0.13, its short per-destination runs of 16-byte stores competing with
// the bulk copies of the packed reads on the same links; streaming reads by a whole grid do not
// have that problem, and the producers run at their single-GPU speed.)
template <class T>
struct SegList {
    const T* seg[MAX_RANKS];       // segment s (peer-mapped for s != me)
    uint64_t cap[MAX_RANKS];       // its capacity (the source's)
    const unsigned long long* n;   // [nr] items in every segment (counters; may exceed the capacity on overflow)
    uint32_t nr;
};
template <class T>
__device__ __forceinline__ uint64_t seg_count(const SegList<T>& L, uint32_t s) {
    const unsigned long long c = L.n[s];
    return c < L.cap[s] ? c : L.cap[s];
}
// pre[s] = logical index of the first item of segment s, pre[nr] = total (block-shared table)
template <class T>
__device__ __forceinline__ void seg_prefix(const SegList<T>& L, unsigned long long* s_pre) {
    if (threadIdx.x == 0) {
        unsigned long long acc = 0;
        for (uint32_t s = 0; s < L.nr; ++s) {
            s_pre[s] = acc;
            acc += seg_count(L, s);
        }
        s_pre[L.nr] = acc;
    }
    __syncthreads();
}
// logical item i < pre[nr]
template <class T>
__device__ __forceinline__ T seg_item(const SegList<T>& L, const unsigned long long* s_pre, uint64_t i) {
    uint32_t s = 0;
    while (s + 1 < L.nr && i >= s_pre[s + 1]) ++s;
    return L.seg[s][i - s_pre[s]];
}

}  // namespace ovl
