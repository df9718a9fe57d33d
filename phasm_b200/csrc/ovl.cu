// ovl.cu -- host side of libphasm_ovl.so: the C ABI of include/phasm_ovl.h over the sm_100a
// kernels K0..K9.  No torch, no pybind11, no CPU compute path: every ovl_overlaps* call runs
// the CUDA pipeline or fails.
//
// One pass (reference: ExactOverlapper::overlaps, /root/reference/src/overlapper.cpp:28-150):
//
//   K1 k1_pack            ASCII -> 2-bit granules, "-" entries reverse-complemented on the fly
//   K2 k2_prefix_ascii, k2_insert   prefix anchors of all entries with |b| >= l: hash table
//                         (partitioned by k-mer hash over the GPUs of the job) + bit filter
//   K3 k3_filter, k3_resolve (k3_scan when l < 33)   A positions probe the anchors -> candidates
//   K4 k4a_head, k4b_tail packed-word comparison of a[p:] with b[0:] -> 64-bit hit keys
//   K6 k6_route           (N > 1 GPUs) hit keys to the rank that owns their A entry
//   K7 k7_*               bucket + sort the hit keys, longest suffix-prefix per pair + all
//                         containments -> rows
//   K9 k9_*               (N > 1 GPUs) rows into rank 0's HBM, status to everyone
//
// The pass is enqueued as a whole and the host waits ONCE, at its end, on 1 GPU and on N: every
// buffer is sized by capacity, every kernel reads its input size from a device counter, and the
// ranks of a multi-GPU job exchange data through peer-mapped arenas and flag words
// (k6_exchange.cuh), not through host-synchronised collectives.  NCCL is only the bootstrap
// channel (IPC handles, ranges) at connect time.
//
// CUB supplies the two exclusive prefix sums of the bucket ordering; everything else is in the
// k*.cuh files next to this one.
#include <algorithm>
#include <cerrno>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include <cub/cub.cuh>
#include <cuda_runtime.h>

#include "../../include/phasm_ovl.h"
#include "k1_pack.cuh"
#include "k2_anchor.cuh"
#include "k3_scan.cuh"
#include "k4_verify.cuh"
#include "k6_exchange.cuh"
#include "k7_bucket.cuh"
#include "k8_filter.cuh"
#include "k10_graph.cuh"
#include "nccl_dyn.h"

static_assert(sizeof(ovl_row) == sizeof(ovl::Row), "row layouts must agree");
static_assert(sizeof(ovl_row) == 20, "ovl_row is 5 x u32");
static_assert(sizeof(ovl::ArenaHdr) % 256 == 0, "arena header is padded to 256 bytes");

namespace {

constexpr uint64_t NO_SRC = ~0ull;  // entry whose bases live on another rank
constexpr size_t SLACK = 16;        // bytes of slack at both ends of the ASCII staging area
constexpr int MAX_CHUNKS = 16;      // pieces of the chunked host -> HBM upload

thread_local std::string g_create_err;

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    template <class T> T* as() const { return static_cast<T*>(p); }
};

struct PinBuf {
    uint8_t* p = nullptr;
    size_t cap = 0;
};

enum Ev { E_H2D0, E_H2D1, E_PACK0, E_ANCHOR0, E_SCAN0, E_VERIFY0, E_SORT0, E_EMIT0, E_GATHER0, E_END, E_D2H0, E_D2H1, E_COUNT };

// device counter block (one 64-bit word each)
enum Ctr {
    C_ERR,     // (entry + 1) of an entry holding a non-ACGT byte
    C_NANC,    // entries with |b| >= l
    C_SUMM, C_SUMB,  // accounting mode: sum m_c, sum 2*ceil(m_c/4)   (adjacent: the verifier takes a pointer to both)
    C_NSURV,   // candidates alive after the verifier's head
    C_TILE_A, C_TILE_B,  // dynamic tile hand-out of k4a / k4b (adjacent)
    C_BIG,     // buckets above BUCKET_MAX keys
    C_FLAGS,   // STF_* bits raised by kernels
    C_NROWS,   // rows of this rank
    C_SEND,    // [L_COUNT][MAX_RANKS] items this rank appended to its segment of every rank's lists (N > 1)
    // [L_COUNT][MAX_RANKS] the inbound segment counts of THIS step, copied out of the arena header by
    // k9_status before the status flag is published (a peer that has seen every flag may already be
    // writing its next step's counts into the header: ADVICE r2)
    C_SEGN = C_SEND + ovl::L_COUNT * ovl::MAX_RANKS,
    C_COUNT = C_SEGN + ovl::L_COUNT * ovl::MAX_RANKS
};

// pinned host mirror read after the one synchronisation of a pass (64-bit words)
constexpr int H_STATUS = C_COUNT;                                             // [MAX_RANKS][STATUS_WORDS]
constexpr int H_STEP = H_STATUS + ovl::MAX_RANKS * ovl::STATUS_WORDS;          // the step id (source of a H2D copy)
constexpr int H_WORDS = H_STEP + 1;

// byte offsets of the pieces of an arena (the header sits at 0); published to the peers
constexpr int PUSH_STREAMS = 8;

struct ArenaLayout {
    uint64_t bytes;
    uint64_t off_pre, off_packed, off_filter, off_table, off_next, off_scan, off_cand, off_inbox, off_rows;
    uint64_t off_in[ovl::L_COUNT];   // N > 1: where this rank pulls the other ranks' segments to (one slot per source)
    uint64_t seg_cap[ovl::L_COUNT];  // items per SEGMENT (one segment per source rank) of the probe / verify / hit list
    uint64_t rows_cap;               // rows (rank 0 of a multi-GPU job: room for the rows of all ranks, for the pull)
};

struct PeerInfo {
    cudaIpcMemHandle_t handle;
    ArenaLayout lay;
    int device;  // CUDA ordinal of the rank's GPU (every rank sees all GPUs of the box)
    int pad;
};

}  // namespace

struct ovl_ctx {
    int device = 0;
    int n_sm = 148;
    cudaStream_t st = nullptr, st_k2 = nullptr, st_push = nullptr, st_sig = nullptr, st_h2d = nullptr;
    cudaStream_t st_pushv[PUSH_STREAMS] = {};  // [0] == st_push: the bulk pushes of a pass run on several copy engines
    cudaEvent_t ev_piece[PUSH_STREAMS] = {};
    cudaEvent_t ev[E_COUNT] = {};
    cudaEvent_t ev_pack = nullptr, ev_k0 = nullptr, ev_prel = nullptr, ev_k2 = nullptr, ev_push = nullptr, ev_sig = nullptr;
    cudaEvent_t ev_chunk[MAX_CHUNKS] = {};
    cudaEvent_t ev_kp[MAX_CHUNKS] = {};  // K1 piece i has been packed (N > 1: its forward granules may travel)
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;  // ovl_timer_*
    cudaEvent_t ev_dbg[12] = {};                    // OVL_DEBUG=1: time line of the side streams
    int debug = 0;
    std::string err;
    // resident blocks per SM of the persistent (grid-stride) kernels, from the occupancy calculator
    struct {
        int bulk6 = 1, bulk5 = 1, filter6 = 1, filter5 = 1, filter4 = 1, filter3 = 1, filterx = 1, resolve = 1, head = 1, tail = 1, pack = 1;
    } occ;

    // ---- host entry store ----
    PinBuf ascii;            // [SLACK][record, 16-byte aligned]...[SLACK]
    size_t ascii_used = SLACK;
    std::vector<uint64_t> src;   // byte offset of the entry's record, NO_SRC if remote
    std::vector<uint32_t> len;
    std::vector<uint8_t> flags;  // bit0: reverse complement of the record
    std::vector<char> id_blob;
    std::vector<uint64_t> id_off{0};
    std::vector<uint32_t> gran0;  // [M] first granule of every entry (two-region layout)
    std::vector<uint32_t> tile0;  // [M+1] first K1 tile of every entry
    uint64_t bases = 0, max_len = 0;
    bool tables_dirty = true;     // entries changed since the tables were sent
    bool ascii_resident = false;  // the staged records are in HBM
    uint64_t generation = 0;      // bumped whenever the entry table changes
    uint32_t e_lo = 0, e_hi = 0;  // entries whose bases are local (contiguous)
    uint32_t G = 0, G_f = 0;      // granules: all, forward region
    ovl::ScanRanges ranges{};     // the local A granules
    int pairs = 0;                // 1: entry e ^ 1 is the other strand of entry e, for every e

    // ---- device: persistent tables ----
    DevBuf d_ascii, d_src, d_len, d_flags, d_gran0, d_ent, d_tile0, d_tile_ent, d_stile0, d_scan_tiles;
    // ---- device: per-pass working buffers outside the arena ----
    DevBuf d_ctr, d_tmp, d_cand, d_cnt, d_bkt, d_big, d_peerhdr;
    DevBuf d_fmark, d_fkeep, d_foff, d_frows;  // ovl_filter_rows
    DevBuf d_gkeys, d_gseqs, d_gflag, d_goff, d_gedges, d_gmark;  // ovl_build_graph
    uint64_t n_frows = 0;      // rows kept by the last ovl_filter_rows
    bool frows_valid = false;  // ... and they belong to the rows of the last pass
    DevBuf d_rows;                             // VIEW of the arena's row array (not owned)
    // ---- the arena ----
    DevBuf arena;
    ArenaLayout lay{};
    uint64_t arena_key[4] = {~0ull, 0, 0, 0};  // {generation, cand_cap, nranks/stride/mirror, epoch} the arena was planned for
    uint64_t arena_epoch = 0;                  // bumped by every collective "grow and repeat"
    void* peer_base[ovl::MAX_RANKS] = {};
    ArenaLayout peer_lay[ovl::MAX_RANKS] = {};
    uint64_t step = 0;

    PinBuf h_frows, h_rows, h_gedges;
    uint64_t* h_ctr = nullptr;   // pinned, H_WORDS words: counters (incl. segment counts) | status of every rank | step
    uint64_t cand_cap = 0;       // candidate keys one pass may emit
    uint64_t min_cand_cap = 1ull << 20;
    uint64_t cand_per_entry = 96;  // first candidate capacity: this many keys per local entry
    uint64_t n_rows = 0;
    bool rows_fetched = false;
    bool gathered = false;       // N > 1, rank 0: the peers' rows of the last pass have been pulled into this rank's row array
    uint64_t rowcnt_all[ovl::MAX_RANKS] = {};  // rows of every rank in the last pass
    int collect = 0;
    bool cache_l_valid = false;
    uint32_t cache_l = 0;
    uint64_t cache_nanc = 0, cache_npos = 0;
    int opt_stride = 0;         // 0 = auto; else upper bound on the anchor stride (1 = dense scan)
    int opt_serial_verify = 0;  // 1: chunk-by-chunk early-exit verifier instead of head/tail
    int opt_mirror = 1;         // 1: strand-mirror halving when every entry has its twin
    int opt_pack_bpsm = 0;      // K1 blocks per SM; 0 = default: 8 on one GPU (pass 1.078 vs 1.101 ms with 16: profiles/r02i_sweep.jsonl), 16 on N > 1
    int opt_filter_kb = 8192;   // budget of the scan's bit filter (all slices), KiB
    int opt_group_cands = -1;   // group the candidate keys by A entry before K4: -1 = when N > 1, 0 / 1 = never / always
    int opt_k2_bpsm = 0;        // K2 blocks per SM (0 = one thread per key): fewer let it trickle beside K1
    int opt_scan_bulk = 0;      // 1: k3_filter_bulk (cp.async.bulk tile ring), 0: k3_filter_wide (register pipeline)
    int opt_chunks = 8;         // pieces of the host -> HBM upload of ovl_overlaps (1 = one copy)
    int opt_timeout_ms = 20000; // peer wait timeout
    ovl_stats stats{};

    // ---- multi-GPU ----
    ncclComm_t comm = nullptr;
    int rank = 0, nranks = 1;
    std::vector<uint32_t> rank_e;  // [nranks+1] entry range of every rank
};

namespace {

using ovl::ArenaHdr;
using ovl::MAX_RANKS;
using ovl::STATUS_WORDS;

int fail(ovl_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf; else g_create_err = buf;
    return code;
}

#define CK(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess)                                                                \
            return fail(c, e_ == cudaErrorMemoryAllocation ? OVL_E_NOMEM : OVL_E_CUDA,        \
                        "CUDA error %s at %s:%d: %s", cudaGetErrorName(e_), __FILE__, __LINE__, \
                        cudaGetErrorString(e_));                                              \
    } while (0)

#define CKN(call)                                                                             \
    do {                                                                                      \
        ncclResult_t r_ = (call);                                                             \
        if (r_ != ncclSuccess)                                                                \
            return fail(c, OVL_E_NCCL, "NCCL error %d at %s:%d: %s", (int)r_, __FILE__,       \
                        __LINE__, nccl_dyn::api().GetErrorString(r_));                        \
    } while (0)

#define OVL_TRY(expr)                  \
    do {                               \
        int rc_ = (expr);              \
        if (rc_ != OVL_OK) return rc_; \
    } while (0)

int dev_ensure(ovl_ctx* c, DevBuf& b, size_t bytes, bool slack = true) {
    if (bytes <= b.cap) return OVL_OK;
    if (b.p) CK(cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
    size_t want = slack ? bytes + bytes / 8 + 256 : bytes;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess && want != bytes) {
        (void)cudaGetLastError();
        want = bytes;
        e = cudaMalloc(&b.p, want);
    }
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        b.p = nullptr;
        return fail(c, OVL_E_NOMEM, "cudaMalloc of %zu bytes failed: %s", want, cudaGetErrorString(e));
    }
    b.cap = want;
    return OVL_OK;
}

int pin_ensure(ovl_ctx* c, PinBuf& b, size_t bytes, size_t keep) {
    if (bytes <= b.cap) return OVL_OK;
    size_t want = b.cap ? b.cap : 4096;
    while (want < bytes) want += want / 2 + 4096;
    uint8_t* np = nullptr;
    cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&np), want, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return fail(c, OVL_E_NOMEM, "cudaHostAlloc of %zu bytes failed: %s", want, cudaGetErrorString(e));
    }
    if (b.p) {
        if (keep) memcpy(np, b.p, keep);
        cudaFreeHost(b.p);
    }
    b.p = np;
    b.cap = want;
    return OVL_OK;
}

inline int ceil_log2(uint64_t x) {
    int b = 0;
    while ((1ull << b) < x) ++b;
    return b;
}

inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }
inline uint64_t align256(uint64_t x) { return (x + 255) & ~uint64_t(255); }

int push_entry(ovl_ctx* c, const char* id, size_t id_len, const char* sfx, uint64_t src, uint64_t n,
               uint8_t fl) {
    if (n > 0xfffffff0ull) return fail(c, OVL_E_LIMIT, "sequence of %llu bases exceeds the 32-bit position limit", (unsigned long long)n);
    if (c->len.size() >= 0xfffffff0ull) return fail(c, OVL_E_LIMIT, "too many entries");
    c->src.push_back(src);
    c->len.push_back((uint32_t)n);
    c->flags.push_back(fl);
    if (id_len) c->id_blob.insert(c->id_blob.end(), id, id + id_len);
    if (sfx) c->id_blob.push_back(*sfx);
    c->id_off.push_back(c->id_blob.size());
    c->bases += n;
    if (n > c->max_len) c->max_len = n;
    c->tables_dirty = true;
    c->ascii_resident = false;
    return OVL_OK;
}

// copy one record into the pinned staging area; returns its byte offset
int stage_record(ovl_ctx* c, const char* seq, size_t n, uint64_t* off) {
    const size_t need = c->ascii_used + align16(n) + SLACK;
    OVL_TRY(pin_ensure(c, c->ascii, need, c->ascii_used));
    *off = c->ascii_used;
    if (n) memcpy(c->ascii.p + c->ascii_used, seq, n);
    c->ascii_used += align16(n);
    return OVL_OK;
}

int set_dev(ovl_ctx* c) {
    CK(cudaSetDevice(c->device));
    return OVL_OK;
}

float ev_ms(ovl_ctx* c, int a, int b) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev[a], c->ev[b]) != cudaSuccess) {
        (void)cudaGetLastError();
        return 0.f;
    }
    return ms;
}

// ---- arena accessors ---------------------------------------------------------------------
template <class T> T* at(void* base, uint64_t off) { return reinterpret_cast<T*>(static_cast<uint8_t*>(base) + off); }
ArenaHdr* hdr_of(ovl_ctx* c, int r) { return static_cast<ArenaHdr*>(c->peer_base[r]); }
ArenaHdr* my_hdr(ovl_ctx* c) { return static_cast<ArenaHdr*>(c->arena.p); }

// ---- sizing rules (the same function sizes the arena for the worst case and one pass) --------
// log2 slots of one table partition holding ~n_keys / nr keys at a load factor <= 0.5
int table_bits(uint64_t n_keys, int nr) {
    const double per = (double)n_keys / nr;
    const uint64_t want = (uint64_t)(2.0 * (per + 6.0 * sqrt(per + 1.0) + 64.0));
    int b = ceil_log2(want);
    return b < 6 ? 6 : b;
}
// log2 bits of one slice of the sparse-path bitmap: the most bits per key out of {32, 16, 8, 4}
// that keeps the whole bitmap (nr slices) inside the budget -- the filter has to stay L2-resident
// while the packed reads stream through (round 1: 32 bits/key unconditionally, 47 % L2 hit rate)
int slice_bits(uint64_t n_keys, int nr, uint64_t budget_bytes) {
    const uint64_t per = n_keys / nr + 1;
    int best = 0;
    for (int bpk = 32; bpk >= 4; bpk >>= 1) {
        int b = ceil_log2(per * bpk);
        if (b < 15) b = 15;
        best = b;
        if (((uint64_t)nr << b) / 8 <= budget_bytes) break;
    }
    if (best > 31) best = 31;
    return best;
}
int dense_filter_bits(uint64_t n_anc, bool smem) {
    int b = smem ? ceil_log2(n_anc) : ceil_log2(2 * n_anc);
    if (b < 5) b = 5;
    if (smem && b > 15) b = 15;
    if (b > 27) b = 27;
    return b;
}
int max_stride_log2(uint32_t M) {
    int s = 6;
    while (s > 0 && (((uint64_t)M << s) >= 0x7ffffff0ull)) --s;
    return s;
}

// all-gather a small host blob through the bootstrap communicator (collective; synchronises)
int allgather_host(ovl_ctx* c, const void* mine, size_t n, void* all) {
    auto& N = nccl_dyn::api();
    const size_t tot = n * (size_t)c->nranks;
    OVL_TRY(dev_ensure(c, c->d_tmp, tot));
    uint8_t* d = c->d_tmp.as<uint8_t>();
    CK(cudaMemcpyAsync(d + n * c->rank, mine, n, cudaMemcpyHostToDevice, c->st));
    CKN(N.AllGather(d + n * c->rank, d, n, ncclUint8, c->comm, c->st));
    CK(cudaMemcpyAsync(all, d, tot, cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    return OVL_OK;
}

void close_peers(ovl_ctx* c) {
    for (int r = 0; r < MAX_RANKS; ++r) {
        if (c->peer_base[r] && r != c->rank) cudaIpcCloseMemHandle(c->peer_base[r]);
        c->peer_base[r] = nullptr;
    }
    (void)cudaGetLastError();
}

// (Re)plan and allocate the arena for the current entry table and capacities; on N > 1 ranks
// exchange the IPC handles and map every peer's arena.  Collective on N > 1: all ranks reach it in
// the same call (the key changes with the entry table, the rank count, or an overflow every rank
// learned about from the status exchange).
int connect_arena(ovl_ctx* c, bool mirror, uint64_t scan_cap) {
    using namespace ovl;
    const int nr = c->nranks, me = c->rank;
    const uint32_t M = (uint32_t)c->len.size();
    const uint64_t hit_cap = mirror ? 3 * c->cand_cap : c->cand_cap;  // own key + SP twin + CONT twin
    // the three item lists: one segment per source rank.  Items spread about evenly over the
    // destinations (k-mer hash, placement rule, base-balanced A ranges): a segment gets twice its
    // even share, the receiver checks the counts and the pass is repeated larger if one overflows
    // (even: k_pull moves the 8-byte keys in 16-byte units)
    auto seg = [&](uint64_t total) { return nr > 1 ? ((2 * total / nr + 64 + 1) & ~1ull) : total; };
    const uint64_t own_rows = (uint64_t)nr * seg(hit_cap);  // a rank emits at most one row per hit key in its list
    uint64_t rows_total = own_rows;
    if (nr > 1) {
        std::vector<uint64_t> caps(nr);
        OVL_TRY(allgather_host(c, &own_rows, 8, caps.data()));
        if (me == 0) {
            rows_total = 0;
            for (int r = 0; r < nr; ++r) rows_total += caps[r];
        }
    }
    const int slog_max = max_stride_log2(M);
    const uint64_t keys_max = (uint64_t)M << slog_max;
    ArenaLayout L{};
    uint64_t off = sizeof(ArenaHdr);
    auto take = [&](uint64_t bytes) {
        const uint64_t o = off;
        off = align256(off + bytes);
        return o;
    };
    L.off_pre = take((uint64_t)M * 32);
    L.off_packed = take(((uint64_t)c->G + 2) * 16);
    const uint64_t f_sparse = ((uint64_t)nr << slice_bits(keys_max, nr, (uint64_t)c->opt_filter_kb << 10)) / 8;
    const uint64_t f_dense = 4ull << dense_filter_bits(M, false);
    L.off_filter = take(std::max(f_sparse, f_dense));
    L.off_table = take(8ull << table_bits(keys_max, nr));
    L.off_next = take(keys_max * 4);
    L.seg_cap[L_SCAN] = seg(scan_cap);
    L.seg_cap[L_CAND] = seg(c->cand_cap);
    L.seg_cap[L_HITS] = seg(hit_cap);
    if ((uint64_t)nr * L.seg_cap[L_CAND] > 0xfffffff0ull) return fail(c, OVL_E_LIMIT, "more than 2^32 candidate slots on one GPU");
    L.off_scan = take((uint64_t)nr * L.seg_cap[L_SCAN] * 16);
    L.off_cand = take((uint64_t)nr * L.seg_cap[L_CAND] * 8);
    L.off_inbox = take((uint64_t)nr * L.seg_cap[L_HITS] * 8);
    if (nr > 1) {
        L.off_in[L_SCAN] = take((uint64_t)nr * L.seg_cap[L_SCAN] * 16);
        L.off_in[L_CAND] = take((uint64_t)nr * L.seg_cap[L_CAND] * 8);
        L.off_in[L_HITS] = take((uint64_t)nr * L.seg_cap[L_HITS] * 8);
    }
    L.rows_cap = rows_total;
    L.off_rows = take(rows_total * sizeof(Row));
    L.bytes = off;

    // nobody may still map the old arena when it is freed
    close_peers(c);
    if (nr > 1) {
        uint64_t token = 1;
        std::vector<uint64_t> all(nr);
        OVL_TRY(allgather_host(c, &token, 8, all.data()));
    }
    if (c->arena.p) CK(cudaFree(c->arena.p));
    c->arena.p = nullptr;
    c->arena.cap = 0;
    {
        cudaError_t e = cudaMalloc(&c->arena.p, L.bytes);
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            c->arena.p = nullptr;
            return fail(c, OVL_E_NOMEM, "cudaMalloc of the %llu-byte arena failed: %s", (unsigned long long)L.bytes, cudaGetErrorString(e));
        }
        c->arena.cap = L.bytes;
    }
    CK(cudaMemsetAsync(c->arena.p, 0, sizeof(ArenaHdr), c->st));
    CK(cudaMemsetAsync(at<uint8_t>(c->arena.p, L.off_packed), 0, ((uint64_t)c->G + 2) * 16, c->st));
    CK(cudaStreamSynchronize(c->st));
    c->lay = L;
    c->step = 0;
    c->peer_base[me] = c->arena.p;
    c->peer_lay[me] = L;
    if (nr > 1) {
        PeerInfo mine{};
        CK(cudaIpcGetMemHandle(&mine.handle, c->arena.p));
        mine.lay = L;
        mine.device = c->device;
        std::vector<PeerInfo> all(nr);
        OVL_TRY(allgather_host(c, &mine, sizeof mine, all.data()));
        for (int r = 0; r < nr; ++r) {
            if (r == me) continue;
            void* p = nullptr;
            if (all[r].device != c->device) {  // copy-engine pushes want explicit peer access; "already enabled" is fine
                (void)cudaDeviceEnablePeerAccess(all[r].device, 0);
                (void)cudaGetLastError();
            }
            cudaError_t e = cudaIpcOpenMemHandle(&p, all[r].handle, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                (void)cudaGetLastError();
                return fail(c, OVL_E_CUDA, "cannot map the arena of rank %d over CUDA IPC: %s (all GPUs of the job must be visible to every rank and peer access must be possible)", r, cudaGetErrorString(e));
            }
            c->peer_base[r] = p;
            c->peer_lay[r] = all[r].lay;
        }
        // everyone has mapped everyone before anyone starts a pass
        uint64_t token = 2;
        std::vector<uint64_t> tk(nr);
        OVL_TRY(allgather_host(c, &token, 8, tk.data()));
    }
    // device table of the peers' headers (k_signal / publish kernels)
    OVL_TRY(dev_ensure(c, c->d_peerhdr, sizeof(void*) * MAX_RANKS, false));
    CK(cudaMemcpyAsync(c->d_peerhdr.p, c->peer_base, sizeof(void*) * MAX_RANKS, cudaMemcpyHostToDevice, c->st));
    CK(cudaStreamSynchronize(c->st));
    c->d_rows.p = at<uint8_t>(c->arena.p, L.off_rows);
    c->d_rows.cap = L.rows_cap * sizeof(Row);
    return OVL_OK;
}

// host -> HBM: the entry tables and the maps derived from them (only when the entries changed)
int upload_tables(ovl_ctx* c) {
    using namespace ovl;
    if (!c->tables_dirty) return OVL_OK;
    const size_t M = c->len.size();
    try {
        c->gran0.resize(M);
        c->tile0.resize(M + 1);
    } catch (const std::bad_alloc&) {
        return fail(c, OVL_E_NOMEM, "out of host memory");
    }
    auto is_twin = [&](size_t e) { return e > 0 && (c->flags[e] & 1) && !(c->flags[e - 1] & 1); };
    // pass 1: forward region = every entry that is not a twin, in entry order
    uint64_t gf = 0, nt = 0;
    uint32_t lo = (uint32_t)M, hi = 0;
    for (size_t e = 0; e < M; ++e) {
        c->tile0[e] = (uint32_t)nt;
        const uint64_t ge = (c->len[e] + 63u) / 64u;
        if (!is_twin(e)) {
            if (gf + ge >= 0xfffffff0ull) return fail(c, OVL_E_LIMIT, "more than 2^32 granules");
            c->gran0[e] = (uint32_t)gf;
            gf += ge;
            // K1 tiles: local non-twin entries (a twin is written by its forward entry's tiles)
            if (c->src[e] != NO_SRC) nt += (4 * ge + PACK_TILE_UNITS - 1) / PACK_TILE_UNITS;
        }
        if (c->src[e] != NO_SRC) {
            if (e < lo) lo = (uint32_t)e;
            hi = (uint32_t)e + 1;
        }
    }
    if (M) c->tile0[M] = (uint32_t)nt;
    // pass 2: twin region behind it
    uint64_t g = gf;
    for (size_t e = 0; e < M; ++e)
        if (is_twin(e)) {
            c->gran0[e] = (uint32_t)g;
            g += (c->len[e] + 63u) / 64u;
        }
    if (g >= 0xfffffff0ull) return fail(c, OVL_E_LIMIT, "more than 2^32 granules (%llu)", (unsigned long long)g);
    if (lo > hi) lo = hi = 0;
    for (uint32_t e = lo; e < hi; ++e)
        if (c->src[e] == NO_SRC) return fail(c, OVL_E_ARG, "local entries must be one contiguous range (entry %u is remote)", e);
    if (lo < hi && is_twin(lo)) return fail(c, OVL_E_ARG, "a read and its reverse-complement entry must be local to the same rank (entry %u)", lo);
    if (hi < M && is_twin(hi)) return fail(c, OVL_E_ARG, "a read and its reverse-complement entry must be local to the same rank (entry %u)", hi);
    c->e_lo = lo;
    c->e_hi = hi;
    c->G = (uint32_t)g;
    c->G_f = (uint32_t)gf;
    if (c->nranks == 1 && (lo != 0 || hi != M) && M) return fail(c, OVL_E_ARG, "remote entries need ovl_comm_init");
    // strand-mirror halving needs entry e ^ 1 to be the other strand of entry e, for every e
    bool pairs = M % 2 == 0 && M > 0;
    for (size_t e = 0; pairs && e < M; e += 2) pairs = !(c->flags[e] & 1) && (c->flags[e + 1] & 1) && c->len[e] == c->len[e + 1];
    c->pairs = pairs ? 1 : 0;

    // the local granule ranges (what travels / what is scanned) and the scan tiles of the local entries
    ScanRanges R{};
    std::vector<uint32_t> stile0;
    try {
        stile0.resize((size_t)(hi - lo) + 1);
    } catch (const std::bad_alloc&) {
        return fail(c, OVL_E_NOMEM, "out of host memory");
    }
    {
        bool seen[2] = {false, false};
        uint64_t ntile = 0;
        for (uint32_t e = lo; e < hi; ++e) {
            const int r = is_twin(e) ? 1 : 0;
            const uint32_t ge = (c->len[e] + 63u) / 64u;
            if (!seen[r]) {
                R.g_begin[r] = c->gran0[e];
                seen[r] = true;
            }
            R.g_end[r] = c->gran0[e] + ge;
            stile0[e - lo] = (uint32_t)ntile;
            ntile += (ge + 31) / 32;
        }
        if (ntile >= 0xfffffff0ull) return fail(c, OVL_E_LIMIT, "too many scan tiles");
        stile0[hi - lo] = (uint32_t)ntile;
        R.n_tiles = (uint32_t)ntile;
    }
    c->ranges = R;

    cudaStream_t st = c->st;
    if (M) {
        OVL_TRY(dev_ensure(c, c->d_src, M * 8));
        OVL_TRY(dev_ensure(c, c->d_len, M * 4));
        OVL_TRY(dev_ensure(c, c->d_flags, M));
        OVL_TRY(dev_ensure(c, c->d_gran0, M * 4));
        OVL_TRY(dev_ensure(c, c->d_ent, M * 8));
        OVL_TRY(dev_ensure(c, c->d_tile0, (M + 1) * 4));
        OVL_TRY(dev_ensure(c, c->d_tile_ent, (nt + 1) * 4));
        OVL_TRY(dev_ensure(c, c->d_stile0, stile0.size() * 4));
        OVL_TRY(dev_ensure(c, c->d_scan_tiles, ((size_t)R.n_tiles + 1) * 16));
        CK(cudaMemcpyAsync(c->d_tile0.p, c->tile0.data(), (M + 1) * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(c->d_src.p, c->src.data(), M * 8, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(c->d_len.p, c->len.data(), M * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(c->d_flags.p, c->flags.data(), M, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(c->d_gran0.p, c->gran0.data(), M * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(c->d_stile0.p, stile0.data(), stile0.size() * 4, cudaMemcpyHostToDevice, st));
        k0_ent_table<<<(unsigned)((M + 255) / 256), 256, 0, st>>>(c->d_gran0.as<uint32_t>(), c->d_len.as<uint32_t>(), c->d_ent.as<uint2>(), (uint32_t)M);
        if (hi > lo) {
            k0_tile_ent<<<c->n_sm * 4, 256, 0, st>>>(c->d_tile0.as<uint32_t>(), c->d_tile_ent.as<uint32_t>(), lo, hi);
            k0_scan_tiles<<<c->n_sm * 4, 256, 0, st>>>(c->d_stile0.as<uint32_t>(), c->d_gran0.as<uint32_t>(), c->d_len.as<uint32_t>(),
                                                       c->d_scan_tiles.as<uint4>(), lo, hi);
        }
        CK(cudaGetLastError());
    }
    CK(cudaStreamSynchronize(st));  // the pageable vectors above must not change under the copy
    c->tables_dirty = false;
    ++c->generation;
    c->cache_l_valid = false;
    c->cand_cap = 0;
    return OVL_OK;
}

// host -> HBM: the staged ASCII records in one copy (the chunked form lives in run_pass)
int upload_ascii(ovl_ctx* c) {
    const size_t ab = c->ascii_used + SLACK;
    CK(cudaEventRecord(c->ev[E_H2D0], c->st));
    if (!c->len.empty() && c->ascii.p) {
        OVL_TRY(dev_ensure(c, c->d_ascii, ab));
        CK(cudaMemcpyAsync(c->d_ascii.p, c->ascii.p, ab, cudaMemcpyHostToDevice, c->st));
    }
    CK(cudaEventRecord(c->ev[E_H2D1], c->st));
    CK(cudaStreamSynchronize(c->st));
    c->stats.ascii_bytes = c->len.empty() ? 0 : ab;
    c->stats.ms_h2d = ev_ms(c, E_H2D0, E_H2D1);
    c->ascii_resident = true;
    return OVL_OK;
}

// 8-byte copy of the current step id into flag word `kind` of peer r (stream ordered behind the
// data it announces; runs on a copy engine, needs no SM)
int signal_copy(ovl_ctx* c, int r, int kind, cudaStream_t s) {
    unsigned long long* dst = &hdr_of(c, r)->flag[kind][c->rank];
    CK(cudaMemcpyAsync(dst, &my_hdr(c)->step_word, 8, cudaMemcpyDefault, s));
    return OVL_OK;
}

int wait_flags(ovl_ctx* c, int kind, cudaStream_t s, uint32_t* launches, uint32_t mask = 0xffffffffu) {
    ovl::k_wait<<<1, 32, 0, s>>>(my_hdr(c)->flag[kind], (uint32_t)c->nranks, (uint32_t)c->rank, mask, c->step,
                                 (unsigned long long)c->opt_timeout_ms * 1000000ull, &my_hdr(c)->abort);
    ++*launches;
    CK(cudaGetLastError());
    return OVL_OK;
}

int signal_all(ovl_ctx* c, int kind, cudaStream_t s, uint32_t* launches) {
    ovl::k_signal<<<1, 32, 0, s>>>(c->d_peerhdr.as<ArenaHdr*>(), (uint32_t)c->nranks, (uint32_t)c->rank, kind, c->step);
    ++*launches;
    CK(cudaGetLastError());
    return OVL_OK;
}

// small publish kernels of the gather / status phase -------------------------------------------
__global__ void __launch_bounds__(32) k9_publish_status(ArenaHdr* const* __restrict__ peer_hdr, uint32_t nr, uint32_t me,
                                                        unsigned long long step) {
    const uint32_t s = threadIdx.x;
    if (s < nr && s != me) {
        const unsigned long long* src = peer_hdr[me]->status[me];
        unsigned long long* dst = peer_hdr[s]->status[me];
        for (int i = 0; i < STATUS_WORDS; ++i) dst[i] = src[i];
        __threadfence_system();
        ovl::st_release_sys(&peer_hdr[s]->flag[ovl::F_STATUS][me], step);
    }
}

// -----------------------------------------------------------------------------------------
// One pass over the entry store.  chunked: the ASCII records are sent host -> HBM in pieces on
// the copy stream and packed piece by piece under the copy (the ovl_overlaps route); otherwise
// they are resident already (ovl_upload).
int run_pass(ovl_ctx* c, uint32_t min_length, bool chunked) {
    using namespace ovl;
    const uint32_t l = min_length < 1 ? 1u : min_length;  // min_length 0 behaves as 1
    const int k = l < 32 ? (int)l : 32;
    const uint32_t M = (uint32_t)c->len.size();
    const int nr = c->nranks, me = c->rank;
    ovl_stats& S = c->stats;
    const float keep_h2d = S.ms_h2d;
    const uint64_t keep_ab = S.ascii_bytes;
    memset(&S, 0, sizeof S);
    S.ms_h2d = keep_h2d;
    S.ascii_bytes = keep_ab;
    S.entries = M;
    S.bases = c->bases;
    S.kmer = (uint32_t)k;
    c->n_rows = 0;
    c->frows_valid = false;
    c->rows_fetched = false;
    c->gathered = false;
    if (M == 0) return OVL_OK;  // the reference segfaults here; we return no rows
    if (nr > MAX_RANKS) return fail(c, OVL_E_LIMIT, "more than %d ranks", MAX_RANKS);
    if (nr > 1 && (int)c->rank_e.size() != nr + 1) return fail(c, OVL_E_ARG, "call ovl_comm_sync_ranges before ovl_overlaps_device");
    const bool mirror = c->opt_mirror && c->pairs == 1;
    S.mirror = mirror ? 1u : 0u;
    MirrorRule mrule{};
    mrule.on = mirror ? 1 : 0;
    mrule.map.n = nr;
    if (nr > 1)
        for (int r = 0; r <= nr; ++r) mrule.map.e[r] = c->rank_e[r];
    // the peers whose forward strands this rank compares against (verifier_rank, ovl_common.cuh)
    uint32_t need_mask = 0;
    for (int s = 0; s < nr; ++s)
        if (ovl::needs_reads_of(nr, me, s)) need_mask |= 1u << s;
    S.granules = c->G;
    KeyFmt fmt;
    fmt.bits_e = ceil_log2((uint64_t)M < 2 ? 2 : M);
    fmt.bits_p = ceil_log2(c->max_len + 1 < 2 ? 2 : c->max_len + 1);
    if (fmt.total() > 64)
        return fail(c, OVL_E_LIMIT, "candidate key needs %d bits (entries %u, longest %llu): over the 64-bit key layout",
                    fmt.total(), M, (unsigned long long)c->max_len);
    // host-known sizes (no device round trip): anchors, scan positions
    if (!c->cache_l_valid || c->cache_l != l) {
        uint64_t na = 0, np = 0;
        for (uint32_t e = 0; e < M; ++e) na += c->len[e] >= l;
        for (uint32_t e = c->e_lo; e < c->e_hi; ++e)
            if (c->len[e] >= l) np += c->len[e] - l + 1;
        c->cache_nanc = na;
        c->cache_npos = np;
        c->cache_l = l;
        c->cache_l_valid = true;
    }
    const uint64_t n_anc = c->cache_nanc;
    S.anchors = n_anc;
    S.scan_positions = c->cache_npos;
    // anchor stride: largest power of two <= min(64, l-k+1); 1 = canonical dense scan.  The table is
    // partitioned over the ranks, so its size per GPU does not depend on N and the largest stride
    // (fewest probes) is the choice; it is only lowered while the partition would outgrow L2 by far
    // and the scan of a smaller stride is still the cheaper side (DRAM-resident table: ~3x the
    // per-key cost; per-probe cost of the filter pass: ~1/8 of a resident insert).
    int slog = 0;
    {
        uint32_t smax = l - (uint32_t)k + 1;  // >= 1
        if (smax > 64) smax = 64;
        if (c->opt_stride > 0 && (uint32_t)c->opt_stride < smax) smax = (uint32_t)c->opt_stride;
        if (c->collect) smax = 1;  // byte accounting is defined on the canonical candidates
        while ((2u << slog) <= smax) ++slog;
        const int cap = max_stride_log2(M);
        if (slog > cap) slog = cap;
        if (c->opt_stride == 0 && slog > 1) {
            const double l2 = 96.0 * 1048576.0;
            const double gran = (double)(c->ranges.g_end[0] - c->ranges.g_begin[0]) + (double)(c->ranges.g_end[1] - c->ranges.g_begin[1]);
            int best = slog;
            double best_cost = 1e300;
            for (int sl = slog; sl >= 1; --sl) {
                const double keys = (double)n_anc * (double)(1u << sl) / nr;
                const double tbytes = 8.0 * (double)(1ull << table_bits((uint64_t)n_anc << sl, nr));
                const double ins = keys * (tbytes > l2 ? 3.0 : 1.0);
                const double probes = gran * (double)(64u >> sl) * 0.125;
                if (ins + probes < best_cost) {
                    best_cost = ins + probes;
                    best = sl;
                }
            }
            slog = best;
        }
    }
    const bool dense = slog == 0;
    S.anchor_stride = 1u << slog;
    const uint64_t n_keys = n_anc << slog;
    const int tbits = table_bits(n_keys, nr);
    const bool smem_filter = dense && n_anc <= (1u << 17);
    const int bm_bits = dense ? dense_filter_bits(n_anc, smem_filter)
                              : slice_bits(n_keys, nr, (uint64_t)c->opt_filter_kb << 10);
    const uint32_t n_local = c->e_hi - c->e_lo;
    const ScanRanges& R = c->ranges;
    const uint64_t n_gran_local = (uint64_t)(R.g_end[0] - R.g_begin[0]) + (R.g_end[1] - R.g_begin[1]);
    const uint64_t probes = n_gran_local * (uint64_t)(64u >> slog);
    // N > 1: a candidate is found by the rank that owns its k-mer, compared by verifier_rank() of its
    // two entries (and, with mirror halving, yields its twins' keys too) and emitted by the rank that
    // owns its A entry: survivors, candidate keys and hit keys go straight from the producing kernel
    // into the consumer rank's list (ovl_common.cuh: Sink / SegList)
    const bool have_anchors = n_anc != 0;

    cudaStream_t st = c->st, sk = c->st_k2, sp = c->st_push, sg = c->st_sig;
    const int nsm = c->n_sm;
    const uint32_t kblocks = (uint32_t)nsm * 8;

    for (int attempt = 0;; ++attempt) {
        if (c->cand_cap == 0) {
            uint64_t cap = c->cand_per_entry * n_local;
            if (cap < c->min_cand_cap) cap = c->min_cand_cap;
            c->cand_cap = cap;
        }
        if (c->cand_cap > 0x7fffffffull) return fail(c, OVL_E_LIMIT, "more than 2^31 candidates on one GPU");
        const uint64_t scan_cap = dense ? 1 : std::min<uint64_t>(probes + 32, 2 * c->cand_cap + probes / 8 + 4096);
        // ---- the arena (and, on N > 1, the peer mappings) -----------------------------------
        const uint64_t key[4] = {c->generation, c->cand_cap, ((uint64_t)nr << 8) | ((uint64_t)slog << 1) | (uint64_t)mirror, c->arena_epoch};
        if (memcmp(key, c->arena_key, sizeof key) != 0 || !c->arena.p) {
            OVL_TRY(connect_arena(c, mirror, scan_cap));
            memcpy(c->arena_key, key, sizeof key);
        }
        const ArenaLayout& L = c->lay;
        ArenaHdr* H = my_hdr(c);
        uint8_t* A = c->arena.as<uint8_t>();
        uint4* d_pre = at<uint4>(A, L.off_pre);
        uint4* d_packed = at<uint4>(A, L.off_packed);
        uint32_t* d_filter = at<uint32_t>(A, L.off_filter);
        Slot* d_table = at<Slot>(A, L.off_table);
        uint32_t* d_next = at<uint32_t>(A, L.off_next);
        Row* d_rows = at<Row>(A, L.off_rows);
        uint64_t hits_room = 0, cand_room = 0;  // hit keys / candidate keys this rank can receive (= rows it can emit / keys it verifies)
        for (int s2 = 0; s2 < nr; ++s2) {
            hits_room += std::min(L.seg_cap[L_HITS], c->peer_lay[s2].seg_cap[L_HITS]);
            cand_room += std::min(L.seg_cap[L_CAND], c->peer_lay[s2].seg_cap[L_CAND]);
        }
        OVL_TRY(dev_ensure(c, c->d_cand, std::max(hits_room, cand_room) * 8, false));  // the candidate keys grouped by A entry (N > 1), then the bucketed / sorted hit keys
        OVL_TRY(dev_ensure(c, c->d_cnt, cand_room * 4, false));   // survivor list of the verifier's head
        OVL_TRY(dev_ensure(c, c->d_big, (size_t)M * 4 + 4, false));
        const size_t M1 = (size_t)M + 1;
        OVL_TRY(dev_ensure(c, c->d_bkt, M1 * (8 + 8 + 4 + 4 + 4)));
        size_t tb = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tb, (uint32_t*)nullptr, (uint64_t*)nullptr, (int)M1, st));
        OVL_TRY(dev_ensure(c, c->d_tmp, tb));
        if (chunked) OVL_TRY(dev_ensure(c, c->d_ascii, c->ascii_used + SLACK));

        uint64_t* ctr = c->d_ctr.as<uint64_t>();
        uint32_t launches = 0, lib = 0;
        ++c->step;
        uint64_t* h_step = c->h_ctr + H_STEP;
        *h_step = c->step;

        // ---- start: counters, step word -----------------------------------------------------
        if (chunked) CK(cudaEventRecord(c->ev[E_H2D0], st));
        CK(cudaEventRecord(c->ev[E_PACK0], st));
        CK(cudaMemsetAsync(ctr, 0, C_COUNT * sizeof(uint64_t), st));
        CK(cudaMemsetAsync(H->seg_n, 0, sizeof H->seg_n, st));
        CK(cudaMemcpyAsync(&H->step_word, h_step, 8, cudaMemcpyHostToDevice, st));
        CK(cudaEventRecord(c->ev_k0, st));
        CK(cudaStreamWaitEvent(sk, c->ev_k0, 0));

        // ---- K1 (+ the prefix table K2 needs), whole or piece by piece under the upload ----------
        // resident reads on N > 1 GPUs: K1 in 4 pieces, so that the forward granules of a piece start
        // travelling to the peers while the next piece is packed
        const int n_chunks = chunked ? std::max(1, std::min(c->opt_chunks, MAX_CHUNKS)) : (nr > 1 ? 4 : 1);
        uint32_t piece_e[MAX_CHUNKS + 1];  // entry boundaries of the pieces
        piece_e[0] = c->e_lo;
        if (!chunked && n_local) {
            // the prefix table of ALL local entries on the main stream AHEAD of K1 -- it is what K2 (and, on
            // N GPUs, every peer's K2) waits for: 13 us of work that took until +0.08..0.14 ms when it had to
            // find SM slots beside the persistent K1 grid
            k2_prefix_ascii<<<(n_local + 127) / 128, 128, 0, st>>>(
                c->d_ascii.as<uint8_t>(), c->d_src.as<uint64_t>(), c->d_len.as<uint32_t>(), c->d_flags.as<uint8_t>(),
                d_pre, c->e_lo, c->e_hi);
            ++launches;
            CK(cudaEventRecord(c->ev_prel, st));
            CK(cudaStreamWaitEvent(sk, c->ev_prel, 0));
        }
        PackParams K;
        K.ascii = c->d_ascii.as<uint8_t>();
        K.src = c->d_src.as<uint64_t>();
        K.ent = c->d_ent.as<uint2>();
        K.flags = c->d_flags.as<uint8_t>();
        K.tile0 = c->d_tile0.as<uint32_t>();
        K.tile_ent = c->d_tile_ent.as<uint32_t>();
        K.packed32 = reinterpret_cast<uint32_t*>(d_packed);
        K.M = M;
        K.err = reinterpret_cast<uint32_t*>(ctr + C_ERR);
        uint32_t e_prev = c->e_lo;
        size_t byte_prev = 0;
        for (int ch = 0; ch < n_chunks; ++ch) {
            // entries of this piece: an even share of the local entries, cut on a record boundary
            uint32_t e_next = ch + 1 == n_chunks ? c->e_hi : c->e_lo + (uint32_t)(((uint64_t)n_local * (ch + 1)) / n_chunks);
            while (e_next < c->e_hi && (c->flags[e_next] & 1)) ++e_next;  // keep a twin with its forward entry
            if (chunked) {
                const size_t byte_next = e_next >= c->e_hi ? c->ascii_used + SLACK : (size_t)c->src[e_next];
                if (byte_next > byte_prev && c->ascii.p)
                    CK(cudaMemcpyAsync(c->d_ascii.as<uint8_t>() + byte_prev, c->ascii.p + byte_prev, byte_next - byte_prev,
                                       cudaMemcpyHostToDevice, c->st_h2d));
                byte_prev = std::max(byte_prev, byte_next);
                CK(cudaEventRecord(c->ev_chunk[ch], c->st_h2d));
                CK(cudaStreamWaitEvent(st, c->ev_chunk[ch], 0));
                CK(cudaStreamWaitEvent(sk, c->ev_chunk[ch], 0));
            }
            if (e_next > e_prev) {
                if (chunked) {  // on K2's stream, beside this piece's K1
                    k2_prefix_ascii<<<(e_next - e_prev + 127) / 128, 128, 0, sk>>>(
                        c->d_ascii.as<uint8_t>(), c->d_src.as<uint64_t>(), c->d_len.as<uint32_t>(), c->d_flags.as<uint8_t>(),
                        d_pre, e_prev, e_next);
                    ++launches;
                }
                const uint32_t t_lo = c->tile0[e_prev], t_hi = c->tile0[e_next];
                if (t_hi > t_lo) {
                    K.t_begin = t_lo;
                    K.t_end = t_hi;
                    uint32_t blocks = (t_hi - t_lo + 7) / 8;
                    const int bpsm = c->opt_pack_bpsm > 0 ? c->opt_pack_bpsm : (nr == 1 ? 8 : 16);
                    if (blocks > (uint32_t)(nsm * bpsm)) blocks = (uint32_t)(nsm * bpsm);
                    k1_pack<<<blocks, 256, 0, st>>>(K);
                    ++launches;
                }
                CK(cudaGetLastError());
            }
            e_prev = e_next;
            piece_e[ch + 1] = e_next;
            if (nr > 1) CK(cudaEventRecord(c->ev_kp[ch], st));
        }
        if (chunked) {
            CK(cudaEventRecord(c->ev[E_H2D1], c->st_h2d));
            c->ascii_resident = true;
            S.ascii_bytes = c->ascii_used + SLACK;
        }
        if (chunked || n_local == 0) CK(cudaEventRecord(c->ev_prel, sk));
        CK(cudaEventRecord(c->ev_pack, st));
        // ---- N > 1: push the prefix slice (small, urgent) and the forward region slice (bulk) -------
        if (nr > 1) {
            CK(cudaStreamWaitEvent(sg, c->ev_prel, 0));
            const uint64_t pre_bytes = (uint64_t)n_local * 32;
            if (pre_bytes) {
                PushTargets PT{};
                PT.nr = (uint32_t)nr;
                PT.me = (uint32_t)me;
                for (int r = 0; r < nr; ++r) PT.dst[r] = at<uint4>(c->peer_base[r], c->peer_lay[r].off_pre + (uint64_t)c->e_lo * 32);
                k_push<<<dim3(16, nr - 1), 256, 0, sg>>>(d_pre + 2ull * c->e_lo, PT, pre_bytes / 16);
                ++launches;
                CK(cudaGetLastError());
            }
            OVL_TRY(signal_all(c, F_PRE, sg, &launches));
            if (c->debug) CK(cudaEventRecord(c->ev_dbg[2], sg));
            // bulk: my slice of the forward region to the ranks that compare against my reads (half
            // of the peers), in ring order (at any moment a rank receives from one sender per stream)
            // and in pieces over several streams so that more than one copy engine drives NVLink
            int dst[MAX_RANKS], nd = 0;
            for (int j = 1; j < nr; ++j) {
                const int r = (me - j + nr) % nr;
                if (needs_reads_of(nr, r, me)) dst[nd++] = r;
            }
            // piece q = the forward granules of K1 piece q (twins stay with their forward entry, so a piece
            // is a contiguous granule range of the forward region); it leaves as soon as it is packed
            auto fwd_gran = [&](uint32_t e) -> uint64_t { return e < c->e_hi ? (uint64_t)c->gran0[e] : (uint64_t)R.g_end[0]; };
            // every destination has its own group of streams (so its copies never queue behind another
            // destination's), every piece is split over the streams of the group (several copy engines
            // per destination), and the group's first stream signals when all of them have drained
            // (measured on this box, profiles/r02o_p2p_bandwidth.txt: ONE stream drives a copy engine at
            // 760 GB/s in each direction at once; splitting over 8 streams gives 650 GB/s, and the 32 small
            // copies of the first form of this loop ran at ~260 GB/s)
            const int spd = 1;
            for (int i = 0; i < nd; ++i) {
                const int r = dst[i];
                for (int q = 0; q < n_chunks; ++q) {
                    const uint64_t g0 = fwd_gran(piece_e[q]), g1 = fwd_gran(piece_e[q + 1]);
                    for (int t = 0; t < spd; ++t) {
                        cudaStream_t sq = c->st_pushv[(i * spd + t) % PUSH_STREAMS];
                        CK(cudaStreamWaitEvent(sq, c->ev_kp[q], 0));
                        const uint64_t h0 = g0 + (g1 - g0) * t / spd, h1 = g0 + (g1 - g0) * (t + 1) / spd;
                        if (h1 > h0)
                            CK(cudaMemcpyAsync(at<uint8_t>(c->peer_base[r], c->peer_lay[r].off_packed + h0 * 16),
                                               A + L.off_packed + h0 * 16, (h1 - h0) * 16, cudaMemcpyDefault, sq));
                    }
                }
                cudaStream_t first = c->st_pushv[(i * spd) % PUSH_STREAMS];
                for (int t = 1; t < spd; ++t) {
                    CK(cudaEventRecord(c->ev_piece[t], c->st_pushv[(i * spd + t) % PUSH_STREAMS]));
                    CK(cudaStreamWaitEvent(first, c->ev_piece[t], 0));
                }
                OVL_TRY(signal_copy(c, r, F_FWD, first));
            }
            for (int i = 1; i < PUSH_STREAMS; ++i) {  // ev_push = every push stream drained
                CK(cudaEventRecord(c->ev_piece[i], c->st_pushv[i]));
                CK(cudaStreamWaitEvent(sp, c->ev_piece[i], 0));
            }
            CK(cudaEventRecord(c->ev_push, sp));
            if (c->debug) CK(cudaEventRecord(c->ev_dbg[4], sp));
        }
        CK(cudaEventRecord(c->ev[E_ANCHOR0], st));

        // ---- K2: this rank's partition of the anchor table, on its own stream under K1 -------------
        TableView T{};
        T.nr = (uint32_t)nr;
        T.tbits = tbits;
        for (int r = 0; r < nr; ++r) {
            T.tbl[r] = at<Slot>(c->peer_base[r], c->peer_lay[r].off_table);
            T.nxt[r] = at<uint32_t>(c->peer_base[r], c->peer_lay[r].off_next);
        }
        if (have_anchors) {
            if (nr > 1) OVL_TRY(wait_flags(c, F_PRE, sk, &launches));
            if (c->debug) CK(cudaEventRecord(c->ev_dbg[0], sk));
            const size_t table_bytes = sizeof(Slot) << tbits;
            CK(cudaMemsetAsync(d_table, 0xff, table_bytes, sk));  // EMPTY_SLOT
            const uint64_t slice_bytes = dense ? (4ull << bm_bits) : ((1ull << bm_bits) / 8);
            uint8_t* my_slice = reinterpret_cast<uint8_t*>(d_filter) + (dense ? 0 : slice_bytes * me);
            CK(cudaMemsetAsync(my_slice, 0, slice_bytes, sk));
            AnchorParams P;
            P.pre64 = reinterpret_cast<const uint64_t*>(d_pre);
            P.ent = c->d_ent.as<uint2>();
            P.M = M;
            P.l = l;
            P.k = k;
            P.stride_log2 = slog;
            P.table = d_table;
            P.tbits = tbits;
            P.next = d_next;
            P.bitmap = d_filter;
            P.bm_bits = bm_bits;
            P.dense = dense ? 1 : 0;
            P.me = (uint32_t)me;
            P.nr = (uint32_t)nr;
            P.flags = reinterpret_cast<unsigned long long*>(ctr + C_FLAGS);
            const uint64_t nthreads = (uint64_t)M << slog;
            unsigned k2_blocks = (unsigned)((nthreads + 255) / 256);
            if (c->opt_k2_bpsm > 0 && k2_blocks > (unsigned)(nsm * c->opt_k2_bpsm)) k2_blocks = (unsigned)(nsm * c->opt_k2_bpsm);
            k2_insert<<<k2_blocks, 256, 0, sk>>>(P);
            ++launches;
            CK(cudaGetLastError());
            CK(cudaEventRecord(c->ev_k2, sk));
            if (c->debug) CK(cudaEventRecord(c->ev_dbg[1], sk));
            if (nr > 1) {  // my filter slice (sparse path) and "my partition is ready" to everyone
                CK(cudaStreamWaitEvent(sg, c->ev_k2, 0));
                if (!dense) {
                    PushTargets PT{};
                    PT.nr = (uint32_t)nr;
                    PT.me = (uint32_t)me;
                    for (int r = 0; r < nr; ++r) PT.dst[r] = at<uint4>(c->peer_base[r], c->peer_lay[r].off_filter + slice_bytes * me);
                    k_push<<<dim3(32, nr - 1), 256, 0, sg>>>(reinterpret_cast<const uint4*>(my_slice), PT, slice_bytes / 16);
                    ++launches;
                    CK(cudaGetLastError());
                }
                OVL_TRY(signal_all(c, F_FILT, sg, &launches));
                CK(cudaEventRecord(c->ev_sig, sg));
                if (c->debug) CK(cudaEventRecord(c->ev_dbg[3], sg));
            }
            CK(cudaStreamWaitEvent(st, c->ev_k2, 0));
            if (nr > 1) OVL_TRY(wait_flags(c, F_FILT, st, &launches));
        } else {
            CK(cudaStreamWaitEvent(st, c->ev_prel, 0));
        }
        CK(cudaEventRecord(c->ev[E_SCAN0], st));

        // ---- the three item lists of this pass: where this rank appends (its segment in every
        //      rank's list), and its own list as the consumers see it ---------------------------------
        const uint64_t off_list[L_COUNT] = {L.off_scan, L.off_cand, L.off_inbox};
        const size_t item_size[L_COUNT] = {sizeof(Surv), 8, 8};
        // producer side: the segment for destination r is in MY arena (slot r of my list region)
        auto fill_sink = [&](auto& sink, int kind) {
            using T = std::remove_pointer_t<std::remove_reference_t<decltype(sink.seg[0])>>;
            sink.nr = (uint32_t)nr;
            for (int r = 0; r < nr; ++r) {
                sink.seg[r] = at<T>(A, off_list[kind] + (uint64_t)r * L.seg_cap[kind] * item_size[kind]);
                sink.cap[r] = L.seg_cap[kind];
            }
            // 1 GPU: the only segment's counter is the list's own; N GPUs: counters in this rank's
            // memory, published with the phase signal
            sink.cnt = nr > 1 ? reinterpret_cast<unsigned long long*>(ctr + C_SEND + kind * MAX_RANKS) : &H->seg_n[kind][0];
        };
        // consumer side: segment s = slot `me` of rank s's list region, read in place (peer mapping)
        auto fill_list = [&](auto& list, int kind) {
            using T = std::remove_const_t<std::remove_pointer_t<std::remove_reference_t<decltype(list.seg[0])>>>;
            for (int s2 = 0; s2 < nr; ++s2) {
                if (s2 == me) {  // my own segment for myself: in place
                    list.seg[s2] = at<T>(A, off_list[kind] + (uint64_t)me * L.seg_cap[kind] * item_size[kind]);
                    list.cap[s2] = L.seg_cap[kind];
                } else {         // slot s2 of my inbox, filled by k_pull
                    list.seg[s2] = at<T>(A, L.off_in[kind] + (uint64_t)s2 * L.seg_cap[kind] * item_size[kind]);
                    list.cap[s2] = std::min(L.seg_cap[kind], c->peer_lay[s2].seg_cap[kind]);
                }
            }
            list.n = H->seg_n[kind];
            list.nr = (uint32_t)nr;
        };
        // phase done: my counts of list `kind` to every rank, then wait for everybody's
        auto exchange = [&](int flag, int kind) -> int {
            if (nr == 1) return OVL_OK;
            k_signal_counts<<<1, 32, 0, st>>>(c->d_peerhdr.as<ArenaHdr*>(), (uint32_t)nr, (uint32_t)me, flag, kind,
                                              reinterpret_cast<const unsigned long long*>(ctr + C_SEND + kind * MAX_RANKS), c->step);
            ++launches;
            CK(cudaGetLastError());
            OVL_TRY(wait_flags(c, flag, st, &launches));
            // pull the segments the others produced for me into my inbox
            PullParams PP{};
            PP.nr = (uint32_t)nr;
            PP.me = (uint32_t)me;
            PP.item_bytes = (uint32_t)item_size[kind];
            PP.n = H->seg_n[kind];
            for (int s2 = 0; s2 < nr; ++s2) {
                const ArenaLayout& PL = c->peer_lay[s2];
                const uint64_t o = kind == L_SCAN ? PL.off_scan : kind == L_CAND ? PL.off_cand : PL.off_inbox;
                PP.src[s2] = at<uint4>(c->peer_base[s2], o + (uint64_t)me * PL.seg_cap[kind] * item_size[kind]);
                PP.dst[s2] = at<uint4>(A, L.off_in[kind] + (uint64_t)s2 * L.seg_cap[kind] * item_size[kind]);
                PP.cap[s2] = std::min(L.seg_cap[kind], PL.seg_cap[kind]);
            }
            k_pull<<<(unsigned)(nsm * 6), 256, 0, st>>>(PP);
            ++launches;
            CK(cudaGetLastError());
            return OVL_OK;
        };

        // ---- K3: scan the local A entries -----------------------------------------------------------
        // (a rank without local reads still takes part in every exchange and resolves the probes of
        // the others against its partition of the table)
        if (have_anchors) {
            ScanParams P;
            P.packed = d_packed;
            P.tiles = c->d_scan_tiles.as<uint4>();
            P.n_tiles = R.n_tiles;
            P.ent = c->d_ent.as<uint2>();
            P.l = l;
            P.k = k;
            P.stride_log2 = slog;
            P.filter = d_filter;
            P.bm_bits = bm_bits;
            P.T = T;
            P.fmt = fmt;
            fill_sink(P.surv, L_SCAN);
            fill_list(P.probes, L_SCAN);
            fill_sink(P.sink, L_CAND);
            P.mirror = mrule;
            if (!dense) {
                if (R.n_tiles) {
                    // persistent grids: exactly the blocks that are resident at once (a static tile split
                    // over 8 blocks per SM with 5 resident ran 1.6 waves: ncu r02e)
                    const uint32_t want = (R.n_tiles + 7) / 8;
                    auto grid = [&](int resident) { return std::min<uint32_t>(want, (uint32_t)(nsm * resident)); };
                    switch (slog) {
                        case 6:
                            if (c->opt_scan_bulk) k3_filter_bulk<6><<<grid(c->occ.bulk6), 256, 0, st>>>(P);
                            else if (nr > 1) k3_filter_wide<6, true><<<grid(c->occ.filter6), 256, 0, st>>>(P);
                            else k3_filter_wide<6, false><<<grid(c->occ.filter6), 256, 0, st>>>(P);
                            break;
                        case 5:
                            if (c->opt_scan_bulk) k3_filter_bulk<5><<<grid(c->occ.bulk5), 256, 0, st>>>(P);
                            else if (nr > 1) k3_filter_wide<5, true><<<grid(c->occ.filter5), 256, 0, st>>>(P);
                            else k3_filter_wide<5, false><<<grid(c->occ.filter5), 256, 0, st>>>(P);
                            break;
                        case 4: k3_filter<4><<<grid(c->occ.filter4), 256, 0, st>>>(P); break;
                        case 3: k3_filter<3><<<grid(c->occ.filter3), 256, 0, st>>>(P); break;
                        default: k3_filter<-1><<<grid(c->occ.filterx), 256, 0, st>>>(P); break;
                    }
                    ++launches;
                    CK(cudaGetLastError());
                }
                if (c->debug) CK(cudaEventRecord(c->ev_dbg[6], st));
                OVL_TRY(exchange(F_SCAN, L_SCAN));  // the survivors went straight into their owners' probe lists
                if (c->debug) CK(cudaEventRecord(c->ev_dbg[5], st));
                k3_resolve<<<(uint32_t)(nsm * c->occ.resolve), 256, 0, st>>>(P);
                ++launches;
                if (c->debug) CK(cudaEventRecord(c->ev_dbg[8], st));
            } else if (R.n_tiles) {
                if (smem_filter)
                    k3_scan<true><<<nsm, 1024, 32 * WBUF * 8 + (4u << bm_bits), st>>>(P);
                else
                    k3_scan<false><<<nsm, 1024, 32 * WBUF * 8, st>>>(P);
                ++launches;
            }
            CK(cudaGetLastError());
            OVL_TRY(exchange(F_CAND, L_CAND));  // the candidate keys went straight into their verifiers' lists
        }
        CK(cudaEventRecord(c->ev[E_VERIFY0], st));

        // ---- K4: verify.  N > 1: the forward strands of the peers this rank compares against must
        //      have landed.  With strand-mirror halving the compared candidate of a twin pair is the
        //      one whose remote entry is a forward strand, so nothing else is needed; without it the
        //      reverse-complement entries of those peers are rebuilt here -------------------------------
        if (nr > 1) {
            OVL_TRY(wait_flags(c, F_FWD, st, &launches, need_mask));
            if (c->debug) CK(cudaEventRecord(c->ev_dbg[9], st));
            if (!mirror && c->G > c->G_f)
                for (int s = 0; s < nr; ++s) {
                    const uint32_t e0 = c->rank_e[s], e1 = c->rank_e[s + 1];
                    if (!((need_mask >> s) & 1u) || e1 <= e0) continue;
                    k1_rc<<<e1 - e0, 128, 0, st>>>(c->d_ent.as<uint2>(), c->d_flags.as<uint8_t>(),
                                                   reinterpret_cast<uint32_t*>(d_packed), e0, e1, M);
                    ++launches;
                }
            CK(cudaGetLastError());
        }
        if (have_anchors) {
            if (c->collect || c->opt_serial_verify) {
                VerifyParams V;
                fill_list(V.keys, L_CAND);
                V.packed = d_packed;
                V.ent = c->d_ent.as<uint2>();
                V.fmt = fmt;
                fill_sink(V.hits, L_HITS);
                V.counters = reinterpret_cast<unsigned long long*>(ctr + C_SUMM);
                V.collect = c->collect;
                V.mirror = mrule;
                k4_verify_serial<<<kblocks, 256, 0, st>>>(V);
                ++launches;
            } else {
                Verify2Params V;
                fill_list(V.keys, L_CAND);
                const bool group = c->opt_group_cands < 0 ? nr > 1 : c->opt_group_cands != 0;
                if (group) {
                    // N > 1: the candidates reach their verifier in short runs (a source warp's survivors
                    // are split over the k-mer owners, and their candidates again over the verifiers), so
                    // the A operand of consecutive candidates is no longer the same entry and stops
                    // coming from L1 (N = 2: verify 0.545 -> 0.65 ms).  Group the keys by A entry first:
                    // the K7 histogram / scatter pair over the candidate list.
                    uint64_t* g_off = c->d_bkt.as<uint64_t>();
                    uint32_t* g_cnt = reinterpret_cast<uint32_t*>(g_off + 2 * M1);
                    BucketParams G{};
                    fill_list(G.hits, L_CAND);
                    G.fmt = fmt;
                    G.a0 = 0;
                    G.M = M;
                    G.cnt = g_cnt;
                    G.off = g_off;
                    G.cur = g_cnt + M1;
                    G.sorted = c->d_cand.as<uint64_t>();
                    CK(cudaMemsetAsync(g_cnt, 0, M1 * 8, st));  // cnt, cur
                    k7_hist<<<kblocks, 256, 0, st>>>(G);
                    size_t tbg = c->d_tmp.cap;
                    CK(cub::DeviceScan::ExclusiveSum(c->d_tmp.p, tbg, g_cnt, g_off, (int)M1, st));
                    k7_scatter<<<kblocks, 256, 0, st>>>(G);
                    launches += 2;
                    lib += 1;
                    CK(cudaGetLastError());
                    if (c->debug) CK(cudaEventRecord(c->ev_dbg[7], st));
                    V.keys.seg[0] = c->d_cand.as<uint64_t>();
                    V.keys.n = reinterpret_cast<const unsigned long long*>(g_off + M);  // total = last element of the exclusive sum
                    V.keys.cap[0] = cand_room;
                    V.keys.nr = 1;
                }
                V.packed64 = reinterpret_cast<const uint64_t*>(d_packed);
                V.ent = c->d_ent.as<uint2>();
                V.fmt = fmt;
                V.surv = c->d_cnt.as<uint32_t>();
                V.n_surv = reinterpret_cast<unsigned long long*>(ctr + C_NSURV);
                fill_sink(V.hits, L_HITS);
                V.mirror = mrule;
                V.tile_ctr = reinterpret_cast<unsigned long long*>(ctr + C_TILE_A);
                k4a_head<<<kblocks, 256, 0, st>>>(V);
                k4b_tail<<<kblocks, 256, 0, st>>>(V);
                launches += 2;
            }
            CK(cudaGetLastError());
        }
        CK(cudaEventRecord(c->ev[E_SORT0], st));
        // the hit keys (a matching candidate's own and its twins') went straight into the hit list of
        // the rank that owns their A entry
        if (have_anchors) OVL_TRY(exchange(F_HITS, L_HITS));
        if (c->debug && nr > 1) CK(cudaEventRecord(c->ev_dbg[10], st));

        // ---- K7: order the hit keys by (a, b, p, flags) and emit rows: k7_bucket.cuh ------------------
        uint64_t* b_off = c->d_bkt.as<uint64_t>();
        uint64_t* b_rowoff = b_off + M1;
        uint32_t* b_cnt = reinterpret_cast<uint32_t*>(b_rowoff + M1);
        uint32_t* b_cur = b_cnt + M1;
        uint32_t* b_rowcnt = b_cur + M1;
        BucketParams B;
        fill_list(B.hits, L_HITS);
        B.fmt = fmt;
        // bucket arrays over THIS rank's A range only (every hit key routed here has its a in it)
        const uint32_t Mb = n_local;
        const size_t Mb1 = (size_t)Mb + 1;
        B.a0 = c->e_lo;
        B.M = Mb;
        B.cnt = b_cnt;
        B.off = b_off;
        B.cur = b_cur;
        B.sorted = c->d_cand.as<uint64_t>();
        B.rowcnt = b_rowcnt;
        B.rowoff = b_rowoff;
        B.big = reinterpret_cast<unsigned long long*>(ctr + C_BIG);
        B.big_list = c->d_big.as<uint32_t>();
        B.ent = c->d_ent.as<uint2>();
        B.rows = d_rows;
        B.rows_cap = hits_room < L.rows_cap ? hits_room : L.rows_cap;  // this rank's share
        uint32_t wblocks = (Mb + 7) / 8;
        if (wblocks < 1) wblocks = 1;
        if (wblocks > (uint32_t)nsm * 16) wblocks = (uint32_t)nsm * 16;
        CK(cudaMemsetAsync(b_cnt, 0, M1 * 12, st));  // cnt, cur, rowcnt
        k7_hist<<<kblocks, 256, 0, st>>>(B);
        tb = c->d_tmp.cap;
        CK(cub::DeviceScan::ExclusiveSum(c->d_tmp.p, tb, b_cnt, b_off, (int)Mb1, st));
        k7_scatter<<<kblocks, 256, 0, st>>>(B);
        k7_sort<<<wblocks, 256, 0, st>>>(B);
        k7_sort_big<<<nsm, 1024, 0, st>>>(B);
        CK(cudaEventRecord(c->ev[E_EMIT0], st));
        tb = c->d_tmp.cap;
        CK(cub::DeviceScan::ExclusiveSum(c->d_tmp.p, tb, b_rowcnt, b_rowoff, (int)Mb1, st));
        k7_emit<<<wblocks, 256, 0, st>>>(B);
        launches += 5;
        lib += 2;
        CK(cudaGetLastError());
        CK(cudaEventRecord(c->ev[E_GATHER0], st));

        // ---- end of the data path.  The rows stay where they were emitted: sharded by A range over
        //      the ranks' HBM like the reads (rank order = canonical order).  A consumer that wants
        //      them in one place pulls them afterwards (ovl_fetch_rows / ovl_comm_gather_rows_device
        //      on rank 0); the row counts travel in the status vectors below ------------------------
        const unsigned long long* d_nrows = reinterpret_cast<const unsigned long long*>(b_rowoff + Mb);
        if (nr > 1) {
            // never leave a push in flight behind the end of the pass
            CK(cudaStreamWaitEvent(st, c->ev_push, 0));
            if (have_anchors) CK(cudaStreamWaitEvent(st, c->ev_sig, 0));
        }
        CK(cudaGetLastError());
        CK(cudaEventRecord(c->ev[E_END], st));

        // ---- status to everyone (one vector per rank), then the single host synchronisation --------
        {
            StatusParams SP;
            SP.st = H->status[me];
            SP.err = reinterpret_cast<const unsigned long long*>(ctr + C_ERR);
            SP.flags = reinterpret_cast<const unsigned long long*>(ctr + C_FLAGS);
            SP.abort_word = &H->abort;
            SP.seg_n = &H->seg_n[0][0];
            SP.seg_snap = reinterpret_cast<unsigned long long*>(ctr + C_SEGN);
            for (int kk = 0; kk < L_COUNT; ++kk)
                for (int s2 = 0; s2 < nr; ++s2) SP.seg_cap[kk][s2] = std::min(L.seg_cap[kk], c->peer_lay[s2].seg_cap[kk]);
            SP.nr = (uint32_t)nr;
            SP.n_surv = reinterpret_cast<const unsigned long long*>(ctr + C_NSURV);
            SP.cap_surv = cand_room;
            SP.n_rows = d_nrows;
            SP.cap_rows = B.rows_cap;
            k9_status<<<1, 32, 0, st>>>(SP);
            ++launches;
        }
        if (nr > 1) {
            k9_publish_status<<<1, 32, 0, st>>>(c->d_peerhdr.as<ArenaHdr*>(), (uint32_t)nr, (uint32_t)me, c->step);
            ++launches;
            OVL_TRY(wait_flags(c, F_STATUS, st, &launches));
        }
        CK(cudaGetLastError());
        uint64_t* h_status = c->h_ctr + H_STATUS;
        uint64_t* h_segn = c->h_ctr + C_SEGN;  // k9_status's snapshot, part of the counter block
        CK(cudaMemcpyAsync(c->h_ctr, ctr, C_COUNT * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(h_status, H->status, sizeof H->status, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));  // the one synchronisation of a pass
        CK(cudaStreamSynchronize(sk));
        if (chunked) {
            CK(cudaStreamSynchronize(c->st_h2d));
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, c->ev[E_H2D0], c->ev[E_H2D1]) == cudaSuccess) S.ms_h2d = ms;
            (void)cudaGetLastError();
        }

        // ---- what every rank now knows about every rank ------------------------------------------------
        S.launches += launches;
        S.lib_launches += lib;
        int bad_rank = -1;
        uint64_t bad_entry = 0, flags_any = 0;
        for (int r = 0; r < nr; ++r) {
            const uint64_t* s = h_status + (size_t)r * STATUS_WORDS;
            if (s[ST_ERR] && bad_rank < 0) {
                bad_rank = r;
                bad_entry = s[ST_ERR] - 1;
            }
            flags_any |= s[ST_FLAGS];
        }
        if (flags_any & STF_TIMEOUT) {
            c->arena_key[0] = ~0ull;  // flags and counters are in an unknown state: re-connect before the next pass
            return fail(c, OVL_E_NCCL, "multi-GPU pass timed out waiting for a peer (step %llu): a rank died or the ranks' calls diverged",
                        (unsigned long long)c->step);
        }
        if (bad_rank >= 0) {  // every rank raises, none is left waiting in a collective
            const uint32_t e = (uint32_t)bad_entry;
            if (bad_rank == me && e < M)
                return fail(c, OVL_E_ALPHABET, "entry %u ('%.*s') holds a byte outside upper-case ACGT", e,
                            (int)(c->id_off[e + 1] - c->id_off[e]), c->id_blob.data() + c->id_off[e]);
            return fail(c, OVL_E_ALPHABET, "entry %u (held by rank %d) holds a byte outside upper-case ACGT", e, bad_rank);
        }
        if (flags_any & STF_TABLE) return fail(c, OVL_E_LIMIT, "anchor table partition overflow: internal error");
        if (flags_any & STF_OVERFLOW) {
            // some rank ran out of room: every rank sees that in the status vectors and repeats the
            // pass; the ranks that overflowed grow first (the arenas are re-published either way)
            if (attempt >= 3) return fail(c, OVL_E_LIMIT, "candidate buffers keep overflowing");
            ++S.retries;
            // the fullest buffer of ANY rank, in 1/1024 of its capacity: every capacity scales with cand_cap,
            // and a cross-rank segment is clipped at min(source's, receiver's) capacity while only the
            // receiver sees it overflow -- so the source has to grow by the receiver's factor as well
            // (a skewed overflow never converged when every rank grew by its own fill only)
            uint64_t fill = 0;
            for (int r = 0; r < nr; ++r) fill = std::max<uint64_t>(fill, h_status[(size_t)r * STATUS_WORDS + ST_FILL]);
            if (fill > 1024) {
                c->cand_cap = (uint64_t)(1.07 * (double)c->cand_cap * (double)fill / 1024.0) + 1024;
                c->min_cand_cap = std::max(c->min_cand_cap, c->cand_cap);
            }
            ++c->arena_epoch;
            continue;
        }
        // items this rank produced (sent) and holds (received), per list
        auto sent = [&](int kind, bool remote_only) {
            uint64_t t = 0;
            if (nr == 1) return remote_only ? 0ull : (unsigned long long)h_segn[kind * MAX_RANKS];
            for (int d = 0; d < nr; ++d)
                if (!remote_only || d != me) t += c->h_ctr[C_SEND + kind * MAX_RANKS + d];
            return (unsigned long long)t;
        };
        auto held = [&](int kind, bool remote_only) {
            uint64_t t = 0;
            for (int r = 0; r < nr; ++r)
                if (!remote_only || r != me) t += h_segn[kind * MAX_RANKS + r];
            return (unsigned long long)t;
        };
        S.scan_survivors = sent(L_SCAN, false);
        S.candidates = sent(L_CAND, false);   // found by this rank
        S.verified = held(L_CAND, false);     // compared by this rank
        S.hits = held(L_HITS, false);         // keys in this rank's hit list
        S.hits_sent = sent(L_HITS, true);
        S.hits_received = held(L_HITS, true);
        S.compared_bases = c->h_ctr[C_SUMM];
        S.operand_bytes = c->h_ctr[C_SUMB];
        S.big_buckets = (uint32_t)c->h_ctr[C_BIG];
        uint64_t total = 0;
        for (int r = 0; r < nr; ++r) {
            c->rowcnt_all[r] = h_status[(size_t)r * STATUS_WORDS + ST_NROWS];
            total += c->rowcnt_all[r];
        }
        S.rows_local = c->rowcnt_all[me];
        // N > 1: rank 0 reports the job's row count; the rows of the other ranks are still in their
        // HBM (gather_rows_pull brings them over when somebody asks for them)
        c->n_rows = nr > 1 ? (me == 0 ? total : 0) : total;
        c->gathered = nr == 1;
        S.rows = c->n_rows;
        break;
    }
    if (c->debug) {
        auto rel = [&](cudaEvent_t e) {
            float ms = -1.f;
            if (cudaEventElapsedTime(&ms, c->ev[E_PACK0], e) != cudaSuccess) {
                (void)cudaGetLastError();
                ms = -1.f;
            }
            return ms;
        };
        const bool mg = c->nranks > 1;
        fprintf(stderr, "[ovl rank %d step %llu] +ms: pack_end %.3f | sk: pre_wait_end %.3f insert_end %.3f | sg: pre_pushed %.3f filt_pushed %.3f | "
                        "sp: fwd_pushed %.3f | st: scan0 %.3f filter_end %.3f probes_in %.3f resolve_end %.3f cands_in(verify0) %.3f fwd_in %.3f grouped %.3f sort0 %.3f "
                        "hits_in %.3f emit0 %.3f end %.3f\n",
                c->rank, (unsigned long long)c->step, rel(c->ev[E_ANCHOR0]), rel(c->ev_dbg[0]), rel(c->ev_dbg[1]),
                mg ? rel(c->ev_dbg[2]) : -1.f, mg ? rel(c->ev_dbg[3]) : -1.f, mg ? rel(c->ev_dbg[4]) : -1.f, rel(c->ev[E_SCAN0]),
                rel(c->ev_dbg[6]), mg ? rel(c->ev_dbg[5]) : -1.f, rel(c->ev_dbg[8]), rel(c->ev[E_VERIFY0]), mg ? rel(c->ev_dbg[9]) : -1.f, mg ? rel(c->ev_dbg[7]) : -1.f,
                rel(c->ev[E_SORT0]), mg ? rel(c->ev_dbg[10]) : -1.f, rel(c->ev[E_EMIT0]), rel(c->ev[E_END]));
    }
    S.ms_pack = ev_ms(c, E_PACK0, E_ANCHOR0);
    S.ms_anchor = ev_ms(c, E_ANCHOR0, E_SCAN0);
    S.ms_scan = ev_ms(c, E_SCAN0, E_VERIFY0);
    S.ms_verify = ev_ms(c, E_VERIFY0, E_SORT0);
    S.ms_sort = ev_ms(c, E_SORT0, E_EMIT0);
    S.ms_emit = ev_ms(c, E_EMIT0, E_GATHER0);
    S.ms_gather = 0.f;  // set by gather_rows_pull when somebody asks for all rows on rank 0
    S.ms_device = ev_ms(c, E_PACK0, E_END);
    if (chunked) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, c->ev[E_H2D1], c->ev[E_END]) == cudaSuccess) S.ms_exposed = ms;
        (void)cudaGetLastError();
    } else {
        S.ms_exposed = S.ms_device;
    }
    return OVL_OK;
}

// Rank 0 of a multi-GPU job: pull the rows of the other ranks (still in their HBM, at the start of
// their row arrays) behind this rank's own rows, over NVLink with the copy engines.  Rank order =
// canonical row order.  Valid until any rank starts its next pass or changes its entry store; the
// other ranks do not take part.
int gather_rows_pull(ovl_ctx* c) {
    if (c->nranks == 1 || c->gathered) return OVL_OK;
    c->gathered = true;
    if (c->rank != 0) return OVL_OK;
    uint64_t off = c->rowcnt_all[0];
    CK(cudaEventRecord(c->ev_dbg[6], c->st));
    CK(cudaEventRecord(c->ev_pack, c->st));
    int used = 0;
    for (int r = 1; r < c->nranks; ++r) {
        const uint64_t n = c->rowcnt_all[r];
        if (n) {
            if (off + n > c->lay.rows_cap) return fail(c, OVL_E_LIMIT, "row gather overflow: internal error");
            cudaStream_t s = c->st_pushv[used % PUSH_STREAMS];
            if (used < PUSH_STREAMS) CK(cudaStreamWaitEvent(s, c->ev_pack, 0));
            CK(cudaMemcpyAsync(c->d_rows.as<uint8_t>() + off * sizeof(ovl::Row),
                               at<uint8_t>(c->peer_base[r], c->peer_lay[r].off_rows), n * sizeof(ovl::Row), cudaMemcpyDefault, s));
            ++used;
        }
        off += n;
    }
    for (int i = 0; i < std::min(used, PUSH_STREAMS); ++i) {
        CK(cudaEventRecord(c->ev_piece[i], c->st_pushv[i]));
        CK(cudaStreamWaitEvent(c->st, c->ev_piece[i], 0));
    }
    CK(cudaEventRecord(c->ev_dbg[7], c->st));
    CK(cudaStreamSynchronize(c->st));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, c->ev_dbg[6], c->ev_dbg[7]));
    c->stats.ms_gather = ms;
    return OVL_OK;
}

}  // namespace

// =========================================================================================
extern "C" {

int ovl_create(int device, ovl_ctx** out) {
    ovl_ctx* c = nullptr;
    if (!out) return fail(c, OVL_E_ARG, "ovl_create: out is NULL");
    *out = nullptr;
    // A pass keeps 9 streams busy (main, K2, signals, 4 push streams, H2D); with the default of 8
    // hardware queues two of them may share one and serialise.  Only effective when this is the
    // first CUDA call of the process; callers that initialise CUDA earlier set it themselves.
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        (void)cudaGetLastError();
        return fail(c, OVL_E_CUDA, "no usable CUDA device (%s); this library has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    if (device < 0 || device >= n) return fail(c, OVL_E_ARG, "device %d out of range (0..%d)", device, n - 1);
    ovl_ctx* x = new (std::nothrow) ovl_ctx();
    if (!x) return fail(c, OVL_E_NOMEM, "out of host memory");
    x->device = device;
    c = x;
    auto bail = [&](int code) {
        g_create_err = x->err;
        ovl_destroy(x);
        return code;
    };
    cudaDeviceProp prop;
    if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
        fail(c, OVL_E_CUDA, "cannot select CUDA device %d: %s", device, cudaGetErrorString(cudaGetLastError()));
        return bail(OVL_E_CUDA);
    }
    if (prop.major < 10) {
        fail(c, OVL_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor);
        return bail(OVL_E_CUDA);
    }
    x->n_sm = prop.multiProcessorCount;
    auto init = [&]() -> int {
        int lo_pri = 0, hi_pri = 0;
        CK(cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));
        CK(cudaStreamCreateWithFlags(&x->st, cudaStreamNonBlocking));
        // K2 is latency bound and small; at high priority its blocks take the SM slots K1's retiring
        // blocks free, so it really runs UNDER K1 (at equal priority it ran behind it: r02f time line)
        CK(cudaStreamCreateWithPriority(&x->st_k2, cudaStreamNonBlocking, hi_pri));
        for (int i = 0; i < PUSH_STREAMS; ++i) {
            CK(cudaStreamCreateWithPriority(&x->st_pushv[i], cudaStreamNonBlocking, hi_pri));
            CK(cudaEventCreateWithFlags(&x->ev_piece[i], cudaEventDisableTiming));
        }
        x->st_push = x->st_pushv[0];
        CK(cudaStreamCreateWithPriority(&x->st_sig, cudaStreamNonBlocking, hi_pri));
        CK(cudaStreamCreateWithFlags(&x->st_h2d, cudaStreamNonBlocking));
        cudaEvent_t* plain[] = {&x->ev_pack, &x->ev_k0, &x->ev_prel, &x->ev_k2, &x->ev_push, &x->ev_sig};
        for (cudaEvent_t* e2 : plain) CK(cudaEventCreateWithFlags(e2, cudaEventDisableTiming));
        for (int i = 0; i < MAX_CHUNKS; ++i) CK(cudaEventCreateWithFlags(&x->ev_chunk[i], cudaEventDisableTiming));
        for (int i = 0; i < MAX_CHUNKS; ++i) CK(cudaEventCreateWithFlags(&x->ev_kp[i], cudaEventDisableTiming));
        for (int i = 0; i < E_COUNT; ++i) CK(cudaEventCreate(&x->ev[i]));
        CK(cudaEventCreate(&x->ev_t0));
        CK(cudaEventCreate(&x->ev_t1));
        for (int i = 0; i < 12; ++i) CK(cudaEventCreate(&x->ev_dbg[i]));
        x->debug = getenv("OVL_DEBUG") ? atoi(getenv("OVL_DEBUG")) : 0;
        if (getenv("OVL_SCAN_BULK")) x->opt_scan_bulk = atoi(getenv("OVL_SCAN_BULK")) != 0;  // A/B switch for whole test runs
        const size_t hwords = H_WORDS;
        CK(cudaHostAlloc(reinterpret_cast<void**>(&x->h_ctr), hwords * sizeof(uint64_t), cudaHostAllocDefault));
        memset(x->h_ctr, 0, hwords * sizeof(uint64_t));
        OVL_TRY(dev_ensure(c, x->d_ctr, C_COUNT * sizeof(uint64_t), false));
        CK(cudaFuncSetAttribute(ovl::k3_scan<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * ovl::WBUF * 8 + 128 * 1024));
        CK(cudaFuncSetAttribute(ovl::k3_scan<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * ovl::WBUF * 8));
        auto resident = [&](auto kernel, int& out_n) -> int {
            int nb = 0;
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, 256, 0));
            out_n = nb < 1 ? 1 : nb;
            return OVL_OK;
        };
        OVL_TRY(resident(ovl::k3_filter_bulk<6>, x->occ.bulk6));
        OVL_TRY(resident(ovl::k3_filter_bulk<5>, x->occ.bulk5));
        OVL_TRY(resident(ovl::k3_filter_wide<6, true>, x->occ.filter6));
        OVL_TRY(resident(ovl::k3_filter_wide<5, true>, x->occ.filter5));
        OVL_TRY(resident(ovl::k3_filter<4>, x->occ.filter4));
        OVL_TRY(resident(ovl::k3_filter<3>, x->occ.filter3));
        OVL_TRY(resident(ovl::k3_filter<-1>, x->occ.filterx));
        OVL_TRY(resident(ovl::k3_resolve, x->occ.resolve));
        OVL_TRY(resident(ovl::k4a_head, x->occ.head));
        OVL_TRY(resident(ovl::k4b_tail, x->occ.tail));
        OVL_TRY(resident(ovl::k1_pack, x->occ.pack));
        return OVL_OK;
    };
    int rc = init();
    if (rc != OVL_OK) return bail(rc);
    *out = x;
    return OVL_OK;
}

void ovl_destroy(ovl_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    close_peers(c);
    if (c->comm) nccl_dyn::api().CommDestroy(c->comm);
    DevBuf* bufs[] = {&c->d_ascii, &c->d_src, &c->d_len, &c->d_flags, &c->d_gran0, &c->d_ent, &c->d_tile0,
                      &c->d_tile_ent, &c->d_stile0, &c->d_scan_tiles, &c->d_ctr, &c->d_tmp, &c->d_cand,
                      &c->d_cnt, &c->d_bkt, &c->d_big, &c->d_peerhdr, &c->d_fmark, &c->d_fkeep, &c->d_foff, &c->d_frows,
                      &c->d_gkeys, &c->d_gseqs, &c->d_gflag, &c->d_goff, &c->d_gedges, &c->d_gmark, &c->arena};
    for (DevBuf* b : bufs)
        if (b->p) cudaFree(b->p);
    if (c->ascii.p) cudaFreeHost(c->ascii.p);
    if (c->h_rows.p) cudaFreeHost(c->h_rows.p);
    if (c->h_frows.p) cudaFreeHost(c->h_frows.p);
    if (c->h_gedges.p) cudaFreeHost(c->h_gedges.p);
    if (c->h_ctr) cudaFreeHost(c->h_ctr);
    for (int i = 0; i < E_COUNT; ++i)
        if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    for (int i = 0; i < MAX_CHUNKS; ++i) {
        if (c->ev_chunk[i]) cudaEventDestroy(c->ev_chunk[i]);
        if (c->ev_kp[i]) cudaEventDestroy(c->ev_kp[i]);
    }
    cudaEvent_t evs[] = {c->ev_pack, c->ev_k0, c->ev_prel, c->ev_k2, c->ev_push, c->ev_sig, c->ev_t0, c->ev_t1,
                         c->ev_dbg[0], c->ev_dbg[1], c->ev_dbg[2], c->ev_dbg[3], c->ev_dbg[4], c->ev_dbg[5], c->ev_dbg[6], c->ev_dbg[7],
                         c->ev_dbg[8], c->ev_dbg[9], c->ev_dbg[10], c->ev_dbg[11]};
    for (cudaEvent_t e : evs)
        if (e) cudaEventDestroy(e);
    cudaStream_t sts[] = {c->st, c->st_k2, c->st_sig, c->st_h2d};
    for (cudaStream_t s : sts)
        if (s) cudaStreamDestroy(s);
    for (int i = 0; i < PUSH_STREAMS; ++i) {
        if (c->st_pushv[i]) cudaStreamDestroy(c->st_pushv[i]);
        if (c->ev_piece[i]) cudaEventDestroy(c->ev_piece[i]);
    }
    (void)cudaGetLastError();
    delete c;
}

const char* ovl_last_error(const ovl_ctx* c) { return c ? c->err.c_str() : g_create_err.c_str(); }

int ovl_reserve(ovl_ctx* c, uint64_t seq_bytes, uint32_t n_entries) {
    if (!c) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    try {
        c->src.reserve(n_entries);
        c->len.reserve(n_entries);
        c->flags.reserve(n_entries);
        c->id_off.reserve((size_t)n_entries + 1);
    } catch (const std::bad_alloc&) {
        return fail(c, OVL_E_NOMEM, "out of host memory");
    }
    return pin_ensure(c, c->ascii, c->ascii_used + seq_bytes + 16ull * n_entries + SLACK, c->ascii_used);
}

int ovl_add_sequence(ovl_ctx* c, const char* id, size_t id_len, const char* seq, size_t seq_len) {
    if (!c || (!id && id_len) || (!seq && seq_len)) return fail(c, OVL_E_ARG, "ovl_add_sequence: NULL argument");
    OVL_TRY(set_dev(c));
    try {
        uint64_t off = 0;
        OVL_TRY(stage_record(c, seq, seq_len, &off));
        return push_entry(c, id, id_len, nullptr, off, seq_len, 0);
    } catch (const std::bad_alloc&) {
        return fail(c, OVL_E_NOMEM, "out of host memory");
    }
}

int ovl_add_read(ovl_ctx* c, const char* name, size_t name_len, const char* seq, size_t seq_len) {
    if (!c || (!name && name_len) || (!seq && seq_len)) return fail(c, OVL_E_ARG, "ovl_add_read: NULL argument");
    OVL_TRY(set_dev(c));
    try {
        uint64_t off = 0;
        OVL_TRY(stage_record(c, seq, seq_len, &off));
        OVL_TRY(push_entry(c, name, name_len, "+", off, seq_len, 0));
        return push_entry(c, name, name_len, "-", off, seq_len, 1);
    } catch (const std::bad_alloc&) {
        return fail(c, OVL_E_NOMEM, "out of host memory");
    }
}

int ovl_add_bulk(ovl_ctx* c, const uint8_t* buf, const uint64_t* offsets, uint32_t n, const char* names,
                 const uint64_t* name_offsets, int both_strands) {
    if (!c || (n && (!buf || !offsets)) || (names && !name_offsets))
        return fail(c, OVL_E_ARG, "ovl_add_bulk: NULL argument");
    if (!n) return OVL_OK;
    OVL_TRY(set_dev(c));
    try {
        OVL_TRY(ovl_reserve(c, offsets[n] - offsets[0], (uint32_t)(c->len.size() + (both_strands ? 2ull * n : n))));
        for (uint32_t i = 0; i < n; ++i) {
            if (offsets[i + 1] < offsets[i]) return fail(c, OVL_E_ARG, "ovl_add_bulk: offsets not monotone at %u", i);
            const size_t sl = (size_t)(offsets[i + 1] - offsets[i]);
            const char* nm = names ? names + name_offsets[i] : nullptr;
            const size_t nl = names ? (size_t)(name_offsets[i + 1] - name_offsets[i]) : 0;
            uint64_t off = 0;
            OVL_TRY(stage_record(c, reinterpret_cast<const char*>(buf + offsets[i]), sl, &off));
            if (both_strands) {
                OVL_TRY(push_entry(c, nm, nl, "+", off, sl, 0));
                OVL_TRY(push_entry(c, nm, nl, "-", off, sl, 1));
            } else {
                OVL_TRY(push_entry(c, nm, nl, nullptr, off, sl, 0));
            }
        }
    } catch (const std::bad_alloc&) {
        return fail(c, OVL_E_NOMEM, "out of host memory");
    }
    return OVL_OK;
}

int ovl_add_remote(ovl_ctx* c, const uint32_t* lengths, uint32_t n, int both_strands) {
    if (!c || (n && !lengths)) return fail(c, OVL_E_ARG, "ovl_add_remote: NULL argument");
    try {
        for (uint32_t i = 0; i < n; ++i) {
            OVL_TRY(push_entry(c, nullptr, 0, both_strands ? "+" : nullptr, NO_SRC, lengths[i], 0));
            if (both_strands) OVL_TRY(push_entry(c, nullptr, 0, "-", NO_SRC, lengths[i], 1));
        }
    } catch (const std::bad_alloc&) {
        return fail(c, OVL_E_NOMEM, "out of host memory");
    }
    return OVL_OK;
}

int ovl_clear(ovl_ctx* c) {
    if (!c) return OVL_E_ARG;
    c->ascii_used = SLACK;
    c->src.clear();
    c->len.clear();
    c->flags.clear();
    c->id_blob.clear();
    c->id_off.assign(1, 0);
    c->gran0.clear();
    c->tile0.clear();
    c->bases = c->max_len = 0;
    c->tables_dirty = true;
    c->ascii_resident = false;
    c->n_rows = 0;
    c->frows_valid = false;
    c->rows_fetched = false;
    c->gathered = false;
    c->e_lo = c->e_hi = 0;
    c->G = c->G_f = 0;
    return OVL_OK;
}

int ovl_entry_count(const ovl_ctx* c, uint32_t* m) {
    if (!c || !m) return OVL_E_ARG;
    *m = (uint32_t)c->len.size();
    return OVL_OK;
}

int ovl_entry_id(const ovl_ctx* c, uint32_t idx, const char** id, size_t* id_len) {
    if (!c || !id || !id_len || idx >= c->len.size()) return OVL_E_ARG;
    *id = c->id_blob.data() + c->id_off[idx];
    *id_len = (size_t)(c->id_off[idx + 1] - c->id_off[idx]);
    return OVL_OK;
}

int ovl_entry_length(const ovl_ctx* c, uint32_t idx, uint32_t* len) {
    if (!c || !len || idx >= c->len.size()) return OVL_E_ARG;
    *len = c->len[idx];
    return OVL_OK;
}

int ovl_set_collect_stats(ovl_ctx* c, int on) {
    if (!c) return OVL_E_ARG;
    c->collect = on ? 1 : 0;
    return OVL_OK;
}

int ovl_set_option(ovl_ctx* c, const char* name, int64_t value) {
    if (!c || !name) return OVL_E_ARG;
    if (!strcmp(name, "anchor_stride")) {
        if (value < 0 || value > 64 || (value & (value - 1))) return fail(c, OVL_E_ARG, "anchor_stride must be 0 (auto) or a power of two <= 64");
        c->opt_stride = (int)value;
        return OVL_OK;
    }
    if (!strcmp(name, "serial_verify")) {
        c->opt_serial_verify = value != 0;
        return OVL_OK;
    }
    if (!strcmp(name, "pack_blocks_per_sm")) {
        if (value < 0 || value > 32) return fail(c, OVL_E_ARG, "pack_blocks_per_sm must be in 0..32 (0 = default)");
        c->opt_pack_bpsm = (int)value;
        return OVL_OK;
    }
    if (!strcmp(name, "mirror")) {
        c->opt_mirror = value != 0;
        return OVL_OK;
    }
    if (!strcmp(name, "filter_kb")) {
        if (value < 4 || value > (1 << 22)) return fail(c, OVL_E_ARG, "filter_kb must be in 4..4194304");
        c->opt_filter_kb = (int)value;
        c->arena_key[0] = ~0ull;  // the arena is sized with it
        return OVL_OK;
    }
    if (!strcmp(name, "group_cands")) {
        c->opt_group_cands = value < 0 ? -1 : (value != 0);
        return OVL_OK;
    }
    if (!strcmp(name, "k2_blocks_per_sm")) {
        if (value < 0 || value > 64) return fail(c, OVL_E_ARG, "k2_blocks_per_sm must be in 0..64");
        c->opt_k2_bpsm = (int)value;
        return OVL_OK;
    }
    if (!strcmp(name, "scan_bulk")) {  // 1: bulk-async tile ring in the stride-32/64 filter pass, 0: register pipeline
        c->opt_scan_bulk = value != 0;
        return OVL_OK;
    }
    if (!strcmp(name, "upload_chunks")) {
        if (value < 1 || value > MAX_CHUNKS) return fail(c, OVL_E_ARG, "upload_chunks must be in 1..%d", MAX_CHUNKS);
        c->opt_chunks = (int)value;
        return OVL_OK;
    }
    if (!strcmp(name, "min_cand_cap")) {  // test hook: a small first capacity forces the overflow -> grow -> repeat route
        if (value < 32) return fail(c, OVL_E_ARG, "min_cand_cap must be >= 32");
        c->min_cand_cap = (uint64_t)value;
        c->cand_cap = 0;
        return OVL_OK;
    }
    if (!strcmp(name, "cand_per_entry")) {  // first candidate capacity per local entry (default 96; config 2 needs ~41)
        if (value < 1 || value > 100000) return fail(c, OVL_E_ARG, "cand_per_entry must be in 1..100000");
        c->cand_per_entry = (uint64_t)value;
        c->cand_cap = 0;
        return OVL_OK;
    }
    if (!strcmp(name, "peer_timeout_ms")) {
        if (value < 1 || value > 600000) return fail(c, OVL_E_ARG, "peer_timeout_ms must be in 1..600000");
        c->opt_timeout_ms = (int)value;
        return OVL_OK;
    }
    return fail(c, OVL_E_ARG, "unknown option '%s'", name);
}

int ovl_get_stats(const ovl_ctx* c, ovl_stats* out) {
    if (!c || !out) return OVL_E_ARG;
    *out = c->stats;
    return OVL_OK;
}

int ovl_timer_start(ovl_ctx* c) {
    if (!c) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    CK(cudaEventRecord(c->ev_t0, c->st));
    return OVL_OK;
}

int ovl_timer_stop(ovl_ctx* c, float* ms) {
    if (!c || !ms) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    CK(cudaEventRecord(c->ev_t1, c->st));
    CK(cudaEventSynchronize(c->ev_t1));
    CK(cudaEventElapsedTime(ms, c->ev_t0, c->ev_t1));
    return OVL_OK;
}

// -----------------------------------------------------------------------------------------
// host -> HBM.  The entry tables travel only when entries were added since the last call; the
// ASCII records only when they are not resident (ovl_overlaps re-sends them on every call: host
// buffers in, like the reference rebuilds its index on every call, src/overlapper.cpp:33-36).
int ovl_upload(ovl_ctx* c) {
    if (!c) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    OVL_TRY(upload_tables(c));
    if (!c->ascii_resident) OVL_TRY(upload_ascii(c));
    return OVL_OK;
}

int ovl_overlaps_device(ovl_ctx* c, uint32_t min_length, uint64_t* n_rows_out) {
    if (!c) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    if (n_rows_out) *n_rows_out = 0;
    if (c->tables_dirty || !c->ascii_resident) return fail(c, OVL_E_ARG, "ovl_overlaps_device: call ovl_upload first");
    OVL_TRY(run_pass(c, min_length, false));
    if (n_rows_out) *n_rows_out = c->n_rows;
    return OVL_OK;
}

int ovl_fetch_rows(ovl_ctx* c, const ovl_row** rows, uint64_t* n_rows) {
    if (!c || !rows || !n_rows) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    *rows = nullptr;
    *n_rows = c->n_rows;
    if (!c->n_rows) return OVL_OK;
    if (!c->rows_fetched) {
        OVL_TRY(gather_rows_pull(c));
        const size_t bytes = c->n_rows * sizeof(ovl_row);
        OVL_TRY(pin_ensure(c, c->h_rows, bytes, 0));
        CK(cudaEventRecord(c->ev[E_D2H0], c->st));
        CK(cudaMemcpyAsync(c->h_rows.p, c->d_rows.p, bytes, cudaMemcpyDeviceToHost, c->st));
        CK(cudaEventRecord(c->ev[E_D2H1], c->st));
        CK(cudaStreamSynchronize(c->st));
        c->stats.ms_d2h = ev_ms(c, E_D2H0, E_D2H1);
        c->rows_fetched = true;
    }
    *rows = reinterpret_cast<const ovl_row*>(c->h_rows.p);
    return OVL_OK;
}

int ovl_overlaps(ovl_ctx* c, uint32_t min_length, const ovl_row** rows, uint64_t* n_rows) {
    if (!c || !rows || !n_rows) return OVL_E_ARG;
    *rows = nullptr;
    *n_rows = 0;
    OVL_TRY(set_dev(c));
    // host buffers in: the records are re-sent on every call, in pieces, and packed under the copy
    OVL_TRY(upload_tables(c));
    OVL_TRY(run_pass(c, min_length, true));
    OVL_TRY(ovl_fetch_rows(c, rows, n_rows));
    c->stats.ms_total = c->stats.ms_h2d + c->stats.ms_exposed + c->stats.ms_d2h;
    return OVL_OK;
}

// Multi-GPU form of ovl_overlaps for a consumer that stays sharded: host buffers in on every rank
// (chunked upload under the pack kernel), the pass (rows of all ranks also land in rank 0's HBM),
// then every rank brings ITS OWN rows to its host over its own PCIe link.  Concatenating the
// ranks' arrays in rank order gives the canonical row order.
int ovl_overlaps_local(ovl_ctx* c, uint32_t min_length, const ovl_row** rows, uint64_t* n_rows) {
    if (!c || !rows || !n_rows) return OVL_E_ARG;
    *rows = nullptr;
    *n_rows = 0;
    OVL_TRY(set_dev(c));
    OVL_TRY(upload_tables(c));
    OVL_TRY(run_pass(c, min_length, true));
    OVL_TRY(ovl_fetch_local_rows(c, rows, n_rows));
    c->stats.ms_total = c->stats.ms_h2d + c->stats.ms_exposed + c->stats.ms_d2h;
    return OVL_OK;
}

// -----------------------------------------------------------------------------------------
int ovl_comm_unique_id(uint8_t id_out[128]) {
    ovl_ctx* c = nullptr;
    if (!id_out) return OVL_E_ARG;
    if (!nccl_dyn::load()) return fail(c, OVL_E_NCCL, "cannot load libnccl: %s", nccl_dyn::load_error());
    ncclUniqueId id;
    static_assert(sizeof(ncclUniqueId) == 128, "NCCL unique id is 128 bytes");
    CKN(nccl_dyn::api().GetUniqueId(&id));
    memcpy(id_out, &id, 128);
    return OVL_OK;
}

int ovl_comm_init(ovl_ctx* c, const uint8_t id[128], int rank, int nranks) {
    if (!c || !id || nranks < 1 || rank < 0 || rank >= nranks) return fail(c, OVL_E_ARG, "ovl_comm_init: bad argument");
    if (nranks > ovl::MAX_RANKS) return fail(c, OVL_E_LIMIT, "more than %d ranks", ovl::MAX_RANKS);
    OVL_TRY(set_dev(c));
    if (!nccl_dyn::load()) return fail(c, OVL_E_NCCL, "cannot load libnccl: %s", nccl_dyn::load_error());
    close_peers(c);
    if (c->comm) {
        nccl_dyn::api().CommDestroy(c->comm);
        c->comm = nullptr;
    }
    ncclUniqueId uid;
    memcpy(&uid, id, 128);
    CKN(nccl_dyn::api().CommInitRank(&c->comm, nranks, uid, rank));
    c->rank = rank;
    c->nranks = nranks;
    c->tables_dirty = true;
    c->arena_key[0] = ~0ull;
    return OVL_OK;
}

// Exchange every rank's local entry range; must tile [0, M) in rank order.
int ovl_comm_sync_ranges(ovl_ctx* c) {
    if (!c) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    c->rank_e.assign((size_t)c->nranks + 1, 0);
    const uint32_t M = (uint32_t)c->len.size();
    if (c->nranks == 1) {
        c->rank_e[1] = M;
        return OVL_OK;
    }
    if (c->tables_dirty) return fail(c, OVL_E_ARG, "ovl_comm_sync_ranges: call ovl_upload first");
    uint32_t mine[2] = {c->e_lo, c->e_hi};
    std::vector<uint32_t> all(2 * (size_t)c->nranks);
    OVL_TRY(allgather_host(c, mine, 8, all.data()));
    uint32_t pos = 0;
    for (int r = 0; r < c->nranks; ++r) {
        uint32_t lo = all[2 * r], hi = all[2 * r + 1];
        if (hi == lo) lo = hi = pos;  // a rank may own nothing
        if (lo != pos) return fail(c, OVL_E_ARG, "rank %d owns entries [%u,%u) but [%u,..) was expected: ranges must tile the entry table in rank order", r, lo, hi, pos);
        c->rank_e[r] = lo;
        pos = hi;
    }
    if (pos != M) return fail(c, OVL_E_ARG, "ranks cover %u of %u entries", pos, M);
    c->rank_e[c->nranks] = M;
    c->arena_key[0] = ~0ull;  // (re)connect at the next pass
    return OVL_OK;
}

// Rank 0: pull the rows of the other ranks of the last pass into this rank's HBM (rank order ==
// canonical order) and report the job's row count; other ranks: no-op.  Not collective.
int ovl_comm_gather_rows_device(ovl_ctx* c, uint64_t* n_rows) {
    if (!c) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    OVL_TRY(gather_rows_pull(c));
    if (n_rows) *n_rows = c->n_rows;
    return OVL_OK;
}

int ovl_comm_gather_rows(ovl_ctx* c, const ovl_row** rows, uint64_t* n_rows) {
    if (!c || !rows || !n_rows) return OVL_E_ARG;
    return ovl_fetch_rows(c, rows, n_rows);
}

// This rank's own rows of the last pass (rank 0: the first rows_local rows of the gathered array;
// other ranks: their row array), device -> pinned host.  Lets every rank of a multi-GPU job bring
// its share to its host over its own PCIe link.
int ovl_fetch_local_rows(ovl_ctx* c, const ovl_row** rows, uint64_t* n_rows) {
    if (!c || !rows || !n_rows) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    *rows = nullptr;
    const uint64_t n = c->stats.rows_local;
    *n_rows = n;
    if (!n) return OVL_OK;
    const size_t bytes = n * sizeof(ovl_row);
    OVL_TRY(pin_ensure(c, c->h_rows, bytes, 0));
    CK(cudaEventRecord(c->ev[E_D2H0], c->st));
    CK(cudaMemcpyAsync(c->h_rows.p, c->d_rows.p, bytes, cudaMemcpyDeviceToHost, c->st));
    CK(cudaEventRecord(c->ev[E_D2H1], c->st));
    CK(cudaStreamSynchronize(c->st));
    c->stats.ms_d2h = ev_ms(c, E_D2H0, E_D2H1);
    c->rows_fetched = false;  // h_rows now holds the local share only
    *rows = reinterpret_cast<const ovl_row*>(c->h_rows.p);
    return OVL_OK;
}

// -----------------------------------------------------------------------------------------
int ovl_debug_packed(ovl_ctx* c, uint32_t idx, uint8_t* out, size_t out_cap, size_t* out_len) {
    if (!c || !out_len || idx >= c->len.size()) return fail(c, OVL_E_ARG, "ovl_debug_packed: bad argument");
    OVL_TRY(set_dev(c));
    if (c->tables_dirty || !c->arena.p || c->gran0.size() != c->len.size()) return fail(c, OVL_E_ARG, "ovl_debug_packed: nothing packed yet");
    const size_t bytes = (size_t)((c->len[idx] + 63u) / 64u) * 16;
    *out_len = bytes;
    if (!bytes) return OVL_OK;
    if (!out || out_cap < bytes) return fail(c, OVL_E_ARG, "ovl_debug_packed: buffer too small (%zu needed)", bytes);
    CK(cudaMemcpy(out, c->arena.as<uint8_t>() + c->lay.off_packed + (size_t)c->gran0[idx] * 16, bytes, cudaMemcpyDeviceToHost));
    return OVL_OK;
}

int ovl_debug_mirror_items(int n_ranks, const uint32_t* rank_e, uint32_t a, uint32_t la, uint32_t b, uint32_t lb,
                           uint32_t p, uint32_t* items, int* verifier) {
    if (n_ranks < 1 || n_ranks > ovl::MAX_RANKS || !rank_e || !items || !verifier || p >= la) return OVL_E_ARG;
    ovl::MirrorRule R{};
    R.on = 1;
    R.map.n = n_ranks;
    for (int r = 0; r <= n_ranks; ++r) R.map.e[r] = rank_e[r];
    *items = ovl::mirror_items(R, a, la, b, lb, p);
    *verifier = ovl::verifier_rank(n_ranks, ovl::owner_of(R.map, a), ovl::owner_of(R.map, b), a, b);
    return OVL_OK;
}

}  // extern "C"

#include "ovl_io.inl"
