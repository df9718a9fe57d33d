// ovl.cu -- host side of libphasm_ovl.so: the C ABI of include/phasm_ovl.h over the sm_100a
// kernels K0..K5.  No torch, no pybind11, no CPU compute path: every ovl_overlaps* call runs
// the CUDA pipeline or fails.
//
// Pipeline of one ovl_overlaps_device() call (reference: ExactOverlapper::overlaps,
// /root/reference/src/overlapper.cpp:28-150):
//
//   K0 k0_gran_entry   granule -> entry map
//   K1 k1_pack         ASCII -> 2-bit granules, "-" entries reverse-complemented on the fly
//   K2 select/keys/sort/build   prefix anchors of all entries with |b| >= l, hash table + bloom
//   K3 k3_scan         every A position with |a|-p >= l probes the anchors -> candidates
//   K4 k4_verify       warp per candidate compares a[p:] with b[0:] -> 64-bit hit keys
//   K5 sort/count/scan/emit     longest suffix-prefix per pair + all containments -> rows
//
// Device-wide sort / select / prefix-sum come from CUB (ships with CUDA 12.9); everything else
// is in the k*.cuh files next to this one.
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
#include "k5_emit.cuh"
#include "k6_exchange.cuh"
#include "k7_bucket.cuh"
#include "k8_filter.cuh"
#include "nccl_dyn.h"

static_assert(sizeof(ovl_row) == sizeof(ovl::Row), "row layouts must agree");
static_assert(sizeof(ovl_row) == 20, "ovl_row is 5 x u32");

namespace {

constexpr uint64_t NO_SRC = ~0ull;  // entry whose bases live on another rank
constexpr size_t SLACK = 16;        // bytes of slack at both ends of the ASCII staging area

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

enum Ev { E_H2D0, E_PACK0, E_ANCHOR0, E_SCAN0, E_VERIFY0, E_SORT0, E_EMIT0, E_END, E_D2H0, E_D2H1, E_COUNT };

// device counter block (one 64-bit word each)
enum Ctr { C_ERR, C_NANC, C_NCAND, C_NHITS, C_SUMM, C_SUMB, C_NSURV, C_TILE_A, C_TILE_B, C_BIG, C_COUNT };  // h_ctr has 2 spare words behind

}  // namespace

struct ovl_ctx {
    int device = 0;
    int n_sm = 148;
    cudaStream_t st = nullptr, st_comm = nullptr, st_k2 = nullptr;
    cudaEvent_t ev[E_COUNT] = {};
    cudaEvent_t ev_comm = nullptr, ev_pack = nullptr, ev_pre = nullptr;
    cudaEvent_t ev_k0 = nullptr, ev_prel = nullptr, ev_k2 = nullptr;
    cudaEvent_t ev_dbg[4] = {};  // OVL_DEBUG=1: timing of the replication on the comm stream  // fork/join of the anchor-build stream
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;  // ovl_timer_*
    std::string err;

    // ---- host entry store ----
    PinBuf ascii;            // [SLACK][record, 16-byte aligned]...[SLACK]
    size_t ascii_used = SLACK;
    std::vector<uint64_t> src;   // byte offset of the entry's record, NO_SRC if remote
    std::vector<uint32_t> len;
    std::vector<uint8_t> flags;  // bit0: reverse complement of the record
    std::vector<char> id_blob;
    std::vector<uint64_t> id_off{0};
    std::vector<uint32_t> gran0;
    std::vector<uint32_t> tile0;  // [M+1] first K1 tile of every entry
    uint64_t bases = 0, max_len = 0;
    bool uploaded = false;
    uint32_t e_lo = 0, e_hi = 0;  // entries whose bases are local (contiguous)

    // ---- device ----
    DevBuf d_ascii, d_src, d_len, d_flags, d_gran0, d_gran_ent, d_ent, d_packed, d_tile0, d_tile_ent, d_xchg, d_pre;
    DevBuf d_table, d_filter, d_next, d_ctr, d_tmp;
    DevBuf d_cand, d_cand2, d_cnt, d_offs, d_rows, d_cnts, d_bkt, d_fwd;
    DevBuf d_fmark, d_fkeep, d_foff, d_frows;  // ovl_filter_rows
    DevBuf d_gath;                             // row gather slots (multi-GPU)
    PinBuf h_frows;
    uint64_t* h_ctr = nullptr;  // pinned mirror of d_ctr
    uint64_t cand_cap = 0;  // candidate keys one pass may emit (d_cand/d_cand2 hold the derived hit capacity)
    uint64_t n_rows = 0;
    PinBuf h_rows;
    bool rows_fetched = false;
    bool gathered = false;  // ovl_comm_gather_rows_device already ran for the current rows
    int collect = 0;
    // host-side per-call scans of the entry table, cached between calls (reset by ovl_upload)
    int cache_pairs = -1;            // 1: entry e^1 is the other strand of entry e, for every e
    uint32_t cache_l = 0;
    bool cache_l_valid = false;
    uint64_t cache_nanc = 0, cache_npos = 0;
    int opt_stride = 0;  // 0 = auto; else upper bound on the anchor stride (1 = dense scan)
    int opt_serial_verify = 0;  // 1: chunk-by-chunk early-exit verifier instead of head/tail
    int opt_mirror = 1;         // 1: strand-mirror halving when every entry has its twin
    int opt_pack_bpsm = 16;     // K1 blocks per SM (fewer leave room for the concurrent anchor build)
    ovl_stats stats{};

    // ---- multi-GPU ----
    ncclComm_t comm = nullptr;
    int rank = 0, nranks = 1;
    std::vector<uint32_t> rank_e;  // [nranks+1] entry range of every rank
    PinBuf h_gather;
};

namespace {

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
    c->uploaded = false;
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

// read the device counter block (synchronises the compute stream)
int read_ctr(ovl_ctx* c) {
    CK(cudaMemcpyAsync(c->h_ctr, c->d_ctr.p, C_COUNT * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    return OVL_OK;
}

// Multi-GPU mirror mode: partition the n_hits keys in d_cand2 by owner rank, ship every run to
// its owner (grouped ncclSend/ncclRecv), leave this rank's keys (own + received, unsorted) in
// d_cand2 and their number in *n_hits.  Collective: every rank must call it.
int exchange_hits(ovl_ctx* c, const ovl::KeyFmt& fmt, uint64_t* n_hits, uint32_t* launches, uint32_t* lib) {
    using namespace ovl;
    (void)lib;
    auto& N = nccl_dyn::api();
    const int R = c->nranks, me = c->rank;
    if (R > MAX_RANKS) return fail(c, OVL_E_LIMIT, "more than %d ranks", MAX_RANKS);
    if ((int)c->rank_e.size() != R + 1) return fail(c, OVL_E_ARG, "call ovl_comm_sync_ranges before ovl_overlaps_device");
    cudaStream_t st = c->st;
    RankMap map;
    map.n = R;
    for (int r = 0; r <= R; ++r) map.e[r] = c->rank_e[r];
    const uint64_t n = *n_hits;
    // device scratch: counts[R] | cursor[R] | matrix[R*R]
    OVL_TRY(dev_ensure(c, c->d_xchg, 8ull * (2 * R + (size_t)R * R), false));
    unsigned long long* d_counts = c->d_xchg.as<unsigned long long>();
    unsigned long long* d_cursor = d_counts + R;
    unsigned long long* d_matrix = d_cursor + R;
    CK(cudaMemsetAsync(d_counts, 0, 16ull * R, st));
    OVL_TRY(dev_ensure(c, c->d_cand, (n ? n : 1) * 8, false));  // partition output (the candidate buffer is free after the verifier)
    if (n) {
        unsigned blocks = (unsigned)((n + 255) / 256);
        if (blocks > (unsigned)c->n_sm * 8) blocks = (unsigned)c->n_sm * 8;
        k6_count<<<blocks, 256, 0, st>>>(c->d_cand2.as<uint64_t>(), n, fmt, map, d_counts);
        k6_scatter<<<blocks, 256, 0, st>>>(c->d_cand2.as<uint64_t>(), n, fmt, map, d_counts, d_cursor, c->d_cand.as<uint64_t>());
        *launches += 2;
        CK(cudaGetLastError());
    }
    CKN(N.AllGather(d_counts, d_matrix, (size_t)R, ncclUint64, c->comm, st));
    std::vector<unsigned long long> mat((size_t)R * R);
    CK(cudaMemcpyAsync(mat.data(), d_matrix, 8ull * R * R, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    // mat[s*R + d] = keys rank s holds for rank d
    uint64_t total = 0;
    std::vector<uint64_t> soff(R + 1, 0), roff(R + 1, 0);
    for (int r = 0; r < R; ++r) {
        soff[r + 1] = soff[r] + mat[(size_t)me * R + r];
        roff[r + 1] = roff[r] + mat[(size_t)r * R + me];
    }
    total = roff[R];
    if (soff[R] != n) return fail(c, OVL_E_NCCL, "hit exchange: partition counted %llu of %llu keys", (unsigned long long)soff[R], (unsigned long long)n);
    OVL_TRY(dev_ensure(c, c->d_cand2, (total ? total : 1) * 8));  // old contents live in d_cand now
    uint64_t* src = c->d_cand.as<uint64_t>();
    uint64_t* dst = c->d_cand2.as<uint64_t>();
    const uint64_t own = mat[(size_t)me * R + me];
    if (own) CK(cudaMemcpyAsync(dst + roff[me], src + soff[me], own * 8, cudaMemcpyDeviceToDevice, st));
    CKN(N.GroupStart());
    for (int r = 0; r < R; ++r) {
        if (r == me) continue;
        const uint64_t ns = mat[(size_t)me * R + r], nr = mat[(size_t)r * R + me];
        if (ns) CKN(N.Send(src + soff[r], ns, ncclUint64, r, c->comm, st));
        if (nr) CKN(N.Recv(dst + roff[r], nr, ncclUint64, r, c->comm, st));
    }
    CKN(N.GroupEnd());
    c->stats.hits_sent = n - own;
    c->stats.hits_received = total - own;
    *n_hits = total;
    return OVL_OK;
}

}  // namespace

// =========================================================================================
extern "C" {

int ovl_create(int device, ovl_ctx** out) {
    ovl_ctx* c = nullptr;
    if (!out) return fail(c, OVL_E_ARG, "ovl_create: out is NULL");
    *out = nullptr;
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
        CK(cudaStreamCreateWithFlags(&x->st, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&x->st_comm, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&x->st_k2, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&x->ev_k0, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&x->ev_prel, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&x->ev_k2, cudaEventDisableTiming));
        for (int i = 0; i < 4; ++i) CK(cudaEventCreate(&x->ev_dbg[i]));
        for (int i = 0; i < E_COUNT; ++i) CK(cudaEventCreate(&x->ev[i]));
        CK(cudaEventCreateWithFlags(&x->ev_comm, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&x->ev_pack, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&x->ev_pre, cudaEventDisableTiming));
        CK(cudaEventCreate(&x->ev_t0));
        CK(cudaEventCreate(&x->ev_t1));
        CK(cudaHostAlloc(reinterpret_cast<void**>(&x->h_ctr), (C_COUNT + 2) * sizeof(uint64_t), cudaHostAllocDefault));
        OVL_TRY(dev_ensure(c, x->d_ctr, C_COUNT * sizeof(uint64_t), false));
        CK(cudaFuncSetAttribute(ovl::k3_scan<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * ovl::WBUF * 8 + 128 * 1024));
        CK(cudaFuncSetAttribute(ovl::k3_scan<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * ovl::WBUF * 8));
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
    if (c->comm) nccl_dyn::api().CommDestroy(c->comm);
    DevBuf* bufs[] = {&c->d_ascii, &c->d_src, &c->d_len, &c->d_flags, &c->d_gran0, &c->d_gran_ent,
                      &c->d_ent, &c->d_packed, &c->d_table, &c->d_filter, &c->d_next, &c->d_ctr,
                      &c->d_tmp, &c->d_cand, &c->d_cand2, &c->d_cnt, &c->d_offs, &c->d_rows, &c->d_cnts,
                      &c->d_tile0, &c->d_tile_ent, &c->d_xchg, &c->d_pre, &c->d_bkt, &c->d_fwd, &c->d_fmark, &c->d_fkeep, &c->d_foff, &c->d_frows, &c->d_gath};
    for (DevBuf* b : bufs)
        if (b->p) cudaFree(b->p);
    if (c->ascii.p) cudaFreeHost(c->ascii.p);
    if (c->h_rows.p) cudaFreeHost(c->h_rows.p);
    if (c->h_gather.p) cudaFreeHost(c->h_gather.p);
    if (c->h_frows.p) cudaFreeHost(c->h_frows.p);
    if (c->h_ctr) cudaFreeHost(c->h_ctr);
    for (int i = 0; i < E_COUNT; ++i)
        if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    if (c->ev_comm) cudaEventDestroy(c->ev_comm);
    if (c->ev_pack) cudaEventDestroy(c->ev_pack);
    if (c->ev_pre) cudaEventDestroy(c->ev_pre);
    if (c->ev_k0) cudaEventDestroy(c->ev_k0);
    if (c->ev_prel) cudaEventDestroy(c->ev_prel);
    if (c->ev_k2) cudaEventDestroy(c->ev_k2);
    for (int i = 0; i < 4; ++i)
        if (c->ev_dbg[i]) cudaEventDestroy(c->ev_dbg[i]);
    if (c->st_k2) cudaStreamDestroy(c->st_k2);
    if (c->ev_t0) cudaEventDestroy(c->ev_t0);
    if (c->ev_t1) cudaEventDestroy(c->ev_t1);
    if (c->st) cudaStreamDestroy(c->st);
    if (c->st_comm) cudaStreamDestroy(c->st_comm);
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
    c->uploaded = false;
    c->n_rows = 0;
    c->rows_fetched = false;
    c->gathered = false;
    c->e_lo = c->e_hi = 0;
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
        if (value < 1 || value > 32) return fail(c, OVL_E_ARG, "pack_blocks_per_sm must be in 1..32");
        c->opt_pack_bpsm = (int)value;
        return OVL_OK;
    }
    if (!strcmp(name, "mirror")) {
        c->opt_mirror = value != 0;
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
// host -> HBM: ASCII records of the local entries + the entry table
int ovl_upload(ovl_ctx* c) {
    if (!c) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    const size_t M = c->len.size();
    // local entry range (contiguous) and granule prefix sum
    uint32_t lo = (uint32_t)M, hi = 0;
    try {
        c->gran0.resize(M + 1);
        c->tile0.resize(M + 1);
    } catch (const std::bad_alloc&) {
        return fail(c, OVL_E_NOMEM, "out of host memory");
    }
    uint64_t g = 0, nt = 0;
    for (size_t e = 0; e < M; ++e) {
        c->gran0[e] = (uint32_t)g;
        c->tile0[e] = (uint32_t)nt;
        const uint64_t ge = (c->len[e] + 63u) / 64u;
        g += ge;
        // K1 tiles: forward local entries only (a twin is written by its forward entry's tiles)
        if (c->src[e] != NO_SRC && !(c->flags[e] & 1)) nt += (4 * ge + ovl::PACK_TILE_UNITS - 1) / ovl::PACK_TILE_UNITS;
        if (c->src[e] != NO_SRC) {
            if (e < lo) lo = (uint32_t)e;
            hi = (uint32_t)e + 1;
        }
    }
    if (g >= 0xffffffffull) return fail(c, OVL_E_LIMIT, "more than 2^32 granules (%llu)", (unsigned long long)g);
    if (M) {
        c->gran0[M] = (uint32_t)g;
        c->tile0[M] = (uint32_t)nt;
    }
    if (lo > hi) lo = hi = 0;
    for (uint32_t e = lo; e < hi; ++e)
        if (c->src[e] == NO_SRC) return fail(c, OVL_E_ARG, "local entries must be one contiguous range (entry %u is remote)", e);
    c->e_lo = lo;
    c->e_hi = hi;
    if (c->nranks == 1 && (lo != 0 || hi != M)) return fail(c, OVL_E_ARG, "remote entries need ovl_comm_init");

    CK(cudaEventRecord(c->ev[E_H2D0], c->st));
    if (M) {
        const size_t ab = c->ascii_used + SLACK;
        OVL_TRY(dev_ensure(c, c->d_ascii, ab));
        OVL_TRY(dev_ensure(c, c->d_src, M * 8));
        OVL_TRY(dev_ensure(c, c->d_len, M * 4));
        OVL_TRY(dev_ensure(c, c->d_flags, M));
        OVL_TRY(dev_ensure(c, c->d_gran0, (M + 1) * 4));
        OVL_TRY(dev_ensure(c, c->d_gran_ent, (g + 1) * 4));
        OVL_TRY(dev_ensure(c, c->d_ent, M * 8));
        OVL_TRY(dev_ensure(c, c->d_packed, (g + 2) * 16));
        OVL_TRY(dev_ensure(c, c->d_tile0, (M + 1) * 4));
        OVL_TRY(dev_ensure(c, c->d_tile_ent, (nt + 1) * 4));
        CK(cudaMemcpyAsync(c->d_tile0.p, c->tile0.data(), (M + 1) * 4, cudaMemcpyHostToDevice, c->st));
        if (c->ascii.p) CK(cudaMemcpyAsync(c->d_ascii.p, c->ascii.p, ab, cudaMemcpyHostToDevice, c->st));
        CK(cudaMemcpyAsync(c->d_src.p, c->src.data(), M * 8, cudaMemcpyHostToDevice, c->st));
        CK(cudaMemcpyAsync(c->d_len.p, c->len.data(), M * 4, cudaMemcpyHostToDevice, c->st));
        CK(cudaMemcpyAsync(c->d_flags.p, c->flags.data(), M, cudaMemcpyHostToDevice, c->st));
        CK(cudaMemcpyAsync(c->d_gran0.p, c->gran0.data(), (M + 1) * 4, cudaMemcpyHostToDevice, c->st));
        c->stats.ascii_bytes = ab;
    }
    CK(cudaEventRecord(c->ev[E_PACK0], c->st));
    CK(cudaStreamSynchronize(c->st));  // the pageable vectors above must not change under the copy
    c->stats.ms_h2d = ev_ms(c, E_H2D0, E_PACK0);
    c->uploaded = true;
    c->cache_pairs = -1;
    c->cache_l_valid = false;
    return OVL_OK;
}

// -----------------------------------------------------------------------------------------
int ovl_overlaps_device(ovl_ctx* c, uint32_t min_length, uint64_t* n_rows_out) {
    using namespace ovl;
    if (!c) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    if (!c->uploaded) return fail(c, OVL_E_ARG, "ovl_overlaps_device: call ovl_upload first");
    const uint32_t l = min_length < 1 ? 1u : min_length;  // min_length 0 behaves as 1
    const int k = l < 32 ? (int)l : 32;
    const uint32_t M = (uint32_t)c->len.size();
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
    c->rows_fetched = false;
    c->gathered = false;
    if (n_rows_out) *n_rows_out = 0;
    if (M == 0) return OVL_OK;  // the reference segfaults here; we return no rows
    // strand-mirror halving needs entry e ^ 1 to be the other strand of entry e, for every e
    if (c->cache_pairs < 0) {
        bool pairs = M % 2 == 0;
        for (uint32_t e = 0; pairs && e < M; e += 2) pairs = !(c->flags[e] & 1) && (c->flags[e + 1] & 1);
        c->cache_pairs = pairs ? 1 : 0;
    }
    const bool mirror = c->opt_mirror && c->cache_pairs == 1;
    S.mirror = mirror ? 1u : 0u;
    ovl::MirrorRule mrule;
    mrule.on = mirror ? 1 : 0;

    const uint32_t G = c->gran0[M];
    S.granules = G;
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
    const uint64_t n_anc = c->cache_nanc, npos = c->cache_npos;
    S.anchors = n_anc;
    S.scan_positions = npos;
    // anchor stride: largest power of two <= min(64, l-k+1); 1 = canonical dense scan
    int slog = 0;
    {
        uint32_t smax = l - (uint32_t)k + 1;  // >= 1
        if (smax > 64) smax = 64;
        if (c->opt_stride > 0 && (uint32_t)c->opt_stride < smax) smax = (uint32_t)c->opt_stride;
        if (c->collect) smax = 1;  // byte accounting is defined on the canonical candidates
        while ((2u << slog) <= smax) ++slog;
        while (slog > 0 && (((uint64_t)M << slog) > 0xfffffff0ull)) --slog;
        // Auto mode: the table build costs ~ (anchor entries x stride) inserts on EVERY rank (and
        // much more per insert once the table outgrows L2), the scan ~ (own granules x 64/stride)
        // filter probes.  Pick the stride with the smallest modelled sum; both tables are B200
        // measurements on config 2 (55 200 anchor entries, 12.9 M granules; N=1 and N=2 columns of
        // profiles/r01i_stride_sweep.txt).  A dense scan (stride 1) is only taken when l < 33
        // forces it.
        if (c->opt_stride == 0 && slog > 1) {
            const uint64_t n_b = n_anc;
            const double gl = (double)(c->gran0[c->e_hi] - c->gran0[c->e_lo]) / 12.9e6;
            const double al = (double)n_b / 55200.0;
            static const double scan_ms[7] = {0, 1.45, 0.98, 0.56, 0.40, 0.34, 0.27};      // stride 2^i
            static const double anchor_ms[7] = {0, 0.01, 0.02, 0.03, 0.06, 0.13, 0.17};    // stride 2^i: what K2 adds beyond the concurrent K1
            int best = slog;
            double best_cost = 1e30;
            for (int sl = slog; sl >= 1; --sl) {
                const double keys = (double)n_b * (double)(1u << sl);
                const double table_bytes = 8.0 * (double)(1ull << ceil_log2((uint64_t)(2 * keys) + 1));
                double a_ms = al * anchor_ms[sl];
                if (table_bytes > 64.0 * 1048576.0) a_ms *= 2.5;  // table no longer L2-resident
                if (c->nranks > 1) a_ms *= 1.4;  // the concurrent replication of the reads competes for L2
                const double cost = a_ms + gl * scan_ms[sl];
                if (cost < best_cost) {
                    best_cost = cost;
                    best = sl;
                }
            }
            slog = best;
        }
    }
    const bool dense = slog == 0;
    S.anchor_stride = 1u << slog;
    cudaStream_t st = c->st;
    uint64_t* ctr = c->d_ctr.as<uint64_t>();
    uint32_t launches = 0, lib = 0;
    const int nsm = c->n_sm;

    CK(cudaEventRecord(c->ev[E_PACK0], st));
    CK(cudaMemsetAsync(ctr, 0, C_COUNT * sizeof(uint64_t), st));
    // ---- K0 + K1: pack the local entries ------------------------------------------------
    const uint32_t g_lo = c->gran0[c->e_lo], g_hi = c->gran0[c->e_hi];
    // multi-GPU with paired entries: the ranks exchange compact FORWARD granules only (k1_expand)
    const bool fwd_x = c->nranks > 1 && c->cache_pairs == 1;
    if (fwd_x) OVL_TRY(dev_ensure(c, c->d_fwd, ((size_t)G / 2 + 2) * 16));
    k0_ent_table<<<(M + 255) / 256, 256, 0, st>>>(c->d_gran0.as<uint32_t>(), c->d_len.as<uint32_t>(), c->d_ent.as<uint2>(), M);
    ++launches;
    // ---- fork: the anchor-build stream.  The prefix table (K2's only input) is packed straight
    //      from the ASCII records, so K2 does not wait for K1 -------------------------------------
    cudaStream_t sk = c->st_k2;
    CK(cudaEventRecord(c->ev_k0, st));
    CK(cudaStreamWaitEvent(sk, c->ev_k0, 0));
    OVL_TRY(dev_ensure(c, c->d_pre, (size_t)M * 32));
    if (c->e_hi > c->e_lo) {
        k2_prefix_ascii<<<(c->e_hi - c->e_lo + 127) / 128, 128, 0, sk>>>(
            c->d_ascii.as<uint8_t>(), c->d_src.as<uint64_t>(), c->d_len.as<uint32_t>(), c->d_flags.as<uint8_t>(),
            c->d_pre.as<uint4>(), c->e_lo, c->e_hi);
        ++launches;
        CK(cudaGetLastError());
    }
    CK(cudaEventRecord(c->ev_prel, sk));
    cudaEvent_t k2_last = c->ev_prel;  // last event on the anchor-build stream: every exit path joins it
    if (c->e_hi > c->e_lo) {
        k0_gran_entry<<<nsm * 4, 256, 0, st>>>(c->d_gran0.as<uint32_t>(), c->d_gran_ent.as<uint32_t>(),
                                               c->d_tile0.as<uint32_t>(), c->d_tile_ent.as<uint32_t>(), c->e_lo, c->e_hi);
        ++launches;
        const uint32_t t_lo = c->tile0[c->e_lo], t_hi = c->tile0[c->e_hi];
        if (t_hi > t_lo) {
            PackParams K;
            K.ascii = c->d_ascii.as<uint8_t>();
            K.src = c->d_src.as<uint64_t>();
            K.ent = c->d_ent.as<uint2>();
            K.flags = c->d_flags.as<uint8_t>();
            K.tile0 = c->d_tile0.as<uint32_t>();
            K.tile_ent = c->d_tile_ent.as<uint32_t>();
            K.packed32 = c->d_packed.as<uint32_t>();
            K.fwd32 = fwd_x ? c->d_fwd.as<uint32_t>() : nullptr;
            K.M = M;
            K.t_begin = t_lo;
            K.t_end = t_hi;
            K.err = reinterpret_cast<uint32_t*>(ctr + C_ERR);
            uint32_t blocks = (t_hi - t_lo + 7) / 8;
            if (blocks > (uint32_t)(nsm * c->opt_pack_bpsm)) blocks = (uint32_t)(nsm * c->opt_pack_bpsm);
            k1_pack<<<blocks, 256, 0, st>>>(K);
            ++launches;
        }
    }
    CK(cudaMemsetAsync(c->d_packed.as<uint4>() + G, 0, 32, st));  // terminator granules
    CK(cudaGetLastError());
    // ---- multi-GPU: replicate on the comm stream (1) the prefix table, which K2 waits for, then
    //      (2) the packed granules of every rank, which only K4 needs: that transfer runs under
    //      K2 and K3 -------------------------------------------------------------------------
    bool wait_packed = false;
    if (c->nranks > 1) {
        auto& N = nccl_dyn::api();
        if ((int)c->rank_e.size() != c->nranks + 1) return fail(c, OVL_E_ARG, "call ovl_comm_sync_ranges before ovl_overlaps_device");
        CK(cudaStreamWaitEvent(c->st_comm, c->ev_prel, 0));
        CKN(N.GroupStart());
        for (int r = 0; r < c->nranks; ++r) {
            const uint32_t a = c->rank_e[r], b = c->rank_e[r + 1];
            if (b > a) {
                uint4* p = c->d_pre.as<uint4>() + 2ull * a;
                CKN(N.Broadcast(p, p, (size_t)(b - a) * 32, ncclUint8, r, c->comm, c->st_comm));
            }
        }
        CKN(N.GroupEnd());
        CK(cudaEventRecord(c->ev_pre, c->st_comm));
        CK(cudaStreamWaitEvent(sk, c->ev_pre, 0));  // K2 reads every rank's prefixes
        CK(cudaEventRecord(c->ev_pack, st));        // K1 done: the packed granules may travel
        CK(cudaStreamWaitEvent(c->st_comm, c->ev_pack, 0));
        CK(cudaEventRecord(c->ev_dbg[0], c->st_comm));
        if (fwd_x) {
            // equal shares (the usual case): one in-place all-gather; else one broadcast per rank
            bool equal = true;
            const uint32_t cnt0 = (c->gran0[c->rank_e[1]] - c->gran0[c->rank_e[0]]) / 2;
            for (int r = 0; r < c->nranks; ++r)
                equal = equal && (c->gran0[c->rank_e[r + 1]] - c->gran0[c->rank_e[r]]) / 2 == cnt0;
            if (equal && cnt0) {
                uint4* all = c->d_fwd.as<uint4>();
                CKN(N.AllGather(all + (size_t)c->rank * cnt0, all, (size_t)cnt0 * 16, ncclUint8, c->comm, c->st_comm));
            } else {
                CKN(N.GroupStart());
                for (int r = 0; r < c->nranks; ++r) {
                    const uint32_t a = c->gran0[c->rank_e[r]] / 2, b = c->gran0[c->rank_e[r + 1]] / 2;
                    if (b > a) {
                        uint4* p = c->d_fwd.as<uint4>() + a;
                        CKN(N.Broadcast(p, p, (size_t)(b - a) * 16, ncclUint8, r, c->comm, c->st_comm));
                    }
                }
                CKN(N.GroupEnd());
            }
            CK(cudaEventRecord(c->ev_dbg[1], c->st_comm));
            // both strands of the remote reads, rebuilt locally: below and above the own range
            const uint32_t pr[2][2] = {{0u, c->e_lo / 2}, {c->e_hi / 2, M / 2}};
            for (int i = 0; i < 2; ++i)
                if (pr[i][1] > pr[i][0]) {
                    k1_expand<<<pr[i][1] - pr[i][0], 128, 0, c->st_comm>>>(
                        c->d_fwd.as<uint4>(), c->d_gran0.as<uint32_t>(), c->d_len.as<uint32_t>(),
                        c->d_packed.as<uint32_t>(), pr[i][0], pr[i][1]);
                    ++launches;
                }
            CK(cudaGetLastError());
        } else {
            CKN(N.GroupStart());
            for (int r = 0; r < c->nranks; ++r) {
                const uint32_t a = c->gran0[c->rank_e[r]], b = c->gran0[c->rank_e[r + 1]];
                if (b > a) {
                    uint4* p = c->d_packed.as<uint4>() + a;
                    CKN(N.Broadcast(p, p, (size_t)(b - a) * 16, ncclUint8, r, c->comm, c->st_comm));
                }
            }
            CKN(N.GroupEnd());
        }
        CK(cudaEventRecord(c->ev_dbg[2], c->st_comm));
        CK(cudaEventRecord(c->ev_comm, c->st_comm));
        wait_packed = true;
    }
    CK(cudaEventRecord(c->ev[E_ANCHOR0], st));

    auto stage_times = [&]() {
        S.launches = launches;
        S.lib_launches = lib;
        S.ms_pack = ev_ms(c, E_PACK0, E_ANCHOR0);
        S.ms_anchor = ev_ms(c, E_ANCHOR0, E_SCAN0);
        S.ms_scan = ev_ms(c, E_SCAN0, E_VERIFY0);
        S.ms_verify = ev_ms(c, E_VERIFY0, E_SORT0);
        S.ms_sort = ev_ms(c, E_SORT0, E_EMIT0);
        S.ms_emit = ev_ms(c, E_EMIT0, E_END);
        S.ms_device = ev_ms(c, E_PACK0, E_END);
    };
    auto check_alphabet = [&]() -> int {
        if (c->h_ctr[C_ERR]) {
            const uint32_t e = (uint32_t)c->h_ctr[C_ERR] - 1;
            return fail(c, OVL_E_ALPHABET, "entry %u ('%.*s') holds a byte outside upper-case ACGT", e,
                        (int)(c->id_off[e + 1] - c->id_off[e]), c->id_blob.data() + c->id_off[e]);
        }
        return OVL_OK;
    };
    // record the remaining stage events from `from` on, wait, fill the stage times
    auto finish_from = [&](int from) -> int {
        if (wait_packed) CK(cudaStreamWaitEvent(st, c->ev_comm, 0));  // never leave the replication in flight
        CK(cudaStreamWaitEvent(st, k2_last, 0));                      // ... nor the anchor-build stream
        for (int e = from; e <= E_END; ++e) CK(cudaEventRecord(c->ev[e], st));
        OVL_TRY(read_ctr(c));
        OVL_TRY(check_alphabet());
        stage_times();
        return OVL_OK;
    };
    if (n_anc == 0) return finish_from(E_SCAN0);  // same decision on every rank
    // multi-GPU + mirror: hits are exchanged between ranks, so a rank without own work must
    // still walk through the collective
    const bool xchg = mirror && c->nranks > 1;
    const bool have_a = g_hi > g_lo;
    if (!have_a && !xchg) return finish_from(E_SCAN0);

    // ---- K2: anchor table ---------------------------------------------------------------------
    const uint64_t n_keys = n_anc << slog;
    const int tbits = ceil_log2(2 * n_keys) < 6 ? 6 : ceil_log2(2 * n_keys);
    if (tbits > 31) return fail(c, OVL_E_LIMIT, "anchor table of 2^%d slots is over the layout limit", tbits);
    const bool smem_filter = dense && n_anc <= (1u << 17);
    int bm_bits;  // dense: log2 words of the blocked bloom filter; sparse: log2 bits of the bitmap
    if (dense) {
        bm_bits = smem_filter ? ceil_log2(n_anc) : ceil_log2(2 * n_anc);
        if (bm_bits < 5) bm_bits = 5;
        if (smem_filter && bm_bits > 15) bm_bits = 15;
        if (bm_bits > 27) bm_bits = 27;
    } else {
        bm_bits = ceil_log2(32 * n_keys);
        if (bm_bits < 15) bm_bits = 15;
        if (bm_bits > 31) bm_bits = 31;
    }
    if (have_a) {
        const size_t filter_bytes = dense ? (4ull << bm_bits) : ((1ull << bm_bits) / 8);
        const size_t table_bytes = sizeof(Slot) * (1ull << tbits);
        OVL_TRY(dev_ensure(c, c->d_table, table_bytes));
        OVL_TRY(dev_ensure(c, c->d_filter, filter_bytes));
        OVL_TRY(dev_ensure(c, c->d_next, ((uint64_t)M << slog) * 4));
        CK(cudaMemsetAsync(c->d_table.p, 0xff, table_bytes, sk));  // EMPTY_SLOT
        CK(cudaMemsetAsync(c->d_filter.p, 0, filter_bytes, sk));
        {
            AnchorParams A;
            A.pre64 = reinterpret_cast<const uint64_t*>(c->d_pre.p);
            A.ent = c->d_ent.as<uint2>();
            A.M = M;
            A.l = l;
            A.k = k;
            A.stride_log2 = slog;
            A.table = c->d_table.as<Slot>();
            A.tbits = tbits;
            A.next = c->d_next.as<uint32_t>();
            A.bitmap = c->d_filter.as<uint32_t>();
            A.bm_bits = bm_bits;
            A.dense = dense ? 1 : 0;
            A.n_anchors = reinterpret_cast<unsigned long long*>(ctr + C_NANC);
            const uint64_t nthreads = (uint64_t)M << slog;
            k2_insert<<<(unsigned)((nthreads + 255) / 256), 256, 0, sk>>>(A);
            ++launches;
            CK(cudaGetLastError());
        }
        // ---- join: the scan needs the table ------------------------------------------------------
        CK(cudaEventRecord(c->ev_k2, sk));
        k2_last = c->ev_k2;
        CK(cudaStreamWaitEvent(st, c->ev_k2, 0));
        CK(cudaEventRecord(c->ev[E_SCAN0], st));

    } else {
        CK(cudaStreamWaitEvent(st, k2_last, 0));
        CK(cudaEventRecord(c->ev[E_SCAN0], st));
    }

    // ---- K3 .. K5 without host round trips ---------------------------------------------------
    // Every buffer is sized by CAPACITY and every kernel reads its input size from the device
    // counter block, so a single-GPU pass synchronises with the host once, at its end.  Only the
    // accounting / serial-verifier modes and the multi-GPU hit exchange (NCCL needs the counts)
    // read counters in between.  A candidate overflow is detected at the end: the buffers grow and
    // the pass from K3 on is repeated.
    const bool host_n = c->collect || c->opt_serial_verify || xchg;
    if (c->cand_cap == 0) {
        uint64_t cap = 96ull * (c->e_hi - c->e_lo);
        if (cap < (1ull << 20)) cap = 1ull << 20;
        c->cand_cap = cap;
    }
    uint64_t n_cand = 0, n_hits = 0, n_rows = 0, hit_cap = 0;
    const uint32_t kblocks = (uint32_t)nsm * 8;
    for (int attempt = 0;; ++attempt) {
        if (c->cand_cap > 0x7fffffffull) return fail(c, OVL_E_LIMIT, "more than 2^31 candidates on one GPU");
        hit_cap = mirror ? 3 * c->cand_cap : c->cand_cap;  // own key + SP twin + CONT twin
        OVL_TRY(dev_ensure(c, c->d_cand, hit_cap * 8, false));   // candidates, later the sorted hit keys
        OVL_TRY(dev_ensure(c, c->d_cand2, hit_cap * 8, false));  // hit keys
        OVL_TRY(dev_ensure(c, c->d_cnt, c->cand_cap * 4, false));  // survivor list
        OVL_TRY(dev_ensure(c, c->d_rows, hit_cap * sizeof(Row), false));
        if (attempt) CK(cudaMemsetAsync(ctr + C_NCAND, 0, (C_COUNT - C_NCAND) * sizeof(uint64_t), st));
        if (have_a) {
            ScanParams P;
            P.packed = c->d_packed.as<uint4>();
            P.gran_ent = c->d_gran_ent.as<uint32_t>();
            P.ent = c->d_ent.as<uint2>();
            P.g_begin = g_lo;
            P.g_end = g_hi;
            P.l = l;
            P.k = k;
            P.stride_log2 = slog;
            P.filter = c->d_filter.as<uint32_t>();
            P.bm_bits = bm_bits;
            P.table = c->d_table.as<Slot>();
            P.tbits = tbits;
            P.next = c->d_next.as<uint32_t>();
            P.fmt = fmt;
            P.sink.out = c->d_cand.as<uint64_t>();
            P.sink.n_out = reinterpret_cast<unsigned long long*>(ctr + C_NCAND);
            P.sink.cap = c->cand_cap;
            P.mirror = mrule;
            if (!dense) {
                const uint32_t tiles = (g_hi - g_lo + 31) / 32;
                uint32_t blocks = (tiles + 7) / 8;
                if (blocks > (uint32_t)nsm * 8) blocks = (uint32_t)nsm * 8;
                switch (slog) {
                    case 6: k3_probe<6><<<blocks, 256, 0, st>>>(P); break;
                    case 5: k3_probe<5><<<blocks, 256, 0, st>>>(P); break;
                    case 4: k3_probe<4><<<blocks, 256, 0, st>>>(P); break;
                    case 3: k3_probe<3><<<blocks, 256, 0, st>>>(P); break;
                    default: k3_probe<-1><<<blocks, 256, 0, st>>>(P); break;
                }
            } else if (smem_filter) {
                k3_scan<true><<<nsm, 1024, 32 * WBUF * 8 + (4u << bm_bits), st>>>(P);
            } else {
                k3_scan<false><<<nsm, 1024, 32 * WBUF * 8, st>>>(P);
            }
            ++launches;
            CK(cudaGetLastError());
        }
        if (host_n) {
            OVL_TRY(read_ctr(c));
            OVL_TRY(check_alphabet());
            n_cand = c->h_ctr[C_NCAND];
            if (n_cand > c->cand_cap) {
                if (attempt >= 2) return fail(c, OVL_E_LIMIT, "candidate buffer overflow persists (%llu)", (unsigned long long)n_cand);
                ++S.retries;
                c->cand_cap = n_cand + n_cand / 16;
                continue;
            }
        }

        // ---- K4: verify -------------------------------------------------------------------
        CK(cudaEventRecord(c->ev[E_VERIFY0], st));
        if (wait_packed) CK(cudaStreamWaitEvent(st, c->ev_comm, 0));  // the operands of other ranks' entries
        CandSink hits;
        hits.out = c->d_cand2.as<uint64_t>();
        hits.n_out = reinterpret_cast<unsigned long long*>(ctr + C_NHITS);
        hits.cap = hit_cap;
        if (c->collect || c->opt_serial_verify) {
            if (n_cand) {
                VerifyParams V;
                V.keys = c->d_cand.as<uint64_t>();
                V.n = n_cand;
                V.packed = c->d_packed.as<uint4>();
                V.ent = c->d_ent.as<uint2>();
                V.fmt = fmt;
                V.hits = hits;
                V.counters = reinterpret_cast<unsigned long long*>(ctr + C_SUMM);
                V.collect = c->collect;
                V.mirror = mrule;
                uint64_t blocks = (n_cand + 255) / 256;
                if (blocks > kblocks) blocks = kblocks;
                k4_verify_serial<<<(unsigned)blocks, 256, 0, st>>>(V);
                ++launches;
            }
        } else {
            Verify2Params V;
            V.keys = c->d_cand.as<uint64_t>();
            V.n_ptr = reinterpret_cast<const unsigned long long*>(ctr + C_NCAND);
            V.cap = c->cand_cap;
            V.packed64 = reinterpret_cast<const uint64_t*>(c->d_packed.p);
            V.ent = c->d_ent.as<uint2>();
            V.fmt = fmt;
            V.surv = c->d_cnt.as<uint32_t>();
            V.n_surv = reinterpret_cast<unsigned long long*>(ctr + C_NSURV);
            V.hits = hits;
            V.mirror = mrule;
            V.tile_ctr = reinterpret_cast<unsigned long long*>(ctr + C_TILE_A);
            k4a_head<<<kblocks, 256, 0, st>>>(V);
            k4b_tail<<<kblocks, 256, 0, st>>>(V);
            launches += 2;
        }
        CK(cudaGetLastError());
        CK(cudaEventRecord(c->ev[E_SORT0], st));
        uint64_t keys_cap = hit_cap;  // capacity of d_cand2 in keys, as the bucket kernels see it
        if (host_n) {
            OVL_TRY(read_ctr(c));
            n_hits = c->h_ctr[C_NHITS];
            S.compared_bases = c->h_ctr[C_SUMM];
            S.operand_bytes = c->h_ctr[C_SUMB];
            if (n_hits > hit_cap) return fail(c, OVL_E_LIMIT, "hit list overflow (%llu > %llu): internal error", (unsigned long long)n_hits, (unsigned long long)hit_cap);
            // multi-GPU mirror mode: every hit key goes to the rank that owns its A entry
            if (xchg) {
                OVL_TRY(exchange_hits(c, fmt, &n_hits, &launches, &lib));
                if (n_hits > 0xfffffff0ull) return fail(c, OVL_E_LIMIT, "more than 2^32 hit keys on one GPU");
                c->h_ctr[C_NHITS] = n_hits;
                CK(cudaMemcpyAsync(ctr + C_NHITS, &c->h_ctr[C_NHITS], 8, cudaMemcpyHostToDevice, st));
                keys_cap = c->d_cand2.cap / 8;
                OVL_TRY(dev_ensure(c, c->d_cand, n_hits * 8, false));
                OVL_TRY(dev_ensure(c, c->d_rows, n_hits * sizeof(Row), false));
            }
        }

        // ---- order the hit keys by (a, b, p, flags) and emit rows: k7_bucket.cuh ------------------
        const size_t M1 = (size_t)M + 1;
        OVL_TRY(dev_ensure(c, c->d_bkt, M1 * (8 + 8 + 4 + 4 + 4)));
        uint64_t* b_off = c->d_bkt.as<uint64_t>();
        uint64_t* b_rowoff = b_off + M1;
        uint32_t* b_cnt = reinterpret_cast<uint32_t*>(b_rowoff + M1);
        uint32_t* b_cur = b_cnt + M1;
        uint32_t* b_rowcnt = b_cur + M1;
        size_t tb = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tb, b_cnt, b_off, (int)M1, st));
        OVL_TRY(dev_ensure(c, c->d_tmp, tb));
        BucketParams B;
        B.hits = c->d_cand2.as<uint64_t>();
        B.n_ptr = reinterpret_cast<const unsigned long long*>(ctr + C_NHITS);
        B.cap = keys_cap;
        B.fmt = fmt;
        B.M = M;
        B.cnt = b_cnt;
        B.off = b_off;
        B.cur = b_cur;
        B.sorted = c->d_cand.as<uint64_t>();
        B.rowcnt = b_rowcnt;
        B.rowoff = b_rowoff;
        B.big = reinterpret_cast<unsigned long long*>(ctr + C_BIG);
        B.ent = c->d_ent.as<uint2>();
        B.rows = c->d_rows.as<Row>();
        B.rows_cap = c->d_rows.cap / sizeof(Row);
        uint32_t wblocks = (M + 7) / 8;
        if (wblocks > (uint32_t)nsm * 16) wblocks = (uint32_t)nsm * 16;
        CK(cudaMemsetAsync(b_cnt, 0, M1 * 12, st));  // cnt, cur, rowcnt
        k7_hist<<<kblocks, 256, 0, st>>>(B);
        tb = c->d_tmp.cap;
        CK(cub::DeviceScan::ExclusiveSum(c->d_tmp.p, tb, b_cnt, b_off, (int)M1, st));
        k7_scatter<<<kblocks, 256, 0, st>>>(B);
        k7_sort<<<wblocks, 256, 0, st>>>(B);
        CK(cudaEventRecord(c->ev[E_EMIT0], st));
        tb = c->d_tmp.cap;
        CK(cub::DeviceScan::ExclusiveSum(c->d_tmp.p, tb, b_rowcnt, b_rowoff, (int)M1, st));
        k7_emit<<<wblocks, 256, 0, st>>>(B);
        launches += 4;
        lib += 2;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(&c->h_ctr[C_COUNT], b_rowoff + M, 8, cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(c->ev[E_END], st));
        OVL_TRY(read_ctr(c));  // the one synchronisation of a single-GPU pass
        OVL_TRY(check_alphabet());
        if (!host_n) {
            n_cand = c->h_ctr[C_NCAND];
            if (n_cand > c->cand_cap) {
                if (attempt >= 2) return fail(c, OVL_E_LIMIT, "candidate buffer overflow persists (%llu)", (unsigned long long)n_cand);
                ++S.retries;
                c->cand_cap = n_cand + n_cand / 16;
                continue;
            }
            n_hits = c->h_ctr[C_NHITS];
            if (n_hits > hit_cap) return fail(c, OVL_E_LIMIT, "hit list overflow (%llu > %llu): internal error", (unsigned long long)n_hits, (unsigned long long)hit_cap);
        }
        n_rows = c->h_ctr[C_COUNT];
        break;
    }
    S.candidates = n_cand;
    S.verified = n_cand;
    S.hits = n_hits;

    // ---- an A entry with more than BUCKET_MAX hits: redo the ordering with the device-wide radix
    //      sort (library call) and the per-key K5 kernels ------------------------------------------
    if (c->h_ctr[C_BIG] && n_hits) {
        S.radix_fallback = 1;
        uint64_t* keys = c->d_cand.as<uint64_t>();
        size_t tb = 0;
        CK(cub::DeviceRadixSort::SortKeys(nullptr, tb, c->d_cand2.as<uint64_t>(), keys, (int)n_hits, 0, fmt.total(), st));
        OVL_TRY(dev_ensure(c, c->d_tmp, tb));
        tb = c->d_tmp.cap;
        CK(cub::DeviceRadixSort::SortKeys(c->d_tmp.p, tb, c->d_cand2.as<uint64_t>(), keys, (int)n_hits, 0, fmt.total(), st));
        ++lib;
        OVL_TRY(dev_ensure(c, c->d_cnt, n_hits * 4));
        OVL_TRY(dev_ensure(c, c->d_offs, (n_hits + 1) * 8));
        const unsigned nb = (unsigned)((n_hits + 255) / 256);
        k5_count<<<nb, 256, 0, st>>>(keys, n_hits, fmt, c->d_cnt.as<uint32_t>());
        ++launches;
        tb = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tb, c->d_cnt.as<uint32_t>(), c->d_offs.as<uint64_t>(), (int)n_hits, st));
        OVL_TRY(dev_ensure(c, c->d_tmp, tb));
        tb = c->d_tmp.cap;
        CK(cub::DeviceScan::ExclusiveSum(c->d_tmp.p, tb, c->d_cnt.as<uint32_t>(), c->d_offs.as<uint64_t>(), (int)n_hits, st));
        ++lib;
        CK(cudaMemcpyAsync(&c->h_ctr[C_COUNT], c->d_offs.as<uint64_t>() + (n_hits - 1), 8, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&c->h_ctr[C_COUNT + 1], c->d_cnt.as<uint32_t>() + (n_hits - 1), 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        n_rows = c->h_ctr[C_COUNT] + (c->h_ctr[C_COUNT + 1] & 0xffffffffull);
        if (n_rows) {
            OVL_TRY(dev_ensure(c, c->d_rows, n_rows * sizeof(Row)));
            k5_emit<<<nb, 256, 0, st>>>(keys, n_hits, fmt, c->d_offs.as<uint64_t>(), c->d_cnt.as<uint32_t>(),
                                        c->d_ent.as<uint2>(), c->d_rows.as<Row>());
            ++launches;
            CK(cudaGetLastError());
        }
        CK(cudaEventRecord(c->ev[E_END], st));
        CK(cudaStreamSynchronize(st));
    }
    c->n_rows = n_rows;
    S.rows = n_rows;
    stage_times();
    if (wait_packed && getenv("OVL_DEBUG")) {
        float t_pk = 0, t_bc = 0, t_ex = 0, t_v = 0;
        cudaEventElapsedTime(&t_pk, c->ev[E_PACK0], c->ev_dbg[0]);
        if (fwd_x) cudaEventElapsedTime(&t_bc, c->ev_dbg[0], c->ev_dbg[1]);
        cudaEventElapsedTime(&t_ex, fwd_x ? c->ev_dbg[1] : c->ev_dbg[0], c->ev_dbg[2]);
        cudaEventElapsedTime(&t_v, c->ev[E_PACK0], c->ev[E_VERIFY0]);
        (void)cudaGetLastError();
        fprintf(stderr, "[ovl rank %d] comm stream: starts +%.3f ms, broadcast %.3f ms, %s %.3f ms; verify stage starts +%.3f ms\n",
                c->rank, t_pk, t_bc, fwd_x ? "expand" : "replicate", t_ex, t_v);
    }
    if (n_rows_out) *n_rows_out = n_rows;
    return OVL_OK;
}

int ovl_fetch_rows(ovl_ctx* c, const ovl_row** rows, uint64_t* n_rows) {
    if (!c || !rows || !n_rows) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    *rows = nullptr;
    *n_rows = c->n_rows;
    if (!c->n_rows) return OVL_OK;
    if (!c->rows_fetched) {
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
    // host buffers in: the entry store is re-sent on every call, like the reference rebuilds
    // its index on every call (src/overlapper.cpp:33-36)
    OVL_TRY(ovl_upload(c));
    uint64_t n = 0;
    OVL_TRY(ovl_overlaps_device(c, min_length, &n));
    OVL_TRY(ovl_fetch_rows(c, rows, n_rows));
    c->stats.ms_total = c->stats.ms_h2d + c->stats.ms_device + c->stats.ms_d2h;
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
    OVL_TRY(set_dev(c));
    if (!nccl_dyn::load()) return fail(c, OVL_E_NCCL, "cannot load libnccl: %s", nccl_dyn::load_error());
    if (c->comm) {
        nccl_dyn::api().CommDestroy(c->comm);
        c->comm = nullptr;
    }
    ncclUniqueId uid;
    memcpy(&uid, id, 128);
    CKN(nccl_dyn::api().CommInitRank(&c->comm, nranks, uid, rank));
    c->rank = rank;
    c->nranks = nranks;
    c->uploaded = false;
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
    if (!c->uploaded) return fail(c, OVL_E_ARG, "ovl_comm_sync_ranges: call ovl_upload first");
    auto& N = nccl_dyn::api();
    OVL_TRY(dev_ensure(c, c->d_tmp, 8ull * (c->nranks + 1)));
    uint32_t mine[2] = {c->e_lo, c->e_hi};
    uint32_t* d = c->d_tmp.as<uint32_t>();
    CK(cudaMemcpyAsync(d + 2 * c->rank, mine, 8, cudaMemcpyHostToDevice, c->st));
    CKN(N.AllGather(d + 2 * c->rank, d, 2, ncclUint32, c->comm, c->st));
    std::vector<uint32_t> all(2 * (size_t)c->nranks);
    CK(cudaMemcpyAsync(all.data(), d, 8ull * c->nranks, cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
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
    return OVL_OK;
}

// Gather every rank's rows into rank 0's HBM (rank order == canonical order); rows stay on the
// device.  Idempotent per ovl_overlaps_device call.
int ovl_comm_gather_rows_device(ovl_ctx* c, uint64_t* n_rows) {
    if (!c) return OVL_E_ARG;
    OVL_TRY(set_dev(c));
    if (n_rows) *n_rows = 0;
    if (c->nranks == 1 || c->gathered) {
        if (n_rows) *n_rows = c->n_rows;
        return OVL_OK;
    }
    auto& N = nccl_dyn::api();
    const int R = c->nranks;
    // counts
    OVL_TRY(dev_ensure(c, c->d_cnts, 8ull * R));
    uint64_t* d = c->d_cnts.as<uint64_t>();
    c->h_ctr[C_COUNT] = c->n_rows;
    CK(cudaMemcpyAsync(d + c->rank, &c->h_ctr[C_COUNT], 8, cudaMemcpyHostToDevice, c->st));
    CKN(N.AllGather(d + c->rank, d, 1, ncclUint64, c->comm, c->st));
    std::vector<uint64_t> cnt(R);
    CK(cudaMemcpyAsync(cnt.data(), d, 8ull * R, cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    uint64_t total = 0;
    for (int r = 0; r < R; ++r) total += cnt[r];
    uint64_t maxc = 0;
    for (int r = 0; r < R; ++r) maxc = cnt[r] > maxc ? cnt[r] : maxc;
    // Rows travel by ONE in-place all-gather of equal, padded slots (a collective uses every NCCL
    // channel; the grouped point-to-point receives it replaces reached ~40 GB/s per peer, 1.1 ms for
    // the 32 MB of a 2-rank run).  Above 8 GB of slots fall back to point-to-point.
    const bool by_allgather = maxc > 0 && (uint64_t)R * maxc * sizeof(ovl_row) <= (8ull << 30);
    if (by_allgather) {
        OVL_TRY(dev_ensure(c, c->d_gath, (size_t)R * maxc * sizeof(ovl_row)));
        ovl_row* G = c->d_gath.as<ovl_row>();
        if (c->n_rows) CK(cudaMemcpyAsync(G + (size_t)c->rank * maxc, c->d_rows.p, c->n_rows * sizeof(ovl_row), cudaMemcpyDeviceToDevice, c->st));
        CKN(N.AllGather(G + (size_t)c->rank * maxc, G, (size_t)maxc * sizeof(ovl_row), ncclUint8, c->comm, c->st));
    }
    if (c->rank == 0) {
        // grow d_rows to hold everything, keeping the local rows at the front
        if (total * sizeof(ovl_row) > c->d_rows.cap) {
            DevBuf nb;
            OVL_TRY(dev_ensure(c, nb, total * sizeof(ovl_row)));
            if (c->n_rows) CK(cudaMemcpyAsync(nb.p, c->d_rows.p, c->n_rows * sizeof(ovl_row), cudaMemcpyDeviceToDevice, c->st));
            CK(cudaStreamSynchronize(c->st));
            if (c->d_rows.p) CK(cudaFree(c->d_rows.p));
            c->d_rows = nb;
        }
        uint64_t off = cnt[0];
        if (by_allgather) {
            const ovl_row* G = c->d_gath.as<ovl_row>();
            for (int r = 1; r < R; ++r) {
                if (cnt[r]) CK(cudaMemcpyAsync(c->d_rows.as<ovl_row>() + off, G + (size_t)r * maxc, cnt[r] * sizeof(ovl_row), cudaMemcpyDeviceToDevice, c->st));
                off += cnt[r];
            }
        } else {
            CKN(N.GroupStart());
            for (int r = 1; r < R; ++r) {
                if (cnt[r]) CKN(N.Recv(c->d_rows.as<ovl_row>() + off, cnt[r] * sizeof(ovl_row), ncclUint8, r, c->comm, c->st));
                off += cnt[r];
            }
            CKN(N.GroupEnd());
        }
        c->n_rows = total;
        c->stats.rows = total;
    } else {
        if (c->n_rows && !by_allgather) {
            CKN(N.GroupStart());
            CKN(N.Send(c->d_rows.p, c->n_rows * sizeof(ovl_row), ncclUint8, 0, c->comm, c->st));
            CKN(N.GroupEnd());
        }
        c->n_rows = 0;
    }
    CK(cudaStreamSynchronize(c->st));
    c->rows_fetched = false;
    c->gathered = true;
    if (n_rows) *n_rows = c->n_rows;
    return OVL_OK;
}

int ovl_comm_gather_rows(ovl_ctx* c, const ovl_row** rows, uint64_t* n_rows) {
    if (!c || !rows || !n_rows) return OVL_E_ARG;
    *rows = nullptr;
    *n_rows = 0;
    OVL_TRY(ovl_comm_gather_rows_device(c, nullptr));
    return ovl_fetch_rows(c, rows, n_rows);
}

// -----------------------------------------------------------------------------------------
int ovl_debug_packed(ovl_ctx* c, uint32_t idx, uint8_t* out, size_t out_cap, size_t* out_len) {
    if (!c || !out_len || idx >= c->len.size()) return fail(c, OVL_E_ARG, "ovl_debug_packed: bad argument");
    OVL_TRY(set_dev(c));
    if (!c->uploaded || c->gran0.size() != c->len.size() + 1) return fail(c, OVL_E_ARG, "ovl_debug_packed: nothing packed yet");
    const size_t bytes = (size_t)(c->gran0[idx + 1] - c->gran0[idx]) * 16;
    *out_len = bytes;
    if (!bytes) return OVL_OK;
    if (!out || out_cap < bytes) return fail(c, OVL_E_ARG, "ovl_debug_packed: buffer too small (%zu needed)", bytes);
    CK(cudaMemcpy(out, c->d_packed.as<uint8_t>() + (size_t)c->gran0[idx] * 16, bytes, cudaMemcpyDeviceToHost));
    return OVL_OK;
}

}  // extern "C"

#include "ovl_io.inl"
