// k3_scan.cuh -- K3: streaming k-mer scan of the A entries against the anchor table.
//
// Replaces the suffix-tree DFS push phase (/root/reference/src/overlapper.cpp:38-62): instead
// of walking a suffix tree to learn which suffixes of A end at a node on the path to B's leaf,
// A positions look their k-mer up among the indexed B k-mers (k2_anchor.cuh).
//
// Mapping: a warp owns a tile of 32 consecutive granules (2048 bases) of ONE local entry (one
// 16-byte tile record from k0_scan_tiles tells it which); each lane loads its granule with ONE
// coalesced 128-bit load (512 B per warp) and takes the following word from the next lane by
// shuffle (the k-mer carry across granules).  Everything about the entry is warp-uniform.
//
// The sparse scan (stride >= 2; stride 64 whenever l >= 95) is split in two kernels, because the
// two halves want different shapes:
//
//   k3_filter   streams the packed reads (0.25 B/base per strand) and tests the k-mer at every
//               stride-th position against the L2-resident one-bit-per-key bitmap.  The dependent
//               chain of a tile is granule -> bitmap word, nothing else; the few positions that
//               pass are appended as 16-byte SURVIVOR records {hash, a, q}.
//   k3_resolve  one thread per survivor: slot in the owner's table partition (a peer-mapped
//               pointer when the owner is another GPU: one NVLink round trip for a singleton
//               key) -> anchor list -> candidate keys, with every lane of a warp busy.
//
// (The single-kernel form of round 1 kept a warp on its tile through filter -> slot -> list with
// ~30 % of the lanes holding a hit: 6 us per tile, 12 % of the HBM roofline.)
//
//   k3_scan     (stride 1: l < 33, or the canonical-candidate accounting pass): every position
//               rolls its k-mer and tests a blocked bloom filter held in shared memory; filter
//               hits are resolved in place, warp-converged, against the table.  ~14 integer
//               instructions per position: issue bound.
//
// Candidates are 64-bit keys [a|b|p|00]; they are staged per warp in shared memory and
// flushed with one global atomic per ~224 keys.
#pragma once
#include "k1_pack.cuh"
#include "k2_anchor.cuh"
#include "k6_exchange.cuh"
#include "ovl_common.cuh"

namespace ovl {

// survivor of the filter pass: x = k-mer hash, y = (a << 32) | q
typedef ulonglong2 Surv;
typedef Sink<Surv> SurvSink;

struct ScanParams {
    const uint4* packed;        // granules
    const uint4* tiles;         // [n_tiles] {first granule, entry, length, index of the first granule in the entry}
    uint32_t n_tiles;
    const uint2* ent;           // [M] {gran0, len}
    uint32_t l;                 // min overlap length (>= 1)
    int k;                      // anchor length
    int stride_log2;
    const uint32_t* filter;     // sparse: bitmap words; dense: bloom words
    int bm_bits;                // sparse: log2 bits of one slice; dense: log2 words
    TableView T;
    KeyFmt fmt;
    SurvSink surv;              // k3_filter: survivors to the rank that owns their k-mer
    SegList<Surv> probes;       // k3_resolve: the survivors routed to this rank
    CandSink sink;              // candidate keys to the rank that compares them
    MirrorRule mirror;          // on: emit only the candidates that verify for themselves
};

// candidates of one (a, q) probe whose key resolved to list pointer `ptr` (warp-converged walk)
__device__ __forceinline__ uint32_t cand_dest(const ScanParams& P, uint64_t key) {
    const uint32_t a = P.fmt.a(key), b = P.fmt.b(key);
    return (uint32_t)verifier_rank(P.mirror.map.n, owner_of(P.mirror.map, a), owner_of(P.mirror.map, b), a, b);
}

__device__ __forceinline__ void walk_list(const ScanParams& P, uint32_t owner, uint32_t ptr, uint32_t a, uint32_t la,
                                          uint32_t q, uint64_t* buf, uint32_t& nbuf, uint32_t lane, RouteScratch* rs) {
    const int sl = P.stride_log2;
    const uint32_t smask = (1u << sl) - 1u;
    while (__any_sync(FULL, ptr != NIL)) {
        bool have = false;
        uint64_t key = 0;
        if (ptr != NIL) {
            const uint32_t id = ptr & ID_MASK;
            const uint32_t b = id >> sl, delta = id & smask;
            const uint32_t nxt = (ptr & HAS_NEXT) ? __ldg(P.T.nxt[owner] + id) : NIL;
            // candidate p = q - delta: needs p >= 0, |a| - p >= l, and b != a
            // (identity by index: overlapper.cpp:72-74,101-104)
            if (q >= delta && b != a && la - (q - delta) >= P.l) {
                have = !P.mirror.on || mirror_items(P.mirror, a, la, b, __ldg(P.ent + b).y, q - delta) != 0;
                key = P.fmt.make(a, b, q - delta);
            }
            ptr = nxt;
        }
        wbuf_push(buf, nbuf, P.sink, lane, have, key, rs, [&](uint64_t k2) { return cand_dest(P, k2); });
    }
}

// ---- sparse path, pass 1 ------------------------------------------------------------------
// Stride 64 / 32 (one or two probes per lane, the k-mers sit at fixed word offsets): a three-stage
// software pipeline per warp.  In iteration i the warp
//   (1) requests the tile record of tile i+3 and the granule of tile i+2,
//   (2) hashes the k-mers of tile i+1 (its granule was requested one iteration ago) and requests
//       their bitmap words -- all of them before any is looked at,
//   (3) tests the bitmap words of tile i (requested one iteration ago) and places the survivors with
//       one ballot per probe.
// No load is consumed in the iteration that issued it (ncu r02e, previous two-stage form: 38 % of
// the stall samples sat on the two bitmap words, which were also loaded one after the other).
__device__ __forceinline__ uint32_t bitmap_word(uint64_t h, uint32_t nr, int slice_log2) {
    return (owner_of_hash(h, nr) << (slice_log2 - 5)) | (((uint32_t)h & ((1u << slice_log2) - 1u)) >> 5);
}

template <int SL, bool MULTI>
__global__ void __launch_bounds__(256, 4) k3_filter_wide(const ScanParams P) {
    static_assert(SL == 5 || SL == 6, "one or two probes per granule");
    constexpr int NP = SL == 6 ? 1 : 2;
    constexpr uint32_t SCAP = WBUF / 2;
    __shared__ Surv s_buf[8][SCAP];
    __shared__ RouteScratch s_rs[8];
    RouteScratch* rs = &s_rs[threadIdx.x >> 5];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t ltmask = (1u << lane) - 1u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t n_tiles = P.n_tiles;
    const uint64_t kmask = kmer_mask(P.k);
    const uint32_t nr = MULTI ? P.T.nr : 1u;  // one GPU: the owner arithmetic folds away
    auto surv_dest = [nr](const Surv& r) { return owner_of_hash(r.x, nr); };
    const uint32_t ll = P.l;
    const int bmb = P.bm_bits;
    Surv* buf = s_buf[threadIdx.x >> 5];
    uint32_t nbuf = 0;

    const uint4 zero4 = make_uint4(0, 0, 0, 0);
    auto load_rec = [&](uint32_t t) -> uint4 { return t < n_tiles ? __ldg(P.tiles + t) : zero4; };
    auto load_gran = [&](const uint4& rec) -> uint4 {  // a record of length 0 (no tile) loads nothing
        return rec.w + lane < ((rec.z + 63u) >> 6) ? ldg_stream(P.packed + rec.x + lane) : zero4;
    };
    // stage (2): hashes and bitmap words of one tile
    uint64_t h_cur[NP], h_nx[NP];
    uint32_t bw_cur[NP], bw_nx[NP];
    uint32_t a_cur = 0, q_cur = 0, nv_cur = 0, a_nx, q_nx, nv_nx;
    auto hash_tile = [&](const uint4& rec, const uint4& v, uint64_t (&h)[NP], uint32_t (&bw)[NP], uint32_t& a, uint32_t& q,
                         uint32_t& nv) {
        const uint32_t L = rec.z;
        const uint32_t u = rec.w + lane;  // granule of this lane inside the entry
        a = rec.y;
        q = u * GRAN_BASES;
        // probes q + 32 j that can start a row: p = q - delta, delta < stride, needs |a| - p >= l, i.e.
        // q <= |a| - l + stride - 1 (stride <= l - k + 1, so the k-mer lies inside the read; the exact
        // test per candidate is in k3_resolve)
        nv = 0;
        const uint32_t qmax = L - ll + ((1u << SL) - 1u);  // only used when L >= ll
        if (u < ((L + 63u) >> 6) && L >= ll && q <= qmax) nv = ((qmax - q) >> SL) + 1u;
        h[0] = mix64(((uint64_t)v.x | ((uint64_t)v.y << 32)) & kmask);
        if (NP == 2) h[NP - 1] = mix64(((uint64_t)v.z | ((uint64_t)v.w << 32)) & kmask);
#pragma unroll
        for (int j = 0; j < NP; ++j)  // lanes without a probe read word 0 (one broadcast line)
            bw[j] = __ldg(P.filter + ((uint32_t)j < nv ? bitmap_word(h[j], nr, bmb) : 0u));
    };

    uint4 rec1 = load_rec(warp + nwarps), rec2 = load_rec(warp + 2 * nwarps);
    uint4 v1 = load_gran(rec1);
    {
        const uint4 rec0 = load_rec(warp);
        const uint4 v0 = load_gran(rec0);
        hash_tile(rec0, v0, h_cur, bw_cur, a_cur, q_cur, nv_cur);
    }
    for (uint32_t tile = warp; tile < n_tiles; tile += nwarps) {
        const uint4 rec3 = load_rec(tile + 3 * nwarps);
        const uint4 v2 = load_gran(rec2);
        hash_tile(rec1, v1, h_nx, bw_nx, a_nx, q_nx, nv_nx);

        // stage (3): this tile's survivors
        uint32_t m[NP], total = 0;
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            const bool pass = (uint32_t)j < nv_cur && ((bw_cur[j] >> ((uint32_t)h_cur[j] & 31u)) & 1u);
            m[j] = __ballot_sync(FULL, pass);
            total += __popc(m[j]);
        }
        if (total) {  // warp-uniform
            if (nbuf + total > SCAP) wbuf_flush(buf, nbuf, P.surv, lane, rs, surv_dest);
            uint32_t o = nbuf;
#pragma unroll
            for (int j = 0; j < NP; ++j) {
                if ((m[j] >> lane) & 1u)
                    buf[o + __popc(m[j] & ltmask)] = make_ulonglong2(h_cur[j], ((uint64_t)a_cur << 32) | (q_cur + 32u * j));
                o += __popc(m[j]);
            }
            nbuf += total;
            __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            h_cur[j] = h_nx[j];
            bw_cur[j] = bw_nx[j];
        }
        a_cur = a_nx;
        q_cur = q_nx;
        nv_cur = nv_nx;
        rec1 = rec2;
        v1 = v2;
        rec2 = rec3;
    }
    if (nbuf) wbuf_flush(buf, nbuf, P.surv, lane, rs, surv_dest);
}

// ---- the same pass with bulk-async tile staging (stride 64 / 32; the default) -------------------
// A warp owns a CONTIGUOUS range of tiles (so its granule stream and its tile records are contiguous
// in HBM) and keeps a ring of BULK_STAGES tiles of 512 B in shared memory: one lane issues a
// `cp.async.bulk` global -> shared copy per tile (SASS UBLKCP) whose completion is counted on the
// stage's mbarrier (expect_tx), BULK_STAGES tiles ahead of the tile being hashed; the lanes read their
// granule with one LDS.128.  The granule loads cost no registers and no per-lane address arithmetic,
// and 3 KB per warp are in flight whatever the compiler does with the loop (ncu r02f, register
// pipeline: 5.3 long-scoreboard stall cycles per issued instruction, 55 % of the issue slots used).
// The tile records come 32 at a time, one coalesced load per lane; the bitmap words of tile i are
// requested while tile i-1's are tested, as in k3_filter_wide.
constexpr int BULK_STAGES = 6;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!done);
}

template <int SL>
__global__ void __launch_bounds__(256, 4) k3_filter_bulk(const ScanParams P) {
    static_assert(SL == 5 || SL == 6, "one or two probes per granule");
    constexpr int NP = SL == 6 ? 1 : 2;
    constexpr int S = BULK_STAGES;
    constexpr uint32_t SCAP = WBUF / 2;
    __shared__ __align__(128) uint4 s_ring[8][S][32];
    __shared__ __align__(8) unsigned long long s_bar[8][S];
    __shared__ Surv s_buf[8][SCAP];
    __shared__ RouteScratch s_rs[8];
    const uint32_t wib = threadIdx.x >> 5;
    RouteScratch* rs = &s_rs[wib];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t ltmask = (1u << lane) - 1u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t per = (P.n_tiles + nwarps - 1) / nwarps;
    const uint32_t t_begin = min(P.n_tiles, warp * per), t_end = min(P.n_tiles, t_begin + per);
    const uint32_t n = t_end - t_begin;  // tiles of this warp
    const uint64_t kmask = kmer_mask(P.k);
    const uint32_t nr = P.T.nr;
    auto surv_dest = [nr](const Surv& r) { return owner_of_hash(r.x, nr); };
    const uint32_t ll = P.l;
    const int bmb = P.bm_bits;
    Surv* buf = s_buf[wib];
    uint32_t nbuf = 0;
    const uint32_t ring0 = smem_u32(&s_ring[wib][0][0]), bar0 = smem_u32(&s_bar[wib][0]);

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) mbar_init(bar0 + 8u * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    // tile records, 32 at a time: lane j holds the record of tile t_begin + 32 * blk + j
    const uint4 zero4 = make_uint4(0, 0, 0, 0);
    auto load_recs = [&](uint32_t blk) -> uint4 {
        const uint32_t i = 32u * blk + lane;
        return i < n ? __ldg(P.tiles + t_begin + i) : zero4;
    };
    uint4 rec_cur = load_recs(0), rec_nxt = load_recs(1);
    uint32_t blk = 0;
    // issue the copy of this warp's tile i (i < n) into its stage
    auto issue = [&](uint32_t i) {
        const bool in_cur = (i >> 5) == blk;  // else it belongs to the next block of records (S <= 32)
        const uint32_t src_lane = i & 31u;
        const uint32_t x = __shfl_sync(FULL, in_cur ? rec_cur.x : rec_nxt.x, src_lane);
        const uint32_t z = __shfl_sync(FULL, in_cur ? rec_cur.z : rec_nxt.z, src_lane);
        const uint32_t w = __shfl_sync(FULL, in_cur ? rec_cur.w : rec_nxt.w, src_lane);
        if (lane == 0) {
            const uint32_t ng = ((z + 63u) >> 6) - w;  // granules of the entry from this tile on (>= 1)
            const uint32_t bytes = 16u * (ng < 32u ? ng : 32u);
            const uint32_t st = i % S;
            mbar_expect_tx(bar0 + 8u * st, bytes);
            bulk_g2s(ring0 + 512u * st, P.packed + x, bytes, bar0 + 8u * st);
        }
    };
    for (uint32_t i = 0; i < (uint32_t)S && i < n; ++i) issue(i);

    uint64_t h_prev[NP];
    uint32_t bw_prev[NP];
    uint32_t a_prev = 0, q_prev = 0, nv_prev = 0;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
        h_prev[j] = 0;
        bw_prev[j] = 0;
    }
    for (uint32_t i = 0; i <= n; ++i) {
        uint64_t h[NP];
        uint32_t bw[NP];
        uint32_t a = 0, q = 0, nv = 0;
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            h[j] = 0;
            bw[j] = 0;
        }
        if (i < n) {  // warp-uniform
            if ((i >> 5) != blk) {  // entering the next block of records
                ++blk;
                rec_cur = rec_nxt;
                rec_nxt = load_recs(blk + 1);
            }
            const uint32_t src_lane = i & 31u;
            a = __shfl_sync(FULL, rec_cur.y, src_lane);
            const uint32_t L = __shfl_sync(FULL, rec_cur.z, src_lane);
            const uint32_t u = __shfl_sync(FULL, rec_cur.w, src_lane) + lane;  // granule of this lane inside the entry
            const uint32_t st = i % S;
            mbar_wait(bar0 + 8u * st, (i / S) & 1u);
            uint4 v = zero4;
            const bool inr = u < ((L + 63u) >> 6);
            if (inr) v = s_ring[wib][st][lane];
            q = u * GRAN_BASES;
            // probes q + 32 j that can start a row: p = q - delta, delta < stride, needs |a| - p >= l, i.e.
            // q <= |a| - l + stride - 1 (the exact test per candidate is in k3_resolve)
            const uint32_t qmax = L - ll + ((1u << SL) - 1u);  // only used when L >= ll
            if (inr && L >= ll && q <= qmax) nv = ((qmax - q) >> SL) + 1u;
            h[0] = mix64(((uint64_t)v.x | ((uint64_t)v.y << 32)) & kmask);
            if (NP == 2) h[NP - 1] = mix64(((uint64_t)v.z | ((uint64_t)v.w << 32)) & kmask);
#pragma unroll
            for (int j = 0; j < NP; ++j)  // lanes without a probe read word 0 (one broadcast line)
                bw[j] = __ldg(P.filter + ((uint32_t)j < nv ? bitmap_word(h[j], nr, bmb) : 0u));
            __syncwarp();  // every lane has its granule: the stage can be refilled
            if (i + S < n) issue(i + S);
        }
        // the previous tile's survivors (its bitmap words were requested one iteration ago)
        uint32_t m[NP], total = 0;
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            const bool pass = (uint32_t)j < nv_prev && ((bw_prev[j] >> ((uint32_t)h_prev[j] & 31u)) & 1u);
            m[j] = __ballot_sync(FULL, pass);
            total += __popc(m[j]);
        }
        if (total) {  // warp-uniform
            if (nbuf + total > SCAP) wbuf_flush(buf, nbuf, P.surv, lane, rs, surv_dest);
            uint32_t o = nbuf;
#pragma unroll
            for (int j = 0; j < NP; ++j) {
                if ((m[j] >> lane) & 1u)
                    buf[o + __popc(m[j] & ltmask)] = make_ulonglong2(h_prev[j], ((uint64_t)a_prev << 32) | (q_prev + 32u * j));
                o += __popc(m[j]);
            }
            nbuf += total;
            __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            h_prev[j] = h[j];
            bw_prev[j] = bw[j];
        }
        a_prev = a;
        q_prev = q;
        nv_prev = nv;
    }
    if (nbuf) wbuf_flush(buf, nbuf, P.surv, lane, rs, surv_dest);
}

// Smaller strides (SL = 4, 3: 4 or 8 probes per lane at fixed offsets, loops unrolled; SL < 0: the
// stride is read from the parameters): two tiles in flight per warp, one ballot round per probe.
template <int SL>
__global__ void __launch_bounds__(256) k3_filter(const ScanParams P) {
    __shared__ Surv s_buf[8][WBUF / 2];
    __shared__ RouteScratch s_rs[8];
    constexpr uint32_t SCAP = WBUF / 2;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wib = threadIdx.x >> 5;
    RouteScratch* rs = &s_rs[wib];
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t n_tiles = P.n_tiles;
    const uint64_t kmask = kmer_mask(P.k);
    const uint64_t* packed64 = reinterpret_cast<const uint64_t*>(P.packed);
    const int sl = SL >= 0 ? SL : P.stride_log2;
    const uint32_t n_probe = GRAN_BASES >> sl;
    const uint32_t nr = P.T.nr;
    auto surv_dest = [nr](const Surv& r) { return owner_of_hash(r.x, nr); };
    Surv* buf = s_buf[wib];
    uint32_t nbuf = 0;

    // pipeline: rec_cur / v_cur = this tile; rec_nx = the next tile's record (its granule is
    // requested at the top of the iteration); the record after that is requested there too
    uint4 rec_cur = make_uint4(0, 0, 0, 0), rec_nx = rec_cur, v_cur = rec_cur;
    if (warp < n_tiles) {
        rec_cur = __ldg(P.tiles + warp);
        if (warp + nwarps < n_tiles) rec_nx = __ldg(P.tiles + warp + nwarps);
        if (rec_cur.w + lane < ((rec_cur.z + 63u) >> 6)) v_cur = ldg_stream(P.packed + rec_cur.x + lane);
    }
    for (uint32_t tile = warp; tile < n_tiles; tile += nwarps) {
        const uint4 rec = rec_cur;
        const uint4 v = v_cur;
        uint4 rec_n2 = make_uint4(0, 0, 0, 0);
        if (tile + 2 * nwarps < n_tiles) rec_n2 = __ldg(P.tiles + tile + 2 * nwarps);
        v_cur = make_uint4(0, 0, 0, 0);
        if (tile + nwarps < n_tiles && rec_nx.w + lane < ((rec_nx.z + 63u) >> 6))
            v_cur = ldg_stream(P.packed + rec_nx.x + lane);
        rec_cur = rec_nx;
        rec_nx = rec_n2;

        const uint32_t a = rec.y, L = rec.z;
        const uint32_t u = rec.w + lane;  // granule of this lane inside the entry
        const bool inr = u < ((L + 63u) >> 6);
        const uint32_t pos0 = u * GRAN_BASES;
        const uint64_t w0 = (uint64_t)v.x | ((uint64_t)v.y << 32);
        const uint64_t w1 = (uint64_t)v.z | ((uint64_t)v.w << 32);
        uint64_t w2 = 0;
        if (sl < 5) {  // k-mers that start in the second word run into the next granule
            w2 = __shfl_down_sync(FULL, w0, 1);
            if (lane == 31) w2 = inr ? packed64[2ull * (rec.x + lane + 1)] : 0;  // a spare granule follows the array
        }
        // probes q = pos0 + off that can start a row: |a| - p >= l for some p = q - delta,
        // delta < stride, i.e. q <= |a| - l + stride - 1 (the exact test is in k3_resolve)
        uint32_t nv = 0;
        const uint32_t qmax = L - P.l + ((1u << sl) - 1u);  // only used when L >= l
        if (inr && L >= P.l && pos0 <= qmax) nv = ((qmax - pos0) >> sl) + 1u;
        if (nv > n_probe) nv = n_probe;

        // ---- 4 or more probes per lane: one ballot round per probe ----
        uint32_t hitmask = 0;
        if (nv) {
#pragma unroll 8
            for (uint32_t j = 0; j < n_probe; ++j) {
                const uint32_t off = j << sl;
                const uint64_t x = (off < 32 ? kmer_at(w0, w1, off) : kmer_at(w1, w2, off - 32)) & kmask;
                const uint64_t bit = bitmap_bit(mix64(x), nr, P.bm_bits);
                hitmask |= ((__ldg(P.filter + (bit >> 5)) >> (bit & 31)) & 1u) << j;
            }
            hitmask = nv >= 32 ? hitmask : (hitmask & ((1u << nv) - 1u));
        }
        while (__any_sync(FULL, hitmask != 0)) {
            Surv r = make_ulonglong2(0, 0);
            const bool have = hitmask != 0;
            if (have) {
                const uint32_t j = __ffs(hitmask) - 1;
                hitmask &= hitmask - 1;
                const uint32_t off = j << sl;
                const uint64_t x = (off < 32 ? kmer_at(w0, w1, off) : kmer_at(w1, w2, off - 32)) & kmask;
                r.x = mix64(x);
                r.y = ((uint64_t)a << 32) | (pos0 + off);
            }
            const unsigned m = __ballot_sync(FULL, have);
            if (nbuf + 32 > SCAP) wbuf_flush(buf, nbuf, P.surv, lane, rs, surv_dest);
            if (have) buf[nbuf + __popc(m & ((1u << lane) - 1u))] = r;
            nbuf += __popc(m);
            __syncwarp();
        }
    }
    if (nbuf) wbuf_flush(buf, nbuf, P.surv, lane, rs, surv_dest);
}

// ---- sparse path, pass 2 ------------------------------------------------------------------
__global__ void __launch_bounds__(256) k3_resolve(const ScanParams P) {
    __shared__ uint64_t s_buf[8][WBUF];
    __shared__ RouteScratch s_rs[8];
    const uint32_t lane = threadIdx.x & 31;
    uint64_t* buf = s_buf[threadIdx.x >> 5];
    RouteScratch* rs = &s_rs[threadIdx.x >> 5];
    uint32_t nbuf = 0;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    for (uint32_t seg = 0; seg < P.probes.nr; ++seg) {  // the survivors of every source rank for this rank's k-mers
        const uint64_t n = seg_count(P.probes, seg);
        const Surv* __restrict__ list = P.probes.seg[seg];  // in the source's arena: streamed over NVLink
        const uint64_t first = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~31ull;
        // one record ahead: the (possibly remote) load of the next item is in flight while this one is resolved
        Surv nxt = make_ulonglong2(0, 0);
        if (first + lane < n) nxt = list[first + lane];
        for (uint64_t i0 = first; i0 < n; i0 += nthreads) {
            const uint64_t i = i0 + lane;
            const Surv rec = nxt;
            if (i + nthreads < n) nxt = list[i + nthreads];
            uint32_t ptr = NIL, owner = 0, a = 0, q = 0, la = 0;
            if (i < n) {
                a = (uint32_t)(rec.y >> 32);
                q = (uint32_t)rec.y;
                ptr = table_find(P.T, rec.x, owner);
                la = __ldg(P.ent + a).y;
            }
            walk_list(P, owner, ptr, a, la, q, buf, nbuf, lane, rs);
        }
    }
    if (nbuf) wbuf_flush(buf, nbuf, P.sink, lane, rs, [&](uint64_t k2) { return cand_dest(P, k2); });
}

// ---- dense path (stride 1) -----------------------------------------------------------------
template <bool SMEM_FILTER>
__global__ void __launch_bounds__(1024, 1) k3_scan(const ScanParams P) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    __shared__ RouteScratch s_rs[32];
    RouteScratch* rs = &s_rs[threadIdx.x >> 5];
    uint64_t* s_bufs = reinterpret_cast<uint64_t*>(s_raw);               // [32][WBUF]
    uint32_t* s_filter = reinterpret_cast<uint32_t*>(s_raw + 32 * WBUF * 8);
    const uint32_t* filt;
    if (SMEM_FILTER) {
        const uint32_t nw = 1u << P.bm_bits;
        for (uint32_t i = threadIdx.x; i < nw; i += blockDim.x) s_filter[i] = P.filter[i];
        __syncthreads();
        filt = s_filter;
    } else {
        filt = P.filter;
    }
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t n_tiles = P.n_tiles;
    const uint64_t kmask = kmer_mask(P.k);
    const uint64_t* packed64 = reinterpret_cast<const uint64_t*>(P.packed);
    const int fwb = P.bm_bits;
    uint64_t* buf = s_bufs + (threadIdx.x >> 5) * WBUF;
    uint32_t nbuf = 0;

    for (uint32_t tile = warp; tile < n_tiles; tile += nwarps) {
        const uint4 rec = __ldg(P.tiles + tile);
        const uint32_t a = rec.y, la = rec.z;
        const uint32_t g = rec.x + lane;
        const bool inr = rec.w + lane < ((la + 63u) >> 6);
        const uint32_t pos0 = (rec.w + lane) * GRAN_BASES;
        uint64_t w0 = 0, w1 = 0;
        int nvalid = 0;
        if (inr) {
            const uint4 v = ldg_stream(P.packed + g);
            w0 = (uint64_t)v.x | ((uint64_t)v.y << 32);
            w1 = (uint64_t)v.z | ((uint64_t)v.w << 32);
            // positions p in [pos0, pos0+64) with L - p >= l
            const int64_t nv = (int64_t)la - (int64_t)P.l + 1 - (int64_t)pos0;
            nvalid = nv <= 0 ? 0 : (nv >= GRAN_BASES ? GRAN_BASES : (int)nv);
        }
        if (!__any_sync(FULL, nvalid > 0)) continue;
        uint64_t w2 = __shfl_down_sync(FULL, w0, 1);
        if (lane == 31) w2 = inr ? packed64[2ull * (g + 1)] : 0;  // a spare granule follows the array

        uint64_t hits = 0;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const uint32_t h = kmer_hash(kmer_at(w0, w1, j) & kmask);
            const uint32_t fb = filter_bits(h);
            if ((filt[filter_word(h, fwb)] & fb) == fb) hits |= 1ull << j;
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const uint32_t h = kmer_hash(kmer_at(w1, w2, j) & kmask);
            const uint32_t fb = filter_bits(h);
            if ((filt[filter_word(h, fwb)] & fb) == fb) hits |= 1ull << (32 + j);
        }
        hits &= nvalid >= 64 ? ~0ull : ((1ull << nvalid) - 1ull);

        // ---- resolve filter hits against the table, one position per lane per round ----
        // stride 1: the anchor id IS b.  The list may hold a itself (always at p == 0, and at
        // p != 0 for periodic reads): walk_list drops it (identity by index,
        // overlapper.cpp:72-74,101-104)
        while (__any_sync(FULL, hits != 0)) {
            uint32_t ptr = NIL, owner = 0, p = 0;
            if (hits) {
                const int j = __ffsll((long long)hits) - 1;
                hits &= hits - 1;
                p = pos0 + j;
                const uint64_t x = (j < 32 ? kmer_at(w0, w1, j) : kmer_at(w1, w2, j - 32)) & kmask;
                ptr = table_find(P.T, mix64(x), owner);
            }
            walk_list(P, owner, ptr, a, la, p, buf, nbuf, lane, rs);
        }
    }
    if (nbuf) wbuf_flush(buf, nbuf, P.sink, lane, rs, [&](uint64_t k2) { return cand_dest(P, k2); });
}

}  // namespace ovl
