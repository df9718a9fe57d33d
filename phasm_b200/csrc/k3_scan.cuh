// k3_scan.cuh -- K3: streaming k-mer scan of the A entries against the anchor table.
//
// Replaces the suffix-tree DFS push phase (/root/reference/src/overlapper.cpp:38-62): instead
// of walking a suffix tree to learn which suffixes of A end at a node on the path to B's leaf,
// A positions look their k-mer up among the indexed B k-mers (k2_anchor.cuh).
//
// Mapping (both kernels): a warp owns a tile of 32 consecutive granules (2048 bases); each lane
// loads its granule with ONE coalesced 128-bit load (512 B per warp) and takes the following
// word from the next lane by shuffle (the k-mer carry across granules).
//
//   k3_probe  (stride >= 2; stride 64 whenever l >= 95): only the positions q = 0 mod stride are
//             probed -- 64/stride k-mers per granule, word-aligned for stride 32/64 -- against
//             an L2-resident one-bit-per-key bitmap, then the hash table.  The kernel streams
//             the packed reads at 0.25 B/base and is HBM/L2-latency bound, not issue bound.
//   k3_scan   (stride 1: l < 33, or the canonical-candidate accounting pass): every position
//             rolls its k-mer and tests a blocked bloom filter held in shared memory; filter
//             hits are resolved afterwards, warp-converged, against the table.  ~14 integer
//             instructions per position: issue bound.
//
// Candidates are 64-bit keys [a|b|p|00]; they are staged per warp in shared memory and
// flushed with one global atomic per ~224 keys.
#pragma once
#include "k2_anchor.cuh"
#include "ovl_common.cuh"

namespace ovl {

struct ScanParams {
    const uint4* packed;        // granules (G + 2, last two zero)
    const uint32_t* gran_ent;   // [G]
    const uint2* ent;           // [M] {gran0, len}
    uint32_t g_begin, g_end;    // granule range of the A entries this launch scans
    uint32_t l;                 // min overlap length (>= 1)
    int k;                      // anchor length
    int stride_log2;
    const uint32_t* filter;     // sparse: bitmap words; dense: bloom words
    int bm_bits;                // sparse: log2 bits; dense: log2 words
    const Slot* table;
    int tbits;
    const uint32_t* next;
    KeyFmt fmt;
    CandSink sink;
    MirrorRule mirror;          // on: emit only the candidates that verify for themselves
};

// ---- sparse path ------------------------------------------------------------------------
// SL >= 0: the stride is the compile-time constant 2^SL (6, 5, 4 and 3 are instantiated: 1, 2, 4 or
// 8 k-mers per granule at fixed offsets, loops fully unrolled); SL < 0: read it from the parameters.
// The warp's NEXT tile (packed granule + entry index) is loaded before the current one is
// processed, and the filter words are requested before the entry record is needed, so the
// dependent chain per tile is filter -> table slot -> anchor list, all L2 hits (ncu r01f: 60 % of
// the stall samples of the previous form waited on exactly these loads one after the other).
template <int SL>
__global__ void __launch_bounds__(256) k3_probe(const ScanParams P) {
    __shared__ uint64_t s_buf[8][WBUF];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wib = threadIdx.x >> 5;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t n_gran = P.g_end - P.g_begin;
    const uint32_t n_tiles = (n_gran + 31) / 32;
    const uint64_t kmask = kmer_mask(P.k);
    const uint64_t* packed64 = reinterpret_cast<const uint64_t*>(P.packed);
    const int sl = SL >= 0 ? SL : P.stride_log2;
    const uint32_t stride = 1u << sl;
    const uint32_t n_probe = GRAN_BASES >> sl;
    uint64_t* buf = s_buf[wib];
    uint32_t nbuf = 0;

    uint4 v_nx = make_uint4(0, 0, 0, 0);
    uint32_t a_nx = 0;
    if (warp < n_tiles) {
        const uint32_t g = P.g_begin + warp * 32 + lane;
        if (g < P.g_end) {
            v_nx = ldg_stream(P.packed + g);
            a_nx = __ldg(P.gran_ent + g);
        }
    }
    for (uint32_t tile = warp; tile < n_tiles; tile += nwarps) {
        const uint32_t g = P.g_begin + tile * 32 + lane;
        const bool inr = g < P.g_end;
        const uint4 v = v_nx;
        const uint32_t a = a_nx;
        {  // prefetch the next tile of this warp
            const uint32_t gn = g + nwarps * 32;
            if (tile + nwarps < n_tiles && gn < P.g_end) {
                v_nx = ldg_stream(P.packed + gn);
                a_nx = __ldg(P.gran_ent + gn);
            }
        }
        const uint64_t w0 = (uint64_t)v.x | ((uint64_t)v.y << 32);
        const uint64_t w1 = (uint64_t)v.z | ((uint64_t)v.w << 32);
        uint2 m = make_uint2(0, 0);
        if (inr) m = __ldg(P.ent + a);
        uint64_t w2 = 0;
        if (sl < 5) {  // k-mers that start in the second word run into the next granule
            w2 = __shfl_down_sync(FULL, w0, 1);
            if (lane == 31) w2 = inr ? packed64[2ull * (g + 1)] : 0;
        }
        // phase 1: filter bit of every probed position of this granule (independent loads, issued
        // before the entry record is looked at)
        uint32_t fbits = 0;
        if (inr) {
#pragma unroll 8
            for (uint32_t j = 0; j < n_probe; ++j) {
                const uint32_t off = j << sl;
                const uint64_t x = (off < 32 ? kmer_at(w0, w1, off) : kmer_at(w1, w2, off - 32)) & kmask;
                const uint32_t bit = bitmap_bit(mix64(x), P.bm_bits);
                fbits |= ((__ldg(P.filter + (bit >> 5)) >> (bit & 31)) & 1u) << j;
            }
        }
        const uint32_t L = m.y;
        const uint32_t pos0 = (g - m.x) * GRAN_BASES;
        uint32_t hitmask = 0;
        if (inr && L >= P.l && pos0 + (uint32_t)P.k <= L) {
            const uint32_t nv = ((L - (uint32_t)P.k - pos0) >> sl) + 1u;  // probes with pos0 + off + k <= L
            hitmask = nv >= 32 ? fbits : (fbits & ((1u << nv) - 1u));
        }
        // phase 2: resolve the (few) filter hits against the table, one position per lane per round
        while (__any_sync(FULL, hitmask != 0)) {
            uint32_t node = NIL, q = 0;
            if (hitmask) {
                const uint32_t j = __ffs(hitmask) - 1;
                hitmask &= hitmask - 1;
                const uint32_t off = j << sl;
                q = pos0 + off;
                const uint64_t x = (off < 32 ? kmer_at(w0, w1, off) : kmer_at(w1, w2, off - 32)) & kmask;
                node = table_find(P.table, P.tbits, mix64(x));
            }
            while (__any_sync(FULL, node != NIL)) {
                bool have = false;
                uint64_t key = 0;
                if (node != NIL) {
                    const uint32_t b = node >> sl;
                    const uint32_t delta = node & (stride - 1u);
                    const uint32_t nxt = __ldg(P.next + node);
                    // candidate p = q - delta: needs p >= 0, |a| - p >= l, and b != a
                    // (identity by index: overlapper.cpp:72-74,101-104)
                    if (q >= delta && b != a && L - (q - delta) >= P.l) {
                        have = !P.mirror.on || mirror_items(P.mirror, a, L, b, __ldg(P.ent + b).y, q - delta) != 0;
                        key = P.fmt.make(a, b, q - delta);
                    }
                    node = nxt;
                }
                wbuf_push(buf, nbuf, P.sink, lane, have, key);
            }
        }
    }
    if (nbuf) wbuf_flush(buf, nbuf, P.sink, lane);
}

// ---- dense path (stride 1) -----------------------------------------------------------------
template <bool SMEM_FILTER>
__global__ void __launch_bounds__(1024, 1) k3_scan(const ScanParams P) {
    extern __shared__ __align__(16) unsigned char s_raw[];
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
    const uint32_t n_gran = P.g_end - P.g_begin;
    const uint32_t n_tiles = (n_gran + 31) / 32;
    const uint64_t kmask = kmer_mask(P.k);
    const uint64_t* packed64 = reinterpret_cast<const uint64_t*>(P.packed);
    const int fwb = P.bm_bits;
    uint64_t* buf = s_bufs + (threadIdx.x >> 5) * WBUF;
    uint32_t nbuf = 0;

    for (uint32_t tile = warp; tile < n_tiles; tile += nwarps) {
        const uint32_t g = P.g_begin + tile * 32 + lane;
        const bool inr = g < P.g_end;
        uint64_t w0 = 0, w1 = 0;
        uint32_t a = 0, pos0 = 0;
        int nvalid = 0;
        if (inr) {
            const uint4 v = ldg_stream(P.packed + g);
            w0 = (uint64_t)v.x | ((uint64_t)v.y << 32);
            w1 = (uint64_t)v.z | ((uint64_t)v.w << 32);
            a = __ldg(P.gran_ent + g);
            const uint2 m = __ldg(P.ent + a);
            pos0 = (g - m.x) * GRAN_BASES;
            // positions p in [pos0, pos0+64) with L - p >= l
            const int64_t nv = (int64_t)m.y - (int64_t)P.l + 1 - (int64_t)pos0;
            nvalid = nv <= 0 ? 0 : (nv >= GRAN_BASES ? GRAN_BASES : (int)nv);
        }
        if (!__any_sync(FULL, nvalid > 0)) continue;
        uint64_t w2 = __shfl_down_sync(FULL, w0, 1);
        if (lane == 31) w2 = inr ? packed64[2ull * (g + 1)] : 0;  // granule G exists (zero pad)

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
        while (__any_sync(FULL, hits != 0)) {
            uint32_t node = NIL, p = 0;
            if (hits) {
                const int j = __ffsll((long long)hits) - 1;
                hits &= hits - 1;
                p = pos0 + j;
                const uint64_t x = (j < 32 ? kmer_at(w0, w1, j) : kmer_at(w1, w2, j - 32)) & kmask;
                node = table_find(P.table, P.tbits, mix64(x));
            }
            while (__any_sync(FULL, node != NIL)) {
                bool have = false;
                uint64_t key = 0;
                if (node != NIL) {
                    // stride 1: the anchor id IS b.  The list may hold a itself (always at
                    // p == 0, and at p != 0 for periodic reads): identity by index,
                    // overlapper.cpp:72-74,101-104
                    if (node != a) {
                        have = !P.mirror.on ||
                               mirror_items(P.mirror, a, __ldg(P.ent + a).y, node, __ldg(P.ent + node).y, p) != 0;
                        key = P.fmt.make(a, node, p);
                    }
                    node = __ldg(P.next + node);
                }
                wbuf_push(buf, nbuf, P.sink, lane, have, key);
            }
        }
    }
    if (nbuf) wbuf_flush(buf, nbuf, P.sink, lane);
}

}  // namespace ovl
