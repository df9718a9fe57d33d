// k3_scan.cuh -- K3: streaming k-mer scan of every A position against the anchor table.
//
// Replaces the suffix-tree DFS push phase (/root/reference/src/overlapper.cpp:38-62): instead
// of walking a suffix tree to learn which suffixes of A end at a node on the path to B's leaf,
// every position p of every entry A with |A|-p >= l looks its k-mer up among the B prefixes.
//
// Mapping: a warp owns a tile of 32 consecutive granules (2048 bases); each lane loads its
// granule with ONE coalesced 128-bit load (512 B per warp), takes the following word from the
// next lane by shuffle (the hash carry across granules), and tests its 64 positions against a
// blocked bloom filter (shared memory when it fits, else L2-resident global).  Filter hits are
// collected in a 64-bit mask and resolved afterwards, warp-converged, against the hash table;
// candidates are appended with one warp-aggregated atomic per round.
//
// Roofline: nominally HBM (0.25 B per scanned base + 4 B per granule for gran_ent), in practice
// integer-issue bound (~14 instructions per position); see DESIGN.md.
#pragma once
#include "ovl_common.cuh"

namespace ovl {

struct ScanParams {
    const uint4* packed;        // granules (G + 1, last one zero)
    const uint32_t* gran_ent;   // [G]
    const uint32_t* gran0;      // [M+1]
    const uint32_t* len;        // [M]
    uint32_t g_begin, g_end;    // granule range of the A entries this launch scans
    uint32_t l;                 // min overlap length (>= 1)
    int k;                      // anchor length
    const uint32_t* filter;     // 2^fw_bits words
    int fw_bits;
    const Slot* table;
    int tbits;
    const uint32_t* anc_ent;    // key-sorted B list
    Cand* cand;
    unsigned long long* n_cand; // total candidates found (may exceed cap)
    uint64_t cap;
};

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

template <bool SMEM_FILTER>
__global__ void __launch_bounds__(1024, 1) k3_scan(const ScanParams P) {
    extern __shared__ uint32_t s_filter[];
    const uint32_t* filt;
    if (SMEM_FILTER) {
        const uint32_t nw = 1u << P.fw_bits;
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
    const int fwb = P.fw_bits;

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
            const uint32_t L = __ldg(P.len + a);
            pos0 = (g - __ldg(P.gran0 + a)) * GRAN_BASES;
            // positions p in [pos0, pos0+64) with L - p >= l
            const int64_t nv = (int64_t)L - (int64_t)P.l + 1 - (int64_t)pos0;
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
            uint32_t start = 0, n_emit = 0, p = 0;
            if (hits) {
                const int j = __ffsll((long long)hits) - 1;
                hits &= hits - 1;
                p = pos0 + j;
                const uint64_t x = (j < 32 ? kmer_at(w0, w1, j) : kmer_at(w1, w2, j - 32)) & kmask;
                const uint32_t tmask = (1u << P.tbits) - 1u;
                uint32_t s = slot_hash(kmer_hash(x)) >> (32 - P.tbits);
                for (;;) {
                    const Slot sl = P.table[s];
                    if (sl.count == 0) break;
                    if (sl.key == x) {
                        start = sl.start;
                        // the list may hold a itself (always at p == 0, and at p != 0 for
                        // periodic reads); those records are written too and dropped by the
                        // verifier (identity by index: overlapper.cpp:72-74,101-104)
                        n_emit = sl.count;
                        break;
                    }
                    s = (s + 1) & tmask;
                }
            }
            // warp-aggregated slot allocation
            uint32_t incl = n_emit;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, incl, d);
                if (lane >= d) incl += t;
            }
            const uint32_t total = __shfl_sync(FULL, incl, 31);
            if (total == 0) continue;
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(P.n_cand, (unsigned long long)total);
            base = __shfl_sync(FULL, base, 0);
            if (n_emit) {
                uint64_t o = base + (incl - n_emit);
                for (uint32_t i = 0; i < n_emit; ++i, ++o)
                    if (o < P.cap) P.cand[o] = Cand{a, p, __ldg(P.anc_ent + start + i), 0u};
            }
        }
    }
}

}  // namespace ovl
