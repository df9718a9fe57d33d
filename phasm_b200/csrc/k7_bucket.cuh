// k7_bucket.cuh -- ordering of the verified hit keys and row emission without a host round trip.
//
// The hit keys [a | b | p | flags] have to be ordered by (a, b, p, flags) before the
// select-longest rule (key_rows_prev below) can run (R1/R2, /root/reference/src/overlapper.cpp:64-115).
// A device-wide radix sort does that in 6 passes of a library kernel and needs the hit count on
// the host first.  The keys are far from random, though: an entry `a` owns a few dozen hits, so
//
//   k7_hist      count the hits of every A entry                      (one L2 atomic per hit)
//   (exclusive sum over the M+1 counts: CUB DeviceScan, size known on the host)
//   k7_scatter   move every key into its entry's bucket              (one L2 atomic per hit)
//   k7_sort      one warp per entry: bitonic sort of the bucket in shared memory (<= BUCKET_MAX
//                keys), sorted keys written back, rows of the entry counted (key_rows rule)
//   (exclusive sum over the M+1 row counts)
//   k7_emit      one warp per entry: rows written at the entry's row offset
//
// leaves the keys globally sorted and the rows in canonical (a,b,astart,aend,bend) order --
// with every size read from device counters: the whole overlap pass then needs ONE host
// synchronisation, at its end.  An entry with more than BUCKET_MAX hits (a high-copy repeat, an
// amplicon pile-up) goes on a short list that k7_sort_big orders afterwards, one thread block per
// bucket, with an in-place all-ascending bitonic network of arbitrary length: no fallback path,
// no library sort, no extra synchronisation.
//
// Roofline: HBM/L2, 8 B read + 8 B written per hit in scatter and sort, 20 B written per row.
#pragma once
#include "ovl_common.cuh"

namespace ovl {

constexpr uint32_t BUCKET_MAX = 256;  // keys one warp sorts in shared memory

struct BucketParams {
    SegList<uint64_t> hits;           // unsorted hit keys: the hit list of this rank (sizes are device counters)
    KeyFmt fmt;
    uint32_t a0;                      // first entry of this rank's A range: every hit key routed here has a in [a0, a0 + M)
    uint32_t M;                       // entries of the A range (the arrays below are indexed by a - a0: K7 does not grow with the ranks)
    uint32_t* cnt;                    // [M+1] hits per entry (zeroed by the host), cnt[M] stays 0
    const uint64_t* off;              // [M+1] exclusive sum of cnt
    uint32_t* cur;                    // [M] scatter cursors (zeroed by the host)
    uint64_t* sorted;                 // bucketed, then sorted keys
    uint32_t* rowcnt;                 // [M+1] rows per entry, rowcnt[M] = 0
    const uint64_t* rowoff;           // [M+1] exclusive sum of rowcnt
    unsigned long long* big;          // number of buckets with more than BUCKET_MAX keys
    uint32_t* big_list;               // [M] their entries
    const uint2* ent;
    Row* rows;
    uint64_t rows_cap;
};

// f(key, valid) for every key of the hit list (segment by segment, grid-stride inside); all 32 lanes
// of a warp call f together (valid = false past the end), so f may use warp collectives
template <class F>
__device__ __forceinline__ void for_each_hit(const BucketParams& P, F f) {
    const uint32_t lane = threadIdx.x & 31;
    for (uint32_t seg = 0; seg < P.hits.nr; ++seg) {
        const uint64_t n = seg_count(P.hits, seg);
        const uint64_t* __restrict__ list = P.hits.seg[seg];  // in the source's arena: streamed over NVLink
        for (uint64_t i0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~31ull; i0 < n; i0 += (uint64_t)gridDim.x * blockDim.x) {
            const uint64_t i = i0 + lane;
            f(i < n ? list[i] : 0ull, i < n);
        }
    }
}

// The hit list is far from random: a verifier warp appends runs of keys of ONE A entry (plus their
// strand-mirror twins), so neighbouring lanes mostly want the same counter.  One atomic per distinct
// entry per warp (match_any) instead of one per key: ncu r02f showed k7_hist / k7_scatter at 130-180
// long-scoreboard stall cycles per issue, i.e. same-address L2 atomics serialising.
__global__ void __launch_bounds__(256) k7_hist(const BucketParams P) {
    const int sa = 2 + P.fmt.bits_p + P.fmt.bits_e;
    const uint32_t lane = threadIdx.x & 31;
    for_each_hit(P, [&](uint64_t key, bool valid) {
        uint32_t a = valid ? (uint32_t)(key >> sa) - P.a0 : 0xffffffffu;
        if (a >= P.M) a = 0xffffffffu;  // cannot happen (keys are routed by owner); never index out of range
        const unsigned grp = __match_any_sync(FULL, a);
        if (a != 0xffffffffu && lane == (uint32_t)(__ffs(grp) - 1)) atomicAdd(P.cnt + a, (uint32_t)__popc(grp));
    });
}

__global__ void __launch_bounds__(256) k7_scatter(const BucketParams P) {
    const int sa = 2 + P.fmt.bits_p + P.fmt.bits_e;
    const uint32_t lane = threadIdx.x & 31;
    for_each_hit(P, [&](uint64_t key, bool valid) {
        uint32_t a = valid ? (uint32_t)(key >> sa) - P.a0 : 0xffffffffu;
        if (a >= P.M) a = 0xffffffffu;
        const bool live = a != 0xffffffffu;
        const unsigned grp = __match_any_sync(FULL, a);
        const int leader = __ffs(grp) - 1;
        uint32_t base = 0;
        if (live && lane == (uint32_t)leader) base = atomicAdd(P.cur + a, (uint32_t)__popc(grp));
        base = __shfl_sync(FULL, base, leader);
        if (live) P.sorted[P.off[a] + base + __popc(grp & ((1u << lane) - 1u))] = key;
    });
}

// rows of key i of a sorted run whose predecessor is `prev` (have_prev = false at a run start)
__device__ __forceinline__ uint32_t key_rows_prev(uint64_t key, uint64_t prev, bool have_prev, int pair_shift) {
    uint32_t c = (key & HIT_CONT) ? 1u : 0u;
    if (key & HIT_SP) c += !have_prev || ((prev >> pair_shift) != (key >> pair_shift)) || !(prev & HIT_SP);
    return c;
}

__global__ void __launch_bounds__(256) k7_sort(const BucketParams P) {
    __shared__ uint64_t s_keys[8][BUCKET_MAX];
    uint64_t* s = s_keys[threadIdx.x >> 5];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const int pair_shift = P.fmt.bits_p + 2;
    for (uint32_t a = warp; a < P.M; a += nwarps) {
        const uint64_t lo = P.off[a], hi = P.off[a + 1];
        const uint32_t n = (uint32_t)(hi - lo);
        if (n == 0) {
            if (lane == 0) P.rowcnt[a] = 0;
            continue;
        }
        if (hi - lo > BUCKET_MAX) {  // left to k7_sort_big (which also counts its rows)
            if (lane == 0) P.big_list[atomicAdd(P.big, 1ull)] = a;
            continue;
        }
        uint32_t np2 = 32;
        while (np2 < n) np2 <<= 1;
        for (uint32_t i = lane; i < np2; i += 32) s[i] = i < n ? P.sorted[lo + i] : ~0ull;
        __syncwarp();
        if (n > 1) {
            for (uint32_t k = 2; k <= np2; k <<= 1) {
                for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                    for (uint32_t t = lane; t < (np2 >> 1); t += 32) {
                        const uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                        const uint32_t x = i | j;
                        const uint64_t u = s[i], v = s[x];
                        if ((u > v) == ((i & k) == 0)) {
                            s[i] = v;
                            s[x] = u;
                        }
                    }
                    __syncwarp();
                }
            }
        }
        uint32_t c = 0;
        for (uint32_t i = lane; i < n; i += 32) {
            const uint64_t key = s[i];
            P.sorted[lo + i] = key;
            c += key_rows_prev(key, i ? s[i - 1] : 0ull, i != 0, pair_shift);
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(FULL, c, d);
        if (lane == 0) P.rowcnt[a] = c;
        __syncwarp();
    }
}

// Buckets above BUCKET_MAX keys: one block per bucket, in place in global memory (the bucket is a
// few thousand keys: L1/L2 resident).  All comparators of this bitonic network are ascending
// (first step of a merge pairs i with i ^ (k-1), the others i with i ^ j), so a length that is
// not a power of two needs no padding: a comparator whose partner lies past the end is skipped,
// as if +infinity sat there.
__global__ void __launch_bounds__(1024) k7_sort_big(const BucketParams P) {
    __shared__ uint32_t s_cnt[32];
    const unsigned long long nbig = *P.big;
    const int pair_shift = P.fmt.bits_p + 2;
    for (unsigned long long bi = blockIdx.x; bi < nbig; bi += gridDim.x) {
        const uint32_t a = P.big_list[bi];
        const uint64_t lo = P.off[a];
        const uint64_t n = P.off[a + 1] - lo;
        uint64_t* s = P.sorted + lo;
        uint64_t np2 = 1;
        while (np2 < n) np2 <<= 1;
        for (uint64_t k = 2; k <= np2; k <<= 1) {
            for (uint64_t j = k >> 1; j > 0; j >>= 1) {
                const bool flip = j == (k >> 1);
                for (uint64_t t = threadIdx.x; t < (np2 >> 1); t += blockDim.x) {
                    const uint64_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // bit j clear
                    const uint64_t x = flip ? (i ^ (k - 1)) : (i | j);
                    if (x < n) {  // i < x always
                        const uint64_t u = s[i], v = s[x];
                        if (u > v) {
                            s[i] = v;
                            s[x] = u;
                        }
                    }
                }
                __syncthreads();
            }
        }
        uint32_t c = 0;
        for (uint64_t i = threadIdx.x; i < n; i += blockDim.x)
            c += key_rows_prev(s[i], i ? s[i - 1] : 0ull, i != 0, pair_shift);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(FULL, c, d);
        if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = c;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t tot = 0;
            for (uint32_t w = 0; w < (blockDim.x >> 5); ++w) tot += s_cnt[w];
            P.rowcnt[a] = tot;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k7_emit(const BucketParams P) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const int pair_shift = P.fmt.bits_p + 2;
    for (uint32_t a = warp; a < P.M; a += nwarps) {
        const uint64_t lo = P.off[a], hi = P.off[a + 1];
        if (hi == lo) continue;
        const uint64_t n = hi - lo;
        const uint32_t ea = P.a0 + a;  // the entry
        const uint32_t la = __ldg(P.ent + ea).y;
        uint64_t o = P.rowoff[a];
        for (uint64_t i0 = 0; i0 < n; i0 += 32) {
            const uint64_t i = i0 + lane;
            uint64_t key = 0;
            uint32_t c = 0;
            if (i < n) {
                key = P.sorted[lo + i];
                c = key_rows_prev(key, i ? P.sorted[lo + i - 1] : 0ull, i != 0, pair_shift);
            }
            uint32_t inc = c;  // inclusive warp scan
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t v = __shfl_up_sync(FULL, inc, d);
                if (lane >= (uint32_t)d) inc += v;
            }
            if (c) {
                uint64_t w = o + inc - c;
                const uint32_t p = P.fmt.p(key), b = P.fmt.b(key);
                if ((key & HIT_CONT) && w < P.rows_cap) {
                    const uint32_t lb = __ldg(P.ent + b).y;
                    P.rows[w] = Row{ea, b, p, p + lb, lb};  // R2
                    ++w;
                }
                if (w < o + inc && w < P.rows_cap) P.rows[w] = Row{ea, b, p, la, la - p};  // R1 (longest)
            }
            o += __shfl_sync(FULL, inc, 31);
        }
    }
}

}  // namespace ovl
