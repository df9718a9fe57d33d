// k8_filter.cuh -- optional post-pass: the alignment filters of the reference's `layout` stage
// over the emitted rows.
//
// Reference: phasm/cli/assembler.py:76-98 builds  [ContainedReads(), MinReadLength?, MinOverlapLength?,
// MaxOverhang(abs, rel)]  and keeps a row iff  all(f(row) for f in filters)  (short-circuit, in that
// order), rows streaming in GFA2 E-line order.  phasm/filter.py:36-122, phasm/alignments.py:237-258.
//
// Every row of this library has brange = (0, bend), so with la = |a|, lb = |b|:
//   classify:  A_CONTAINED  iff astart <= 0 and la-aend <= lb-bend
//              B_CONTAINED  iff (else) la-aend >= lb-bend          (astart >= 0 always holds)
//              OVERLAP_AB   otherwise                              (OVERLAP_BA cannot occur)
//   overlap_length = max(aend-astart, bend),   overhang = min(astart, 0) + min(la-aend, lb-bend)
//
// ContainedReads is the only stateful filter that matters (MinReadLength's set only ever holds
// reads its own length test rejects anyway): row i marks an entry when it is of a contained type,
// and a non-contained row i is dropped iff a or b was marked by a row j < i.  That streaming rule
// is order dependent but not sequential: firstmark[e] = min { i : row i marks e } by atomicMin,
// then row i is dropped iff firstmark[a] < i or firstmark[b] < i.
//
// Roofline: HBM, 20 B per row read twice + 20 B per kept row written.
#pragma once
#include "ovl_common.cuh"

namespace ovl {

struct FilterParams {
    const Row* rows;
    uint64_t n;
    const uint2* ent;              // {gran0, len}
    uint32_t* firstmark;           // [M], 0xffffffff = never marked
    uint32_t* keep;                // [n] 1 = row survives
    uint32_t contained_reads, min_read_length, min_overlap_length, max_overhang, collapse_pairs;
    long long max_overhang_abs;
    double max_overhang_rel;
    unsigned long long* counters;  // [0] rows of a contained type
};

enum RowClass : uint32_t { ROW_A_CONTAINED = 0, ROW_B_CONTAINED = 1, ROW_OVERLAP_AB = 2 };

__device__ __forceinline__ uint32_t row_class(const Row& r, uint32_t la, uint32_t lb) {
    const long long ta = (long long)la - (long long)r.aend, tb = (long long)lb - (long long)r.bend;
    if (r.astart == 0 && ta <= tb) return ROW_A_CONTAINED;
    if (ta >= tb) return ROW_B_CONTAINED;
    return ROW_OVERLAP_AB;
}

__global__ void __launch_bounds__(256) k8_mark(const FilterParams P) {
    unsigned long long nc = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.n; i += (uint64_t)gridDim.x * blockDim.x) {
        const Row r = P.rows[i];
        const uint32_t cls = row_class(r, __ldg(P.ent + r.a).y, __ldg(P.ent + r.b).y);
        if (cls == ROW_A_CONTAINED) atomicMin(P.firstmark + r.a, (uint32_t)i);
        if (cls == ROW_B_CONTAINED) atomicMin(P.firstmark + r.b, (uint32_t)i);
        nc += cls != ROW_OVERLAP_AB;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) nc += __shfl_xor_sync(FULL, nc, d);
    if ((threadIdx.x & 31) == 0 && nc) atomicAdd(P.counters, nc);
}

__global__ void __launch_bounds__(256) k8_keep(const FilterParams P) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.n; i += (uint64_t)gridDim.x * blockDim.x) {
        const Row r = P.rows[i];
        const uint32_t la = __ldg(P.ent + r.a).y, lb = __ldg(P.ent + r.b).y;
        bool keep = true;
        if (P.contained_reads) {
            keep = row_class(r, la, lb) == ROW_OVERLAP_AB && __ldg(P.firstmark + r.a) >= (uint32_t)i &&
                   __ldg(P.firstmark + r.b) >= (uint32_t)i;
        }
        if (keep && P.min_read_length) keep = la >= P.min_read_length && lb >= P.min_read_length;
        const uint32_t ovl_len = max(r.aend - r.astart, r.bend);
        if (keep && P.min_overlap_length) keep = ovl_len >= P.min_overlap_length;
        if (keep && P.max_overhang) {
            const long long ta = (long long)la - (long long)r.aend, tb = (long long)lb - (long long)r.bend;
            const long long overhang = (ta < tb ? ta : tb);  // + min(astart, bstart) = 0
            const double rel = P.max_overhang_rel * (double)ovl_len;
            const double thr = (double)P.max_overhang_abs < rel ? (double)P.max_overhang_abs : rel;  // min(abs, rel*len)
            keep = (double)overhang <= thr;
        }
        P.keep[i] = keep ? 1 : 0;
    }
}

// of the kept rows of one ordered pair (adjacent in canonical order) only the last one stays
__global__ void __launch_bounds__(256) k8_collapse(const FilterParams P, uint32_t* __restrict__ keep2) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t k = P.keep[i];
        if (k) {
            const Row r = P.rows[i];
            for (uint64_t j = i + 1; j < P.n; ++j) {
                const Row q = P.rows[j];
                if (q.a != r.a || q.b != r.b) break;
                if (P.keep[j]) {
                    k = 0;
                    break;
                }
            }
        }
        keep2[i] = k;
    }
}

__global__ void __launch_bounds__(256) k8_compact(const Row* __restrict__ rows, uint64_t n, const uint32_t* __restrict__ keep,
                                                  const uint64_t* __restrict__ off, Row* __restrict__ out) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        if (keep[i]) out[off[i]] = rows[i];
}

__global__ void __launch_bounds__(256) k8_count_marked(const uint32_t* __restrict__ firstmark, uint32_t M,
                                                       unsigned long long* __restrict__ out, uint32_t unmarked) {
    unsigned long long c = 0;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < M; e += gridDim.x * blockDim.x) c += firstmark[e] != unmarked;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(FULL, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

}  // namespace ovl
