// k5_emit.cuh -- K5: select-longest + deterministic compaction of verified candidates into rows.
//
// Replaces the two emission sites of the reference DFS:
//   R1  /root/reference/src/overlapper.cpp:64-91   only the TOP of A's stack (= the longest
//       suffix of A that is a prefix of B) is emitted per ordered pair (A,B);
//   R2  /root/reference/src/overlapper.cpp:93-115  every occurrence of the whole B inside A.
//
// Input: the keys [a | b | p | cont | sp] of the verified candidates, radix-sorted ascending, so
// hits of one (a,b) pair are adjacent and ordered by p.  rem = |a|-p falls as p grows, hence
// inside a pair the flags go  CONT-only ... (CONT|SP at rem == |b|) ... SP-only: the first hit
// carrying SP is the smallest p, i.e. the longest suffix-prefix overlap.  A hit with both flags
// yields the same row twice -- that is the reference's duplicate row.
//
// Because rows of one key share astart = p and keys are sorted by (a,b,p), the emitted rows are
// already in (a,b,astart,aend,bend) order: the output is deterministic and canonical.
//
// Roofline: HBM, 8 B read per hit (x2: count pass + emit pass) + 20 B written per row.
#pragma once
#include "ovl_common.cuh"

namespace ovl {

__device__ __forceinline__ uint32_t key_rows(const uint64_t* __restrict__ keys, uint64_t i,
                                             uint64_t key, int pair_shift) {
    uint32_t c = (key & HIT_CONT) ? 1u : 0u;
    if (key & HIT_SP) {
        bool first = true;
        if (i) {
            const uint64_t prev = keys[i - 1];
            first = ((prev >> pair_shift) != (key >> pair_shift)) || !(prev & HIT_SP);
        }
        c += first;
    }
    return c;
}

__global__ void __launch_bounds__(256)
k5_count(const uint64_t* __restrict__ keys, uint64_t n, KeyFmt fmt, uint32_t* __restrict__ cnt) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const uint64_t key = keys[i];
        cnt[i] = key_rows(keys, i, key, fmt.bits_p + 2);
    }
}

__global__ void __launch_bounds__(256)
k5_emit(const uint64_t* __restrict__ keys, uint64_t n, KeyFmt fmt,
        const uint64_t* __restrict__ offs, const uint32_t* __restrict__ cnt,
        const uint2* __restrict__ ent, Row* __restrict__ rows) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t c = cnt[i];
    if (!c) return;
    const uint64_t key = keys[i];
    const uint32_t p = fmt.p(key), b = fmt.b(key), a = fmt.a(key);
    const uint32_t la = __ldg(ent + a).y, lb = __ldg(ent + b).y;
    uint64_t o = offs[i];
    const uint64_t end = o + c;
    if (key & HIT_CONT) rows[o++] = Row{a, b, p, p + lb, lb};  // R2
    if (o < end) rows[o] = Row{a, b, p, la, la - p};           // R1 (longest)
}

}  // namespace ovl
