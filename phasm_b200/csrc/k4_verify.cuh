// k4_verify.cuh -- K4: warp-cooperative packed-word verifier.
//
// Replaces the suffix-tree emission tests of the reference: a candidate (a,p,b) is a row source
// iff a[p : p+n] == b[0 : n] with n = min(|a|-p, |b|).  rem = |a|-p <= |b| is the suffix-prefix
// case (R1, /root/reference/src/overlapper.cpp:64-91), rem >= |b| the containment case (R2,
// overlapper.cpp:93-115); rem == |b| is both, which is where the reference's duplicate rows
// come from.  Candidates with a == b are dropped (identity by index, overlapper.cpp:72-74).
//
// One warp per candidate.  Per iteration a lane compares two 64-bit words (64 bases): B comes
// as one aligned 128-bit load, A is re-aligned from three 64-bit words with funnel shifts.
// The warp leaves at the first 2048-base chunk holding a mismatch.
//
// Roofline: HBM/L2.  Algorithmic bytes: 2*ceil(m_c/4) per candidate (both operands, m_c bases
// up to and including the first mismatch), SURVEY.md 8(d).
#pragma once
#include "ovl_common.cuh"

namespace ovl {

struct VerifyParams {
    const Cand* cand;
    const unsigned long long* n_cand;  // device counter written by K3
    uint64_t cap;
    const uint64_t* packed64;
    const uint32_t* gran0;
    const uint32_t* len;
    KeyFmt fmt;
    uint64_t* hits;                    // 64-bit hit keys, capacity == cap
    unsigned long long* n_hits;
    unsigned long long* counters;      // [0] self candidates, [1] sum m_c, [2] sum 2*ceil(m_c/4)
    int collect;
};

__device__ __forceinline__ uint64_t low_bases_mask(int64_t nb) {  // nb low bases of a word
    return nb >= 32 ? ~0ull : (nb <= 0 ? 0ull : ((1ull << (2 * nb)) - 1ull));
}

__global__ void __launch_bounds__(256) k4_verify(const VerifyParams P) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    unsigned long long nc = *P.n_cand;
    if (nc > P.cap) nc = P.cap;
    unsigned long long self_cnt = 0, sum_m = 0, sum_b = 0;

    for (uint64_t c = warp; c < nc; c += nwarps) {
        const uint4 cv = *reinterpret_cast<const uint4*>(P.cand + c);
        const uint32_t a = cv.x, p = cv.y, b = cv.z;
        if (a == b) {
            ++self_cnt;
            continue;
        }
        const uint32_t la = __ldg(P.len + a), lb = __ldg(P.len + b);
        const uint32_t rem = la - p;
        const uint32_t n = rem < lb ? rem : lb;
        const uint64_t wa = 2ull * __ldg(P.gran0 + a) + (p >> 5);
        const uint64_t wb = 2ull * __ldg(P.gran0 + b);
        const int s = 2 * (int)(p & 31);
        const uint32_t nw = (n + 31) >> 5;
        bool ok = true;
        uint32_t first_bad = n;  // base index of the first mismatch (stats only)
        for (uint32_t i0 = 0; i0 < nw; i0 += 64) {
            const uint32_t i = i0 + 2 * lane;
            uint64_t diff0 = 0, diff1 = 0;
            if (i < nw) {
                const uint4 bv = *reinterpret_cast<const uint4*>(P.packed64 + wb + i);
                const uint64_t b0 = (uint64_t)bv.x | ((uint64_t)bv.y << 32);
                const uint64_t b1 = (uint64_t)bv.z | ((uint64_t)bv.w << 32);
                const uint64_t a0 = P.packed64[wa + i];
                const uint64_t a1 = P.packed64[wa + i + 1];
                const uint64_t a2 = P.packed64[wa + i + 2];
                const uint64_t x0 = s ? ((a0 >> s) | (a1 << (64 - s))) : a0;
                const uint64_t x1 = s ? ((a1 >> s) | (a2 << (64 - s))) : a1;
                diff0 = (x0 ^ b0) & low_bases_mask((int64_t)n - 32ll * i);
                diff1 = (x1 ^ b1) & low_bases_mask((int64_t)n - 32ll * (i + 1));
            }
            const unsigned bad = __ballot_sync(FULL, (diff0 | diff1) != 0);
            if (bad) {
                ok = false;
                if (P.collect) {
                    const int src = __ffs(bad) - 1;  // lowest lane == lowest base index
                    uint32_t fb = 0;
                    if (lane == (uint32_t)src)
                        fb = diff0 ? 32 * i + ((__ffsll((long long)diff0) - 1) >> 1)
                                   : 32 * (i + 1) + ((__ffsll((long long)diff1) - 1) >> 1);
                    first_bad = __shfl_sync(FULL, fb, src);
                }
                break;
            }
        }
        if (P.collect) {
            const uint32_t mc = ok ? n : first_bad + 1;
            sum_m += mc;
            sum_b += 2ull * ((mc + 3) / 4);
        }
        if (ok && lane == 0) {
            const uint64_t fl = (rem <= lb ? HIT_SP : 0) | (rem >= lb ? HIT_CONT : 0);
            const uint64_t key =
                (((((uint64_t)a << P.fmt.bits_e) | b) << P.fmt.bits_p | p) << 2) | fl;
            P.hits[atomicAdd(P.n_hits, 1ull)] = key;
        }
    }
    if (lane == 0) {
        if (self_cnt) atomicAdd(P.counters + 0, self_cnt);
        if (P.collect && sum_m) {
            atomicAdd(P.counters + 1, sum_m);
            atomicAdd(P.counters + 2, sum_b);
        }
    }
}

}  // namespace ovl
