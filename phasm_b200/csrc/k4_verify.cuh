// k4_verify.cuh -- K4: warp-cooperative packed-word verifier.
//
// Replaces the suffix-tree emission tests of the reference: a candidate (a,p,b) is a row source
// iff a[p : p+n] == b[0 : n] with n = min(|a|-p, |b|).  rem = |a|-p <= |b| is the suffix-prefix
// case (R1, /root/reference/src/overlapper.cpp:64-91), rem >= |b| the containment case (R2,
// overlapper.cpp:93-115); rem == |b| is both, which is where the reference's duplicate rows
// come from.
//
// Input: candidate keys radix-sorted by (a, b, p).  A warp takes 32 consecutive keys: every
// lane decodes one key and fetches the two entry records in parallel (one dependent-load chain
// per 32 candidates instead of one per candidate), then the warp walks the 32 candidates; per
// step a lane compares one 64-base granule: B comes as one aligned streaming 128-bit load,
// A as two aligned 128-bit loads (granules i, i+1) funnel-shifted by (p mod 64) bases.
// Because the keys are sorted by a, the 8 warps of a block work on the same few A entries and
// the A operand is served by L1; B streams from L2/HBM.  The first step covers 2048 bases,
// later steps 4096 (two independent load groups in flight).  Verified candidates get their
// flag bits set in place.
//
// Roofline: HBM/L2.  Algorithmic bytes: 2*ceil(m_c/4) per candidate (both operands, m_c bases
// up to and including the first mismatch), SURVEY.md 8(d).
#pragma once
#include "ovl_common.cuh"

namespace ovl {

struct VerifyParams {
    uint64_t* keys;                    // sorted candidate keys, flags set in place
    uint64_t n;
    const uint4* packed;
    const uint2* ent;                  // {gran0, len}
    KeyFmt fmt;
    unsigned long long* n_hits;
    unsigned long long* counters;      // [0] sum m_c, [1] sum 2*ceil(m_c/4)   (collect only)
    int collect;
};

// (hi:lo) >> sh for sh in [0, 128), sh even; lo/hi are 128-bit little-endian granules.
__device__ __forceinline__ uint4 shr256(const uint4 lo, const uint4 hi, uint32_t sh) {
    uint32_t w0 = lo.x, w1 = lo.y, w2 = lo.z, w3 = lo.w, w4 = hi.x, w5 = hi.y, w6 = hi.z, w7 = hi.w;
    if (sh & 64) { w0 = w2; w1 = w3; w2 = w4; w3 = w5; w4 = w6; w5 = w7; }
    if (sh & 32) { w0 = w1; w1 = w2; w2 = w3; w3 = w4; w4 = w5; }
    const uint32_t t = sh & 31;
    return make_uint4(__funnelshift_r(w0, w1, t), __funnelshift_r(w1, w2, t),
                      __funnelshift_r(w2, w3, t), __funnelshift_r(w3, w4, t));
}

// mask of the low nb bases (0 < nb < 64) of a granule
__device__ __forceinline__ uint4 low_bases_mask128(int nb) {
    auto m = [](int n) -> uint32_t { return n >= 16 ? 0xffffffffu : (n <= 0 ? 0u : ((1u << (2 * n)) - 1u)); };
    return make_uint4(m(nb), m(nb - 16), m(nb - 32), m(nb - 48));
}

// difference of granule i of the comparison (a[p+64i ..] vs b[64i ..]), masked to n bases
__device__ __forceinline__ uint4 gran_diff(const uint4* __restrict__ pa, const uint4* __restrict__ pb,
                                           uint32_t i, uint32_t sh, uint32_t n) {
    const uint4 bv = ldg_stream(pb + i);
    const uint4 a0 = __ldg(pa + i);
    const uint4 a1 = __ldg(pa + i + 1);
    const uint4 x = shr256(a0, a1, sh);
    uint4 d = make_uint4(x.x ^ bv.x, x.y ^ bv.y, x.z ^ bv.z, x.w ^ bv.w);
    const int nb = (int)(n - 64u * i);
    if (nb < 64) {
        const uint4 m = low_bases_mask128(nb);
        d.x &= m.x; d.y &= m.y; d.z &= m.z; d.w &= m.w;
    }
    return d;
}

__device__ __forceinline__ uint32_t first_diff_base(const uint4 d) {  // d != 0
    if (d.x) return (__ffs(d.x) - 1) >> 1;
    if (d.y) return 16 + ((__ffs(d.y) - 1) >> 1);
    if (d.z) return 32 + ((__ffs(d.z) - 1) >> 1);
    return 48 + ((__ffs(d.w) - 1) >> 1);
}

__global__ void __launch_bounds__(256) k4_verify(const VerifyParams P) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    const uint64_t n_tiles = (P.n + 31) / 32;
    unsigned long long hit_cnt = 0, sum_m = 0, sum_b = 0;

    for (uint64_t tile = warp; tile < n_tiles; tile += nwarps) {
        const uint64_t c = tile * 32 + lane;
        uint64_t key = 0;
        uint32_t my_ga = 0, my_gb = 0, my_sh = 0, my_n = 0, my_fl = 0;
        if (c < P.n) {
            key = P.keys[c];
            const uint32_t a = P.fmt.a(key), b = P.fmt.b(key), p = P.fmt.p(key);
            const uint2 ma = __ldg(P.ent + a), mb = __ldg(P.ent + b);
            const uint32_t rem = ma.y - p;
            my_n = rem < mb.y ? rem : mb.y;
            my_fl = (rem <= mb.y ? (uint32_t)HIT_SP : 0u) | (rem >= mb.y ? (uint32_t)HIT_CONT : 0u);
            my_ga = ma.x + (p >> 6);
            my_gb = mb.x;
            my_sh = 2u * (p & 63u);
        }
        const uint32_t cnt = (uint32_t)((P.n - tile * 32) < 32 ? (P.n - tile * 32) : 32);
        uint32_t my_ok = 0;
        for (uint32_t j = 0; j < cnt; ++j) {
            const uint32_t ga = __shfl_sync(FULL, my_ga, j);
            const uint32_t gb = __shfl_sync(FULL, my_gb, j);
            const uint32_t sh = __shfl_sync(FULL, my_sh, j);
            const uint32_t n = __shfl_sync(FULL, my_n, j);
            const uint4* pa = P.packed + ga;
            const uint4* pb = P.packed + gb;
            const uint32_t ng = (n + 63) >> 6;
            bool ok = true;
            uint32_t first_bad = n;
            // first step: 32 granules; later steps: 64 granules (two load groups in flight)
            uint32_t i0 = 0;
            while (i0 < ng) {
                uint4 d0 = make_uint4(0, 0, 0, 0), d1 = d0;
                const uint32_t i = i0 + lane;
                const bool two = i0 != 0;
                if (i < ng) d0 = gran_diff(pa, pb, i, sh, n);
                if (two && i + 32 < ng) d1 = gran_diff(pa, pb, i + 32, sh, n);
                const bool b0 = (d0.x | d0.y | d0.z | d0.w) != 0;
                const bool b1 = (d1.x | d1.y | d1.z | d1.w) != 0;
                const unsigned bad0 = __ballot_sync(FULL, b0);
                const unsigned bad1 = __ballot_sync(FULL, b1);
                if (bad0 | bad1) {
                    ok = false;
                    if (P.collect) {
                        const int srcl = bad0 ? __ffs(bad0) - 1 : __ffs(bad1) - 1;  // lowest base index
                        uint32_t fb = 0;
                        if (lane == (uint32_t)srcl)
                            fb = bad0 ? 64u * i + first_diff_base(d0) : 64u * (i + 32) + first_diff_base(d1);
                        first_bad = __shfl_sync(FULL, fb, srcl);
                    }
                    break;
                }
                i0 += two ? 64 : 32;
            }
            if (P.collect) {
                const uint32_t mc = ok ? n : first_bad + 1;
                sum_m += mc;
                sum_b += 2ull * ((mc + 3) / 4);
            }
            if (ok && lane == j) my_ok = 1;
        }
        if (my_ok) {
            P.keys[c] = key | my_fl;
            ++hit_cnt;
        }
    }
    // per-warp totals -> one atomic each
    for (int d = 16; d; d >>= 1) hit_cnt += __shfl_down_sync(FULL, hit_cnt, d);
    if (lane == 0) {
        if (hit_cnt) atomicAdd(P.n_hits, hit_cnt);
        if (P.collect && sum_m) {
            atomicAdd(P.counters + 0, sum_m);
            atomicAdd(P.counters + 1, sum_b);
        }
    }
}

}  // namespace ovl
