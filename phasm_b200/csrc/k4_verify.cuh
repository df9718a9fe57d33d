// k4_verify.cuh -- K4: warp-cooperative packed-word verifier.
//
// Replaces the suffix-tree emission tests of the reference: a candidate (a,p,b) is a row source
// iff a[p : p+n] == b[0 : n] with n = min(|a|-p, |b|).  rem = |a|-p <= |b| is the suffix-prefix
// case (R1, /root/reference/src/overlapper.cpp:64-91), rem >= |b| the containment case (R2,
// overlapper.cpp:93-115); rem == |b| is both, which is where the reference's duplicate rows
// come from.
//
// Input: the candidate keys as the scan flushed them (per-warp runs of one A entry).  Output:
// the keys of the candidates that verified, flag bits set, appended to the hit list.
//
// k4_verify_serial (the accounting form: exact m_c per candidate): a warp takes 32 consecutive
// keys, every lane decodes one key and fetches the two entry records in parallel, then the warp
// walks the 32 candidates; per step a lane compares one 64-base granule: B comes as one aligned
// streaming 128-bit load, A as two aligned 128-bit loads (granules i, i+1) funnel-shifted by
// (p mod 64) bases.  The first step covers 2048 bases, later steps 4096; the warp leaves a
// candidate at the first step holding a mismatch.
//
// Roofline: HBM/L2.  Algorithmic bytes: 2*ceil(m_c/4) per candidate (both operands, m_c bases
// up to and including the first mismatch), SURVEY.md 8(d).
#pragma once
#include "ovl_common.cuh"

namespace ovl {

struct VerifyParams {
    SegList<uint64_t> keys;            // candidate keys (flags clear), any order: the verify list of this rank
    const uint4* packed;
    const uint2* ent;                  // {gran0, len}
    KeyFmt fmt;
    CandSink hits;                     // verified keys with their flag bits set
    unsigned long long* counters;      // [0] sum m_c, [1] sum 2*ceil(m_c/4)   (collect only)
    int collect;
    MirrorRule mirror;
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

// keys a matching candidate contributes: [0] its own, [1] the SP twin, [2] the CONT twin (0 = none;
// a real key is never 0 because its flag bits are set).  See "Strand-mirror halving" below.
__device__ __forceinline__ void hit_keys(const KeyFmt& fmt, const uint2* __restrict__ ent, const MirrorRule& R,
                                         uint64_t key, uint32_t fl, uint64_t out[3]) {
    out[0] = key | fl;
    out[1] = out[2] = 0;
    if (!R.on) return;
    const uint32_t a = fmt.a(key), b = fmt.b(key), p = fmt.p(key);
    const uint32_t la = __ldg(ent + a).y, lb = __ldg(ent + b).y;
    const uint32_t bt = b ^ 1u, at = a ^ 1u;
    if ((fl & HIT_SP) && a != bt)  // (fl has SP only if this candidate is the one that compares)
        out[1] = fmt.make(bt, at, lb - (la - p)) | HIT_SP;
    if (fl & HIT_CONT) out[2] = fmt.make(at, bt, la - p - lb) | HIT_CONT;
}

// a hit key goes to the rank that owns its A entry (rows are emitted per A entry)
template <class Params>
__device__ __forceinline__ uint32_t hit_dest(const Params& P, uint64_t key) {
    return (uint32_t)owner_of(P.mirror.map, P.fmt.a(key));
}

__global__ void __launch_bounds__(256) k4_verify_serial(const VerifyParams P) {
    __shared__ uint64_t s_buf[8][WBUF];
    __shared__ RouteScratch s_rs[8];
    __shared__ unsigned long long s_pre[MAX_RANKS + 1];
    uint64_t* buf = s_buf[threadIdx.x >> 5];
    RouteScratch* rs = &s_rs[threadIdx.x >> 5];
    uint32_t nbuf = 0;
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    seg_prefix(P.keys, s_pre);
    const uint64_t n_all = s_pre[P.keys.nr];
    const uint64_t n_tiles = (n_all + 31) / 32;
    auto dest = [&](uint64_t k2) { return hit_dest(P, k2); };
    unsigned long long sum_m = 0, sum_b = 0;

    for (uint64_t tile = warp; tile < n_tiles; tile += nwarps) {
        const uint64_t c = tile * 32 + lane;
        uint64_t key = 0;
        uint32_t my_ga = 0, my_gb = 0, my_sh = 0, my_n = 0, my_fl = 0;
        if (c < n_all) {
            key = seg_item(P.keys, s_pre, c);
            const uint32_t a = P.fmt.a(key), b = P.fmt.b(key), p = P.fmt.p(key);
            const uint2 ma = __ldg(P.ent + a), mb = __ldg(P.ent + b);
            const uint32_t rem = ma.y - p;
            my_n = rem < mb.y ? rem : mb.y;
            my_fl = mirror_items(P.mirror, a, ma.y, b, mb.y, p);
            my_ga = ma.x + (p >> 6);
            my_gb = mb.x;
            my_sh = 2u * (p & 63u);
        }
        const uint32_t cnt = (uint32_t)((n_all - tile * 32) < 32 ? (n_all - tile * 32) : 32);
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
        uint64_t hk[3] = {0, 0, 0};
        if (my_ok) hit_keys(P.fmt, P.ent, P.mirror, key, my_fl, hk);
#pragma unroll
        for (int i = 0; i < 3; ++i) wbuf_push(buf, nbuf, P.hits, lane, hk[i] != 0, hk[i], rs, dest);
    }
    if (nbuf) wbuf_flush(buf, nbuf, P.hits, lane, rs, dest);
    if (lane == 0) {
        if (P.collect && sum_m) {
            atomicAdd(P.counters + 0, sum_m);
            atomicAdd(P.counters + 1, sum_b);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Two-pass throughput form (the default).  The serial kernel above walks a candidate chunk by
// chunk and leaves at the first mismatch: one dependent HBM round trip per chunk per warp, which
// is latency bound (ncu r01b: 55 % of the stall samples wait on the B load).  Error-free reads
// make the outcome predictable instead: a false candidate (another haplotype, a repeat copy)
// almost always differs within its first 2048 bases, and a candidate that survives them is
// almost always a true overlap that must be compared to its end anyway.  So:
//
//   k4a_head  every candidate's first FIRST_GRAN granules (2048 bases), 4 candidates per warp
//             step with all loads of a lane issued before the first compare.  Candidates that
//             end inside the head are appended to the hit list; the others that match so far go
//             to the survivor list.
//   k4b_tail  one warp per survivor compares the remaining granules, 4 granules per lane in
//             flight, no exit test inside the loop, one vote at the end.
//
// Both read A as one 128-bit + one 64-bit load per granule (load_a below).  Candidates need not be
// sorted: the scan flushes them in per-warp runs of one A entry, which is what keeps the A
// operand in L1.
// Bytes touched beyond the first mismatch are the price: <= 2048 bases per false candidate.
constexpr uint32_t FIRST_GRAN = 32;
constexpr int HEAD_UNROLL = 4;

// Strand-mirror halving.  When every read was added on both strands (entry 2i = read i, entry
// 2i+1 = its reverse complement: assembler.py:39-40) a match statement has an equivalent twin
// on the opposite strands (x' = x ^ 1 is the other strand of x):
//
//   SP   a[p:] == b[:rem]            <=>  b'[|b|-rem:] == a'[:rem]       key (b', a', |b|-rem)
//   CONT a[p:p+|b|] == b             <=>  a'[|a|-p-|b| : |a|-p] == b'    key (a', b', |a|-p-|b|)
//
// Both maps are involutions on (key, type) items, so exactly one item of every twin pair is
// compared (SP: picked by the swap-symmetric rule sp_self(); CONT: the one with a forward) and, if it matches, BOTH
// keys are appended to the hit list.  The scan finds every true hit on its own, so a true hit
// whose item is not the canonical one is always produced by its canonical twin; a false
// candidate's twin is false too.  Half of the operand bytes are never read.
// Multi-GPU: the twin's key may belong to another rank's A range; k6_exchange.cuh ships every
// hit key to the rank that owns its A entry before the sort.
struct Verify2Params {
    SegList<uint64_t> keys;            // candidate keys (flags clear), any order: the verify list of this rank
                                       // (sizes are device counters: no host round trip after the scan)
    const uint64_t* packed64;
    const uint2* ent;                  // {gran0, len}
    KeyFmt fmt;
    uint32_t* surv;                    // k4a -> k4b: logical indices (into the key list) of the candidates alive after the head
    unsigned long long* n_surv;
    CandSink hits;                     // verified keys with their flag bits set
    MirrorRule mirror;                 // on: the scan kept only the self-verifying candidates
    unsigned long long* tile_ctr;      // [0] k4a, [1] k4b: next tile to hand out (zeroed per call)
};

// dynamic tile hand-out: comparisons differ a lot in length, a static grid-stride split leaves
// the last wave unbalanced
__device__ __forceinline__ uint64_t next_tile(unsigned long long* ctr, uint32_t lane) {
    unsigned long long t = 0;
    if (lane == 0) t = atomicAdd(ctr, 1ull);
    return __shfl_sync(FULL, t, 0);
}

struct CandMeta {
    uint32_t ga;            // granule holding a[p]
    uint32_t gb;            // first granule of b
    uint32_t n, sh, fl;     // bases to compare (0: skip), bit offset of a[p] in its granule (0..126, even), HIT_* items
};

__device__ __forceinline__ CandMeta cand_meta(const Verify2Params& P, uint64_t key) {
    CandMeta m;
    const uint32_t a = P.fmt.a(key), b = P.fmt.b(key), p = P.fmt.p(key);
    const uint2 ma = __ldg(P.ent + a), mb = __ldg(P.ent + b);
    const uint32_t rem = ma.y - p;
    m.n = rem < mb.y ? rem : mb.y;
    m.fl = mirror_items(P.mirror, a, ma.y, b, mb.y, p);  // never 0: the scan dropped those
    m.ga = ma.x + (p >> 6);
    m.gb = mb.x;
    m.sh = 2u * (p & 63u);
    return m;
}

// The A operand of compare-granule i is bits [sh, sh+128) of the 256-bit value (G1:G0), G0 =
// granule ga+i, G1 = granule ga+i+1 (32-bit words x0..x7).  Only words x[v .. v+4], v = sh/32, are
// needed, and they always lie in ONE whole granule plus ONE aligned half of the other:
//
//   v = 0,1 (sh < 64):  big = G0 (x0..x3), small = low half of G1  (x4, x5)
//   v = 2,3 (sh >= 64): big = G1 (x4..x7), small = high half of G0 (x2, x3)
//
// so a lane issues one 128-bit and one 64-bit load for A (both addresses are plain arithmetic on
// sh: no branch in front of the loads) and one streaming 128-bit load for B: 12 L1 wavefronts per
// 32 granules instead of the 16 of three 64-bit A loads (ncu r01f: k4a_head sat at 88 % of the
// LSU wavefront rate).  The re-alignment is a warp-uniform 4-way switch on v and four funnel
// shifts by t = sh mod 32.
__device__ __forceinline__ void load_a(const uint4* __restrict__ g0, uint32_t hv, uint4& big, uint2& small) {
    big = __ldg(g0 + hv);
    small = __ldg(reinterpret_cast<const uint2*>(g0 + 1 - hv) + hv);
}

template <int V>
__device__ __forceinline__ uint4 align_a(const uint4 b, const uint2 s, uint32_t t) {
    if (V == 0)
        return make_uint4(__funnelshift_r(b.x, b.y, t), __funnelshift_r(b.y, b.z, t), __funnelshift_r(b.z, b.w, t),
                          __funnelshift_r(b.w, s.x, t));
    if (V == 1)
        return make_uint4(__funnelshift_r(b.y, b.z, t), __funnelshift_r(b.z, b.w, t), __funnelshift_r(b.w, s.x, t),
                          __funnelshift_r(s.x, s.y, t));
    if (V == 2)
        return make_uint4(__funnelshift_r(s.x, s.y, t), __funnelshift_r(s.y, b.x, t), __funnelshift_r(b.x, b.y, t),
                          __funnelshift_r(b.y, b.z, t));
    return make_uint4(__funnelshift_r(s.y, b.x, t), __funnelshift_r(b.x, b.y, t), __funnelshift_r(b.y, b.z, t),
                      __funnelshift_r(b.z, b.w, t));
}

__device__ __forceinline__ uint4 align_a_dyn(const uint4 b, const uint2 s, uint32_t sh) {
    const uint32_t t = sh & 31u;
    switch (sh >> 5) {  // warp-uniform
        case 0: return align_a<0>(b, s, t);
        case 1: return align_a<1>(b, s, t);
        case 2: return align_a<2>(b, s, t);
        default: return align_a<3>(b, s, t);
    }
}

__global__ void __launch_bounds__(256, 4) k4a_head(const Verify2Params P) {
    __shared__ uint64_t s_buf[8][WBUF];
    __shared__ RouteScratch s_rs[8];
    __shared__ unsigned long long s_pre[MAX_RANKS + 1];
    const uint32_t lane = threadIdx.x & 31;
    seg_prefix(P.keys, s_pre);
    const uint64_t n_all = s_pre[P.keys.nr];
    const uint64_t n_tiles = (n_all + 31) / 32;
    uint64_t* buf = s_buf[threadIdx.x >> 5];
    RouteScratch* rs = &s_rs[threadIdx.x >> 5];
    auto dest = [&](uint64_t k2) { return hit_dest(P, k2); };
    uint32_t nbuf = 0;
    const uint4* packed128 = reinterpret_cast<const uint4*>(P.packed64);

    for (uint64_t tile = next_tile(P.tile_ctr, lane); tile < n_tiles; tile = next_tile(P.tile_ctr, lane)) {
        const uint64_t c = tile * 32 + lane;
        uint64_t key = 0;
        CandMeta my = {0, 0, 0, 0, 0};  // n == 0 marks "no candidate" (a real one has n >= l >= 1)
        if (c < n_all) {
            key = seg_item(P.keys, s_pre, c);
            my = cand_meta(P, key);
        }
        uint32_t ok_mask = 0;  // bit j: candidate j matched over its head
        for (uint32_t j0 = 0; j0 < 32; j0 += HEAD_UNROLL) {
            if (__shfl_sync(FULL, my.n, j0) == 0) break;  // tile tail (warp-uniform)
            uint4 bv[HEAD_UNROLL], big[HEAD_UNROLL];
            uint2 sm[HEAD_UNROLL];
            uint32_t sh[HEAD_UNROLL], nn[HEAD_UNROLL], gl[HEAD_UNROLL];
#pragma unroll
            for (int u = 0; u < HEAD_UNROLL; ++u) {
                const uint32_t ga = __shfl_sync(FULL, my.ga, j0 + u);
                const uint32_t gb = __shfl_sync(FULL, my.gb, j0 + u);
                sh[u] = __shfl_sync(FULL, my.sh, j0 + u);
                nn[u] = __shfl_sync(FULL, my.n, j0 + u);
                // lanes past the end re-read the last granule (same address as the last live lane: no
                // extra wavefront, no zero-initialised registers, no predicates on the loads)
                // (an empty slot of the group, nn == 0, reads granule 0 of entry 0)
                gl[u] = nn[u] ? min(lane, (nn[u] - 1u) >> 6) : 0u;
                load_a(packed128 + ga + gl[u], sh[u] >> 6, big[u], sm[u]);
                bv[u] = ldg_stream(packed128 + gb + gl[u]);
            }
#pragma unroll
            for (int u = 0; u < HEAD_UNROLL; ++u) {
                const uint4 x = align_a_dyn(big[u], sm[u], sh[u]);
                uint4 d = make_uint4(x.x ^ bv[u].x, x.y ^ bv[u].y, x.z ^ bv[u].z, x.w ^ bv[u].w);
                const int nb = (int)nn[u] - 64 * (int)gl[u];
                if (nb < 64) {
                    const uint4 m = low_bases_mask128(nb);
                    d.x &= m.x; d.y &= m.y; d.z &= m.z; d.w &= m.w;
                }
                const unsigned bad = __ballot_sync(FULL, (d.x | d.y | d.z | d.w) != 0);
                if (!bad && nn[u]) ok_mask |= 1u << (j0 + u);
            }
        }
        const bool ok = (ok_mask >> lane) & 1u;
        const bool done = ok && my.n <= 64u * FIRST_GRAN;
        const bool alive = ok && !done;
        uint64_t hk[3] = {0, 0, 0};
        if (done) hit_keys(P.fmt, P.ent, P.mirror, key, my.fl, hk);
#pragma unroll
        for (int i = 0; i < 3; ++i) wbuf_push(buf, nbuf, P.hits, lane, hk[i] != 0, hk[i], rs, dest);
        const unsigned am = __ballot_sync(FULL, alive);
        if (am) {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(P.n_surv, (unsigned long long)__popc(am));
            base = __shfl_sync(FULL, base, 0);
            if (alive) P.surv[base + __popc(am & ((1u << lane) - 1u))] = (uint32_t)c;
        }
    }
    if (nbuf) wbuf_flush(buf, nbuf, P.hits, lane, rs, dest);
}

// OR of the differences of compare-granules [FIRST_GRAN, ceil(n/64)) for one lane (stride 32,
// TAIL_ROWS rows in flight).  Whole granules need no mask and cost 4 funnel shifts + 4 LOP3
// (acc |= x ^ b) each; the partial last granule (n mod 64 bases) is one extra masked compare by a
// single lane.  (ncu r01f: the previous form, which masked inside the loop, ran at 80 % of the
// ALU pipe -- the verifier was instruction bound, not HBM bound.)  A survivor has
// n > 64*FIRST_GRAN.
constexpr int TAIL_ROWS = 4;  // tail_diff's dispatch below is written for 4

// R rows (32 granules each) of one comparison, all loads issued before the first compare; in the
// last row the lanes past granule `glast` re-read it (same address as a live lane).
template <int V, int R>
__device__ __forceinline__ void tail_rows(const uint4* __restrict__ pa, const uint4* __restrict__ pb, uint32_t g0,
                                          uint32_t glast, uint32_t t, uint4& acc) {
    uint4 bv[R], big[R];
    uint2 sm[R];
#pragma unroll
    for (int u = 0; u < R; ++u) {
        uint32_t g = g0 + 32u * u;
        if (u == R - 1) g = min(g, glast);
        load_a(pa + g, V >> 1, big[u], sm[u]);
        bv[u] = ldg_stream(pb + g);
    }
#pragma unroll
    for (int u = 0; u < R; ++u) {
        const uint4 x = align_a<V>(big[u], sm[u], t);
        acc.x |= x.x ^ bv[u].x; acc.y |= x.y ^ bv[u].y; acc.z |= x.z ^ bv[u].z; acc.w |= x.w ^ bv[u].w;
    }
}

template <int V>
__device__ __forceinline__ uint4 tail_diff(const uint4* __restrict__ pa, const uint4* __restrict__ pb,
                                           uint32_t n, uint32_t t, uint32_t lane) {
    constexpr uint32_t HV = V >> 1;
    const uint32_t nf = n >> 6;  // whole granules (>= FIRST_GRAN for a survivor)
    uint4 acc = make_uint4(0, 0, 0, 0);
#pragma unroll 1
    for (uint32_t base = FIRST_GRAN; base < nf; base += 32u * TAIL_ROWS) {
        const uint32_t rows = (nf - base + 31u) >> 5;  // warp-uniform
        const uint32_t g0 = base + lane, gl = nf - 1u;
        if (rows >= 4) tail_rows<V, 4>(pa, pb, g0, gl, t, acc);
        else if (rows == 3) tail_rows<V, 3>(pa, pb, g0, gl, t, acc);
        else if (rows == 2) tail_rows<V, 2>(pa, pb, g0, gl, t, acc);
        else tail_rows<V, 1>(pa, pb, g0, gl, t, acc);
    }
    const int nb = (int)(n & 63u);
    if (nb && lane == (nf & 31u)) {
        uint4 big;
        uint2 sm;
        load_a(pa + nf, HV, big, sm);
        const uint4 bv = ldg_stream(pb + nf);
        const uint4 x = align_a<V>(big, sm, t);
        const uint4 m = low_bases_mask128(nb);
        acc.x |= (x.x ^ bv.x) & m.x; acc.y |= (x.y ^ bv.y) & m.y; acc.z |= (x.z ^ bv.z) & m.z; acc.w |= (x.w ^ bv.w) & m.w;
    }
    return acc;
}

__global__ void __launch_bounds__(256, 4) k4b_tail(const Verify2Params P) {
    __shared__ uint64_t s_buf[8][WBUF];
    __shared__ RouteScratch s_rs[8];
    __shared__ unsigned long long s_pre[MAX_RANKS + 1];
    RouteScratch* rs = &s_rs[threadIdx.x >> 5];
    auto dest = [&](uint64_t k2) { return hit_dest(P, k2); };
    const uint32_t lane = threadIdx.x & 31;
    seg_prefix(P.keys, s_pre);
    const uint64_t ns = *P.n_surv;
    const uint64_t n_tiles = (ns + 31) / 32;
    uint64_t* buf = s_buf[threadIdx.x >> 5];
    uint32_t nbuf = 0;
    const uint4* packed128 = reinterpret_cast<const uint4*>(P.packed64);
    // A warp takes 32 survivors at a time: every lane fetches and decodes ONE of them (one
    // dependent surv -> key -> entry chain per 32 survivors instead of per survivor), then the
    // warp walks the 32 comparisons, one operand round trip each.
    for (uint64_t tile = next_tile(P.tile_ctr + 1, lane); tile < n_tiles; tile = next_tile(P.tile_ctr + 1, lane)) {
        const uint64_t s = tile * 32 + lane;
        uint64_t key = 0;
        CandMeta my = {0, 0, 0, 0, 0};
        if (s < ns) {
            key = seg_item(P.keys, s_pre, P.surv[s]);
            my = cand_meta(P, key);
        }
        const uint32_t cnt = (uint32_t)((ns - tile * 32) < 32 ? (ns - tile * 32) : 32);
        uint32_t ok_mask = 0;
#pragma unroll 1
        for (uint32_t j = 0; j < cnt; ++j) {
            const uint32_t ga = __shfl_sync(FULL, my.ga, j);
            const uint32_t gb = __shfl_sync(FULL, my.gb, j);
            const uint32_t sh = __shfl_sync(FULL, my.sh, j);
            const uint32_t n = __shfl_sync(FULL, my.n, j);
            const uint4* pa = packed128 + ga;
            const uint4* pb = packed128 + gb;
            const uint32_t t = sh & 31u;
            uint4 acc;
            switch (sh >> 5) {  // warp-uniform
                case 0: acc = tail_diff<0>(pa, pb, n, t, lane); break;
                case 1: acc = tail_diff<1>(pa, pb, n, t, lane); break;
                case 2: acc = tail_diff<2>(pa, pb, n, t, lane); break;
                default: acc = tail_diff<3>(pa, pb, n, t, lane); break;
            }
            const unsigned bad = __ballot_sync(FULL, (acc.x | acc.y | acc.z | acc.w) != 0);
            if (!bad) ok_mask |= 1u << j;
        }
        uint64_t hk[3] = {0, 0, 0};
        if ((ok_mask >> lane) & 1u) hit_keys(P.fmt, P.ent, P.mirror, key, my.fl, hk);
#pragma unroll
        for (int i = 0; i < 3; ++i) wbuf_push(buf, nbuf, P.hits, lane, hk[i] != 0, hk[i], rs, dest);
    }
    if (nbuf) wbuf_flush(buf, nbuf, P.hits, lane, rs, dest);
}

}  // namespace ovl
