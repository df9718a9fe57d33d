// k2_anchor.cuh -- K2: prefix-anchor hash table.
//
// Replaces the reference's whole index construction (generalised suffix array + LCP + child
// table, /root/reference/src/overlapper.cpp:33-36).  Every row of overlaps() has a B with
// B[0:n] == A[p:p+n], n >= l, so for any delta with delta + k <= l the k-mer B[delta:delta+k]
// equals A[p+delta : p+delta+k].  Indexing the k-mers of every B at ALL offsets
// delta in [0, stride) therefore lets the scan probe only the A positions q = 0 mod stride:
// for a row at p the unique delta = (-p) mod stride is found at q = p + delta.
// (stride = 1 is the canonical "first k-mer of B" anchor of SURVEY.md 8(d).)
//
//   k      = min(max(l,1), 32)
//   stride = largest power of two <= min(64, l - k + 1)
//
// The table is open addressing on the packed k-mer with a linked list of anchor ids
// (b * stride + delta) per key: one kernel, no sort.  Which anchor ends up where in a list
// depends on thread timing, but the candidates are radix-sorted before they are verified, so
// the output is deterministic.
//
// Roofline: HBM/L2-latency; tiny (one 16-byte slot + 4 B per anchor).
#pragma once
#include "ovl_common.cuh"

namespace ovl {

struct AnchorParams {
    const uint64_t* packed64;
    const uint2* ent;       // {gran0, len}
    uint32_t M;
    uint32_t l;
    int k;
    int stride_log2;
    Slot* table;            // (1 << tbits) + 1 slots, all {EMPTY_KEY, NIL, 0}
    int tbits;
    uint32_t* next;         // [M << stride_log2]
    uint32_t* bitmap;       // sparse path: 1 << bm_bits bits   (dense path: bloom words)
    int bm_bits;            // sparse: log2 bits; dense: log2 WORDS of the blocked bloom filter
    int dense;              // 1: stride == 1, fill the blocked bloom filter instead of the bitmap
    unsigned long long* n_anchors;  // entries with |b| >= l
};

// One thread per (entry, delta).
__global__ void __launch_bounds__(256) k2_insert(const AnchorParams P) {
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t b = (uint32_t)(idx >> P.stride_log2);
    bool is_anchor0 = false;
    if (b < P.M) {
        const uint2 m = __ldg(P.ent + b);
        if (m.y >= P.l) {
            const uint32_t delta = (uint32_t)idx & ((1u << P.stride_log2) - 1u);
            is_anchor0 = delta == 0;
            const uint64_t* w = P.packed64 + 2ull * m.x + (delta >> 5);
            const uint64_t x = kmer_at(w[0], w[1], delta & 31) & kmer_mask(P.k);
            const uint64_t h = mix64(x);
            uint32_t s;
            if (x == EMPTY_KEY) {
                s = 1u << P.tbits;  // reserved slot
                P.table[s].key = x;
            } else {
                const uint32_t tmask = (1u << P.tbits) - 1u;
                s = slot_of(h, P.tbits);
                for (;;) {
                    const unsigned long long old = atomicCAS(&P.table[s].key, EMPTY_KEY, x);
                    if (old == EMPTY_KEY || old == x) break;
                    s = (s + 1) & tmask;
                }
            }
            P.next[idx] = atomicExch(&P.table[s].head, (uint32_t)idx);
            if (P.dense) {
                const uint32_t h32 = kmer_hash(x);
                atomicOr(&P.bitmap[filter_word(h32, P.bm_bits)], filter_bits(h32));
            } else {
                const uint32_t bit = bitmap_bit(h, P.bm_bits);
                atomicOr(&P.bitmap[bit >> 5], 1u << (bit & 31));
            }
        }
    }
    const unsigned mk = __ballot_sync(FULL, is_anchor0);
    if ((threadIdx.x & 31) == 0 && mk) atomicAdd(P.n_anchors, (unsigned long long)__popc(mk));
}

// Table lookup shared by the scan kernels: head of the anchor list of k-mer x, NIL if absent.
__device__ __forceinline__ uint32_t table_find(const Slot* __restrict__ table, int tbits, uint64_t x,
                                               uint64_t h) {
    if (x == EMPTY_KEY) return table[1u << tbits].head;
    const uint32_t tmask = (1u << tbits) - 1u;
    uint32_t s = slot_of(h, tbits);
    for (;;) {
        const uint4 v = *reinterpret_cast<const uint4*>(table + s);
        const uint64_t key = (uint64_t)v.x | ((uint64_t)v.y << 32);
        if (key == x) return v.z;
        if (key == EMPTY_KEY) return NIL;
        s = (s + 1) & tmask;
    }
}

}  // namespace ovl
