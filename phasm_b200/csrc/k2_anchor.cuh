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
// The table is open addressing on a 32-bit fingerprint of the k-mer with a linked list of anchor ids
// (b * stride + delta) per key: one kernel, no sort.  Which anchor ends up where in a list
// depends on thread timing, but the hit keys are sorted before rows are made, so the output is
// deterministic.  On N GPUs the key space is PARTITIONED by hash (ovl_common.cuh): every rank
// hashes all M * stride keys but inserts only its own 1/N, so the build cost and the table size
// per GPU do not grow with N.
//
// Roofline: HBM/L2-latency; tiny (one 16-byte slot + 4 B per anchor).
#pragma once
#include "ovl_common.cuh"

namespace ovl {

struct AnchorParams {
    const uint64_t* pre64;  // [M][4]: the first two granules (128 bases) of every entry
    const uint2* ent;       // {gran0, len}
    uint32_t M;
    uint32_t l;
    int k;
    int stride_log2;
    Slot* table;            // THIS rank's partition: 1 << tbits slots, all EMPTY_SLOT
    int tbits;
    uint32_t* next;         // [M << stride_log2], written only for nodes that have a successor
    uint32_t* bitmap;       // sparse path: nr slices of 1 << slice_log2 bits (this rank fills its own)
                            // dense path: 1 << bm_bits bloom WORDS (every rank fills all of it)
    int bm_bits;            // sparse: log2 bits of ONE slice; dense: log2 words of the blocked bloom filter
    int dense;              // 1: stride == 1, fill the blocked bloom filter instead of the bitmap
    uint32_t me, nr;        // this rank / ranks: only keys with owner_of_hash() == me are inserted
    unsigned long long* flags;  // STF_TABLE is raised if the partition is full (cannot happen at the load factor ovl.cu sizes it for)
};

// Prefix table: the first two granules (128 bases) of every entry in one compact array.  K2 only
// ever needs bases [0, stride + k - 1) <= [0, 95) of an entry; on N > 1 GPUs this small array
// (32 B per entry) is what the ranks exchange BEFORE the anchor build, so that the big replication
// of the packed reads can overlap K2 and K3 (ovl.cu).
//
// It is packed straight from the staged ASCII records (forward entries read the record's first
// 128 bytes, reverse-complement entries its last 128 backwards): K2 then depends on nothing K1
// writes, so the anchor build runs on its own stream UNDER the pack kernel (K1 is bound by the
// integer pipes, K2 by L2 atomic latency: they overlap almost for free).
__global__ void __launch_bounds__(128) k2_prefix_ascii(const uint8_t* __restrict__ ascii, const uint64_t* __restrict__ src,
                                                       const uint32_t* __restrict__ len, const uint8_t* __restrict__ flags,
                                                       uint4* __restrict__ pre, uint32_t e_begin, uint32_t e_end) {
    const uint32_t e = e_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= e_end) return;
    const uint32_t L = __ldg(len + e);
    const uint8_t* rec = ascii + __ldg(src + e);
    const bool rc = __ldg(flags + e) & 1;
    const uint32_t nb = L < 128u ? L : 128u;
    uint64_t w[4] = {0, 0, 0, 0};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint64_t acc = 0;
#pragma unroll 8
        for (uint32_t j = 0; j < 32; ++j) {
            const uint32_t i = 32u * q + j;
            if (i < nb) {
                const uint32_t ch = rc ? rec[L - 1u - i] : rec[i];
                acc |= (uint64_t)(((ch >> 1) & 3u) ^ (rc ? 2u : 0u)) << (2 * j);
            }
        }
        w[q] = acc;
    }
    pre[2ull * e] = make_uint4((uint32_t)w[0], (uint32_t)(w[0] >> 32), (uint32_t)w[1], (uint32_t)(w[1] >> 32));
    pre[2ull * e + 1] = make_uint4((uint32_t)w[2], (uint32_t)(w[2] >> 32), (uint32_t)w[3], (uint32_t)(w[3] >> 32));
}

// One thread per (entry, delta); a thread whose key belongs to another rank's partition stops
// after the hash.  Insert protocol: an empty slot is claimed with a 64-bit compare-and-swap
// {fingerprint | id}; a slot that already holds the key gets the new id exchanged into its low
// word with the HAS_NEXT bit set, and the previous list pointer (flag included) goes to next[id].
// Slots are only ever claimed, never released, so a reader that sees the key's fingerprint may
// exchange without re-checking.
__global__ void __launch_bounds__(256) k2_insert(const AnchorParams P) {
    // grid-stride: the host may launch fewer blocks than keys / 256 so that the kernel trickles along
    // beside K1 instead of taking its SM slots (k2_blocks_per_sm)
    const uint64_t n_idx = (uint64_t)P.M << P.stride_log2;
    for (uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n_idx; idx += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t b = (uint32_t)(idx >> P.stride_log2);
        const uint2 m = __ldg(P.ent + b);
        if (m.y >= P.l) {
            const uint32_t delta = (uint32_t)idx & ((1u << P.stride_log2) - 1u);
            const uint64_t* w = P.pre64 + 4ull * b + (delta >> 5);  // delta + k <= 95: words 0..2
            const uint64_t x = kmer_at(w[0], w[1], delta & 31) & kmer_mask(P.k);
            const uint64_t h = mix64(x);
            if (P.dense) {  // replicated filter, partitioned table
                const uint32_t h32 = kmer_hash(x);
                atomicOr(&P.bitmap[filter_word(h32, P.bm_bits)], filter_bits(h32));
            }
            if (owner_of_hash(h, P.nr) == P.me) {
                const uint32_t fp = (uint32_t)h;
                const uint32_t tmask = (1u << P.tbits) - 1u;
                uint32_t s = slot_of(h, P.nr, P.tbits);
                uint32_t tries = 0;
                // (Reading the slot before the compare-and-swap is deliberate: an optimistic CAS-first
                // variant measured slower on B200, 0.34 -> 0.43 ms for pack + anchor build.)
                for (;; ++tries) {
                    if (tries > tmask) {
                        atomicOr(P.flags, STF_TABLE);
                        break;
                    }
                    unsigned long long v = *reinterpret_cast<volatile unsigned long long*>(P.table + s);
                    if (v == EMPTY_SLOT) {
                        v = atomicCAS(P.table + s, EMPTY_SLOT, ((unsigned long long)fp << 32) | (uint32_t)idx);
                        if (v == EMPTY_SLOT) break;  // claimed the slot: a singleton so far, nothing else to write
                    }
                    if ((uint32_t)(v >> 32) == fp) {  // the key's slot: push on its list
                        const uint32_t prev = atomicExch(reinterpret_cast<uint32_t*>(P.table + s), (uint32_t)idx | HAS_NEXT);
                        P.next[idx] = prev;
                        break;
                    }
                    s = (s + 1) & tmask;
                }
                if (!P.dense) {
                    const uint64_t bit = bitmap_bit(h, P.nr, P.bm_bits);
                    atomicOr(&P.bitmap[bit >> 5], 1u << (bit & 31));
                }
            }
        }
    }
    // (Round 1 counted the anchor entries here with one atomicAdd per warp on ONE address: 55 000
    // serialised L2 atomics, which is what the kernel's 0.16 ms were -- the time did not change
    // when the stride, and with it the number of keys, was halved.  The host knows the count.)
}

}  // namespace ovl
