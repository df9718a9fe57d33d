// k2_anchor.cuh -- K2: prefix-anchor keys, hash table and bloom filter.
//
// Replaces the reference's whole index construction (generalised suffix array + LCP + child
// table, /root/reference/src/overlapper.cpp:33-36): every row of overlaps() has a B whose first
// k = min(max(l,1),32) bases equal A[p:p+k], so indexing ONE k-mer per entry is sufficient.
//
// Roofline: HBM, but tiny: 8 B key read/written a few times per entry (32 B/entry algorithmic).
#pragma once
#include "ovl_common.cuh"

namespace ovl {

// keys[i] = first k bases of entry anc_ent[i] (entries with len >= l only).
__global__ void __launch_bounds__(256)
k2_keys(const uint64_t* __restrict__ packed64, const uint32_t* __restrict__ gran0,
        const uint32_t* __restrict__ anc_ent, uint64_t* __restrict__ keys, uint32_t n,
        uint64_t kmask) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keys[i] = packed64[2ull * gran0[anc_ent[i]]] & kmask;
}

// One thread per sorted anchor; run heads insert (key -> [start, count)) and set filter bits.
__global__ void __launch_bounds__(256)
k2_build(const uint64_t* __restrict__ keys, uint32_t n, Slot* __restrict__ table, int tbits,
         uint32_t* __restrict__ filter, int fw_bits, unsigned long long* __restrict__ n_keys) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool head = false;
    if (i < n) {
        const uint64_t key = keys[i];
        head = (i == 0) || (keys[i - 1] != key);
        if (head) {
            // run end: gallop, then bisect (runs are length 1 except on repeats / tiny k)
            uint32_t lo = i, step = 1;  // keys[lo] == key
            while (lo + step < n && keys[lo + step] == key) {
                lo += step;
                step <<= 1;
            }
            uint32_t hi = (lo + step < n) ? lo + step : n;  // keys[hi] != key or hi == n
            while (hi - lo > 1) {
                const uint32_t mid = lo + (hi - lo) / 2;
                if (keys[mid] == key) lo = mid; else hi = mid;
            }
            const uint32_t count = hi - i;
            const uint32_t h = kmer_hash(key);
            const uint32_t tmask = (1u << tbits) - 1u;
            uint32_t s = slot_hash(h) >> (32 - tbits);
            const unsigned long long val = ((unsigned long long)count << 32) | i;  // {start, count}
            // all inserted keys are distinct: claim the first empty slot
            while (atomicCAS(reinterpret_cast<unsigned long long*>(&table[s].start), 0ull, val) != 0ull)
                s = (s + 1) & tmask;
            table[s].key = key;
            atomicOr(&filter[filter_word(h, fw_bits)], filter_bits(h));
        }
    }
    const unsigned m = __ballot_sync(FULL, head);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(n_keys, (unsigned long long)__popc(m));
}

}  // namespace ovl
