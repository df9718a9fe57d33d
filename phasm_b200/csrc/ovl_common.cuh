// ovl_common.cuh -- device layout and small helpers shared by the overlapper kernels.
//
// HBM layout (all arrays live for the duration of one ovl_overlaps* call, cached in the ctx):
//
//   ascii     u8     staged sequence bytes, every record 16-byte aligned, 16 B slack at both
//                    ends (the reverse-complement reader touches up to 15 bytes before a
//                    record and 4 after it)
//   packed    uint4  one 16-byte GRANULE = 64 bases, 2 bits/base, base j of a granule in bits
//                    [2j, 2j+2) of the granule's 128-bit little-endian value; every entry
//                    starts on a granule boundary and is zero padded to it; one extra zero
//                    granule terminates the array.  code = (ascii >> 1) & 3  (A0 C1 T2 G3), so
//                    complement = code ^ 2.
//   gran_ent  u32    entry index of every granule
//   gran0     u32    [M+1] first granule of every entry (prefix sum of ceil(len/64))
//   len       u32    [M]   entry lengths in bases
//   src       u64    [M]   byte offset of the entry's source record in `ascii`
//   flags     u8     [M]   bit0: entry is the reverse complement of its source record
//
// Reference semantics restated by these kernels: /root/reference/src/overlapper.cpp:28-150
// (see SURVEY.md 3.3 for the normative R1/R2 statement).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ovl {

constexpr int GRAN_BASES = 64;
constexpr unsigned FULL = 0xffffffffu;

// Candidate: suffix a[p:] against prefix of b.
struct __align__(16) Cand {
    uint32_t a, p, b, pad;
};

// Anchor hash-table slot: key = first k bases of the B entries [start, start+count) of the
// key-sorted anchor list.  count == 0 marks an empty slot.
struct __align__(16) Slot {
    uint64_t key;
    uint32_t start, count;
};

struct Row {
    uint32_t a, b, astart, aend, bend;
};

// Bit layout of a 64-bit hit key: [a | b | p | cont | sp], most significant first.
struct KeyFmt {
    int bits_e;  // bits of an entry index
    int bits_p;  // bits of a position
    __host__ __device__ int total() const { return 2 * bits_e + bits_p + 2; }
};
constexpr uint64_t HIT_SP = 1;    // rem <= |b|: suffix-prefix (R1 candidate)
constexpr uint64_t HIT_CONT = 2;  // rem >= |b|: b wholly inside a (R2 row)

__host__ __device__ __forceinline__ uint64_t kmer_mask(int k) {
    return k >= 32 ? ~0ull : ((1ull << (2 * k)) - 1ull);
}

// 32-bit hash of a packed k-mer: high half of a 64-bit multiplicative hash.
__host__ __device__ __forceinline__ uint32_t kmer_hash(uint64_t x) {
    return (uint32_t)((x * 0x9E3779B97F4A7C15ull) >> 32);
}
// second-level mix for the table slot (slow path only)
__host__ __device__ __forceinline__ uint32_t slot_hash(uint32_t h) {
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 12;
    return h;
}
// Blocked bloom filter: one 32-bit word per key, two bits in it.
__host__ __device__ __forceinline__ uint32_t filter_word(uint32_t h, int fw_bits) {
    return h >> (32 - fw_bits);
}
__host__ __device__ __forceinline__ uint32_t filter_bits(uint32_t h) {
    return (1u << (h & 31)) | (1u << ((h >> 5) & 31));
}

// k-mer starting at base j (0..31) of word w0, continuing into w1.
__device__ __forceinline__ uint64_t kmer_at(uint64_t w0, uint64_t w1, int j) {
    const int s = 2 * j;
    return s ? ((w0 >> s) | (w1 << (64 - s))) : w0;
}

}  // namespace ovl
