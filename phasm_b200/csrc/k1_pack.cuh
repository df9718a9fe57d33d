// k1_pack.cuh -- K0 granule->entry map, K1 ASCII -> 2-bit packing with the reverse complement
// produced in the same pass.
//
// Replaces, on the device, what the reference does on the host per FASTA record:
// `add_sequence(name+"+", seq)` / `add_sequence(name+"-", dinopy.reverse_complement(seq))`
// (/root/reference/phasm/cli/assembler.py:39-40) followed by the byte copy into the
// StringSet (/root/reference/src/overlapper.cpp:22-26).
//
// One thread packs one 16-base UNIT (one aligned 128-bit ASCII load -> one 32-bit word) of a
// forward entry.  If the entry has a reverse-complement twin (flags bit0 of the next entry),
// the same thread also writes one word of the twin: rc word j of a read of L = 16m + r bases
// is the bit-reversed, complemented 32-bit window that starts 2r bits into forward word
// m-j-1, so a thread needs its own word and its right neighbour's, which comes by warp
// shuffle.  Warps are tiled with a one-unit overlap (lane 31 only supplies the neighbour
// word), so the ASCII is read once (+3 %) and never re-read for the second strand.
//
// Roofline: HBM.  Algorithmic bytes per FORWARD base when both strands are added:
// 1.0 (ASCII read once) + 0.5 (2 x 2 bit written).
#pragma once
#include "ovl_common.cuh"

namespace ovl {

// K0: one warp per entry writes the entry index into every granule it owns and the packed
// {first granule, length} pair the scan / verify kernels read.
__global__ void __launch_bounds__(256) k0_gran_entry(const uint32_t* __restrict__ gran0,
                                                     uint32_t* __restrict__ gran_ent,
                                                     uint32_t e_begin, uint32_t e_end) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t e = e_begin + warp; e < e_end; e += nwarps) {
        const uint32_t g0 = gran0[e], g1 = gran0[e + 1];
        for (uint32_t g = g0 + lane; g < g1; g += 32) gran_ent[g] = e;
    }
}

__global__ void __launch_bounds__(256) k0_ent_table(const uint32_t* __restrict__ gran0,
                                                    const uint32_t* __restrict__ len,
                                                    uint2* __restrict__ ent, uint32_t M) {
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < M) ent[e] = make_uint2(gran0[e], len[e]);
}

__device__ __forceinline__ uint32_t low_bytes_mask(int n) {  // n low bytes set, n clamped to 0..4
    return n >= 4 ? 0xffffffffu : (n <= 0 ? 0u : ((1u << (8 * n)) - 1u));
}

// Four ASCII bytes (upper-case ACGT expected) -> 8 bits, byte i in bits [2i,2i+2).
// `bad` accumulates a non-zero value if any byte is outside {A,C,G,T}.
__device__ __forceinline__ uint32_t pack4(uint32_t w, uint32_t& bad) {
    const uint32_t x = w ^ 0x41414141u;  // A->00 C->02 G->06 T->15
    bad |= (x & 0xE8E8E8E8u) | ((x ^ (x >> 4)) & 0x01010101u) |
           ((((x >> 2) & ~(x >> 1)) ^ (x >> 4)) & 0x01010101u);
    uint32_t c = (w >> 1) & 0x03030303u;
    c = c | (c >> 6) | (c >> 12) | (c >> 18);
    return c & 0xffu;
}

__device__ __forceinline__ uint32_t reverse_groups2(uint32_t x) {
    const uint32_t r = __brev(x);
    return ((r >> 1) & 0x55555555u) | ((r & 0x55555555u) << 1);
}

// K1.  Units [u_begin, u_end) are the 16-base units of the granules of the local entries
// (4 per granule); warp w of the grid owns units u_begin + 31*w + [0, 31) and lane 31 only
// computes the neighbour word.  err[0] is set to (entry index + 1) of some entry holding a
// non-ACGT byte.
__global__ void __launch_bounds__(256)
k1_pack(const uint8_t* __restrict__ ascii, const uint64_t* __restrict__ src,
        const uint2* __restrict__ ent, const uint8_t* __restrict__ flags,
        const uint32_t* __restrict__ gran_ent, uint32_t* __restrict__ packed32, uint32_t M,
        uint64_t u_begin, uint64_t u_end, uint32_t* __restrict__ err) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    const uint64_t n_tiles = (u_end - u_begin + 30) / 31;
    for (uint64_t tile = warp; tile < n_tiles; tile += nwarps) {
        const uint64_t t = u_begin + tile * 31 + lane;
        uint32_t W = 0, e = 0, u = 0, L = 0, g0 = 0;
        bool fwd = false;
        if (t < u_end) {
            const uint32_t g = (uint32_t)(t >> 2);
            e = __ldg(gran_ent + g);
            fwd = !(__ldg(flags + e) & 1);
            if (fwd) {
                const uint2 m = __ldg(ent + e);
                g0 = m.x;
                L = m.y;
                u = (uint32_t)(t - 4ull * g0);
                const int nvalid = (int)L - 16 * (int)u;  // may be <= 0 (padding unit)
                if (nvalid > 0) {
                    const uint4 v = ldg_stream(reinterpret_cast<const uint4*>(ascii + __ldg(src + e) + 16ull * u));
                    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
                    uint32_t bad = 0;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        // bytes outside the entry are replaced by 'A' (packs to 0, validates clean)
                        const uint32_t m4 = low_bytes_mask(nvalid - 4 * i);
                        W |= pack4((w[i] & m4) | (0x41414141u & ~m4), bad) << (8 * i);
                    }
                    if (bad) atomicMax(err, e + 1);
                }
            }
        }
        const uint32_t Wn_s = __shfl_down_sync(FULL, W, 1);
        if (!fwd || lane == 31) continue;
        packed32[t] = W;
        if (e + 1 >= M || !(__ldg(flags + e + 1) & 1)) continue;  // no reverse-complement twin
        const uint32_t units = 4u * ((L + 63u) >> 6);
        const uint32_t Wn = (u + 1 < units) ? Wn_s : 0u;        // next unit of the SAME entry
        const uint32_t m = L >> 4, r = L & 15u;
        uint32_t* rc = packed32 + 4ull * __ldg(ent + e + 1).x;
        if (u < m) rc[m - u - 1] = reverse_groups2(__funnelshift_r(W, Wn, 2 * r)) ^ 0xAAAAAAAAu;
        if (u == 0 && r) rc[m] = (reverse_groups2(W << (32 - 2 * r)) ^ 0xAAAAAAAAu) & ((1u << (2 * r)) - 1u);
        if (u >= m + (r ? 1u : 0u)) rc[u] = 0u;
    }
}

}  // namespace ovl
