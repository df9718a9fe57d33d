// k1_pack.cuh -- K0 granule->entry map, K1 ASCII -> 2-bit packing with on-device reverse
// complement.
//
// Replaces, on the device, what the reference does on the host per FASTA record:
// `add_sequence(name+"+", seq)` / `add_sequence(name+"-", dinopy.reverse_complement(seq))`
// (/root/reference/phasm/cli/assembler.py:39-40) followed by the byte copy into the
// StringSet (/root/reference/src/overlapper.cpp:22-26).
//
// Roofline: HBM.  Algorithmic bytes per FORWARD base when both strands are added:
// 1.0 (ASCII read once; the second strand's re-read is an L2 hit by construction: entries
// 2i and 2i+1 are adjacent in the output, hence adjacent in time) + 0.5 (2 x 2 bit written).
#pragma once
#include "ovl_common.cuh"

namespace ovl {

// K0: one warp per entry writes the entry index into every granule it owns.
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

// K1: one thread per 16-base unit (32 bits of output); a warp writes 128 contiguous bytes.
// err[0] is set to (entry index + 1) of some entry holding a non-ACGT byte.
__global__ void __launch_bounds__(256)
k1_pack(const uint8_t* __restrict__ ascii, const uint64_t* __restrict__ src,
        const uint32_t* __restrict__ len, const uint8_t* __restrict__ flags,
        const uint32_t* __restrict__ gran0, const uint32_t* __restrict__ gran_ent,
        uint32_t* __restrict__ packed32, uint64_t u_begin, uint64_t u_end,
        uint32_t* __restrict__ err) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t t = u_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < u_end;
         t += stride) {
        const uint32_t g = (uint32_t)(t >> 2);
        const uint32_t e = gran_ent[g];
        const uint32_t L = len[e];
        const uint32_t u = (uint32_t)(t - 4ull * gran0[e]);
        const int64_t base0 = 16ll * u;
        int nvalid = (int64_t)L - base0 >= 16 ? 16 : (int)((int64_t)L - base0);
        uint32_t out = 0;
        if (nvalid > 0) {
            const bool rc = flags[e] & 1;
            // byte offset of the 16 source bytes; for rc the window is read backwards
            const int64_t o = rc ? (int64_t)src[e] + (int64_t)L - base0 - 16 : (int64_t)src[e] + base0;
            const uint8_t* p = ascii + o;
            const uint32_t sh = (uint32_t)((uintptr_t)p & 3u);
            const uint32_t* q = reinterpret_cast<const uint32_t*>(p - sh);
            uint32_t w[5];
#pragma unroll
            for (int i = 0; i < 5; ++i) w[i] = __ldg(q + i);
            uint32_t bad = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                uint32_t f = __funnelshift_r(w[i], w[i + 1], 8 * sh);
                // bytes outside the entry are replaced by 'A' (packs to 0, validates clean)
                const uint32_t m = rc ? ~low_bytes_mask(16 - nvalid - 4 * i) : low_bytes_mask(nvalid - 4 * i);
                f = (f & m) | (0x41414141u & ~m);
                out |= pack4(f, bad) << (8 * i);
            }
            if (rc) out = reverse_groups2(out) ^ 0xAAAAAAAAu;
            if (nvalid < 16) out &= (1u << (2 * nvalid)) - 1u;
            if (bad) atomicMax(err, e + 1);
        }
        packed32[t] = out;
    }
}

}  // namespace ovl
