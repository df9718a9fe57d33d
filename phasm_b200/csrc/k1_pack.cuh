// k1_pack.cuh -- K0 tile->entry maps, K1 ASCII -> 2-bit packing with the reverse complement
// produced in the same pass, k1_rc (multi-GPU: twin strands of the reads other ranks packed).
//
// Replaces, on the device, what the reference does on the host per FASTA record:
// `add_sequence(name+"+", seq)` / `add_sequence(name+"-", dinopy.reverse_complement(seq))`
// (/root/reference/phasm/cli/assembler.py:39-40) followed by the byte copy into the
// StringSet (/root/reference/src/overlapper.cpp:22-26).
//
// A warp packs one TILE = 128 consecutive 16-base units (2048 bases) of ONE forward entry, so
// all entry metadata is warp-uniform: four coalesced 128-bit ASCII loads per lane are in flight
// before the first is packed (unit 32*i + lane), and every unit becomes one 32-bit word,
// stored coalesced.  If the entry has a reverse-complement twin (flags bit0 of the next
// entry) the same thread also writes one word of the twin: rc word j of a read of
// L = 16m + r bases is the bit-reversed, complemented 32-bit window that starts 2r bits into
// forward word m-j-1, so a thread needs its own word and its right neighbour's, which comes
// by warp shuffle (the tile's last unit re-packs one halo unit).  The ASCII is read once and
// never re-read for the second strand.  The tile -> entry map is written by K0.  The forward
// words go to the forward region of the packed array, the twin's to the twin region (ovl_common.cuh).
//
// Roofline: HBM.  Algorithmic bytes per FORWARD base when both strands are added:
// 1.0 (ASCII read once) + 0.5 (2 x 2 bit written).
#pragma once
#include "ovl_common.cuh"

namespace ovl {

// K0 (run when the entry table is uploaded, not per pass): the maps that tell a warp which
// entry a tile of work belongs to -- tile_ent for K1 (tiles of 2048 bases = 128 sixteen-base units
// inside ONE forward entry) and scan_tiles for the scan (tiles of 32 granules inside ONE entry of
// either strand: one 16-byte record {first granule, entry, length, first granule's index in the
// entry} per tile, so that a warp learns everything about its tile from a single uniform load).
// 16 bytes per 2048 bases instead of the 4 bytes per granule a granule -> entry array costs
// (52 MB of extra scan traffic on config 2 in round 1).
struct ScanRanges {
    uint32_t g_begin[2], g_end[2];  // granule ranges of the local entries: forward-region slice, twin-region slice
    uint32_t n_tiles;               // scan tiles of the local entries
};

__global__ void __launch_bounds__(256) k0_tile_ent(const uint32_t* __restrict__ tile0, uint32_t* __restrict__ tile_ent,
                                                   uint32_t e_begin, uint32_t e_end) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t e = e_begin + warp; e < e_end; e += nwarps) {
        const uint32_t t0 = tile0[e], t1 = tile0[e + 1];
        for (uint32_t t = t0 + lane; t < t1; t += 32) tile_ent[t] = e;
    }
}

// stile0[e - e_begin] = first scan tile of local entry e (prefix sum of ceil(granules / 32))
__global__ void __launch_bounds__(256) k0_scan_tiles(const uint32_t* __restrict__ stile0, const uint32_t* __restrict__ gran0,
                                                     const uint32_t* __restrict__ len, uint4* __restrict__ scan_tiles,
                                                     uint32_t e_begin, uint32_t e_end) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t e = e_begin + warp; e < e_end; e += nwarps) {
        const uint32_t t0 = stile0[e - e_begin], t1 = stile0[e - e_begin + 1];
        const uint32_t g0 = gran0[e], L = len[e];
        for (uint32_t t = t0 + lane; t < t1; t += 32) scan_tiles[t] = make_uint4(g0 + 32u * (t - t0), e, L, 32u * (t - t0));
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

// Four ASCII bytes -> their 2-bit codes in the TOP byte of the result (base i in bits
// [24+2i, 24+2i+2)); `bad` becomes non-zero if a byte is not one of upper-case A, C, G, T.
// code = (byte >> 1) & 3 (A0 C1 T2 G3).  The multiply gathers the four 2-bit fields (disjoint
// bit ranges, so no carries); the check rebuilds the expected letters from the codes with a
// byte permute of "ACTG" and compares them with the input.
__device__ __forceinline__ uint32_t pack4_top(uint32_t w, uint32_t& bad) {
    const uint32_t c = (w >> 1) & 0x03030303u;
    const uint32_t s = c | (c >> 4);                                  // byte0 = c0|c1<<4, byte2 = c2|c3<<4
    const uint32_t sel = __byte_perm(s, 0, 0x4420);                   // c0 | c1<<4 | c2<<8 | c3<<12
    bad |= __byte_perm(0x47544341u /* "ACTG" */, 0, sel) ^ w;
    return c * 0x01041040u;
}

// 16 ASCII bytes (one uint4) -> one 32-bit packed word
__device__ __forceinline__ uint32_t pack16(const uint4 v, uint32_t& bad) {
    const uint32_t t0 = pack4_top(v.x, bad), t1 = pack4_top(v.y, bad);
    const uint32_t t2 = pack4_top(v.z, bad), t3 = pack4_top(v.w, bad);
    return __byte_perm(__byte_perm(t0, t1, 0x0073), __byte_perm(t2, t3, 0x0073), 0x5410);
}

// the same for a unit with only nvalid (1..15) bases inside its entry: the rest reads as 'A'
__device__ __forceinline__ uint32_t pack16_partial(uint4 v, int nvalid, uint32_t& bad) {
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t m4 = low_bytes_mask(nvalid - 4 * i);
        w[i] = (w[i] & m4) | (0x41414141u & ~m4);
    }
    return pack16(make_uint4(w[0], w[1], w[2], w[3]), bad);
}

__device__ __forceinline__ uint32_t reverse_groups2(uint32_t x) {
    const uint32_t r = __brev(x);
    return ((r >> 1) & 0x55555555u) | ((r & 0x55555555u) << 1);
}

constexpr uint32_t PACK_TILE_UNITS = 128;  // 16-base units per warp tile

struct PackParams {
    const uint8_t* ascii;
    const uint64_t* src;       // [M] byte offset of the entry's record
    const uint2* ent;          // [M] {gran0, len}
    const uint8_t* flags;      // [M]
    const uint32_t* tile0;     // [M+1] first tile of every entry (0 tiles for twins / remote entries)
    const uint32_t* tile_ent;  // [n_tiles]
    uint32_t* packed32;
    uint32_t M;
    uint32_t t_begin, t_end;   // tiles of the local entries
    uint32_t* err;             // set to (entry index + 1) of some entry holding a non-ACGT byte
};

// K1: one warp per tile (grid-stride).
__global__ void __launch_bounds__(256) k1_pack(const PackParams P) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t tile = P.t_begin + warp; tile < P.t_end; tile += nwarps) {
        const uint32_t e = __ldg(P.tile_ent + tile);
        const uint2 m_e = __ldg(P.ent + e);
        const uint32_t L = m_e.y;
        const uint32_t units = 4u * ((L + 63u) >> 6);
        const uint32_t u0 = (tile - __ldg(P.tile0 + e)) * PACK_TILE_UNITS;
        const uint4* in = reinterpret_cast<const uint4*>(P.ascii + __ldg(P.src + e));
        uint32_t* out = P.packed32 + 4ull * m_e.x;
        const bool twin = e + 1 < P.M && (__ldg(P.flags + e + 1) & 1);

        // ---- load: 4 (+1 halo) independent 128-bit loads per lane ----
        uint4 v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t u = u0 + 32u * i + lane;
            v[i] = make_uint4(0x41414141u, 0x41414141u, 0x41414141u, 0x41414141u);
            if (16u * u < L) v[i] = ldg_stream(in + u);
        }
        const uint32_t uh = u0 + PACK_TILE_UNITS;  // halo: first unit of the next tile
        uint4 vh = make_uint4(0x41414141u, 0x41414141u, 0x41414141u, 0x41414141u);
        const bool need_halo = twin && lane == 31 && 16u * uh < L;
        if (need_halo) vh = ldg_stream(in + uh);

        // ---- pack ----
        uint32_t W[4], bad = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t u = u0 + 32u * i + lane;
            const int nvalid = (int)L - 16 * (int)u;
            W[i] = nvalid >= 16 ? pack16(v[i], bad) : (nvalid > 0 ? pack16_partial(v[i], nvalid, bad) : 0u);
            if (u < units) out[u] = W[i];
        }
        uint32_t Wh = 0;
        if (need_halo) {
            const int nvalid = (int)L - 16 * (int)uh;
            Wh = nvalid >= 16 ? pack16(vh, bad) : pack16_partial(vh, nvalid, bad);
        }
        if (bad) atomicMax(P.err, e + 1);
        if (!twin) continue;  // warp-uniform

        // ---- reverse-complement twin ----
        const uint32_t m = L >> 4, r = L & 15u;
        uint32_t* rc = P.packed32 + 4ull * __ldg(P.ent + e + 1).x;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t u = u0 + 32u * i + lane;
            // right neighbour: next lane of this round, lane 0 of the next round, or the halo
            uint32_t Wn = __shfl_down_sync(FULL, W[i], 1);
            const uint32_t first_next = __shfl_sync(FULL, i < 3 ? W[i < 3 ? i + 1 : 3] : 0u, 0);
            if (lane == 31) Wn = i < 3 ? first_next : Wh;
            if (u >= units) continue;
            if (u < m) rc[m - u - 1] = reverse_groups2(__funnelshift_r(W[i], Wn, 2 * r)) ^ 0xAAAAAAAAu;
            if (u == 0 && r) rc[m] = (reverse_groups2(W[i] << (32 - 2 * r)) ^ 0xAAAAAAAAu) & ((1u << (2 * r)) - 1u);
            if (u >= m + (r ? 1u : 0u)) rc[u] = 0u;
        }
    }
}

// Multi-GPU: the reverse-complement strand of the REMOTE reads.  Only the forward region of the
// packed array travels between the ranks (half the NVLink volume of replicating both strands,
// and it lands in place: no staging copy); the twin region of the remote reads is rebuilt here.
// One block per entry of [e_begin, e_end): an entry followed by its twin reads its own granules
// and writes the twin's, one thread per granule (same word arithmetic as the twin part of
// k1_pack); all other blocks exit.
__global__ void __launch_bounds__(128) k1_rc(const uint2* __restrict__ ent, const uint8_t* __restrict__ flags,
                                             uint32_t* __restrict__ packed32, uint32_t e_begin, uint32_t e_end,
                                             uint32_t M) {
    const uint32_t e = e_begin + blockIdx.x;
    if (e >= e_end || e + 1 >= M) return;
    if ((__ldg(flags + e) & 1) || !(__ldg(flags + e + 1) & 1)) return;
    const uint2 m_e = __ldg(ent + e);
    const uint32_t L = m_e.y;
    const uint32_t ng = (L + 63u) >> 6, units = 4u * ng;
    const uint4* src = reinterpret_cast<const uint4*>(packed32) + m_e.x;
    uint32_t* rc = packed32 + 4ull * __ldg(ent + e + 1).x;
    const uint32_t m = L >> 4, r = L & 15u;
    for (uint32_t ig = threadIdx.x; ig < ng; ig += blockDim.x) {
        const uint4 v = src[ig];
        const uint32_t wn = ig + 1 < ng ? reinterpret_cast<const uint32_t*>(src + ig + 1)[0] : 0u;
        const uint32_t W[5] = {v.x, v.y, v.z, v.w, wn};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t u = 4u * ig + k;
            if (u >= units) continue;
            if (u < m) rc[m - u - 1] = reverse_groups2(__funnelshift_r(W[k], W[k + 1], 2 * r)) ^ 0xAAAAAAAAu;
            if (u == 0 && r) rc[m] = (reverse_groups2(W[k] << (32 - 2 * r)) ^ 0xAAAAAAAAu) & ((1u << (2 * r)) - 1u);
            if (u >= m + (r ? 1u : 0u)) rc[u] = 0u;
        }
    }
}

}  // namespace ovl
