// k6_exchange.cuh -- multi-GPU hit exchange (only with strand-mirror halving on N > 1 ranks).
//
// With mirror halving a rank compares one candidate of each strand-mirror pair and produces BOTH
// hit keys; the twin's A entry may belong to another rank's read range.  Rows are emitted by the
// rank that owns A (that is what keeps "smallest p per (a,b)" local), so the hit
// keys are partitioned by owner here and shipped with grouped ncclSend/ncclRecv over NVLink --
// the one real exchange step of the path (8 bytes per hit; no operand bytes move).
//
//   k6_count    per destination rank: how many of my hit keys it owns
//   k6_scatter  stable-less partition of the keys into per-destination runs
//
// Order inside a run is irrelevant: the receiver radix-sorts everything it owns before K5.
#pragma once
#include "ovl_common.cuh"

namespace ovl {

constexpr int MAX_RANKS = 64;

struct RankMap {
    uint32_t e[MAX_RANKS + 1];  // rank r owns entries [e[r], e[r+1])
    int n;
};

__device__ __forceinline__ int owner_of(const RankMap& R, uint32_t a) {
    int r = 0;
    while (r + 1 < R.n && a >= R.e[r + 1]) ++r;
    return r;
}

__global__ void __launch_bounds__(256)
k6_count(const uint64_t* __restrict__ keys, uint64_t n, KeyFmt fmt, const RankMap R,
         unsigned long long* __restrict__ counts) {
    __shared__ unsigned int s_cnt[MAX_RANKS];
    for (int i = threadIdx.x; i < R.n; i += blockDim.x) s_cnt[i] = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    for (uint64_t i0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~31ull; i0 < n;
         i0 += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = i0 + lane;
        const int dest = i < n ? owner_of(R, fmt.a(keys[i])) : -1;
        for (int r = 0; r < R.n; ++r) {
            const unsigned m = __ballot_sync(FULL, dest == r);
            if (m && lane == 0) atomicAdd(&s_cnt[r], __popc(m));
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < R.n; i += blockDim.x)
        if (s_cnt[i]) atomicAdd(counts + i, (unsigned long long)s_cnt[i]);
}

// counts[r] is final; cursor[r] starts at 0.  Run r begins at sum(counts[0..r)).
__global__ void __launch_bounds__(256)
k6_scatter(const uint64_t* __restrict__ keys, uint64_t n, KeyFmt fmt, const RankMap R,
           const unsigned long long* __restrict__ counts, unsigned long long* __restrict__ cursor,
           uint64_t* __restrict__ out) {
    __shared__ unsigned long long s_off[MAX_RANKS];
    if (threadIdx.x == 0) {
        unsigned long long acc = 0;
        for (int r = 0; r < R.n; ++r) {
            s_off[r] = acc;
            acc += counts[r];
        }
    }
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    for (uint64_t i0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~31ull; i0 < n;
         i0 += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = i0 + lane;
        uint64_t key = 0;
        int dest = -1;
        if (i < n) {
            key = keys[i];
            dest = owner_of(R, fmt.a(key));
        }
        for (int r = 0; r < R.n; ++r) {
            const unsigned m = __ballot_sync(FULL, dest == r);
            if (!m) continue;
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(cursor + r, (unsigned long long)__popc(m));
            base = __shfl_sync(FULL, base, 0);
            if (dest == r) out[s_off[r] + base + __popc(m & ((1u << lane) - 1u))] = key;
        }
    }
}

}  // namespace ovl
