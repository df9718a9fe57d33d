// k10_graph.cuh -- optional post-pass: the assembly-graph edge list of the reference's `layout`
// stage over the (filtered) rows (graph part).
//
// Reference: build_assembly_graph (phasm/assembly_graph.py:136-179), fed by the filtered
// LocalAlignment stream of phasm/cli/assembler.py:93-100.  For every alignment, in stream order:
//
//   OVERLAP_AB:  add_edge(a,   b,   weight = astart - bstart,                     overlap_len)
//                add_edge(b^1, a^1, weight = (|b| - bend) - (|a| - aend),         overlap_len)
//   OVERLAP_BA:  add_edge(b,   a,   weight = bstart - astart,                     overlap_len)
//                add_edge(a^1, b^1, weight = (|a| - aend) - (|b| - bend),         overlap_len)
//   A_CONTAINED / B_CONTAINED: nothing
//
// with classify / get_overlap_length of phasm/alignments.py:237-258, node x^1 = OrientedRead.reverse()
// (the other strand of the same read: entries come in strand pairs, assembler.py:39-40).  The graph
// is a networkx DiGraph: adding an edge that exists REPLACES its attributes (the last alignment of an
// ordered node pair wins) and keeps its place in the adjacency order (the first one decides that); a
// node exists as soon as it is an end of some edge.
//
// Device form: every OVERLAP row yields two (key = u<<32|v, seq = 2*row + which) items; the items are
// radix-sorted on the key (stable, so the seq of equal keys stay ascending); the LAST item of a run
// of equal keys gives the edge its attributes, the FIRST its insertion rank.  Output: one record
// per edge, sorted by (u, v) = a CSR adjacency; `first_seq` lets a consumer rebuild networkx's
// insertion order.  Roofline: HBM (20 B per row read twice, 12 B per item sorted, 24 B per edge).
#pragma once
#include "k8_filter.cuh"
#include "ovl_common.cuh"

namespace ovl {

struct Edge {
    uint32_t u, v;
    int32_t edge_len;       // the reference's `weight`
    uint32_t overlap_len;
    uint64_t first_seq;     // 2 * (index of the row that first added the edge) + (0: its first add_edge, 1: its second)
};

struct GraphParams {
    const Row* rows;
    uint64_t n;
    const uint2* ent;              // {gran0, len}
    uint32_t* flag;                // [n] 1 = the row is of an OVERLAP type
    const uint64_t* off;           // [n] exclusive sum of flag
    uint64_t* keys;                // [2 * n_ovl]
    uint32_t* seqs;                // [2 * n_ovl]
    uint32_t* last;                // [2 * n_ovl] 1 = last item of its run
    const uint64_t* eoff;          // exclusive sum of last
    Edge* edges;
    uint32_t* node_mark;           // [M] 1 = node is an end of some edge
    uint64_t n_items;
};

// alignment type with brange = (bstart, bend); rows of this library have bstart == 0
enum AlnClass : uint32_t { ALN_A_CONTAINED = 0, ALN_B_CONTAINED = 1, ALN_OVERLAP_AB = 2, ALN_OVERLAP_BA = 3 };

__host__ __device__ __forceinline__ uint32_t aln_class(uint32_t astart, uint32_t aend, uint32_t bstart, uint32_t bend,
                                                       uint32_t la, uint32_t lb) {
    const long long ta = (long long)la - (long long)aend, tb = (long long)lb - (long long)bend;
    if (astart <= bstart && ta <= tb) return ALN_A_CONTAINED;
    if (astart >= bstart && ta >= tb) return ALN_B_CONTAINED;
    return astart >= bstart ? ALN_OVERLAP_AB : ALN_OVERLAP_BA;
}

// the two add_edge calls of one OVERLAP alignment (which = 0, 1), in the reference's order
__host__ __device__ __forceinline__ void aln_edge(uint32_t cls, int which, uint32_t a, uint32_t b, uint32_t astart,
                                                  uint32_t aend, uint32_t bstart, uint32_t bend, uint32_t la, uint32_t lb,
                                                  uint32_t& u, uint32_t& v, int32_t& w) {
    const long long ta = (long long)la - (long long)aend, tb = (long long)lb - (long long)bend;
    if (cls == ALN_OVERLAP_AB) {
        if (which == 0) { u = a; v = b; w = (int32_t)((long long)astart - (long long)bstart); }
        else { u = b ^ 1u; v = a ^ 1u; w = (int32_t)(tb - ta); }
    } else {
        if (which == 0) { u = b; v = a; w = (int32_t)((long long)bstart - (long long)astart); }
        else { u = a ^ 1u; v = b ^ 1u; w = (int32_t)(ta - tb); }
    }
}

__global__ void __launch_bounds__(256) k10_flag(const GraphParams P) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.n; i += (uint64_t)gridDim.x * blockDim.x) {
        const Row r = P.rows[i];
        const uint32_t cls = aln_class(r.astart, r.aend, 0u, r.bend, __ldg(P.ent + r.a).y, __ldg(P.ent + r.b).y);
        P.flag[i] = cls >= ALN_OVERLAP_AB ? 1u : 0u;
    }
}

__global__ void __launch_bounds__(256) k10_items(const GraphParams P) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.n; i += (uint64_t)gridDim.x * blockDim.x) {
        if (!P.flag[i]) continue;
        const Row r = P.rows[i];
        const uint32_t la = __ldg(P.ent + r.a).y, lb = __ldg(P.ent + r.b).y;
        const uint32_t cls = aln_class(r.astart, r.aend, 0u, r.bend, la, lb);
        const uint64_t o = 2ull * P.off[i];
#pragma unroll
        for (int which = 0; which < 2; ++which) {
            uint32_t u, v;
            int32_t w;
            aln_edge(cls, which, r.a, r.b, r.astart, r.aend, 0u, r.bend, la, lb, u, v, w);
            P.keys[o + which] = ((uint64_t)u << 32) | v;
            P.seqs[o + which] = (uint32_t)(2ull * i + which);
            P.node_mark[u] = 1u;  // benign race: every writer stores 1
            P.node_mark[v] = 1u;
        }
    }
}

// after the sort: last[i] = item i closes its run of equal keys
__global__ void __launch_bounds__(256) k10_runs(const GraphParams P) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.n_items; i += (uint64_t)gridDim.x * blockDim.x)
        P.last[i] = (i + 1 == P.n_items || P.keys[i + 1] != P.keys[i]) ? 1u : 0u;
}

__global__ void __launch_bounds__(256) k10_edges(const GraphParams P) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.n_items; i += (uint64_t)gridDim.x * blockDim.x) {
        if (!P.last[i]) continue;
        const uint64_t key = P.keys[i];
        uint64_t j = i;  // first item of the run (runs are as long as an ordered pair has OVERLAP rows: short)
        while (j > 0 && P.keys[j - 1] == key) --j;
        const uint32_t seq = P.seqs[i];
        const Row r = P.rows[seq >> 1];
        const uint32_t la = __ldg(P.ent + r.a).y, lb = __ldg(P.ent + r.b).y;
        const uint32_t cls = aln_class(r.astart, r.aend, 0u, r.bend, la, lb);
        Edge e;
        aln_edge(cls, (int)(seq & 1u), r.a, r.b, r.astart, r.aend, 0u, r.bend, la, lb, e.u, e.v, e.edge_len);
        e.overlap_len = max(r.aend - r.astart, r.bend);
        e.first_seq = P.seqs[j];
        P.edges[P.eoff[i]] = e;
    }
}

}  // namespace ovl
