// ovl_static.cu -- ovl_filter_rows_static: the row filter of k8_filter.cuh / ovl_filter_rows for
// callers that must not allocate.  (include/phasm_ovl.h has the contract.)
//
// Who this is for.  ovl_filter_rows grows its working buffers with cudaMalloc on first use (and
// whenever a job is larger than the last one).  Some hosts of this library cannot tolerate that:
//   * a pipeline that has handed all of the GPU's memory to its own pool (a caching allocator, an
//     arena carved up at start-up): a cudaMalloc from a plug-in either fails or fragments what the
//     pool believes it owns;
//   * a latency-bound loop: cudaMalloc / cudaFree synchronise the whole device, so every stream of
//     the process stalls behind one call that happens to need a bigger buffer;
//   * a process running under a memory budget that is checked once, at load time.
// For them this file offers the same filter with NO allocation at call time.  The caller hands in
// device pointers for the rows, the entry lengths and the results; everything in between (first-mark
// table, keep flags, tile offsets, counters) lives in ONE file-scope __device__ array below, which
// is part of the library image: each GPU that loads the library holds those OVL_STATIC_SCRATCH_BYTES
// from module load until the process exits, used or not.  The price is a fixed maximum job size.
//
// The array's name is only meaningful to device code.  On the host it is a stand-in symbol, not a
// device address, so the host side asks the runtime for the real address once per device
// (cudaGetSymbolAddress, with that device current) and passes it to the kernels as a plain pointer.
//
// Sharing.  There is exactly one scratch array per GPU, shared by every call from every thread and
// every stream, and it keeps whatever the last call left in it.  Each call therefore clears the
// parts that must start clean (first-mark table, counters) as its own first kernel, and two calls
// that can run on the same GPU at the same time (different streams, no ordering between them)
// race on it.  Calls on one stream are ordered by the stream and are safe back to back.
//
// Self-contained translation unit: nothing here touches ovl.cu's context, buffers or kernels.
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/phasm_ovl.h"

namespace ovl_static {

struct Row {
    uint32_t a, b, astart, aend, bend;
};
static_assert(sizeof(Row) == sizeof(ovl_row) && sizeof(Row) == 20, "row record layout");

constexpr uint32_t MAX_ROWS = OVL_STATIC_MAX_ROWS;
constexpr uint32_t MAX_ENTRIES = OVL_STATIC_MAX_ENTRIES;
constexpr uint32_t TILE_ROWS = 2048;  // rows per block of the compaction: 256 threads x 8 consecutive rows
constexpr uint32_t MAX_TILES = MAX_ROWS / TILE_ROWS;
static_assert(MAX_ROWS % TILE_ROWS == 0 && MAX_TILES <= 1024, "one 1024-thread block scans the tile counts");
constexpr uint32_t UNMARKED = 0xffffffffu;

// scratch layout, byte offsets into g_scratch (every part 256-byte aligned)
constexpr size_t OFF_FIRSTMARK = 0;                                         // uint32[MAX_ENTRIES]
constexpr size_t OFF_KEEP = OFF_FIRSTMARK + (size_t)MAX_ENTRIES * 4;        // uint8[MAX_ROWS]
constexpr size_t OFF_KEEP2 = OFF_KEEP + (size_t)MAX_ROWS;                   // uint8[MAX_ROWS]
constexpr size_t OFF_TILES = OFF_KEEP2 + (size_t)MAX_ROWS;                  // uint32[MAX_TILES]
constexpr size_t OFF_CTR = OFF_TILES + (size_t)MAX_TILES * 4;               // uint64[4] (+ padding)
constexpr size_t SCRATCH_BYTES = OFF_CTR + 256;
static_assert(SCRATCH_BYTES == OVL_STATIC_SCRATCH_BYTES, "the header states the size of the static scratch");
static_assert(OFF_KEEP % 256 == 0 && OFF_KEEP2 % 256 == 0 && OFF_TILES % 256 == 0 && OFF_CTR % 256 == 0, "alignment");

// The scratch itself.  Lives in the library's device image; one copy per GPU that loads it.
static __device__ __align__(256) unsigned char g_scratch[SCRATCH_BYTES];

enum Counter { C_CONTAINED = 0, C_MARKED = 1, C_BAD = 2 };

struct Params {
    const Row* rows;
    uint32_t n;
    const uint32_t* len;  // [m] entry lengths
    uint32_t m;
    uint32_t* firstmark;      // scratch
    uint8_t* keep;            // scratch
    uint8_t* keep2;           // scratch
    uint32_t* tiles;          // scratch
    unsigned long long* ctr;  // scratch
    uint32_t contained_reads, min_read_length, min_overlap_length, max_overhang, collapse_pairs;
    long long max_overhang_abs;
    double max_overhang_rel;
    Row* out;
    ovl_static_counts* counts;
};

enum RowClass : uint32_t { ROW_A_CONTAINED = 0, ROW_B_CONTAINED = 1, ROW_OVERLAP_AB = 2 };

// LocalAlignment.classify with brange = (0, bend), as in k8_filter.cuh
__device__ __forceinline__ uint32_t row_class(const Row& r, uint32_t la, uint32_t lb) {
    const long long ta = (long long)la - (long long)r.aend, tb = (long long)lb - (long long)r.bend;
    if (r.astart == 0 && ta <= tb) return ROW_A_CONTAINED;
    if (ta >= tb) return ROW_B_CONTAINED;
    return ROW_OVERLAP_AB;
}

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// What the previous call left behind must not leak into this one.
__global__ void __launch_bounds__(256) ks_clear(const Params P) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    for (uint32_t e = t; e < P.m; e += gridDim.x * blockDim.x) P.firstmark[e] = UNMARKED;
    if (t < 4) P.ctr[t] = 0;
}

// ContainedReads, first half: firstmark[e] = the first row that marks entry e as contained.
__global__ void __launch_bounds__(256) ks_mark(const Params P) {
    unsigned long long n_cont = 0, n_new = 0, n_bad = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < P.n; i += gridDim.x * blockDim.x) {
        const Row r = P.rows[i];
        if (r.a >= P.m || r.b >= P.m) {
            ++n_bad;
            continue;
        }
        const uint32_t cls = row_class(r, __ldg(P.len + r.a), __ldg(P.len + r.b));
        if (cls == ROW_A_CONTAINED) n_new += atomicMin(P.firstmark + r.a, i) == UNMARKED;
        if (cls == ROW_B_CONTAINED) n_new += atomicMin(P.firstmark + r.b, i) == UNMARKED;
        n_cont += cls != ROW_OVERLAP_AB;
    }
    n_cont = warp_sum(n_cont);
    n_new = warp_sum(n_new);
    n_bad = warp_sum(n_bad);
    if ((threadIdx.x & 31) == 0) {
        if (n_cont) atomicAdd(P.ctr + C_CONTAINED, n_cont);
        if (n_new) atomicAdd(P.ctr + C_MARKED, n_new);
        if (n_bad) atomicAdd(P.ctr + C_BAD, n_bad);
    }
}

// The filter chain of phasm/cli/assembler.py:76-86 on one row (same arithmetic as k8_keep).
__global__ void __launch_bounds__(256) ks_keep(const Params P, int count_bad) {
    unsigned long long n_bad = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < P.n; i += gridDim.x * blockDim.x) {
        const Row r = P.rows[i];
        bool keep = r.a < P.m && r.b < P.m;
        if (!keep) {
            ++n_bad;
        } else {
            const uint32_t la = __ldg(P.len + r.a), lb = __ldg(P.len + r.b);
            if (P.contained_reads)
                keep = row_class(r, la, lb) == ROW_OVERLAP_AB && P.firstmark[r.a] >= i && P.firstmark[r.b] >= i;
            if (keep && P.min_read_length) keep = la >= P.min_read_length && lb >= P.min_read_length;
            const uint32_t ovl_len = max(r.aend - r.astart, r.bend);
            if (keep && P.min_overlap_length) keep = ovl_len >= P.min_overlap_length;
            if (keep && P.max_overhang) {
                const long long ta = (long long)la - (long long)r.aend, tb = (long long)lb - (long long)r.bend;
                const long long overhang = ta < tb ? ta : tb;
                const double rel = P.max_overhang_rel * (double)ovl_len;
                const double thr = (double)P.max_overhang_abs < rel ? (double)P.max_overhang_abs : rel;
                keep = (double)overhang <= thr;
            }
        }
        P.keep[i] = keep ? 1 : 0;
    }
    if (count_bad) {  // ks_mark did not run, so nobody has counted the out-of-range rows yet
        n_bad = warp_sum(n_bad);
        if ((threadIdx.x & 31) == 0 && n_bad) atomicAdd(P.ctr + C_BAD, n_bad);
    }
}

// Of the kept rows of one ordered pair (adjacent in canonical order) only the last one stays.
__global__ void __launch_bounds__(256) ks_collapse(const Params P) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < P.n; i += gridDim.x * blockDim.x) {
        uint8_t k = P.keep[i];
        if (k) {
            const Row r = P.rows[i];
            for (uint32_t j = i + 1; j < P.n; ++j) {
                const Row q = P.rows[j];
                if (q.a != r.a || q.b != r.b) break;
                if (P.keep[j]) {
                    k = 0;
                    break;
                }
            }
        }
        P.keep2[i] = k;
    }
}

// Order-preserving compaction in three small steps (no library scan, so no temporary storage to size).
__global__ void __launch_bounds__(256) ks_tile_count(const Params P, const uint8_t* __restrict__ flag) {
    __shared__ uint32_t s[8];
    const uint32_t base = blockIdx.x * TILE_ROWS + threadIdx.x * 8;
    uint32_t c = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j)
        if (base + j < P.n) c += flag[base + j];
    c = (uint32_t)warp_sum(c);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += s[w];
        P.tiles[blockIdx.x] = t;
    }
}

// One block: tile counts -> tile offsets (in place), and the four result counters.
__global__ void __launch_bounds__(1024) ks_tile_scan(const Params P, uint32_t n_tiles) {
    __shared__ uint32_t s[1024];
    const uint32_t t = threadIdx.x;
    const uint32_t mine = t < n_tiles ? P.tiles[t] : 0;
    s[t] = mine;
    __syncthreads();
    for (uint32_t d = 1; d < 1024; d <<= 1) {
        const uint32_t v = t >= d ? s[t - d] : 0;
        __syncthreads();
        s[t] += v;
        __syncthreads();
    }
    if (t < n_tiles) P.tiles[t] = s[t] - mine;
    if (t == 0) {
        P.counts->rows_kept = s[1023];
        P.counts->rows_contained = P.ctr[C_CONTAINED];
        P.counts->entries_removed = P.ctr[C_MARKED];
        P.counts->rows_bad = P.ctr[C_BAD];
    }
}

__global__ void __launch_bounds__(256) ks_scatter(const Params P, const uint8_t* __restrict__ flag) {
    __shared__ uint32_t s[8];
    const uint32_t base = blockIdx.x * TILE_ROWS + threadIdx.x * 8;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t f = 0, c = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j)
        if (base + j < P.n && flag[base + j]) {
            f |= 1u << j;
            ++c;
        }
    uint32_t incl = c;  // inclusive scan of the per-thread counts over the warp
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    if (lane == 31) s[warp] = incl;
    __syncthreads();
    uint32_t off = P.tiles[blockIdx.x] + incl - c;
    for (uint32_t w = 0; w < warp; ++w) off += s[w];
#pragma unroll
    for (int j = 0; j < 8; ++j)
        if (f & (1u << j)) P.out[off++] = P.rows[base + j];
}

// ---- host side --------------------------------------------------------------------------------
namespace {
constexpr int MAX_DEVICES = 64;
struct DeviceState {
    std::atomic<int> ready{0};
    unsigned char* scratch = nullptr;  // the DEVICE address of g_scratch on this GPU
    int n_sm = 0;
};
DeviceState g_dev[MAX_DEVICES];
std::mutex g_setup_mutex;

int fail(char* err, size_t cap, int code, const char* fmt, ...) {
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

// Once per device, with that device current: turn the host-side stand-in `g_scratch` into the
// address the kernels can use.
int setup_device(int device, char* err, size_t cap) {
    DeviceState& D = g_dev[device];
    if (D.ready.load(std::memory_order_acquire)) return OVL_OK;
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (D.ready.load(std::memory_order_relaxed)) return OVL_OK;
    int major = 0, n_sm = 0;
    cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess)
        return fail(err, cap, OVL_E_CUDA, "ovl_filter_rows_static: cannot query device %d: %s", device, cudaGetErrorString(e));
    if (major < 10)
        return fail(err, cap, OVL_E_CUDA, "ovl_filter_rows_static: device %d is sm_%dx; this library is built for sm_100a only", device, major);
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_scratch);
    if (e != cudaSuccess || !p)
        return fail(err, cap, OVL_E_CUDA, "ovl_filter_rows_static: cudaGetSymbolAddress of the static scratch failed on device %d: %s",
                    device, cudaGetErrorString(e));
    D.scratch = static_cast<unsigned char*>(p);
    D.n_sm = n_sm;
    D.ready.store(1, std::memory_order_release);
    return OVL_OK;
}

int run(int device, cudaStream_t st, const ovl_row* d_rows, uint64_t n_rows, const uint32_t* d_entry_len, uint32_t n_entries,
        const ovl_filter_opts* o, ovl_row* d_rows_out, ovl_static_counts* d_counts, char* err, size_t cap) {
    int rc = setup_device(device, err, cap);
    if (rc != OVL_OK) return rc;
    const DeviceState& D = g_dev[device];
    Params P{};
    P.rows = reinterpret_cast<const Row*>(d_rows);
    P.n = (uint32_t)n_rows;
    P.len = d_entry_len;
    P.m = n_entries;
    P.firstmark = reinterpret_cast<uint32_t*>(D.scratch + OFF_FIRSTMARK);
    P.keep = D.scratch + OFF_KEEP;
    P.keep2 = D.scratch + OFF_KEEP2;
    P.tiles = reinterpret_cast<uint32_t*>(D.scratch + OFF_TILES);
    P.ctr = reinterpret_cast<unsigned long long*>(D.scratch + OFF_CTR);
    P.contained_reads = o->contained_reads;
    P.min_read_length = o->min_read_length;
    P.min_overlap_length = o->min_overlap_length;
    P.max_overhang = o->max_overhang;
    P.collapse_pairs = o->collapse_pairs;
    P.max_overhang_abs = (long long)o->max_overhang_abs;
    P.max_overhang_rel = o->max_overhang_rel;
    P.out = reinterpret_cast<Row*>(d_rows_out);
    P.counts = d_counts;

    const unsigned cap_blocks = (unsigned)D.n_sm * 16;
    unsigned row_blocks = (P.n + 255) / 256;
    if (row_blocks > cap_blocks) row_blocks = cap_blocks;
    unsigned ent_blocks = (P.m + 255) / 256;
    if (ent_blocks > cap_blocks) ent_blocks = cap_blocks;
    if (ent_blocks == 0) ent_blocks = 1;
    const uint32_t n_tiles = (P.n + TILE_ROWS - 1) / TILE_ROWS;

    ks_clear<<<ent_blocks, 256, 0, st>>>(P);
    const uint8_t* flag = P.keep;
    if (P.n) {
        if (o->contained_reads) ks_mark<<<row_blocks, 256, 0, st>>>(P);
        ks_keep<<<row_blocks, 256, 0, st>>>(P, o->contained_reads ? 0 : 1);
        if (o->collapse_pairs) {
            ks_collapse<<<row_blocks, 256, 0, st>>>(P);
            flag = P.keep2;
        }
        ks_tile_count<<<n_tiles, 256, 0, st>>>(P, flag);
    }
    ks_tile_scan<<<1, 1024, 0, st>>>(P, n_tiles);
    if (P.n) ks_scatter<<<n_tiles, 256, 0, st>>>(P, flag);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        return fail(err, cap, OVL_E_CUDA, "ovl_filter_rows_static: kernel launch failed on device %d: %s (%s)", device,
                    cudaGetErrorName(e), cudaGetErrorString(e));
    return OVL_OK;
}

}  // namespace
}  // namespace ovl_static

extern "C" int ovl_filter_rows_static(int device, void* stream, const ovl_row* d_rows, uint64_t n_rows,
                                      const uint32_t* d_entry_len, uint32_t n_entries, const ovl_filter_opts* opts,
                                      ovl_row* d_rows_out, ovl_static_counts* d_counts, char* err, size_t err_cap) {
    using namespace ovl_static;
    if (err && err_cap) err[0] = 0;
    if (!opts || !d_counts || (n_rows && (!d_rows || !d_rows_out)) || (n_entries && !d_entry_len))
        return fail(err, err_cap, OVL_E_ARG, "ovl_filter_rows_static: NULL argument");
    if (device < 0 || device >= MAX_DEVICES)
        return fail(err, err_cap, OVL_E_ARG, "ovl_filter_rows_static: device %d out of range", device);
    if (n_rows > MAX_ROWS)
        return fail(err, err_cap, OVL_E_LIMIT,
                    "ovl_filter_rows_static: %llu rows, the static scratch holds at most OVL_STATIC_MAX_ROWS = %u "
                    "(use ovl_filter_rows, which allocates, or split the job)",
                    (unsigned long long)n_rows, MAX_ROWS);
    if (n_entries > MAX_ENTRIES)
        return fail(err, err_cap, OVL_E_LIMIT,
                    "ovl_filter_rows_static: %u entries, the static scratch holds at most OVL_STATIC_MAX_ENTRIES = %u "
                    "(use ovl_filter_rows, which allocates)",
                    n_entries, MAX_ENTRIES);
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
    if (e != cudaSuccess)
        return fail(err, err_cap, OVL_E_CUDA, "ovl_filter_rows_static: cannot select CUDA device %d: %s", device, cudaGetErrorString(e));
    const int rc = run(device, static_cast<cudaStream_t>(stream), d_rows, n_rows, d_entry_len, n_entries, opts, d_rows_out,
                       d_counts, err, err_cap);
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    return rc;
}
