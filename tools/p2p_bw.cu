// p2p_bw.cu -- NVLink peer-to-peer bandwidth on this box, the ways the multi-GPU path could move the
// packed reads: copy-engine copies over 1..8 streams, SM stores into peer memory (push), SM loads
// from peer memory (pull), one direction and both directions at once.  Measurement tool only.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/p2p_bw tools/p2p_bw.cu && tools/p2p_bw
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <vector>

#define CK(x)                                                                      \
    do {                                                                           \
        cudaError_t e_ = (x);                                                      \
        if (e_ != cudaSuccess) {                                                   \
            fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
            exit(1);                                                               \
        }                                                                          \
    } while (0)

__global__ void k_copy(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n16) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

int main() {
    int nd = 0;
    CK(cudaGetDeviceCount(&nd));
    if (nd < 2) {
        printf("needs 2 GPUs\n");
        return 0;
    }
    const size_t bytes = 256ull << 20;
    uint4 *a[2], *b[2];
    for (int d = 0; d < 2; ++d) {
        CK(cudaSetDevice(d));
        CK(cudaDeviceEnablePeerAccess(1 - d, 0));
        CK(cudaMalloc(&a[d], bytes));
        CK(cudaMalloc(&b[d], bytes));
        CK(cudaMemset(a[d], 1, bytes));
        CK(cudaMemset(b[d], 2, bytes));
    }
    const int S = 8;
    cudaStream_t st[2][S];
    cudaEvent_t e0[2], e1[2];
    for (int d = 0; d < 2; ++d) {
        CK(cudaSetDevice(d));
        for (int i = 0; i < S; ++i) CK(cudaStreamCreateWithFlags(&st[d][i], cudaStreamNonBlocking));
        CK(cudaEventCreate(&e0[d]));
        CK(cudaEventCreate(&e1[d]));
    }
    auto sync_all = [&]() {
        for (int d = 0; d < 2; ++d) {
            CK(cudaSetDevice(d));
            CK(cudaDeviceSynchronize());
        }
    };
    // mode 0: copy engines, ns streams; mode 1: SM push kernel with nb blocks; mode 2: SM pull kernel
    auto run = [&](int mode, int ns, int nb, bool bidir, const char* what) {
        float best = 1e9f;
        for (int rep = 0; rep < 4; ++rep) {
            sync_all();
            for (int d = 0; d < (bidir ? 2 : 1); ++d) {
                CK(cudaSetDevice(d));
                CK(cudaEventRecord(e0[d], st[d][0]));
                for (int i = 1; i < ns; ++i) CK(cudaStreamWaitEvent(st[d][i], e0[d], 0));
                for (int i = 0; i < ns; ++i) {
                    const size_t o = bytes / 16 * i / ns, n = bytes / 16 * (i + 1) / ns - o;
                    if (mode == 0) CK(cudaMemcpyAsync(b[1 - d] + o, a[d] + o, n * 16, cudaMemcpyDefault, st[d][i]));
                    if (mode == 1) k_copy<<<nb / ns > 0 ? nb / ns : 1, 256, 0, st[d][i]>>>(a[d] + o, b[1 - d] + o, n);   // local read, remote write
                    if (mode == 2) k_copy<<<nb / ns > 0 ? nb / ns : 1, 256, 0, st[d][i]>>>(a[1 - d] + o, b[d] + o, n);   // remote read, local write
                }
                static cudaEvent_t ev[2][S];
                static bool made = false;
                if (!made) {
                    for (int dd = 0; dd < 2; ++dd) {
                        CK(cudaSetDevice(dd));
                        for (int i = 0; i < S; ++i) CK(cudaEventCreateWithFlags(&ev[dd][i], cudaEventDisableTiming));
                    }
                    made = true;
                    CK(cudaSetDevice(d));
                }
                for (int i = 1; i < ns; ++i) {
                    CK(cudaEventRecord(ev[d][i], st[d][i]));
                    CK(cudaStreamWaitEvent(st[d][0], ev[d][i], 0));
                }
                CK(cudaEventRecord(e1[d], st[d][0]));
            }
            sync_all();
            float ms = 0, m2 = 0;
            CK(cudaEventElapsedTime(&ms, e0[0], e1[0]));
            if (bidir) {
                CK(cudaEventElapsedTime(&m2, e0[1], e1[1]));
                if (m2 > ms) ms = m2;
            }
            if (ms < best) best = ms;
        }
        printf("%-46s %2d streams %4d blocks %s: %7.3f ms  %7.1f GB/s per direction\n", what, ns, nb, bidir ? "both dirs" : "one dir  ", best,
               bytes / best / 1e6);
    };
    for (int bidir = 0; bidir < 2; ++bidir) {
        for (int ns : {1, 2, 4, 8}) run(0, ns, 0, bidir, "copy engines (cudaMemcpyAsync peer)");
        for (int nb : {32, 148, 592, 1184}) run(1, 1, nb, bidir, "SM push (local loads, peer stores, 16 B/lane)");
        for (int nb : {148, 592, 1184, 2368}) run(2, 1, nb, bidir, "SM pull (peer loads, local stores, 16 B/lane)");
    }
    return 0;
}
