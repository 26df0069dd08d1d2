// Microbenchmark: cost of __match_any_sync vs. a ballot-built peer mask (7 digit bits) on sm_100a.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o match_vs_ballot match_vs_ballot.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int iters, int bits, uint32_t mask) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t d = in[tid];
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
        uint32_t peers;
        if (MODE == 0) {
            peers = __match_any_sync(0xffffffffu, d);
        } else {
            peers = 0xffffffffu;
            for (int kbit = 0; kbit < bits; ++kbit) {
                const bool bit = (d >> kbit) & 1u;
                const uint32_t b = __ballot_sync(0xffffffffu, bit);
                peers &= bit ? b : ~b;
            }
        }
        acc += __popc(peers);
        d = (d * 1664525u + 1013904223u + (peers & 0u)) & mask;
    }
    out[tid] = acc;
}

int main() {
    const int blocks = 148 * 8, threads = 256, n = blocks * threads, iters = 2048;
    uint32_t *in, *out;
    cudaMalloc(&in, n * 4); cudaMalloc(&out, n * 4);
    uint32_t* h = new uint32_t[n];
    for (int i = 0; i < n; ++i) h[i] = (uint32_t)(i * 2654435761u >> 7) & 127u;
    cudaMemcpy(in, h, n * 4, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 8; ++mode) {
        const uint32_t mask = mode < 2 ? 127u : mode < 4 ? 7u : mode < 6 ? 1u : 0u;
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            if ((mode & 1) == 0) k<0><<<blocks, threads>>>(in, out, iters, 7, mask); else k<1><<<blocks, threads>>>(in, out, iters, 7, mask);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            const double warp_ops = (double)n / 32 * iters;
            if (rep) printf("mask %u %s: %.3f ms, %.2f ns per warp-op per SM (=> cycles at 1.965 GHz: %.1f)\n", mask, (mode & 1) ? "ballot x7" : "match_any", ms,
                            ms * 1e6 / (warp_ops / 148), ms * 1e6 / (warp_ops / 148) * 1.965);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
