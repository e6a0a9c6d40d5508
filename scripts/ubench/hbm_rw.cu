// HBM bandwidth of one GPU by direction: read-only, write-only and copy over a 4 GiB buffer (>> the 126 MB L2).
// The step kernels write 5x what they read (state 32 B in; state 32 B + obs / mask / record 144 B out per env), so the
// write-only ceiling is the one their traffic meets, not only the copy figure of MEASURED_PEAKS.json.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/ubench/_bin/hbm_rw scripts/ubench/hbm_rw.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_write(uint4 *dst, size_t n, unsigned v) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, step = (size_t)gridDim.x * blockDim.x;
    const uint4 x = make_uint4(v, v + 1, v + 2, v + 3);
    for (; i < n; i += step) dst[i] = x;
}
__global__ void k_read(const uint4 *src, size_t n, unsigned *sink) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, step = (size_t)gridDim.x * blockDim.x;
    unsigned acc = 0;
    for (; i < n; i += step) { const uint4 x = src[i]; acc ^= x.x ^ x.y ^ x.z ^ x.w; }
    if (acc == 0x13572468u) sink[0] = acc;
}
__global__ void k_copy(const uint4 *src, uint4 *dst, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, step = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += step) dst[i] = src[i];
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const size_t bytes = 4ull << 30, n = bytes / 16;
    uint4 *a, *b; unsigned *sink;
    if (cudaMalloc(&a, bytes) != cudaSuccess || cudaMalloc(&b, bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMalloc(&sink, 64);
    cudaMemset(a, 1, bytes); cudaMemset(b, 2, bytes);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = p.multiProcessorCount * 16, block = 512;
    printf("%s: 4 GiB buffers, grid %d x %d, best of 5\n", p.name, grid, block);
    for (int kind = 0; kind < 4; kind++) {
        float best = 1e30f;
        for (int r = 0; r < 6; r++) {
            cudaEventRecord(e0);
            if (kind == 0) k_write<<<grid, block>>>(a, n, r);
            if (kind == 1) k_read<<<grid, block>>>(a, n, sink);
            if (kind == 2) k_copy<<<grid, block>>>(a, b, n);
            if (kind == 3) cudaMemsetAsync(a, r, bytes);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (r > 0 && ms < best) best = ms;
        }
        const double moved = kind == 2 ? 2.0 * bytes : 1.0 * bytes;
        const char *names[4] = {"write only (st.v4)", "read only (ld.v4)", "copy (read + write)", "cudaMemset"};
        printf("%-22s %8.3f ms  %7.1f GB/s\n", names[kind], best, moved / (best * 1e-3) / 1e9);
    }
    return 0;
}
