// Issue rate of the integer instructions the Othello kernels are made of, alone and paired, on one GPU.
// Each thread runs ITER x 8 independent dependency chains of one instruction kind (inline PTX so that nvcc keeps the
// form), 8 CTAs of 256 threads per SM.  Output: warp-instructions per clock per SM for every kind.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/ubench/_bin/int_pipes scripts/ubench/int_pipes.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned u32; typedef unsigned long long u64;
#define ITER 512
enum { K_LOP3, K_SHF, K_IMAD, K_IMADHI, K_IMADWIDE, K_IADD3, K_MIX_LOP_IMAD, K_MIX_LOP_WIDE, K_MIX_LOP_HI, K_MIX_LOP_SHF, K_COUNT };
template <int K> __global__ void __launch_bounds__(256) probe(u32 *sink, u32 seed, u32 mul) {
    u32 a[8]; u64 w[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { a[j] = seed + threadIdx.x * 8 + j; w[j] = a[j]; }
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (K == K_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(seed), "r"(mul));
            if (K == K_SHF) asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(a[j]) : "r"(seed));
            if (K == K_IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(mul), "r"(seed));
            if (K == K_IMADHI) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[j]) : "r"(mul));
            if (K == K_IMADWIDE) asm volatile("{.reg .b32 lo, hi; mul.wide.u32 %0, %1, %2; mov.b64 {lo, hi}, %0; xor.b32 %1, lo, hi;}" : "+l"(w[j]), "+r"(a[j]) : "r"(mul));
            if (K == K_IADD3) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[j]) : "r"(seed));
            if (K == K_MIX_LOP_IMAD) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(seed), "r"(mul)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[(j + 4) & 7]) : "r"(mul), "r"(seed)); }
            if (K == K_MIX_LOP_WIDE) {      // 3 LOP3 (two independent, one folding the product) per IMAD.WIDE
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(seed), "r"(mul));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[(j + 3) & 7]) : "r"(seed), "r"(mul));
                asm volatile("{.reg .b32 lo, hi; mul.wide.u32 %0, %1, %2; mov.b64 {lo, hi}, %0; xor.b32 %1, lo, hi;}" : "+l"(w[j]), "+r"(a[(j + 5) & 7]) : "r"(mul));
            }
            if (K == K_MIX_LOP_HI) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(seed), "r"(mul)); asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[(j + 4) & 7]) : "r"(mul)); }
            if (K == K_MIX_LOP_SHF) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(seed), "r"(mul)); asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(a[(j + 4) & 7]) : "r"(seed)); }
        }
    }
    u32 x = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) x ^= a[j] ^ (u32)w[j] ^ (u32)(w[j] >> 32);
    if (x == 0x12345678u) sink[0] = x;
}
template <int K> static void run(const char *name, int per_iter, u32 *sink, int sms, double mhz) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<K><<<sms * 8, 256>>>(sink, 3, 0x9E3779B1u);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    probe<K><<<sms * 8, 256>>>(sink, 3, 0x9E3779B1u);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double warp_inst = (double)sms * 8 * 8 * ITER * 8 * per_iter;     // CTAs * warps * iterations * chains * instructions
    double clocks = ms * 1e-3 * mhz * 1e6;
    printf("%-22s %8.3f ms  %6.3f warp-inst / clk / SM   (%.3g thread-inst/s)\n", name, ms, warp_inst / clocks / sms, warp_inst * 32 / (ms * 1e-3));
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount; int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double mhz = khz / 1e3;
    printf("%s, %d SMs, %.0f MHz (nominal max; rates assume it)\n", p.name, sms, mhz);
    u32 *sink; cudaMalloc(&sink, 64);
    run<K_LOP3>("LOP3", 1, sink, sms, mhz);
    run<K_SHF>("SHF", 1, sink, sms, mhz);
    run<K_IADD3>("IADD3", 1, sink, sms, mhz);
    run<K_IMAD>("IMAD", 1, sink, sms, mhz);
    run<K_IMADHI>("IMAD.HI", 1, sink, sms, mhz);
    run<K_IMADWIDE>("IMAD.WIDE (+1 LOP3)", 2, sink, sms, mhz);
    run<K_MIX_LOP_IMAD>("LOP3 + IMAD", 2, sink, sms, mhz);
    run<K_MIX_LOP_WIDE>("3 LOP3 + IMAD.WIDE", 4, sink, sms, mhz);
    run<K_MIX_LOP_HI>("LOP3 + IMAD.HI", 2, sink, sms, mhz);
    run<K_MIX_LOP_SHF>("LOP3 + SHF", 2, sink, sms, mhz);
    return 0;
}
