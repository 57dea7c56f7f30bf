// Development aid: issue rate of the integer instructions the DP cell uses (warp-instructions per clock per SM).
#include <cstdio>
#include <cuda_runtime.h>
#define CHAINS 8
template <int OP>
__global__ void k(int iters, unsigned seed, unsigned* sink, long long* clk) {
    unsigned a[CHAINS];
    for (int i = 0; i < CHAINS; i++) a[i] = seed + threadIdx.x * 7 + i;
    unsigned kk = seed | 0x00030003u, c = seed * 3;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
            for (int i = 0; i < CHAINS; i++) {
                if (OP == 0) a[i] = __viaddmax_s32(a[i], kk, c);
                if (OP == 1) a[i] = __viaddmax_u16x2(a[i], kk, c);
                if (OP == 2) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(kk), "r"(c));
                if (OP == 3) asm volatile("mad.lo.u32 %0, %0, 3, %1;" : "+r"(a[i]) : "r"(c));
                if (OP == 4) a[i] = __vimax3_u16x2(a[i], kk, c + i);
                if (OP == 5) a[i] = __vimax3_s32(a[i], kk, c + i);
                if (OP == 6) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
            }
    }
    long long t1 = clock64();
    unsigned r = 0;
    for (int i = 0; i < CHAINS; i++) r ^= a[i];
    if (r == 0x12345678) sink[0] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
}
template <int OP> void run(const char* name) {
    unsigned* sink; long long* clk;
    cudaMalloc(&sink, 4); cudaMalloc(&clk, 8);
    int iters = 4096, blocks = 148, threads = 1024;
    k<OP><<<blocks, threads>>>(16, 1, sink, clk);
    k<OP><<<blocks, threads>>>(iters, 1, sink, clk);
    long long h; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    double winstr = (double)iters * 4 * CHAINS * (threads / 32);
    printf("%-22s %.3f warp-instr/clk/SM  (%.1f lanes/clk/SM)\n", name, winstr / h, winstr / h * 32);
}
int main() {
    run<0>("VIADDMNMX.S32"); run<1>("VIADDMNMX.U16x2"); run<2>("LOP3"); run<3>("IMAD"); run<4>("VIMNMX.U16x2"); run<5>("VIMNMX3.S32"); run<6>("IADD");
    return 0;
}
