// Integer-pipe throughput probes for sm_100a: how many warp-instructions per clock per SM each
// SASS instruction the profile DP uses (or could use) sustains, alone and mixed.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_int ubench_int.cu && ./ubench_int
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
#define CH 8

template <int OP>
__device__ __forceinline__ void op(int &a, int b, int c) {
    if (OP == 0) a = a + b;                                     // IADD / IMAD.IADD (compiler's choice)
    if (OP == 2) a = min(min(a, b), c + a);                     // VIMNMX3 + add
    if (OP == 3) a = __dp4a(b, c, a);                           // IDP.4A
    if (OP == 4) a = __funnelshift_l((unsigned)b, (unsigned)a, 1);  // SHF.L.W
    if (OP == 5) a = __viaddmin_s32(a, b, c);                   // VIADDMNMX
    if (OP == 6) a = __vimin3_s32(a, b, c);                     // VIMNMX3
    if (OP == 7) a = (int)(((unsigned)b >> 16) + (unsigned)a);  // LEA.HI
    if (OP == 8) a = a * b + c;                                 // IMAD
    if (OP == 9) a = __byte_perm(a, b, c);                      // PRMT
    if (OP == 10) a = (a & b) ^ c;                              // LOP3
    if (OP == 11) a = __umulhi((unsigned)a, (unsigned)b);       // IMAD.HI
    if (OP == 12) a = __viaddmin_s16x2(a, b, c);                // VIADDMNMX.S16x2
    if (OP == 13) a = __vimin3_s16x2(a, b, c);                  // VIMNMX3.S16x2
    if (OP == 14) a = __dp2a_lo(b, c, a);                       // IDP.2A
}

template <int OPA, int OPB, int NA, int NB>
__global__ void __launch_bounds__(256) probe(int *out, int b, int c, long long *clk) {
    int a[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) a[i] = threadIdx.x + i;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < NA; ++k)
#pragma unroll
            for (int i = 0; i < CH; ++i) op<OPA>(a[i], b, c);
#pragma unroll
        for (int k = 0; k < NB; ++k)
#pragma unroll
            for (int i = 0; i < CH; ++i) op<OPB>(a[i], b, c);
    }
    long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <int OPA, int OPB, int NA, int NB>
static void run(const char *name, int *d_out, long long *d_clk, int sms) {
    const int blocks = sms * 4, threads = 256;  // 8 warps per SMSP
    probe<OPA, OPB, NA, NB><<<blocks, threads>>>(d_out, 3, 5, d_clk);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<OPA, OPB, NA, NB><<<blocks, threads>>>(d_out, 3, 5, d_clk);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    long long clk = 0;
    cudaMemcpy(&clk, d_clk, sizeof(clk), cudaMemcpyDeviceToHost);
    const double winst = (double)ITERS * CH * (NA + NB) * (threads / 32) * 4;  // warp-instructions per SM
    printf("%-34s %8.3f ms  %6.3f warp-inst/clk/SM (by clock64)  %8.1f Gthread-inst/s\n", name, ms,
           winst / (double)clk, (double)ITERS * CH * (NA + NB) * blocks * threads / (ms * 1e-3) / 1e9);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int *d_out;
    long long *d_clk;
    cudaMalloc(&d_out, (size_t)p.multiProcessorCount * 4 * 256 * sizeof(int));
    cudaMalloc(&d_clk, sizeof(long long));
    printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
    const int sms = p.multiProcessorCount;
    run<0, 0, 1, 0>("add (a+b)", d_out, d_clk, sms);
    run<3, 3, 1, 0>("IDP.4A", d_out, d_clk, sms);
    run<14, 14, 1, 0>("IDP.2A", d_out, d_clk, sms);
    run<4, 4, 1, 0>("SHF.L.W", d_out, d_clk, sms);
    run<5, 5, 1, 0>("VIADDMNMX", d_out, d_clk, sms);
    run<6, 6, 1, 0>("VIMNMX3", d_out, d_clk, sms);
    run<7, 7, 1, 0>("LEA.HI", d_out, d_clk, sms);
    run<8, 8, 1, 0>("IMAD", d_out, d_clk, sms);
    run<9, 9, 1, 0>("PRMT", d_out, d_clk, sms);
    run<10, 10, 1, 0>("LOP3", d_out, d_clk, sms);
    run<11, 11, 1, 0>("IMAD.HI", d_out, d_clk, sms);
    run<12, 12, 1, 0>("VIADDMNMX.S16x2", d_out, d_clk, sms);
    run<13, 13, 1, 0>("VIMNMX3.S16x2", d_out, d_clk, sms);
    run<3, 4, 1, 1>("IDP.4A + SHF", d_out, d_clk, sms);
    run<3, 6, 1, 1>("IDP.4A + VIMNMX3", d_out, d_clk, sms);
    run<8, 6, 1, 1>("IMAD + VIMNMX3", d_out, d_clk, sms);
    run<8, 3, 1, 1>("IMAD + IDP.4A", d_out, d_clk, sms);
    run<6, 4, 1, 1>("VIMNMX3 + SHF", d_out, d_clk, sms);
    run<5, 8, 1, 1>("VIADDMNMX + IMAD", d_out, d_clk, sms);
    run<7, 8, 1, 1>("LEA.HI + IMAD", d_out, d_clk, sms);
    run<11, 6, 1, 1>("IMAD.HI + VIMNMX3", d_out, d_clk, sms);
    return 0;
}
