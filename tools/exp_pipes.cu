// Micro-benchmark: issue rate of candidate instructions on sm_100a (per SM, warp instructions per cycle).
// Not part of the product.  Each kernel runs NI independent chains per thread, 1024 threads per SM.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#define ITER 4096
template <int OP>
__global__ void __launch_bounds__(1024) k(uint32_t *out, uint32_t a, uint32_t b)
{
    uint32_t x[8];
    uint64_t w[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { x[i] = threadIdx.x * 7 + i; w[i] = x[i]; }
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) x[i] = x[i] * a + b;                                        // IMAD
            if (OP == 1) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"((uint32_t)w[i]), "r"(a));   // IMAD.WIDE
            if (OP == 2) x[i] = __byte_perm(x[i], b, 0x5432);                         // PRMT
            if (OP == 3) x[i] = __umulhi(x[i], a) + b;                                // IMAD.HI
            if (OP == 4) x[i] = __funnelshift_l(x[i], b, 16);                         // SHF
            if (OP == 5) x[i] = __vimin3_s16x2(x[i], a, b);                           // VIMNMX3
            if (OP == 6) x[i] = __viaddmin_s16x2(x[i], a, b);                         // VIADDMNMX
            if (OP == 7) { x[i] = x[i] * a + b; x[(i + 1) & 7] = __byte_perm(x[(i + 1) & 7], b, 0x5432); }   // IMAD + PRMT pair
            if (OP == 8) { asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"((uint32_t)w[i]), "r"(a)); x[i] = __viaddmin_s16x2(x[i], a, b); }  // WIDE + VIADDMNMX
            if (OP == 9) x[i] = (x[i] & a) | b;                                        // LOP3
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += x[i] + (uint32_t)w[i] + (uint32_t)(w[i] >> 32);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP>
void run(const char *name, int per_iter, uint32_t *out, int sms, double mhz)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP><<<sms, 1024>>>(out, 3, 5);
    cudaEventRecord(e0);
    k<OP><<<sms, 1024>>>(out, 3, 5);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double winstr = (double)ITER * 8 * per_iter * 32;   // warp instructions per SM (32 warps)
    printf("%-22s %.3f ms  -> %.2f warp-instr/cycle/SM (%.2f per SMSP)\n", name, ms, winstr / (ms * 1e-3 * mhz * 1e6), winstr / (ms * 1e-3 * mhz * 1e6) / 4);
}
int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    uint32_t *out; cudaMalloc(&out, p.multiProcessorCount * 1024 * 4);
    const double mhz = 1965.0;
    run<0>("IMAD", 1, out, p.multiProcessorCount, mhz);
    run<1>("IMAD.WIDE", 1, out, p.multiProcessorCount, mhz);
    run<2>("PRMT", 1, out, p.multiProcessorCount, mhz);
    run<3>("IMAD.HI (+add)", 1, out, p.multiProcessorCount, mhz);
    run<4>("SHF", 1, out, p.multiProcessorCount, mhz);
    run<5>("VIMNMX3", 1, out, p.multiProcessorCount, mhz);
    run<6>("VIADDMNMX", 1, out, p.multiProcessorCount, mhz);
    run<7>("IMAD + PRMT", 2, out, p.multiProcessorCount, mhz);
    run<8>("IMAD.WIDE + VIADDMNMX", 2, out, p.multiProcessorCount, mhz);
    run<9>("LOP3", 1, out, p.multiProcessorCount, mhz);
    return 0;
}
