// svb_bench.cu — measurement helper: packed 16-bit integer ALU peak of the device (the roofline denominator for the
// aggregation stage, which MEASURED_PEAKS.json does not carry; SURVEY.md §8d asks the builder to measure it).
//
// Each thread runs 8 independent dependency chains of the two instructions the sweep kernel is made of
// (VIADDMNMX.S16x2 = __viaddmin_s16x2 and VIMNMX.S16x2 = __vmins2), all SMs, enough resident warps to cover the
// 4-cycle ALU latency.  Reported: packed instructions per second per thread-lane summed over the device; one packed
// instruction is two 16-bit lane operations.
#include "svb_common.cuh"

template <int MIX>
__global__ void __launch_bounds__(256) k_alu_peak(uint32_t *out, uint32_t a, uint32_t b, int iters)
{
    uint32_t x[8];
#pragma unroll
    for (int k = 0; k < 8; k++) x[k] = threadIdx.x * 0x00010001u + k;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (MIX == 0) x[k] = __viaddmin_s16x2(x[k], a, b);
                else if (MIX == 1) x[k] = __vmins2(x[k] ^ a, b);      // LOP3 + VIMNMX: both ALU pipe (counted as 2)
                else x[k] = (r & 1) ? __viaddmin_s16x2(x[k], a, b) : __vmaxs2(x[k], a);
            }
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s ^= x[k];
    if (s == 0x12345678u) out[0] = s;   // never true in practice; keeps the chains alive
}

extern "C" int svb_measure_packed_alu_peak(svb_context *ctx, double *packed_instr_per_s, double *sm_clock_mhz, void *stream)
{
    SVB_REQUIRE(ctx && packed_instr_per_s, SVB_ERR_INVALID, "measure_packed_alu_peak: null argument");
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t *d = nullptr;
    SVB_CUDA_CHECK(cudaMalloc(&d, 64));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 4096, grid = ctx->num_sms * 8;
    double best = 0;
    for (int mix = 0; mix < 3; mix++) {
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0, st);
            if (mix == 0) k_alu_peak<0><<<grid, 256, 0, st>>>(d, 0x00030003u, 0x7FFF7FFFu, iters);
            else if (mix == 1) k_alu_peak<1><<<grid, 256, 0, st>>>(d, 0x00030003u, 0x7FFF7FFFu, iters);
            else k_alu_peak<2><<<grid, 256, 0, st>>>(d, 0x00030003u, 0x7FFF7FFFu, iters);
            cudaEventRecord(e1, st);
            SVB_CUDA_CHECK(cudaEventSynchronize(e1));
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            const double per_thread = (double)iters * 4 * 8 * (mix == 1 ? 2 : 1);
            const double rate = per_thread * 256.0 * grid / (ms * 1e-3);
            if (rep > 0 && rate > best) best = rate;
        }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(d);
    *packed_instr_per_s = best;
    if (sm_clock_mhz) {
        int khz = 0;
        cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device);
        *sm_clock_mhz = khz / 1000.0;
    }
    return SVB_OK;
}
