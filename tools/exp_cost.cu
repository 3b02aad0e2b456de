// Timing harness for k_planes / k_cost_fused variants at a BASELINE geometry.  Not part of the product.
//   exp_cost [B] [H] [W] [D] [blockSize] [channels]
// Prints the best-of-N kernel time and a 64-bit hash of C (frame 0 and frame B-1), so that variants built with
// different -D switches can be compared for speed AND bit-identity in one gpurun call.
#include "../stereo_vision_b200/csrc/svb_api.cu"
#include "../stereo_vision_b200/csrc/svb_cost.cu"
#include <vector>
int svb_launch_aggregate(svb_context *, const SgbmPlan &, char *, size_t, int, int, cudaStream_t) { return 0; }
int svb_launch_lrcheck(const SgbmPlan &, char *, size_t, int, cudaStream_t) { return 0; }
int svb_launch_aggregate_concurrent(svb_context *, const SgbmPlan &, char *, size_t, int, cudaStream_t, cudaEvent_t) { return 0; }
int svb_launch_cost_aggregate_3way(svb_context *, const SgbmPlan &, char *, size_t, int, cudaStream_t) { return 0; }
int svb_launch_median3(const int16_t *, int16_t *, int, int, size_t, size_t, size_t, size_t, int, cudaStream_t) { return 0; }
int svb_launch_speckle(int16_t *, int, int, size_t, size_t, int, int, int, int, int32_t *, size_t, cudaStream_t) { return 0; }

__global__ void k_hash(const uint32_t *p, size_t n, unsigned long long *out)
{
    unsigned long long h = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        h += (unsigned long long)p[i] * (2654435761ull + 2 * (i % 1000003)) ^ (i * 0x9E3779B97F4A7C15ull);
    atomicAdd(out, h);
}

int main(int argc, char **argv)
{
    const int B = argc > 1 ? atoi(argv[1]) : 32, H = argc > 2 ? atoi(argv[2]) : 1080, W = argc > 3 ? atoi(argv[3]) : 1920;
    const int D = argc > 4 ? atoi(argv[4]) : 256, bs = argc > 5 ? atoi(argv[5]) : 5, ch = argc > 6 ? atoi(argv[6]) : 1;
    svb_sgbm_params p; memset(&p, 0, sizeof p);
    p.minDisparity = 0; p.numDisparities = D; p.blockSize = bs; p.P1 = 8 * ch * bs * bs; p.P2 = 32 * ch * bs * bs;
    p.disp12MaxDiff = 1; p.preFilterCap = 31; p.uniquenessRatio = 10; p.speckleWindowSize = 100; p.speckleRange = 2; p.mode = 1;
    SgbmPlan pl;
    if (svb_make_plan(&p, H, W, ch, B, 0, &pl)) { printf("plan: %s\n", svb_last_error()); return 1; }
    const size_t img = (size_t)H * W * ch;
    std::vector<uint8_t> hl(img * B), hr(img * B);
    uint32_t s = 12345;
    for (int b = 0; b < B; b++)
        for (size_t i = 0; i < img; i++) {
            s = s * 1664525u + 1013904223u;
            const size_t x = (i / ch) % W, y = (i / ch) / W;
            const int base = (int)(128 + 90 * sinf(0.013f * x + 0.02f * y + b) + 30 * sinf(0.21f * x));
            hl[b * img + i] = (uint8_t)min(255, max(0, base + (int)(s >> 28)));
            hr[b * img + i] = (uint8_t)min(255, max(0, (int)(128 + 90 * sinf(0.013f * (x + 37) + 0.02f * y + b) + 30 * sinf(0.21f * (x + 37))) + (int)((s >> 20) & 15)));
        }
    uint8_t *dl, *dr; char *ws; unsigned long long *dh;
    cudaMalloc(&dl, img * B); cudaMalloc(&dr, img * B); cudaMalloc(&dh, 16);
    if (cudaMalloc(&ws, pl.per_frame_bytes * B) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMemcpy(dl, hl.data(), img * B, cudaMemcpyHostToDevice);
    cudaMemcpy(dr, hr.data(), img * B, cudaMemcpyHostToDevice);
    cudaMemset(ws, 0, pl.per_frame_bytes * B);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best[2] = {1e9f, 1e9f};
    for (int rep = 0; rep < 5; rep++) {
        float ms;
        cudaEventRecord(e0);
        if (svb_launch_planes(pl, dl, dr, (size_t)W * ch, img, ws, pl.per_frame_bytes, B, 0)) { printf("planes: %s\n", svb_last_error()); return 1; }
        cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1); best[0] = fminf(best[0], ms);
        cudaEventRecord(e0);
        if (svb_launch_cost(pl, ws, pl.per_frame_bytes, B, 0, pl.H, 0)) { printf("cost: %s\n", svb_last_error()); return 1; }
        cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1); best[1] = fminf(best[1], ms);
    }
    const size_t vol = (size_t)H * pl.W1 * pl.Dp * 2;
    unsigned long long hh[2] = {0, 0};
    cudaMemset(dh, 0, 16);
    k_hash<<<1024, 256>>>((const uint32_t *)(ws + pl.off_C), vol / 4, dh);
    k_hash<<<1024, 256>>>((const uint32_t *)(ws + (size_t)(B - 1) * pl.per_frame_bytes + pl.off_C), vol / 4, dh + 1);
    cudaMemcpy(hh, dh, 16, cudaMemcpyDeviceToHost);
    const cudaError_t e = cudaDeviceSynchronize();
    printf("B=%d %dx%d D=%d bs=%d ch=%d: planes %.3f ms, cost %.3f ms (%.1f GB/s of C written), hash %016llx %016llx (%s)\n", B, W, H, D, bs,
           ch, best[0], best[1], vol * B / best[1] / 1e6, hh[0], hh[1], cudaGetErrorString(e));
    return 0;
}
