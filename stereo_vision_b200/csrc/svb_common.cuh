// svb_common.cuh — shared host/device helpers for libsvb (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <mutex>

#include "../../include/svb.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libsvb is written for sm_100a (B200) only"
#endif

void svb_set_error(const char *fmt, ...);

#define SVB_CUDA_CHECK(expr)                                                                   \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            svb_set_error("CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__, __LINE__, \
                          cudaGetErrorString(_e));                                             \
            return SVB_ERR_CUDA;                                                               \
        }                                                                                      \
    } while (0)

#define SVB_REQUIRE(cond, code, ...)      \
    do {                                  \
        if (!(cond)) {                    \
            svb_set_error(__VA_ARGS__);   \
            return (code);                \
        }                                 \
    } while (0)

#define SVB_NUM_STAGES 7
struct svb_context {
    int device;
    int num_sms;
    void *ws;          // grow-only cached workspace
    size_t ws_bytes;
    unsigned int *tickets;  // device counters for the sweep scheduler
    // One StereoSGBM job owns the tickets / progress counters / workspace at a time.  Host threads are serialised by
    // `mu` while they enqueue; a job enqueued on a different stream than the previous one first waits (on the device)
    // for the event recorded behind the previous job, so overlapping calls are ordered instead of corrupting each other.
    std::mutex mu;
    cudaEvent_t job_done;
    cudaStream_t job_stream;
    int job_valid;
    int latency_off;        // svb_set_latency_form(ctx, 0): always run the two sweeps back to back
    cudaStream_t aux;       // second stream of the latency form (forked from / joined to the caller's stream by events)
    cudaEvent_t ev_fork, ev_join;
    int timing;             // record per-stage events in svb_sgbm_compute
    cudaEvent_t ev[SVB_NUM_STAGES + 1];
    int ev_valid;
};

// ---- resolved SGBM plan (host) ----
struct SgbmPlan {
    svb_sgbm_params p;   // effective parameters
    int H, W, ch;
    int minX1, maxX1, W1, D, Dp, E, G;  // E elems/lane, G lanes/group; Dp = E*G
    int R;               // block radius
    int ftzero;
    int invalid16;
    int max_cost;
    int safe;            // packed S accumulation cannot overflow
    int nplanes;
    // sweep geometry
    int NG;              // row-groups per CTA (rows per band)
    int nbands;
    // workspace layout (bytes, per frame)
    size_t off_lplanes, off_rplanes, off_C, off_S, off_edge, off_prog, off_dpre, off_key, off_dtmp, off_labels;
    // Latency form (few pairs in flight: B * nbands <= SVB_LATENCY_MAX_BANDS, modes 0 and 1): the reverse sweep runs
    // concurrently with the forward one on a second stream, writes its own sum S2 through its own band-edge buffer, and
    // k_select_sum adds the two volumes and selects.  Costs one more volume + edge buffer per pair, halves the wavefront
    // time a single pair waits for.
    int latency;
    size_t off_S2, off_edge2;
    size_t per_frame_bytes;
    int Wr;              // right-plane row length (elements) incl. padding
};

// unclamped prefetch reach of k_sweep around the C / S volumes: (lag_last + AGG_UNROLL + PFD) * Dp * 2 bytes <= 62 KB
#define SVB_SWEEP_MARGIN (64 << 10)
#define SVB_LATENCY_MAX_BANDS 74    // 2 x 74 bands: both sweeps of a job this small fit the GPU's 148 SMs side by side

// B: pairs processed together (decides the latency form: SgbmPlan::latency)
int svb_make_plan(const svb_sgbm_params *raw, int H, int W, int ch, int B, int latency_ok, SgbmPlan *plan);

static inline size_t svb_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// kernels' host launchers (defined in the .cu files)
int svb_launch_planes(const SgbmPlan &pl, const uint8_t *left, const uint8_t *right, size_t pitch, size_t istride,
                      char *ws, size_t ws_stride, int B, cudaStream_t st);
int svb_launch_cost(const SgbmPlan &pl, char *ws, size_t ws_stride, int B, int y0, int y1, cudaStream_t st);
// passes: bit 0 = pass 1 (4 forward paths), bit 1 = pass 2 (remaining paths + selection)
int svb_launch_aggregate(svb_context *ctx, const SgbmPlan &pl, char *ws, size_t ws_stride, int B, int passes,
                         cudaStream_t st);
// latency form: both sweeps concurrently (st + ctx->aux), then the sum + selection kernel
int svb_launch_aggregate_concurrent(svb_context *ctx, const SgbmPlan &pl, char *ws, size_t ws_stride, int B, cudaStream_t st,
                                    cudaEvent_t mark_pass1);
// MODE_SGBM_3WAY: cost volume + aggregation + selection, stripe by stripe
int svb_launch_cost_aggregate_3way(svb_context *ctx, const SgbmPlan &pl, char *ws, size_t ws_stride, int B, cudaStream_t st);
int svb_launch_lrcheck(const SgbmPlan &pl, char *ws, size_t ws_stride, int B, cudaStream_t st);
int svb_launch_median3(const int16_t *src, int16_t *dst, int H, int W, size_t spitch, size_t sstride, size_t dpitch,
                       size_t dstride, int B, cudaStream_t st);
int svb_launch_speckle(int16_t *img, int H, int W, size_t pitch, size_t stride, int B, int newVal, int maxSize,
                       int maxDiff, int32_t *labels, size_t lstride, cudaStream_t st);
