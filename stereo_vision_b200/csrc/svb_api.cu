// svb_api.cu — context, parameter resolution, workspace plan and the StereoSGBM.compute entry points.
#include "svb_common.cuh"
#include <algorithm>

static thread_local char g_err[512] = "";

void svb_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

extern "C" const char *svb_last_error(void) { return g_err; }
extern "C" const char *svb_version(void) { return "svb 0.1 (sm_100a)"; }

extern "C" int svb_create(int device_id, svb_context **out)
{
    SVB_REQUIRE(out, SVB_ERR_INVALID, "svb_create: null out pointer");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        svb_set_error("svb_create: no CUDA device available (%s) — this engine has no CPU fallback",
                      e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        return SVB_ERR_CUDA;
    }
    SVB_REQUIRE(device_id >= 0 && device_id < n, SVB_ERR_INVALID, "svb_create: device %d out of range (%d devices)",
                device_id, n);
    SVB_CUDA_CHECK(cudaSetDevice(device_id));
    cudaDeviceProp prop;
    SVB_CUDA_CHECK(cudaGetDeviceProperties(&prop, device_id));
    SVB_REQUIRE(prop.major >= 10, SVB_ERR_UNSUPPORTED, "svb_create: device %s is sm_%d%d; libsvb is built for sm_100a only",
                prop.name, prop.major, prop.minor);
    svb_context *c = new svb_context();
    c->device = device_id;
    c->num_sms = prop.multiProcessorCount;
    c->ws = nullptr;
    c->ws_bytes = 0;
    c->tickets = nullptr;
    c->timing = 0;
    c->ev_valid = 0;
    c->job_valid = 0;
    c->job_stream = nullptr;
    c->latency_off = 0;
    cudaStreamCreateWithFlags(&c->aux, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->job_done, cudaEventDisableTiming);
    for (int i = 0; i <= SVB_NUM_STAGES; i++) cudaEventCreate(&c->ev[i]);
    e = cudaMalloc(&c->tickets, 16 * sizeof(unsigned));
    if (e != cudaSuccess) {
        delete c;
        svb_set_error("svb_create: cudaMalloc failed: %s", cudaGetErrorString(e));
        return SVB_ERR_CUDA;
    }
    *out = c;
    return SVB_OK;
}

extern "C" int svb_destroy(svb_context *ctx)
{
    if (!ctx) return SVB_OK;
    cudaSetDevice(ctx->device);
    if (ctx->ws) cudaFree(ctx->ws);
    if (ctx->tickets) cudaFree(ctx->tickets);
    for (int i = 0; i <= SVB_NUM_STAGES; i++) cudaEventDestroy(ctx->ev[i]);
    cudaEventDestroy(ctx->job_done);
    cudaEventDestroy(ctx->ev_fork);
    cudaEventDestroy(ctx->ev_join);
    cudaStreamDestroy(ctx->aux);
    delete ctx;
    return SVB_OK;
}

// SURVEY.md §3.2.1
extern "C" int svb_sgbm_resolve_params(const svb_sgbm_params *raw, svb_sgbm_params *e)
{
    SVB_REQUIRE(raw && e, SVB_ERR_INVALID, "resolve_params: null argument");
    *e = *raw;
    e->blockSize = raw->blockSize > 0 ? raw->blockSize : 5;
    e->P1 = raw->P1 > 0 ? raw->P1 : 2;
    e->P2 = std::max(raw->P2 > 0 ? raw->P2 : 5, e->P1 + 1);
    e->preFilterCap = std::max(raw->preFilterCap, 15) | 1;
    e->uniquenessRatio = raw->uniquenessRatio >= 0 ? raw->uniquenessRatio : 10;
    e->disp12MaxDiff = raw->disp12MaxDiff > 0 ? raw->disp12MaxDiff : 1;
    return SVB_OK;
}

static int next_pow2(int v)
{
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

int svb_make_plan(const svb_sgbm_params *raw, int H, int W, int ch, int B, int latency_ok, SgbmPlan *pl)
{
    SVB_REQUIRE(raw && pl, SVB_ERR_INVALID, "sgbm: null argument");
    memset(pl, 0, sizeof *pl);
    svb_sgbm_resolve_params(raw, &pl->p);
    const svb_sgbm_params &p = pl->p;
    SVB_REQUIRE(ch == 1 || ch == 3, SVB_ERR_INVALID, "sgbm: images must be CV_8UC1 or CV_8UC3 (channels=%d)", ch);
    SVB_REQUIRE(H > 0 && W > 0, SVB_ERR_INVALID, "sgbm: bad image size %dx%d", W, H);
    SVB_REQUIRE(p.numDisparities > 0, SVB_ERR_INVALID, "sgbm: numDisparities must be positive (got %d)", p.numDisparities);
    SVB_REQUIRE(p.mode >= 0 && p.mode <= 3, SVB_ERR_INVALID, "sgbm: unknown mode %d", p.mode);
    pl->H = H; pl->W = W; pl->ch = ch;
    const int minD = p.minDisparity, D = p.numDisparities, maxD = minD + D;
    pl->D = D;
    pl->minX1 = std::max(maxD, 0);
    pl->maxX1 = W + std::min(minD, 0);
    pl->W1 = pl->maxX1 - pl->minX1;
    pl->R = p.blockSize / 2;
    // cv2 4.13: "width - (minDisparity+numDisparities) > blockSize/2" else error -2
    SVB_REQUIRE(W - maxD > pl->R && pl->W1 > 0, SVB_ERR_INVALID,
                "sgbm: image width %d is too small for minDisparity+numDisparities=%d and blockSize=%d", W, maxD,
                p.blockSize);
    SVB_REQUIRE(pl->R <= 5, SVB_ERR_UNSUPPORTED, "sgbm: blockSize %d > 11 is outside this engine's contract", p.blockSize);
    SVB_REQUIRE(D <= 512, SVB_ERR_UNSUPPORTED, "sgbm: numDisparities %d > 512 is outside this engine's contract", D);
    SVB_REQUIRE(W <= 65535 && H <= 65535, SVB_ERR_UNSUPPORTED, "sgbm: image larger than 65535 px");
    if (p.mode == 2) {   // MODE_SGBM_3WAY: 4 row stripes; cv2 behaves differently (not modelled) when they are shorter than their overlap
        const int stripe = (H + 3) / 4, overlap = p.blockSize / 2 + 1 + (stripe + 9) / 10;
        SVB_REQUIRE(overlap <= stripe, SVB_ERR_UNSUPPORTED,
                    "sgbm: MODE_SGBM_3WAY needs row stripes (%d rows) at least as long as their overlap (%d): image too short", stripe, overlap);
    }
    pl->ftzero = p.preFilterCap;
    pl->invalid16 = (minD - 1) * 16;
    pl->nplanes = 2 * ch;
    const int bs = 2 * pl->R + 1;
    pl->max_cost = bs * bs * ch * (2 * pl->ftzero + 63);
    SVB_REQUIRE(pl->max_cost + p.P2 <= 32767 && p.P1 + p.P2 <= 32767, SVB_ERR_UNSUPPORTED,
                "sgbm: blockSize^2*channels*(2*preFilterCap+63)+P2 = %d exceeds int16 (cv2's cost type overflows too)",
                pl->max_cost + p.P2);
    pl->safe = (pl->max_cost + p.P2 <= 8191);
    if (D <= 256) {
        pl->E = 8;
        pl->G = std::max(2, next_pow2((D + 7) / 8));
    } else {
        pl->E = 16;
        pl->G = 32;
    }
    pl->Dp = pl->E * pl->G;
    pl->NG = pl->G == 32 ? (pl->E == 16 ? 7 : 15) : 16;   // rows per band (must match the k_sweep instantiation)
    pl->nbands = (H + pl->NG - 1) / pl->NG;
    // Latency form (SgbmPlan::latency): small jobs only.  At Dp = 64 / 128 it also switches to the wide-lane layout — a
    // whole warp per row with 2 / 4 disparities per lane instead of 8 / 16 lanes with 8: a single pair is bound by the
    // dependent instruction stream of one warp per scheduler, and a quarter of the per-lane work shortens it.
    pl->latency = 0;
#ifndef SVB_NO_LATENCY_FORM
    if (latency_ok && (p.mode == 0 || p.mode == 1)) {
        const bool wide = pl->Dp == 64 || pl->Dp == 128;
        const int ng = wide ? 15 : pl->NG, nb = (H + ng - 1) / ng;
        if ((long long)B * nb <= SVB_LATENCY_MAX_BANDS) {
            pl->latency = 1;
            if (wide) { pl->G = 32; pl->E = pl->Dp / 32; pl->NG = ng; pl->nbands = nb; }
        }
    }
#endif
    pl->Wr = (int)svb_align_up((size_t)W + pl->Dp + 8, 8);
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = svb_align_up(off + std::max<size_t>(bytes, 16), 256); return o; };
    const size_t vol = (size_t)H * pl->W1 * pl->Dp * 2;
    pl->off_lplanes = take((size_t)H * pl->nplanes * pl->W1 * 4);   // [H][channel][W1] uint2 (one word per plane)
    pl->off_rplanes = take((size_t)H * pl->nplanes * 24 * (pl->Wr + 8) * 2 + 64);   // [H][plane][3 arrays][Wr windows] uint4 (fits: 48*Wr <= 48*(Wr+8))
    // the sweep kernels address C and S as running pointers that are not clamped to the row: their ramp-up /
    // ramp-down prefetches touch (never use) up to SVB_SWEEP_MARGIN bytes either side of each volume
    take(SVB_SWEEP_MARGIN);
    pl->off_C = take(vol);
    take(SVB_SWEEP_MARGIN);
    pl->off_S = take(vol);
    take(SVB_SWEEP_MARGIN);
    const size_t edge_bytes = (size_t)std::max(pl->nbands - 1, 1) * pl->W1 * 3 * pl->Dp * 2;
    pl->off_edge = take(edge_bytes);
    if (pl->latency) {
        pl->off_S2 = take(vol);
        take(SVB_SWEEP_MARGIN);
        pl->off_edge2 = take(edge_bytes);
    }
    pl->off_prog = take(2 * sizeof(unsigned) * pl->nbands);
    pl->off_dpre = take((size_t)H * W * 2);
    pl->off_key = take((size_t)H * W * 4);
    pl->off_dtmp = take((size_t)H * W * 2);
    pl->off_labels = take((size_t)H * W * 8);
    pl->per_frame_bytes = svb_align_up(off, 256);
    return SVB_OK;
}

extern "C" int svb_sgbm_get_geometry(const svb_sgbm_params *raw, int W, int channels, svb_sgbm_geometry *g)
{
    SVB_REQUIRE(g, SVB_ERR_INVALID, "get_geometry: null argument");
    SgbmPlan pl;
    int rc = svb_make_plan(raw, 4096, W, channels, 1, 0, &pl);   // the height only matters to the workspace (and to mode 2's stripe check)
    if (rc) return rc;
    g->minX1 = pl.minX1; g->maxX1 = pl.maxX1; g->width1 = pl.W1; g->D = pl.D; g->Dp = pl.Dp;
    g->invalid16 = pl.invalid16; g->max_cost = pl.max_cost; g->safe_packed = pl.safe;
    return SVB_OK;
}

extern "C" int svb_sgbm_workspace_bytes(const svb_sgbm_params *raw, int H, int W, int channels, int B, size_t *bytes)
{
    SVB_REQUIRE(bytes && B > 0, SVB_ERR_INVALID, "workspace_bytes: bad argument");
    SgbmPlan pl;
    int rc = svb_make_plan(raw, H, W, channels, B, 1, &pl);   // (the latency form's layout is the larger one)
    if (rc) return rc;
    *bytes = pl.per_frame_bytes * (size_t)B;
    return SVB_OK;
}

static int get_workspace(svb_context *ctx, size_t need, void *user_ws, size_t user_bytes, char **out)
{
    if (user_ws) {
        SVB_REQUIRE(user_bytes >= need, SVB_ERR_WORKSPACE, "sgbm: workspace too small (%zu < %zu bytes)", user_bytes, need);
        SVB_REQUIRE(((uintptr_t)user_ws & 255) == 0, SVB_ERR_INVALID, "sgbm: workspace must be 256-byte aligned");
        *out = (char *)user_ws;
        return SVB_OK;
    }
    if (ctx->ws_bytes < need) {
        SVB_CUDA_CHECK(cudaDeviceSynchronize());
        if (ctx->ws) cudaFree(ctx->ws);
        ctx->ws = nullptr;
        ctx->ws_bytes = 0;
        SVB_CUDA_CHECK(cudaMalloc(&ctx->ws, need));
        ctx->ws_bytes = need;
    }
    *out = (char *)ctx->ws;
    return SVB_OK;
}

// Begin / end of a job that uses the context's tickets and (possibly) its cached workspace: see svb_context::mu.
static int job_begin(svb_context *ctx, cudaStream_t st)
{
    if (ctx->job_valid && ctx->job_stream != st) SVB_CUDA_CHECK(cudaStreamWaitEvent(st, ctx->job_done, 0));
    return SVB_OK;
}
static int job_end(svb_context *ctx, cudaStream_t st)
{
    SVB_CUDA_CHECK(cudaEventRecord(ctx->job_done, st));
    ctx->job_stream = st;
    ctx->job_valid = 1;
    return SVB_OK;
}

static int run_front(svb_context *ctx, const SgbmPlan &pl, const uint8_t *left, const uint8_t *right, size_t pitch,
                     size_t istride, int B, char *ws, cudaStream_t st)
{
    int rc;
    if ((rc = svb_launch_planes(pl, left, right, pitch, istride, ws, pl.per_frame_bytes, B, st))) return rc;
    if ((rc = svb_launch_cost(pl, ws, pl.per_frame_bytes, B, 0, pl.H, st))) return rc;
    return SVB_OK;
}

extern "C" int svb_sgbm_compute(svb_context *ctx, const svb_sgbm_params *raw, const uint8_t *left, const uint8_t *right,
                                int H, int W, int channels, size_t pitch_bytes, size_t image_stride_bytes, int B,
                                int16_t *disp, size_t disp_pitch_elems, size_t disp_stride_elems, void *ws_user,
                                size_t ws_bytes, void *stream)
{
    SVB_REQUIRE(ctx && left && right && disp, SVB_ERR_INVALID, "sgbm_compute: null argument");
    SVB_REQUIRE(B > 0, SVB_ERR_INVALID, "sgbm_compute: batch must be positive");
    SgbmPlan pl;
    int rc = svb_make_plan(raw, H, W, channels, B, !ctx->latency_off, &pl);
    if (rc) return rc;
    SVB_REQUIRE(pitch_bytes >= (size_t)W * channels && disp_pitch_elems >= (size_t)W, SVB_ERR_INVALID,
                "sgbm_compute: pitch smaller than a row");
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lock(ctx->mu);
    char *ws;
    if ((rc = get_workspace(ctx, pl.per_frame_bytes * (size_t)B, ws_user, ws_bytes, &ws))) return rc;
    if ((rc = job_begin(ctx, st))) return rc;
#define SVB_MARK(i) do { if (ctx->timing) cudaEventRecord(ctx->ev[i], st); } while (0)
    ctx->ev_valid = 0;
    SVB_MARK(0);
    if ((rc = svb_launch_planes(pl, left, right, pitch_bytes, image_stride_bytes, ws, pl.per_frame_bytes, B, st))) return rc;
    SVB_MARK(1);
    if (pl.p.mode == 2) {   // stripes: cost + both sweeps per stripe (the stage times report them all under pass 2)
        SVB_MARK(2);
        SVB_MARK(3);
        if ((rc = svb_launch_cost_aggregate_3way(ctx, pl, ws, pl.per_frame_bytes, B, st))) return rc;
    } else {
        if ((rc = svb_launch_cost(pl, ws, pl.per_frame_bytes, B, 0, pl.H, st))) return rc;
        SVB_MARK(2);
        if (pl.latency) {   // both sweeps side by side, then sum + selection (mark 3 = end of the forward sweep)
            if ((rc = svb_launch_aggregate_concurrent(ctx, pl, ws, pl.per_frame_bytes, B, st, ctx->timing ? ctx->ev[3] : nullptr)))
                return rc;
        } else {
            if ((rc = svb_launch_aggregate(ctx, pl, ws, pl.per_frame_bytes, B, 1, st))) return rc;
            SVB_MARK(3);
            if ((rc = svb_launch_aggregate(ctx, pl, ws, pl.per_frame_bytes, B, 2, st))) return rc;
        }
    }
    SVB_MARK(4);
    if ((rc = svb_launch_lrcheck(pl, ws, pl.per_frame_bytes, B, st))) return rc;
    if ((rc = svb_launch_median3((const int16_t *)(ws + pl.off_dtmp), disp, H, W, W, pl.per_frame_bytes / 2,
                                 disp_pitch_elems, disp_stride_elems, B, st)))
        return rc;
    SVB_MARK(5);
    if (pl.p.speckleWindowSize > 0)
        if ((rc = svb_launch_speckle(disp, H, W, disp_pitch_elems, disp_stride_elems, B, pl.invalid16,
                                     pl.p.speckleWindowSize, 16 * pl.p.speckleRange, (int32_t *)(ws + pl.off_labels),
                                     pl.per_frame_bytes / 4, st)))
            return rc;
    SVB_MARK(6);
    if (ctx->timing) ctx->ev_valid = 1;
    return job_end(ctx, st);
}

// Per-stage device times (ms) of the last svb_sgbm_compute on this context, measured with CUDA events recorded
// on the caller's stream: [planes, cost, aggregate pass 1, aggregate pass 2 + selection, LR check + median,
// speckle].  Blocks until that call has finished.  Enable with svb_set_timing(ctx, 1).
extern "C" int svb_set_latency_form(svb_context *ctx, int enable)
{
    SVB_REQUIRE(ctx, SVB_ERR_INVALID, "set_latency_form: null context");
    std::lock_guard<std::mutex> lock(ctx->mu);
    ctx->latency_off = !enable;
    return SVB_OK;
}
extern "C" int svb_set_timing(svb_context *ctx, int enable)
{
    SVB_REQUIRE(ctx, SVB_ERR_INVALID, "set_timing: null context");
    ctx->timing = enable;
    ctx->ev_valid = 0;
    return SVB_OK;
}
extern "C" int svb_sgbm_last_stage_ms(svb_context *ctx, float *ms, int n)
{
    SVB_REQUIRE(ctx && ms && n >= 6, SVB_ERR_INVALID, "last_stage_ms: need room for 6 floats");
    SVB_REQUIRE(ctx->ev_valid, SVB_ERR_INVALID, "last_stage_ms: no timed svb_sgbm_compute on this context");
    SVB_CUDA_CHECK(cudaEventSynchronize(ctx->ev[6]));
    for (int i = 0; i < 6; i++) SVB_CUDA_CHECK(cudaEventElapsedTime(&ms[i], ctx->ev[i], ctx->ev[i + 1]));
    return SVB_OK;
}

extern "C" int svb_sgbm_debug_stages(svb_context *ctx, const svb_sgbm_params *raw, const uint8_t *left,
                                     const uint8_t *right, int H, int W, int channels, size_t pitch_bytes, int16_t *Cvol,
                                     int16_t *Svol, int16_t *disp_raw, void *stream)
{
    SVB_REQUIRE(ctx && left && right, SVB_ERR_INVALID, "sgbm_debug_stages: null argument");
    SgbmPlan pl;
    int rc = svb_make_plan(raw, H, W, channels, 1, 0, &pl);   // the taps come from the back-to-back schedule
    if (rc) return rc;
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lock(ctx->mu);
    char *ws;
    if ((rc = get_workspace(ctx, pl.per_frame_bytes, nullptr, 0, &ws))) return rc;
    if ((rc = job_begin(ctx, st))) return rc;
    if ((rc = run_front(ctx, pl, left, right, pitch_bytes, 0, 1, ws, st))) return rc;
    const size_t vol = (size_t)H * pl.W1 * pl.Dp * 2;
    if (Cvol) SVB_CUDA_CHECK(cudaMemcpyAsync(Cvol, ws + pl.off_C, vol, cudaMemcpyDeviceToDevice, st));
    if (Svol || disp_raw) {
        // the volume holds S after pass 1 only (pass 2 keeps the final S in registers): that is what Svol taps; the
        // remaining directions are checked through disp_raw
        if ((rc = svb_launch_aggregate(ctx, pl, ws, pl.per_frame_bytes, 1, 3, st))) return rc;
        if (Svol) SVB_CUDA_CHECK(cudaMemcpyAsync(Svol, ws + pl.off_S, vol, cudaMemcpyDeviceToDevice, st));
        if (disp_raw) {
            if ((rc = svb_launch_lrcheck(pl, ws, pl.per_frame_bytes, 1, st))) return rc;
            SVB_CUDA_CHECK(cudaMemcpyAsync(disp_raw, ws + pl.off_dtmp, (size_t)H * W * 2, cudaMemcpyDeviceToDevice, st));
        }
    }
    return job_end(ctx, st);
}
