// svb_post.cu — left-right check, 3x3 median (int16, replicate border) and filterSpeckles (GPU connected
// components by union-find).  Semantics: SURVEY.md §3.2.5 (second loop) and §3.2.6.  All HBM-bound, tiny.
#include "svb_common.cuh"

// ---------------- LR check ----------------
// dpre : int16 [H][W]  pre-check disparity (valid only for x in [minX1,maxX1))
// keys : u32   [H][W]  min-cost scatter keys  (minS<<16 | 0xFFFF-x_src), 0xFFFFFFFF = empty
__global__ void __launch_bounds__(256) k_lrcheck(const int16_t *__restrict__ dpre, const uint32_t *__restrict__ keys,
                                                 int16_t *__restrict__ out, int H, int W, int minX1, int maxX1, int minD,
                                                 int invalid16, int d12, size_t fstride_dpre, size_t fstride_keys,
                                                 size_t fstride_out)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    dpre += (size_t)b * fstride_dpre;
    keys += (size_t)b * fstride_keys;
    out += (size_t)b * fstride_out;
    const size_t row = (size_t)y * W;
    int d1 = invalid16;
    if (x >= minX1 && x < maxX1) {
        d1 = dpre[row + x];
        if (d1 != invalid16) {
            const int _d = d1 >> 4, d_ = (d1 + 15) >> 4;
            const int _x = x - _d, x_ = x - d_;
            bool bad = true;
            if (0 <= _x && _x < W) {
                uint32_t k = keys[row + _x];
                int d2 = (k == 0xFFFFFFFFu) ? invalid16 : (0xFFFF - (int)(k & 0xFFFFu)) - _x;
                bad = (d2 >= minD) && (abs(d2 - _d) > d12);
            } else
                bad = false;
            if (bad) {
                if (0 <= x_ && x_ < W) {
                    uint32_t k = keys[row + x_];
                    int d2 = (k == 0xFFFFFFFFu) ? invalid16 : (0xFFFF - (int)(k & 0xFFFFu)) - x_;
                    bad = (d2 >= minD) && (abs(d2 - d_) > d12);
                } else
                    bad = false;
            }
            if (bad) d1 = invalid16;
        }
    }
    out[row + x] = (int16_t)d1;
}

int svb_launch_lrcheck(const SgbmPlan &pl, char *ws, size_t ws_stride, int B, cudaStream_t st)
{
    dim3 grid((pl.W + 255) / 256, pl.H, B);
    k_lrcheck<<<grid, 256, 0, st>>>((const int16_t *)(ws + pl.off_dpre), (const uint32_t *)(ws + pl.off_key),
                                    (int16_t *)(ws + pl.off_dtmp), pl.H, pl.W, pl.minX1, pl.maxX1, pl.p.minDisparity,
                                    pl.invalid16, pl.p.disp12MaxDiff, ws_stride / 2, ws_stride / 4, ws_stride / 2);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

// ---------------- median 3x3 ----------------
__device__ __forceinline__ void cswap(int &a, int &b)
{
    int lo = min(a, b), hi = max(a, b);
    a = lo;
    b = hi;
}

__global__ void __launch_bounds__(256) k_median3(const int16_t *__restrict__ src, int16_t *__restrict__ dst, int H,
                                                 int W, size_t spitch, size_t sstride, size_t dpitch, size_t dstride)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    src += (size_t)b * sstride;
    const int xm = max(x - 1, 0), xp = min(x + 1, W - 1);
    const int16_t *r0 = src + (size_t)max(y - 1, 0) * spitch, *r1 = src + (size_t)y * spitch,
                  *r2 = src + (size_t)min(y + 1, H - 1) * spitch;
    int p0 = r0[xm], p1 = r0[x], p2 = r0[xp], p3 = r1[xm], p4 = r1[x], p5 = r1[xp], p6 = r2[xm], p7 = r2[x], p8 = r2[xp];
    // 19-comparator median-of-9 network
    cswap(p1, p2); cswap(p4, p5); cswap(p7, p8); cswap(p0, p1); cswap(p3, p4); cswap(p6, p7);
    cswap(p1, p2); cswap(p4, p5); cswap(p7, p8); cswap(p0, p3); cswap(p5, p8); cswap(p4, p7);
    cswap(p3, p6); cswap(p1, p4); cswap(p2, p5); cswap(p4, p7); cswap(p4, p2); cswap(p6, p4);
    cswap(p4, p2);
    dst[(size_t)b * dstride + (size_t)y * dpitch + x] = (int16_t)p4;
}

int svb_launch_median3(const int16_t *src, int16_t *dst, int H, int W, size_t spitch, size_t sstride, size_t dpitch,
                       size_t dstride, int B, cudaStream_t st)
{
    dim3 grid((W + 255) / 256, H, B);
    k_median3<<<grid, 256, 0, st>>>(src, dst, H, W, spitch, sstride, dpitch, dstride);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

extern "C" int svb_median3_s16(svb_context *ctx, const int16_t *src, int16_t *dst, int H, int W, size_t pitch_elems,
                               int B, size_t stride_elems, void *stream)
{
    SVB_REQUIRE(ctx && src && dst && src != dst, SVB_ERR_INVALID, "median3: null or aliased argument");
    SVB_REQUIRE(H > 0 && W > 0 && B > 0 && pitch_elems >= (size_t)W, SVB_ERR_INVALID, "median3: bad size");
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    return svb_launch_median3(src, dst, H, W, pitch_elems, stride_elems, pitch_elems, stride_elems, B,
                              (cudaStream_t)stream);
}

// ---------------- speckle filter: union-find connected components ----------------
__device__ __forceinline__ int uf_find(const int *__restrict__ L, int a)
{
    int p = L[a];
    while (p != a) {
        a = p;
        p = L[a];
    }
    return a;
}
__device__ __forceinline__ void uf_union(int *L, int a, int b)
{
    while (true) {
        a = uf_find(L, a);
        b = uf_find(L, b);
        if (a == b) return;
        if (a < b) { int t = a; a = b; b = t; }   // a > b: hook a under b
        int old = atomicMin(&L[a], b);
        if (old == a) return;
        a = old;
    }
}

// Phase 1: per-row runs.  label = index of the first pixel of the horizontal run inside the CTA's 256-pixel segment, or
// — for the run that continues from the previous segment — of that segment's last pixel, whose own label points
// further left: every label points left inside the same run, which is all the union-find of the later phases needs
// (chains are at most W/256 hops long).  One ballot per warp + a look-back over the CTA's 8 warps instead of a
// per-pixel backward walk.
__global__ void __launch_bounds__(256) k_ccl_init(const int16_t *__restrict__ img, int *__restrict__ L, int H, int W,
                                                  size_t pitch, size_t stride, int newVal, int maxDiff, size_t lstride)
{
    __shared__ int s_start[8];   // per warp: x of the last run start in its 32 pixels, -1 if all 32 continue a run
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y, b = blockIdx.z;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int16_t *row = img + (size_t)b * stride + (size_t)y * pitch;
    int *Lr = L + (size_t)b * lstride + (size_t)y * W;
    const bool in = x < W;
    const int v = in ? row[x] : newVal;
    int u = __shfl_up_sync(0xffffffffu, v, 1);
    if (lane == 0) u = (in && x > 0) ? row[x - 1] : newVal;
    const bool valid = in && v != newVal;
    const bool linked = valid && u != newVal && abs(u - v) <= maxDiff;     // same run as pixel x-1
    const unsigned brk = __ballot_sync(0xffffffffu, !linked);              // bit i: a run starts (or no run is) at lane i
    if (lane == 0) s_start[wid] = brk ? x + 31 - __clz(brk) : -1;
    __syncthreads();
    if (!in) return;
    if (!valid) { Lr[x] = -1; return; }
    const unsigned below = brk & (0xffffffffu >> (31 - lane));             // run starts at or below this lane
    int xs;
    if (below) xs = x - lane + (31 - __clz(below));
    else {   // the run continues from the pixels on the left: last start among the CTA's earlier warps, else the
             // previous CTA's last pixel
        xs = (int)(blockIdx.x * blockDim.x) - 1;
        for (int w = wid - 1; w >= 0; w--)
            if (s_start[w] >= 0) { xs = s_start[w]; break; }
    }
    Lr[x] = y * W + xs;
}

__global__ void __launch_bounds__(256) k_ccl_merge(const int16_t *__restrict__ img, int *__restrict__ L, int H, int W,
                                                   size_t pitch, size_t stride, int newVal, int maxDiff, size_t lstride)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const int16_t *im = img + (size_t)b * stride;
    int *Lb = L + (size_t)b * lstride;
    const int v = im[(size_t)y * pitch + x];
    if (v == newVal) return;
    const int i = y * W + x;
    // horizontal links are implicit in the run labels written by k_ccl_init (every label points left inside
    // the same run), so only vertical links need a union.
    if (y > 0) {
        int u = im[(size_t)(y - 1) * pitch + x];
        if (u != newVal && abs(u - v) <= maxDiff) {
            // skip if the left neighbour made the same vertical link (both pairs connected horizontally)
            bool dup = false;
            if (x > 0) {
                int vl = im[(size_t)y * pitch + x - 1], ul = im[(size_t)(y - 1) * pitch + x - 1];
                dup = vl != newVal && ul != newVal && abs(vl - v) <= maxDiff && abs(ul - u) <= maxDiff &&
                      abs(ul - vl) <= maxDiff;
            }
            if (!dup) uf_union(Lb, i, i - W);
        }
    }
}

__global__ void __launch_bounds__(256) k_ccl_count(int *__restrict__ L, int H, int W, size_t lstride)
{
    __shared__ int s_root[8], s_cnt[8];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.z;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int *Lb = L + (size_t)b * lstride;
    int *Sz = Lb + (size_t)H * W;
    int r = -1;
    if (i < H * W && Lb[i] >= 0) {
        r = uf_find(Lb, i);
        Lb[i] = r;   // benign race: only ever shortens chains towards the same root
    }
    // Two-level aggregated count: lanes that share a root elect one leader (large components would otherwise
    // serialise ~W*H same-address atomics in L2); the group of each warp's first labelled lane is then merged across
    // the CTA's 8 warps, so that a component covering the whole CTA costs one atomic per 256 pixels, not eight.
    const unsigned lab = __ballot_sync(0xffffffffu, r >= 0);
    const unsigned same = __match_any_sync(0xffffffffu, r);
    const int first = lab ? __ffs(lab) - 1 : 0;
    const int r0 = __shfl_sync(0xffffffffu, r, first);
    if (lane == 0) { s_root[wid] = lab ? r0 : -1; s_cnt[wid] = 0; }
    __syncwarp();
    if (r >= 0 && (int)(__ffs(same) - 1) == lane) {
        if (r == r0) s_cnt[wid] = __popc(same);
        else atomicAdd(&Sz[r], __popc(same));
    }
    __syncthreads();
    if (threadIdx.x < 8 && s_root[threadIdx.x] >= 0) {
        const int w = threadIdx.x, rw = s_root[w];
        bool lead = true;
        for (int k = 0; k < w; k++) lead = lead && s_root[k] != rw;
        if (lead) {
            int c = 0;
            for (int k = w; k < 8; k++) c += (s_root[k] == rw) ? s_cnt[k] : 0;
            atomicAdd(&Sz[rw], c);
        }
    }
}

__global__ void __launch_bounds__(256) k_ccl_paint(int16_t *__restrict__ img, const int *__restrict__ L, int H, int W,
                                                   size_t pitch, size_t stride, int newVal, int maxSize, size_t lstride)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const int *Lb = L + (size_t)b * lstride;
    const int *Sz = Lb + (size_t)H * W;
    const int i = y * W + x;
    int l = Lb[i];
    if (l < 0) return;
    int r = uf_find(Lb, l);
    if (Sz[r] <= maxSize) img[(size_t)b * stride + (size_t)y * pitch + x] = (int16_t)newVal;
}

int svb_launch_speckle(int16_t *img, int H, int W, size_t pitch, size_t stride, int B, int newVal, int maxSize,
                       int maxDiff, int32_t *labels, size_t lstride, cudaStream_t st)
{
    dim3 grid((W + 255) / 256, H, B);
    SVB_CUDA_CHECK(cudaMemset2DAsync(labels + (size_t)H * W, lstride * sizeof(int), 0, sizeof(int) * (size_t)H * W, B, st));
    k_ccl_init<<<grid, 256, 0, st>>>(img, labels, H, W, pitch, stride, newVal, maxDiff, lstride);
    k_ccl_merge<<<grid, 256, 0, st>>>(img, labels, H, W, pitch, stride, newVal, maxDiff, lstride);
    dim3 g1((H * W + 255) / 256, 1, B);
    k_ccl_count<<<g1, 256, 0, st>>>(labels, H, W, lstride);
    k_ccl_paint<<<grid, 256, 0, st>>>(img, labels, H, W, pitch, stride, newVal, maxSize, lstride);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

extern "C" int svb_filter_speckles_s16(svb_context *ctx, int16_t *img, int H, int W, size_t pitch_elems, int B,
                                       size_t stride_elems, int newVal, int maxSpeckleSize, int maxDiff,
                                       int32_t *labels_ws, size_t labels_ws_elems, void *stream)
{
    SVB_REQUIRE(ctx && img && labels_ws, SVB_ERR_INVALID, "filter_speckles: null argument");
    SVB_REQUIRE(H > 0 && W > 0 && B > 0 && pitch_elems >= (size_t)W, SVB_ERR_INVALID, "filter_speckles: bad size");
    SVB_REQUIRE((size_t)H * W < (size_t)1 << 30, SVB_ERR_INVALID, "filter_speckles: image too large");
    SVB_REQUIRE(labels_ws_elems >= (size_t)2 * H * W * B, SVB_ERR_WORKSPACE,
                "filter_speckles: labels workspace needs %zu int32 elements", (size_t)2 * H * W * B);
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    return svb_launch_speckle(img, H, W, pitch_elems, stride_elems, B, newVal, maxSpeckleSize, maxDiff, labels_ws,
                              (size_t)2 * H * W, (cudaStream_t)stream);
}

// ---- filterSpeckles without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops) ----
// The label / component-size scratch of the CCL kernels lives in a file-scope device array of fixed size: every GPU
// that loads libsvb.so holds SVB_NOALLOC_SPECKLE_MAX_PIXELS * 8 bytes (16 MiB) until the process exits.  The array's
// name is only a host-side stand-in here: the device address is resolved once per device, with that device current,
// by cudaGetSymbolAddress, and that pointer is what the kernels get.
static __device__ __align__(256) int32_t g_speckle_scratch[2 * (size_t)SVB_NOALLOC_SPECKLE_MAX_PIXELS];

#define SVB_MAX_DEVICES 64
static int32_t *g_speckle_scratch_ptr[SVB_MAX_DEVICES];   // per device, 0 = not resolved yet
#include <mutex>
static std::mutex g_speckle_scratch_mu;

static int speckle_scratch_for_device(int device, int32_t **out)
{
    SVB_REQUIRE(device >= 0 && device < SVB_MAX_DEVICES, SVB_ERR_INVALID, "filter_speckles_noalloc: device %d out of range", device);
    std::lock_guard<std::mutex> lk(g_speckle_scratch_mu);
    if (!g_speckle_scratch_ptr[device]) {
        int32_t *p = nullptr;
        SVB_CUDA_CHECK(cudaGetSymbolAddress((void **)&p, g_speckle_scratch));   // current device == `device` (caller)
        SVB_REQUIRE(p != nullptr, SVB_ERR_CUDA, "filter_speckles_noalloc: cudaGetSymbolAddress returned a null pointer");
        g_speckle_scratch_ptr[device] = p;
    }
    *out = g_speckle_scratch_ptr[device];
    return SVB_OK;
}

extern "C" int svb_filter_speckles_s16_noalloc(svb_context *ctx, int16_t *img, int H, int W, size_t pitch_elems, int B,
                                               size_t stride_elems, int newVal, int maxSpeckleSize, int maxDiff,
                                               void *stream)
{
    SVB_REQUIRE(ctx && img, SVB_ERR_INVALID, "filter_speckles_noalloc: null argument");
    SVB_REQUIRE(H > 0 && W > 0 && B > 0 && pitch_elems >= (size_t)W, SVB_ERR_INVALID, "filter_speckles_noalloc: bad size");
    SVB_REQUIRE((size_t)H * W * B <= (size_t)SVB_NOALLOC_SPECKLE_MAX_PIXELS, SVB_ERR_WORKSPACE,
                "filter_speckles_noalloc: B*H*W = %zu pixels exceeds the static scratch (SVB_NOALLOC_SPECKLE_MAX_PIXELS = %d); "
                "use svb_filter_speckles_s16 with a caller-owned workspace",
                (size_t)H * W * B, SVB_NOALLOC_SPECKLE_MAX_PIXELS);
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    int32_t *scratch;
    int rc = speckle_scratch_for_device(ctx->device, &scratch);
    if (rc) return rc;
    // svb_launch_speckle clears the component-size half of the scratch on the stream and k_ccl_init writes every
    // label, so nothing of the previous call's contents is read
    return svb_launch_speckle(img, H, W, pitch_elems, stride_elems, B, newVal, maxSpeckleSize, maxDiff, scratch,
                              (size_t)2 * H * W, (cudaStream_t)stream);
}
