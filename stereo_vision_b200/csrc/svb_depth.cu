// svb_depth.cu — reprojectImageTo3D and the fused reprojection + per-box median depth
// (SURVEY.md §8a row a10).  f64 evaluation, f32 results.  HBM-bound: (2|4 + 12) B/px for reproject.
#include "svb_common.cuh"
#include <float.h>
#include <math.h>

struct QMat { double q[16]; };

__device__ __forceinline__ unsigned long long dkey(double v)
{
    unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double dkey_inv(unsigned long long k)
{
    unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}

template <typename T>
__global__ void __launch_bounds__(256) k_disp_min(const T *__restrict__ disp, size_t n, size_t stride,
                                                  unsigned long long *__restrict__ out)
{
    const T *d = disp + (size_t)blockIdx.y * stride;
    double m = DBL_MAX;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        m = fmin(m, (double)d[i]);
    for (int o = 16; o; o >>= 1) m = fmin(m, __shfl_xor_sync(0xffffffffu, m, o));
    __shared__ double sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < 8; i++) m = fmin(m, sm[i]);
        atomicMin(out + blockIdx.y, dkey(m));
    }
}

// ((a*x + b*y) + c*d) + e with separately rounded operations (no FMA contraction): the same f64 sequence as
// the CPU oracle, so the f32 results agree bit for bit with it (and within 1 ulp of cv2).
__device__ __forceinline__ double dot4(const double *q, double x, double y, double d)
{
    return __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(q[0], x), __dmul_rn(q[1], y)), __dmul_rn(q[2], d)), q[3]);
}

__device__ __forceinline__ void reproject_px(const QMat &Q, double x, double y, double d, float &X, float &Y, float &Z)
{
    const double *q = Q.q;
    double Xh = dot4(q + 0, x, y, d), Yh = dot4(q + 4, x, y, d), Zh = dot4(q + 8, x, y, d), Wh = dot4(q + 12, x, y, d);
    double iw = __ddiv_rn(1.0, Wh);
    X = (float)__dmul_rn(Xh, iw);
    Y = (float)__dmul_rn(Yh, iw);
    Z = (float)__dmul_rn(Zh, iw);
}

template <typename T>
__global__ void __launch_bounds__(256) k_reproject(const T *__restrict__ disp, int H, int W, QMat Q, int handle_missing,
                                                   const unsigned long long *__restrict__ minkey,
                                                   float *__restrict__ xyz)
{
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y = blockIdx.y, b = blockIdx.z;
    if (x0 >= W) return;
    const size_t base = ((size_t)b * H + y) * W;
    double minv = 0.0;
    if (handle_missing) minv = dkey_inv(minkey[b]);
    float o[12];
    const int n = min(4, W - x0);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (i < n) {
            double d = (double)disp[base + x0 + i];
            reproject_px(Q, (double)(x0 + i), (double)y, d, o[3 * i], o[3 * i + 1], o[3 * i + 2]);
            if (handle_missing && fabs(d - minv) <= (double)FLT_EPSILON) o[3 * i + 2] = 10000.0f;
        }
    }
    float *dst = xyz + (base + x0) * 3;
    if (n == 4 && (((uintptr_t)dst) & 15) == 0) {
        float4 *d4 = reinterpret_cast<float4 *>(dst);
        __stcs(d4 + 0, make_float4(o[0], o[1], o[2], o[3]));
        __stcs(d4 + 1, make_float4(o[4], o[5], o[6], o[7]));
        __stcs(d4 + 2, make_float4(o[8], o[9], o[10], o[11]));
    } else {
        for (int i = 0; i < 3 * n; i++) dst[i] = o[i];
    }
}

template <typename T>
static int reproject_typed(const T *disp, int H, int W, const QMat &Q, int hm, float *xyz, int B,
                           unsigned long long *minws, cudaStream_t st)
{
    if (hm) {
        SVB_CUDA_CHECK(cudaMemsetAsync(minws, 0xFF, sizeof(unsigned long long) * B, st));
        dim3 g(296, B);
        k_disp_min<T><<<g, 256, 0, st>>>(disp, (size_t)H * W, (size_t)H * W, minws);
    }
    dim3 grid(((W + 3) / 4 + 63) / 64, H, B);
    k_reproject<T><<<grid, 64, 0, st>>>(disp, H, W, Q, hm, minws, xyz);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

extern "C" int svb_reproject_to_3d(svb_context *ctx, const void *disp, int dtype, int H, int W, const double *Qh,
                                   int handle_missing, float *xyz, int B, double *min_ws, void *stream)
{
    SVB_REQUIRE(ctx && disp && Qh && xyz, SVB_ERR_INVALID, "reproject_to_3d: null argument");
    SVB_REQUIRE(dtype >= 0 && dtype <= 3, SVB_ERR_INVALID,
                "reproject_to_3d: disparity must be CV_8UC1, CV_16SC1, CV_32SC1 or CV_32FC1");
    SVB_REQUIRE(H > 0 && W > 0 && B > 0, SVB_ERR_INVALID, "reproject_to_3d: bad size");
    SVB_REQUIRE(!handle_missing || min_ws, SVB_ERR_INVALID, "reproject_to_3d: handle_missing needs min_ws");
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    QMat Q;
    memcpy(Q.q, Qh, sizeof Q.q);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long *mw = reinterpret_cast<unsigned long long *>(min_ws);
    switch (dtype) {
    case 0: return reproject_typed((const uint8_t *)disp, H, W, Q, handle_missing, xyz, B, mw, st);
    case 1: return reproject_typed((const int16_t *)disp, H, W, Q, handle_missing, xyz, B, mw, st);
    case 2: return reproject_typed((const int32_t *)disp, H, W, Q, handle_missing, xyz, B, mw, st);
    default: return reproject_typed((const float *)disp, H, W, Q, handle_missing, xyz, B, mw, st);
    }
}

// ---------------- fused reprojection + per-box median ----------------
__device__ __forceinline__ unsigned fkey(float f)
{
    unsigned b = __float_as_uint(f);
    return (b >> 31) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float fkey_inv(unsigned k)
{
    return __uint_as_float((k >> 31) ? (k & 0x7fffffffu) : ~k);
}

#define BOX_THREADS 512
#define BOX_WARPS (BOX_THREADS / 32)

struct BoxGeom { int x1, y1, bw, bh; };

// key of pixel i of the box, or false if invalid
__device__ __forceinline__ bool box_key(const int16_t *__restrict__ disp, int W, const QMat &Q, int invalid16,
                                        const BoxGeom &g, int i, unsigned &key)
{
    const int yy = g.y1 + i / g.bw, xx = g.x1 + i % g.bw;
    const int dv = disp[(size_t)yy * W + xx];
    if (dv == invalid16) return false;
    const double d = (double)((float)dv / 16.0f);
    const double *q = Q.q;
    const double Zh = dot4(q + 8, (double)xx, (double)yy, d), Wh = dot4(q + 12, (double)xx, (double)yy, d);
    const float z = (float)__dmul_rn(Zh, __ddiv_rn(1.0, Wh));
    if (!isfinite(z)) return false;
    key = fkey(z);
    return true;
}

__global__ void __launch_bounds__(BOX_THREADS) k_box_median(const int16_t *__restrict__ disp16, int H, int W,
                                                            size_t stride, QMat Q, int invalid16,
                                                            const float *__restrict__ boxes,
                                                            const int32_t *__restrict__ box_image,
                                                            float *__restrict__ out)
{
    __shared__ unsigned hist[BOX_WARPS][256];
    __shared__ unsigned tot[256];
    __shared__ unsigned s_prefix, s_rank, s_n, s_eq;
    const int b = blockIdx.x;
    const int tid = threadIdx.x, wid = tid >> 5;
    const int16_t *disp = disp16 + (size_t)(box_image ? box_image[b] : 0) * stride;
    BoxGeom g;
    {
        int x1 = (int)floorf(boxes[b * 4 + 0]), y1 = (int)floorf(boxes[b * 4 + 1]);
        int x2 = (int)ceilf(boxes[b * 4 + 2]), y2 = (int)ceilf(boxes[b * 4 + 3]);
        x1 = max(0, min(W, x1)); x2 = max(0, min(W, x2)); y1 = max(0, min(H, y1)); y2 = max(0, min(H, y2));
        g.x1 = x1; g.y1 = y1; g.bw = max(x2 - x1, 0); g.bh = max(y2 - y1, 0);
    }
    const int area = g.bw * g.bh;
    // Radix select, 8 bits per pass.  The first pass (top byte, no prefix) also counts the valid pixels; the lower
    // median comes out after four passes together with the number of keys below / equal to it, so the upper median
    // (even counts) is either the same value or the smallest key above it: one more pass instead of four.
    unsigned prefix = 0, mask = 0, rank = 0, n = 0, below = 0, eq = 0;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = tid; i < BOX_WARPS * 256; i += BOX_THREADS) (&hist[0][0])[i] = 0;
        __syncthreads();
        for (int i = tid; i < area; i += BOX_THREADS) {
            unsigned k;
            if (box_key(disp, W, Q, invalid16, g, i, k) && (k & mask) == prefix)
                atomicAdd(&hist[wid][(k >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (tid < 256) {
            unsigned t = 0;
#pragma unroll
            for (int w = 0; w < BOX_WARPS; w++) t += hist[w][tid];
            tot[tid] = t;
        }
        __syncthreads();
        if (tid == 0) {
            unsigned r = rank;
            if (shift == 24) {   // total count -> rank of the lower median
                unsigned c = 0;
                for (int bin = 0; bin < 256; bin++) c += tot[bin];
                s_n = c;
                r = c ? (c - 1) / 2 : 0;
            }
            unsigned cum = 0;
            int bin = 0;
            for (; bin < 255; bin++) {
                if (cum + tot[bin] > r) break;
                cum += tot[bin];
            }
            s_prefix = prefix | ((unsigned)bin << shift);
            s_rank = r - cum;
            s_eq = tot[bin];
        }
        __syncthreads();
        if (shift == 24) {
            n = s_n;
            if (n == 0) {
                if (tid == 0) out[b] = __int_as_float(0x7fc00000);
                return;
            }
            rank = (n - 1) / 2;
        }
        below += rank - s_rank;      // keys smaller than every key that still matches the prefix
        prefix = s_prefix;
        rank = s_rank;
        eq = s_eq;
        mask |= 255u << shift;
        __syncthreads();
    }
    const float lo = fkey_inv(prefix);
    float hi = lo;
    if ((n & 1u) == 0 && n / 2 >= below + eq) {   // the upper median is the next larger key
        if (tid == 0) s_prefix = 0xFFFFFFFFu;
        __syncthreads();
        unsigned m = 0xFFFFFFFFu;
        for (int i = tid; i < area; i += BOX_THREADS) {
            unsigned k;
            if (box_key(disp, W, Q, invalid16, g, i, k) && k > prefix) m = min(m, k);
        }
        m = __reduce_min_sync(0xffffffffu, m);
        if ((tid & 31) == 0) atomicMin(&s_prefix, m);
        __syncthreads();
        hi = fkey_inv(s_prefix);
    }
    if (tid == 0) out[b] = (n & 1u) ? lo : __fmul_rn(__fadd_rn(lo, hi), 0.5f);
}

extern "C" int svb_box_median_depth(svb_context *ctx, const int16_t *disp16, int H, int W, size_t stride_elems,
                                    const double *Qh, int invalid16, const float *boxes, const int32_t *box_image,
                                    int nboxes, float *out, void *stream)
{
    SVB_REQUIRE(ctx && disp16 && Qh && out, SVB_ERR_INVALID, "box_median_depth: null argument");
    SVB_REQUIRE(H > 0 && W > 0 && nboxes >= 0, SVB_ERR_INVALID, "box_median_depth: bad size");
    if (nboxes == 0) return SVB_OK;
    SVB_REQUIRE(boxes, SVB_ERR_INVALID, "box_median_depth: null boxes");
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    QMat Q;
    memcpy(Q.q, Qh, sizeof Q.q);
    k_box_median<<<nboxes, BOX_THREADS, 0, (cudaStream_t)stream>>>(disp16, H, W, stride_elems, Q, invalid16, boxes,
                                                                  box_image, out);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}
