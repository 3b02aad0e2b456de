// svb_sparse.cu — SURVEY.md §8(f) rows 1, 3, 4: what the reference runs per detection, and the ingest step in front
// of the matcher.
//   k_undistort_points   <- cv2.undistortPoints(pts, K, dist, R, P)      ball_drop.py:239-240, 3dpose.py:538-549,
//                                                                         checkerboard_distance.py:83-86
//   k_triangulate_points <- cv2.triangulatePoints(P1, P2, pts1, pts2)     ball_drop.py:254, 3dpose.py:553,
//                                                                         checkerboard_distance.py:95
//   k_bgr2gray           <- cv2.cvtColor(img, COLOR_BGR2GRAY)             (host-side today)
// One thread per point, f64 throughout (the reference's own precision); batches of thousands of detections /
// landmarks per call are the intended use — a single point is launch-latency-bound like any GPU call.
#include "svb_common.cuh"

struct UndistortArgs {
    double k[14];
    double imt[9];  // inverse tilt (Scheimpflug) projection, identity when k[12] = k[13] = 0
    int has_tilt;
    double fx, fy, cx, cy;
    double RR[9];   // P[:, :3] * R
};

template <typename T>
__global__ void k_undistort_points(const T *__restrict__ src, T *__restrict__ dst, int n, const UndistortArgs a)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double x = ((double)src[2 * i] - a.cx) / a.fx, y = ((double)src[2 * i + 1] - a.cy) / a.fy;
    if (a.has_tilt) {   // compensate the tilt first, as cv2 does
        const double vx = a.imt[0] * x + a.imt[1] * y + a.imt[2], vy = a.imt[3] * x + a.imt[4] * y + a.imt[5];
        const double vz = a.imt[6] * x + a.imt[7] * y + a.imt[8];
        const double ip = vz != 0.0 ? 1. / vz : 1.;
        x = ip * vx; y = ip * vy;
    }
    const double x0 = x, y0 = y;
    const double *k = a.k;
#pragma unroll 1
    for (int it = 0; it < 5; it++) {   // cv2 default TermCriteria(COUNT, 5, 0.01): exactly five iterations
        const double r2 = x * x + y * y;
        const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
        if (icdist < 0) { x = x0; y = y0; break; }
        const double dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
        const double dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
        x = (x0 - dx) * icdist;
        y = (y0 - dy) * icdist;
    }
    const double X = a.RR[0] * x + a.RR[1] * y + a.RR[2], Y = a.RR[3] * x + a.RR[4] * y + a.RR[5];
    const double iw = 1. / (a.RR[6] * x + a.RR[7] * y + a.RR[8]);
    dst[2 * i] = (T)(X * iw);
    dst[2 * i + 1] = (T)(Y * iw);
}

struct TriArgs { double P1[12], P2[12]; };

// DLT: right singular vector of the smallest singular value of the 4x4 matrix, by one-sided Jacobi (Hestenes)
template <typename T>
__global__ void k_triangulate_points(const T *__restrict__ pa, const T *__restrict__ pb, size_t astride, size_t bstride,
                                     size_t apt, size_t bpt, T *__restrict__ out, int n, const TriArgs t)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double ax = pa[i * apt], ay = pa[i * apt + astride], bx = pb[i * bpt], by = pb[i * bpt + bstride];
    double A[4][4], V[4][4];
#pragma unroll
    for (int c = 0; c < 4; c++) {
        A[0][c] = ax * t.P1[8 + c] - t.P1[c];
        A[1][c] = ay * t.P1[8 + c] - t.P1[4 + c];
        A[2][c] = bx * t.P2[8 + c] - t.P2[c];
        A[3][c] = by * t.P2[8 + c] - t.P2[4 + c];
#pragma unroll
        for (int r = 0; r < 4; r++) V[r][c] = r == c ? 1. : 0.;
    }
#pragma unroll 1
    for (int sweep = 0; sweep < 30; sweep++) {
        bool rotated = false;
#pragma unroll
        for (int p = 0; p < 3; p++)
#pragma unroll
            for (int q = p + 1; q < 4; q++) {
                double al = 0, be = 0, ga = 0;
#pragma unroll
                for (int r = 0; r < 4; r++) { al += A[r][p] * A[r][p]; be += A[r][q] * A[r][q]; ga += A[r][p] * A[r][q]; }
                if (fabs(ga) <= 1e-15 * sqrt(al * be) || ga == 0) continue;
                rotated = true;
                const double zeta = (be - al) / (2 * ga);
                const double tt = (zeta >= 0 ? 1. : -1.) / (fabs(zeta) + sqrt(1 + zeta * zeta));
                const double cs = 1. / sqrt(1 + tt * tt), sn = cs * tt;
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    const double ap = A[r][p], aq = A[r][q];
                    A[r][p] = cs * ap - sn * aq; A[r][q] = sn * ap + cs * aq;
                    const double vp = V[r][p], vq = V[r][q];
                    V[r][p] = cs * vp - sn * vq; V[r][q] = sn * vp + cs * vq;
                }
            }
        if (!rotated) break;
    }
    double bn = 1e300, v0 = 0, v1 = 0, v2 = 0, v3 = 0;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        double s = 0;
#pragma unroll
        for (int r = 0; r < 4; r++) s += A[r][c] * A[r][c];
        if (s < bn) { bn = s; v0 = V[0][c]; v1 = V[1][c]; v2 = V[2][c]; v3 = V[3][c]; }
    }
    // 4 x N, row-major (cv2's output layout)
    out[i] = (T)v0; out[(size_t)n + i] = (T)v1; out[2 * (size_t)n + i] = (T)v2; out[3 * (size_t)n + i] = (T)v3;
}

__device__ __forceinline__ uint32_t gray15(uint32_t bl, uint32_t g, uint32_t r)
{
    return (bl * 3735u + g * 19235u + r * 9798u + 16384u) >> 15;   // cv2's 15-bit fixed point
}
// VEC: one thread = 4 pixels = three aligned 32-bit loads and one 32-bit store (rows 4-byte aligned, W % 4 == 0)
template <bool VEC>
__global__ void __launch_bounds__(256) k_bgr2gray(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, int H, int W,
                                                   size_t spitch, size_t sstride, size_t dpitch, size_t dstride)
{
    const int x = (blockIdx.x * blockDim.x + threadIdx.x) * (VEC ? 4 : 1), y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const uint8_t *p = src + (size_t)b * sstride + (size_t)y * spitch + 3 * x;
    uint8_t *q = dst + (size_t)b * dstride + (size_t)y * dpitch + x;
    if (VEC) {
        const uint32_t *pw = reinterpret_cast<const uint32_t *>(p);
        const uint32_t w0 = __ldcs(pw), w1 = __ldcs(pw + 1), w2 = __ldcs(pw + 2);   // B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
        const uint32_t g0 = gray15(w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u);
        const uint32_t g1 = gray15(w0 >> 24, w1 & 255u, (w1 >> 8) & 255u);
        const uint32_t g2 = gray15((w1 >> 16) & 255u, w1 >> 24, w2 & 255u);
        const uint32_t g3 = gray15((w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24);
        *reinterpret_cast<uint32_t *>(q) = g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
    } else {
        *q = (uint8_t)gray15(p[0], p[1], p[2]);
    }
}

extern "C" int svb_undistort_points(svb_context *ctx, const void *src, void *dst, int dtype, int n, const double *K,
                                    const double *dist, int ndist, const double *R, const double *P33, void *stream)
{
    SVB_REQUIRE(ctx && src && dst && K, SVB_ERR_INVALID, "undistort_points: null argument");
    SVB_REQUIRE(dtype == 3 || dtype == 4, SVB_ERR_INVALID, "undistort_points: points must be float32 (3) or float64 (4)");
    SVB_REQUIRE(n >= 0, SVB_ERR_INVALID, "undistort_points: negative point count");
    SVB_REQUIRE(ndist == 0 || ndist == 4 || ndist == 5 || ndist == 8 || ndist == 12 || ndist == 14, SVB_ERR_INVALID,
                "undistort_points: distCoeffs must have 4, 5, 8, 12 or 14 elements (got %d)", ndist);
    if (n == 0) return SVB_OK;
    UndistortArgs a;
    memset(&a, 0, sizeof a);
    for (int i = 0; i < ndist && i < 14; i++) a.k[i] = dist[i];
    a.fx = K[0]; a.fy = K[4]; a.cx = K[2]; a.cy = K[5];
    a.has_tilt = a.k[12] != 0.0 || a.k[13] != 0.0;
    if (a.has_tilt) {   // inverse of the tilt projection: matRotXY^T * invProjZ (same construction as the map kernel's)
        const double ctx_ = cos(a.k[12]), stx = sin(a.k[12]), cty = cos(a.k[13]), sty = sin(a.k[13]);
        const double rot[9] = {cty, sty * stx, -sty * ctx_, 0, ctx_, stx, sty, -cty * stx, cty * ctx_};
        const double inv = 1. / rot[8];
        const double ipz[9] = {inv, 0, inv * rot[2], 0, inv, inv * rot[5], 0, 0, 1};
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                double s = 0;
                for (int t = 0; t < 3; t++) s += rot[t * 3 + i] * ipz[t * 3 + j];
                a.imt[i * 3 + j] = s;
            }
    }
    double RR[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (R) memcpy(RR, R, sizeof RR);
    if (P33) {
        double t[9];
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) t[3 * r + c] = P33[3 * r] * RR[c] + P33[3 * r + 1] * RR[3 + c] + P33[3 * r + 2] * RR[6 + c];
        memcpy(RR, t, sizeof RR);
    }
    memcpy(a.RR, RR, sizeof RR);
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = (n + 127) / 128;
    if (dtype == 3) k_undistort_points<float><<<grid, 128, 0, st>>>((const float *)src, (float *)dst, n, a);
    else k_undistort_points<double><<<grid, 128, 0, st>>>((const double *)src, (double *)dst, n, a);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

extern "C" int svb_triangulate_points(svb_context *ctx, const double *P1, const double *P2, const void *pts1,
                                      const void *pts2, int dtype, int n, int layout, void *out4xn, void *stream)
{
    SVB_REQUIRE(ctx && P1 && P2 && pts1 && pts2 && out4xn, SVB_ERR_INVALID, "triangulate_points: null argument");
    SVB_REQUIRE(dtype == 3 || dtype == 4, SVB_ERR_INVALID, "triangulate_points: points must be float32 (3) or float64 (4)");
    SVB_REQUIRE(layout == 0 || layout == 1, SVB_ERR_INVALID, "triangulate_points: layout 0 = 2xN, 1 = Nx2");
    if (n <= 0) return n == 0 ? SVB_OK : SVB_ERR_INVALID;
    TriArgs t;
    memcpy(t.P1, P1, sizeof t.P1);
    memcpy(t.P2, P2, sizeof t.P2);
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = (n + 63) / 64;
    const size_t cstride = layout == 0 ? (size_t)n : 1, pt = layout == 0 ? 1 : 2;
    if (dtype == 3)
        k_triangulate_points<float><<<grid, 64, 0, st>>>((const float *)pts1, (const float *)pts2, cstride, cstride, pt, pt, (float *)out4xn, n, t);
    else
        k_triangulate_points<double><<<grid, 64, 0, st>>>((const double *)pts1, (const double *)pts2, cstride, cstride, pt, pt, (double *)out4xn, n, t);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

extern "C" int svb_bgr2gray_u8(svb_context *ctx, const uint8_t *src, uint8_t *dst, int H, int W, size_t spitch,
                               size_t sstride, size_t dpitch, size_t dstride, int B, void *stream)
{
    SVB_REQUIRE(ctx && src && dst, SVB_ERR_INVALID, "bgr2gray: null argument");
    SVB_REQUIRE(H > 0 && W > 0 && B > 0 && spitch >= (size_t)3 * W && dpitch >= (size_t)W, SVB_ERR_INVALID, "bgr2gray: bad size");
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    const bool vec = W % 4 == 0 && ((uintptr_t)src | (uintptr_t)dst | spitch | sstride | dpitch | dstride) % 4 == 0;
    if (vec) {
        dim3 grid((W / 4 + 127) / 128, H, B);
        k_bgr2gray<true><<<grid, 128, 0, (cudaStream_t)stream>>>(src, dst, H, W, spitch, sstride, dpitch, dstride);
    } else {
        dim3 grid((W + 255) / 256, H, B);
        k_bgr2gray<false><<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, H, W, spitch, sstride, dpitch, dstride);
    }
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}
