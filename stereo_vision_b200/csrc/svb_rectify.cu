// svb_rectify.cu — initUndistortRectifyMap (f64) and fixed-point INTER_LINEAR remap (u8, C=1|3).
// Semantics: SURVEY.md §3.2.7 / §3.2.8 (cv2 4.13.0).  Both kernels are HBM-bound streaming kernels:
//   maps : writes 8 B/px                      remap: reads 8 B/px of maps + gathers C B/px, writes C B/px
#include "svb_common.cuh"
#include <math.h>

struct MapParams {
    double iR[9];
    double k[14];
    double mt[9];
    double fx, fy, cx, cy;
    int has_tilt;
};

__device__ __forceinline__ void map_uv(const MapParams &mp, int j, int i, double &u, double &v)
{
    const double *iR = mp.iR;
    const double *k = mp.k;
    // no FMA contraction: cv2 evaluates these as separate multiplies/adds
    double X = __dadd_rn(__dadd_rn(__dmul_rn(iR[0], (double)j), __dmul_rn(iR[1], (double)i)), iR[2]);
    double Y = __dadd_rn(__dadd_rn(__dmul_rn(iR[3], (double)j), __dmul_rn(iR[4], (double)i)), iR[5]);
    double Wc = __dadd_rn(__dadd_rn(__dmul_rn(iR[6], (double)j), __dmul_rn(iR[7], (double)i)), iR[8]);
    double w = __ddiv_rn(1.0, Wc);
    double x = __dmul_rn(X, w), y = __dmul_rn(Y, w);
    double x2 = __dmul_rn(x, x), y2 = __dmul_rn(y, y);
    double r2 = __dadd_rn(x2, y2);
    double _2xy = __dmul_rn(__dmul_rn(2.0, x), y);
    const double k1 = k[0], k2 = k[1], p1 = k[2], p2 = k[3], k3 = k[4], k4 = k[5], k5 = k[6], k6 = k[7];
    const double s1 = k[8], s2 = k[9], s3 = k[10], s4 = k[11];
    double num = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k3, r2), k2), r2), k1), r2));
    double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k6, r2), k5), r2), k4), r2));
    double kr = __ddiv_rn(num, den);
    double r4 = __dmul_rn(r2, r2);
    double xd = __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(x, kr), __dmul_rn(p1, _2xy)),
                                              __dmul_rn(p2, __dadd_rn(r2, __dmul_rn(2.0, x2)))),
                                    __dmul_rn(s1, r2)),
                          __dmul_rn(s2, r4));
    double yd = __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(y, kr), __dmul_rn(p1, __dadd_rn(r2, __dmul_rn(2.0, y2)))),
                                              __dmul_rn(p2, _2xy)),
                                    __dmul_rn(s3, r2)),
                          __dmul_rn(s4, r4));
    if (mp.has_tilt) {
        const double *mt = mp.mt;
        double vx = __dadd_rn(__dadd_rn(__dmul_rn(mt[0], xd), __dmul_rn(mt[1], yd)), mt[2]);
        double vy = __dadd_rn(__dadd_rn(__dmul_rn(mt[3], xd), __dmul_rn(mt[4], yd)), mt[5]);
        double vw = __dadd_rn(__dadd_rn(__dmul_rn(mt[6], xd), __dmul_rn(mt[7], yd)), mt[8]);
        double invp = vw != 0.0 ? __ddiv_rn(1.0, vw) : 1.0;
        u = __dadd_rn(__dmul_rn(__dmul_rn(mp.fx, invp), vx), mp.cx);
        v = __dadd_rn(__dmul_rn(__dmul_rn(mp.fy, invp), vy), mp.cy);
    } else {
        u = __dadd_rn(__dmul_rn(mp.fx, xd), mp.cx);
        v = __dadd_rn(__dmul_rn(mp.fy, yd), mp.cy);
    }
}

// cvRound(double) on x86 (cvtsd2si): round-half-even; NaN / out of range -> 0x80000000
__device__ __forceinline__ int cv_round_f64(double v) { return (fabs(v) < 2147483648.0) ? __double2int_rn(v) : (int)0x80000000; }

// FIXED = false: CV_32FC1 maps.  FIXED = true: CV_16SC2 + CV_16UC1 (integer part as short2, 5+5-bit fractions),
// rounded from the f64 coordinate as cv2 does — the form cv2.undistort() uses internally.
template <bool FIXED>
__global__ void __launch_bounds__(256) k_init_undistort_rectify_map(MapParams mp, int W, int H, void *__restrict__ m1,
                                                                   void *__restrict__ m2)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= W) return;
    double u, v;
    map_uv(mp, j, i, u, v);
    const size_t o = (size_t)i * W + j;
    if (FIXED) {
        const int iu = cv_round_f64(__dmul_rn(u, 32.0)), iv = cv_round_f64(__dmul_rn(v, 32.0));
        reinterpret_cast<short2 *>(m1)[o] = make_short2((short)(iu >> 5), (short)(iv >> 5));
        reinterpret_cast<uint16_t *>(m2)[o] = (uint16_t)((iv & 31) * 32 + (iu & 31));
    } else {
        reinterpret_cast<float *>(m1)[o] = (float)u;
        reinterpret_cast<float *>(m2)[o] = (float)v;
    }
}

static int inv3x3(const double *A, double *Ai)
{
    double M[3][6];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            M[i][j] = A[i * 3 + j];
            M[i][j + 3] = i == j;
        }
    for (int c = 0; c < 3; c++) {
        int piv = c;
        for (int r = c + 1; r < 3; r++)
            if (fabs(M[r][c]) > fabs(M[piv][c])) piv = r;
        if (M[piv][c] == 0.0) return -1;
        if (piv != c)
            for (int j = 0; j < 6; j++) {
                double t = M[c][j];
                M[c][j] = M[piv][j];
                M[piv][j] = t;
            }
        double d = M[c][c];
        for (int j = 0; j < 6; j++) M[c][j] /= d;
        for (int r = 0; r < 3; r++)
            if (r != c) {
                double f = M[r][c];
                for (int j = 0; j < 6; j++) M[r][j] -= f * M[c][j];
            }
    }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) Ai[i * 3 + j] = M[i][j + 3];
    return 0;
}

static int launch_maps(svb_context *ctx, const double *K, const double *dist, int ndist, const double *R,
                       const double *P33, int W, int H, void *mapx, void *mapy, bool fixed, void *stream)
{
    SVB_REQUIRE(ctx && K && mapx && mapy, SVB_ERR_INVALID, "init_undistort_rectify_map: null argument");
    SVB_REQUIRE(W > 0 && H > 0, SVB_ERR_INVALID, "init_undistort_rectify_map: bad size %dx%d", W, H);
    SVB_REQUIRE(ndist == 0 || ndist == 4 || ndist == 5 || ndist == 8 || ndist == 12 || ndist == 14, SVB_ERR_INVALID,
                "init_undistort_rectify_map: distCoeffs must have 4, 5, 8, 12 or 14 elements (got %d)", ndist);
    MapParams mp;
    memset(&mp, 0, sizeof mp);
    for (int i = 0; i < ndist; i++) mp.k[i] = dist[i];
    double Rm[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, A[9], PR[9];
    if (R) memcpy(Rm, R, sizeof Rm);
    memcpy(A, P33 ? P33 : K, sizeof A);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double s = 0;
            for (int t = 0; t < 3; t++) s += A[i * 3 + t] * Rm[t * 3 + j];
            PR[i * 3 + j] = s;
        }
    SVB_REQUIRE(inv3x3(PR, mp.iR) == 0, SVB_ERR_INVALID, "init_undistort_rectify_map: P*R is singular");
    mp.fx = K[0]; mp.fy = K[4]; mp.cx = K[2]; mp.cy = K[5];
    const double tx = mp.k[12], ty = mp.k[13];
    mp.has_tilt = (tx != 0.0 || ty != 0.0);
    double mt[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (mp.has_tilt) {
        double ctx_ = cos(tx), stx = sin(tx), cty = cos(ty), sty = sin(ty);
        double rot[9] = {cty, sty * stx, -sty * ctx_, 0, ctx_, stx, sty, -cty * stx, cty * ctx_};
        double pz[9] = {rot[8], 0, -rot[2], 0, rot[8], -rot[5], 0, 0, 1};
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                double s = 0;
                for (int t = 0; t < 3; t++) s += pz[i * 3 + t] * rot[t * 3 + j];
                mt[i * 3 + j] = s;
            }
    }
    memcpy(mp.mt, mt, sizeof mt);
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    dim3 grid((W + 255) / 256, H);
    if (fixed)
        k_init_undistort_rectify_map<true><<<grid, 256, 0, (cudaStream_t)stream>>>(mp, W, H, mapx, mapy);
    else
        k_init_undistort_rectify_map<false><<<grid, 256, 0, (cudaStream_t)stream>>>(mp, W, H, mapx, mapy);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

extern "C" int svb_init_undistort_rectify_map(svb_context *ctx, const double *K, const double *dist, int ndist,
                                              const double *R, const double *P33, int W, int H, float *mapx,
                                              float *mapy, void *stream)
{
    return launch_maps(ctx, K, dist, ndist, R, P33, W, H, mapx, mapy, false, stream);
}

extern "C" int svb_init_undistort_rectify_map_fixed(svb_context *ctx, const double *K, const double *dist, int ndist,
                                                    const double *R, const double *P33, int W, int H, int16_t *map1,
                                                    uint16_t *map2, void *stream)
{
    return launch_maps(ctx, K, dist, ndist, R, P33, W, H, map1, map2, true, stream);
}

// ---------------- remap ----------------
__device__ __forceinline__ int cv_round_x86(float v)
{
    // cvtps2dq: round-half-even; NaN / |v| >= 2^31 -> 0x80000000
    return (fabsf(v) < 2147483648.0f) ? __float2int_rn(v) : (int)0x80000000;
}
__device__ __forceinline__ int sat16(int v) { return max(-32768, min(32767, v)); }

template <int C, bool WORDS = false>
__device__ __forceinline__ void remap_fixed(const uint8_t *__restrict__ src, int sH, int sW, size_t spitch, int ix, int iy,
                                            int fx, int fy, uint8_t *o)
{
    const int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32, w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
    const bool y0 = (unsigned)iy < (unsigned)sH, y1 = (unsigned)(iy + 1) < (unsigned)sH;
    const bool x0 = (unsigned)ix < (unsigned)sW, x1 = (unsigned)(ix + 1) < (unsigned)sW;
    const uint8_t *r0 = src + (size_t)(y0 ? iy : 0) * spitch;
    const uint8_t *r1 = src + (size_t)(y1 ? iy + 1 : 0) * spitch;
    const int xa = (x0 ? ix : 0) * C, xb = (x1 ? ix + 1 : 0) * C;
    if (WORDS) {
        // interior pixel: the 2*C bytes of both taps of a row are consecutive -> aligned 32-bit loads + funnel shifts
        // instead of 2*C byte loads per row.  The caller guarantees 4-byte aligned rows; the last words of a row are
        // left to the byte path.
        constexpr int NWORD = C == 1 ? 2 : 3;
        const int a0 = xa & ~3, sh = (xa & 3) * 8;
        if (y0 && y1 && x0 && x1 && a0 + 4 * NWORD <= sW * C) {
            const uint32_t *p0 = reinterpret_cast<const uint32_t *>(r0 + a0);
            const uint32_t *p1 = reinterpret_cast<const uint32_t *>(r1 + a0);
            const uint32_t a = __ldg(p0), b = __ldg(p0 + 1), c2 = NWORD > 2 ? __ldg(p0 + 2) : 0u;
            const uint32_t d = __ldg(p1), e = __ldg(p1 + 1), f2 = NWORD > 2 ? __ldg(p1 + 2) : 0u;
            const uint32_t lo0 = __funnelshift_r(a, b, sh), hi0 = __funnelshift_r(b, c2, sh);   // bytes 0..3 / 4..7 from xa
            const uint32_t lo1 = __funnelshift_r(d, e, sh), hi1 = __funnelshift_r(e, f2, sh);
            const unsigned long long q0 = ((unsigned long long)hi0 << 32) | lo0, q1 = ((unsigned long long)hi1 << 32) | lo1;
#pragma unroll
            for (int c = 0; c < C; c++) {
                const int t00 = (int)((q0 >> (8 * c)) & 255u), t01 = (int)((q0 >> (8 * (c + C))) & 255u);
                const int t10 = (int)((q1 >> (8 * c)) & 255u), t11 = (int)((q1 >> (8 * (c + C))) & 255u);
                o[c] = (uint8_t)((w00 * t00 + w01 * t01 + w10 * t10 + w11 * t11 + 16384) >> 15);
            }
            return;
        }
    }
#pragma unroll
    for (int c = 0; c < C; c++) {
        int t00 = (y0 && x0) ? (int)__ldg(r0 + xa + c) : 0;
        int t01 = (y0 && x1) ? (int)__ldg(r0 + xb + c) : 0;
        int t10 = (y1 && x0) ? (int)__ldg(r1 + xa + c) : 0;
        int t11 = (y1 && x1) ? (int)__ldg(r1 + xb + c) : 0;
        o[c] = (uint8_t)((w00 * t00 + w01 * t01 + w10 * t10 + w11 * t11 + 16384) >> 15);
    }
}

template <int C, bool WORDS = false>
__device__ __forceinline__ void remap_px(const uint8_t *__restrict__ src, int sH, int sW, size_t spitch, float mx,
                                         float my, uint8_t *o)
{
    const int sx = cv_round_x86(mx * 32.0f), sy = cv_round_x86(my * 32.0f);
    remap_fixed<C, WORDS>(src, sH, sW, spitch, sat16(sx >> 5), sat16(sy >> 5), sx & 31, sy & 31, o);
}

// one thread = 4 consecutive destination pixels (vector map loads, word stores when aligned)
template <int C>
__global__ void __launch_bounds__(256) k_remap_u8(const uint8_t *__restrict__ src, int sH, int sW, size_t spitch,
                                                  size_t sstride, const float *__restrict__ mapx,
                                                  const float *__restrict__ mapy, uint8_t *__restrict__ dst, int dH,
                                                  int dW, size_t dpitch, size_t dstride, int vec_ok)
{
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y = blockIdx.y;
    const int b = blockIdx.z;
    if (x4 >= dW) return;
    src += (size_t)b * sstride;
    uint8_t *drow = dst + (size_t)b * dstride + (size_t)y * dpitch;
    const size_t mi = (size_t)y * dW + x4;
    if (vec_ok && x4 + 4 <= dW) {
        // default caching: the maps are shared by the B frames of a batch and stay L2-resident between them
        const float4 mx = __ldg(reinterpret_cast<const float4 *>(mapx + mi));
        const float4 my = __ldg(reinterpret_cast<const float4 *>(mapy + mi));
        __align__(4) uint8_t o[4 * C];
        if (vec_ok & 2) {   // source rows 4-byte aligned: word gathers
            remap_px<C, true>(src, sH, sW, spitch, mx.x, my.x, o + 0 * C);
            remap_px<C, true>(src, sH, sW, spitch, mx.y, my.y, o + 1 * C);
            remap_px<C, true>(src, sH, sW, spitch, mx.z, my.z, o + 2 * C);
            remap_px<C, true>(src, sH, sW, spitch, mx.w, my.w, o + 3 * C);
        } else {
            remap_px<C>(src, sH, sW, spitch, mx.x, my.x, o + 0 * C);
            remap_px<C>(src, sH, sW, spitch, mx.y, my.y, o + 1 * C);
            remap_px<C>(src, sH, sW, spitch, mx.z, my.z, o + 2 * C);
            remap_px<C>(src, sH, sW, spitch, mx.w, my.w, o + 3 * C);
        }
        uint32_t *dw = reinterpret_cast<uint32_t *>(drow + (size_t)x4 * C);
        const uint32_t *ow = reinterpret_cast<const uint32_t *>(o);
#pragma unroll
        for (int i = 0; i < C; i++) dw[i] = ow[i];
    } else {
        for (int x = x4; x < min(x4 + 4, dW); x++) {
            uint8_t o[C];
            remap_px<C>(src, sH, sW, spitch, mapx[(size_t)y * dW + x], mapy[(size_t)y * dW + x], o);
            for (int c = 0; c < C; c++) drow[(size_t)x * C + c] = o[c];
        }
    }
}

// (Measured and not adopted: an interleaved mapping — thread t of a 128-thread CTA gathers pixels t, t+128, t+256, t+384 of
// a 512-pixel run, results transposed through shared memory — so that one gather instruction of a warp touches 3-4
// cache lines instead of 12: 93.5 us per 8 x 1080p BGR frames against 85.3 us for the mapping below.)
extern "C" int svb_remap_u8(svb_context *ctx, const uint8_t *src, int sH, int sW, int C, size_t spitch, size_t sstride,
                            const float *mapx, const float *mapy, uint8_t *dst, int dH, int dW, size_t dpitch,
                            size_t dstride, int B, void *stream)
{
    SVB_REQUIRE(ctx && src && mapx && mapy && dst, SVB_ERR_INVALID, "remap: null argument");
    SVB_REQUIRE(C == 1 || C == 3, SVB_ERR_INVALID, "remap: only 1- or 3-channel u8 images (got C=%d)", C);
    SVB_REQUIRE(sH > 0 && sW > 0 && dH > 0 && dW > 0 && B > 0, SVB_ERR_INVALID, "remap: bad size");
    SVB_REQUIRE(sH < 32768 && sW < 32768, SVB_ERR_INVALID, "remap: source larger than SHRT_MAX (cv2 asserts too)");
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    int vec_ok = (dW % 4 == 0) && (((uintptr_t)mapx | (uintptr_t)mapy) % 16 == 0) && ((uintptr_t)dst % 4 == 0) &&
                 (dpitch % 4 == 0) && (dstride % 4 == 0);
    if (vec_ok && (uintptr_t)src % 4 == 0 && spitch % 4 == 0 && sstride % 4 == 0) vec_ok |= 2;   // word gathers
    dim3 grid(((dW + 3) / 4 + 127) / 128, dH, B);
    if (C == 1)
        k_remap_u8<1><<<grid, 128, 0, (cudaStream_t)stream>>>(src, sH, sW, spitch, sstride, mapx, mapy, dst, dH, dW,
                                                             dpitch, dstride, vec_ok);
    else
        k_remap_u8<3><<<grid, 128, 0, (cudaStream_t)stream>>>(src, sH, sW, spitch, sstride, mapx, mapy, dst, dH, dW,
                                                             dpitch, dstride, vec_ok);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

// remap with CV_16SC2 + CV_16UC1 fixed-point maps (one thread per destination pixel; the second half of cv2.undistort)
template <int C>
__global__ void __launch_bounds__(256) k_remap_u8_fixed(const uint8_t *__restrict__ src, int sH, int sW, size_t spitch,
                                                        size_t sstride, const short2 *__restrict__ map1,
                                                        const uint16_t *__restrict__ map2, uint8_t *__restrict__ dst,
                                                        int dW, size_t dpitch, size_t dstride)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= dW) return;
    const short2 xy = __ldg(map1 + (size_t)y * dW + x);
    const int f = __ldg(map2 + (size_t)y * dW + x) & 1023;
    uint8_t o[C];
    remap_fixed<C>(src + (size_t)blockIdx.z * sstride, sH, sW, spitch, xy.x, xy.y, f & 31, f >> 5, o);
    uint8_t *d = dst + (size_t)blockIdx.z * dstride + (size_t)y * dpitch + (size_t)x * C;
#pragma unroll
    for (int c = 0; c < C; c++) d[c] = o[c];
}

extern "C" int svb_remap_u8_fixed(svb_context *ctx, const uint8_t *src, int sH, int sW, int C, size_t spitch,
                                  size_t sstride, const int16_t *map1, const uint16_t *map2, uint8_t *dst, int dH, int dW,
                                  size_t dpitch, size_t dstride, int B, void *stream)
{
    SVB_REQUIRE(ctx && src && map1 && map2 && dst, SVB_ERR_INVALID, "remap: null argument");
    SVB_REQUIRE(C == 1 || C == 3, SVB_ERR_INVALID, "remap: only 1- or 3-channel u8 images (got C=%d)", C);
    SVB_REQUIRE(sH > 0 && sW > 0 && dH > 0 && dW > 0 && B > 0, SVB_ERR_INVALID, "remap: bad size");
    SVB_REQUIRE(sH < 32768 && sW < 32768, SVB_ERR_INVALID, "remap: source larger than SHRT_MAX (cv2 asserts too)");
    SVB_CUDA_CHECK(cudaSetDevice(ctx->device));
    dim3 grid((dW + 255) / 256, dH, B);
    const short2 *m1 = reinterpret_cast<const short2 *>(map1);
    if (C == 1)
        k_remap_u8_fixed<1><<<grid, 256, 0, (cudaStream_t)stream>>>(src, sH, sW, spitch, sstride, m1, map2, dst, dW, dpitch, dstride);
    else
        k_remap_u8_fixed<3><<<grid, 256, 0, (cudaStream_t)stream>>>(src, sH, sW, spitch, sstride, m1, map2, dst, dW, dpitch, dstride);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}
