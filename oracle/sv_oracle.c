/*
 * sv_oracle.c — CPU ORACLE (test infrastructure, NOT product code).
 *
 * Plain-C restatement of the dense stereo hot path the reference scripts drive
 * through OpenCV (cv2 4.13.0, un-vendored third-party dependency of
 * jacksonroah/stereo-vision; reference call sites:
 *   stereo_calibration/scripts/checkerboard_distance.py:49-57,71-72  (stereoRectify,
 *       initUndistortRectifyMap, remap)
 *   calibration_tests/calibrate_v4/distance.py:130-157               (rectify_images)
 *   calibration_tests/calibrate_v4/extrinsic.py:312-324,371-378      (Q, maps, remap)
 *   stereo_calibration/scripts/ball_drop.py:89-160,225-259           (box -> 3-D point)
 * ).  The algorithm itself lives in OpenCV's calib3d/imgproc modules whose source is
 * not on this box; every function below follows the bit-exact specification that
 * SURVEY.md §3.2 recovered by probing the cv2 4.13.0 binary, and is PINNED by
 * tests/test_oracle_vs_golden.py against golden vectors generated from that binary
 * (tests/golden/make_golden.py).  The reference repo itself holds no golden vectors for
 * this path — so parity is pinned to cv2 4.13.0 outputs, not to
 * reference-owned fixtures.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product path (stereo_vision_b200) never does.
 *
 * Build: see oracle/Makefile  (gcc -O3 -fopenmp -shared -fPIC).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define SVO_MAX_COST 32767
#define SVO_DISP_SHIFT 4
#define SVO_DISP_SCALE 16

typedef struct {
    int minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff, preFilterCap,
        uniquenessRatio, speckleWindowSize, speckleRange, mode;
} svo_params;

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* SURVEY.md §3.2.1 — StereoSGBM_create parameter resolution (effective values). */
void svo_resolve_params(const svo_params *in, svo_params *out)
{
    *out = *in;
    out->blockSize = in->blockSize > 0 ? in->blockSize : 5;
    out->P1 = in->P1 > 0 ? in->P1 : 2;
    out->P2 = imax(in->P2 > 0 ? in->P2 : 5, out->P1 + 1);
    out->preFilterCap = imax(in->preFilterCap, 15) | 1; /* ftzero */
    out->uniquenessRatio = in->uniquenessRatio >= 0 ? in->uniquenessRatio : 10;
    out->disp12MaxDiff = in->disp12MaxDiff > 0 ? in->disp12MaxDiff : 1;
}

/* ---- stage 1 helpers: gradient / intensity planes ---- */
static void build_planes(const uint8_t *img, int H, int W, int C, size_t pitch, int ch, int ftzero,
                         uint8_t *grad, uint8_t *inten)
{
    for (int y = 0; y < H; y++) {
        const uint8_t *rc = img + (size_t)y * pitch;
        const uint8_t *rn = img + (size_t)imax(y - 1, 0) * pitch;
        const uint8_t *rs = img + (size_t)imin(y + 1, H - 1) * pitch;
        uint8_t *g = grad + (size_t)y * W, *it = inten + (size_t)y * W;
        g[0] = g[W - 1] = (uint8_t)ftzero;
        it[0] = it[W - 1] = (uint8_t)ftzero;
        for (int x = 1; x < W - 1; x++) {
            int gv = 2 * ((int)rc[(x + 1) * C + ch] - (int)rc[(x - 1) * C + ch]) +
                     ((int)rn[(x + 1) * C + ch] - (int)rn[(x - 1) * C + ch]) +
                     ((int)rs[(x + 1) * C + ch] - (int)rs[(x - 1) * C + ch]);
            g[x] = (uint8_t)(iclamp(gv, -ftzero, ftzero) + ftzero);
            it[x] = rc[x * C + ch];
        }
    }
}

/* p, p0=min(p,pl,pr), p1=max(p,pl,pr) with half-pixel interpolants (integer floor) */
static void bt_minmax(const uint8_t *p, int W, uint8_t *lo, uint8_t *hi)
{
    for (int x = 0; x < W; x++) {
        int v = p[x];
        int pl = x > 0 ? (v + p[x - 1]) / 2 : v;
        int pr = x < W - 1 ? (v + p[x + 1]) / 2 : v;
        lo[x] = (uint8_t)imin(v, imin(pl, pr));
        hi[x] = (uint8_t)imax(v, imax(pl, pr));
    }
}

/*
 * Cost volume C[y][xi][d] (int16), xi in [0,width1), d in [0,D)   — SURVEY.md §3.2.2-3.2.3.
 * Caller allocates H*width1*D int16.  Returns 0, or <0 on bad geometry.
 */
static int cost_volume_rows(const uint8_t *left, const uint8_t *right, int H, int W, int C, size_t pitch,
                            const svo_params *praw, int y0, int y1, int16_t *Cvol);
int svo_sgbm_cost_volume(const uint8_t *left, const uint8_t *right, int H, int W, int C, size_t pitch,
                         const svo_params *praw, int16_t *Cvol)
{
    return cost_volume_rows(left, right, H, W, C, pitch, praw, 0, H, Cvol);
}
/* Rows [y0, y1) of the cost volume (stored from row 0 of Cvol).  The block window is clamped to row y0 at the top
 * and to the IMAGE's last row at the bottom: this is how MODE_SGBM_3WAY runs every stripe (probed on cv2 4.13.0);
 * with y0 = 0, y1 = H it is the ordinary volume.  The Sobel / intensity planes always come from the whole image. */
static int cost_volume_rows(const uint8_t *left, const uint8_t *right, int H, int W, int C, size_t pitch,
                            const svo_params *praw, int y0, int y1, int16_t *Cvol)
{
    svo_params p;
    svo_resolve_params(praw, &p);
    const int minD = p.minDisparity, D = p.numDisparities, maxD = minD + D;
    const int minX1 = imax(maxD, 0), maxX1 = W + imin(minD, 0), width1 = maxX1 - minX1;
    const int ftzero = p.preFilterCap;
    const int r = p.blockSize / 2;
    if (D <= 0 || width1 <= 0) return -1;
    const int nplanes = 2 * C;
    size_t plane_sz = (size_t)H * W;
    uint8_t *Lp = malloc(plane_sz * nplanes * 3), *Rp = malloc(plane_sz * nplanes * 3);
    if (!Lp || !Rp) { free(Lp); free(Rp); return -2; }
    /* layout: [plane][3: val, lo, hi][H][W] */
    for (int ch = 0; ch < C; ch++) {
        for (int side = 0; side < 2; side++) {
            const uint8_t *img = side ? right : left;
            uint8_t *base = side ? Rp : Lp;
            uint8_t *g = base + (size_t)(2 * ch) * 3 * plane_sz;
            uint8_t *it = base + (size_t)(2 * ch + 1) * 3 * plane_sz;
            build_planes(img, H, W, C, pitch, ch, ftzero, g, it);
            for (int y = 0; y < H; y++) {
                bt_minmax(g + (size_t)y * W, W, g + plane_sz + (size_t)y * W, g + 2 * plane_sz + (size_t)y * W);
                bt_minmax(it + (size_t)y * W, W, it + plane_sz + (size_t)y * W, it + 2 * plane_sz + (size_t)y * W);
            }
        }
    }
    const int b0 = y0, b1 = H;   /* box clamp range */
    const int py0 = imax(b0, y0 - r), py1 = imin(b1, y1 + r);
    size_t rowsz = (size_t)width1 * D;
    int16_t *pix = malloc(sizeof(int16_t) * rowsz * H);
    int16_t *hs = malloc(sizeof(int16_t) * rowsz * H);
    if (!pix || !hs) { free(Lp); free(Rp); free(pix); free(hs); return -2; }
#pragma omp parallel for schedule(static)
    for (int y = py0; y < py1; y++) {
        int16_t *prow = pix + (size_t)y * rowsz;
        memset(prow, 0, sizeof(int16_t) * rowsz);
        for (int pl = 0; pl < nplanes; pl++) {
            const int shift = (pl & 1) ? 2 : 0;
            const uint8_t *lu = Lp + (size_t)pl * 3 * plane_sz + (size_t)y * W;
            const uint8_t *lu0 = lu + plane_sz, *lu1 = lu + 2 * plane_sz;
            const uint8_t *rv = Rp + (size_t)pl * 3 * plane_sz + (size_t)y * W;
            const uint8_t *rv0 = rv + plane_sz, *rv1 = rv + 2 * plane_sz;
            for (int xi = 0; xi < width1; xi++) {
                const int x = xi + minX1;
                const int u = lu[x], u0 = lu0[x], u1 = lu1[x];
                int16_t *o = prow + (size_t)xi * D;
                for (int d = 0; d < D; d++) {
                    const int xr = x - (d + minD);
                    const int v = rv[xr], v0 = rv0[xr], v1 = rv1[xr];
                    const int c0 = imax(0, imax(u - v1, v0 - u));
                    const int c1 = imax(0, imax(v - u1, u0 - v));
                    o[d] = (int16_t)(o[d] + (imin(c0, c1) >> shift));
                }
            }
        }
    }
    /* horizontal box sum with clamp in cost space */
#pragma omp parallel for schedule(static)
    for (int y = py0; y < py1; y++) {
        const int16_t *prow = pix + (size_t)y * rowsz;
        int16_t *hrow = hs + (size_t)y * rowsz;
        for (int xi = 0; xi < width1; xi++)
            for (int d = 0; d < D; d++) {
                int s = 0;
                for (int dx = -r; dx <= r; dx++) s += prow[(size_t)iclamp(xi + dx, 0, width1 - 1) * D + d];
                hrow[(size_t)xi * D + d] = (int16_t)s;
            }
    }
    /* vertical box sum with clamp */
#pragma omp parallel for schedule(static)
    for (int y = y0; y < y1; y++) {
        int16_t *crow = Cvol + (size_t)(y - y0) * rowsz;
        for (size_t i = 0; i < rowsz; i++) {
            int s = 0;
            for (int dy = -r; dy <= r; dy++) s += hs[(size_t)iclamp(y + dy, b0, b1 - 1) * rowsz + i];
            crow[i] = (int16_t)s; /* wraps like cv2's int16 cost type if bs>=15 */
        }
    }
    /* MODE_HH4 (cv2 4.13.0, probed: tests/golden/make_golden.py mode-3 cases): rows whose block window leaves the
     * image at the bottom (y + r > H-1) never get a cost — the volume stays 0 there. */
    if (p.mode == 3 && r > 0)
        for (int y = imax(H - r, y0); y < y1; y++) memset(Cvol + (size_t)(y - y0) * rowsz, 0, rowsz * sizeof(int16_t));
    free(Lp); free(Rp); free(pix); free(hs);
    return 0;
}

/* one scanline-recurrence step: SURVEY.md §3.2.4. Lp==NULL means "outside grid" (L=0,minL=0). */
static inline int path_step(const int16_t *Cp, const int16_t *Lp, int minLp, int D, int P1, int P2, int16_t *Lout)
{
    int minL = SVO_MAX_COST;
    if (!Lp) {
        for (int d = 0; d < D; d++) {
            /* min(0, MAX or 0 + P1, 0+P2) - 0 = 0  ->  L = C */
            int L = Cp[d];
            Lout[d] = (int16_t)L;
            if (L < minL) minL = L;
        }
        return minL;
    }
    const int delta = minLp + P2;
    for (int d = 0; d < D; d++) {
        int a = Lp[d];
        int b = (d > 0 ? Lp[d - 1] : SVO_MAX_COST) + P1;
        int c = (d < D - 1 ? Lp[d + 1] : SVO_MAX_COST) + P1;
        int L = Cp[d] + imin(imin(a, b), imin(c, delta)) - minLp;
        Lout[d] = (int16_t)L;
        if (L < minL) minL = L;
    }
    return minL;
}

static inline void sat_add_row(int16_t *S, const int16_t *L, int D)
{
    for (int d = 0; d < D; d++) {
        int s = (int)S[d] + (int)L[d];
        S[d] = (int16_t)(s > 32767 ? 32767 : s);
    }
}

/* aggregate one direction (predecessor offset ox,oy) into S (saturating). */
static void aggregate_dir(const int16_t *Cvol, int16_t *S, int H, int W1, int D, int ox, int oy, int P1, int P2)
{
    const size_t rowsz = (size_t)W1 * D;
    if (oy == 0) {
#pragma omp parallel
        {
            int16_t *La = malloc(sizeof(int16_t) * D), *Lb = malloc(sizeof(int16_t) * D);
#pragma omp for schedule(static)
            for (int y = 0; y < H; y++) {
                int16_t *prev = NULL, *cur = La;
                int minLp = 0;
                for (int k = 0; k < W1; k++) {
                    int x = ox < 0 ? k : W1 - 1 - k;
                    const int16_t *Cp = Cvol + (size_t)y * rowsz + (size_t)x * D;
                    minLp = path_step(Cp, prev, minLp, D, P1, P2, cur);
                    sat_add_row(S + (size_t)y * rowsz + (size_t)x * D, cur, D);
                    prev = cur;
                    cur = (cur == La) ? Lb : La;
                }
            }
            free(La); free(Lb);
        }
        return;
    }
    int16_t *rowA = malloc(sizeof(int16_t) * rowsz), *rowB = malloc(sizeof(int16_t) * rowsz);
    int *minA = malloc(sizeof(int) * W1), *minB = malloc(sizeof(int) * W1);
    int16_t *prevL = NULL, *curL = rowA;
    int *prevM = NULL, *curM = minA;
    for (int k = 0; k < H; k++) {
        int y = oy < 0 ? k : H - 1 - k;
#pragma omp parallel for schedule(static)
        for (int x = 0; x < W1; x++) {
            int xp = x + ox;
            const int16_t *Lp = (prevL && xp >= 0 && xp < W1) ? prevL + (size_t)xp * D : NULL;
            int mp = Lp ? prevM[xp] : 0;
            const int16_t *Cp = Cvol + (size_t)y * rowsz + (size_t)x * D;
            curM[x] = path_step(Cp, Lp, mp, D, P1, P2, curL + (size_t)x * D);
            sat_add_row(S + (size_t)y * rowsz + (size_t)x * D, curL + (size_t)x * D, D);
        }
        prevL = curL; prevM = curM;
        curL = (curL == rowA) ? rowB : rowA;
        curM = (curM == minA) ? minB : minA;
    }
    free(rowA); free(rowB); free(minA); free(minB);
}

/* Aggregated volume S (sat int16) for mode 0 (5 paths) or 1 (8 paths). */
int svo_sgbm_aggregate_dirs(const int16_t *Cvol, int H, int W1, int D, int P1, int P2, int ndirs, int16_t *S)
{
    memset(S, 0, sizeof(int16_t) * (size_t)H * W1 * D);
    static const int dirs[8][2] = {{-1, 0}, {-1, -1}, {0, -1}, {1, -1}, {1, 0}, {1, 1}, {0, 1}, {-1, 1}};
    for (int i = 0; i < ndirs && i < 8; i++) aggregate_dir(Cvol, S, H, W1, D, dirs[i][0], dirs[i][1], P1, P2);
    return 0;
}
/* any subset of the 8 directions: bit i of mask = dirs[i] (0 left, 1 up-left, 2 up, 3 up-right, 4 right, 5 down-right,
 * 6 down, 7 down-left predecessor). */
int svo_sgbm_aggregate_mask(const int16_t *Cvol, int H, int W1, int D, int P1, int P2, int mask, int16_t *S)
{
    memset(S, 0, sizeof(int16_t) * (size_t)H * W1 * D);
    static const int dirs[8][2] = {{-1, 0}, {-1, -1}, {0, -1}, {1, -1}, {1, 0}, {1, 1}, {0, 1}, {-1, 1}};
    for (int i = 0; i < 8; i++)
        if (mask >> i & 1) aggregate_dir(Cvol, S, H, W1, D, dirs[i][0], dirs[i][1], P1, P2);
    return 0;
}
int svo_sgbm_aggregate(const int16_t *Cvol, int H, int W1, int D, int P1, int P2, int mode, int16_t *S)
{
    /* MODE_HH4: left, up, right, down (4 paths); MODE_HH: all 8; MODE_SGBM: left, up-left, up, up-right, right */
    if (mode == 3) return svo_sgbm_aggregate_mask(Cvol, H, W1, D, P1, P2, 0x55, S);
    return svo_sgbm_aggregate_dirs(Cvol, H, W1, D, P1, P2, mode == 1 ? 8 : 5, S);
}

/* SURVEY.md §3.2.5 — WTA, uniqueness, subpixel, disp2 scatter, LR check; one image. */
static void select_impl(const int16_t *S, int H, int W, int W1, int D, const svo_params *p, int three_way, int16_t *disp);
void svo_sgbm_select(const int16_t *S, int H, int W, int W1, int D, const svo_params *p /*resolved*/, int16_t *disp)
{
    select_impl(S, H, W, W1, D, p, 0, disp);
}
/* three_way: MODE_SGBM_3WAY's selection (cv2 4.13.0, probed with tests/golden/make_golden_3way.py) differs from the
 * other modes' in two details: the winner among equal minima (8-lane SIMD search, see below) and the uniqueness test
 * (a precomputed integer threshold, skipped altogether when uniquenessRatio == 0). */
static void select_impl(const int16_t *S, int H, int W, int W1, int D, const svo_params *p, int three_way, int16_t *disp)
{
    const int minD = p->minDisparity, maxD = minD + D;
    const int minX1 = imax(maxD, 0), maxX1 = W + imin(minD, 0);
    const int INVALID = (minD - 1) * SVO_DISP_SCALE;
    const int uniq = p->uniquenessRatio, d12 = p->disp12MaxDiff;
    const size_t rowsz = (size_t)W1 * D;
#pragma omp parallel
    {
        int *disp2 = malloc(sizeof(int) * W), *disp2cost = malloc(sizeof(int) * W);
#pragma omp for schedule(static)
        for (int y = 0; y < H; y++) {
            int16_t *d1 = disp + (size_t)y * W;
            for (int x = 0; x < W; x++) { d1[x] = (int16_t)INVALID; disp2[x] = INVALID; disp2cost[x] = SVO_MAX_COST; }
            for (int xi = W1 - 1; xi >= 0; xi--) {
                const int16_t *Sp = S + (size_t)y * rowsz + (size_t)xi * D;
                /* cv2 starts from minS = MAX_COST, bestDisp = -1 and uses strict <: a pixel whose costs are all
                 * saturated keeps bestDisp = -1 (handled below; ends as INVALID). */
                int minS = SVO_MAX_COST, bestD = -1;
                for (int d = 0; d < D; d++)
                    if (Sp[d] < minS) { minS = Sp[d]; bestD = d; }
                if (three_way && bestD >= 0) {
                    /* the winner is searched with 8-lane SIMD: every lane (d mod 8) keeps the LAST d that attains its
                     * minimum, then the smallest such d among the lanes holding the global minimum wins */
                    const int NL = 8;
                    int best = -1;
                    for (int j = 0; j < NL && j < D; j++) {
                        int last = -1;
                        for (int d = j; d < D; d += NL)
                            if (Sp[d] == minS) last = d;
                        if (last >= 0 && (best < 0 || last < best)) best = last;
                    }
                    bestD = best;
                }
                int d;
                for (d = 0; d < D; d++)
                    if (three_way) {   /* S(d) <= floor(minS * 100 / (100 - uniq)), and no test at all for uniq == 0 */
                        if (uniq > 0 && (int)Sp[d] <= (minS * 100) / (100 - uniq) && abs(bestD - d) > 1) break;
                    } else if ((int)Sp[d] * (100 - uniq) < minS * 100 && abs(bestD - d) > 1) break;
                if (d < D) continue;
                if (bestD < 0) {
                    /* all costs saturated: cv2 proceeds with bestD=-1 (x2 = x+1-minD ...). Mirror: */
                    int x2 = xi + minX1 - bestD - minD;
                    if (x2 >= 0 && x2 < W && disp2cost[x2] > minS) { disp2cost[x2] = minS; disp2[x2] = bestD + minD; }
                    d1[xi + minX1] = (int16_t)(bestD * SVO_DISP_SCALE + minD * SVO_DISP_SCALE);
                    continue;
                }
                int x2 = xi + minX1 - bestD - minD;
                if (disp2cost[x2] > minS) { disp2cost[x2] = minS; disp2[x2] = bestD + minD; }
                int d16;
                if (0 < bestD && bestD < D - 1) {
                    int denom2 = imax((int)Sp[bestD - 1] + (int)Sp[bestD + 1] - 2 * (int)Sp[bestD], 1);
                    d16 = bestD * SVO_DISP_SCALE + (((int)Sp[bestD - 1] - (int)Sp[bestD + 1]) * SVO_DISP_SCALE + denom2) / (denom2 * 2);
                } else
                    d16 = bestD * SVO_DISP_SCALE;
                d1[xi + minX1] = (int16_t)(d16 + minD * SVO_DISP_SCALE);
            }
            for (int x = minX1; x < maxX1; x++) {
                int dv = d1[x];
                if (dv == INVALID) continue;
                int _d = dv >> SVO_DISP_SHIFT, d_ = (dv + SVO_DISP_SCALE - 1) >> SVO_DISP_SHIFT;
                int _x = x - _d, x_ = x - d_;
                if (0 <= _x && _x < W && disp2[_x] >= minD && abs(disp2[_x] - _d) > d12 &&
                    0 <= x_ && x_ < W && disp2[x_] >= minD && abs(disp2[x_] - d_) > d12)
                    d1[x] = (int16_t)INVALID;
            }
        }
        free(disp2); free(disp2cost);
    }
}

/* SURVEY.md §3.2.6.1 — 3x3 median on int16, BORDER_REPLICATE. src != dst. */
void svo_median3_s16(const int16_t *src, int H, int W, int16_t *dst)
{
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            int16_t v[9];
            int n = 0;
            for (int dy = -1; dy <= 1; dy++)
                for (int dx = -1; dx <= 1; dx++)
                    v[n++] = src[(size_t)iclamp(y + dy, 0, H - 1) * W + iclamp(x + dx, 0, W - 1)];
            for (int i = 1; i < 9; i++) { /* insertion sort */
                int16_t k = v[i]; int j = i - 1;
                while (j >= 0 && v[j] > k) { v[j + 1] = v[j]; j--; }
                v[j + 1] = k;
            }
            dst[(size_t)y * W + x] = v[4];
        }
}

/* SURVEY.md §3.2.6.2 — filterSpeckles, 4-connectivity, neighbour-to-neighbour predicate. In place. */
int svo_filter_speckles_s16(int16_t *img, int H, int W, int newVal, int maxSpeckleSize, int maxDiff)
{
    size_t n = (size_t)H * W;
    int32_t *label = calloc(n, sizeof(int32_t));
    int32_t *stack = malloc(n * sizeof(int32_t));
    if (!label || !stack) { free(label); free(stack); return -2; }
    int cur = 0;
    for (size_t s = 0; s < n; s++) {
        if (label[s] || img[s] == newVal) continue;
        cur++;
        size_t sp = 0, count = 0;
        stack[sp++] = (int32_t)s; label[s] = cur;
        size_t first = 0; /* we need the member list to repaint: reuse stack region as queue */
        (void)first;
        /* BFS using stack as a queue (members stay in stack[0..count)) */
        size_t head = 0;
        while (head < sp) {
            int32_t q = stack[head++];
            count++;
            int y = q / W, x = q % W;
            int v = img[q];
            static const int nb[4][2] = {{1, 0}, {-1, 0}, {0, 1}, {0, -1}};
            for (int k = 0; k < 4; k++) {
                int xx = x + nb[k][0], yy = y + nb[k][1];
                if (xx < 0 || xx >= W || yy < 0 || yy >= H) continue;
                size_t t = (size_t)yy * W + xx;
                if (label[t] || img[t] == newVal) continue;
                if (abs((int)img[t] - v) > maxDiff) continue;
                label[t] = cur; stack[sp++] = (int32_t)t;
            }
        }
        if ((int)count <= maxSpeckleSize)
            for (size_t i = 0; i < sp; i++) img[stack[i]] = (int16_t)newVal;
        /* NOTE: painted pixels keep their label so they are never revisited; since the predicate is
         * symmetric, components are independent of traversal order. */
    }
    free(label); free(stack);
    return 0;
}

static void sgbm_3way_raw(const uint8_t *left, const uint8_t *right, int H, int W, int C, size_t pitch,
                          const svo_params *praw, int16_t *disp_raw);
/* Whole StereoSGBM.compute(): planes -> C -> S -> select -> median3 -> speckle.    */
int svo_sgbm_compute(const uint8_t *left, const uint8_t *right, int H, int W, int C, size_t pitch,
                     const svo_params *praw, int16_t *disp)
{
    svo_params p;
    svo_resolve_params(praw, &p);
    if (p.mode < 0 || p.mode > 3) return -3;
    const int minD = p.minDisparity, D = p.numDisparities, maxD = minD + D;
    const int minX1 = imax(maxD, 0), maxX1 = W + imin(minD, 0), W1 = maxX1 - minX1;
    if (D <= 0) return -1;
    if (W1 <= 0) return -1;
    if (p.mode == 2) {
        const int stripe = (H + 3) / 4;
        if (p.blockSize / 2 + 1 + (int)ceil(0.1 * stripe) > stripe) return -3;
        int16_t *raw = malloc((size_t)H * W * sizeof(int16_t));
        if (!raw) return -2;
        sgbm_3way_raw(left, right, H, W, C, pitch, praw, raw);
        svo_median3_s16(raw, H, W, disp);
        if (p.speckleWindowSize > 0)
            svo_filter_speckles_s16(disp, H, W, (minD - 1) * SVO_DISP_SCALE, p.speckleWindowSize,
                                    SVO_DISP_SCALE * p.speckleRange);
        free(raw);
        return 0;
    }
    size_t vol = (size_t)H * W1 * D;
    int16_t *Cvol = malloc(vol * sizeof(int16_t)), *S = malloc(vol * sizeof(int16_t));
    int16_t *tmp = malloc((size_t)H * W * sizeof(int16_t));
    if (!Cvol || !S || !tmp) { free(Cvol); free(S); free(tmp); return -2; }
    int rc = svo_sgbm_cost_volume(left, right, H, W, C, pitch, praw, Cvol);
    if (rc) { free(Cvol); free(S); free(tmp); return rc; }
    svo_sgbm_aggregate(Cvol, H, W1, D, p.P1, p.P2, p.mode, S);
    svo_sgbm_select(S, H, W, W1, D, &p, tmp);
    svo_median3_s16(tmp, H, W, disp);
    if (p.speckleWindowSize > 0)
        svo_filter_speckles_s16(disp, H, W, (minD - 1) * SVO_DISP_SCALE, p.speckleWindowSize,
                                SVO_DISP_SCALE * p.speckleRange);
    free(Cvol); free(S); free(tmp);
    return 0;
}

/* MODE_SGBM_3WAY (cv2 4.13.0, probed — tests/golden/make_golden_3way.py): the image is cut into 4 row stripes of
 * ceil(H/4) rows; stripe n > 0 starts `overlap` = blockSize/2 + 1 + ceil(0.1 * stripe) rows early and only its own
 * rows are kept.  Inside a stripe: block sums clamped to the stripe's first row at the top (and to the image at the
 * bottom), three paths — from the left, from above (restarted at the stripe's first row) and from the right — with the
 * ordinary recurrence, then the 3-way selection (select_impl).  Images so short that overlap > stripe behave
 * differently in cv2 (not modelled: svo_sgbm_compute returns -3 there, and so does the engine). */
static void sgbm_3way_raw(const uint8_t *left, const uint8_t *right, int H, int W, int C, size_t pitch,
                          const svo_params *praw, int16_t *disp_raw)
{
    svo_params p;
    svo_resolve_params(praw, &p);
    const int minD = p.minDisparity, D = p.numDisparities, maxD = minD + D;
    const int minX1 = imax(maxD, 0), maxX1 = W + imin(minD, 0), W1 = maxX1 - minX1;
    const int nstripes = 4;
    const int stripe = (H + nstripes - 1) / nstripes;
    const int overlap = p.blockSize / 2 + 1 + (int)ceil(0.1 * stripe);
    for (int n = 0; n < nstripes; n++) {
        const int ys = n == 0 ? 0 : imax(imin(n * stripe - overlap, H), 0);
        const int ye = n == nstripes - 1 ? H : imin((n + 1) * stripe, H);
        const int keep0 = n == 0 ? 0 : imin(n * stripe, H);
        if (ye <= ys || keep0 >= ye) continue;
        const int hs = ye - ys;
        size_t vol = (size_t)hs * W1 * D;
        int16_t *Cvol = malloc(vol * sizeof(int16_t)), *S = malloc(vol * sizeof(int16_t));
        int16_t *tmp = malloc((size_t)hs * W * sizeof(int16_t));
        cost_volume_rows(left, right, H, W, C, pitch, praw, ys, ye, Cvol);
        svo_sgbm_aggregate_mask(Cvol, hs, W1, D, p.P1, p.P2, 0x15, S);   /* left, up, right predecessors */
        select_impl(S, hs, W, W1, D, &p, 1, tmp);
        memcpy(disp_raw + (size_t)keep0 * W, tmp + (size_t)(keep0 - ys) * W, (size_t)(ye - keep0) * W * sizeof(int16_t));
        free(Cvol); free(S); free(tmp);
    }
}

/* debug tap: rows [ys, ye) of the cost volume with 3-way clamping, and the 3-path S of that stripe */
int svo_3way_stripe_debug(const uint8_t *left, const uint8_t *right, int H, int W, int C, size_t pitch,
                          const svo_params *praw, int ys, int ye, int16_t *Cvol, int16_t *S)
{
    svo_params p;
    svo_resolve_params(praw, &p);
    const int minD = p.minDisparity, D = p.numDisparities, maxD = minD + D;
    const int minX1 = imax(maxD, 0), maxX1 = W + imin(minD, 0), W1 = maxX1 - minX1;
    cost_volume_rows(left, right, H, W, C, pitch, praw, ys, ye, Cvol);
    svo_sgbm_aggregate_mask(Cvol, ye - ys, W1, D, p.P1, p.P2, 0x15, S);
    return 0;
}

/* pre-median / pre-speckle disparity, for stage-level diffs */
int svo_sgbm_compute_raw(const uint8_t *left, const uint8_t *right, int H, int W, int C, size_t pitch,
                         const svo_params *praw, int16_t *disp_raw, int16_t *Cvol_out, int16_t *S_out)
{
    svo_params p;
    svo_resolve_params(praw, &p);
    const int minD = p.minDisparity, D = p.numDisparities, maxD = minD + D;
    const int minX1 = imax(maxD, 0), maxX1 = W + imin(minD, 0), W1 = maxX1 - minX1;
    if (D <= 0 || W1 <= 0) return -1;
    size_t vol = (size_t)H * W1 * D;
    int16_t *Cvol = Cvol_out ? Cvol_out : malloc(vol * sizeof(int16_t));
    int16_t *S = S_out ? S_out : malloc(vol * sizeof(int16_t));
    int rc = svo_sgbm_cost_volume(left, right, H, W, C, pitch, praw, Cvol);
    if (!rc) {
        svo_sgbm_aggregate(Cvol, H, W1, D, p.P1, p.P2, p.mode, S);
        svo_sgbm_select(S, H, W, W1, D, &p, disp_raw);
    }
    if (!Cvol_out) free(Cvol);
    if (!S_out) free(S);
    return rc;
}

/* ---- remap, SURVEY.md §3.2.7 ---- */
static inline int cv_round_x86(float v)
{
    /* cvtps2dq semantics: round-half-even; NaN / out of int32 range -> INT_MIN */
    if (!(v >= -2147483648.0f && v < 2147483648.0f)) return INT32_MIN;
    return (int)lrintf(v); /* default rounding mode = nearest-even */
}
static inline int sat16(int v) { return v < -32768 ? -32768 : (v > 32767 ? 32767 : v); }

void svo_remap_u8(const uint8_t *src, int sH, int sW, int C, size_t spitch, const float *mapx, const float *mapy,
                  int dH, int dW, uint8_t *dst, size_t dpitch)
{
#pragma omp parallel for schedule(static)
    for (int y = 0; y < dH; y++)
        for (int x = 0; x < dW; x++) {
            int sx = cv_round_x86(mapx[(size_t)y * dW + x] * 32.0f);
            int sy = cv_round_x86(mapy[(size_t)y * dW + x] * 32.0f);
            int ix = sat16(sx >> 5), iy = sat16(sy >> 5), fx = sx & 31, fy = sy & 31;
            int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32, w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
            for (int c = 0; c < C; c++) {
#define TAP(yy, xx) (((yy) >= 0 && (yy) < sH && (xx) >= 0 && (xx) < sW) ? (int)src[(size_t)(yy) * spitch + (size_t)(xx) * C + c] : 0)
                int v = w00 * TAP(iy, ix) + w01 * TAP(iy, ix + 1) + w10 * TAP(iy + 1, ix) + w11 * TAP(iy + 1, ix + 1);
#undef TAP
                dst[(size_t)y * dpitch + (size_t)x * C + c] = (uint8_t)((v + 16384) >> 15);
            }
        }
}

/* ---- initUndistortRectifyMap, SURVEY.md §3.2.8 ---- */
static int inv3x3(const double *A, double *Ai)
{
    /* Gauss-Jordan with partial pivoting (cv2 uses LU; agreement checked after the x32 quantisation) */
    double M[3][6];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) { M[i][j] = A[i * 3 + j]; M[i][j + 3] = i == j; }
    for (int c = 0; c < 3; c++) {
        int piv = c;
        for (int r = c + 1; r < 3; r++) if (fabs(M[r][c]) > fabs(M[piv][c])) piv = r;
        if (M[piv][c] == 0.0) return -1;
        if (piv != c) for (int j = 0; j < 6; j++) { double t = M[c][j]; M[c][j] = M[piv][j]; M[piv][j] = t; }
        double d = M[c][c];
        for (int j = 0; j < 6; j++) M[c][j] /= d;
        for (int r = 0; r < 3; r++) if (r != c) {
            double f = M[r][c];
            for (int j = 0; j < 6; j++) M[r][j] -= f * M[c][j];
        }
    }
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) Ai[i * 3 + j] = M[i][j + 3];
    return 0;
}

/* K: 3x3, dist: ndist in {0,4,5,8,12,14}, R: 3x3 (NULL=I), P: 3x3 "new camera matrix" part (row-major 9,
 * i.e. P[:3,:3]; NULL = K). Outputs f32 H*W maps. */
int svo_init_undistort_rectify_map(const double *K, const double *dist, int ndist, const double *R, const double *P33,
                                   int W, int H, float *mapx, float *mapy)
{
    double k[14] = {0};
    for (int i = 0; i < ndist && i < 14; i++) k[i] = dist[i];
    const double k1 = k[0], k2 = k[1], p1 = k[2], p2 = k[3], k3 = k[4], k4 = k[5], k5 = k[6], k6 = k[7];
    const double s1 = k[8], s2 = k[9], s3 = k[10], s4 = k[11], tx = k[12], ty = k[13];
    double Rm[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, A[9], PR[9], iR[9];
    if (R) memcpy(Rm, R, sizeof Rm);
    memcpy(A, P33 ? P33 : K, sizeof A);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double s = 0;
            for (int t = 0; t < 3; t++) s += A[i * 3 + t] * Rm[t * 3 + j];
            PR[i * 3 + j] = s;
        }
    if (inv3x3(PR, iR)) return -1;
    const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    /* tilt (Scheimpflug) matrix, identity when tx=ty=0 */
    double mt[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (tx != 0.0 || ty != 0.0) {
        double ctx = cos(tx), stx = sin(tx), cty = cos(ty), sty = sin(ty);
        double rot[9] = {cty, sty * stx, -sty * ctx, 0, ctx, stx, sty, -cty * stx, cty * ctx};
        double pz[9] = {rot[8], 0, -rot[2], 0, rot[8], -rot[5], 0, 0, 1};
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                double s = 0;
                for (int t = 0; t < 3; t++) s += pz[i * 3 + t] * rot[t * 3 + j];
                mt[i * 3 + j] = s;
            }
    }
#pragma omp parallel for schedule(static)
    for (int i = 0; i < H; i++)
        for (int j = 0; j < W; j++) {
            double X = iR[0] * j + iR[1] * i + iR[2];
            double Y = iR[3] * j + iR[4] * i + iR[5];
            double Wc = iR[6] * j + iR[7] * i + iR[8];
            double w = 1.0 / Wc, x = X * w, y = Y * w;
            double x2 = x * x, y2 = y * y, r2 = x2 + y2, _2xy = 2 * x * y;
            double kr = (1 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1 + ((k6 * r2 + k5) * r2 + k4) * r2);
            double xd = x * kr + p1 * _2xy + p2 * (r2 + 2 * x2) + s1 * r2 + s2 * r2 * r2;
            double yd = y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy + s3 * r2 + s4 * r2 * r2;
            double vx = mt[0] * xd + mt[1] * yd + mt[2];
            double vy = mt[3] * xd + mt[4] * yd + mt[5];
            double vw = mt[6] * xd + mt[7] * yd + mt[8];
            double invp = vw != 0.0 ? 1.0 / vw : 1.0;
            mapx[(size_t)i * W + j] = (float)(fx * invp * vx + cx);
            mapy[(size_t)i * W + j] = (float)(fy * invp * vy + cy);
        }
    return 0;
}

/* ---- reprojectImageTo3D, SURVEY.md §3.2.9 (f64 evaluation, f32 out; tol 1e-5 rel vs cv2) ----
 * dtype: 0=u8, 1=s16, 2=s32, 3=f32.  The raw value is used (no /16), as cv2 does. */
void svo_reproject_to_3d(const void *disp, int dtype, int H, int W, const double *Q, int handleMissing, float *xyz)
{
    double minv = DBL_MAX;
    size_t n = (size_t)H * W;
    if (handleMissing)
        for (size_t i = 0; i < n; i++) {
            double d = dtype == 0 ? ((const uint8_t *)disp)[i] : dtype == 1 ? ((const int16_t *)disp)[i]
                     : dtype == 2 ? ((const int32_t *)disp)[i] : (double)((const float *)disp)[i];
            if (d < minv) minv = d;
        }
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            size_t i = (size_t)y * W + x;
            double d = dtype == 0 ? ((const uint8_t *)disp)[i] : dtype == 1 ? ((const int16_t *)disp)[i]
                     : dtype == 2 ? ((const int32_t *)disp)[i] : (double)((const float *)disp)[i];
            double X = Q[0] * x + Q[1] * y + Q[2] * d + Q[3];
            double Y = Q[4] * x + Q[5] * y + Q[6] * d + Q[7];
            double Z = Q[8] * x + Q[9] * y + Q[10] * d + Q[11];
            double Wh = Q[12] * x + Q[13] * y + Q[14] * d + Q[15];
            double iw = 1.0 / Wh;
            float fx_ = (float)(X * iw), fy_ = (float)(Y * iw), fz_ = (float)(Z * iw);
            if (handleMissing && fabs(d - minv) <= FLT_EPSILON) fz_ = 10000.0f;
            xyz[i * 3 + 0] = fx_; xyz[i * 3 + 1] = fy_; xyz[i * 3 + 2] = fz_;
        }
}

/* ---- per-box median depth (NEW op; semantics defined in SURVEY.md §8a row a10) ----
 * disp16: int16 disparity x16 as StereoSGBM.compute returns; Z evaluated as
 * reprojectImageTo3D(disp16.astype(f32)/16, Q)[...,2].  boxes: N x 4 f32 xyxy in left-image pixels.
 * valid <=> disp16 != invalid16 and isfinite(Z).  out[n] = np.median semantics in f32, NaN if none. */
static int cmp_f32(const void *a, const void *b)
{
    float x = *(const float *)a, y = *(const float *)b;
    return x < y ? -1 : x > y;
}
void svo_box_median_depth(const int16_t *disp16, int H, int W, const double *Q, int invalid16, const float *boxes,
                          int N, float *out)
{
    for (int b = 0; b < N; b++) {
        int x1 = (int)floorf(boxes[b * 4 + 0]), y1 = (int)floorf(boxes[b * 4 + 1]);
        int x2 = (int)ceilf(boxes[b * 4 + 2]), y2 = (int)ceilf(boxes[b * 4 + 3]);
        x1 = iclamp(x1, 0, W); x2 = iclamp(x2, 0, W); y1 = iclamp(y1, 0, H); y2 = iclamp(y2, 0, H);
        size_t cap = (size_t)imax(x2 - x1, 0) * imax(y2 - y1, 0), n = 0;
        float *v = malloc(sizeof(float) * (cap ? cap : 1));
        for (int y = y1; y < y2; y++)
            for (int x = x1; x < x2; x++) {
                int dv = disp16[(size_t)y * W + x];
                if (dv == invalid16) continue;
                double d = (double)((float)dv / 16.0f);
                double Z = Q[8] * x + Q[9] * y + Q[10] * d + Q[11];
                double Wh = Q[12] * x + Q[13] * y + Q[14] * d + Q[15];
                float z = (float)(Z * (1.0 / Wh));
                if (!isfinite(z)) continue;
                v[n++] = z;
            }
        if (!n) out[b] = NAN;
        else {
            qsort(v, n, sizeof(float), cmp_f32);
            out[b] = (n & 1) ? v[n / 2] : (v[n / 2 - 1] + v[n / 2]) * 0.5f;
        }
        free(v);
    }
}

int svo_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void svo_set_num_threads(int n)
{
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ================= SURVEY.md §8(f) "next" rows: the sparse 3-D path the reference runs per detection ===========
 * cv2.undistortPoints + cv2.triangulatePoints as called by ball_drop.py:225-259 (calculate_3d_position),
 * checkerboard_distance.py:83-103 and 3dpose.py:530-561; cv2.cvtColor(BGR2GRAY) in front of the matcher.
 * Semantics recovered by probing the cv2 4.13.0 binary (tests/golden/make_golden.py pins them):
 *   undistortPoints: x=(u-cx)/fx, y=(v-cy)/fy; exactly 5 fixed-point iterations of the inverse distortion in
 *     double (default TermCriteria(COUNT,5,0.01)); then (P[:, :3]*R)*[x y 1], divide.  Bit-exact in f64.
 *   triangulatePoints: per point the 4x4 DLT matrix [x1*P1[2]-P1[0]; y1*P1[2]-P1[1]; x2*P2[2]-P2[0]; y2*P2[2]-P2[1]],
 *     right singular vector of the smallest singular value (sign arbitrary, callers divide by w).
 *   BGR2GRAY (u8): (B*3735 + G*19235 + R*9798 + 16384) >> 15   (4.13 uses 15-bit coefficients).
 */
void svo_undistort_points(const double *src, int n, const double *K, const double *dist, int ndist, const double *R,
                          const double *P33, double *dst)
{
    double k[14] = {0};
    for (int i = 0; i < ndist && i < 14; i++) k[i] = dist[i];
    const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    double RR[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (R) memcpy(RR, R, sizeof RR);
    if (P33) {
        double t[9];
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) t[3 * r + c] = P33[3 * r] * RR[c] + P33[3 * r + 1] * RR[3 + c] + P33[3 * r + 2] * RR[6 + c];
        memcpy(RR, t, sizeof RR);
    }
    double imt[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    const int has_tilt = k[12] != 0.0 || k[13] != 0.0;
    if (has_tilt) {   /* inverse tilt projection = matRotXY^T * invProjZ */
        const double ctx_ = cos(k[12]), stx = sin(k[12]), cty = cos(k[13]), sty = sin(k[13]);
        const double rot[9] = {cty, sty * stx, -sty * ctx_, 0, ctx_, stx, sty, -cty * stx, cty * ctx_};
        const double inv = 1. / rot[8];
        const double ipz[9] = {inv, 0, inv * rot[2], 0, inv, inv * rot[5], 0, 0, 1};
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                double s = 0;
                for (int t = 0; t < 3; t++) s += rot[t * 3 + i] * ipz[t * 3 + j];
                imt[i * 3 + j] = s;
            }
    }
    for (int i = 0; i < n; i++) {
        double x = (src[2 * i] - cx) / fx, y = (src[2 * i + 1] - cy) / fy;
        if (has_tilt) {
            const double vx = imt[0] * x + imt[1] * y + imt[2], vy = imt[3] * x + imt[4] * y + imt[5];
            const double vz = imt[6] * x + imt[7] * y + imt[8];
            const double ip = vz != 0.0 ? 1. / vz : 1.;
            x = ip * vx; y = ip * vy;
        }
        const double x0 = x, y0 = y;
        for (int it = 0; it < 5; it++) {
            const double r2 = x * x + y * y;
            double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = x0; y = y0; break; }
            const double dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - dx) * icdist;
            y = (y0 - dy) * icdist;
        }
        const double X = RR[0] * x + RR[1] * y + RR[2], Y = RR[3] * x + RR[4] * y + RR[5], Wc = RR[6] * x + RR[7] * y + RR[8];
        const double iw = 1. / Wc;
        dst[2 * i] = X * iw;
        dst[2 * i + 1] = Y * iw;
    }
}

/* one-sided Jacobi (Hestenes) on the 4x4 DLT matrix; out = V column of the smallest singular value */
void svo_triangulate_points(const double *P1, const double *P2, const double *a, const double *b, int n, double *out4)
{
    for (int i = 0; i < n; i++) {
        double A[4][4], V[4][4];
        for (int c = 0; c < 4; c++) {
            A[0][c] = a[2 * i] * P1[8 + c] - P1[c];
            A[1][c] = a[2 * i + 1] * P1[8 + c] - P1[4 + c];
            A[2][c] = b[2 * i] * P2[8 + c] - P2[c];
            A[3][c] = b[2 * i + 1] * P2[8 + c] - P2[4 + c];
            for (int r = 0; r < 4; r++) V[r][c] = r == c;
        }
        for (int sweep = 0; sweep < 30; sweep++) {
            int rotated = 0;
            for (int p = 0; p < 3; p++)
                for (int q = p + 1; q < 4; q++) {
                    double al = 0, be = 0, ga = 0;
                    for (int r = 0; r < 4; r++) { al += A[r][p] * A[r][p]; be += A[r][q] * A[r][q]; ga += A[r][p] * A[r][q]; }
                    if (fabs(ga) <= 1e-15 * sqrt(al * be) || ga == 0) continue;
                    rotated = 1;
                    const double zeta = (be - al) / (2 * ga);
                    const double t = (zeta >= 0 ? 1. : -1.) / (fabs(zeta) + sqrt(1 + zeta * zeta));
                    const double cs = 1. / sqrt(1 + t * t), sn = cs * t;
                    for (int r = 0; r < 4; r++) {
                        const double ap = A[r][p], aq = A[r][q];
                        A[r][p] = cs * ap - sn * aq; A[r][q] = sn * ap + cs * aq;
                        const double vp = V[r][p], vq = V[r][q];
                        V[r][p] = cs * vp - sn * vq; V[r][q] = sn * vp + cs * vq;
                    }
                }
            if (!rotated) break;
        }
        int best = 0; double bn = DBL_MAX;
        for (int c = 0; c < 4; c++) {
            double s = 0;
            for (int r = 0; r < 4; r++) s += A[r][c] * A[r][c];
            if (s < bn) { bn = s; best = c; }
        }
        for (int r = 0; r < 4; r++) out4[4 * i + r] = V[r][best];
    }
}

void svo_bgr2gray(const uint8_t *src, int n, uint8_t *dst)
{
    for (int i = 0; i < n; i++)
        dst[i] = (uint8_t)((src[3 * i] * 3735 + src[3 * i + 1] * 19235 + src[3 * i + 2] * 9798 + 16384) >> 15);
}
