// svb_cost.cu — matching cost: preFilterCap-clipped Sobel-x + intensity planes, Birchfield-Tomasi sampling,
// blockSize box sum.  Semantics: SURVEY.md §3.2.2-3.2.3.  Output: C[y][xi][Dp] int16 (packed pairs, d fastest).
//
//   k_planes : u8 images -> per-pixel BT operands, already widened to packed s16x2 so the cost kernel's inner
//              loop is pure DPX/VIADD work:
//                left  LP[y][pl][xi]          = uint4 { (u,u), (255-u,255-u), (u0,u0), (255-u1,255-u1) }
//                right RP[y][pl][copy 8][arr 4][Wr+8] = int16 arrays V, 255-V, V0, 255-V1 stored REVERSED in x
//                                               (index j = W-1-x); copy c holds element j at position j-c+8, so a
//                                               run of 8 disparities starting at any j is one aligned LDG.128.
//   k_cost_fused : BT pixel cost for 8 disparities per cell (4 packed registers), 2R+1 box sum in x through a
//              shared-memory pix row, running box sum in y through a shared-memory ring -> C (written once),
//              padded disparities overwritten with CPAD
//
// BT in biased form (no negative operands):  with a' = 255-a,
//   c0+255 = max(255, u+v1', v0+u')   c1+255 = max(255, v+u1', u0+v')   cost = max(min(X,Y)-255, 0)
//   X = max(u+v1', v0+u'), Y = max(v+u1', u0+v')            -> 2 VIADD + 2 VIADDMNMX + 1 VIMNMX + 1 VIADDMNMX.RELU
#include "svb_common.cuh"
#include <algorithm>


__device__ __forceinline__ int grad_at(const uint8_t *rc, const uint8_t *rn, const uint8_t *rs, int x, int W, int C,
                                       int ch, int ftzero)
{
    if (x < 1 || x > W - 2) return ftzero;
    int g = 2 * ((int)rc[(x + 1) * C + ch] - (int)rc[(x - 1) * C + ch]) + ((int)rn[(x + 1) * C + ch] - (int)rn[(x - 1) * C + ch]) +
            ((int)rs[(x + 1) * C + ch] - (int)rs[(x - 1) * C + ch]);
    return max(-ftzero, min(ftzero, g)) + ftzero;
}
__device__ __forceinline__ int inten_at(const uint8_t *rc, int x, int W, int C, int ch, int ftzero)
{
    if (x < 1 || x > W - 2) return ftzero;
    return rc[x * C + ch];
}
__device__ __forceinline__ void bt_triplet(int pm, int p, int pp, bool has_l, bool has_r, int &lo, int &hi)
{
    int pl = has_l ? (p + pm) >> 1 : p;
    int pr = has_r ? (p + pp) >> 1 : p;
    lo = min(p, min(pl, pr));
    hi = max(p, max(pl, pr));
}
__device__ __forceinline__ uint32_t dup16(int v) { return (uint32_t)v * 0x10001u; }

// One CTA per image row.  Per plane: plane values of both images are staged in shared memory once, the left BT
// operands go out as one uint4 per column, the right operands are built reversed in x in shared memory and then
// written as 8 copies, copy c shifted so that a disparity run starting at reversed index j0 with j0 % 8 == c is
// 16-byte aligned: the cost kernel reads every right operand with one coalesced LDG.128 per lane.
//   RP[y][plane][copy c][array a][Wr8],  element j of array a sits at position j - c + 8 of copy c.
#define PLANES_THREADS 256
__global__ void __launch_bounds__(PLANES_THREADS) k_planes(const uint8_t *__restrict__ left, const uint8_t *__restrict__ right,
                                                           int H, int W, int C, size_t pitch, size_t istride, int ftzero,
                                                           int minX1, int W1, int Wr, uint4 *__restrict__ LP,
                                                           int16_t *__restrict__ RP, size_t ws_stride)
{
    extern __shared__ __align__(16) unsigned char psm[];
    const int Wa = Wr + 16;                                   // staged array length (8 zeros in front, >= 8 behind)
    int16_t *A = reinterpret_cast<int16_t *>(psm);            // [4][Wa]
    uint8_t *pvL = psm + (size_t)4 * Wa * 2;                  // [W]
    uint8_t *pvR = pvL + ((W + 15) & ~15);                    // [W]
    const int y = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const int NPL = 2 * C;
    const int Wr8 = Wr + 8;
    left += (size_t)b * istride;
    right += (size_t)b * istride;
    LP = reinterpret_cast<uint4 *>(reinterpret_cast<char *>(LP) + (size_t)b * ws_stride);
    RP = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(RP) + (size_t)b * ws_stride);
    const int yn = max(y - 1, 0), ys = min(y + 1, H - 1);
    const uint8_t *lc = left + (size_t)y * pitch, *ln = left + (size_t)yn * pitch, *ls = left + (size_t)ys * pitch;
    const uint8_t *rc = right + (size_t)y * pitch, *rn = right + (size_t)yn * pitch, *rs = right + (size_t)ys * pitch;
    for (int pl = 0; pl < NPL; pl++) {
        const int ch = pl >> 1;
        __syncthreads();   // previous plane's readers are done with A / pv
        for (int x = tid; x < W; x += PLANES_THREADS) {
            pvL[x] = (uint8_t)((pl & 1) ? inten_at(lc, x, W, C, ch, ftzero) : grad_at(lc, ln, ls, x, W, C, ch, ftzero));
            pvR[x] = (uint8_t)((pl & 1) ? inten_at(rc, x, W, C, ch, ftzero) : grad_at(rc, rn, rs, x, W, C, ch, ftzero));
        }
        __syncthreads();
        // left operands
        for (int xi = tid; xi < W1; xi += PLANES_THREADS) {
            const int x = xi + minX1;
            int lo, hi;
            const int p = pvL[x];
            bt_triplet(x > 0 ? pvL[x - 1] : 0, p, x < W - 1 ? pvL[x + 1] : 0, x > 0, x < W - 1, lo, hi);
            LP[((size_t)y * NPL + pl) * W1 + xi] = make_uint4(dup16(p), dup16(255 - p), dup16(lo), dup16(255 - hi));
        }
        // right operands, reversed: staged index j + 8 <-> x = W-1-j
        for (int i = tid; i < Wa; i += PLANES_THREADS) {
            const int x = W - 1 - (i - 8);
            int v = 0, vn = 0, v0 = 0, v1n = 0;
            if (i >= 8 && x >= 0) {
                int lo, hi;
                const int p = pvR[x];
                bt_triplet(x > 0 ? pvR[x - 1] : 0, p, x < W - 1 ? pvR[x + 1] : 0, x > 0, x < W - 1, lo, hi);
                v = p; vn = 255 - p; v0 = lo; v1n = 255 - hi;
            }
            A[i] = (int16_t)v; A[Wa + i] = (int16_t)vn; A[2 * Wa + i] = (int16_t)v0; A[3 * Wa + i] = (int16_t)v1n;
        }
        __syncthreads();
        // 8 shifted copies, 16 bytes per store
        const int nq = Wr8 / 8;
        int16_t *outp = RP + ((size_t)y * NPL + pl) * 32 * Wr8;
        for (int i = tid; i < 4 * nq; i += PLANES_THREADS) {
            const int a = i / nq, q = i - a * nq;
            const uint4 va = *reinterpret_cast<const uint4 *>(A + (size_t)a * Wa + 8 * q);
            const uint4 vb = *reinterpret_cast<const uint4 *>(A + (size_t)a * Wa + 8 * q + 8);
            const uint32_t w[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
#pragma unroll
            for (int c = 0; c < 8; c++) {
                uint32_t o[4];
#pragma unroll
                for (int k = 0; k < 4; k++)
                    o[k] = (c & 1) ? __byte_perm(w[c / 2 + k], w[c / 2 + k + 1], 0x5432) : w[c / 2 + k];
                *reinterpret_cast<uint4 *>(outp + ((size_t)c * 4 + a) * Wr8 + 8 * q) = make_uint4(o[0], o[1], o[2], o[3]);
            }
        }
    }
}

int svb_launch_planes(const SgbmPlan &pl, const uint8_t *left, const uint8_t *right, size_t pitch, size_t istride,
                      char *ws, size_t ws_stride, int B, cudaStream_t st)
{
    dim3 grid(pl.H, B);
    const size_t smem = (size_t)4 * (pl.Wr + 16) * 2 + 2 * (size_t)((pl.W + 15) & ~15);
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        SVB_CUDA_CHECK(cudaFuncSetAttribute(k_planes, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    k_planes<<<grid, PLANES_THREADS, smem, st>>>(left, right, pl.H, pl.W, pl.ch, pitch, istride, pl.ftzero, pl.minX1,
                                                 pl.W1, pl.Wr, (uint4 *)(ws + pl.off_lplanes),
                                                 (int16_t *)(ws + pl.off_rplanes), ws_stride);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

// ---- fused cost kernel: BT pixel cost -> horizontal box sum -> vertical running box sum -> C ----
// One CTA owns a tile of TX = TXP-2R output columns x all disparities and marches down a segment of rows.
// Per input row r (clamped to the image: cost-space border replication):
//   phase 1 (window-mapped): the right-image operands of a cell (column xc, disparities 8l..8l+7) are the 8 reversed
//            pixels starting at w = jb0 - xc + 8l, so all cells on the diagonal xc - 8l = const share them.  A thread
//            owns one window w: it loads the 4 operand arrays x NPL planes once (aligned LDG.128 from copy w%8,
//            prefetched one row ahead) and computes every cell of its diagonal inside the tile (<= TXP/8+1 cells),
//            parking the pixel costs in a double-buffered shared-memory row (column stride Dp+8: conflict-free).
//   phase 2 (column-mapped): lane l of column block cb walks CPB adjacent output columns with a running sum of the
//            2R+1 neighbouring pix columns (Hs), swaps Hs into the cell's private slot of a (2R+1)-row ring in
//            shared memory and updates the vertical running sum acc += Hs(r) - Hs(r-(2R+1)); from the (2R+1)-th
//            row on it writes C[r-R][x][8l..8l+8).
// The horizontal sums never touch HBM; C is written exactly once (2 B per pixel-disparity).
#define COST_THREADS 352
#define COST_CPB_MAX 6   // output columns per thread (phase 2)

template <int NPL>
struct CostWindow { uint4 v[NPL][4]; };   // [plane][V, 255-V, V0, 255-V1]

template <int NPL>
__device__ __forceinline__ void load_window(CostWindow<NPL> &cw, const int16_t *__restrict__ RProw, int Wr8, int w)
{
    const int cpy = w & 7;
    const int16_t *rb = RProw + (size_t)cpy * 4 * Wr8 + (w - cpy + 8);
#pragma unroll
    for (int pl = 0; pl < NPL; pl++)
#pragma unroll
        for (int a = 0; a < 4; a++)
            cw.v[pl][a] = __ldg(reinterpret_cast<const uint4 *>(rb + ((size_t)pl * 32 + a) * Wr8));
}

__device__ __forceinline__ uint32_t u4get(const uint4 &q, int i) { return i == 0 ? q.x : i == 1 ? q.y : i == 2 ? q.z : q.w; }

template <int NPL>
__device__ __forceinline__ void pix_cell(const CostWindow<NPL> &cw, const uint4 *__restrict__ LPx, int W1, uint32_t (&acc)[4])
{
#pragma unroll
    for (int i = 0; i < 4; i++) acc[i] = 0;
#pragma unroll
    for (int pl = 0; pl < NPL; pl++) {
        const uint4 Lw = __ldg(LPx + (size_t)pl * W1);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t Bv = __vadd2(u4get(cw.v[pl][2], i), Lw.y);                 // v0 + (255-u)
            const uint32_t X = __viaddmax_s16x2(Lw.x, u4get(cw.v[pl][3], i), Bv);     // max(u + 255-v1, .)
            const uint32_t Ev = __vadd2(Lw.z, u4get(cw.v[pl][1], i));                 // u0 + (255-v)
            const uint32_t Y = __viaddmax_s16x2(u4get(cw.v[pl][0], i), Lw.w, Ev);     // max(v + 255-u1, .)
            uint32_t c = __viaddmax_s16x2_relu(__vmins2(X, Y), 0xFF01FF01u, 0u);      // max(min-255, 0)
            if (pl & 1) c = (c >> 2) & 0x3FFF3FFFu;
            acc[i] = __vadd2(acc[i], c);
        }
    }
}

template <int NPL, int R>
__global__ void __launch_bounds__(COST_THREADS) k_cost_fused(const uint4 *__restrict__ LP, const int16_t *__restrict__ RP,
                                                             int16_t *__restrict__ Cv, int H, int W1, int Wr, int Dp,
                                                             int D, int G, int jb0, int TXP, int rows_per_seg,
                                                             uint32_t cpad, int zrow, size_t ws_stride)
{
    constexpr int BS = 2 * R + 1;
    extern __shared__ __align__(16) unsigned char csm[];
    const int TX = TXP - 2 * R;
    const int PS = Dp + 8;                                                 // pix column stride (elements)
    int16_t *pixbuf = reinterpret_cast<int16_t *>(csm);                    // [2][TXP][PS]
    int16_t *ring = pixbuf + (size_t)2 * TXP * PS;                         // [BS][TX][Dp]
    const int b = blockIdx.z, tid = threadIdx.x;
    const int x0 = blockIdx.x * TX;
    const int ya = blockIdx.y * rows_per_seg, yb = min(ya + rows_per_seg, H);
    const int Wr8 = Wr + 8;
    LP = reinterpret_cast<const uint4 *>(reinterpret_cast<const char *>(LP) + (size_t)b * ws_stride);
    RP = reinterpret_cast<const int16_t *>(reinterpret_cast<const char *>(RP) + (size_t)b * ws_stride);
    Cv = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(Cv) + (size_t)b * ws_stride);
    {   // ring starts at zero: "subtract the row that leaves the window" is then unconditional
        uint4 *rz = reinterpret_cast<uint4 *>(ring);
        const int n = BS * TX * Dp / 8;
        for (int i = tid; i < n; i += COST_THREADS) rz[i] = make_uint4(0, 0, 0, 0);
    }
    // ---- phase-1 role: window w = wmin + tid; valid pix columns of the tile are [xlo, xhi] ----
    const int xlo = max(x0 - R, 0), xhi = min(x0 - R + TXP - 1, W1 - 1);
    const int w = jb0 - xhi + tid;
    const bool hasWin = tid < (xhi - xlo) + 8 * (G - 1) + 1;
    // cells of this window: xc = jb0 - w + 8l in [xlo, xhi], l in [0, G)
    const int xb = jb0 - w;                                 // column of l = 0 (<= xhi)
    const int l_lo = max(0, (xlo - xb + 7) >> 3), l_hi = hasWin ? min(G - 1, (xhi - xb) >> 3) : -1;
    // ---- phase-2 role: lane l of column block cb ----
    const int lane = tid % G, cb = tid / G, NCB = COST_THREADS / G;
    const int CPB = (TX + NCB - 1) / NCB;
    const int c0 = cb * CPB;                                // first output column (tile-local) of this thread
    uint32_t keep[4];
#pragma unroll
    for (int k = 0; k < 4; k++)
        keep[k] = ((lane * 8 + 2 * k < D) ? 0x0000FFFFu : 0u) | ((lane * 8 + 2 * k + 1 < D) ? 0xFFFF0000u : 0u);
    uint32_t acc[COST_CPB_MAX][4];
#pragma unroll
    for (int k = 0; k < COST_CPB_MAX; k++)
#pragma unroll
        for (int q = 0; q < 4; q++) acc[k][q] = 0;
    int slot = 0;
    const int r0 = ya - R, r1 = yb - 1 + R;
    CostWindow<NPL> cur, nxt;
    if (hasWin) load_window<NPL>(cur, RP + (size_t)max(0, min(H - 1, r0)) * NPL * 32 * Wr8, Wr8, w);
    for (int r = r0; r <= r1; r++) {
        const int rc = max(0, min(H - 1, r));
        int16_t *pb = pixbuf + (size_t)((r - r0) & 1) * TXP * PS;
        // ---- phase 1 ----
        {
            const uint4 *LProw = LP + (size_t)rc * NPL * W1;
            for (int l = l_lo; l <= l_hi; l++) {
                const int xc = xb + 8 * l;
                uint32_t p[4];
                pix_cell<NPL>(cur, LProw + xc, W1, p);
                *reinterpret_cast<uint4 *>(pb + (size_t)(xc - (x0 - R)) * PS + l * 8) = make_uint4(p[0], p[1], p[2], p[3]);
            }
            if (hasWin && r < r1) load_window<NPL>(nxt, RP + (size_t)max(0, min(H - 1, r + 1)) * NPL * 32 * Wr8, Wr8, w);
        }
        __syncthreads();
        // ---- phase 2 ----
        const bool emit = r >= ya + R;
        const int y = r - R;
        const uint32_t live = y < zrow ? 0xFFFFFFFFu : 0u;   // MODE_HH4: rows whose window leaves the image stay 0
        if (c0 < TX && x0 + c0 < W1) {
            // pix column of output column c, offset dx: tile-local index c+dx, clamped to the valid columns
            auto pixcol = [&](int ci) -> const int16_t * {
                const int x = max(xlo, min(xhi, x0 - R + ci));
                return pb + (size_t)(x - (x0 - R)) * PS + lane * 8;
            };
            uint32_t hs[4] = {0, 0, 0, 0};
#pragma unroll
            for (int dx = 0; dx < BS; dx++) {
                const uint4 v = *reinterpret_cast<const uint4 *>(pixcol(c0 + dx));
                hs[0] += v.x; hs[1] += v.y; hs[2] += v.z; hs[3] += v.w;   // per-half sums < 2^15: no carry
            }
#pragma unroll
            for (int k = 0; k < COST_CPB_MAX; k++) {
                const int c = c0 + k;
                if (k < CPB && c < TX && x0 + c < W1) {
                    if (k > 0) {   // slide the window one column to the right
                        const uint4 vin = *reinterpret_cast<const uint4 *>(pixcol(c + BS - 1));
                        const uint4 vout = *reinterpret_cast<const uint4 *>(pixcol(c - 1));
                        hs[0] += vin.x - vout.x; hs[1] += vin.y - vout.y; hs[2] += vin.z - vout.z; hs[3] += vin.w - vout.w;
                    }
                    uint4 *cell = reinterpret_cast<uint4 *>(ring + ((size_t)slot * TX + c) * Dp + lane * 8);
                    const uint4 old = *cell;
                    *cell = make_uint4(hs[0], hs[1], hs[2], hs[3]);
                    acc[k][0] += hs[0] - old.x; acc[k][1] += hs[1] - old.y;
                    acc[k][2] += hs[2] - old.z; acc[k][3] += hs[3] - old.w;     // acc >= old per half: no borrow
                    if (emit) {
                        uint4 o;
                        o.x = (acc[k][0] & keep[0] & live) | (cpad & ~keep[0]);
                        o.y = (acc[k][1] & keep[1] & live) | (cpad & ~keep[1]);
                        o.z = (acc[k][2] & keep[2] & live) | (cpad & ~keep[2]);
                        o.w = (acc[k][3] & keep[3] & live) | (cpad & ~keep[3]);
                        *reinterpret_cast<uint4 *>(Cv + ((size_t)y * W1 + x0 + c) * Dp + lane * 8) = o;
                    }
                }
            }
        }
        slot = slot + 1 == BS ? 0 : slot + 1;
        cur = nxt;
    }
}

struct CostGeom { int TXP, TX, ntiles, rows_per_seg, nseg; size_t smem; };
static CostGeom cost_geom(const SgbmPlan &pl)
{
    CostGeom g;
    const int Gc = pl.Dp / 8, NCB = COST_THREADS / Gc, BS = 2 * pl.R + 1;
    const size_t budget = 200 * 1024;
    auto bytes = [&](int txp) { return (size_t)2 * ((size_t)2 * txp * (pl.Dp + 8) + (size_t)BS * (txp - 2 * pl.R) * pl.Dp); };
    // widest tile whose windows fit the CTA (TXP + 8(G-1) <= threads), whose column blocks fit COST_CPB_MAX and
    // whose buffers fit shared memory
    int TXP = std::min(COST_THREADS - 8 * (Gc - 1), NCB * COST_CPB_MAX + 2 * pl.R);
    TXP = std::min(TXP, pl.W1 + 2 * pl.R);
    while (TXP > 2 * pl.R + 1 && bytes(TXP) > budget) TXP--;
    g.TXP = TXP; g.TX = TXP - 2 * pl.R;
    g.ntiles = (pl.W1 + g.TX - 1) / g.TX;
    g.nseg = std::max(1, std::min(8, pl.H / 32));
    g.rows_per_seg = (pl.H + g.nseg - 1) / g.nseg;
    g.nseg = (pl.H + g.rows_per_seg - 1) / g.rows_per_seg;
    g.smem = bytes(TXP);
    return g;
}

template <int NPL, int R>
static int launch_cost_fused(const SgbmPlan &pl, const CostGeom &g, char *ws, size_t ws_stride, int B, cudaStream_t st)
{
    auto kern = k_cost_fused<NPL, R>;
    static size_t configured = 0;
    if (g.smem > configured) {
        SVB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
        configured = g.smem;
    }
    const int jb0 = pl.W - 1 - pl.minX1 + pl.p.minDisparity;
    const uint32_t cpad = (uint32_t)(32767 - pl.p.P2) * 0x10001u;
    dim3 grid(g.ntiles, g.nseg, B);
    kern<<<grid, COST_THREADS, g.smem, st>>>((const uint4 *)(ws + pl.off_lplanes), (const int16_t *)(ws + pl.off_rplanes),
                                             (int16_t *)(ws + pl.off_C), pl.H, pl.W1, pl.Wr, pl.Dp, pl.D, pl.Dp / 8, jb0,
                                             g.TXP, g.rows_per_seg, cpad, pl.p.mode == 3 ? pl.H - pl.R : pl.H, ws_stride);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

template <int NPL>
static int launch_cost_r(const SgbmPlan &pl, const CostGeom &g, char *ws, size_t ws_stride, int B, cudaStream_t st)
{
    switch (pl.R) {
    case 0: return launch_cost_fused<NPL, 0>(pl, g, ws, ws_stride, B, st);
    case 1: return launch_cost_fused<NPL, 1>(pl, g, ws, ws_stride, B, st);
    case 2: return launch_cost_fused<NPL, 2>(pl, g, ws, ws_stride, B, st);
    case 3: return launch_cost_fused<NPL, 3>(pl, g, ws, ws_stride, B, st);
    case 4: return launch_cost_fused<NPL, 4>(pl, g, ws, ws_stride, B, st);
    default: return launch_cost_fused<NPL, 5>(pl, g, ws, ws_stride, B, st);
    }
}

int svb_launch_cost(const SgbmPlan &pl, char *ws, size_t ws_stride, int B, cudaStream_t st)
{
    const CostGeom g = cost_geom(pl);
    SVB_REQUIRE(g.smem <= 227 * 1024, SVB_ERR_UNSUPPORTED, "sgbm: cost tile does not fit shared memory (%zu bytes)", g.smem);
    return pl.nplanes == 2 ? launch_cost_r<2>(pl, g, ws, ws_stride, B, st) : launch_cost_r<6>(pl, g, ws, ws_stride, B, st);
}
