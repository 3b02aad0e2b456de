// svb_cost.cu — matching cost: preFilterCap-clipped Sobel-x + intensity planes, Birchfield-Tomasi sampling,
// blockSize box sum.  Semantics: SURVEY.md §3.2.2-3.2.3.  Output: C[y][xi][Dp] int16 (packed pairs, d fastest).
//
//   k_planes : u8 images -> per-pixel BT operands, already widened to packed s16x2 so the cost kernel's inner
//              loop is pure DPX/VIADD work:
//                left  LP[y][pl][xi]          = uint4 { (u,u), (255-u,255-u), (u0,u0), (255-u1,255-u1) }
//                right RP[y][pl][copy 8][arr 4][Wr+8] = int16 arrays V, 255-V, V0, 255-V1 stored REVERSED in x
//                                               (index j = W-1-x); copy c holds element j at position j-c+8, so a
//                                               run of 8 disparities starting at any j is one aligned LDG.128.
//   k_cost_h : BT pixel cost for 8 disparities per thread (4 packed registers), running horizontal box sum
//              along x in registers -> Hs[y][xi][Dp]
//   k_cost_v : running vertical box sum of Hs -> C, padded disparities overwritten with CPAD
//
// BT in biased form (no negative operands):  with a' = 255-a,
//   c0+255 = max(255, u+v1', v0+u')   c1+255 = max(255, v+u1', u0+v')   cost = max(min(X,Y)-255, 0)
//   X = max(u+v1', v0+u'), Y = max(v+u1', u0+v')            -> 2 VIADD + 2 VIADDMNMX + 1 VIMNMX + 1 VIADDMNMX.RELU
#include "svb_common.cuh"

#define COST_TXR 32   // x positions per running-sum run

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

// pixel cost for 8 disparities (4 packed regs) at cost column xc of row y
template <int NPL>
__device__ __forceinline__ void pix8(const uint4 *__restrict__ LProw, const int16_t *__restrict__ RProw, int W1, int Wr8,
                                     int xc, int jb0, int lane8, uint32_t (&acc)[4])
{
    // reversed right index of disparity 0 of this lane: jb = (W-1-x+minD) with x = xc+minX1  ->  jb0 - xc
    const int j0 = jb0 - xc + lane8;
    const int cpy = j0 & 7;                       // copy whose position of j0 is 16-byte aligned
    const int16_t *rb = RProw + (size_t)cpy * 4 * Wr8 + (j0 - cpy + 8);
#pragma unroll
    for (int i = 0; i < 4; i++) acc[i] = 0;
#pragma unroll
    for (int pl = 0; pl < NPL; pl++) {
        const uint4 Lw = __ldg(LProw + (size_t)pl * W1 + xc);
        const int16_t *rp = rb + (size_t)pl * 32 * Wr8;
        const uint4 V = __ldg(reinterpret_cast<const uint4 *>(rp));
        const uint4 Vn = __ldg(reinterpret_cast<const uint4 *>(rp + Wr8));
        const uint4 V0 = __ldg(reinterpret_cast<const uint4 *>(rp + 2 * Wr8));
        const uint4 V1n = __ldg(reinterpret_cast<const uint4 *>(rp + 3 * Wr8));
        const uint32_t v[4] = {V.x, V.y, V.z, V.w}, vn[4] = {Vn.x, Vn.y, Vn.z, Vn.w};
        const uint32_t v0[4] = {V0.x, V0.y, V0.z, V0.w}, v1n[4] = {V1n.x, V1n.y, V1n.z, V1n.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t Bv = __vadd2(v0[i], Lw.y);                 // v0 + (255-u)
            const uint32_t X = __viaddmax_s16x2(Lw.x, v1n[i], Bv);    // max(u + 255-v1, .)
            const uint32_t Ev = __vadd2(Lw.z, vn[i]);                 // u0 + (255-v)
            const uint32_t Y = __viaddmax_s16x2(v[i], Lw.w, Ev);      // max(v + 255-u1, .)
            uint32_t c = __viaddmax_s16x2_relu(__vmins2(X, Y), 0xFF01FF01u, 0u);   // max(min-255, 0)
            if (pl & 1) c = (c >> 2) & 0x3FFF3FFFu;
            acc[i] = __vadd2(acc[i], c);
        }
    }
}

// grid: (ceil(W1/TXC), H, B); block: 256 threads = G lanes x NRUN runs; run = COST_TXR consecutive columns
template <int NPL, int R>
__global__ void __launch_bounds__(256) k_cost_h(const uint4 *__restrict__ LP, const int16_t *__restrict__ RP,
                                                int16_t *__restrict__ Hs, int H, int W1, int Wr, int Dp, int G, int jb0,
                                                size_t ws_stride)
{
    constexpr int BS = 2 * R + 1;
    const int y = blockIdx.y, b = blockIdx.z;
    const int lane = threadIdx.x % G, run = threadIdx.x / G;
    const int nrun = 256 / G;
    const int xa = (blockIdx.x * nrun + run) * COST_TXR;
    if (xa >= W1) return;
    const int xe = min(xa + COST_TXR, W1);
    LP = reinterpret_cast<const uint4 *>(reinterpret_cast<const char *>(LP) + (size_t)b * ws_stride);
    RP = reinterpret_cast<const int16_t *>(reinterpret_cast<const char *>(RP) + (size_t)b * ws_stride);
    Hs = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(Hs) + (size_t)b * ws_stride);
    const uint4 *LProw = LP + (size_t)y * NPL * W1;
    const int Wr8 = Wr + 8;
    const int16_t *RProw = RP + (size_t)y * NPL * 32 * Wr8;
    uint4 *out = reinterpret_cast<uint4 *>(Hs + ((size_t)y * W1) * Dp + lane * 8);
    const size_t ostride = Dp / 8;   // uint4 per column
    uint32_t ring[BS][4];
    uint32_t sum[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < BS; i++)
#pragma unroll
        for (int k = 0; k < 4; k++) ring[i][k] = 0;
    for (int k0 = xa - R; k0 < xe + R; k0 += BS) {
#pragma unroll
        for (int i = 0; i < BS; i++) {
            const int k = k0 + i;
            const int xc = max(0, min(W1 - 1, k));
            uint32_t p[4];
            pix8<NPL>(LProw, RProw, W1, Wr8, xc, jb0, lane * 8, p);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                sum[q] = sum[q] + p[q] - ring[i][q];   // per-half: no carry/borrow (sum>=ring, totals < 2^15)
                ring[i][q] = p[q];
            }
            const int xo = k - R;
            if (xo >= xa && xo < xe) out[(size_t)xo * ostride] = make_uint4(sum[0], sum[1], sum[2], sum[3]);
        }
    }
}

// vertical running sum; thread = (column xi, 8 disparities); band of TY rows per CTA in y
#define COST_TY 40
template <int R>
__global__ void __launch_bounds__(256) k_cost_v(const int16_t *__restrict__ Hs, int16_t *__restrict__ Cv, int H, int W1,
                                                int Dp, int D, uint32_t cpad, size_t ws_stride)
{
    const int b = blockIdx.z;
    const size_t rowq = (size_t)W1 * Dp / 8;   // uint4 per row
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= rowq) return;
    const uint4 *in = reinterpret_cast<const uint4 *>(reinterpret_cast<const char *>(Hs) + (size_t)b * ws_stride) + t;
    uint4 *out = reinterpret_cast<uint4 *>(reinterpret_cast<char *>(Cv) + (size_t)b * ws_stride) + t;
    const int d0 = (int)(t % (Dp / 8)) * 8;
    uint32_t keep[4];
#pragma unroll
    for (int k = 0; k < 4; k++)
        keep[k] = ((d0 + 2 * k < D) ? 0x0000FFFFu : 0u) | ((d0 + 2 * k + 1 < D) ? 0xFFFF0000u : 0u);
    const int y0 = blockIdx.y * COST_TY, y1 = min(y0 + COST_TY, H);
    uint32_t v[4] = {0, 0, 0, 0};
#pragma unroll
    for (int dy = -R; dy <= R; dy++) {
        const uint4 a = __ldg(in + (size_t)max(0, min(H - 1, y0 + dy)) * rowq);
        v[0] += a.x; v[1] += a.y; v[2] += a.z; v[3] += a.w;
    }
    for (int y = y0; y < y1; y++) {
        uint4 o;
        o.x = (v[0] & keep[0]) | (cpad & ~keep[0]);
        o.y = (v[1] & keep[1]) | (cpad & ~keep[1]);
        o.z = (v[2] & keep[2]) | (cpad & ~keep[2]);
        o.w = (v[3] & keep[3]) | (cpad & ~keep[3]);
        out[(size_t)y * rowq] = o;
        const uint4 a = __ldg(in + (size_t)min(H - 1, y + R + 1) * rowq);
        const uint4 s = __ldg(in + (size_t)max(0, y - R) * rowq);
        v[0] = v[0] + a.x - s.x; v[1] = v[1] + a.y - s.y; v[2] = v[2] + a.z - s.z; v[3] = v[3] + a.w - s.w;
    }
}

template <int NPL>
static void launch_cost_h(int R, dim3 grid, cudaStream_t st, const uint4 *LP, const int16_t *RP, int16_t *Hs, int H,
                          int W1, int Wr, int Dp, int G, int jb0, size_t ws_stride)
{
    switch (R) {
    case 0: k_cost_h<NPL, 0><<<grid, 256, 0, st>>>(LP, RP, Hs, H, W1, Wr, Dp, G, jb0, ws_stride); break;
    case 1: k_cost_h<NPL, 1><<<grid, 256, 0, st>>>(LP, RP, Hs, H, W1, Wr, Dp, G, jb0, ws_stride); break;
    case 2: k_cost_h<NPL, 2><<<grid, 256, 0, st>>>(LP, RP, Hs, H, W1, Wr, Dp, G, jb0, ws_stride); break;
    case 3: k_cost_h<NPL, 3><<<grid, 256, 0, st>>>(LP, RP, Hs, H, W1, Wr, Dp, G, jb0, ws_stride); break;
    case 4: k_cost_h<NPL, 4><<<grid, 256, 0, st>>>(LP, RP, Hs, H, W1, Wr, Dp, G, jb0, ws_stride); break;
    default: k_cost_h<NPL, 5><<<grid, 256, 0, st>>>(LP, RP, Hs, H, W1, Wr, Dp, G, jb0, ws_stride); break;
    }
}

int svb_launch_cost(const SgbmPlan &pl, char *ws, size_t ws_stride, int B, cudaStream_t st)
{
    const uint4 *LP = (const uint4 *)(ws + pl.off_lplanes);
    const int16_t *RP = (const int16_t *)(ws + pl.off_rplanes);
    int16_t *Hs = (int16_t *)(ws + pl.off_S);   // S volume doubles as the horizontal-sum scratch
    int16_t *Cv = (int16_t *)(ws + pl.off_C);
    const int Gc = pl.Dp / 8;   // cost threads hold 8 disparities each
    const int nrun = 256 / Gc;
    const int jb0 = pl.W - 1 - pl.minX1 + pl.p.minDisparity;
    dim3 grid((pl.W1 + nrun * COST_TXR - 1) / (nrun * COST_TXR), pl.H, B);
    if (pl.nplanes == 2)
        launch_cost_h<2>(pl.R, grid, st, LP, RP, Hs, pl.H, pl.W1, pl.Wr, pl.Dp, Gc, jb0, ws_stride);
    else
        launch_cost_h<6>(pl.R, grid, st, LP, RP, Hs, pl.H, pl.W1, pl.Wr, pl.Dp, Gc, jb0, ws_stride);
    SVB_CUDA_CHECK(cudaGetLastError());
    const size_t rowq = (size_t)pl.W1 * pl.Dp / 8;
    dim3 g2((unsigned)((rowq + 255) / 256), (pl.H + COST_TY - 1) / COST_TY, B);
    const uint32_t cpad = (uint32_t)(32767 - pl.p.P2) * 0x10001u;
    switch (pl.R) {
    case 0: k_cost_v<0><<<g2, 256, 0, st>>>(Hs, Cv, pl.H, pl.W1, pl.Dp, pl.D, cpad, ws_stride); break;
    case 1: k_cost_v<1><<<g2, 256, 0, st>>>(Hs, Cv, pl.H, pl.W1, pl.Dp, pl.D, cpad, ws_stride); break;
    case 2: k_cost_v<2><<<g2, 256, 0, st>>>(Hs, Cv, pl.H, pl.W1, pl.Dp, pl.D, cpad, ws_stride); break;
    case 3: k_cost_v<3><<<g2, 256, 0, st>>>(Hs, Cv, pl.H, pl.W1, pl.Dp, pl.D, cpad, ws_stride); break;
    case 4: k_cost_v<4><<<g2, 256, 0, st>>>(Hs, Cv, pl.H, pl.W1, pl.Dp, pl.D, cpad, ws_stride); break;
    default: k_cost_v<5><<<g2, 256, 0, st>>>(Hs, Cv, pl.H, pl.W1, pl.Dp, pl.D, cpad, ws_stride); break;
    }
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}
