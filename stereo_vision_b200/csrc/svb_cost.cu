// svb_cost.cu — matching cost: preFilterCap-clipped Sobel-x + intensity planes, Birchfield-Tomasi sampling,
// blockSize box sum.  Semantics: SURVEY.md §3.2.2-3.2.3.  Output: C[y][xi][Dp] int16 (packed pairs, d fastest).
//
//   k_planes : u8 images -> per-pixel BT operands, already widened to packed s16x2 so the cost kernel's inner
//              loop is pure DPX/VIADD work:
//                left  LP[y][pl][xi]          = bytes (u, 255-u, u0, 255-u1) of plane pl of channel pl/2: LP[y][ch][xi] = uint2 {gradient, intensity};
//                                               the cost kernel stages a row of them in shared memory and widens them
//                                               to (u,u), (255-u,255-u), (u0+255,u0+255), (255-u1,255-u1) per cell
//                right RP[y][pl][arr 3][w < Wr] = uint4: elements j = w..w+7 of the int16 arrays V, V0, 255-V1
//                                               stored REVERSED in x (index j = W-1-x): the 8 disparities of any
//                                               window w are one aligned LDG.128, and consecutive windows (consecutive
//                                               lanes of the cost kernel) are consecutive 16-byte words.
//   k_cost_fused : BT pixel cost for 8 disparities per cell (4 packed registers), 2R+1 box sum in x through a
//              shared-memory pix row, running box sum in y through a shared-memory ring -> C (written once),
//              padded disparities overwritten with CPAD
//
// BT in biased form (no negative operands):  with a' = 255-a,
//   c0+255 = max(255, u+v1', v0+u')   c1+255 = max(255, v+u1', u0+v')   cost = max(min(X,Y)-255, 0)
//   X = max(u+v1', v0+u'), Y = max(v+u1', u0+v')            -> 2 VIADD + 2 VIADDMNMX + 1 VIMNMX + 1 VIADDMNMX.RELU
//   (u0+v' is evaluated as (u0+255) - v with the 255 folded into the stored left operand: the right image needs no
//   255-V array)
#include "svb_common.cuh"
#include <algorithm>
#include <cstdlib>


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
// written once per window: RP[y][plane][array a][w] = elements w..w+7, so the cost kernel reads every right operand
// with one LDG.128 per lane, 128 contiguous bytes per quarter warp.
#define PLANES_THREADS 256
__global__ void __launch_bounds__(PLANES_THREADS) k_planes(const uint8_t *__restrict__ left, const uint8_t *__restrict__ right,
                                                           int H, int W, int C, size_t pitch, size_t istride, int ftzero,
                                                           int minX1, int W1, int Wr, uint32_t *__restrict__ LP,
                                                           int16_t *__restrict__ RP, size_t ws_stride)
{
    extern __shared__ __align__(16) unsigned char psm[];
    const int Wa = Wr + 16;                                   // staged array length (8 zeros in front, >= 8 behind)
    int16_t *A = reinterpret_cast<int16_t *>(psm);            // [3][Wa]: V, V0, 255-V1
    uint8_t *pvL = psm + (size_t)3 * Wa * 2;                  // [W]
    uint8_t *pvR = pvL + ((W + 15) & ~15);                    // [W]
    const int y = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const int NPL = 2 * C;
    left += (size_t)b * istride;
    right += (size_t)b * istride;
    LP = reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(LP) + (size_t)b * ws_stride);
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
            // one word per column and plane; the two planes of a channel are the .x / .y of one uint2
            LP[(((size_t)y * (NPL / 2) + (pl >> 1)) * W1 + xi) * 2 + (pl & 1)] =
                (uint32_t)p | ((uint32_t)(255 - p) << 8) | ((uint32_t)lo << 16) | ((uint32_t)(255 - hi) << 24);
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
            A[i] = (int16_t)v; A[Wa + i] = (int16_t)v0; A[2 * Wa + i] = (int16_t)v1n;
        }
        __syncthreads();
        // one uint4 per window (8 consecutive reversed elements), consecutive windows by consecutive threads:
        // 5 aligned words from the staged array, shifted by half a word for odd windows
        uint4 *outp = reinterpret_cast<uint4 *>(RP) + ((size_t)y * NPL + pl) * 3 * Wr;
        for (int i = tid; i < 3 * Wr; i += PLANES_THREADS) {
            const int a = i / Wr, w = i - a * Wr;
            const uint32_t *src = reinterpret_cast<const uint32_t *>(A + (size_t)a * Wa) + ((w + 8) >> 1);
            const uint32_t sel = (w & 1) ? 0x5432u : 0x3210u;
            const uint32_t w0 = src[0], w1 = src[1], w2 = src[2], w3 = src[3], w4 = src[4];
            outp[i] = make_uint4(__byte_perm(w0, w1, sel), __byte_perm(w1, w2, sel), __byte_perm(w2, w3, sel),
                                 __byte_perm(w3, w4, sel));
        }
    }
}

int svb_launch_planes(const SgbmPlan &pl, const uint8_t *left, const uint8_t *right, size_t pitch, size_t istride,
                      char *ws, size_t ws_stride, int B, cudaStream_t st)
{
    dim3 grid(pl.H, B);
    const size_t smem = (size_t)3 * (pl.Wr + 16) * 2 + 2 * (size_t)((pl.W + 15) & ~15);
    if (smem > 48 * 1024)   // per device and function: set on every launch (cheap), correct with several devices
        SVB_CUDA_CHECK(cudaFuncSetAttribute(k_planes, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_planes<<<grid, PLANES_THREADS, smem, st>>>(left, right, pl.H, pl.W, pl.ch, pitch, istride, pl.ftzero, pl.minX1,
                                                 pl.W1, pl.Wr, (uint32_t *)(ws + pl.off_lplanes),
                                                 (int16_t *)(ws + pl.off_rplanes), ws_stride);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

// ---- fused cost kernel: BT pixel cost -> horizontal box sum -> vertical running box sum -> C ----
// The right-image operands of a cell (column x, disparities 8l..8l+7) are the 8 reversed pixels starting at window
// w = jb0 - x + 8l: all cells on the diagonal x - 8l = const share them.  A CTA therefore owns a PARALLELOGRAM of the
// (x, d) plane: NW consecutive windows x all G = Dp/8 disparity groups, i.e. for group l the NW columns
// x = T0 - R + 8l + ci, ci in [0, NW).  Every window has exactly G cells, so phase 1 is perfectly balanced, and for a
// fixed l the tile is a contiguous run of columns, which is all the box sums need.  Per input row r (clamped to the
// image: cost-space border replication):
//   phase 1: thread (ci, part) holds window ci's operands in registers (3 arrays x NPL planes, aligned LDG.128 from
//            copy w%8, prefetched one row ahead) and computes its CPT = G/NPART cells l = part*CPT.., parking the pixel
//            costs in a double-buffered shared-memory row pix[ci][l] (column stride Dp+8: conflict-free).
//   phase 2: lane l of column block cb owns CPB adjacent output columns of group l: running sum of the 2R+1
//            neighbouring pix columns (clamped per lane to the image: border replication), swapped into the cell's slot
//            of a (2R+1)-row ring in shared memory, vertical running sum acc += Hs(r) - Hs(r-(2R+1)); from the
//            (2R+1)-th row on it writes C[r-R][x][8l..8l+8), x = T0 + 8l + c.
// The horizontal sums never touch HBM; C is written exactly once (2 B per pixel-disparity).
#define COST_ROLE 256    // threads per role (phase 1 / phase 2)
#define COST_CPB_MAX 8   // output columns per thread (phase 2)
#define COST_CPT_MAX 8   // cells per thread (phase 1)
#ifndef COST_P2_COLS
#define COST_P2_COLS 4   // columns per phase-2 pass (register footprint of the consumer role)
#endif

template <int NPL>
struct CostWindow { uint4 v[NPL][3]; };   // [plane][V, V0, 255-V1]

// rb: this thread's window in copy w%8 of plane 0 / array 0 of the row; (plane, array) strides are CTA-uniform
template <int NPL>
__device__ __forceinline__ void load_window(CostWindow<NPL> &cw, const char *__restrict__ rb, uint32_t arr_bytes)
{
#pragma unroll
    for (int pl = 0; pl < NPL; pl++)
#pragma unroll
        for (int a = 0; a < 3; a++)
            cw.v[pl][a] = __ldg(reinterpret_cast<const uint4 *>(rb + (size_t)((pl * 3 + a) * arr_bytes)));
}

__device__ __forceinline__ uint32_t u4get(const uint4 &q, int i) { return i == 0 ? q.x : i == 1 ? q.y : i == 2 ? q.z : q.w; }

__device__ __forceinline__ uint4 ldg_nc_pred(const void *p, bool pred)
{
    uint4 v;   // volatile: the asm loads/stores keep program order, which is how the kernel software-pipelines them
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %5, 0;\n\t@p ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];\n\t}"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "r"((int)pred));
    return v;
}
__device__ __forceinline__ void sts128_pred(uint32_t saddr, const uint32_t (&v)[4], bool pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %5, 0;\n\t@p st.shared.v4.u32 [%0], {%1, %2, %3, %4};\n\t}"
                 ::"r"(saddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"((int)pred) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t saddr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t saddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// BT cost of one cell (8 disparities = 4 packed registers) summed over the planes; plain `+` where no half can carry
// The plain 32-bit adds / subtracts between packed halves (no carry or borrow by construction) are written as
// x * one + y / x * mone + y with one == 1, mone == 0xFFFFFFFF opaque kernel arguments: ptxas keeps them as IMADs on
// the FMA pipe instead of IADD3 / VIADD on the ALU pipe, which the packed min/max instructions need (COST_FMA).
#ifndef COST_FMA
#define COST_FMA 1
#endif
#if COST_FMA
#define CADD(x, y, one) ((x) * (one) + (y))
#define CSUB(y, x, mone) ((x) * (mone) + (y))          // y - x
#else
#define CADD(x, y, one) ((x) + (y))
#define CSUB(y, x, mone) ((y) - (x))
#endif
template <int NPL>
__device__ __forceinline__ void pix_cell(const CostWindow<NPL> &cw, const uint4 (&Lw)[NPL], uint32_t (&acc)[4],
                                         uint32_t one, uint32_t mone)
{
#pragma unroll
    for (int pl = 0; pl < NPL; pl++) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t Bv = CADD(u4get(cw.v[pl][1], i), Lw[pl].y, one);               // v0 + (255-u)
            const uint32_t X = __viaddmax_s16x2(Lw[pl].x, u4get(cw.v[pl][2], i), Bv);     // max(u + 255-v1, .)
            const uint32_t Ev = CSUB(Lw[pl].z, u4get(cw.v[pl][0], i), mone);              // (u0+255) - v: no borrow
            const uint32_t Y = __viaddmax_s16x2(u4get(cw.v[pl][0], i), Lw[pl].w, Ev);     // max(v + 255-u1, .)
            uint32_t c = __viaddmax_s16x2_relu(__vmins2(X, Y), 0xFF01FF01u, 0u);          // max(min-255, 0)
#if COST_FMA
            // c >> 2 per half (c <= 255 per half): c * 64 puts half h's bits 2..7 into byte 2h+1 — one IMAD on the FMA
            // pipe + one PRMT instead of SHF + LOP3 on the ALU pipe
            if (pl & 1) c = __byte_perm(c * (one << 6), 0u, 0x4341);
#else
            if (pl & 1) c = (c >> 2) & 0x3FFF3FFFu;
#endif
            acc[i] = pl ? CADD(c, acc[i], one) : c;
        }
    }
}

struct CostArgs {
    const uint2 *LP; const int16_t *RP; int16_t *Cv;
    int H, W1, Wr, Dp, D, G, jb0, NW, NPART, CPT, NCB, rows_per_seg, zrow, y0, y1;
    uint32_t cpad;
    uint32_t one, mone;      // == 1, == 0xFFFFFFFF (opaque multipliers: x * one + y stays an IMAD on the FMA pipe)
    size_t ws_stride;
    // phase-2 job of every thread: (column block << 8) | disparity group, 0xFFFF = none.  Threads are assigned so that
    // the 8 lanes of a quarter warp own 8 consecutive groups at the SAME columns (see cost_jobs)
    uint16_t job[COST_ROLE];
};

// phase 2 for one thread, one input row, NC adjacent output columns starting at tile column c0 + kofs.  One code path
// for every lane (border lanes clamp their pix columns and mask their stores): all consumer warps take the same time,
// which is what the producer/consumer pipeline wants.  All shared-memory reads are issued up front (the asm loads keep
// program order): the ring cells that leave the window, then the NC + BS - 1 pix columns.
struct CostP2 {
    uint32_t pbs;        // pix row buffer (shared address) + lane*16
    uint32_t PS2;        // pix column stride in bytes
    uint32_t ringcell;   // ring slot cell of column c0 (shared address, + lane*16)
    uint32_t Dp2;
    int c0, cilo, cihi, lane, D;
    uint32_t cmask, cpad, one, mone;
    bool live, padded;
};
__device__ __forceinline__ uint4 lds128_pred(uint32_t saddr, bool pred)
{
    uint4 v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %5, 0;\n\t@p ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];\n\t}"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr), "r"((int)pred));
    return v;
}
__device__ __forceinline__ void sts128_pred4(uint32_t saddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d, bool pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %5, 0;\n\t@p st.shared.v4.u32 [%0], {%1, %2, %3, %4};\n\t}"
                 ::"r"(saddr), "r"(a), "r"(b), "r"(c), "r"(d), "r"((int)pred) : "memory");
}
__device__ __forceinline__ void stg128_pred(void *p, uint32_t a, uint32_t b, uint32_t c, uint32_t d, bool pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %5, 0;\n\t@p st.global.v4.u32 [%0], {%1, %2, %3, %4};\n\t}"
                 ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d), "r"((int)pred) : "memory");
}
template <int BS, int NC, bool EMIT>
__device__ __forceinline__ void cost_phase2(const CostP2 &q, int kofs, int16_t *cout, uint32_t (*acc)[4])
{
    const uint32_t cmask = q.cmask >> kofs;
    const uint32_t ringcell = q.ringcell + (uint32_t)kofs * q.Dp2;
    uint4 old[NC];
#pragma unroll
    for (int k = 0; k < NC; k++) old[k] = lds128_pred(ringcell + (uint32_t)k * q.Dp2, cmask >> k & 1);
    uint4 pv[NC + BS - 1];
#pragma unroll
    for (int i = 0; i < NC + BS - 1; i++)
        pv[i] = lds128(q.pbs + (uint32_t)min(max(q.c0 + kofs + i, q.cilo), q.cihi) * q.PS2);
    const uint32_t one = q.one, mone = q.mone;
    uint32_t hs[4] = {pv[0].x, pv[0].y, pv[0].z, pv[0].w};
#pragma unroll
    for (int dx = 1; dx < BS; dx++) {
        hs[0] = CADD(pv[dx].x, hs[0], one); hs[1] = CADD(pv[dx].y, hs[1], one);
        hs[2] = CADD(pv[dx].z, hs[2], one); hs[3] = CADD(pv[dx].w, hs[3], one);
    }
#pragma unroll
    for (int k = 0; k < NC; k++) {
        if (k > 0) {   // slide the window one column to the right (per-half sums < 2^15: no carry / borrow)
            hs[0] = CSUB(CADD(pv[k + BS - 1].x, hs[0], one), pv[k - 1].x, mone);
            hs[1] = CSUB(CADD(pv[k + BS - 1].y, hs[1], one), pv[k - 1].y, mone);
            hs[2] = CSUB(CADD(pv[k + BS - 1].z, hs[2], one), pv[k - 1].z, mone);
            hs[3] = CSUB(CADD(pv[k + BS - 1].w, hs[3], one), pv[k - 1].w, mone);
        }
        const bool v = cmask >> k & 1;      // columns that are not real keep meaningless sums and store nothing
        sts128_pred4(ringcell + (uint32_t)k * q.Dp2, hs[0], hs[1], hs[2], hs[3], v);
        acc[k][0] = CSUB(CADD(hs[0], acc[k][0], one), old[k].x, mone);
        acc[k][1] = CSUB(CADD(hs[1], acc[k][1], one), old[k].y, mone);
        acc[k][2] = CSUB(CADD(hs[2], acc[k][2], one), old[k].z, mone);
        acc[k][3] = CSUB(CADD(hs[3], acc[k][3], one), old[k].w, mone);     // acc + hs >= old per half: no borrow
        if (EMIT) {
            char *cp = reinterpret_cast<char *>(cout) + (size_t)k * q.Dp2;
            if (q.padded) {                                     // padded disparities carry CPAD (CTA-uniform branch)
                uint32_t o[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint32_t keep = ((q.lane * 8 + 2 * j < q.D) ? 0x0000FFFFu : 0u) |
                                          ((q.lane * 8 + 2 * j + 1 < q.D) ? 0xFFFF0000u : 0u);
                    o[j] = (acc[k][j] & keep) | (q.cpad & ~keep);
                }
                stg128_pred(cp, o[0], o[1], o[2], o[3], v);
            } else {
                // straight from the running sums: they are not written again before the next row's phase 2, so the
                // store's source registers are not what the next row's phase 1 starts writing into
                stg128_pred(cp, acc[k][0], acc[k][1], acc[k][2], acc[k][3], v);
            }
        }
    }
}
template <int BS, bool EMIT>
__device__ __forceinline__ void cost_phase2_row(const CostP2 &q, int16_t *cout, uint32_t (&acc)[COST_CPB_MAX][4])
{
    constexpr int NC = COST_P2_COLS;
    if (EMIT) asm volatile("" : "+l"(cout));   // one pointer for the row: keep it out of the per-column blocks
#pragma unroll
    for (int h = 0; h < COST_CPB_MAX / NC; h++)
        cost_phase2<BS, NC, EMIT>(q, h * NC, reinterpret_cast<int16_t *>(reinterpret_cast<char *>(cout) + (size_t)(h * NC) * q.Dp2),
                                  acc + h * NC);
}

// left operands: one word per column and plane (bytes u, 255-u, u0, 255-u1), widened to duplicated s16x2 per cell
__device__ __forceinline__ uint4 widen_left(uint32_t w)
{
    return make_uint4(__byte_perm(w, 0u, 0x4040), __byte_perm(w, 0u, 0x4141), __byte_perm(w, 0u, 0x4242) + 0x00FF00FFu,
                      __byte_perm(w, 0u, 0x4343));
}
__device__ __forceinline__ uint2 lds64(uint32_t saddr)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(saddr));
    return v;
}
__device__ __forceinline__ void cp_async8(uint32_t saddr, const void *gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(saddr), "l"(gmem) : "memory");
}

struct CostTile {   // per-CTA geometry shared by both roles
    int G, Dp, NW, TX, W1, H, PS, T0, ya, yb, r0, r1, NCOL;
    int ylo;            // input rows are clamped to [ylo, H-1] (0 except for the row stripes of MODE_SGBM_3WAY)
    uint32_t smem0, pix_bytes, left0, left_bytes;   // left0: staged left operands [2][NCH][NCOL] uint2
};

// ---- producer role (phase 1): window ci of the tile, cells l = l0 .. l0+CPT-1 ----
// Left operands: the row's NCOL = NW + 8(G-1) columns x NCH channels are staged in shared memory one row ahead with
// cp.async (8 bytes per column and channel); every cell reads its word pair with one LDS.64 and widens it.  No
// left-operand register is live across cells and no global load sits on the cells' critical path (ptxas sinks
// ld.global.nc to its first use when registers are short: the former register double buffer had no distance left).
template <int NPL, int R>
struct CostProducer {
    static constexpr int NCH = NPL / 2;
    // the window registers hold ONE channel (gradient + intensity plane) at a time; for 3-channel input the pixel
    // cost accumulates in the thread's own shared-memory cells, and the next chunk's operands load while one computes
    CostWindow<2> cur;
    const uint2 *LPb;          // this frame's left operands
    const char *rb0;
    size_t rrow_bytes;
    uint32_t arr_bytes, sts0, vmask, lds0, one, mone;
    int tid;
    bool work;

    // stage the left operands of image row rc into buffer lb (all threads; columns clamped to the image)
    __device__ __forceinline__ void stage(const CostTile &t, int rc, int lb) const
    {
        const uint2 *src = LPb + (size_t)rc * NCH * t.W1;
        const uint32_t dst = t.left0 + (uint32_t)lb * t.left_bytes;
        for (int i = tid; i < NCH * t.NCOL; i += COST_ROLE) {
            const int ch = NCH == 1 ? 0 : i / t.NCOL, col = i - ch * t.NCOL;   // (no runtime division for gray input: 14.36 -> 14.01 ms)
            const int x = min(max(t.T0 - R + col, 0), t.W1 - 1);
            cp_async8(dst + 8u * (uint32_t)i, src + (size_t)ch * t.W1 + x);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    __device__ __forceinline__ void init(const CostArgs &a, const CostTile &t, int tid_, int b)
    {
        static_assert(COST_CPT_MAX == 8, "phase 1 is written for 8 cells per thread");
        tid = tid_; one = a.one; mone = a.mone;
        LPb = reinterpret_cast<const uint2 *>(reinterpret_cast<const char *>(a.LP) + (size_t)b * a.ws_stride);
        const char *RP = reinterpret_cast<const char *>(a.RP) + (size_t)b * a.ws_stride;
        const int ci = tid % t.NW, part = tid / t.NW;
        const bool p1 = part < a.NPART;
        const int l0 = part * a.CPT;
        const int wld = min(max(a.jb0 - (t.T0 - R + ci), 0), a.Wr - 8);   // clamped: out-of-range windows own no valid cell
        const int xf = t.T0 - R + ci + 8 * l0;                            // column of the thread's first cell
        vmask = 0;
#pragma unroll
        for (int k = 0; k < COST_CPT_MAX; k++)
            if (p1 && k < a.CPT && (unsigned)(xf + 8 * k) < (unsigned)t.W1) vmask |= 1u << k;
        work = vmask != 0;
        sts0 = t.smem0 + (uint32_t)(ci * t.PS + l0 * 8) * 2u;
        lds0 = t.left0 + 8u * (uint32_t)(ci + 8 * l0);                    // staged word pair of the first cell, channel 0
        arr_bytes = (uint32_t)a.Wr * 16u;                                 // one array of one plane: Wr windows
        rrow_bytes = (size_t)NPL * 3 * arr_bytes;
        rb0 = RP + (size_t)wld * 16;
        const int rc0 = max(t.ylo, min(t.H - 1, t.r0));
        stage(t, rc0, 0);
        if (work) load_window<2>(cur, rb0 + (size_t)rc0 * rrow_bytes, arr_bytes);
        asm volatile("cp.async.wait_group 0;" ::: "memory");             // made visible by the caller's __syncthreads
    }
    // writes pix row r into pix buffer buf (it == r - r0: the staged left operands of row r sit in buffer it & 1)
    __device__ __forceinline__ void row(const CostTile &t, int r, int buf)
    {
        const int it = r - t.r0;
        if (r < t.r1) stage(t, max(t.ylo, min(t.H - 1, r + 1)), (it + 1) & 1);   // next row's left operands
        const uint32_t sts = sts0 + (uint32_t)buf * t.pix_bytes;
        const uint32_t lds = lds0 + (uint32_t)(it & 1) * t.left_bytes;
#pragma unroll 1
        for (int ch = 0; ch < NCH; ch++) {
            const bool last = ch + 1 == NCH;
            if (work) {
                const uint32_t la = lds + 8u * (uint32_t)(ch * t.NCOL);
                uint2 w[COST_CPT_MAX];
#pragma unroll
                for (int k = 0; k < COST_CPT_MAX; k++) w[k] = lds64(la + 64u * k);   // cell k: 8 columns further (cells k >= CPT
                                                                                     // read the slack behind the staged rows)
#pragma unroll
                for (int k = 0; k < COST_CPT_MAX; k++) {
                    const uint4 Lw[2] = {widen_left(w[k].x), widen_left(w[k].y)};
                    uint32_t p[4];
                    pix_cell<2>(cur, Lw, p, one, mone);
                    if (ch > 0) {
                        const uint4 prev = lds128(sts + 16u * k);        // in-bounds even when the cell is not valid
                        p[0] = CADD(prev.x, p[0], one); p[1] = CADD(prev.y, p[1], one);
                        p[2] = CADD(prev.z, p[2], one); p[3] = CADD(prev.w, p[3], one);
                    }
                    sts128_pred(sts + 16u * k, p, vmask >> k & 1);
                }
            }
            if (work && (!last || r < t.r1)) {                            // next chunk's window
                const int rn = max(t.ylo, min(t.H - 1, last ? r + 1 : r)), chn = last ? 0 : ch + 1;
                load_window<2>(cur, rb0 + (size_t)rn * rrow_bytes + (size_t)(chn * 6) * arr_bytes, arr_bytes);
            }
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");             // next row's left operands landed (this thread's)
    }
};

// ---- consumer role (phase 2): lane l of column block cb, COST_CPB_MAX columns ----
template <int R>
struct CostConsumer {
    static constexpr int BS = 2 * R + 1;
    CostP2 q;
    uint32_t acc[COST_CPB_MAX][4];
    uint32_t pix0, ringcell0, slot_bytes;
    int16_t *cout0;
    size_t crow;
    int slot, zrow;

    __device__ __forceinline__ uint32_t ring0(const CostTile &t) const { return t.smem0 + 2u * t.pix_bytes; }
    __device__ __forceinline__ void init(const CostArgs &a, const CostTile &t, int tid, int b)
    {
        int16_t *Cv = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(a.Cv) + (size_t)b * a.ws_stride);
        const unsigned job = a.job[tid];
        const int lane = job & 0xFF, cb = job >> 8;                        // cb == 0xFF (>= NCB): no phase-2 work
        const int c0 = cb * COST_CPB_MAX;
        const int xl = t.T0 + 8 * lane;                                    // x of output column 0 of this lane
        uint32_t cmask = 0;
#pragma unroll
        for (int k = 0; k < COST_CPB_MAX; k++)
            if (cb < a.NCB && c0 + k < t.TX && (unsigned)(xl + c0 + k) < (unsigned)t.W1) cmask |= 1u << k;
        const uint32_t Dp2 = (uint32_t)t.Dp * 2u;
        q.PS2 = (uint32_t)t.PS * 2u; q.Dp2 = Dp2;
        q.c0 = c0; q.lane = lane; q.D = a.D;
        q.cilo = min(max(R - xl, 0), t.NW - 1); q.cihi = min(max(t.W1 - 1 - xl + R, 0), t.NW - 1);
        q.cmask = cmask; q.cpad = a.cpad; q.padded = a.D != t.Dp; q.one = a.one; q.mone = a.mone;
        pix0 = t.smem0 + (uint32_t)lane * 16u;
        ringcell0 = ring0(t) + (uint32_t)(c0 * t.Dp + lane * 8) * 2u;
        slot_bytes = (uint32_t)t.TX * Dp2;
        cout0 = Cv + ((ptrdiff_t)(xl + c0)) * t.Dp + lane * 8;
        crow = (size_t)t.W1 * Dp2;
        slot = 0; zrow = a.zrow;
#pragma unroll
        for (int k = 0; k < COST_CPB_MAX; k++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[k][j] = 0;
    }
    __device__ __forceinline__ void row(const CostTile &t, int r, int buf)
    {
        // keep the per-thread constants opaque inside the loop: hoisting every clamped column index and mask bit out of
        // it costs ~40 registers (spills); recomputing them per row costs a few ALU instructions
        asm volatile("" : "+r"(q.cilo), "+r"(q.cihi), "+r"(q.cmask), "+r"(q.c0));
        if (q.cmask) {
            const int y = r - R;
            q.pbs = pix0 + (uint32_t)buf * t.pix_bytes;
            q.ringcell = ringcell0 + (uint32_t)slot * slot_bytes;
            q.live = y < zrow;
            int16_t *cout = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(cout0) + (ptrdiff_t)y * (ptrdiff_t)crow);
            if (r < t.ya + R) cost_phase2_row<BS, false>(q, cout, acc);
            else if (q.live) cost_phase2_row<BS, true>(q, cout, acc);
            else {   // MODE_HH4: rows whose window leaves the image stay 0 (CTA-uniform; every later row of the image is
                     // dead too, so the running sums are not needed any more)
#pragma unroll
                for (int k = 0; k < COST_CPB_MAX; k++) {
                    uint32_t o[4] = {0u, 0u, 0u, 0u};
                    if (q.padded) {
#pragma unroll
                        for (int j = 0; j < 4; j++)
                            o[j] = q.cpad & ~(((q.lane * 8 + 2 * j < q.D) ? 0x0000FFFFu : 0u) |
                                              ((q.lane * 8 + 2 * j + 1 < q.D) ? 0xFFFF0000u : 0u));
                    }
                    stg128_pred(reinterpret_cast<char *>(cout) + (size_t)k * q.Dp2, o[0], o[1], o[2], o[3], q.cmask >> k & 1);
                }
            }
        }
        slot = slot + 1 == BS ? 0 : slot + 1;
    }
};

// The row loop.  Between two barriers a thread owes phase 2 of row i (reads pix buffer i & 1) and phase 1 of row i+1
// (writes the other buffer) in either order.  Warps w and w + 4 share a scheduler: the lower four run phase 2 first,
// the upper four phase 1 first (their barrier sits between the two), so that every scheduler has one warp in the
// ALU-pipe-heavy BT phase while the other is in the FMA-pipe / shared-memory-heavy box-sum phase (COST_DEPHASE).
#ifndef COST_DEPHASE
#define COST_DEPHASE 1
#endif
template <int NPL, int R>
__device__ __forceinline__ void cost_rows(CostProducer<NPL, R> &P, CostConsumer<R> &C, const CostTile &t)
{
    const int n = t.r1 - t.r0 + 1;
    const int late = (COST_DEPHASE && threadIdx.x >= COST_ROLE / 2) ? 1 : 0;
    P.row(t, t.r0, 0);
    for (int i = 0; i <= n; i++) {
        if (!late) __syncthreads();
        const int ic = i - late;                                  // phase 2 of row ic
        if (ic >= 0 && ic < n) C.row(t, t.r0 + ic, ic & 1);
        if (late) __syncthreads();
        if (i + 1 < n) P.row(t, t.r0 + i + 1, (i + 1) & 1);
    }
}

// Every thread plays both roles, one __syncthreads per row (the pix row and the staged left operands are double-buffered).
// (Measured and not adopted: a second, mask-free instantiation of the row loop for tiles that lie wholly inside the
// image — 105 instructions fewer per row in phase 2, but 254 registers and twice the code: 14.75 ms against 14.26;
// border replication done by phase 1 so that phase 2 needs no per-column clamp — 48 instructions fewer, 255 registers:
// 14.15 ms against 13.95; single-LOP3 predicate tests: neutral.)
template <int NPL, int R, int GT>   // GT: compile-time G (0 = runtime)
__global__ void __launch_bounds__(COST_ROLE, 1) k_cost_fused(const __grid_constant__ CostArgs a)
{
    constexpr int BS = 2 * R + 1;
    extern __shared__ __align__(16) unsigned char csm[];
    CostTile t;
    t.G = GT ? GT : a.G; t.Dp = t.G * 8;
    t.NW = a.NW; t.TX = t.NW - 2 * R; t.W1 = a.W1; t.H = a.H;
    t.PS = t.Dp + 8;                                                       // pix column stride (elements)
    t.smem0 = (uint32_t)__cvta_generic_to_shared(csm);                     // pix [2][NW][PS], ring [BS][TX][Dp], left
    t.pix_bytes = (uint32_t)(t.NW * t.PS) * 2u;
    t.NCOL = t.NW + 8 * (t.G - 1);
    t.left_bytes = (uint32_t)((NPL / 2) * t.NCOL) * 8u;
    t.left0 = t.smem0 + 2u * t.pix_bytes + (uint32_t)(BS * t.TX * t.Dp) * 2u;
    t.T0 = blockIdx.x * t.TX - 8 * (t.G - 1);                              // x of output column 0 of group 0
    t.ylo = a.y0;                                                           // rows [y0, y1) of the volume are produced
    t.ya = a.y0 + blockIdx.y * a.rows_per_seg; t.yb = min(t.ya + a.rows_per_seg, a.y1);
    t.r0 = t.ya - R; t.r1 = t.yb - 1 + R;
    const int b = blockIdx.z;
    const int nring = BS * t.TX * t.Dp / 8;   // ring starts at zero: "subtract the row that leaves the window" is unconditional
    CostProducer<NPL, R> P;
    CostConsumer<R> C;
    for (int i = threadIdx.x; i < nring; i += COST_ROLE) sts128(C.ring0(t) + 16u * i, 0, 0, 0, 0);
    P.init(a, t, threadIdx.x, b);
    C.init(a, t, threadIdx.x, b);
    __syncthreads();                          // the first row's staged left operands are visible
    cost_rows<NPL, R>(P, C, t);
}

struct CostGeom { int NW, TX, NPART, CPT, NCB, ntiles, rows_per_seg, nseg; size_t smem; };

// Phase-2 ownership.  Job (l, cb) = disparity group l, output columns x = T0 + 8*(l + cb) + k, k < 8.  Jobs with the
// same s = l + cb and consecutive l write 16-byte pieces that are contiguous in C (same pixel, consecutive groups), so
// the jobs are handed to threads in runs of 8 consecutive l of one s: a quarter warp then stores one contiguous 128-byte
// run per column instead of 8 pieces in 8 different lines (tid % G / tid / G made every lane of a warp hit its own
// 128-byte line: 32 L1 tag lookups per STG.128).  Full runs first (they stay aligned to quarter warps, which also keeps
// every LDS.128 / STS.128 of phase 2 conflict-free: the shared-memory address is 16*l + const mod 128), ragged ends
// of the parallelogram last.
#ifndef COST_JOBMAP
#define COST_JOBMAP 1
#endif
static void cost_jobs(int G, int NCB, uint16_t *job)
{
    int n = 0;
    for (int i = 0; i < COST_ROLE; i++) job[i] = 0xFFFF;
#if COST_JOBMAP
    for (int pass = 0; pass < 2; pass++)            // pass 0: full runs of 8, pass 1: the rest
        for (int s = 0; s <= G + NCB - 2; s++) {
            const int lo = std::max(0, s - NCB + 1), hi = std::min(G - 1, s);
            for (int l0 = lo; l0 <= hi; l0 += 8) {
                const int len = std::min(8, hi - l0 + 1);
                if ((len == 8) != (pass == 0)) continue;
                for (int l = l0; l < l0 + len && n < COST_ROLE; l++) job[n++] = (uint16_t)(((s - l) << 8) | l);
            }
        }
#else
    for (int cb = 0; cb < NCB; cb++)
        for (int l = 0; l < G && n < COST_ROLE; l++) job[n++] = (uint16_t)((cb << 8) | l);
#endif
}
static CostGeom cost_geom(const SgbmPlan &pl, int nrows)
{
    CostGeom g;
    const int Gc = pl.Dp / 8, BS = 2 * pl.R + 1;
    const size_t budget = (size_t)226 * 1024;
    g.NPART = std::max(1, Gc / 8);
    g.CPT = Gc / g.NPART;
    auto bytes = [&](int nw) {   // pix [2][nw][Dp+8] + ring [BS][nw-2R][Dp] int16, staged left operands [2][NCH][nw+8(G-1)] uint2
        return (size_t)2 * ((size_t)2 * nw * (pl.Dp + 8) + (size_t)BS * (nw - 2 * pl.R) * pl.Dp) +
               (size_t)2 * (pl.nplanes / 2) * (nw + 8 * (Gc - 1)) * 8 +
               (size_t)64 * (COST_CPT_MAX - g.CPT);   // slack: phase 1 reads 8 cells per thread even when CPT < 8
    };
    // widest tile whose windows fit the CTA (phase 1: NW*NPART threads), whose column blocks of COST_CPB_MAX columns
    // fit it too (phase 2: ceil(TX/8)*G threads) and whose buffers fit shared memory
    int NW = std::min(COST_ROLE / g.NPART, pl.W1 + 8 * (Gc - 1) + 2 * pl.R);
    NW = std::min(NW, COST_CPB_MAX * (COST_ROLE / Gc) + 2 * pl.R);
    while (NW > 2 * pl.R + 1 && bytes(NW) > budget) NW--;
    g.NW = NW; g.TX = NW - 2 * pl.R;
    g.NCB = (g.TX + COST_CPB_MAX - 1) / COST_CPB_MAX;
    g.ntiles = (pl.W1 + 8 * (Gc - 1) + g.TX - 1) / g.TX;
    g.nseg = std::max(1, std::min(8, nrows / 32));
    g.rows_per_seg = (nrows + g.nseg - 1) / g.nseg;
    g.nseg = (nrows + g.rows_per_seg - 1) / g.rows_per_seg;
    g.smem = bytes(NW);
    return g;
}

template <int NPL, int R, int GT>
static int launch_cost_fused(const SgbmPlan &pl, const CostGeom &g, char *ws, size_t ws_stride, int B, int y0, int y1,
                             cudaStream_t st)
{
    auto kern = k_cost_fused<NPL, R, GT>;
    SVB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
    CostArgs a;
    a.LP = (const uint2 *)(ws + pl.off_lplanes); a.RP = (const int16_t *)(ws + pl.off_rplanes); a.Cv = (int16_t *)(ws + pl.off_C);
    a.H = pl.H; a.W1 = pl.W1; a.Wr = pl.Wr; a.Dp = pl.Dp; a.D = pl.D; a.G = pl.Dp / 8;
    a.jb0 = pl.W - 1 - pl.minX1 + pl.p.minDisparity;
    a.NW = g.NW; a.NPART = g.NPART; a.CPT = g.CPT; a.NCB = g.NCB; a.rows_per_seg = g.rows_per_seg;
    a.zrow = pl.p.mode == 3 ? pl.H - pl.R : pl.H;
    a.y0 = y0; a.y1 = y1;
    a.cpad = (uint32_t)(32767 - pl.p.P2) * 0x10001u;
    a.one = 1u; a.mone = 0xFFFFFFFFu;
    a.ws_stride = ws_stride;
    cost_jobs(a.G, g.NCB, a.job);
    dim3 grid(g.ntiles, g.nseg, B);
    kern<<<grid, COST_ROLE, g.smem, st>>>(a);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

template <int NPL, int GT>
static int launch_cost_r(const SgbmPlan &pl, const CostGeom &g, char *ws, size_t ws_stride, int B, int y0, int y1,
                         cudaStream_t st)
{
    switch (pl.R) {
    case 0: return launch_cost_fused<NPL, 0, GT>(pl, g, ws, ws_stride, B, y0, y1, st);
    case 1: return launch_cost_fused<NPL, 1, GT>(pl, g, ws, ws_stride, B, y0, y1, st);
    case 2: return launch_cost_fused<NPL, 2, GT>(pl, g, ws, ws_stride, B, y0, y1, st);
    case 3: return launch_cost_fused<NPL, 3, GT>(pl, g, ws, ws_stride, B, y0, y1, st);
    case 4: return launch_cost_fused<NPL, 4, GT>(pl, g, ws, ws_stride, B, y0, y1, st);
    default: return launch_cost_fused<NPL, 5, GT>(pl, g, ws, ws_stride, B, y0, y1, st);
    }
}

// rows [y0, y1) of C; input rows are clamped to [y0, H-1] (the whole image: y0 = 0, y1 = H)
int svb_launch_cost(const SgbmPlan &pl, char *ws, size_t ws_stride, int B, int y0, int y1, cudaStream_t st)
{
    const CostGeom g = cost_geom(pl, y1 - y0);
    SVB_REQUIRE(g.smem <= 227 * 1024 && g.NW > 2 * pl.R + 1, SVB_ERR_UNSUPPORTED,
                "sgbm: cost tile does not fit shared memory (%zu bytes)", g.smem);
    const bool g32 = pl.Dp == 256 && g.CPT == 8;   // the 129..256-disparity case gets compile-time strides
    if (pl.nplanes == 2) return g32 ? launch_cost_r<2, 32>(pl, g, ws, ws_stride, B, y0, y1, st) : launch_cost_r<2, 0>(pl, g, ws, ws_stride, B, y0, y1, st);
    return g32 ? launch_cost_r<6, 32>(pl, g, ws, ws_stride, B, y0, y1, st) : launch_cost_r<6, 0>(pl, g, ws, ws_stride, B, y0, y1, st);
}
