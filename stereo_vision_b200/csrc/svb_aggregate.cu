// svb_aggregate.cu — SGM path aggregation (5-path MODE_SGBM, 8-path MODE_HH) + WTA / uniqueness / subpixel /
// disp2 scatter.  Semantics: SURVEY.md §3.2.4-3.2.5 (cv2 4.13.0), bit-exact.
//
// Design: ONE sweep kernel computes four path directions per pass as a wavefront.
//   * a "row-group" of G lanes owns one image row; each lane holds E=2*NR consecutive disparities as NR packed
//     s16x2 registers.  The group walks its row in x; the horizontal path lives in registers.
//   * row v trails row v-1 by two pixels, so the three vertical/diagonal predecessors (u-1,u,u+1 of row v-1) are
//     already available; inside a CTA (a band of NGRP rows) they travel through a 4-slot shared-memory ring, and
//     between bands through a small global edge buffer with acquire/release progress counters.  Bands take
//     tickets in launch order, so a band only ever waits for one that is already running (no deadlock).
//   * the cost volume C is read exactly once per pass, S is written once (pass 1) and read once (pass 2): the
//     four directions of a pass accumulate in registers.
//   * recurrence in normalised form: with Lc = min(Lprev - minLprev, P2),
//         L(d) = C(d) + min(Lc(d), min(Lc(d-1), Lc(d+1)) + P1)          (== cv2's formula)
//     -> per packed pair: PRMT, VIMNMX, VIADDMNMX, IADD, IADD(S), 1/2 VIMNMX3, VIADDMNMX(normalise+clamp).
//   * pass 2 (reverse sweep) adds its four directions to S from pass 1 and selects the disparity in-register.
#include "svb_common.cuh"

#define AGG_NGRP 16        // rows per band (row-groups per CTA)
#define AGG_SLOTS 4        // ring depth (u-1, u, u+1 live + one being written)
#define AGG_PUB 8          // publish cross-band progress every AGG_PUB pixels

struct AggArgs {
    const int16_t *C;      // [H][W1][Dp]
    int16_t *S;            // [H][W1][Dp]   pass 1: out, pass 2: in
    int16_t *edge;         // [nbands][W1][3][Dp]
    unsigned *prog;        // [nbands]   (this pass)
    unsigned *ticket;      // one counter (this pass)
    int16_t *dpre;         // [H][W]  pass 2 out
    unsigned *keys;        // [H][W]  pass 2 out (atomicMin)
    size_t ws_stride;      // bytes between frames for every pointer above except ticket
    int H, W, W1, D, Dp, minX1, minD;
    int nbands, B;
    int flip;              // 0: rows top-down, x ascending; 1: both reversed
    uint32_t P1d, P2d;     // duplicated halves
    int uniq, invalid16;
    int has_pad;
};

__device__ __forceinline__ unsigned ld_acquire(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(unsigned *p, unsigned v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <int G>
__device__ __forceinline__ uint32_t group_min_u32(uint32_t v)
{
    if (G == 32) return __reduce_min_sync(0xffffffffu, v);
#pragma unroll
    for (int o = G >> 1; o; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

template <int NR>
__device__ __forceinline__ uint32_t lane_min16(const uint32_t (&L)[NR])
{
    uint32_t m = L[0];
#pragma unroll
    for (int k = 1; k + 1 < NR; k += 2) m = __vimin3_s16x2(m, L[k], L[k + 1]);
    if ((NR & 1) == 0) m = __vmins2(m, L[NR - 1]);
    return min(m & 0xFFFFu, m >> 16);
}

// one direction: prev (clamped, normalised) Lc in p[], neighbour words wprev (hi half = Lc[d0-1]) and
// wnext (lo half = Lc[d0+E]).  Produces L[], returns nothing; caller reduces.
template <int NR>
__device__ __forceinline__ void dir_update(const uint32_t (&p)[NR], uint32_t wprev, uint32_t wnext,
                                           const uint32_t (&c)[NR], uint32_t P1d, uint32_t (&L)[NR])
{
    uint32_t m[NR + 1];
    m[0] = __byte_perm(wprev, p[0], 0x5432);
#pragma unroll
    for (int k = 1; k < NR; k++) m[k] = __byte_perm(p[k - 1], p[k], 0x5432);
    m[NR] = __byte_perm(p[NR - 1], wnext, 0x5432);
#pragma unroll
    for (int k = 0; k < NR; k++) {
        const uint32_t nb = __vmins2(m[k], m[k + 1]);
        const uint32_t t = __viaddmin_s16x2(nb, P1d, p[k]);
        L[k] = c[k] + t;   // packed add, no carry: C + t <= 32767 by contract
    }
}

template <int NR, int G, bool SAFE>
__device__ __forceinline__ void finish_dir(uint32_t (&L)[NR], uint32_t (&s)[NR], uint32_t P2d, const uint32_t (&keep)[NR],
                                           int has_pad)
{
    const uint32_t minL = group_min_u32<G>(lane_min16<NR>(L));
    const uint32_t negd = ((0u - minL) & 0xFFFFu) * 0x10001u;
#pragma unroll
    for (int k = 0; k < NR; k++) {
        if (SAFE) s[k] += L[k];
        else s[k] = __viaddmin_u16x2(s[k], L[k], 0x7FFF7FFFu);
        L[k] = __viaddmin_s16x2(L[k], negd, P2d);   // Lc = min(L - minL, P2)
    }
    if (has_pad) {
#pragma unroll
        for (int k = 0; k < NR; k++) L[k] = (L[k] & keep[k]) | (P2d & ~keep[k]);
    }
}

template <int NR> struct VecT;
template <> struct VecT<4> { typedef uint4 T; };
template <int NR>
__device__ __forceinline__ void ld_regs_shared(const int16_t *p, uint32_t (&r)[NR])
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++) {
        const uint4 v = reinterpret_cast<const uint4 *>(p)[q];
        r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
    }
}
template <int NR>
__device__ __forceinline__ void st_regs(int16_t *p, const uint32_t (&r)[NR])
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++)
        reinterpret_cast<uint4 *>(p)[q] = make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
}
template <int NR>
__device__ __forceinline__ void ld_regs_stream(const int16_t *p, uint32_t (&r)[NR])
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++) {
        const uint4 v = __ldcs(reinterpret_cast<const uint4 *>(p) + q);
        r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
    }
}
template <int NR>
__device__ __forceinline__ void ld_regs_cg(const int16_t *p, uint32_t (&r)[NR])
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++) {
        const uint4 v = __ldcg(reinterpret_cast<const uint4 *>(p) + q);
        r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
    }
}
template <int NR>
__device__ __forceinline__ void st_regs_cg(int16_t *p, const uint32_t (&r)[NR])
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++)
        __stcg(reinterpret_cast<uint4 *>(p) + q, make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]));
}
template <int NR>
__device__ __forceinline__ void st_regs_stream(int16_t *p, const uint32_t (&r)[NR])
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++)
        __stcs(reinterpret_cast<uint4 *>(p) + q, make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]));
}

// ---- disparity selection for one pixel (first loop).  s[] = final S, clamped to 0x7FFF,
// padded disparities forced to 0x7FFF.  All lanes of the warp must call (warp collectives inside). ----
template <int NR, int G>
__device__ __forceinline__ void wta_select(const uint32_t (&s)[NR], int lane, const AggArgs &a, int16_t *scratch,
                                           int16_t *dpre, unsigned *keys, int x_full, int y, bool active)
{
    constexpr int E = 2 * NR;
    const uint32_t d0 = (uint32_t)lane * E;
    uint32_t key[E];
#pragma unroll
    for (int k = 0; k < NR; k++) {
        key[2 * k] = __byte_perm(s[k], d0 + 2 * k, 0x1054);       // (S_lo << 16) | d
        key[2 * k + 1] = __byte_perm(s[k], d0 + 2 * k + 1, 0x3254);   // (S_hi << 16) | d
    }
    uint32_t lk = key[0];
#pragma unroll
    for (int k = 1; k < E; k++) lk = min(lk, key[k]);
    const uint32_t best = group_min_u32<G>(lk);
    const int minS = (int)(best >> 16), bestD = (int)(best & 0xFFFFu);
    bool fail = minS >= 32767;   // all costs saturated: cv2 ends with bestDisp=-1 -> INVALID
    if (a.uniq > 0) {
        uint32_t lk2 = 0xFFFFFFFFu;
#pragma unroll
        for (int k = 0; k < E; k++) {
            const uint32_t kk = ((uint32_t)((int)d0 + k - bestD + 1) <= 2u) ? 0xFFFFFFFFu : key[k];
            lk2 = min(lk2, kk);
        }
        const int S2 = (int)(group_min_u32<G>(lk2) >> 16);
        fail = fail || (S2 * (100 - a.uniq) < minS * 100);
    }
    // subpixel taps through shared memory
    st_regs<NR>(scratch + lane * E, s);
    __syncwarp();
    if (lane == 0 && active) {
        int out = a.invalid16;
        if (!fail) {
            const int x2 = x_full - bestD - a.minD;
            atomicMin(keys + (size_t)y * a.W + x2, ((unsigned)minS << 16) | (unsigned)(0xFFFF - x_full));
            int d16 = bestD * 16;
            if (bestD > 0 && bestD < a.D - 1) {
                const int sm1 = scratch[bestD - 1], sp1 = scratch[bestD + 1];
                const int denom2 = max(sm1 + sp1 - 2 * minS, 1);
                d16 += ((sm1 - sp1) * 16 + denom2) / (denom2 * 2);
            }
            out = d16 + a.minD * 16;
        }
        dpre[(size_t)y * a.W + x_full] = (int16_t)out;
    }
    __syncwarp();
}

// NV = 3: wavefront sweep with the three vertical/diagonal predecessors; NV = 0: horizontal path only
// (rows independent; used for the 5th path of MODE_SGBM).
template <int NR, int G, int NV, bool PASS2, bool SAFE>
__global__ void __launch_bounds__(AGG_NGRP *G) k_sweep(const AggArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int E = 2 * NR;
    const int Dp = a.Dp;
    int16_t *ring = reinterpret_cast<int16_t *>(smem_raw);
    int16_t *scratch_base = ring + (NV ? (size_t)AGG_NGRP * AGG_SLOTS * 3 * Dp : 0);
    __shared__ unsigned s_ticket;
    const int tid = threadIdx.x, g = tid / G, lane = tid % G;
    unsigned ticket = blockIdx.x;
    if (NV) {
        if (tid == 0) s_ticket = atomicAdd(a.ticket, 1u);
        __syncthreads();
        ticket = s_ticket;
    }
    const int frame = ticket / a.nbands, band = ticket % a.nbands;
    const size_t fo = (size_t)frame * a.ws_stride;
    const int16_t *Cv = reinterpret_cast<const int16_t *>(reinterpret_cast<const char *>(a.C) + fo);
    int16_t *Sv = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(a.S) + fo);
    int16_t *edge = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(a.edge) + fo);
    unsigned *prog = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(a.prog) + fo);
    int16_t *dpre = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(a.dpre) + fo);
    unsigned *keys = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(a.keys) + fo);

    const int v = band * AGG_NGRP + g;
    const bool rowActive = v < a.H;
    const int y = a.flip ? a.H - 1 - v : v;
    const size_t rowoff = (size_t)(rowActive ? y : 0) * a.W1 * Dp + lane * E;
    const int16_t *Crow = Cv + rowoff;
    int16_t *Srow = Sv + rowoff;
    uint32_t keep[NR];
#pragma unroll
    for (int k = 0; k < NR; k++)
        keep[k] = ((lane * E + 2 * k < a.D) ? 0x0000FFFFu : 0u) | ((lane * E + 2 * k + 1 < a.D) ? 0xFFFF0000u : 0u);
    const uint32_t P1d = a.P1d, P2d = a.P2d;
    int16_t *ring_g = ring + (size_t)g * AGG_SLOTS * 3 * Dp;
    const int16_t *ring_prev = ring + (size_t)(g > 0 ? g - 1 : 0) * AGG_SLOTS * 3 * Dp;
    int16_t *scratch = scratch_base + (size_t)g * Dp;
    const size_t edge_band = (size_t)a.W1 * 3 * Dp;
    const int16_t *edge_in = band > 0 ? edge + (size_t)(band - 1) * edge_band : nullptr;
    int16_t *edge_out = (NV && band + 1 < a.nbands && g == AGG_NGRP - 1) ? edge + (size_t)band * edge_band : nullptr;
    const bool publisher = NV && (band + 1 < a.nbands) && tid == AGG_NGRP * G - 1;

    uint32_t h[NR], cn[NR], sn[NR];
#pragma unroll
    for (int k = 0; k < NR; k++) { h[k] = 0; cn[k] = 0; sn[k] = 0; }
    const int lag = NV ? 2 * g : 0;
    const int nsteps = a.W1 + (NV ? 2 * (AGG_NGRP - 1) : 0);
    unsigned known = 0;
    // prefetch for the first step of group 0
    if (rowActive && lag == 0) {
        const int x0 = a.flip ? a.W1 - 1 : 0;
        ld_regs_stream<NR>(Crow + (size_t)x0 * Dp, cn);
        if (PASS2) ld_regs_stream<NR>(Srow + (size_t)x0 * Dp, sn);
    }
    for (int t = 0; t < nsteps; t++) {
        const int u = t - lag;
        const bool active = rowActive && u >= 0 && u < a.W1;
        if (NV) {
            if (tid == 0 && band > 0 && t < a.W1) {
                const unsigned need = (unsigned)min(t + 2, a.W1);
                while (known < need) known = ld_acquire(prog + band - 1);
            }
            __syncthreads();
            if (publisher) {
                const int done = t - 2 * (AGG_NGRP - 1);
                if (done > 0 && (done % AGG_PUB) == 0) {
                    __threadfence();
                    st_release(prog + band, (unsigned)done);
                }
            }
        }
        const int x = a.flip ? a.W1 - 1 - u : u;
        uint32_t c[NR], s[NR];
#pragma unroll
        for (int k = 0; k < NR; k++) { c[k] = cn[k]; s[k] = PASS2 ? sn[k] : 0u; }
        {   // prefetch next pixel of this row
            const int un = u + 1;
            if (rowActive && un >= 0 && un < a.W1) {
                const int xn = a.flip ? a.W1 - 1 - un : un;
                ld_regs_stream<NR>(Crow + (size_t)xn * Dp, cn);
                if (PASS2) ld_regs_stream<NR>(Srow + (size_t)xn * Dp, sn);
            }
        }
        uint32_t L[NR];
        // ---- horizontal path (registers) ----
        {
            if (u <= 0) {
#pragma unroll
                for (int k = 0; k < NR; k++) h[k] = 0;
            }
            uint32_t wprev = __shfl_up_sync(0xffffffffu, h[NR - 1], 1);
            uint32_t wnext = __shfl_down_sync(0xffffffffu, h[0], 1);
            if (lane == 0) wprev = P2d;
            if (lane == G - 1) wnext = P2d;
            dir_update<NR>(h, wprev, wnext, c, P1d, L);
            finish_dir<NR, G, SAFE>(L, s, P2d, keep, a.has_pad);
#pragma unroll
            for (int k = 0; k < NR; k++) h[k] = L[k];
        }
        // ---- vertical / diagonal paths: predecessors (u-1, u, u+1) of the previous row ----
        if (NV) {
#pragma unroll
            for (int j = 0; j < 3; j++) {
                const int uu = u - 1 + j;
                uint32_t p[NR], wprev = 0, wnext = 0;
#pragma unroll
                for (int k = 0; k < NR; k++) p[k] = 0;
                if (active && v > 0 && uu >= 0 && uu < a.W1) {
                    if (g > 0) {
                        const int16_t *src = ring_prev + (size_t)((uu & (AGG_SLOTS - 1)) * 3 + j) * Dp + lane * E;
                        ld_regs_shared<NR>(src, p);
                        wprev = lane > 0 ? *reinterpret_cast<const uint32_t *>(src - 2) : P2d;
                        wnext = lane < G - 1 ? *reinterpret_cast<const uint32_t *>(src + E) : P2d;
                    } else {
                        const int16_t *src = edge_in + ((size_t)uu * 3 + j) * Dp + lane * E;
                        ld_regs_cg<NR>(src, p);
                        wprev = lane > 0 ? __ldcg(reinterpret_cast<const uint32_t *>(src - 2)) : P2d;
                        wnext = lane < G - 1 ? __ldcg(reinterpret_cast<const uint32_t *>(src + E)) : P2d;
                    }
                }
                dir_update<NR>(p, wprev, wnext, c, P1d, L);
                finish_dir<NR, G, SAFE>(L, s, P2d, keep, a.has_pad);
                if (active) {
                    st_regs<NR>(ring_g + (size_t)((u & (AGG_SLOTS - 1)) * 3 + j) * Dp + lane * E, L);
                    if (edge_out) st_regs_cg<NR>(edge_out + ((size_t)u * 3 + j) * Dp + lane * E, L);
                }
            }
        }
        if (!PASS2) {
            if (active) {
                if (!SAFE) {
#pragma unroll
                    for (int k = 0; k < NR; k++) s[k] = __vminu2(s[k], 0x7FFF7FFFu);
                }
                st_regs_stream<NR>(Srow + (size_t)x * Dp, s);
            }
        } else {
#pragma unroll
            for (int k = 0; k < NR; k++) {
                if (a.has_pad) s[k] |= 0x7FFF7FFFu & ~keep[k];
                s[k] = __vminu2(s[k], 0x7FFF7FFFu);
            }
            wta_select<NR, G>(s, lane, a, scratch, dpre, keys, x + a.minX1, y, active);
        }
    }
    if (NV && (band + 1 < a.nbands)) {
        __syncthreads();
        if (publisher) {
            __threadfence();
            st_release(prog + band, (unsigned)a.W1);
        }
    }
}

// ---------------- host side ----------------
template <int NR, int G, int NV, bool PASS2, bool SAFE>
static int launch_sweep_t(const AggArgs &a, int ngrid, cudaStream_t st)
{
    const size_t smem = (NV ? (size_t)AGG_NGRP * AGG_SLOTS * 3 * a.Dp * 2 : 0) + (PASS2 ? (size_t)AGG_NGRP * a.Dp * 2 : 0);
    auto kern = k_sweep<NR, G, NV, PASS2, SAFE>;
    static size_t configured = 0;
    if (smem > configured) {
        SVB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    kern<<<ngrid, AGG_NGRP * G, smem, st>>>(a);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

template <int NV, bool PASS2, bool SAFE>
static int launch_sweep_g(int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st)
{
    if (NR == 8) return launch_sweep_t<8, 32, NV, PASS2, SAFE>(a, ngrid, st);
    switch (G) {
    case 2: return launch_sweep_t<4, 2, NV, PASS2, SAFE>(a, ngrid, st);
    case 4: return launch_sweep_t<4, 4, NV, PASS2, SAFE>(a, ngrid, st);
    case 8: return launch_sweep_t<4, 8, NV, PASS2, SAFE>(a, ngrid, st);
    case 16: return launch_sweep_t<4, 16, NV, PASS2, SAFE>(a, ngrid, st);
    default: return launch_sweep_t<4, 32, NV, PASS2, SAFE>(a, ngrid, st);
    }
}

template <int NV, bool PASS2>
static int launch_sweep(int safe, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st)
{
    return safe ? launch_sweep_g<NV, PASS2, true>(NR, G, a, ngrid, st) : launch_sweep_g<NV, PASS2, false>(NR, G, a, ngrid, st);
}

int svb_launch_aggregate(svb_context *ctx, const SgbmPlan &pl, char *ws, size_t ws_stride, int B, int passes,
                         cudaStream_t st)
{
    // reset schedulers / progress / scatter keys
    if (passes & 1) {
    SVB_CUDA_CHECK(cudaMemsetAsync(ctx->tickets, 0, 2 * sizeof(unsigned), st));
    for (int b = 0; b < B; b++) {
        SVB_CUDA_CHECK(cudaMemsetAsync(ws + (size_t)b * ws_stride + pl.off_prog, 0, 2 * sizeof(unsigned) * pl.nbands, st));
        SVB_CUDA_CHECK(cudaMemsetAsync(ws + (size_t)b * ws_stride + pl.off_key, 0xFF, sizeof(unsigned) * (size_t)pl.H * pl.W, st));
    }
    }
    AggArgs a;
    memset(&a, 0, sizeof a);
    a.C = (const int16_t *)(ws + pl.off_C);
    a.S = (int16_t *)(ws + pl.off_S);
    a.edge = (int16_t *)(ws + pl.off_edge);
    a.dpre = (int16_t *)(ws + pl.off_dpre);
    a.keys = (unsigned *)(ws + pl.off_key);
    a.ws_stride = ws_stride;
    a.H = pl.H; a.W = pl.W; a.W1 = pl.W1; a.D = pl.D; a.Dp = pl.Dp; a.minX1 = pl.minX1; a.minD = pl.p.minDisparity;
    a.nbands = pl.nbands; a.B = B;
    a.P1d = (uint32_t)pl.p.P1 * 0x10001u;
    a.P2d = (uint32_t)pl.p.P2 * 0x10001u;
    a.uniq = pl.p.uniquenessRatio;
    a.invalid16 = pl.invalid16;
    a.has_pad = pl.Dp != pl.D;
    const int NR = pl.E / 2;
    // pass 1: left, up-left, up, up-right
    a.flip = 0;
    a.prog = (unsigned *)(ws + pl.off_prog);
    a.ticket = ctx->tickets;
    int rc = SVB_OK;
    if (passes & 1) rc = launch_sweep<3, false>(pl.safe, NR, pl.G, a, pl.nbands * B, st);
    if (rc || !(passes & 2)) return rc;
    a.flip = 1;
    if (pl.p.mode == 1) {   // MODE_HH: right, down-right, down, down-left + selection
        a.prog = (unsigned *)(ws + pl.off_prog) + pl.nbands;
        a.ticket = ctx->tickets + 1;
        rc = launch_sweep<3, true>(pl.safe, NR, pl.G, a, pl.nbands * B, st);
    } else {                // MODE_SGBM: 5th path (from the right) + selection, rows independent
        rc = launch_sweep<0, true>(pl.safe, NR, pl.G, a, pl.nbands * B, st);
    }
    return rc;
}
