// svb_aggregate.cu — SGM path aggregation (5-path MODE_SGBM, 8-path MODE_HH) + WTA / uniqueness / subpixel /
// disp2 scatter.  Semantics: SURVEY.md §3.2.4-3.2.5 (cv2 4.13.0), bit-exact.
//
// Design: ONE sweep kernel computes four path directions per pass as a wavefront.
//   * a "row-group" of G lanes owns one image row; each lane holds E=2*NR consecutive disparities as NR packed
//     s16x2 registers.  The group walks its row in x; the horizontal path lives in registers.
//   * the CTA (a band of NGRP rows) advances AGG_PPS = 2 pixels per bar.sync; row v trails row v-1 by AGG_LAG = 3
//     pixels, so the three vertical/diagonal predecessors (u-1,u,u+1 of row v-1) of both pixels are at least one
//     barrier old.  They travel through an AGG_SLOTS = 6-slot shared-memory ring per row whose slot index is the
//     pixel-time mod 6 — the loop is unrolled by AGG_UNROLL = 12 pixel-times, so every ring address is an immediate.
//     Each ring vector is preceded by a 16-byte pad holding P2, which is the d=-1 / d=D halo of the recurrence, so the
//     neighbour taps are plain unconditional LDS.
//   * between bands the three states travel through a global edge buffer.  A dedicated DMA warp per CTA polls the
//     producer band's progress counter (relaxed loads + one acquire fence per advance, see ld_relaxed), prefetches the edge vectors into L2 AGG_L2PF
//     pixels ahead, pulls them into the "row -1" ring with cp.async, and publishes this band's own progress
//     (st.release) — all off the compute warps' critical path.  Bands take tickets in band-major order, so a band only
//     ever waits for one that is already running (no deadlock, no cooperative launch).
//   * C, S and the edge stream are addressed as one pointer per unrolled iteration + compile-time offsets (Dp = G*E and
//     the sweep direction are template constants); the pointers are not clamped to the row (SVB_SWEEP_MARGIN).
//     C (and S in pass 2) are prefetched PFD pixels ahead into registers; C is read exactly once per pass,
//     S written once (pass 1) / read once (pass 2): the four directions of a pass accumulate in registers.
//   * the loop body is branch-free: out-of-grid predecessors are zero vectors selected by pointer, inactive
//     pixel-times store into a trash row, global stores are predicated.
//   * recurrence in normalised form: with Lc = min(Lprev - minLprev, P2),
//         L(d) = C(d) + min(Lc(d), Lc(d-1) + P1, Lc(d+1) + P1)          (== cv2's formula)
//     -> per packed pair: 1.25 PRMT, VIMNMX3, VIADDMNMX (normalise+clamp) on the ALU pipe; the +P1, +C and the add
//        into S are IMADs on the FMA pipe; directions are finished in pairs (finish_pair).
//   * pass 2 (reverse sweep) adds its four directions to S from pass 1; the selection of a pixel (winner, second
//     best, subpixel taps) runs one pixel-time later from a shared-memory copy of S (select), its three scalars are
//     parked in shared memory, and the scalar tail (uniqueness test, subpixel division, disp2 scatter, store) runs
//     lane-parallel every 24 pixels (flush).
#include "svb_common.cuh"
#include <stdlib.h>
#include <algorithm>

#ifndef AGG_MIN3
#define AGG_MIN3 1         // dir_update: one VIMNMX3 per register (0: VIMNMX + VIADDMNMX)
#endif
#ifndef AGG_SELPIPE
#define AGG_SELPIPE 1      // pass 2: the selection of a pixel runs during the next pixel's path updates
#endif
#define AGG_PPS 2          // pixels every row advances per CTA barrier (one pixel per barrier: 9 % / 7 % slower)
#define AGG_LAG 3          // row v trails row v-1 by AGG_LAG pixels (PPS + 1: the u+1 predecessor is one barrier old)
#define AGG_SLOTS 6        // ring depth: a vector is read up to LAG+1 pixel-times after it was written, i.e. two
                           // barriers later, and its slot is rewritten SLOTS/PPS barriers later
#define AGG_ENS 12         // slots of the cross-band "row -1" ring
#define AGG_EPF 8          // cross-band prefetch distance: at pixel-time t pixels <= t+EPF may be issued; the slot of
                           // pixel t+EPF holds pixel t+EPF-ENS, which row 0 may still be reading during the previous
                           // barrier interval unless EPF <= ENS-PPS-2
// Kernels with G == 32 lanes per row (Dp >= 256) use only AGG_SLOTS slots of that ring, phased like a row's own ring
// (pixel pe in slot (pe + SLOTS - LAG) % SLOTS): row 0 then addresses its predecessors exactly like every other row
// (no per-direction select; 3 % off pass 2, whose register allocation is at the limit).  The price is no fetch slack
// (EPF = SLOTS - PPS - 2 = 2 while row 0 needs pixels <= t+PPS: every cp.async has to land within its own barrier
// interval).  That holds for the long intervals of Dp >= 256 because the vectors were prefetched into L2 (AGG_L2PF);
// with the short intervals of Dp = 128 it made pass 1 25 % slower, hence the compile-time switch.
#define AGG_EDGE_LIKE_ROW(G) ((G) == 32 && NR >= 4)   // (NR: the enclosing template parameter)
#define AGG_UNROLL 12      // pixel-times per loop iteration: a multiple of SLOTS, ENS, PPS and the prefetch depths
#ifndef AGG_PUB
#define AGG_PUB 8          // publish cross-band progress every AGG_PUB pixels (a single frame in flight: every AGG_PPS)
#endif
#ifndef AGG_L2PF
#define AGG_L2PF 8         // cross-band L2 prefetch distance (pixels ahead of the cp.async front), 0 = off;
                           // 6..16 measure the same, 32 is 1-2 % slower (lines evicted again), 0 is 11 % slower
#endif
#define AGG_PADE 8         // int16 elements of P2 halo in front of every ring vector
// (Measured and not adopted: a progress poll pipelined three barrier intervals deep instead of one — the staler value costs
// more than the exposed load latency: single pair c3 3.58 -> 3.71 ms per pass, c1 0.93 -> 0.99 ms, pass 2 at B = 32 +4 %;
// prefetch.global.L2 of the C / S streams 16 pixels ahead of the register prefetch — pass 2
// 22.21 -> 21.89 ms in tools/exp_sweep.cu but 22.5 -> 23.7 ms in the product at B = 32, pass 1 16.90 -> 17.28 ms.)

#ifdef EXP_HARNESS
__device__ unsigned long long g_spin[4];
#endif
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
    uint32_t one;          // == 1 (an opaque multiplier: x * one + y stays an IMAD on the FMA pipe)
    uint32_t k16;          // == 65536, likewise (x * k16 + y instead of a PRMT / shift on the ALU pipe)
    int uniq, invalid16;
    int tw;                // MODE_SGBM_3WAY selection rules (winner among equal minima, integer uniqueness threshold)
    int pubmask;           // publish progress when (pixels done & pubmask) < AGG_PPS
    int slack;             // extra producer lead a starving band waits for (0 when one frame is in flight)
};

// Progress poll.  A relaxed (L1-bypassing) load, not ld.acquire: ptxas lowers every gpu-scope acquire/fence to
// LDG + CCTL.IVALL, i.e. an invalidation of the SM's whole L1 on every poll.  The acquire side of the hand-off is one
// fence_acquire() each time the known progress ADVANCES (at most once per AGG_PUB pixels, not once per poll), issued
// before the cp.async.cg reads of the vectors that progress value guards — and only when the fetch front actually needs
// pixels beyond the last acquired value (`acq`): a consumer whose producer is far ahead (the normal case with several
// frames in flight) fences once per producer lead, not once per publication.  (A fence on every advance cost 2.8 % of
// both sweeps at B = 32: tools/exp_sweep.cu, 17.07 -> 17.56 / 21.63 -> 22.24 ms.)
__device__ __forceinline__ unsigned ld_relaxed(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fence_acquire() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ void st_release(unsigned *p, unsigned v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void *gmem)
{
    asm volatile("prefetch.global.L2::evict_last [%0];" ::"l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int G>
__device__ __forceinline__ uint32_t group_min_u32(uint32_t v)
{
#ifdef EXP_NOREDUX
    return v;
#endif
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
// wnext (lo half = Lc[d0+E]).  Produces L[].
template <int NR>
__device__ __forceinline__ void dir_update(const uint32_t (&p)[NR], uint32_t wprev, uint32_t wnext,
                                           const uint32_t (&c)[NR], uint32_t P1d, uint32_t one, uint32_t (&L)[NR])
{
    uint32_t m[NR + 1];
    m[0] = __byte_perm(wprev, p[0], 0x5432);
#pragma unroll
    for (int k = 1; k < NR; k++) m[k] = __byte_perm(p[k - 1], p[k], 0x5432);
    m[NR] = __byte_perm(p[NR - 1], wnext, 0x5432);
#if AGG_MIN3
    // min(Lc(d), Lc(d-1)+P1, Lc(d+1)+P1) as one three-input packed min: the +P1 of the 5 shifted words is a plain
    // 32-bit add (Lc <= P2, so no carry between halves; up to 2*P2 <= 65534, hence the unsigned compare) that can
    // issue on the FMA pipe, and the ALU pipe sees one VIMNMX3 instead of VIMNMX + VIADDMNMX per register.
#pragma unroll
    for (int k = 0; k <= NR; k++) m[k] = m[k] * one + P1d;   // 'one' is a kernel argument == 1: keeps it an IMAD
#pragma unroll
    for (int k = 0; k < NR; k++) L[k] = c[k] + __vimin3_u16x2(p[k], m[k], m[k + 1]);
#else
#pragma unroll
    for (int k = 0; k < NR; k++) {
        const uint32_t nb = __vmins2(m[k], m[k + 1]);
        const uint32_t t = __viaddmin_s16x2(nb, P1d, p[k]);
        L[k] = c[k] + t;   // packed add, no carry: C + t <= 32767 by contract
    }
#endif
}

// add L into S, then normalise L in place: Lc = min(L - min_d L, P2) (padded disparities forced to P2)
template <int NR, int G, bool SAFE, bool PAD>
__device__ __forceinline__ void finish_dir(uint32_t (&L)[NR], uint32_t (&s)[NR], uint32_t P2d, const uint32_t (&keep)[NR])
{
    const uint32_t minL = group_min_u32<G>(lane_min16<NR>(L));
    const uint32_t negd = ((0u - minL) & 0xFFFFu) * 0x10001u;
#pragma unroll
    for (int k = 0; k < NR; k++) {
        if (SAFE) s[k] += L[k];
        else s[k] = __viaddmin_u16x2(s[k], L[k], 0x7FFF7FFFu);
        L[k] = __viaddmin_s16x2(L[k], negd, P2d);
    }
    if (PAD) {
#pragma unroll
        for (int k = 0; k < NR; k++) L[k] = (L[k] & keep[k]) | (P2d & ~keep[k]);
    }
}

template <int G>
__device__ __forceinline__ uint32_t group_max_u32(uint32_t v)
{
#ifdef EXP_NOREDUX
    return v;
#endif
    if (G == 32) return __reduce_max_sync(0xffffffffu, v);
#pragma unroll
    for (int o = G >> 1; o; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// finish_dir for two directions at once.  The two per-lane minima are folded side by side in one register
// (3 ALU instructions instead of 2 x 3), complemented (0xFFFF - m = -m - 1 per half, no borrow) and reduced with
// REDUX.MAX, whose result is then one add and one PRMT away from the packed -minL the normalisation needs
// (instead of move, negate, mask, multiply): 6 ALU-pipe instructions fewer per pixel and a shorter dependent chain.
template <int NR, int G, bool SAFE, bool PAD>
__device__ __forceinline__ void finish_pair(uint32_t (&LA)[NR], uint32_t (&LB)[NR], uint32_t (&s)[NR], uint32_t P2d,
                                            uint32_t one, const uint32_t (&keep)[NR])
{
    uint32_t mA = LA[0], mB = LB[0];
#pragma unroll
    for (int k = 1; k + 1 < NR; k += 2) { mA = __vimin3_s16x2(mA, LA[k], LA[k + 1]); mB = __vimin3_s16x2(mB, LB[k], LB[k + 1]); }
    if ((NR & 1) == 0) { mA = __vmins2(mA, LA[NR - 1]); mB = __vmins2(mB, LB[NR - 1]); }
    const uint32_t z = __vminu2(__byte_perm(mA, mB, 0x5410), __byte_perm(mA, mB, 0x7632));   // (min A, min B) of this lane
    const uint32_t zc = z * (0u - one) + 0xFFFFFFFFu;                  // ~z on the FMA pipe
    const uint32_t rB = group_max_u32<G>(zc);                          // high half: 0xFFFF - min_d L_B
    const uint32_t rA = group_max_u32<G>(zc * (one << 16));            // high half: 0xFFFF - min_d L_A
    const uint32_t negA = __byte_perm(rA * one + 0x10000u, 0u, 0x3232);   // (-minL) mod 2^16 in both halves (IMAD + PRMT)
    const uint32_t negB = __byte_perm(rB * one + 0x10000u, 0u, 0x3232);
#pragma unroll
    for (int k = 0; k < NR; k++) {
        if (SAFE) s[k] += LA[k] + LB[k];
        else s[k] = __viaddmin_u16x2(__viaddmin_u16x2(s[k], LA[k], 0x7FFF7FFFu), LB[k], 0x7FFF7FFFu);
        LA[k] = __viaddmin_s16x2(LA[k], negA, P2d);
        LB[k] = __viaddmin_s16x2(LB[k], negB, P2d);
    }
    if (PAD) {
#pragma unroll
        for (int k = 0; k < NR; k++) {
            LA[k] = (LA[k] & keep[k]) | (P2d & ~keep[k]);
            LB[k] = (LB[k] & keep[k]) | (P2d & ~keep[k]);
        }
    }
}

// Vector moves of a lane's NR packed registers (E = 2*NR disparities): 16-byte pieces for NR = 4 / 8, one 8- or 4-byte
// piece for the wide-lane layouts NR = 2 / 1 (32 lanes per row at Dp = 128 / 64, used by the latency form).
template <int NR>
__device__ __forceinline__ void ld_regs_shared(const int16_t *p, uint32_t (&r)[NR])
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++) {
        const uint4 v = reinterpret_cast<const uint4 *>(p)[q];
        r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
    }
    if (NR == 2) { const uint2 v = *reinterpret_cast<const uint2 *>(p); r[0] = v.x; r[NR - 1] = v.y; }
    if (NR == 1) r[0] = *reinterpret_cast<const uint32_t *>(p);
}
template <int NR>
__device__ __forceinline__ void st_regs(int16_t *p, const uint32_t (&r)[NR])
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++)
        reinterpret_cast<uint4 *>(p)[q] = make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
    if (NR == 2) *reinterpret_cast<uint2 *>(p) = make_uint2(r[0], r[NR - 1]);
    if (NR == 1) *reinterpret_cast<uint32_t *>(p) = r[0];
}
template <int NR>
__device__ __forceinline__ void ld_regs_stream(const int16_t *p, uint32_t (&r)[NR])
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++) {
        const uint4 v = __ldcs(reinterpret_cast<const uint4 *>(p) + q);
        r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
    }
    if (NR == 2) { const uint2 v = __ldcs(reinterpret_cast<const uint2 *>(p)); r[0] = v.x; r[NR - 1] = v.y; }
    if (NR == 1) r[0] = __ldcs(reinterpret_cast<const uint32_t *>(p));
}
// predicated global stores (no branch: keeps the four directions of a step in one basic block)
__device__ __forceinline__ void st_cg_pred(void *p, uint32_t a, uint32_t b, uint32_t c, uint32_t d, bool pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %5, 0;\n\t@p st.global.cg.v4.u32 [%0], {%1, %2, %3, %4};\n\t}"
                 ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d), "r"((int)pred) : "memory");
}
__device__ __forceinline__ void st_cs_pred(void *p, uint32_t a, uint32_t b, uint32_t c, uint32_t d, bool pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %5, 0;\n\t@p st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};\n\t}"
                 ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d), "r"((int)pred) : "memory");
}
template <bool CS>   // 8- and 4-byte forms (CS: streaming, else cache-global)
__device__ __forceinline__ void st_pred2(void *p, uint32_t a, uint32_t b, bool pred)
{
    if (CS) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %3, 0;\n\t@p st.global.cs.v2.u32 [%0], {%1, %2};\n\t}" ::"l"(p), "r"(a), "r"(b), "r"((int)pred) : "memory");
    else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %3, 0;\n\t@p st.global.cg.v2.u32 [%0], {%1, %2};\n\t}" ::"l"(p), "r"(a), "r"(b), "r"((int)pred) : "memory");
}
template <bool CS>
__device__ __forceinline__ void st_pred1(void *p, uint32_t a, bool pred)
{
    if (CS) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %2, 0;\n\t@p st.global.cs.u32 [%0], %1;\n\t}" ::"l"(p), "r"(a), "r"((int)pred) : "memory");
    else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %2, 0;\n\t@p st.global.cg.u32 [%0], %1;\n\t}" ::"l"(p), "r"(a), "r"((int)pred) : "memory");
}
template <int NR>
__device__ __forceinline__ void st_regs_cg_pred(int16_t *p, const uint32_t (&r)[NR], bool pred)
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++)
        st_cg_pred(reinterpret_cast<uint4 *>(p) + q, r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3], pred);
    if (NR == 2) st_pred2<false>(p, r[0], r[NR - 1], pred);
    if (NR == 1) st_pred1<false>(p, r[0], pred);
}
template <int NR>
__device__ __forceinline__ void st_regs_cs_pred(int16_t *p, const uint32_t (&r)[NR], bool pred)
{
#pragma unroll
    for (int q = 0; q < NR / 4; q++)
        st_cs_pred(reinterpret_cast<uint4 *>(p) + q, r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3], pred);
    if (NR == 2) st_pred2<true>(p, r[0], r[NR - 1], pred);
    if (NR == 1) st_pred1<true>(p, r[0], pred);
}

// ---- pass-2 tail, lane-parallel: lane holds one parked pixel (first loop) ----
__device__ __forceinline__ void select_flush(const AggArgs &a, int16_t *dpre, unsigned *keys, int y, int x_full,
                                             uint32_t best, uint32_t s2, uint32_t taps)
{
    const int minS = (int)(best >> 16), bestD = (int)(best & 0xFFFFu);
    bool fail = minS >= 32767;   // all costs saturated: cv2 ends with bestDisp=-1 -> INVALID
    if (a.uniq > 0)   // cv2's 3-way mode compares against a precomputed integer threshold (oracle: select_impl)
        fail = fail || (a.tw ? (int)s2 <= (minS * 100) / (100 - a.uniq) : (int)s2 * (100 - a.uniq) < minS * 100);
    int out = a.invalid16;
    if (!fail) {
        const int x2 = x_full - bestD - a.minD;
        atomicMin(keys + (size_t)y * a.W + x2, ((unsigned)minS << 16) | (unsigned)(0xFFFF - x_full));
        int d16 = bestD * 16;
        if (bestD > 0 && bestD < a.D - 1) {
            const int sm1 = (int)(taps & 0xFFFFu), sp1 = (int)(taps >> 16);
            const int denom2 = max(sm1 + sp1 - 2 * minS, 1);
            d16 += ((sm1 - sp1) * 16 + denom2) / (denom2 * 2);
        }
        out = d16 + a.minD * 16;
    }
    dpre[(size_t)y * a.W + x_full] = (int16_t)out;
}

// Shared-memory geometry of one CTA (int16 elements)
template <int NGRP, int NV, bool PASS2>
struct SweepSmem {
    __host__ __device__ static int vs(int Dp) { return Dp + AGG_PADE; }
    __host__ __device__ static int nvec() { return NV ? NGRP * AGG_SLOTS * 3 + AGG_ENS * 3 + 1 + AGG_SLOTS * 3 : 0; }
    __host__ __device__ static int ring_elems(int Dp) { return NV ? nvec() * vs(Dp) + AGG_PADE : 0; }
    __host__ __device__ static int scratch_elems(int Dp) { return PASS2 ? NGRP * 2 * Dp : 0; }
    __host__ __device__ static int mask_elems(int E) { return PASS2 ? (E + 4) * E : 0; }
    // pass 2: per row, 2*AGG_UNROLL parked pixels x (best key, second-best S, subpixel taps, -) as uint4
    __host__ __device__ static int park_elems() { return PASS2 ? NGRP * 2 * AGG_UNROLL * 8 : 0; }
    __host__ __device__ static size_t bytes(int Dp, int E)
    {
        return 2 * (size_t)(ring_elems(Dp) + scratch_elems(Dp) + mask_elems(E) + park_elems()) + 16;
    }
};

// Per-thread state of a compute lane.  All arrays are indexed by compile-time constants only.
// PASS2: add pass-1 S and select the disparity (else: store S).  REV: walk the image rotated by 180 degrees (the four
// "reverse" path directions).  REV without PASS2 is the latency form of pass 2: it runs concurrently with pass 1 and
// leaves its own four-path sum in a second volume (k_select_sum finishes the job).
template <int NR, int G, int NGRP, int NV, bool PASS2, bool SAFE, bool PAD, int PFD, bool REV = PASS2>
struct SweepLane {
    static constexpr int E = 2 * NR;
    static constexpr int DP = G * E;                // padded disparities of this instantiation (== a.Dp)
    static constexpr int XDIR = REV ? -1 : 1;       // reverse sweeps walk every row right to left (and the rows bottom-up)
    // pass 2: selection one pixel-time behind the path updates (not for the single-path kernel: nothing to overlap with)
    static constexpr bool SELPIPE = AGG_SELPIPE && NV > 0;
    // loop-invariant
    int W1, VS, lag, lane, y, xbase;                // x = xbase + XDIR*u
    bool rowActive, isRow0;
    uint32_t P1d, P2d;
    uint32_t keep[NR];
    const int16_t *prevL;       // ring vectors of row g-1 (data start + lane*E); rowm1 ring for g == 0
    int16_t *ownL;              // ring vectors of row g
    const int16_t *zeroL;
    int16_t *trashL;            // ring-shaped scratch row: where inactive steps park their (meaningless) vectors
    bool edgeRow;               // last row of a band that has a successor
    int16_t *scratch;           // pass 2: 2*Dp per group
    const int16_t *masktab;
    uint4 *park;                // pass 2: this row's parked pixels
    // loop-carried.  The three streams are addressed as (pointer of pixel-time t0) + compile-time offset: the pointers
    // move once per unrolled iteration and are NOT clamped to the row — prefetches of the ramp-up / ramp-down
    // pixel-times read up to (lag_last + AGG_UNROLL + PFD) vectors outside it (values never used; the workspace keeps a
    // margin of SVB_SWEEP_MARGIN bytes around C and S), stores are predicated.
    const int16_t *cptr;        // C  of pixel u = t0 - lag, + lane*E
    int16_t *sptr;              // S  likewise
    int16_t *eptr;              // edge_out vector 0 of pixel u = t0 - lag, + lane*E
    uint4 *parkw;               // park slot of pixel-time t0 (toggles between the two halves)
    uint32_t h[NR];
    static constexpr int PFS = PFD > 2 ? 2 : PFD;   // S (pass 2) is prefetched less deep than C: it is consumed at the
                                                    // end of the step, and its registers are what pass 2 is short of
    uint32_t cbuf[PFD][NR], sbuf[PFS][NR];

    template <int SLOT, int OFF>   // pixel u = t0 - lag + OFF
    __device__ __forceinline__ void prefetch_c() { ld_regs_stream<NR>(cptr + XDIR * OFF * DP, cbuf[SLOT]); }
    template <int SLOT, int OFF>
    __device__ __forceinline__ void prefetch_s() { if (PASS2) ld_regs_stream<NR>(sptr + XDIR * OFF * DP, sbuf[SLOT]); }
    __device__ __forceinline__ void advance()
    {
        cptr += XDIR * AGG_UNROLL * DP;
        sptr += XDIR * AGG_UNROLL * DP;
        eptr += AGG_UNROLL * 3 * DP;
        if (PASS2) parkw = park + ((parkw - park) ^ AGG_UNROLL);
    }

    template <int PH>
    __device__ __forceinline__ void step(const AggArgs &a, int t, int16_t *dpre, unsigned *keys)
    {
        constexpr int PB = PH % PFD;
        const int u = t - lag;
        const bool act = rowActive && (unsigned)u < (unsigned)W1;
        uint32_t c[NR], s[NR], L[NR];
        if (PASS2 && SELPIPE)   // selection of pixel-time t-1 (slot PH-1 of this half, or the last slot of the other half)
            select<(PH + AGG_UNROLL - 1) % AGG_UNROLL>(PH ? parkw + (PH - 1) : park + (((parkw - park) ^ AGG_UNROLL) + AGG_UNROLL - 1), a.k16);
#pragma unroll
        for (int k = 0; k < NR; k++) { c[k] = cbuf[PB][k]; s[k] = (PASS2 && !SAFE) ? sbuf[PH % PFS][k] : 0u; }
        prefetch_c<PB, PH + PFD>();
        if (!SAFE) prefetch_s<PH % PFS, PH + PFS>();
        // ---- horizontal path (registers) ----
        uint32_t Lh[NR];
        {
            if (PH % AGG_LAG == 0 || !NV) {   // the row starts (u == 0) at pixel-time lag = AGG_LAG * g only
                const uint32_t hm = u > 0 ? 0xFFFFFFFFu : 0u;
#pragma unroll
                for (int k = 0; k < NR; k++) h[k] &= hm;
            }
            uint32_t wprev = __shfl_up_sync(0xffffffffu, h[NR - 1], 1, G);
            uint32_t wnext = __shfl_down_sync(0xffffffffu, h[0], 1, G);
            if (lane == 0) wprev = P2d;
            if (lane == G - 1) wnext = P2d;
            dir_update<NR>(h, wprev, wnext, c, P1d, a.one, Lh);
            if (!NV) {
                finish_dir<NR, G, SAFE, PAD>(Lh, s, P2d, keep);
#pragma unroll
                for (int k = 0; k < NR; k++) h[k] = Lh[k];
            }
        }
        // ---- vertical / diagonal paths: predecessors (u-1, u, u+1) of the previous row ----
        if (NV) {
            int16_t *wbase = act ? ownL : trashL;
#ifdef EXP_NOEDGEST
            const bool epred = false;
#else
            const bool epred = act && edgeRow;
#endif
            uint32_t Lv[3][NR];
#pragma unroll
            for (int j = 0; j < 3; j++) {
                if (NV == 1 && j != 1) continue;   // MODE_HH4: the vertical predecessor only
                const int offv = (((PH + AGG_SLOTS - 1 - AGG_LAG + j) % AGG_SLOTS) * 3 + j) * VS;   // written LAG+1-j pixel-times ago
                const int offe = (((PH + AGG_ENS - 1 + j) % AGG_ENS) * 3 + j) * VS;               // pixel u-1+j, u == t
                const int16_t *src = prevL + ((!AGG_EDGE_LIKE_ROW(G) && isRow0) ? offe : offv);
                if (j == 2 && u + 1 >= W1) src = zeroL;
                uint32_t p[NR];
                ld_regs_shared<NR>(src, p);
                const uint32_t wprev = *reinterpret_cast<const uint32_t *>(src - 2);
                const uint32_t wnext = *reinterpret_cast<const uint32_t *>(src + E);
                dir_update<NR>(p, wprev, wnext, c, P1d, a.one, Lv[j]);
                // the directions are finished in pairs: (horizontal, up-left) and (up, up-right); MODE_HH4: (horizontal, up)
                if (j == (NV == 1 ? 1 : 0)) {
                    finish_pair<NR, G, SAFE, PAD>(Lh, Lv[j], s, P2d, a.one, keep);
#pragma unroll
                    for (int k = 0; k < NR; k++) h[k] = Lh[k];
                }
                if (NV == 3 && j == 2) finish_pair<NR, G, SAFE, PAD>(Lv[1], Lv[2], s, P2d, a.one, keep);
                if (NV == 1 || j != 1) {
                    constexpr int J0 = 0;
#pragma unroll
                    for (int jj = (NV == 3 && j == 2) ? 1 : j; jj <= j; jj++) {
                        st_regs<NR>(wbase + ((PH % AGG_SLOTS) * 3 + jj) * VS, Lv[jj]);
                        st_regs_cg_pred<NR>(eptr + (PH * 3 + jj) * DP, Lv[jj], epred);
                    }
                    (void)J0;
                }
            }
        }
        if (!PASS2) {
            if (!SAFE) {
#pragma unroll
                for (int k = 0; k < NR; k++) s[k] = __vminu2(s[k], 0x7FFF7FFFu);
            }
            st_regs_cs_pred<NR>(sptr + XDIR * PH * DP, s, act);
        } else {
            if (SAFE) {   // pass-1 S joins last (plain packed adds commute): its load has the whole step to land
#pragma unroll
                for (int k = 0; k < NR; k++) s[k] += sbuf[PH % PFS][k];
                prefetch_s<PH % PFS, PH + PFS>();
            }
#pragma unroll
            for (int k = 0; k < NR; k++) {
                if (PAD) s[k] |= 0x7FFF7FFFu & ~keep[k];
                s[k] = __vminu2(s[k], 0x7FFF7FFFu);
            }
            // with SELPIPE the selection of this pixel runs one pixel-time later, interleaved with the next pixel's paths
            st_regs<NR>(scratch + (PH & 1) * DP + lane * E, s);
            if (!SELPIPE) select<PH>(parkw + PH, a.k16, a.tw);
        }
    }

    // pass 2: winner / second best / subpixel taps of the pixel whose clamped S sits in scratch half PHP & 1
    template <int PHP>
    __device__ __forceinline__ void select(uint4 *dst, uint32_t k16, int tw = 0)
    {
        const int16_t *sc = scratch + (PHP & 1) * DP;
        __syncwarp();
        uint32_t s[NR];
        ld_regs_shared<NR>(sc + lane * E, s);
        // winner: first d attaining min S  (key = S<<16 | d)
        const uint32_t d0 = (uint32_t)lane * E;
        uint32_t lk = 0xFFFFFFFFu;
#pragma unroll
        for (int k = 0; k < NR; k++) {
            lk = min(lk, s[k] * k16 + (d0 + 2 * k));                   // (S_lo << 16) | d, as an IMAD (FMA pipe)
            lk = min(lk, __byte_perm(s[k], d0 + 2 * k + 1, 0x3254));   // (S_hi << 16) | d
        }
        uint32_t best = group_min_u32<G>(lk);
        if (NV == 0 && tw) {
            // MODE_SGBM_3WAY (cv2's 8-lane SIMD search): in every residue class d mod 8 the LAST d attaining the
            // minimum is kept, then the smallest of those wins.  This lane's elements are d0 + e, d0 a multiple of 8:
            // residue j lives in elements j and j + 8 of every lane.
            const uint32_t minS = best >> 16;
            uint32_t bd = 0xFFFFu;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                uint32_t last = 0;                                   // (d + 1) of this lane's last match in class j
#pragma unroll
                for (int e = j; e < E; e += 8) {
                    const uint32_t v = (e & 1) ? (s[e >> 1] >> 16) : (s[e >> 1] & 0xFFFFu);
                    if (v == minS) last = d0 + e + 1;
                }
                last = group_max_u32<G>(last);
                if (last) bd = min(bd, last - 1);
            }
            best = (minS << 16) | bd;
        }
        const int bestD = (int)(best & 0xFFFFu);
        // second-best S outside bestD-1..bestD+1 (mask table indexed by bestD relative to this lane)
        uint32_t s2;
        {
            const int ti = min(max(bestD - (int)d0 + 2, 0), E + 3);
            uint32_t mk[NR], sm[NR];
            ld_regs_shared<NR>(masktab + ti * E, mk);
#pragma unroll
            for (int k = 0; k < NR; k++) sm[k] = s[k] | mk[k];
            s2 = group_min_u32<G>(lane_min16<NR>(sm));
        }
        // subpixel taps
        const uint32_t sm1 = (uint16_t)sc[max(bestD - 1, 0)];
        const uint32_t sp1 = (uint16_t)sc[min(bestD + 1, DP - 1)];
        // park the pixel's scalars (all group-uniform); the scalar tail runs lane-parallel in flush()
        if (lane == 0) *dst = make_uint4(best, s2, sm1 | (sp1 << 16), 0u);
    }

    // pass 2: uniqueness test, subpixel interpolation, disp2 scatter and store of the pixels parked since the last
    // flush, one pixel per lane: pixel-times tau0 .. tau0+n-1 (park slot = pixel-time mod 2*AGG_UNROLL)
    __device__ __forceinline__ void flush(const AggArgs &a, int16_t *dpre, unsigned *keys, int tau0, int n)
    {
        __syncwarp();
        for (int i = lane; i < n; i += G) {
            const int tau = tau0 + i, u = tau - lag;
            if (rowActive && (unsigned)u < (unsigned)W1) {
                const uint4 q = park[tau % (2 * AGG_UNROLL)];
                select_flush(a, dpre, keys, y, xbase + XDIR * u + a.minX1, q.x, q.y, q.z);
            }
        }
        __syncwarp();
    }
};

#ifdef EXP_NOBAR
__device__ __forceinline__ void cta_barrier() {}
#else
__device__ __forceinline__ void cta_barrier() { asm volatile("bar.sync 0;" ::: "memory"); }
#endif

// NV = 3: wavefront sweep with the three vertical/diagonal predecessors (+ one DMA warp per CTA);
// NV = 1: same machinery, vertical predecessor only (MODE_HH4: 2 + 2 paths);
// NV = 0: horizontal path only (rows independent; the 5th path of MODE_SGBM), no barriers, no DMA warp.
// MINB: minimum CTAs per SM handed to ptxas (0 = unconstrained).  Only the Dp = 128 kernels (G = 16, 288 threads) have a
// second instantiation with MINB = 2: 112 instead of 140-144 registers, two CTAs per SM — faster with many bands in flight
// (c2 MODE_HH at B = 32: pass 1 7.1 -> 5.6 ms, pass 2 10.9 -> 8.2 ms; B = 16: 3.82 -> 3.48 / 6.01 -> 5.13), slower with few
// (B = 12: pass 1 3.10 -> 3.29; B = 8: 2.34 -> 3.14 ms), so the launcher picks by grid size (AGG_MINB2_GRID).
template <int NR, int G, int NGRP, int NV, bool PASS2, bool SAFE, bool PAD, int PFD, bool REV = PASS2, int MINB = 0>
__global__ void __launch_bounds__(NGRP *G + (NV ? 32 : 0), MINB) k_sweep(const __grid_constant__ AggArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    typedef SweepSmem<NGRP, NV, PASS2> SM;
    constexpr int E = 2 * NR;
    constexpr int NCOMP = NGRP * G;
    const int Dp = a.Dp, W1 = a.W1, VS = SM::vs(Dp);
    int16_t *ring = reinterpret_cast<int16_t *>(smem_raw);
    int16_t *scratch_base = ring + SM::ring_elems(Dp);
    int16_t *masktab = scratch_base + SM::scratch_elems(Dp);
    int16_t *park_base = masktab + SM::mask_elems(E);   // 16-byte aligned: every region is a multiple of 8 elements
    __shared__ unsigned s_ticket;
    const int tid = threadIdx.x;
    unsigned ticket = blockIdx.x;
    if (NV) {
        if (tid == 0) s_ticket = atomicAdd(a.ticket, 1u);
        // ring: data 0, pads P2; (the barrier below also publishes s_ticket)
        const int n = SM::ring_elems(Dp);
        const int16_t p2 = (int16_t)(a.P2d & 0xFFFFu);
        for (int i = tid; i < n; i += blockDim.x) ring[i] = (i % VS) < AGG_PADE ? p2 : (int16_t)0;
    }
    if (PASS2) {
        // mask table: entry ti <-> r = ti-2 = bestD - d0; element e is masked (0x7FFF) iff |e - r| <= 1
        for (int i = tid; i < (E + 4) * E; i += blockDim.x) {
            const int r = i / E - 2, e = i % E;
            masktab[i] = (e >= r - 1 && e <= r + 1) ? (int16_t)0x7FFF : (int16_t)0;
        }
    }
    __syncthreads();
    if (NV) ticket = s_ticket;
    // band-major ticket order: consecutive bands of one frame start B tickets apart, so with enough frames in
    // flight a band's producer is already well ahead when the band gets an SM (no SM is held by a waiting band)
    const int frame = ticket % a.B, band = ticket / a.B;
    const size_t fo = (size_t)frame * a.ws_stride;
    const int16_t *Cv = reinterpret_cast<const int16_t *>(reinterpret_cast<const char *>(a.C) + fo);
    int16_t *Sv = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(a.S) + fo);
    int16_t *edge = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(a.edge) + fo);
    unsigned *prog = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(a.prog) + fo);
    int16_t *dpre = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(a.dpre) + fo);
    unsigned *keys = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(a.keys) + fo);
    const size_t edge_band = (size_t)W1 * 3 * Dp;
    int16_t *rowm1 = ring + (size_t)NGRP * AGG_SLOTS * 3 * VS;          // vectors [ENS][3]
    int16_t *zerov = rowm1 + (size_t)AGG_ENS * 3 * VS;
    const int lag_last = AGG_LAG * (NGRP - 1);
    const int nsteps = (W1 + (NV ? lag_last : 0) + AGG_UNROLL - 1) / AGG_UNROLL * AGG_UNROLL;

    if (NV && tid >= NCOMP) {
        // ================= DMA warp: cross-band hand-off =================
        const int lane = tid - NCOMP;
        const bool consumer = band > 0, producer = band + 1 < a.nbands;
        const int16_t *edge_in = edge + (size_t)(consumer ? band - 1 : 0) * edge_band;
        const unsigned *prog_in = prog + (consumer ? band - 1 : 0);
        constexpr int CPV = G * NR / 4;    // 16-byte chunks per vector (Dp / 8), a power of two
        constexpr int NCH = 3 * CPV;
        auto fetch = [&](int pe) {         // pixel pe of the producer band's last row -> row -1 ring
#ifdef EXP_FETCHHOT
            const int16_t *src = edge_in + (size_t)(pe & 63) * 3 * Dp;
#else
            const int16_t *src = edge_in + (size_t)pe * 3 * Dp;
#endif
            const int slot = AGG_EDGE_LIKE_ROW(G) ? (pe + AGG_SLOTS - AGG_LAG) % AGG_SLOTS : pe % AGG_ENS;
            int16_t *dst = rowm1 + (size_t)(slot * 3) * VS + AGG_PADE;
#ifndef EXP_NOFETCH
#pragma unroll
            for (int i0 = 0; i0 < NCH; i0 += 32) {
                const int i = i0 + lane;
                if (NCH % 32 == 0 || i < NCH) cp_async16(dst + (i / CPV) * VS + (i % CPV) * 8, src + i * 8);
            }
#endif
        };
        // Non-blocking hand-off: the progress counter is re-read every step with the load issued one step before its
        // value is used; pixels are pulled up to AGG_EPF ahead of row 0 as far as the producer allows (at most two per
        // step when catching up).  Only when the pixel row 0 needs next is not even published does the warp spin.
        // The edge vectors of a band 20 % of a row behind its producer have left L2 by the time they are needed (the
        // C/S streams of 148 bands pass through it), and under that load an HBM round trip is longer than the two
        // steps wait_group leaves a cp.async: pull every pixel's 6*Dp bytes into L2 AGG_L2PF pixels ahead.
        const int l2lines = (6 * Dp + 127) >> 7;
        auto l2fetch = [&](int pe) {
            if (lane < l2lines) prefetch_l2(reinterpret_cast<const char *>(edge_in + (size_t)pe * 3 * Dp) + lane * 128);
            if (l2lines > 32 && lane + 32 < l2lines)
                prefetch_l2(reinterpret_cast<const char *>(edge_in + (size_t)pe * 3 * Dp) + (lane + 32) * 128);
        };
        int l2pf = consumer ? 0 : W1;
        if (AGG_L2PF > 0 && consumer)
            for (; l2pf < min(AGG_L2PF, W1); l2pf++) l2fetch(l2pf);
        unsigned known = consumer ? 0u : (unsigned)W1, pending = 0;
        unsigned acq = known;              // progress value the last acquire fence covered: pixels < acq may be fetched
        int fetched = consumer ? 0 : W1;   // pixels [0, fetched) are issued
        int f1 = fetched, f2 = fetched;    // 'fetched' one and two steps ago
#ifdef EXP_NOPOLL
        known = (unsigned)W1;
#endif
        if (consumer) pending = ld_relaxed(prog_in);
        for (int t = 0; t < nsteps; t += AGG_PPS) {
            // this iteration runs alongside the compute lanes' pixel-times t-PPS..t-1 and prepares t..t+PPS-1,
            // in which row 0 reads pixels < t+PPS+1
            const int need = min(t + AGG_PPS + 1, W1);
            if (consumer) {
                known = max(known, pending);
                if (known < (unsigned)W1) pending = ld_relaxed(prog_in);
#ifdef EXP_HARNESS
                if ((int)known < need && lane == 0) { atomicAdd(&g_spin[0], 1ull); }
                { long long c0 = clock64();
#endif
                if ((int)known < need) {   // starving: wait until the producer is a.slack pixels further (hysteresis:
                    const int target = min(need + a.slack, W1);   // equal-speed bands would otherwise re-starve at once)
                    while ((int)known < target) known = ld_relaxed(prog_in);
                }
#ifdef EXP_HARNESS
                  if (lane == 0) atomicAdd(&g_spin[1], (unsigned long long)(clock64() - c0)); }
                if (lane == 0) atomicAdd(&g_spin[2], 1ull);
#endif
                constexpr int EPF = AGG_EDGE_LIKE_ROW(G) ? AGG_SLOTS - AGG_PPS - 2 : AGG_EPF;
                const int want = min(t + 1 + EPF, W1);
#ifndef AGG_NOFENCE
                if ((int)acq < want && known > acq) { fence_acquire(); acq = known; }   // orders the vector reads below after the flag read
#else
                acq = known;
#endif
                const int limit = min(want, (int)acq);
#pragma unroll
                for (int i = 0; i < 2 * AGG_PPS; i++)
                    if (fetched < limit) fetch(fetched++);
                if (AGG_L2PF > 0) {
                    const int l2lim = min(fetched + AGG_L2PF, W1);
#pragma unroll
                    for (int i = 0; i < 2 * AGG_PPS; i++)
                        if (l2pf < l2lim) l2fetch(l2pf++);
                }
            }
            cp_async_commit();
            // everything issued up to two iterations ago is complete after wait_group<2>; fall back when that is not enough
            if (f2 >= need) cp_async_wait<2>();
            else if (f1 >= need) cp_async_wait<1>();
            else cp_async_wait<0>();
            f2 = f1; f1 = fetched;
            cta_barrier();
            if (producer && lane == 0) {
                const int done = t - lag_last;   // pixels of the last row finished before this barrier
                if (done > 0 && done < W1 && (done & a.pubmask) < AGG_PPS) st_release(prog + band, (unsigned)done);
            }
        }
        cta_barrier();
        if (producer && lane == 0) st_release(prog + band, (unsigned)W1);
        return;
    }

    // ================= compute lanes =================
    SweepLane<NR, G, NGRP, NV, PASS2, SAFE, PAD, PFD, REV> st;
    const int g = tid / G, lane = tid % G;
    const int v = band * NGRP + g;
    typedef SweepLane<NR, G, NGRP, NV, PASS2, SAFE, PAD, PFD, REV> Lane;
    st.W1 = W1; st.VS = VS; st.lane = lane;
    st.lag = NV ? AGG_LAG * g : 0;
    st.rowActive = v < a.H;
    st.isRow0 = (g == 0);
    st.y = REV ? a.H - 1 - v : v;            // reverse sweeps run on the image rotated by 180 degrees (a.flip == REV)
    st.xbase = REV ? W1 - 1 : 0;
    st.P1d = a.P1d; st.P2d = a.P2d;
#pragma unroll
    for (int k = 0; k < NR; k++)
        st.keep[k] = ((lane * E + 2 * k < a.D) ? 0x0000FFFFu : 0u) | ((lane * E + 2 * k + 1 < a.D) ? 0xFFFF0000u : 0u);
    {
        const ptrdiff_t rowoff = (ptrdiff_t)(st.rowActive ? st.y : 0) * W1 * Dp + lane * E;
        const ptrdiff_t pix0 = (ptrdiff_t)(st.xbase - Lane::XDIR * st.lag) * Dp;   // pixel u = -lag (pixel-time 0)
        st.cptr = Cv + rowoff + pix0;
        st.sptr = Sv + rowoff + pix0;
        st.eptr = edge + (ptrdiff_t)band * edge_band - (ptrdiff_t)st.lag * 3 * Dp + lane * E;
    }
    st.ownL = ring + (size_t)(g * AGG_SLOTS * 3) * VS + AGG_PADE + lane * E;
    st.prevL = (g == 0 ? rowm1 : ring + (size_t)((g - 1) * AGG_SLOTS * 3) * VS) + AGG_PADE + lane * E;
    st.zeroL = zerov + AGG_PADE + lane * E;
    st.trashL = zerov + VS + AGG_PADE + lane * E;
    st.edgeRow = NV && band + 1 < a.nbands && g == NGRP - 1;
    st.scratch = scratch_base + (size_t)g * 2 * Dp;
    st.masktab = masktab;
    st.park = reinterpret_cast<uint4 *>(park_base) + g * 2 * AGG_UNROLL;
    st.parkw = st.park;
#pragma unroll
    for (int k = 0; k < NR; k++) st.h[k] = 0;
    st.template prefetch_c<0, 0>();
    if (PFD > 1) st.template prefetch_c<(1 < PFD ? 1 : 0), 1>();
    if (PFD > 2) st.template prefetch_c<(2 < PFD ? 2 : 0), 2>();
    if (PFD > 3) st.template prefetch_c<(3 < PFD ? 3 : 0), 3>();
    if (PFD > 4) st.template prefetch_c<(4 < PFD ? 4 : 0), 4>();
    if (PFD > 5) st.template prefetch_c<(5 < PFD ? 5 : 0), 5>();
    st.template prefetch_s<0, 0>();
    if (PFD > 1) st.template prefetch_s<(PFD > 1 ? 1 : 0), 1>();

    int ftau = 0;   // pass 2: first pixel-time not flushed yet
    for (int t0 = 0; t0 < nsteps; t0 += AGG_UNROLL) {
#define SVB_STEP(PH)                                   \
    if (NV && (PH) % AGG_PPS == 0) cta_barrier();      \
    st.template step<PH>(a, t0 + PH, dpre, keys);
        SVB_STEP(0) SVB_STEP(1) SVB_STEP(2) SVB_STEP(3) SVB_STEP(4) SVB_STEP(5) SVB_STEP(6) SVB_STEP(7)
#if AGG_UNROLL == 12
        SVB_STEP(8) SVB_STEP(9) SVB_STEP(10) SVB_STEP(11)
#endif
#undef SVB_STEP
        st.advance();
        if (PASS2 && ((t0 / AGG_UNROLL) & 1)) {   // every second iteration the parked pixels get their scalar tail
            const int upto = t0 + AGG_UNROLL - (Lane::SELPIPE ? 1 : 0);   // pixel-times < upto are parked
            st.flush(a, dpre, keys, ftau, upto - ftau);
            ftau = upto;
        }
    }
    if (PASS2) {
        if (Lane::SELPIPE) st.template select<AGG_UNROLL - 1>(st.park + (((st.parkw - st.park) ^ AGG_UNROLL) + AGG_UNROLL - 1), a.k16);
        st.flush(a, dpre, keys, ftau, nsteps - ftau);
    }
    if (NV) cta_barrier();
}

// ---------------- host side ----------------
#ifndef AGG_MINB2_GRID
#define AGG_MINB2_GRID (4 * 148)   // CTAs from which the two-CTAs-per-SM instantiation of the Dp = 128 kernels is used
#endif
template <int NR, int G, int NGRP, int NV, bool PASS2, bool SAFE, bool PAD, bool REV = PASS2, int MINB = 0>
static int launch_sweep_t(const AggArgs &a, int ngrid, cudaStream_t st)
{
#ifndef AGG_PFD2
#define AGG_PFD2 4
#endif
#ifndef AGG_PFD1
#define AGG_PFD1 6         // pass 1 has registers to spare (106 of 128): 6 pixels of C in flight instead of 4, 17.12 -> 16.89 ms
#endif
    constexpr int PFD = NR <= 4 ? (PASS2 ? AGG_PFD2 : AGG_PFD1) : 2;   // C prefetch depth in pixels (pass 2 is short of registers)
    const size_t smem = SweepSmem<NGRP, NV, PASS2>::bytes(a.Dp, 2 * NR);
    auto kern = k_sweep<NR, G, NGRP, NV, PASS2, SAFE, PAD, PFD, REV, MINB>;
    // the opt-in is per device and per function; setting it on every launch is cheap and stays correct with several
    // devices / host threads in one process
    SVB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<ngrid, NGRP * G + (NV ? 32 : 0), smem, st>>>(a);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}

template <int NV, bool PASS2, bool SAFE, bool PAD, bool REV>
static int launch_sweep_g(int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st)
{
    if (NR == 8) return launch_sweep_t<8, 32, 7, NV, PASS2, SAFE, PAD, REV>(a, ngrid, st);
    if (NR < 4) {   // wide-lane layouts of the latency form (a whole warp per row at Dp = 64 / 128): store-S sweeps only
        if constexpr (!PASS2 && NV != 1) {
            if (NR == 1 && G == 32) return launch_sweep_t<1, 32, 15, NV, PASS2, SAFE, PAD, REV>(a, ngrid, st);
            if (NR == 2 && G == 32) return launch_sweep_t<2, 32, 15, NV, PASS2, SAFE, PAD, REV>(a, ngrid, st);
        }
        svb_set_error("sweep: no kernel for NR=%d G=%d in this family", NR, G);
        return SVB_ERR_UNSUPPORTED;
    }
    switch (G) {
    case 2: return launch_sweep_t<4, 2, 16, NV, PASS2, SAFE, PAD, REV>(a, ngrid, st);
    case 4: return launch_sweep_t<4, 4, 16, NV, PASS2, SAFE, PAD, REV>(a, ngrid, st);
    case 8: return launch_sweep_t<4, 8, 16, NV, PASS2, SAFE, PAD, REV>(a, ngrid, st);
    case 16: return (NV && ngrid >= AGG_MINB2_GRID) ? launch_sweep_t<4, 16, 16, NV, PASS2, SAFE, PAD, REV, 2>(a, ngrid, st)
                                                    : launch_sweep_t<4, 16, 16, NV, PASS2, SAFE, PAD, REV>(a, ngrid, st);
    default: return launch_sweep_t<4, 32, 15, NV, PASS2, SAFE, PAD, REV>(a, ngrid, st);
    }
}

template <int NV, bool PASS2, bool REV = PASS2>
static int launch_sweep(int safe, int pad, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st)
{
    if (safe) return pad ? launch_sweep_g<NV, PASS2, true, true, REV>(NR, G, a, ngrid, st)
                         : launch_sweep_g<NV, PASS2, true, false, REV>(NR, G, a, ngrid, st);
    return pad ? launch_sweep_g<NV, PASS2, false, true, REV>(NR, G, a, ngrid, st)
               : launch_sweep_g<NV, PASS2, false, false, REV>(NR, G, a, ngrid, st);
}

// The seven (NV, PASS2, REV) families are compiled as separate translation units (build.py: -DSVB_AGG_PART=0..6) so
// that the 168 kernel instantiations build in parallel; without the define everything lands in one object.
#ifndef SVB_AGG_PART
#define SVB_AGG_PART -1
#endif
#ifndef EXP_HARNESS
#define SVB_SWEEP_DEF(NAME, NV, PASS2)                                                                       \
    int NAME(int safe, int pad, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st)                 \
    {                                                                                                        \
        return launch_sweep<NV, PASS2>(safe, pad, NR, G, a, ngrid, st);                                      \
    }
#define SVB_SWEEP_DEF_REVSTORE(NAME, NV)                                                                     \
    int NAME(int safe, int pad, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st)                 \
    {                                                                                                        \
        return launch_sweep<NV, false, true>(safe, pad, NR, G, a, ngrid, st);                                \
    }
int svb_sweep_nv3_rs(int safe, int pad, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st);
int svb_sweep_nv0_rs(int safe, int pad, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st);
#if SVB_AGG_PART == -1 || SVB_AGG_PART == 5
SVB_SWEEP_DEF_REVSTORE(svb_sweep_nv3_rs, 3)
#endif
#if SVB_AGG_PART == -1 || SVB_AGG_PART == 6
SVB_SWEEP_DEF_REVSTORE(svb_sweep_nv0_rs, 0)
#endif
int svb_sweep_nv3_p1(int safe, int pad, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st);
int svb_sweep_nv3_p2(int safe, int pad, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st);
int svb_sweep_nv1_p1(int safe, int pad, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st);
int svb_sweep_nv1_p2(int safe, int pad, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st);
int svb_sweep_nv0_p2(int safe, int pad, int NR, int G, const AggArgs &a, int ngrid, cudaStream_t st);
#if SVB_AGG_PART == -1 || SVB_AGG_PART == 0
SVB_SWEEP_DEF(svb_sweep_nv3_p1, 3, false)
#endif
#if SVB_AGG_PART == -1 || SVB_AGG_PART == 1
SVB_SWEEP_DEF(svb_sweep_nv3_p2, 3, true)
#endif
#if SVB_AGG_PART == -1 || SVB_AGG_PART == 2
SVB_SWEEP_DEF(svb_sweep_nv1_p1, 1, false)
#endif
#if SVB_AGG_PART == -1 || SVB_AGG_PART == 3
SVB_SWEEP_DEF(svb_sweep_nv1_p2, 1, true)
#endif
#if SVB_AGG_PART == -1 || SVB_AGG_PART == 4
SVB_SWEEP_DEF(svb_sweep_nv0_p2, 0, true)
#endif

#if SVB_AGG_PART == -1 || SVB_AGG_PART == 4
// ---- latency form: S = S1 + S2 (the two concurrent sweeps' sums), then the selection ----
// A group of G lanes owns one pixel at a time (E disparities per lane, as in k_sweep); a warp works on 32 / G pixels side
// by side and walks its share of the row.  Same selection rules as SweepLane::select + select_flush.
struct SelSumArgs {
    const int16_t *S1, *S2;
    AggArgs a;   // geometry, uniqueness, dpre / keys (C, S, edge, prog, ticket unused)
};
template <int NR, int G>
__global__ void __launch_bounds__(256) k_select_sum(const __grid_constant__ SelSumArgs q)
{
    constexpr int E = 2 * NR, DP = G * E, PPW = 32 / G;             // pixels per warp step
    __shared__ __align__(16) int16_t sc[8][PPW][DP];                  // per warp: the summed vectors (subpixel taps)
    const AggArgs &a = q.a;
    const int y = blockIdx.x, b = blockIdx.y;
    const int wid = threadIdx.x >> 5, lw = threadIdx.x & 31, sub = lw / G, lane = lw % G;
    const size_t fo = (size_t)b * a.ws_stride;
    const int16_t *S1 = reinterpret_cast<const int16_t *>(reinterpret_cast<const char *>(q.S1) + fo) + (size_t)y * a.W1 * DP;
    const int16_t *S2 = reinterpret_cast<const int16_t *>(reinterpret_cast<const char *>(q.S2) + fo) + (size_t)y * a.W1 * DP;
    int16_t *dpre = reinterpret_cast<int16_t *>(reinterpret_cast<char *>(a.dpre) + fo);
    unsigned *keys = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(a.keys) + fo);
    const uint32_t d0 = (uint32_t)lane * E;
    uint32_t keep[NR];
#pragma unroll
    for (int k = 0; k < NR; k++)
        keep[k] = ((d0 + 2 * k < (uint32_t)a.D) ? 0x0000FFFFu : 0u) | ((d0 + 2 * k + 1 < (uint32_t)a.D) ? 0xFFFF0000u : 0u);
    for (int x0 = wid * PPW; x0 < a.W1; x0 += 8 * PPW) {
        const int x = x0 + sub;
        const bool act = x < a.W1;
        uint32_t s[NR], t[NR];
        ld_regs_stream<NR>(S1 + (size_t)(act ? x : 0) * DP + d0, s);
        ld_regs_stream<NR>(S2 + (size_t)(act ? x : 0) * DP + d0, t);
#pragma unroll
        for (int k = 0; k < NR; k++) {
            s[k] = __vminu2(s[k] + t[k], 0x7FFF7FFFu);               // both <= 0x7FFF per half: no carry; saturating sum
            s[k] |= 0x7FFF7FFFu & ~keep[k];                          // padded disparities never win
        }
        int16_t *row = sc[wid][sub];
        __syncwarp();
        st_regs<NR>(row + d0, s);
        uint32_t lk = 0xFFFFFFFFu;
#pragma unroll
        for (int k = 0; k < NR; k++) {
            lk = min(lk, (s[k] << 16) | (d0 + 2 * k));
            lk = min(lk, (s[k] & 0xFFFF0000u) | (d0 + 2 * k + 1));
        }
        const uint32_t best = group_min_u32<G>(lk);
        const int bestD = (int)(best & 0xFFFFu);
        uint32_t m2 = 0x7FFFu;                                        // second best outside bestD-1 .. bestD+1
#pragma unroll
        for (int k = 0; k < NR; k++) {
            const int da = (int)d0 + 2 * k - bestD, db = da + 1;
            if (da < -1 || da > 1) m2 = min(m2, s[k] & 0xFFFFu);
            if (db < -1 || db > 1) m2 = min(m2, s[k] >> 16);
        }
        const uint32_t s2 = group_min_u32<G>(m2);
        __syncwarp();
        if (lane == 0 && act) {
            const uint32_t sm1 = (uint16_t)row[max(bestD - 1, 0)], sp1 = (uint16_t)row[min(bestD + 1, DP - 1)];
            select_flush(a, dpre, keys, y, x + a.minX1, best, s2, sm1 | (sp1 << 16));
        }
    }
}

template <int NR, int G>
static int launch_select_sum_t(const SelSumArgs &q, int H, int B, cudaStream_t st)
{
    k_select_sum<NR, G><<<dim3(H, B), 256, 0, st>>>(q);
    SVB_CUDA_CHECK(cudaGetLastError());
    return SVB_OK;
}
static int launch_select_sum(int NR, int G, const SelSumArgs &q, int H, int B, cudaStream_t st)
{
    if (NR == 8) return launch_select_sum_t<8, 32>(q, H, B, st);
    if (NR == 1) return launch_select_sum_t<1, 32>(q, H, B, st);
    if (NR == 2) return launch_select_sum_t<2, 32>(q, H, B, st);
    switch (G) {
    case 2: return launch_select_sum_t<4, 2>(q, H, B, st);
    case 4: return launch_select_sum_t<4, 4>(q, H, B, st);
    case 8: return launch_select_sum_t<4, 8>(q, H, B, st);
    case 16: return launch_select_sum_t<4, 16>(q, H, B, st);
    default: return launch_select_sum_t<4, 32>(q, H, B, st);
    }
}

// Rows [y0, y1) of the volume are aggregated as an image of their own (every mode but MODE_SGBM_3WAY: the whole image).
static int launch_aggregate_rows(svb_context *ctx, const SgbmPlan &pl, char *ws, size_t ws_stride, int B, int passes,
                                 int y0, int y1, cudaStream_t st)
{
    const int hs = y1 - y0, nbands = (hs + pl.NG - 1) / pl.NG;
    const size_t rowC = (size_t)pl.W1 * pl.Dp * 2;
    // reset schedulers / progress / scatter keys (of these rows)
    if (passes & 1) {
        SVB_CUDA_CHECK(cudaMemsetAsync(ctx->tickets, 0, 2 * sizeof(unsigned), st));
        // one strided memset per buffer for the whole batch (pitch = per-frame workspace stride)
        SVB_CUDA_CHECK(cudaMemset2DAsync(ws + pl.off_prog, ws_stride, 0, 2 * sizeof(unsigned) * pl.nbands, B, st));
        SVB_CUDA_CHECK(cudaMemset2DAsync(ws + pl.off_key + sizeof(unsigned) * (size_t)y0 * pl.W, ws_stride, 0xFF,
                                         sizeof(unsigned) * (size_t)hs * pl.W, B, st));
    }
    AggArgs a;
    memset(&a, 0, sizeof a);
    a.C = (const int16_t *)(ws + pl.off_C + (size_t)y0 * rowC);
    a.S = (int16_t *)(ws + pl.off_S + (size_t)y0 * rowC);
    a.edge = (int16_t *)(ws + pl.off_edge);
    a.dpre = (int16_t *)(ws + pl.off_dpre) + (size_t)y0 * pl.W;
    a.keys = (unsigned *)(ws + pl.off_key) + (size_t)y0 * pl.W;
    a.ws_stride = ws_stride;
    a.H = hs; a.W = pl.W; a.W1 = pl.W1; a.D = pl.D; a.Dp = pl.Dp; a.minX1 = pl.minX1; a.minD = pl.p.minDisparity;
    a.nbands = nbands; a.B = B;
    a.P1d = (uint32_t)pl.p.P1 * 0x10001u;
    a.P2d = (uint32_t)pl.p.P2 * 0x10001u;
    a.one = 1;
    a.k16 = 65536;
    a.uniq = pl.p.uniquenessRatio;
    a.invalid16 = pl.invalid16;
    a.tw = pl.p.mode == 2;
    a.pubmask = AGG_PUB - 1;
    a.slack = 0;   // (hysteresis for starving bands: measured no better than 0 since the edges are L2-prefetched)
    const int pad = pl.Dp != pl.D;
    const int NR = pl.E / 2;
    // pass 1: left, up-left, up, up-right
    a.flip = 0;
    a.prog = (unsigned *)(ws + pl.off_prog);
    a.ticket = ctx->tickets;
    int rc = SVB_OK;
    const bool two = pl.p.mode == 3 || pl.p.mode == 2;   // MODE_HH4: left + up, then right + down; 3WAY: left + up, then right
    if (passes & 1)
        rc = two ? svb_sweep_nv1_p1(pl.safe, pad, NR, pl.G, a, nbands * B, st)
                 : svb_sweep_nv3_p1(pl.safe, pad, NR, pl.G, a, nbands * B, st);
    if (rc || !(passes & 2)) return rc;
    a.flip = 1;
    if (pl.p.mode == 1 || pl.p.mode == 3) {   // MODE_HH: right, down-right, down, down-left + selection; MODE_HH4: right, down
        a.prog = (unsigned *)(ws + pl.off_prog) + pl.nbands;
        a.ticket = ctx->tickets + 1;
        rc = pl.p.mode == 3 ? svb_sweep_nv1_p2(pl.safe, pad, NR, pl.G, a, nbands * B, st)
                            : svb_sweep_nv3_p2(pl.safe, pad, NR, pl.G, a, nbands * B, st);
    } else {                // MODE_SGBM: 5th path (from the right) + selection, rows independent; 3WAY: its third path
        rc = svb_sweep_nv0_p2(pl.safe, pad, NR, pl.G, a, nbands * B, st);
    }
    return rc;
}

int svb_launch_aggregate(svb_context *ctx, const SgbmPlan &pl, char *ws, size_t ws_stride, int B, int passes,
                         cudaStream_t st)
{
    return launch_aggregate_rows(ctx, pl, ws, ws_stride, B, passes, 0, pl.H, st);
}

// Latency form (SgbmPlan::latency): the forward sweep on `st`, the reverse sweep on ctx->aux at the same time (own sum
// volume, own band-edge buffer, own tickets / progress counters), joined by events, then k_select_sum.
int svb_launch_aggregate_concurrent(svb_context *ctx, const SgbmPlan &pl, char *ws, size_t ws_stride, int B, cudaStream_t st,
                                    cudaEvent_t mark_pass1)
{
    SVB_CUDA_CHECK(cudaMemsetAsync(ctx->tickets, 0, 2 * sizeof(unsigned), st));
    SVB_CUDA_CHECK(cudaMemset2DAsync(ws + pl.off_prog, ws_stride, 0, 2 * sizeof(unsigned) * pl.nbands, B, st));
    SVB_CUDA_CHECK(cudaMemset2DAsync(ws + pl.off_key, ws_stride, 0xFF, sizeof(unsigned) * (size_t)pl.H * pl.W, B, st));
    AggArgs a;
    memset(&a, 0, sizeof a);
    a.C = (const int16_t *)(ws + pl.off_C);
    a.dpre = (int16_t *)(ws + pl.off_dpre);
    a.keys = (unsigned *)(ws + pl.off_key);
    a.ws_stride = ws_stride;
    a.H = pl.H; a.W = pl.W; a.W1 = pl.W1; a.D = pl.D; a.Dp = pl.Dp; a.minX1 = pl.minX1; a.minD = pl.p.minDisparity;
    a.nbands = pl.nbands; a.B = B;
    a.P1d = (uint32_t)pl.p.P1 * 0x10001u;
    a.P2d = (uint32_t)pl.p.P2 * 0x10001u;
    a.one = 1; a.k16 = 65536;
    a.uniq = pl.p.uniquenessRatio; a.invalid16 = pl.invalid16; a.tw = 0;
    a.pubmask = AGG_PUB - 1;   // (publishing every barrier is slower even for one frame: st.release costs more than the
                               // shorter hand-off gains — tools/exp_sweep.cu, B = 1: 3.54 ms at 8 pixels, 4.53 ms at 2)
    a.slack = 0;
    const int pad = pl.Dp != pl.D, NR = pl.E / 2;
    SVB_CUDA_CHECK(cudaEventRecord(ctx->ev_fork, st));
    SVB_CUDA_CHECK(cudaStreamWaitEvent(ctx->aux, ctx->ev_fork, 0));
    // reverse sweep (aux): MODE_HH right, down-right, down, down-left; MODE_SGBM: the single path from the right
    AggArgs r = a;
    r.flip = 1;
    r.S = (int16_t *)(ws + pl.off_S2);
    r.edge = (int16_t *)(ws + pl.off_edge2);
    r.prog = (unsigned *)(ws + pl.off_prog) + pl.nbands;
    r.ticket = ctx->tickets + 1;
    int rc = pl.p.mode == 1 ? svb_sweep_nv3_rs(pl.safe, pad, NR, pl.G, r, pl.nbands * B, ctx->aux)
                            : svb_sweep_nv0_rs(pl.safe, pad, NR, pl.G, r, pl.nbands * B, ctx->aux);
    if (rc) return rc;
    SVB_CUDA_CHECK(cudaEventRecord(ctx->ev_join, ctx->aux));
    // forward sweep (st): left, up-left, up, up-right
    a.flip = 0;
    a.S = (int16_t *)(ws + pl.off_S);
    a.edge = (int16_t *)(ws + pl.off_edge);
    a.prog = (unsigned *)(ws + pl.off_prog);
    a.ticket = ctx->tickets;
    if ((rc = svb_sweep_nv3_p1(pl.safe, pad, NR, pl.G, a, pl.nbands * B, st))) return rc;
    if (mark_pass1) cudaEventRecord(mark_pass1, st);
    SVB_CUDA_CHECK(cudaStreamWaitEvent(st, ctx->ev_join, 0));
    SelSumArgs q;
    q.S1 = (const int16_t *)(ws + pl.off_S);
    q.S2 = (const int16_t *)(ws + pl.off_S2);
    q.a = a;
    return launch_select_sum(NR, pl.G, q, pl.H, B, st);
}

// MODE_SGBM_3WAY (oracle: sgbm_3way_raw): four row stripes of ceil(H/4) rows, each started `overlap` rows early with
// its paths and block sums restarted there; a stripe keeps only its own rows.  The stripes run last to first, so that
// the rows a stripe computes ahead of its own range are overwritten by their owner afterwards.
int svb_launch_cost_aggregate_3way(svb_context *ctx, const SgbmPlan &pl, char *ws, size_t ws_stride, int B, cudaStream_t st)
{
    const int nstripes = 4, stripe = (pl.H + nstripes - 1) / nstripes;
    const int overlap = pl.p.blockSize / 2 + 1 + (stripe + 9) / 10;   // ceil(0.1 * stripe)
    for (int n = nstripes - 1; n >= 0; n--) {
        const int ys = n == 0 ? 0 : std::max(std::min(n * stripe - overlap, pl.H), 0);
        const int ye = n == nstripes - 1 ? pl.H : std::min((n + 1) * stripe, pl.H);
        const int keep0 = n == 0 ? 0 : std::min(n * stripe, pl.H);
        if (ye <= ys || keep0 >= ye) continue;
        int rc;
        if ((rc = svb_launch_cost(pl, ws, ws_stride, B, ys, ye, st))) return rc;
        if ((rc = launch_aggregate_rows(ctx, pl, ws, ws_stride, B, 3, ys, ye, st))) return rc;
    }
    return SVB_OK;
}
#endif   // SVB_AGG_PART (host dispatch lives in part 4)
#endif   // !EXP_HARNESS
