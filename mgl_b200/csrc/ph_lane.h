/*
 * ph_lane.h -- the PairHMM forward algorithm, one lane's view.
 *
 * Replaces the reference's inner loops (compute_full_prob_core_1X, compute_prob_scalar.cc:5-44, and the AVX2
 * bands of compute_prob_avx{f,d}.cc) for one (read, haplotype) pair.  A group of G lanes (a whole warp, or half a warp
 * with another pair in the other half) takes a band of G * RPL read rows:
 * lane l owns rows l*RPL .. l*RPL + RPL - 1 of the band and sweeps the haplotype columns one per step, one column
 * behind lane l - 1, so at step t it is at column t - l + 1.  Everything that is constant along a row (the five
 * transition probabilities of compute_prob_scalar.cc:73-81 and the two emission values :25-34) sits in the lane's
 * registers for the whole band; per step a lane receives (M, X, Y) of the row above its first row from lane l - 1
 * (one shuffle each) and keeps the previous column of its own rows.  Reads longer than one band go band by band:
 * the last lane parks its bottom row in a carry row (shared memory on the device) that lane 0 reads back in the
 * next band -- the same hand-over the reference's AVX bands do through pRow0/pRow1 (compute_prob_avxf.cc:670-672).
 *
 *   M[r][c] = distm * (M[r-1][c-1] * pMM[r] + (X[r-1][c-1] + Y[r-1][c-1]) * pGapM[r])      compute_prob_scalar.cc:39
 *   Y[r][c] = M[r][c-1] * pMY[r] + Y[r][c-1] * pZZ[r]                                      :41
 *   X[r][c] = M[r-1][c] * pMX[r] + X[r-1][c] * pZZ[r]                                      :43
 *   row 0: M = X = 0, Y = INITIAL_CONSTANT / hapLen (:104,119-125); column 0 of rows >= 1: all 0 (:140-142)
 *   result = sum over columns 1..hapLen of M[readLen][c] + X[readLen][c], left to right (:213,310)
 *
 * PhLane is PH_HD: ph_kernel.cu drives 32 of them in a warp with __shfl_up_sync; tests/emu/ph_emu.cpp drives 32 of
 * them in a loop on the CPU.
 */
#ifndef MGL_B200_PH_LANE_H
#define MGL_B200_PH_LANE_H
#include <stdint.h>

#if defined(__CUDACC__)
#define PH_HD __host__ __device__ __forceinline__
#else
#define PH_HD inline
#endif

namespace mglph {

/* quality tables in the reference's arithmetic (built on the host by ph_tables.cpp from Context.h) */
template <class N>
struct Tables {
    const N *ph2pr;       /* 128 entries: 10^(-q/10)                                   (Context.h:101,137)  */
    const N *mm;          /* matchToMatchProb, triangular, qualities 0..127            (Context.h:47-58)    */
    N initial;            /* INITIAL_CONSTANT: 2^120 float, 2^1020 double              (Context.h:105,141)  */
};

struct PairView {
    const uint8_t *rs, *q, *ins, *del, *gcp;   /* the read's five tracks                                    */
    const uint8_t *hap;
    int rl, hl;
};

template <class N>
struct Top { N m, x, y; };

/*
 * The emission value of one cell: `a` if the read base equals the haplotype base or the haplotype base is 'N', else
 * `b` (compute_prob_scalar.cc:25-34; a read 'N' is folded into b by band_begin).  On the device, in float, this is
 * one compare-with-or and one select; the compiler's own rendering of the C expression is a compare and two selects.
 */
template <class N>
PH_HD N ph_emission(unsigned base, unsigned hb, unsigned hap_n, N a, N b)
{
    return (base == hb || hap_n) ? a : b;
}
#if defined(__CUDA_ARCH__)
template <>
__device__ __forceinline__ float ph_emission<float>(unsigned base, unsigned hb, unsigned hap_n, float a, float b)
{
    float d;
    asm("{\n\t.reg .pred q, p;\n\tsetp.ne.u32 q, %3, 0;\n\tsetp.eq.or.u32 p, %1, %2, q;\n\tselp.f32 %0, %4, %5, p;\n\t}"
        : "=f"(d) : "r"(base), "r"(hb), "r"(hap_n), "f"(a), "f"(b));
    return d;
}
#endif

template <class N, int RPL, int G = 32>
struct PhLane {
    N pMM[RPL], pGapM[RPL], pMX[RPL], pMY[RPL], pZZ[RPL];
    N dmatch[RPL], dmis[RPL];
    unsigned base[RPL];              /* the read base of each row (a byte array would live in local memory) */
    N Mp[RPL], Xp[RPL], Yp[RPL];    /* previous column of my rows                                           */
    Top<N> tprev;                   /* the row above my first row, at my previous column                    */
    Top<N> bottom;                  /* my last row at the column I computed last: what lane l + 1 receives  */
    N result;
    int lane;                       /* inside the group of G lanes                                           */
    bool last;                      /* my bottom row is the read's last row (in this band)                   */

    /*
     * Rows [band * G * RPL + lane * RPL - pad, + RPL) of the read, pad = ph_pad(rl, RPL): the read is pushed down
     * by `pad` rows so that its LAST row is always some lane's bottom row (the sum over the last row then needs no
     * indexed access to the row registers, which would put them in local memory).  The pad rows above the read
     * behave as DP row 0: M = X = 0 and Y = yInitial in every column.
     */
    PH_HD void band_begin(const Tables<N> &tb, const PairView &pv, int band, int lane_, N y_initial)
    {
        lane = lane_;
        const int pad = (RPL - pv.rl % RPL) % RPL;
        last = band * G * RPL + lane * RPL + (RPL - 1) - pad == pv.rl - 1;
        const N one = (N)1.0, three = (N)3.0;
#pragma unroll
        for (int k = 0; k < RPL; k++) {
            const int row = band * G * RPL + lane * RPL + k - pad;     /* 0-based index into the read */
            Mp[k] = Xp[k] = Yp[k] = (N)0;                              /* column 0 of rows >= 1 */
            if (row < 0) {                                             /* a copy of DP row 0: Y carries yInitial along, M and X stay 0 */
                pMM[k] = pGapM[k] = pMX[k] = pMY[k] = dmatch[k] = dmis[k] = (N)0;
                pZZ[k] = one;
                base[k] = 0;
                Yp[k] = y_initial;
            } else if (row < pv.rl) {
                const int qi = pv.ins[row] & 127, qd = pv.del[row] & 127, qc = pv.gcp[row] & 127;
                const int lo = qi <= qd ? qi : qd, hi = qi <= qd ? qd : qi;   /* Context.h:118-128 */
                pMM[k] = tb.mm[((hi * (hi + 1)) >> 1) + lo];
                pGapM[k] = one - tb.ph2pr[qc];
                pMX[k] = tb.ph2pr[qi];
                pMY[k] = tb.ph2pr[qd];
                pZZ[k] = tb.ph2pr[qc];
                const N dq = tb.ph2pr[pv.q[row] & 127];
                dmatch[k] = one - dq;
                base[k] = pv.rs[row];
                /* distm * threeOver (compute_prob_scalar.cc:16,33) -- or, for a read base 'N', which matches every
                   haplotype base (:25), the match value again: the row-constant half of the test costs nothing per cell */
                dmis[k] = base[k] == (unsigned)'N' ? dmatch[k] : dq * (one / three);
            } else {                                                   /* below the read: computes zeros nobody uses */
                pMM[k] = pGapM[k] = pMX[k] = pMY[k] = pZZ[k] = dmatch[k] = dmis[k] = (N)0;
                base[k] = 0;
            }
        }
        /* the row above my first row at column 0: DP row 0 (Y = yInitial) only for lane 0 of band 0 */
        tprev.m = (N)0; tprev.x = (N)0;
        tprev.y = (band == 0 && lane == 0) ? y_initial : (N)0;
        bottom.m = bottom.x = bottom.y = (N)0;
        if (band == 0) result = (N)0;
    }

    /*
     * One step: my column is c = t - lane + 1; `top` is the row above my first row at column c (from lane l - 1's
     * `bottom`, or for lane 0 DP row 0 / the carry row); hb = hap_byte(pv, t), fetched by the caller a step ahead so
     * that the load is not on the step's critical path.  Returns true if the column was inside 1..hl.
     */
    PH_HD bool step(const PairView &pv, int t, const Top<N> &top, unsigned hb)
    {
        const int c = t - lane + 1;
        if (c < 1 || c > pv.hl) return false;
        N dM = tprev.m, dX = tprev.x, dY = tprev.y;
        N uM = top.m, uX = top.x;
        const unsigned hap_n = hb == (unsigned)'N' ? 1u : 0u;
#pragma unroll
        for (int k = 0; k < RPL; k++) {
            const N distm = ph_emission<N>(base[k], hb, hap_n, dmatch[k], dmis[k]);
            const N Mn = distm * (dM * pMM[k] + (dX + dY) * pGapM[k]);
            const N Yn = Mp[k] * pMY[k] + Yp[k] * pZZ[k];
            const N Xn = uM * pMX[k] + uX * pZZ[k];
            dM = Mp[k]; dX = Xp[k]; dY = Yp[k];
            uM = Mn; uX = Xn;
            Mp[k] = Mn; Xp[k] = Xn; Yp[k] = Yn;
        }
        tprev = top;
        bottom.m = Mp[RPL - 1]; bottom.x = Xp[RPL - 1]; bottom.y = Yp[RPL - 1];
        if (last) result += bottom.m + bottom.x;
        return true;
    }
};

/* the haplotype base lane `lane` meets at step t.  Outside columns 1..hl the step is skipped and any byte will do:
   the load is unconditional, so the buffer holding the haplotypes is readable MGLPH_HAP_PAD bytes before its first
   and after its last byte (the skew reaches 31 columns to either side), plus the longest haplotype behind the last
   byte (a group sharing its warp with a longer haplotype keeps stepping). */
#define MGLPH_HAP_PAD 64
PH_HD unsigned ph_hap_byte(const PairView &pv, int t, int lane)
{
    return (unsigned)pv.hap[t - lane];
}

/* rows the read is pushed down by, bands it then needs, and the lane whose bottom row is its last row */
PH_HD int ph_pad(int rl, int rpl) { return (rpl - rl % rpl) % rpl; }
PH_HD int ph_bands(int rl, int rpl, int g = 32) { return (rl + ph_pad(rl, rpl) + g * rpl - 1) / (g * rpl); }
PH_HD int ph_last_lane(int rl, int rpl, int g = 32) { return ((rl + ph_pad(rl, rpl) - 1) % (g * rpl)) / rpl; }

}  // namespace mglph
#endif
