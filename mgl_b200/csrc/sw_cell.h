/*
 * sw_cell.h -- one DP cell of the affine-gap recurrence, in two arithmetics.
 *
 * The recurrence follows the reference's scalar loop
 * (/root/reference/src/main/native/mgl_sw/sw.cpp:54-93):
 *     d  = H[i-1][j-1] + (a == b ? match : mismatch)                 :55
 *     H  = max(d, E, F)                                              :60-71
 *     E' = max(E - ge, H - go)                                       :73-82
 *     F' = max(F - ge, H - go)                                       :84-93
 * The fill computes VALUES only.  Which of d / F / E won, and whether a gap was
 * opened or extended (the reference's per-cell gap length, :60-93), is decided
 * by the traceback, which recomputes the cells along its path from the fill's
 * checkpoints with the reference's comparisons verbatim (sw_trace.h).
 * Base comparison is raw byte equality (:55), so bytes are compared as they
 * come (mismatch iff a - b != 0) -- no 2-bit recoding.
 *
 * S16x2: two pairs per register, unsigned 16-bit halves holding score + bias.
 *        Plain 32-bit adds/subs/IMADs act on both halves at once; that is
 *        exact as long as every half stays inside [0, 65535], which the host
 *        planner guarantees before routing a pair here (sw_plan.cpp).  max3
 *        and add-then-max are the Blackwell DPX instructions VIMNMX3.U16x2 and
 *        VIADDMNMX.U16x2.
 * S32:   one pair per register, signed 32-bit (VIMNMX3.S32, VIADDMNMX.S32).
 */
#ifndef MGL_B200_SW_CELL_H
#define MGL_B200_SW_CELL_H
#include "sw_common.h"

namespace mglsw {

#if defined(__CUDA_ARCH__)
#define SW_ON_DEVICE 1
#else
#define SW_ON_DEVICE 0
#endif

struct S16x2 {
    static constexpr int KIND = KIND_S16X2;
    static constexpr int PAIRS = 2;

    SW_HD static u32 rep(int v) { return ((u32)v & 0xffffu) * 0x00010001u; }
    SW_HD static int half(u32 v, int h) { return (int)((v >> (16 * h)) & 0xffffu); }

    /* per half: 0 - byte (mod 2^16); a lane's alt bytes are stored this way once per strip */
    SW_HD static u32 negb(u32 b) { return ((0u - (b & 0xffffu)) & 0xffffu) | ((0u - (b >> 16)) << 16); }
    /* mismatch indicator, per half: min((a - b) mod 2^16, 1), with nb = negb(b): ONE instruction (VIADDMNMX.U16x2) */
    SW_HD static u32 ne1(u32 a, u32 nb)
    {
#if SW_ON_DEVICE
        return __viaddmin_u16x2(a, nb, 0x00010001u);
#else
        const u32 lo = (a + nb) & 0xffffu, hi = ((a >> 16) + (nb >> 16)) & 0xffffu;
        return (lo ? 1u : 0u) | (hi ? 0x10000u : 0u);
#endif
    }
    SW_HD static u32 max3(u32 a, u32 b, u32 c)
    {
#if SW_ON_DEVICE
        return __vimax3_u16x2(a, b, c);
#else
        u32 lo = a & 0xffffu, hi = a >> 16;
        if ((b & 0xffffu) > lo) lo = b & 0xffffu;
        if ((c & 0xffffu) > lo) lo = c & 0xffffu;
        if ((b >> 16) > hi) hi = b >> 16;
        if ((c >> 16) > hi) hi = c >> 16;
        return lo | (hi << 16);
#endif
    }
    SW_HD static u32 addmax(u32 a, u32 b, u32 c)       /* per half: max(a + b mod 2^16, c) */
    {
#if SW_ON_DEVICE
        return __viaddmax_u16x2(a, b, c);
#else
        u32 lo = (a + b) & 0xffffu, hi = ((a >> 16) + (b >> 16)) & 0xffffu;
        if ((c & 0xffffu) > lo) lo = c & 0xffffu;
        if ((c >> 16) > hi) hi = c >> 16;
        return lo | (hi << 16);
#endif
    }
};

struct S32 {
    static constexpr int KIND = KIND_S32;
    static constexpr int PAIRS = 1;

    SW_HD static u32 rep(int v) { return (u32)v; }
    SW_HD static int half(u32 v, int) { return (int)v; }

    SW_HD static u32 negb(u32 b) { return 0u - b; }
    /* mismatch indicator: min((a - b) mod 2^32, 1) as unsigned, with nb = negb(b): ONE instruction (VIADDMNMX.U32) */
    SW_HD static u32 ne1(u32 a, u32 nb)
    {
#if SW_ON_DEVICE
        return __viaddmin_u32(a, nb, 1u);
#else
        return (a + nb) ? 1u : 0u;
#endif
    }
    SW_HD static u32 max3(u32 a, u32 b, u32 c)
    {
#if SW_ON_DEVICE
        return (u32)__vimax3_s32((int)a, (int)b, (int)c);
#else
        int m = (int)a;
        if ((int)b > m) m = (int)b;
        if ((int)c > m) m = (int)c;
        return (u32)m;
#endif
    }
    SW_HD static u32 addmax(u32 a, u32 b, u32 c)
    {
#if SW_ON_DEVICE
        return (u32)__viaddmax_s32((int)a, (int)b, (int)c);
#else
        const int s = (int)(a + b);
        return (u32)(s > (int)c ? s : (int)c);
#endif
    }
};

template <class T>
SW_HD CellConsts make_consts(const Scores &s, int bias)
{
    CellConsts c;
    c.neg_delta = (u32)(-(s.match - s.mismatch));
    c.neg_ge = T::rep(-s.ge);
    c.match = T::rep(s.match);
    c.one = 1u;
    c.neg_go = 0u - T::rep(s.go);       /* a 32-bit negation of the packed constant: added with a plain 32-bit add */
    c.bias = bias;
    c.s_match = s.match; c.s_mismatch = s.mismatch; c.s_go = s.go; c.s_ge = s.ge;
    return c;
}

/*
 * One cell.  In:  ra packed ref bytes, nbq the lane's alt bytes NEGATED per half (T::negb, once per
 * strip), hdm = H[i-1][j-1] + match (+bias), e = E[i][j], f = F[i][j].
 * Out: e = E[i+1][j], f = F[i][j+1], hm = H[i][j] + match.
 *
 * Seven instructions.  The byte comparison (sw.cpp:55) is one add-then-min: (a - b) mod 2^16 is 0
 * for equal bytes and >= 1 otherwise, so min(a + (-b), 1) is the mismatch indicator (it was xor + min:
 * one instruction more on the ALU pipe, which is the pipe that binds this kernel).  The add-min, max3
 * and the two add-max exist only on the ALU pipe; the three plain adds are written as multiply-adds
 * (by c.neg_delta, and by the runtime 1 in c.one) so that they issue to the FMA pipe -- an SM
 * sub-partition owns one 16-lane pipe of each kind and one issue slot per clock (int_peak.cu
 * measures both): 4 ALU + 3 FMA per cell.
 *
 * Measured and dropped on this form [B200, profiles/r08a, r08b]: (1) taking F off the max3 -> add -> add-max chain
 * (F' = max(F - min(ge, go), max(d, E) - go): one add-max links a row's cells instead of three instructions) costs
 * one instruction more per pipe and was 4 % SLOWER (fill 15.0 against 14.4 ms): the step is not bound by that
 * chain; (2) the two plain adds as three-input IADD3 instead of IMAD: 12 % slower (15.8 against 14.1 ms), IADD3
 * shares the ALU pipe's dispatch.
 */
template <class T>
SW_HD void sw_cell(const CellConsts &c, u32 ra, u32 nbq, u32 hdm, u32 &e, u32 &f, u32 &hm)
{
    const u32 m = T::ne1(ra, nbq);            /* 0 match, 1 mismatch                     */
    const u32 d = m * c.neg_delta + hdm;      /* H_diag + match - m * (match - mismatch) */
    const u32 h = T::max3(d, e, f);
    const u32 hgo = h * c.one + c.neg_go;
    e = T::addmax(e, c.neg_ge, hgo);
    f = T::addmax(f, c.neg_ge, hgo);
    hm = h * c.one + c.match;
}

}  // namespace mglsw
#endif
