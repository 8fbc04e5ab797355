/*
 * sw_cell.h -- one DP cell of the affine-gap recurrence, in two arithmetics.
 *
 * The recurrence and every tie-break follow the reference's scalar loop
 * (/root/reference/src/main/native/mgl_sw/sw.cpp:54-93):
 *     d  = H[i-1][j-1] + (a == b ? match : mismatch)                 :55
 *     H  = d if d >= E and d >= F; else F if F >= E; else E          :60-71
 *     E' = H - go if H - go > E - ge else E - ge   (tie => extend)   :73-82
 *     F' = H - go if H - go > F - ge else F - ge   (tie => extend)   :84-93
 * The reference stores a gap LENGTH per cell; here a cell stores the four
 * predicates from which the traceback recovers the same lengths
 * (sw_layout.h).  Base comparison is raw byte equality (:55), so bytes are
 * compared as they come (x = a ^ b; mismatch iff x != 0) -- no 2-bit recoding.
 *
 * S16x2: two pairs per register, unsigned 16-bit halves holding score + bias.
 *        Plain 32-bit adds/subs/IMADs act on both halves at once; that is
 *        exact as long as every half stays inside [0, 65535], which the host
 *        planner guarantees before routing a pair here (sw_plan.cpp).  max3
 *        and add-then-max are the Blackwell DPX instructions VIMNMX3.U16x2 and
 *        VIADDMNMX.U16x2.
 * S32:   one pair per register, signed 32-bit (VIMNMX3.S32, VIADDMNMX.S32).
 *
 * Predicates without compare instructions: X >= Y  <=>  bit p of
 * (X - Y + 2^p) is set, provided |X - Y| < 2^p.  The four predicates use four
 * adjacent bit positions so that three LOP3 bit-selects merge them into a
 * nibble.  The planner checks the score-dependent bounds on |X - Y|.
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

/*
 * (a & mask) | (b & ~mask) as ONE LOP3.  Written as inline PTX on the device: left to the
 * compiler, a chain of bit-selects whose unselected bits are dead is re-associated into
 * and/or pairs that cost one more LOP3 per cell.
 */
SW_HD u32 bitsel(u32 mask, u32 a, u32 b)
{
#if SW_ON_DEVICE
    u32 r;
    asm("lop3.b32 %0, %1, %2, %3, 0xCA;" : "=r"(r) : "r"(mask), "r"(a), "r"(b));
    return r;
#else
    return (a & mask) | (b & ~mask);
#endif
}

struct S16x2 {
    static constexpr int KIND = KIND_S16X2;
    static constexpr int PAIRS = 2;
    static constexpr u32 SPW = 4;                       /* steps per trace word */
    static constexpr u32 M1 = 0x80008000u, M2 = 0x40004000u, M3 = 0x20002000u, M4 = 0x10001000u;

    SW_HD static u32 rep(int v) { return ((u32)v & 0xffffu) * 0x00010001u; }
    SW_HD static u32 pack(u32 lo, u32 hi) { return (lo & 0xffffu) | (hi << 16); }
    SW_HD static int half(u32 v, int h) { return (int)((v >> (16 * h)) & 0xffffu); }

    SW_HD static u32 min1(u32 x)
    {
#if SW_ON_DEVICE
        return __vminu2(x, 0x00010001u);
#else
        return ((x & 0xffffu) ? 1u : 0u) | ((x >> 16) ? 0x10000u : 0u);
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
    /* append the step's nibble (top nibble of each half of w) to the accumulator */
    SW_HD static u32 push(u32 acc, u32 w) { return bitsel(0xF000F000u, w, acc >> 4); }
};

struct S32 {
    static constexpr int KIND = KIND_S32;
    static constexpr int PAIRS = 1;
    static constexpr u32 SPW = 8;
    static constexpr u32 M1 = 0x80000000u, M2 = 0x40000000u, M3 = 0x20000000u, M4 = 0x10000000u;

    SW_HD static u32 rep(int v) { return (u32)v; }
    SW_HD static u32 pack(u32 lo, u32) { return lo; }
    SW_HD static int half(u32 v, int) { return (int)v; }

    SW_HD static u32 min1(u32 x) { return x < 1u ? x : 1u; }
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
    SW_HD static u32 push(u32 acc, u32 w)
    {
#if SW_ON_DEVICE
        return __funnelshift_l(w, acc, 4);
#else
        return (acc << 4) | (w >> 28);
#endif
    }
};

template <class T>
SW_HD CellConsts make_consts(const Scores &s, int bias)
{
    CellConsts c;
    const int delta = s.match - s.mismatch;
    c.neg_delta = (u32)(-delta);
    c.go = T::rep(s.go);
    c.neg_ge = T::rep(-s.ge);
    c.match = T::rep(s.match);
    c.b1 = T::M1;
    c.b2 = T::M2;
    /* M3, M4 are single-bit-per-half masks, so (M / rep(1)) is the per-half bit value */
    c.b3 = T::M3 + T::rep(s.go - s.ge);
    c.b4 = T::M4 + T::rep(s.go - s.ge);
    c.one = 1u;
    c.mone = 0xffffffffu;
    c.neg_go = 0u - T::rep(s.go);       /* a 32-bit negation of the packed constant: added with a plain 32-bit add */
    c.bias = bias;
    c.s_match = s.match; c.s_go = s.go; c.s_ge = s.ge;
    return c;
}

/*
 * One cell.  In:  ra/bq packed ref/alt bytes, hdm = H[i-1][j-1] + match (+bias),
 * e = E[i][j], f = F[i][j].  Out: e = E[i+1][j], f = F[i][j+1], hm = H[i][j] + match,
 * acc gets the cell's nibble.
 *
 * Pipe balance (ncu, profiles/r01a_fill_full.txt: ALU pipe 95 % busy, FMA pipe 12 %
 * with every add on the ALU).  An SM sub-partition issues one warp instruction
 * per clock and owns one 16-lane ALU pipe (LOP3, SHF, IADD3, VIMNMX*, VIADDMNMX*)
 * and one 16-lane FMA pipe (IMAD); each takes a warp instruction every second
 * clock.  The DPX max / add-max and the bit merges exist only on the ALU pipe,
 * so everything that is a plain add or subtract is written as a multiply-add
 * with a runtime 1 / -1 (c.one, c.mone) and lands on the FMA pipe:
 *     ALU: xor, min, max3, 2 x add-max, 1 predicate (IADD3), 3 bit-selects, shift, bit-select  = 11
 *     FMA: d, 3 x (const - x), 3 predicates, h - go, h + match                               =  9
 * `alu3` moves the third predicate between the pipes; callers alternate it by
 * register index so that both pipes carry about 10.5 and 9.5 instructions per cell.
 */
template <class T>
SW_HD void sw_cell(const CellConsts &c, u32 ra, u32 bq, u32 hdm, u32 &e, u32 &f, u32 &hm, u32 &acc, bool alu3 = true)
{
    const u32 x = ra ^ bq;
    const u32 m = T::min1(x);                 /* 0 match, 1 mismatch                     */
    const u32 d = m * c.neg_delta + hdm;      /* H_diag + match - m * (match - mismatch) */
    const u32 h = T::max3(d, e, f);
    const u32 nh1 = h * c.mone + c.b1;
    const u32 t1 = d * c.one + nh1;           /* d - h + b1: d >= h, diagonal source     */
    const u32 ne2 = e * c.mone + c.b2;
    const u32 t2 = f * c.one + ne2;           /* f - e + b2: F >= E                      */
    u32 t3;                                   /* e - h + b3: E - ge >= H - go, E extends */
    if (alu3) t3 = e - h + c.b3;
    else { const u32 nh3 = h * c.mone + c.b3; t3 = e * c.one + nh3; }
    const u32 nh4 = h * c.mone + c.b4;
    const u32 t4 = f * c.one + nh4;           /* f - h + b4: F - ge >= H - go, F extends */
    const u32 hgo = h * c.one + c.neg_go;
    e = T::addmax(e, c.neg_ge, hgo);
    f = T::addmax(f, c.neg_ge, hgo);
    hm = h * c.one + c.match;
    u32 w = bitsel(T::M3, t3, t4);
    w = bitsel(T::M2, t2, w);
    w = bitsel(T::M1, t1, w);
    acc = T::push(acc, w);
}

}  // namespace mglsw
#endif
