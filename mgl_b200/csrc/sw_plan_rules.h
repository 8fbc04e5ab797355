/*
 * sw_plan_rules.h -- the per-pair planning rules, shared by the host planner (sw_plan.cpp), the device planner
 * (sw_plan_dev.h / .cu) and the emulator: which arithmetic a pair may use, which kernel shape it gets, and the key
 * that buckets a chunk.  SW_HD: the same code runs as CUDA device code and as host code.
 *
 * No counterpart in the reference, whose call is one pair with int32 cells throughout
 * (/root/reference/src/main/native/mgl_sw/sw.cpp:5-146); these rules exist so that a pair only runs in 16-bit halves
 * when every value the reference would compute for it (H, E, F of sw.cpp:55-93, plus the boundary ramps :29-49) fits.
 */
#ifndef MGL_B200_SW_PLAN_RULES_H
#define MGL_B200_SW_PLAN_RULES_H
#include "sw_common.h"
#include "sw_config.h"
#include "sw_layout.h"

namespace mglsw {

SW_HD long long rules_abs(long long v) { return v < 0 ? -v : v; }

/* margin for the garbage rows a lane computes before its row 1 and after its last row */
SW_HD long long ramp_margin(const Scores &s, int G)
{
    return (long long)(G + 4) * ((long long)s.match + rules_abs(s.mismatch) + s.go + s.ge);
}

/* conservative bounds of every value a cell of a (tl x ql) pair can hold (true scores, unbiased) */
SW_HD void value_bounds(const Scores &s, u32 tl, u32 ql, int G, long long &lo, long long &hi)
{
    const long long m = ramp_margin(s, G);
    lo = -(3ll * s.go + (long long)s.ge * ((long long)tl + ql + 2) + rules_abs(s.mismatch)) - m;
    hi = (long long)s.match * ((long long)(tl < ql ? tl : ql) + 2) + m;
}

/* can this pair be filled with unsigned 16-bit halves? */
SW_HD bool s16_bounds_hd(const Scores &s, u32 tl, u32 ql, long long &lo, long long &hi)
{
    value_bounds(s, tl, ql, MGLSW_G_SHORT, lo, hi);
    if (s.match - s.mismatch >= 32768) return false;
    if (tl > 65535u) return false;                         /* the fused last-column maximum packs (value, row) in 32 bits */
    return hi - lo < 65000;
}

/* lengths a pair may have at all, and the 32-bit headroom of its values */
SW_HD bool pair_ok(const Scores &s, u32 tl, u32 ql)
{
    if (tl == 0 || ql == 0 || tl > (1u << 24) || ql > (1u << 24)) return false;
    long long lo, hi;
    value_bounds(s, tl, ql, MGLSW_G_LONG, lo, hi);
    return !(-lo >= (1ll << 30) || hi >= (1ll << 30));
}

SW_HD bool s16_candidate(const Scores &s, bool allow_s16, u32 tl, u32 ql, long long &need_bias)
{
    long long lo, hi;
    if (allow_s16 && ql <= (u32)MGLSW_G_SHORT * MGLSW_KMAX && s16_bounds_hd(s, tl, ql, lo, hi)) { need_bias = -lo; return true; }
    need_bias = 0;
    return false;
}

SW_HD long long s16_bias_of(long long need_bias) { return need_bias < 49152 ? need_bias : 49152; }

/* with the chunk's bias fixed: does this candidate stay inside 16 bits? */
SW_HD bool s16_fits(const Scores &s, u32 tl, u32 ql, long long bias16)
{
    long long lo, hi;
    s16_bounds_hd(s, tl, ql, lo, hi);
    return -lo <= bias16 && hi <= 65535 - bias16;
}

/* kernel shape of a pair: lane group, strips, registers per lane */
SW_HD void pair_shape(u32 ql, u32 &G, u32 &nstrips, u32 &K)
{
    if (ql <= (u32)MGLSW_G_SHORT * MGLSW_KMAX) { G = MGLSW_G_SHORT; nstrips = 1; }
    else { G = MGLSW_G_LONG; nstrips = ceil_div_u32(ql, (u32)MGLSW_G_LONG * MGLSW_KMAX); }
    K = regs_per_lane(ql, nstrips * G);
}

/*
 * Bucket key, 30 bits, to be sorted DESCENDING (heavy work launches first):
 *   [29:22] launch key: arithmetic (1 = 32-bit), lane group (1 = 32 lanes), K (5 bits), strips pipelined over a block
 *   [21]    INDEL-type boundary
 *   [20:13] alt length (lane groups of 8: exact, <= 192) or strip count (whole-warp units: clipped, order only)
 *   [12:0]  ref length / 8, clipped: a warp's units share their row-block count, which moves in steps of 8 rows
 * Pairs that agree on bits [29:13] (MGLSW_KEY_PAIR_SHIFT) may share a u16x2 unit: same kernel, same boundary kind,
 * same alt length.  30 bits sort in five 6-bit radix passes (sw_plan_dev.cu).
 */
#define MGLSW_KEY_BITS 30
#define MGLSW_KEY_LAUNCH_SHIFT 22
#define MGLSW_KEY_PAIR_SHIFT 13
SW_HD u32 pair_key(int kind, u32 tl, u32 ql, int strategy)
{
    u32 G, nstrips, K;
    pair_shape(ql, G, nstrips, K);
    const u32 indel = (strategy & (OS_INDEL | OS_LEADING_INDEL)) ? 1u : 0u;
    const u32 qlq = G == (u32)MGLSW_G_SHORT ? ql : (nstrips < 255u ? nstrips : 255u);
    const u32 tlq = (tl >> 3) < 8191u ? (tl >> 3) : 8191u;
    u32 key = kind == KIND_S32 ? 1u : 0u;
    key = (key << 1) | (G == (u32)MGLSW_G_LONG ? 1u : 0u);
    key = (key << 5) | (K & 31u);
    key = (key << 1) | (nstrips >= (u32)MGLSW_PIPE_MIN_STRIPS ? 1u : 0u);
    key = (key << 1) | indel;
    key = (key << 8) | (qlq & 255u);
    key = (key << 13) | tlq;
    return key;
}
/* the form that is sorted ASCENDING */
SW_HD u32 pair_sort_key(int kind, u32 tl, u32 ql, int strategy) { return ((1u << MGLSW_KEY_BITS) - 1u) - pair_key(kind, tl, ql, strategy); }
SW_HD u32 sort_key_to_key(u32 skey) { return ((1u << MGLSW_KEY_BITS) - 1u) - skey; }
SW_HD int key_kind(u32 key) { return ((key >> 29) & 1) ? KIND_S32 : KIND_S16X2; }
SW_HD int key_G(u32 key) { return ((key >> 28) & 1) ? MGLSW_G_LONG : MGLSW_G_SHORT; }
SW_HD int key_K(u32 key) { return (int)((key >> 23) & 31); }
SW_HD int key_pipe(u32 key) { return (int)((key >> 22) & 1); }

/*
 * Which sorted positions open a unit.  32-bit arithmetic: every position.  u16x2: position q opens a unit if it is
 * the first of its run (pairs equal on the bits above MGLSW_KEY_PAIR_SHIFT) or q is even; an odd position that is not
 * a run head rides as the second pair of the unit q - 1 opened.  A rule of q and its left neighbour alone -- no scan
 * over runs -- at the price of at most two single-pair units per run instead of one.
 */
SW_HD bool opens_unit(int kind, u32 q, bool run_head) { return kind == KIND_S32 || run_head || (q & 1u) == 0u; }

/* the geometry fields of a unit that follow from its alt length and launch */
SW_HD void unit_geometry(UnitDesc &u, u32 G)
{
    u.nstrips = G == (u32)MGLSW_G_SHORT ? 1u : ceil_div_u32(u.ql, (u32)MGLSW_G_LONG * MGLSW_KMAX);
    const u32 slots = u.nstrips * G;
    u.col_base = u.ql / slots;
    u.col_rem = u.ql % slots;
    u.carry_rows = (u.tl[0] > u.tl[1] ? u.tl[0] : u.tl[1]) + 1;
}

/* words of the captured last row of an active unit */
SW_HD u64 unit_lastrow_words(const UnitDesc &u, int kind, u32 G, u32 K) { return (u64)(kind == KIND_S16X2 ? 2 : 1) * u.nstrips * G * K; }

}  // namespace mglsw
#endif
