/*
 * sw_layout.h -- where every DP cell lives: which lane register computes it
 * and which trace word / nibble records it.  Shared by the fill kernel (writer),
 * the traceback kernel (reader), the host planner (sizes) and the emulator.
 *
 * Geometry.  A lane group of G lanes owns one unit.  The alt (query) axis is cut
 * into nstrips * G lane slots; slot s holds col_base + 1 columns if
 * s < col_rem, else col_base, so that with K = ceil(ql / slots) registers per
 * lane every lane has K or K - 1 real columns (K - 1 == 0 means the lane only
 * forwards its neighbour's values).  Lane l of a strip works on ref row
 * i = t - l at step t (a one-step skew per lane), t = 0 .. tlmax + G - 1.
 *
 * Trace.  4 bits per cell (the reference stores gap
 * LENGTHS, sw.cpp:60-71; direction alone is not enough):
 *     bit3  H came from the diagonal        (sw.cpp:60)
 *     bit2  F >= E, i.e. horizontal wins    (sw.cpp:64)
 *     bit1  E was extended out of this cell (sw.cpp:73-82, tie => extend)
 *     bit0  F was extended out of this cell (sw.cpp:84-93, tie => extend)
 * A lane accumulates one nibble per step per register and the warp flushes
 * every SPW steps with 16-byte stores: region [strip][row block][k / 4][32 lanes]
 * [k % 4], i.e. each store instruction of the warp writes 512 contiguous bytes.
 */
#ifndef MGL_B200_SW_LAYOUT_H
#define MGL_B200_SW_LAYOUT_H
#include "sw_common.h"

namespace mglsw {

SW_HD u32 ceil_div_u32(u32 a, u32 b) { return (a + b - 1) / b; }

/* registers per lane for an alt length spread over `slots` lane slots */
SW_HD u32 regs_per_lane(u32 ql, u32 slots) { return ceil_div_u32(ql, slots); }

/* 16-byte chunks a lane flushes per row block */
SW_HD u32 chunks_per_lane(u32 K) { return (K + 3) / 4; }

/* trace words of one strip of one warp */
SW_HD u64 strip_trace_words(u32 rb_count, u32 K) { return (u64)rb_count * chunks_per_lane(K) * 128u; }

/* number of steps (and row-block count) for a warp iterating tlmax rows */
SW_HD u32 fill_steps(u32 tlmax, u32 G) { return tlmax + G; }
SW_HD u32 row_blocks(u32 tlmax, u32 G, u32 spw) { return ceil_div_u32(fill_steps(tlmax, G), spw); }

/* columns of lane slot `slot` and its first column (0-based) */
SW_HD u32 slot_cols(const UnitDesc &u, u32 slot) { return u.col_base + (slot < u.col_rem ? 1u : 0u); }
SW_HD u32 slot_first_col(const UnitDesc &u, u32 slot)
{
    return slot < u.col_rem ? slot * (u.col_base + 1) : u.col_rem * (u.col_base + 1) + (slot - u.col_rem) * u.col_base;
}

/* 0-based column -> (slot, register) */
SW_HD void col_to_slot(const UnitDesc &u, u32 col, u32 &slot, u32 &reg)
{
    const u32 wide = u.col_rem * (u.col_base + 1);
    if (col < wide) { slot = col / (u.col_base + 1); reg = col - slot * (u.col_base + 1); }
    else { const u32 c = col - wide; slot = u.col_rem + c / u.col_base; reg = c - (slot - u.col_rem) * u.col_base; }
}

/*
 * Word index (relative to the start of the trace buffer) and bit position of
 * the nibble of cell (row i >= 1, slot, reg).  G, K, SPW are the launch's.
 * `half` selects the pair inside a 16x2 word; ignored when SPW == 8.
 */
SW_HD u64 trace_word_index(const UnitDesc &u, u32 G, u32 K, u32 spw, u32 i, u32 slot, u32 reg)
{
    const u32 strip = slot / G, lane = slot - strip * G;
    const u32 t = i + lane;
    const u32 rb = t / spw;
    const u32 Q = chunks_per_lane(K);
    return u.trace_off + (u64)strip * strip_trace_words(u.rb_count, K)
         + ((u64)(rb * Q + (reg >> 2)) * 32u + u.group * G + lane) * 4u + (reg & 3u);
}
SW_HD u32 trace_nibble_shift(u32 G, u32 spw, u32 i, u32 slot, u32 half)
{
    const u32 lane = slot % G;
    const u32 s = (i + lane) % spw;
    /* 16x2: right-shift accumulation inside each half, step s ends at bits 4s of the half.
       32:   left-shift accumulation, step s ends at bits 4 * (7 - s).                      */
    return spw == 4 ? half * 16u + 4u * s : 4u * (7u - s);
}

}  // namespace mglsw
#endif
