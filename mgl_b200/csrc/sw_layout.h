/*
 * sw_layout.h -- where every DP cell lives: which lane register computes it,
 * and where the checkpoints that let the traceback recompute it are stored.
 * Shared by the fill kernel (writer), the traceback kernel (reader), the host
 * planner (sizes) and the emulator.
 *
 * Geometry.  A lane group of G lanes owns one unit.  The alt (query) axis is cut
 * into nstrips * G lane slots; slot s holds col_base + 1 columns if
 * s < col_rem, else col_base, so that with K = ceil(ql / slots) registers per
 * lane every lane has K or K - 1 real columns (K - 1 == 0 means the lane only
 * forwards its neighbour's values).  Lane l of a strip works on ref row
 * i = t - l at step t (a one-step skew per lane), t = 0 .. steps - 1.
 *
 * Checkpoints instead of a trace.  The reference keeps a full int32 backtrack
 * matrix (sw.cpp:264-265, sw_avx.cpp:33-34) that its traceback reads along one
 * path of O(tl + ql) cells.  Here the fill stores no per-cell record at all:
 * it parks the DP state on a grid, and the traceback recomputes, with the
 * reference's predicates (sw.cpp:60-93), only the grid blocks its path crosses.
 *   column checkpoints  every step, every lane: (H + match, F) leaving the lane's register
 *                       KM - 1 ("mid") and leaving its last real register ("end", the pair
 *                       that is shuffled to the next lane): one 16-byte store per lane per
 *                       step, [strip][t][32 lanes][4 words] -- 512 contiguous bytes per warp.
 *   row checkpoints     after every MGLSW_RB-th step, every lane: (H + match, E) of all K
 *                       registers, [strip][row block][chunk][32 lanes][4 words].
 * A block is (lane slot, half of its registers, row block): at most ceil(K / 2) columns by
 * MGLSW_RB rows.  About 0.65 B per cell at K = 20, and the path of a 150-column alignment
 * crosses ~25 blocks, i.e. ~4 % of the matrix is computed twice.
 */
#ifndef MGL_B200_SW_LAYOUT_H
#define MGL_B200_SW_LAYOUT_H
#include "sw_common.h"

#define MGLSW_RB 8u              /* steps per row block (row checkpoint period) */

namespace mglsw {

SW_HD u32 ceil_div_u32(u32 a, u32 b) { return (a + b - 1) / b; }

/* registers per lane for an alt length spread over `slots` lane slots */
SW_HD u32 regs_per_lane(u32 ql, u32 slots) { return ceil_div_u32(ql, slots); }

/* registers in the first half of a lane: the mid column checkpoint sits after register KM - 1 */
SW_HD u32 regs_mid(u32 K) { return (K + 1) / 2; }

/*
 * One lane's row checkpoint: four sections of sec_chunks(K) 16-byte chunks each --
 * H + match of registers [0, KM), H + match of [KM, K), E of [0, KM), E of [KM, K) -- so that
 * the registers of either half of the lane start on a chunk boundary and the traceback
 * reads them with 16-byte loads.
 */
SW_HD u32 sec_chunks(u32 K) { return (regs_mid(K) + 3) / 4; }
SW_HD u32 row_chunks(u32 K) { return 4 * sec_chunks(K); }

/* number of steps (a multiple of the row block) for a warp iterating tlmax rows */
SW_HD u32 row_blocks(u32 tlmax, u32 G) { return ceil_div_u32(tlmax + G, MGLSW_RB); }

/* checkpoint words of one warp: column checkpoints of every strip, then row checkpoints */
SW_HD u64 warp_col_words(u32 nstrips, u32 rb_count) { return (u64)nstrips * rb_count * MGLSW_RB * 128u; }
SW_HD u64 warp_row_words(u32 nstrips, u32 rb_count, u32 K) { return (u64)nstrips * rb_count * row_chunks(K) * 128u; }

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
 * Word index (relative to the checkpoint buffer) of the column checkpoint lane `lane` of group
 * u.group stored at step t of strip `strip`: 4 words (h_mid, f_mid, h_end, f_end).
 */
SW_HD u64 col_ckpt_index(const UnitDesc &u, u32 G, u32 strip, u32 t, u32 lane)
{
    return u.col_off + (((u64)strip * u.rb_count * MGLSW_RB + t) * 32u + u.group * G + lane) * 4u;
}

/* Word index of chunk c (0 .. row_chunks(K) - 1) of the row checkpoint stored after the last step of row block rb. */
SW_HD u64 row_ckpt_index(const UnitDesc &u, u32 G, u32 K, u32 strip, u32 rb, u32 lane, u32 c)
{
    return u.row_off + ((((u64)strip * u.rb_count + rb) * row_chunks(K) + c) * 32u + u.group * G + lane) * 4u;
}

}  // namespace mglsw
#endif
