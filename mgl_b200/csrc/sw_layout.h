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
 *                       that is shuffled to the next lane): 16 bytes per lane per step.
 *   row checkpoints     after every MGLSW_RB-th step, every lane: (H + match, E) of all K
 *                       registers.
 * A block is (lane slot, half of its registers, row block): at most ceil(K / 2) columns by
 * MGLSW_RB rows; the path of a 150-column alignment crosses ~35 blocks, i.e. ~4 % of the
 * matrix is computed twice.  About 1 B per cell at K = 20.
 *
 * Storage is READER-major: the traceback is one thread per pair chasing its own path, so what
 * one block needs must be contiguous (random 8-byte reads cost a 32-byte sector, mostly wasted:
 * the first layout made the traceback DRAM-bound at 8x read amplification).  Per warp, strip
 * and row block there is one region of 32 lanes x lane_stride16(K) 16-byte units; a lane's
 * slice holds its MGLSW_RB column checkpoints (128 B) followed by its row checkpoint in the
 * order H[0,KM) E[0,KM) H[KM,K) E[KM,K), each section padded to whole 16-byte chunks.  The
 * fill stages a region in shared memory and copies it out with 512-byte warp stores; the
 * odd lane stride keeps both the 16-byte shared-memory stores and the copy conflict-free.
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

/* 16-byte chunks per section (H or E of one half of the lane's registers) of a row checkpoint */
SW_HD u32 sec_chunks(u32 K) { return (regs_mid(K) + 3) / 4; }

/* 16-byte units of one lane's slice of a region: MGLSW_RB column checkpoints + 4 sections, made odd */
SW_HD u32 lane_stride16(u32 K) { const u32 s = MGLSW_RB + 4 * sec_chunks(K); return s | 1u; }

/* words of one region (one warp, one strip, one row block) */
SW_HD u32 region_words(u32 K) { return lane_stride16(K) * 128u; }

/* number of row blocks for a warp iterating tlmax rows; steps = row blocks * MGLSW_RB */
SW_HD u32 row_blocks(u32 tlmax, u32 G) { return ceil_div_u32(tlmax + G, MGLSW_RB); }

/* checkpoint words of one warp */
SW_HD u64 warp_ckpt_words(u32 nstrips, u32 rb_count, u32 K) { return (u64)nstrips * rb_count * region_words(K); }

/* word offset, inside a lane's slice, of the column checkpoint of step t and of row-checkpoint chunk q of section (sub, is_e) */
SW_HD u32 slice_col_word(u32 t) { return (t % MGLSW_RB) * 4u; }
SW_HD u32 slice_row_word(u32 K, u32 sub, u32 is_e, u32 q) { return MGLSW_RB * 4u + ((sub * 2u + is_e) * sec_chunks(K) + q) * 4u; }

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

/* word index (relative to the checkpoint buffer) of the slice of lane `lane` of u's group in region (strip, rb) */
SW_HD u64 slice_index(const UnitDesc &u, u32 G, u32 K, u32 strip, u32 rb, u32 lane)
{
    return u.ckpt_off + ((u64)strip * u.rb_count + rb) * region_words(K) + (u64)(u.group * G + lane) * lane_stride16(K) * 4u;
}

}  // namespace mglsw
#endif
