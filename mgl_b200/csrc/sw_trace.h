/*
 * sw_trace.h -- endpoint selection and traceback of one pair, from the 4-bit
 * trace the fill wrote.  One thread per pair on the GPU (sw_trace_kernel.cu);
 * SW_HD so the emulator under tests/emu can run the identical code.
 *
 * Restates, with the same tie-breaks (paths relative to
 * /root/reference/src/main/native/mgl_sw/):
 *   last-column argmax, ">=" with ascending row            sw.cpp:100-104  (fused into the fill, read back here)
 *   last-row scan, ">" or equal-and-nearer-the-diagonal    sw.cpp:116-127
 *   traceback start cell per strategy                      sw.cpp:155-170
 *   walk with k-cell gap jumps and run merging             sw.cpp:173-214
 *   overhang post-processing and offset                    sw.cpp:224-248
 *   zero-length elements dropped when printed              sw.cpp:251-252
 * The reference reads a gap length k from its int32 matrix; here k is recovered
 * by following the extend bits: lenF[i][j] = 1 + number of consecutive cells
 * (i, j-1), (i, j-2), ... whose F-extend bit is set (the bit of cell (i, c) says
 * F[i][c+1] extended F[i][c], sw.cpp:84-93); likewise lenE with rows.
 */
#ifndef MGL_B200_SW_TRACE_H
#define MGL_B200_SW_TRACE_H
#include "sw_cell.h"
#include "sw_fill.h"
#include "sw_layout.h"

namespace mglsw {

struct TraceCursor {
    const UnitDesc *u;
    const u32 *trace;
    u32 G, K, spw, half;
    u32 slot, reg;           /* position of column J                                  */

    SW_HD void seek(u32 J)   /* J >= 1 */
    {
        col_to_slot(*u, J - 1, slot, reg);
    }
    SW_HD void left()        /* J -> J - 1 (caller guarantees J - 1 >= 1) */
    {
        if (reg > 0) { reg--; return; }
        slot--;
        reg = slot_cols(*u, slot) - 1;
    }
    SW_HD u32 nibble(u32 I) const
    {
        const u64 w = trace_word_index(*u, G, K, spw, I, slot, reg);
#if SW_ON_DEVICE
        const u32 v = __ldcg(trace + w);
#else
        const u32 v = trace[w];
#endif
        return (v >> trace_nibble_shift(G, spw, I, slot, half)) & 15u;
    }
};

struct PairResult {
    int offset;
    u32 n_ops;
    int score;               /* biased score of the start cell (diagnostic)            */
};

/*
 * ops_end points one past the pair's scratch region; elements are written
 * downwards so that [ops_end - n_ops, ops_end) reads left to right.
 */
template <class T>
SW_HD PairResult traceback_pair(const UnitDesc &u, u32 half, u32 G, u32 K, const u32 *trace,
                                const u32 *colbest4, const u32 *lastrow, int strategy, u32 *ops_end)
{
    const int tl = (int)u.tl[half], ql = (int)u.ql;
    const u32 nslots = u.nstrips * G;

    /* last column: sw.cpp:100-104; the running maximum was kept by the fill (sw_fill.h), same "+ match + bias" as lastrow */
#if SW_ON_DEVICE
    const int mqe = (int)__ldcg(colbest4 + 2 * half), mqe_t = (int)__ldcg(colbest4 + 2 * half + 1);
#else
    const int mqe = (int)colbest4[2 * half], mqe_t = (int)colbest4[2 * half + 1];
#endif

    int I, J, run = 0, start_score = mqe;
    if (strategy == OS_INDEL) { I = tl; J = ql; }
    else if (strategy != OS_LEADING_INDEL) {
        /* last row: sw.cpp:116-127 */
        int best = mqe, best_t = mqe_t, best_q = ql, seg = 0;
        const u32 *row = lastrow + u.lastrow_off + (size_t)half * nslots * K;
        u32 slot = 0, reg = 0;
        for (int j = 1; j <= ql; j++) {
#if SW_ON_DEVICE
            const int s = T::half(__ldcg(row + (size_t)slot * K + reg), (int)half);
#else
            const int s = T::half(row[(size_t)slot * K + reg], (int)half);
#endif
            const int da = tl - j < 0 ? j - tl : tl - j;
            const int db = best_t - best_q < 0 ? best_q - best_t : best_t - best_q;
            if (s > best || (s == best && da < db)) { best = s; best_t = tl; best_q = j; seg = ql - j; }
            if (++reg >= slot_cols(u, slot)) { reg = 0; slot++; }
        }
        I = best_t; J = best_q; run = seg; start_score = best;
    } else { I = mqe_t; J = ql; }

    u32 *p = ops_end;
#define SW_EMIT(op, len) do { if ((len) > 0) *--p = ((u32)(len) << 4) | (u32)(op); } while (0)

    if (run > 0 && strategy == OS_SOFTCLIP) { SW_EMIT(OP_S, run); run = 0; }   /* sw.cpp:173-176 */

    TraceCursor cur;
    cur.u = &u; cur.trace = trace; cur.G = G; cur.K = K; cur.spw = T::SPW; cur.half = half;
    cur.seek((u32)J);

    u32 state = OP_M;
    do {                                                                        /* sw.cpp:182-214 */
        const u32 nb = cur.nibble((u32)I);
        u32 next; int steplen = 1;
        if (nb & 8u) {                       /* diagonal */
            next = OP_M; I--; J--;
            if (J >= 1) cur.left();
        } else if (nb & 4u) {                /* horizontal gap: insertion, length lenF[I][J] */
            next = OP_I;
            int c = J - 1;
            while (c >= 1) {
                cur.left();
                if (!(cur.nibble((u32)I) & 1u)) break;
                steplen++; c--;
            }
            /* cursor now sits on column c (>= 1) or is stale when c == 0 */
            J -= steplen;
            if (J >= 1 && J != c) cur.seek((u32)J);
        } else {                             /* vertical gap: deletion, length lenE[I][J] */
            next = OP_D;
            int r = I - 1;
            while (r >= 1 && (cur.nibble((u32)r) & 2u)) { steplen++; r--; }
            I -= steplen;
        }
        if (next == state) run += steplen;
        else { SW_EMIT(state, run); run = steplen; state = next; }
    } while (I > 0 && J > 0);

    PairResult res;
    if (strategy == OS_SOFTCLIP) {                                              /* sw.cpp:225-229 */
        SW_EMIT(state, run);
        SW_EMIT(OP_S, J);
        res.offset = I;
    } else if (strategy == OS_IGNORE) {                                         /* sw.cpp:230-233 */
        SW_EMIT(state, run + J);
        res.offset = I - J;
    } else {                                                                    /* sw.cpp:234-248 */
        SW_EMIT(state, run);
        if (I > 0) SW_EMIT(OP_D, I);
        else if (J > 0) SW_EMIT(OP_I, J);
        res.offset = 0;
    }
#undef SW_EMIT
    res.n_ops = (u32)(ops_end - p);
    res.score = start_score;
    return res;
}

}  // namespace mglsw
#endif
