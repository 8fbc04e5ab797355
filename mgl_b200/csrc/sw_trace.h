/*
 * sw_trace.h -- endpoint selection and traceback of one pair, recomputing the
 * cells on the path from the fill's checkpoints.  One thread per pair on the
 * GPU (sw_trace_kernel.cu); SW_HD so the emulator under tests/emu can run the
 * identical code.
 *
 * Restates, with the same tie-breaks (paths relative to
 * /root/reference/src/main/native/mgl_sw/):
 *   cell predicates (source of H, open vs extend)          sw.cpp:60-93   (Replay::build, verbatim comparisons)
 *   last-column argmax, ">=" with ascending row            sw.cpp:100-104 (fused into the fill, read back here)
 *   last-row scan, ">" or equal-and-nearer-the-diagonal    sw.cpp:116-127
 *   traceback start cell per strategy                      sw.cpp:155-170
 *   walk with k-cell gap jumps and run merging             sw.cpp:173-214
 *   overhang post-processing and offset                    sw.cpp:224-248
 *   zero-length elements dropped when printed              sw.cpp:251-252
 * The reference reads a gap length k from its int32 matrix; here a cell's record
 * is four bits (diag-source, F-source, E-extended, F-extended) and k is recovered
 * by following the extend bits: lenF[i][j] = 1 + number of consecutive cells
 * (i, j-1), (i, j-2), ... whose F-extend bit is set (the bit of cell (i, c) says
 * F[i][c+1] extended F[i][c], sw.cpp:84-93); likewise lenE with rows.
 */
#ifndef MGL_B200_SW_TRACE_H
#define MGL_B200_SW_TRACE_H
#include "sw_cell.h"
#include "sw_fill.h"
#include "sw_layout.h"
#include "sw_config.h"

#define MGLSW_KH ((MGLSW_KMAX + 1) / 2)      /* columns of a block: at most half a lane's registers */
#ifndef MGLSW_HOP_WIDE_MIN_STRIPS
#define MGLSW_HOP_WIDE_MIN_STRIPS MGLSW_PIPE_MIN_STRIPS   /* pairs of at least this many column strips (= the pipelined fill launches) hop MGLSW_HOP_BLOCKS blocks at a time */
#endif
#define MGLSW_GAP_HOP_BLOCKS 8                 /* row blocks one deletion hop tries at once (Replay::ehop) */
#define MGLSW_GAP_HOP_LANES 4                  /* lanes one insertion hop tries at once (Replay::fhop) */
#ifndef MGLSW_HOP_SHORT
#define MGLSW_HOP_SHORT 2                     /* ... pairs of lane groups of 8 (the round kernels); other whole-warp units: 1 */
#endif
#ifndef MGLSW_HOP_BLOCKS
#define MGLSW_HOP_BLOCKS 4                    /* blocks one diagonal hop tries to cross per memory round trip */
#endif
/* 64-bit words of one pair's replay buffer: MGLSW_RB rows of 4-bit records, then the H values on the block's left
   boundary column (MGLSW_RB + 1 rows) and top boundary row (MGLSW_KH columns) as 32-bit integers */
#define MGLSW_TRACE_BUF (MGLSW_RB + (MGLSW_RB + 1 + MGLSW_KH + 1) / 2)

namespace mglsw {

#if SW_ON_DEVICE
#define SW_LDCG(p) __ldcg(p)
#else
#define SW_LDCG(p) (*(p))
#endif

struct W4 { u32 x, y, z, w; };
struct W2 { u32 x, y; };
SW_HD W4 load_w4(const u32 *p)
{
#if SW_ON_DEVICE
    const uint4 v = __ldcg(reinterpret_cast<const uint4 *>(p));
    W4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r;
#else
    W4 r; r.x = p[0]; r.y = p[1]; r.z = p[2]; r.w = p[3]; return r;
#endif
}
SW_HD W2 load_w2(const u32 *p)
{
#if SW_ON_DEVICE
    const uint2 v = __ldcg(reinterpret_cast<const uint2 *>(p));
    W2 r; r.x = v.x; r.y = v.y; return r;
#else
    W2 r; r.x = p[0]; r.y = p[1]; return r;
#endif
}

/*
 * The block of DP cells around the traceback's current position, recomputed from the
 * checkpoints that bound it: rows ia..ib of columns ja..jb (1-based, inclusive), one
 * 4-bit record per cell.  A block never extends below / right of the cell that asked for
 * it: the traceback only moves up and left.
 */
/* n / d for n * d < 2^32 as one multiply: m = ceil(2^32 / d) (d >= 2; for d <= 1 the quotient is n itself and m is unused) */
SW_HD u32 magic_u32(u32 d) { return d > 1u ? 0xffffffffu / d + 1u : 0u; }
SW_HD u32 div_magic(u32 n, u32 d, u32 m)
{
#if SW_ON_DEVICE
    return d > 1u ? __umulhi(n, m) : n;
#else
    return d > 1u ? (u32)(((u64)n * m) >> 32) : n;
#endif
}

struct Replay {
    /* the unit's geometry, copied out of its descriptor once */

    /* 0-based column -> (lane slot, register): sw_layout.h col_to_slot, the unit's divisors as multiplications */
    SW_HD void slot_of(u32 col, u32 &slot, u32 &reg) const
    {
        const u32 wide = cr * (cb + 1);
        if (col < wide) { slot = div_magic(col, cb + 1u, m_cb1); reg = col - slot * (cb + 1); }
        else { const u32 c2 = col - wide; slot = cr + div_magic(c2, cb, m_cb); reg = c2 - (slot - cr) * cb; }
    }

    const u32 *base;                     /* checkpoint regions of the warp, at this unit's lane group */
    const uint8_t *ref, *alt;            /* this pair's sequences */
    u32 G, K, KM, SC, LSW, RW, half, rbc; /* SC = sec_chunks(K), LSW / RW = words per lane slice / per region, rbc = row blocks per strip */
    u32 cb, cr;                          /* col_base, col_rem */
    u32 m_cb1, m_cb, gsh;                /* column -> lane slot without a division: ceil(2^32 / (cb + 1)), ceil(2^32 / cb), log2(G) */
    u32 swords;                          /* words of one strip's regions = rbc * RW */
    int match, mismatch, go, ge, bias, indel;
    u32 sh, msk;                         /* this pair's half of a checkpoint word: bits [sh, sh + 16) of a u16x2 unit, the whole word of an s32 unit */
    int ia, ib, ja, jb;
    u64 *nib;                            /* row i - ia at nib[(i - ia) * nstride]: nibble of column j at bits 4 (j - ja); */
    u32 nstride;                         /* shared memory on the device (one column per thread), a plain array in the emulator */
    u32 builds, cells;                   /* statistics */

    SW_HD void init(const UnitDesc &u, u32 G_, u32 K_, u32 half_, bool s32, const u32 *ckpt, const uint8_t *arena, const CellConsts &c,
                    u64 *nib_, u32 nstride_)
    {
        nib = nib_; nstride = nstride_;
        sh = s32 ? 0u : 16u * half_; msk = s32 ? 0xffffffffu : 0xffffu;
        G = G_; K = K_; KM = regs_mid(K_); SC = sec_chunks(K_); half = half_; rbc = u.rb_count;
        cb = u.col_base; cr = u.col_rem;
        m_cb1 = magic_u32(cb + 1u); m_cb = magic_u32(cb); gsh = G_ == 32u ? 5u : 3u;
        LSW = lane_stride16(K_) * 4u; RW = region_words(K_);
        swords = rbc * RW;
        base = ckpt + u.ckpt_off + (size_t)u.group * G_ * LSW;
        ref = arena + u.ref_off[half_]; alt = arena + u.alt_off[half_];
        match = c.s_match; mismatch = c.s_mismatch; go = c.s_go; ge = c.s_ge; bias = c.bias;
        indel = (u.flags & UNIT_INDEL) ? 1 : 0;
        ia = 1; ib = 0; ja = 1; jb = 0; builds = 0; cells = 0;
    }

    SW_HD int bnd(int x) const { return boundary_score(x, indel, go, ge); }
    SW_HD int hval(u32 w) const { return (int)((w >> sh) & msk) - bias - match; }   /* stored H + match + bias -> H */
    SW_HD int gval(u32 w) const { return (int)((w >> sh) & msk) - bias; }           /* stored E / F + bias      */

    /* H on the boundary of the block built last: entry r = H[ia - 1 + r][ja - 1], entry MGLSW_RB + 1 + kk = H[ia - 1][ja + kk] */
    SW_HD void put_h(u32 idx, int v) { reinterpret_cast<u32 *>(nib + (size_t)(MGLSW_RB + idx / 2u) * nstride)[idx & 1u] = (u32)v; }
    SW_HD int get_h(u32 idx) const { return (int)reinterpret_cast<const u32 *>(nib + (size_t)(MGLSW_RB + idx / 2u) * nstride)[idx & 1u]; }
    /* after the walk left the block through its top or left edge: H of the cell it stands on, if that is a boundary cell */
    SW_HD bool boundary_h(int i, int j, int &h) const
    {
        if (ib < ia) return false;
        if (j == ja - 1 && i >= ia - 1 && i <= ib) { h = get_h((u32)(i - (ia - 1))); return true; }
        if (i == ia - 1 && j >= ja && j <= jb) { h = get_h(MGLSW_RB + 1u + (u32)(j - ja)); return true; }
        return false;
    }

    /*
     * Diagonal hop.  H[i][j] >= H[i-1][j-1] + s(i, j) holds at every cell (H is the maximum of three
     * candidates, sw.cpp:55-71), with equality exactly when the reference's traceback takes the
     * diagonal there (its test d >= E && d >= F, sw.cpp:60, is d == H).  Summed along L diagonal
     * cells: H[I][J] - H[I-L][J-L] >= sum of the L substitution scores, with equality iff EVERY one
     * of the L cells is a diagonal step.  So when H is known at (I, J) and at the next checkpointed
     * cell up the diagonal -- the column left of J's block (every row of it is stored by the fill), or
     * row 0 / column 0 -- one 8-byte load and L byte compares decide L traceback steps, and no block
     * is recomputed.
     * hop<NB> tries NB blocks in a row: the checkpointed cells up the diagonal follow from the unit's
     * geometry alone, so their loads and the sequence bytes of the whole stretch are in flight together --
     * one memory round trip where block-by-block hops take one each (a clean stretch of a 150-column
     * alignment is ~12 blocks).  Returns the steps taken over the leading clean blocks (0: the very first
     * block holds a gap, or a tie that goes another way -- the caller then replays the block) and H at the
     * cell reached; `blocked` says that the block after those is known not to be clean, so the caller
     * replays it without asking again.  NB = MGLSW_HOP_BLOCKS costs three times the instructions of NB = 1
     * whether it gets anywhere or not: it pays on long paths, where the chain of round trips is what the
     * pair waits for (sw_traceback_kernel picks per warp).
     */
    template <int NB>
    SW_HD int hop(int I, int J, int hcur, int &hnext, bool &blocked) const
    {
        int Lk[NB], hv[NB];
        const u32 *ap[NB];
        int Ik = I, Jk = J, S = 0;
#pragma unroll
        for (int k = 0; k < NB; k++) {
            Lk[k] = 0; hv[k] = 0; ap[k] = nullptr;
            if (Ik > 0 && Jk > 0) {
                u32 slot, reg;
                slot_of((u32)Jk - 1u, slot, reg);
                const u32 strip = slot >> gsh, lane = slot - (strip << gsh);
                const bool sub = reg >= KM;
                const u32 k_lo = sub ? KM : 0u;
                const int jc = (int)(slot < cr ? slot * (cb + 1) : cr * (cb + 1) + (slot - cr) * cb) + (int)k_lo;   /* = ja - 1 of Jk's block */
                int L = Jk - jc;
                if (L > Ik) L = Ik;
                const int it = Ik - L, jt = Jk - L;
                if (it == 0) hv[k] = bnd(jt);                               /* row 0, sw.cpp:29-35 */
                else if (jt == 0) hv[k] = bnd(it);                          /* column 0, sw.cpp:47-49 */
                else {
                    u32 cs = strip, cl = lane, word = 0;
                    if (!sub) { word = 2; if (lane > 0) cl = lane - 1; else { cl = G - 1; cs = strip - 1; } }
                    const u32 tt = (u32)it + cl;                            /* the step at which lane cl worked on row it */
                    ap[k] = base + (size_t)cs * swords + (size_t)(cl * LSW + word + (tt % MGLSW_RB) * 4u) + (size_t)(tt / MGLSW_RB) * RW;
                }
                Lk[k] = L; S += L; Ik = it; Jk = jt;
            }
        }
        u32 raw[NB];
#pragma unroll
        for (int k = 0; k < NB; k++) raw[k] = ap[k] ? load_w2(ap[k]).x : 0u;
        /* mismatches of the whole stretch, ref[I-S .. I) against alt[J-S .. J): bit u of mm = bytes u of the window of
           NB * MGLSW_KH bytes that ENDS at (I, J) differ (the arena is padded in front by more than that) */
        constexpr int WIN = NB * MGLSW_KH;
        static_assert(NB >= 1 && WIN <= 48 && WIN % 4 == 0 && WIN + 4 <= (int)MGLSW_ARENA_FRONT_PAD, "mismatch window: one 64-bit mask, inside the arena's front padding");
        u64 mm = 0;
#if SW_ON_DEVICE
        {
            const uint8_t *pr = ref + (I - WIN), *pa = alt + (J - WIN);
            const u32 *wr = reinterpret_cast<const u32 *>(reinterpret_cast<size_t>(pr) & ~(size_t)3);
            const u32 *wa = reinterpret_cast<const u32 *>(reinterpret_cast<size_t>(pa) & ~(size_t)3);
            const u32 sr = ((u32)reinterpret_cast<size_t>(pr) & 3u) * 8u, sa = ((u32)reinterpret_cast<size_t>(pa) & 3u) * 8u;
            u32 rw[WIN / 4 + 1], aw[WIN / 4 + 1];
#pragma unroll
            for (int w = 0; w <= WIN / 4; w++) { rw[w] = __ldg(wr + w); aw[w] = __ldg(wa + w); }
#pragma unroll
            for (int w = 0; w < WIN / 4; w++) {
                const u32 x = __vcmpne4(__funnelshift_r(rw[w], rw[w + 1], sr), __funnelshift_r(aw[w], aw[w + 1], sa));
                mm |= (u64)(((x & 0x01010101u) * 0x10204080u) >> 28) << (4 * w);     /* one bit per byte */
            }
        }
#else
        for (int t = 0; t < S; t++) if (ref[I - 1 - t] != alt[J - 1 - t]) mm |= 1ull << (WIN - 1 - t);
#endif
        int h = hcur, adv = 0, pos = WIN;
        blocked = false;
#pragma unroll
        for (int k = 0; k < NB; k++) {
            const int L = Lk[k];
            if (L == 0 || blocked) continue;
            pos -= L;
            const int ht = ap[k] ? hval(raw[k]) : hv[k];
#if SW_ON_DEVICE
            const int nm = __popcll((mm >> pos) & ((1ull << L) - 1ull));
#else
            int nm = 0;
            for (int t = 0; t < L; t++) nm += (int)((mm >> (pos + t)) & 1ull);
#endif
            if (h - ht == L * match - nm * (match - mismatch)) { h = ht; adv += L; }
            else blocked = true;
        }
        hnext = h;
        return adv;
    }

    /*
     * Gap hops.  Inside a gap the walk only asks whether the gap was EXTENDED through each cell (sw.cpp:73-93), and the
     * recurrence gives the same kind of inequality as for the diagonal: E[i+1][j] = max(E[i][j] - ge, H[i][j] - go)
     * >= E[i][j] - ge, with equality exactly when the reference extends there (a tie extends, :75,:86).  Summed down a
     * block's rows: E leaving the block == E entering it - rows * ge iff EVERY cell of the column extended -- and both
     * values are row checkpoints (E of every register after each row block).  The same across a block's columns with
     * F and the column checkpoints (F entering each half-lane block, every row).  Two 4-byte loads decide a whole
     * block of a long deletion / insertion (the overhang gaps of INDEL alignments are hundreds of cells: config 3 rebuilt
     * 17 blocks per pair, most of them along one gap), no block is recomputed.
     * ehop: the walk stands at (r, J) inside a deletion, r the LAST row of a row block: the rows (a whole number of row
     * blocks, up to MGLSW_GAP_HOP_BLOCKS of them, all of rows >= 2) the deletion runs through at column J, or 0.
     */
    SW_HD int ehop(int r, int J) const
    {
        u32 slot, reg;
        slot_of((u32)J - 1u, slot, reg);
        const u32 strip = slot >> gsh, lane = slot - (strip << gsh);
        const u32 t = (u32)r + lane;
        if ((t % MGLSW_RB) != MGLSW_RB - 1u) return 0;
        const u32 rb = t / MGLSW_RB;
        /* the topmost block a hop may cross starts at row >= 2 (its checkpoint above exists, and the walk is left at r >= 1) */
        const u32 rb_min = (lane + 2u + MGLSW_RB - 1u) / MGLSW_RB;
        if (rb < rb_min || rb < 1u) return 0;
        const bool sub = reg >= KM;
        const u32 kk = reg - (sub ? KM : 0u);
        const u32 *ep = base + ((size_t)strip * rbc + rb) * RW + (size_t)(lane * LSW + MGLSW_RB * 4u + ((sub ? 2u * SC : 0u) + SC) * 4u + kk);
        /* the sums telescope: m blocks at once need the same two values, m blocks apart; the far one first, one block as the fallback */
        u32 m = rb - rb_min + 1u;
        if (m > (u32)MGLSW_GAP_HOP_BLOCKS) m = MGLSW_GAP_HOP_BLOCKS;
#if SW_ON_DEVICE
        const u32 bot = __ldcg(ep), far = __ldcg(ep - (size_t)m * RW), top = __ldcg(ep - RW);
#else
        const u32 bot = ep[0], far = *(ep - (size_t)m * RW), top = *(ep - RW);
#endif
        const int vb = (int)((bot >> sh) & msk);
        if ((int)((far >> sh) & msk) - vb == (int)(m * MGLSW_RB) * ge) return (int)(m * MGLSW_RB);
        return (int)((top >> sh) & msk) - vb == (int)MGLSW_RB * ge ? (int)MGLSW_RB : 0;
    }
    /* fhop: the walk stands at (I, cc) inside an insertion, cc the LAST column of a half-lane block that starts at column
       >= 2: the block's width if the insertion runs through all of its columns in row I, else 0 */
    SW_HD int fhop(int I, int cc) const
    {
        u32 slot, reg;
        slot_of((u32)cc - 1u, slot, reg);
        const u32 strip = slot >> gsh, lane = slot - (strip << gsh);
        const u32 n = cb + (slot < cr ? 1u : 0u);
        const bool sub = reg >= KM;
        if (n < KM || reg != (sub ? n - 1u : KM - 1u)) return 0;                /* not the block's last column (or a lane too narrow to have a mid checkpoint) */
        const u32 k_lo = sub ? KM : 0u;
        const int ja_ = (int)(slot < cr ? slot * (cb + 1) : cr * (cb + 1) + (slot - cr) * cb) + 1 + (int)k_lo;
        if (ja_ < 2) return 0;
        /* F entering the block: the end checkpoint of the lane to the left (first half) or this lane's mid checkpoint (second half);
           F leaving it: this lane's mid checkpoint (first half) or its end checkpoint (second half) */
        u32 cs = strip, cl = lane, wl = 1u;
        if (!sub) { wl = 3u; if (lane > 0) cl = lane - 1; else { cl = G - 1; cs = strip - 1; } }
        const u32 tl_ = (u32)I + cl, tr_ = (u32)I + lane;
        const u32 *lp = base + (size_t)cs * swords + (size_t)(cl * LSW + (tl_ % MGLSW_RB) * 4u + wl) + (size_t)(tl_ / MGLSW_RB) * RW;
        const u32 *rp_ = base + (size_t)strip * swords + (size_t)(lane * LSW + (tr_ % MGLSW_RB) * 4u + (sub ? 3u : 1u)) + (size_t)(tr_ / MGLSW_RB) * RW;
#if SW_ON_DEVICE
        const u32 fl = __ldcg(lp), fr = __ldcg(rp_);
#else
        const u32 fl = *lp, fr = *rp_;
#endif
        const int vr = (int)((fr >> sh) & msk);
        /* the sums telescope here too: first the stretch back to the first column of the lane MGLSW_GAP_HOP_LANES to the left (F entering
           it is the end checkpoint of the lane before that one), then this block alone */
        const u32 nslot_row = slot - strip * G;                                  /* lanes to the left inside this strip */
        if (nslot_row >= 2u) {
            const u32 back = nslot_row - 1u < (u32)MGLSW_GAP_HOP_LANES ? nslot_row - 1u : (u32)MGLSW_GAP_HOP_LANES;
            const u32 ts = slot - back, tcl = ts - 1u - strip * G;               /* target lane slot, and the lane whose end checkpoint holds F entering it */
            const int tja = (int)(ts < cr ? ts * (cb + 1) : cr * (cb + 1) + (ts - cr) * cb) + 1;
            const u32 tt_ = (u32)I + tcl;
            const u32 *fp = base + (size_t)strip * swords + (size_t)(tcl * LSW + (tt_ % MGLSW_RB) * 4u + 3u) + (size_t)(tt_ / MGLSW_RB) * RW;
#if SW_ON_DEVICE
            const u32 ff = __ldcg(fp);
#else
            const u32 ff = *fp;
#endif
            const int wide = cc - tja + 1;
            if (tja >= 2 && (int)((ff >> sh) & msk) - vr == wide * ge) return wide;
        }
        const int width = (int)(reg - k_lo) + 1;
        return (int)((fl >> sh) & msk) - vr == width * ge ? width : 0;
    }

    SW_HD void build(int I, int J)
    {
        /* (slot, register) of column J: sw_layout.h col_to_slot */
        u32 slot, reg;
        slot_of((u32)J - 1u, slot, reg);
        const u32 strip = slot >> gsh, lane = slot - (strip << gsh);
        const bool sub = reg >= KM;
        const u32 k_lo = sub ? KM : 0u;
        const int ja_ = (int)(slot < cr ? slot * (cb + 1) : cr * (cb + 1) + (slot - cr) * cb) + 1 + (int)k_lo;
        const int cnt = (int)(reg - k_lo) + 1;
        const u32 rb = ((u32)I + lane) / MGLSW_RB;
        int ia_ = (int)(rb * MGLSW_RB) - (int)lane;
        if (ia_ < 1) ia_ = 1;
        const int rows = I - ia_ + 1;                              /* 1 .. MGLSW_RB */

        /* where the left boundary (H[i][ja_ - 1] + match, F[i][ja_]) of row i lives: a column checkpoint of step i + cl */
        u32 cs = strip, cl = lane, word = 0;
        if (!sub) { word = 2; if (lane > 0) cl = lane - 1; else { cl = G - 1; cs = strip - 1; } }
        const bool edge = ja_ == 1;                                /* matrix column 0: no checkpoint, sw.cpp:47-49 and :22-26,38 */
        /* slice of lane cl in region (cs, step / MGLSW_RB), entry step % MGLSW_RB, words word..word+1 */
        const u32 *lp = base + (size_t)cs * swords + (size_t)(cl * LSW + word);

        /*
         * Every load of the block first, unconditional (indices clamped into the block, dead
         * ones discarded afterwards) so that they are all in flight together: one memory round
         * trip per block instead of one per row.
         */
        W2 lraw[MGLSW_RB + 1];
        u32 rraw[MGLSW_RB + 1];
        const u32 *lp_safe = edge ? base : lp;                     /* column 0 has no checkpoint: read anything, use nothing */
#pragma unroll
        for (int r = 0; r <= (int)MGLSW_RB; r++) {                 /* entry r is row ia_ - 1 + r; entry 0 only yields the corner */
            int ic = ia_ - 1 + r;
            ic = ic < 1 ? 1 : (ic > I ? I : ic);
            const u32 tt = edge ? 0u : (u32)ic + cl;               /* the step at which lane cl worked on row ic */
            lraw[r] = load_w2(lp_safe + (size_t)(tt / MGLSW_RB) * RW + (tt % MGLSW_RB) * 4u);
            rraw[r] = SW_LDG(ref + (ic - 1));
        }
        W4 hraw[MGLSW_KH / 4], eraw[MGLSW_KH / 4];
        {
            const u32 rbm = rb >= 1 ? rb - 1 : 0u;
            /* this lane's row checkpoint of the previous row block: H section of this half, E section right behind it */
            const u32 *rp0 = base + ((size_t)strip * rbc + rbm) * RW + (size_t)lane * LSW + MGLSW_RB * 4u + (sub ? 2u * SC : 0u) * 4u;
#pragma unroll
            for (int q = 0; q < MGLSW_KH / 4; q++) {
                const u32 qc = (u32)q < SC ? (u32)q : SC - 1;      /* stay inside the section */
                hraw[q] = load_w4(rp0 + qc * 4u);
                eraw[q] = load_w4(rp0 + (qc + SC) * 4u);
            }
        }
        u32 ab[MGLSW_KH];
#pragma unroll
        for (int kk = 0; kk < MGLSW_KH; kk++) {
            const int jc = ja_ + (kk < cnt ? kk : cnt - 1);
            ab[kk] = SW_LDG(alt + (jc - 1));
        }

        int hlv[MGLSW_RB + 1], fv[MGLSW_RB + 1];
#pragma unroll
        for (int r = 0; r <= (int)MGLSW_RB; r++) {
            const int i = ia_ - 1 + r;
            int hl, fin;
            if (i < 1) { hl = bnd(ja_ - 1); fin = 0; }            /* row 0: H[0][ja_ - 1] */
            else if (edge) { hl = bnd(i); fin = hl - go; }
            else { hl = hval(lraw[r].x); fin = gval(lraw[r].y); }
            hlv[r] = hl; fv[r] = fin;
        }
        int Hp[MGLSW_KH], E[MGLSW_KH];
#pragma unroll
        for (int q = 0; q < MGLSW_KH / 4; q++) {
            Hp[4 * q] = hval(hraw[q].x); Hp[4 * q + 1] = hval(hraw[q].y); Hp[4 * q + 2] = hval(hraw[q].z); Hp[4 * q + 3] = hval(hraw[q].w);
            E[4 * q] = gval(eraw[q].x); E[4 * q + 1] = gval(eraw[q].y); E[4 * q + 2] = gval(eraw[q].z); E[4 * q + 3] = gval(eraw[q].w);
        }
#pragma unroll
        for (int kk = 0; kk < MGLSW_KH; kk++) {
            if (ia_ == 1) { const int b = bnd(ja_ + kk); Hp[kk] = b; E[kk] = b - go; }  /* rows 0 / 1 seeds: sw.cpp:10-18,29-35 */
        }

        /* keep H of the block's left boundary column and top boundary row: the walk needs it when it leaves the block */
#pragma unroll
        for (int r = 0; r <= (int)MGLSW_RB; r++) put_h((u32)r, hlv[r]);
#pragma unroll
        for (int kk = 0; kk < MGLSW_KH; kk++) put_h(MGLSW_RB + 1u + (u32)kk, Hp[kk]);

        /* unrolled over the (at most MGLSW_RB) rows so that the boundary values stay in registers */
#pragma unroll
        for (int r = 1; r <= (int)MGLSW_RB; r++) {
            if (r > rows) break;
            int diag = hlv[r - 1];                                 /* H[i - 1][ja_ - 1] */
            int f = fv[r];
            const u32 rbyte = rraw[r];
            u32 lo = 0, hi = 0;
#pragma unroll
            for (int kk = 0; kk < MGLSW_KH; kk++) {
                /* columns beyond cnt compute garbage that nobody reads (the block ends at column J) */
                const int d = diag + (rbyte == ab[kk] ? match : mismatch);              /* sw.cpp:55 */
                const int e = E[kk];
                int h; u32 nb;
                if (d >= e && d >= f) { h = d; nb = 8u; }                               /* :60-63 */
                else if (f >= e)      { h = f; nb = 4u; }                               /* :64-67 */
                else                  { h = e; nb = 0u; }                               /* :68-71 */
                const int open = h - go;
                if (open > e - ge) E[kk] = open; else { E[kk] = e - ge; nb |= 2u; }     /* :73-82, tie => extend */
                if (open > f - ge) f = open;     else { f = f - ge;     nb |= 1u; }     /* :84-93 */
                diag = Hp[kk];
                Hp[kk] = h;
                if (kk < 8) lo |= nb << (4 * kk); else hi |= nb << (4 * (kk - 8));
            }
            nib[(size_t)(r - 1) * nstride] = (u64)lo | ((u64)hi << 32);
        }
        ia = ia_; ib = I; ja = ja_; jb = J;
        builds++;
        cells += (u32)(rows * cnt);
    }

    SW_HD bool inside(int i, int j) const { return i >= ia && i <= ib && j >= ja && j <= jb; }
    SW_HD u32 at(int i, int j) const { return (u32)(nib[(size_t)(i - ia) * nstride] >> (4 * (j - ja))) & 15u; }
};

struct PairResult {
    int offset;
    u32 n_ops;
    int score;               /* biased score of the start cell (diagnostic)            */
    u32 builds, hops;        /* blocks recomputed / diagonal hops taken (diagnostic)   */
};

/*
 * The cell the traceback starts from (sw.cpp:155-170), from what the fill left behind: the running
 * last-column maximum (sw.cpp:100-104, kept by sw_fill.h) and the captured last row (scanned here as
 * sw.cpp:116-127 does).  `score` is H of that cell in the fill's representation (+ match + bias);
 * `seg` the reference's seg_length.  Also the whole result of the score-only entry point.
 * select_start is the scan as the reference writes it, one thread per pair: what the emulator runs.  On the device a
 * group of 8 lanes does the same selection (start_cell_group, sw_trace_kernel.cu), checked against the reference's
 * results by the GPU parity tests.
 */
struct StartCell { int i, j, seg, score; };

template <class T>
SW_HD StartCell select_start(const UnitDesc &u, u32 half, u32 G, u32 K, const u32 *colbest4, const u32 *lastrow, int strategy)
{
    const int tl = (int)u.tl[half], ql = (int)u.ql;
    const u32 nslots = u.nstrips * G;
    const u32 *row = lastrow + u.lastrow_off + (size_t)half * nslots * K;
#if SW_ON_DEVICE
    const int mqe = (int)__ldcg(colbest4 + 2 * half), mqe_t = (int)__ldcg(colbest4 + 2 * half + 1);
#else
    const int mqe = (int)colbest4[2 * half], mqe_t = (int)colbest4[2 * half + 1];
#endif
    StartCell sc;
    if (strategy == OS_INDEL) {
        /* the corner (tl, ql): its H is the last real entry of the captured row */
        u32 slot, reg;
        col_to_slot(u, (u32)ql - 1u, slot, reg);
#if SW_ON_DEVICE
        const u32 w = __ldcg(row + (size_t)slot * K + reg);
#else
        const u32 w = row[(size_t)slot * K + reg];
#endif
        sc.i = tl; sc.j = ql; sc.seg = 0; sc.score = T::half(w, (int)half);
    } else if (strategy != OS_LEADING_INDEL) {
        /* last row: sw.cpp:116-127 */
        int best = mqe, best_t = mqe_t, best_q = ql, seg = 0;
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
        sc.i = best_t; sc.j = best_q; sc.seg = seg; sc.score = best;
    } else { sc.i = mqe_t; sc.j = ql; sc.seg = 0; sc.score = mqe; }
    return sc;
}

/* the end of the walk at (I, J) with I == 0 or J == 0: overhang post-processing and offset, sw.cpp:224-248 */
SW_HD u32 *finish_walk(int strategy, u32 state, int run, int I, int J, u32 *p, int &offset)
{
#define SW_EMIT(op, len) do { if ((len) > 0) *--p = ((u32)(len) << 4) | (u32)(op); } while (0)
    if (strategy == OS_SOFTCLIP) {                                              /* sw.cpp:225-229 */
        SW_EMIT(state, run);
        SW_EMIT(OP_S, J);
        offset = I;
    } else if (strategy == OS_IGNORE) {                                         /* sw.cpp:230-233 */
        SW_EMIT(state, run + J);
        offset = I - J;
    } else {                                                                    /* sw.cpp:234-248 */
        SW_EMIT(state, run);
        if (I > 0) SW_EMIT(OP_D, I);
        else if (J > 0) SW_EMIT(OP_I, J);
        offset = 0;
    }
#undef SW_EMIT
    return p;
}

/*
 * The whole path as ONE diagonal hop (Replay::hop with L = min(I, J)): from the start cell (I, J) with H known (the
 * score that selected it) to row 0 or column 0, where H is the boundary score (sw.cpp:29-35, :47-49).
 * H[I][J] - H[I-L][J-L] == the sum of the L substitution scores holds iff every cell on the way is a diagonal step
 * of the reference's traceback (the argument above Replay::hop), i.e. iff the alignment has no gap at all -- then the
 * CIGAR is known without reading a single checkpoint.  `nm` = mismatches among ref[I-L .. I) vs alt[J-L .. J),
 * counted by the caller (a loop here in the emulator, a lane group in sw_start_kernel).  Returns false, writing
 * nothing, when the path has a gap: the pair then goes through traceback_pair.
 */
/* can any mismatch count make clean_path succeed?  (the score deficit of the diagonal must be a whole number of mismatches,
   at most its length) -- lets the caller skip counting: with config 3's scores 349 of 350 gapped pairs fail here */
SW_HD bool clean_path_possible(const StartCell &st, const CellConsts &c, int indel)
{
    const int L = st.i < st.j ? st.i : st.j;
    const int it = st.i - L, jt = st.j - L;
    const int ht = boundary_score(it == 0 ? jt : it, indel, c.s_go, c.s_ge);
    const int need = L * c.s_match - (st.score - c.bias - c.s_match - ht);       /* = mismatches x (match - mismatch) on a clean path */
    const int per = c.s_match - c.s_mismatch;
    return need >= 0 && need % per == 0 && need / per <= L;
}

SW_HD bool clean_path(const StartCell &st, int strategy, const CellConsts &c, int indel, int nm, u32 *ops_end, PairResult &res)
{
    const int L = st.i < st.j ? st.i : st.j;
    const int it = st.i - L, jt = st.j - L;
    const int ht = boundary_score(it == 0 ? jt : it, indel, c.s_go, c.s_ge);
    const int hcur = st.score - c.bias - c.s_match;
    if (hcur - ht != L * c.s_match - nm * (c.s_match - c.s_mismatch)) return false;
    u32 *p = ops_end;
    int run = st.seg;
    if (run > 0 && strategy == OS_SOFTCLIP) { *--p = ((u32)run << 4) | (u32)OP_S; run = 0; }   /* sw.cpp:173-176 */
    p = finish_walk(strategy, OP_M, run + L, it, jt, p, res.offset);
    res.n_ops = (u32)(ops_end - p);
    res.score = st.score;
    res.builds = 0; res.hops = 1;
    return true;
}

/*
 * The walk of one pair (sw.cpp:173-214) as an object with two alternating phases, so that a warp can drive 32 of them
 * around ONE replay site: hop_phase() runs ahead while H is known (each lane on its own), build_phase() rebuilds the
 * block around the current cell and walks inside it (the lanes that need it, together).  The walk is a state machine:
 *   mode 0  at cell (I, J), about to read which of diagonal / F / E produced H
 *   mode 1  inside an insertion, at cell (I, cc): does F extend through it?
 *   mode 2  inside a deletion,  at cell (r, J):  does E extend through it?
 * ops_end points one past the pair's scratch region; elements are written downwards so that [ops_end - n_ops, ops_end)
 * reads left to right.
 */
struct Walker {
    Replay rp;
    u32 *p, *ops_end;
    int I, J, run, strategy, start_score;
    u32 state, next, hops;
    int mode, steplen, cc, r;
    int hcur;                /* H at (I, J), when known: at the start cell (the score that selected it) and on checkpointed cells */
    bool hk, done;

#define SW_EMIT(op, len) do { if ((len) > 0) *--p = ((u32)(len) << 4) | (u32)(op); } while (0)
    SW_HD void begin(const UnitDesc &u, u32 half, bool s32, u32 G, u32 K, const u32 *ckpt, const uint8_t *arena, const CellConsts &c,
                     const StartCell &st, int strategy_, u32 *ops_end_, u64 *nib, u32 nstride)
    {
        I = st.i; J = st.j; run = st.seg; strategy = strategy_; start_score = st.score;
        p = ops_end = ops_end_;
        if (run > 0 && strategy == OS_SOFTCLIP) { SW_EMIT(OP_S, run); run = 0; }   /* sw.cpp:173-176 */
        rp.init(u, G, K, half, s32, ckpt, arena, c, nib, nstride);
        state = OP_M; next = OP_M; hops = 0;
        mode = 0; steplen = 1; cc = 0; r = 0;
        hcur = start_score - c.bias - c.s_match;
        hk = true; done = false;
    }

    /*
     * While H is known, cross blocks with diagonal hops (Replay::hop): a clean stretch of the alignment costs one
     * memory round trip and a few byte compares per NB blocks (Replay::hop<NB>; the caller's choice, the same for every
     * lane of a warp so that they keep running the same code).  Returns when the pair is done or the block around
     * (I, J) has to be rebuilt.
     */
    template <int NB, bool GAPS = true>
    SW_HD void hop_phase()
    {
        /* inside a gap: whole blocks of it, as long as the checkpoints say it runs through them (Replay::ehop / fhop).
           GAPS false: the caller's pairs are short (lane groups of 8, at most 192 columns) and gaps of whole blocks rare: it
           leaves the code out and rebuilds those blocks */
        if (GAPS && mode == 2) { for (int h; (h = rp.ehop(r, J)) != 0;) { steplen += h; r -= h; hops++; } return; }
        if (GAPS && mode == 1) { for (int w; (w = rp.fhop(I, cc)) != 0;) { steplen += w; cc -= w; hops++; } return; }
        while (hk && !done) {
            int hn;
            bool blocked = false;
            const int L = rp.hop<NB>(I, J, hcur, hn, blocked);
            if (L == 0) break;
            if (state == OP_M) run += L;
            else { SW_EMIT(state, run); run = L; state = OP_M; }
            I -= L; J -= L; hcur = hn; hops++;
            if (!(I > 0 && J > 0)) done = true;
            if (blocked) break;                       /* the next block is known to hold a gap: rebuild it */
        }
    }

    /* recompute the block around the current cell and walk inside it */
    SW_HD void build_phase()
    {
        rp.build(mode == 2 ? r : I, mode == 1 ? cc : J);
        for (;;) {
            const int qi = mode == 2 ? r : I, qj = mode == 1 ? cc : J;
            if (!rp.inside(qi, qj)) break;
            const u32 nb = rp.at(qi, qj);
            bool commit = false;
            if (mode == 0) {
                steplen = 1;
                if (nb & 8u) { next = OP_M; I--; J--; commit = true; }           /* diagonal */
                else if (nb & 4u) { next = OP_I; cc = J - 1; mode = 1; }          /* insertion of length lenF[I][J] */
                else { next = OP_D; r = I - 1; mode = 2; }                        /* deletion of length lenE[I][J] */
            } else if (mode == 1) {
                if (nb & 1u) { steplen++; cc--; } else { J -= steplen; mode = 0; commit = true; }
            } else {
                if (nb & 2u) { steplen++; r--; } else { I -= steplen; mode = 0; commit = true; }
            }
            /* a gap that reaches the matrix edge ends there */
            if (mode == 1 && cc < 1) { J -= steplen; mode = 0; commit = true; }
            if (mode == 2 && r < 1) { I -= steplen; mode = 0; commit = true; }
            if (commit) {
                if (next == state) run += steplen;
                else { SW_EMIT(state, run); run = steplen; state = next; }
                if (!(I > 0 && J > 0)) { done = true; break; }
            }
        }
        /* left the block between two cells (not inside a gap): H is known again if we stand on its boundary */
        hk = !done && mode == 0 && rp.boundary_h(I, J, hcur);
    }
#undef SW_EMIT

    /*
     * Between two rebuilt blocks a walk is twelve words (WalkState): the unit's geometry is read again from its
     * descriptor, the block's records are not needed once the walk has left it.  resume() = begin() with the saved
     * scalars; what the walk has emitted so far is in the pair's scratch region already.
     */
    SW_HD void save(u32 sid, u32 out[12]) const
    {
        out[0] = sid; out[1] = (u32)I; out[2] = (u32)J; out[3] = (u32)run;
        out[4] = state | (next << 8) | ((u32)mode << 16) | ((hk ? 1u : 0u) << 24);
        out[5] = (u32)steplen; out[6] = (u32)cc; out[7] = (u32)r;
        out[8] = (u32)hcur; out[9] = (u32)(ops_end - p); out[10] = hops; out[11] = (u32)start_score;
    }
    SW_HD void resume(const UnitDesc &u, u32 half, bool s32, u32 G, u32 K, const u32 *ckpt, const uint8_t *arena, const CellConsts &c,
                      const u32 in[12], int strategy_, u32 *ops_end_, u64 *nib, u32 nstride)
    {
        rp.init(u, G, K, half, s32, ckpt, arena, c, nib, nstride);
        strategy = strategy_; ops_end = ops_end_;
        I = (int)in[1]; J = (int)in[2]; run = (int)in[3];
        state = in[4] & 0xffu; next = (in[4] >> 8) & 0xffu; mode = (int)((in[4] >> 16) & 0xffu); hk = ((in[4] >> 24) & 1u) != 0;
        steplen = (int)in[5]; cc = (int)in[6]; r = (int)in[7];
        hcur = (int)in[8]; p = ops_end_ - in[9]; hops = in[10]; start_score = (int)in[11];
        done = false;
    }

    SW_HD PairResult finish()
    {
        PairResult res;
        p = finish_walk(strategy, state, run, I, J, p, res.offset);
        res.n_ops = (u32)(ops_end - p);
        res.score = start_score;
        res.builds = rp.builds; res.hops = hops;
        return res;
    }
};

/* one pair from start cell to CIGAR: what the emulator runs; the device drives Walkers itself (sw_traceback_kernel) */
/* through_state: the walk is put away after every rebuilt block and taken up again by a fresh Walker, the way the
   device's round kernels hand it from launch to launch (sw_trace_round_kernel) */
SW_HD PairResult traceback_pair(const UnitDesc &u, u32 half, bool s32, u32 G, u32 K, const u32 *ckpt, const uint8_t *arena,
                                const CellConsts &c, const StartCell &st, int strategy, u32 *ops_end, u64 *nib, u32 nstride,
                                bool through_state = false)
{
    Walker w;
    w.begin(u, half, s32, G, K, ckpt, arena, c, st, strategy, ops_end, nib, nstride);
    /* the hop widths the device uses for this kind of unit (sw_trace_kernel.cu) */
    const int nb = u.nstrips >= MGLSW_HOP_WIDE_MIN_STRIPS ? MGLSW_HOP_BLOCKS : (G == (u32)MGLSW_G_SHORT ? MGLSW_HOP_SHORT : 1);
    auto hops = [&]() { if (nb == MGLSW_HOP_BLOCKS) w.hop_phase<MGLSW_HOP_BLOCKS>(); else if (nb == MGLSW_HOP_SHORT) w.hop_phase<MGLSW_HOP_SHORT>(); else w.hop_phase<1>(); };
    hops();
    while (!w.done) {
        if (through_state) {
            u32 rec[12];
            w.save(0u, rec);
            Walker v;
            v.resume(u, half, s32, G, K, ckpt, arena, c, rec, strategy, ops_end, nib, nstride);
            v.rp.builds = w.rp.builds;
            w = v;
        }
        w.build_phase();
        hops();
    }
    return w.finish();
}

}  // namespace mglsw
#endif
