/*
 * sw_fill.h -- the matrix fill, one lane's view.
 *
 * Replaces the reference's fill loops (calculateMatrix sw.cpp:5-146 and its
 * AVX2 twin calculateMatrix_avx sw_avx.cpp:110-322) with a strip-mined
 * wavefront: a group of G lanes sweeps the ref rows, each lane holding K
 * adjacent alt columns in registers (H of the previous row, E and the alt
 * bytes).  It keeps no per-cell record: it parks (H, F) of two columns per
 * lane every step and (H, E) of every column every MGLSW_RB steps
 * (sw_layout.h), from which the traceback recomputes the cells on its path.  Lane l is one row behind lane
 * l - 1; the two values that cross a lane boundary (H and F of the lane's last
 * real column) travel by warp shuffle.  Alt lengths beyond G * K columns are
 * processed as several strips (G == 32 only); the last lane of a strip parks
 * (H, F) per row in a carry column in HBM which lane 0 of the next strip reads
 * back -- the GPU analogue of the score[]/step[] band carry in
 * sw_avx.cpp:36-49,193-198.
 *
 * Endpoint selection is fused into the fill: the lane that owns column ql keeps
 * the running last-column maximum with the reference's tie-break (">=" while the
 * row ascends, sw.cpp:100-104) and stores (value, row) once, at row tl; the
 * registers of row tl are captured for the last-row scan (sw.cpp:116-127).
 *
 * FillLane is SW_HD: the CUDA kernel (sw_fill_kernel.cuh) drives 32 of them in
 * a warp with __shfl_up_sync; tests/emu drives 32 of them in a loop on the CPU.
 */
#ifndef MGL_B200_SW_FILL_H
#define MGL_B200_SW_FILL_H
#include "sw_cell.h"
#include "sw_layout.h"

namespace mglsw {

#if SW_ON_DEVICE
#define SW_LDG(p) __ldg(p)
#else
#define SW_LDG(p) (*(p))
#endif

struct Carry { u32 h, f; };   /* 8 bytes: H[i][last col] + match, F[i][last col + 1] (both + bias) */

/* bytes readable before arena[0] and after the last sequence byte (see FillLane::fetch_ref) */
#define MGLSW_ARENA_FRONT_PAD 64u
#define MGLSW_REF_AHEAD 4u       /* steps between loading a ref byte and using it: covers an L2 round trip */
SW_HD u64 arena_tail_pad(u32 max_ref_len) { return (u64)max_ref_len + 64u; }

/* device buffers of one fill launch */
struct FillArgs {
    const uint8_t *arena;     /* raw sequence bytes; readable on [-FRONT_PAD, bytes + arena_tail_pad(max tl)) */
    const UnitDesc *units;    /* units of this launch, (32 / G) per warp              */
    u32 *ckpt;                /* column and row checkpoints (word offsets in UnitDesc) */
    Carry *carry;             /* strip carry columns (G == 32 launches only)          */
    u32 *lastrow;             /* captured last rows                                   */
    u32 *colbest;             /* per unit 4 words: (value, row) of the last-column maximum of each half */
    u32 n_warps;
    CellConsts c;
};

template <class T, int G, int K>
struct FillLane {
    static constexpr int SC = ((K + 1) / 2 + 3) / 4;   /* chunks per section of a row checkpoint */
    static constexpr int Q2 = 4 * SC;                  /* 16-byte chunks of a row checkpoint     */
    static constexpr int LS16 = (MGLSW_RB + Q2) | 1;   /* lane_stride16(K)                       */
    static constexpr int KM = (K + 1) / 2;       /* registers before the mid checkpoint   */
    static constexpr bool STRIPS = (G == 32);    /* lane groups of 8 always cover the alt in one strip */
    u32 e[K], hm[K], bq[K];
    u32 hleft;                /* H[i-1][first col - 1] + match: diagonal input of register 0 */
    u32 h_out, f_out;         /* this step's values for the right-hand neighbour      */
    u32 f_mid;                /* F leaving register KM - 1 at the lane's latest row    */
    u32 rq0[MGLSW_REF_AHEAD], rq1[MGLSW_REF_AHEAD];   /* ref bytes of the next steps' rows, loaded MGLSW_REF_AHEAD steps early, merged at use */
    Carry cin_next;           /* carry-in of the next step's row (strips > 0, lane 0) */
    u32 bh, bf, bdec;         /* column-0 boundary of the next step's row (lane 0), and its per-row decrement */
    u32 best0, best1;         /* running last-column maximum: S16x2 (value << 16 | row) per half; S32 value, row */

    /* lane constants */
    const uint8_t *rp0, *rp1; /* rp[t] = ref byte of row t - lane                     */
    const uint8_t *alt0, *alt1;
    u32 tl0, tl1, own_rows, ql;
    u32 lane;                 /* lane inside the group                                */
    u32 wlane;                /* lane inside the warp                                 */
    u32 n;                    /* real columns of this lane in the current strip       */
    u32 j0;                   /* 1-based first column of this lane                    */
    u32 slot, nslots;
    u32 indel, active, strip, nstrips, rb_count;
    u32 col_base, col_rem;
    u32 *ckpt_base;           /* the warp's checkpoint regions in HBM                 */
    u32 *ckpt_cur;            /* ... of the current strip                             */
    u32 *stage;               /* device: the warp's staging region in shared memory; emulator: unused */
    Carry *carry_cols;        /* two columns of carry_rows entries                    */
    u32 carry_rows;
    u32 *lastrow;             /* PAIRS regions of nslots * K words                    */
    u32 *colbest;             /* this unit's 4 words                                  */

    SW_HD void setup(const FillArgs &a, const UnitDesc &u, u32 warp_lane, u32 unit_index)
    {
        wlane = warp_lane;
        lane = warp_lane % G;
        const uint8_t *ref0 = a.arena + u.ref_off[0];
        alt0 = a.arena + u.alt_off[0];
        const int two = (T::PAIRS == 2 && u.pair[1] != NO_PAIR);
        const uint8_t *ref1 = two ? a.arena + u.ref_off[1] : ref0;
        alt1 = two ? a.arena + u.alt_off[1] : alt0;
        /*
         * Row i of pair p is ref_p[i - 1].  Rows outside 1..tl are computed and thrown away (the
         * lane skew, a shorter partner in the same unit or warp); their bytes are read where they
         * fall, inside the padding the engine keeps around the arena, instead of being clamped.
         */
        rp0 = ref0 - 1 - (int)lane;
        rp1 = ref1 - 1 - (int)lane;
        tl0 = u.tl[0];
        tl1 = two ? u.tl[1] : u.tl[0];
        own_rows = u.carry_rows - 1;
        ql = u.ql;
        nstrips = STRIPS ? u.nstrips : 1u;
        nslots = nstrips * G;
        col_base = u.col_base;
        col_rem = u.col_rem;
        rb_count = u.rb_count;
        indel = (u.flags & UNIT_INDEL) ? 1u : 0u;
        active = (u.flags & UNIT_ACTIVE) ? 1u : 0u;
        ckpt_base = a.ckpt + u.ckpt_off;
        ckpt_cur = ckpt_base;
        stage = nullptr;
        carry_cols = a.carry + u.carry_off;
        carry_rows = u.carry_rows;
        lastrow = a.lastrow + u.lastrow_off;
        colbest = a.colbest + (size_t)unit_index * 4u;
        h_out = f_out = 0;
    }

    SW_HD int bnd(const CellConsts &c, int x) const { return boundary_score(x, (int)indel, c.s_go, c.s_ge); }

    /* the ref bytes of step t into the last queue slot; the slots move up one step at a time */
    SW_HD void fetch_ref(u32 t)
    {
#pragma unroll
        for (int q = 0; q + 1 < (int)MGLSW_REF_AHEAD; q++) { rq0[q] = rq0[q + 1]; rq1[q] = rq1[q + 1]; }
        rq0[MGLSW_REF_AHEAD - 1] = SW_LDG(rp0 + t);
        rq1[MGLSW_REF_AHEAD - 1] = T::PAIRS == 2 ? (u32)SW_LDG(rp1 + t) : 0u;
    }
    SW_HD Carry fetch_carry(int i) const
    {
        Carry z; z.h = 0; z.f = 0;
        if (!STRIPS || strip == 0 || lane != 0) return z;
        int ii = i < 1 ? 1 : (i > (int)own_rows ? (int)own_rows : i);
        const Carry *col = carry_cols + (size_t)((strip - 1) & 1u) * carry_rows;
#if SW_ON_DEVICE
        const uint2 v = __ldcg(reinterpret_cast<const uint2 *>(col + ii));
        z.h = v.x; z.f = v.y;
        return z;
#else
        return col[ii];
#endif
    }

    /* state of the row above row 1 (sw.cpp:10-40): E seeds, H[0][j], and the diagonal of register 0 */
    SW_HD void reset_rows(const CellConsts &c)
    {
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int b = bnd(c, (int)j0 + k);
            e[k] = T::rep(c.bias + b - c.s_go);
            hm[k] = T::rep(c.bias + b + c.s_match);
        }
        hleft = T::rep(c.bias + c.s_match + bnd(c, (int)j0 - 1));
        best0 = T::PAIRS == 2 ? 0u : 0x80000000u;
        best1 = 0u;
    }

    SW_HD void strip_begin(const CellConsts &c, u32 s)
    {
        strip = s;
        ckpt_cur = ckpt_base + (size_t)s * rb_count * (LS16 * 128u);
        slot = s * G + lane;
        n = col_base + (slot < col_rem ? 1u : 0u);
        j0 = 1u + (slot < col_rem ? slot * (col_base + 1) : col_rem * (col_base + 1) + (slot - col_rem) * col_base);
#pragma unroll
        for (int k = 0; k < K; k++) {
            u32 j = j0 + (u32)k;                      /* 1-based column */
            if (j > ql) j = ql;                       /* virtual register: any byte will do */
            u32 b = SW_LDG(alt0 + (j - 1));
            if (T::PAIRS == 2) b |= (u32)SW_LDG(alt1 + (j - 1)) << 16;
            bq[k] = T::negb(b);                       /* stored negated: the cell's comparison is one add-then-min (sw_cell.h) */
        }
        reset_rows(c);
        h_out = T::rep(c.bias);
        f_out = T::rep(c.bias);
        f_mid = T::rep(c.bias);
        for (u32 q = 0; q < MGLSW_REF_AHEAD; q++) fetch_ref(q);
        cin_next = fetch_carry(-(int)lane);
        /*
         * Column 0 as lane 0 of strip 0 sees it at step t (row t): H[t][0] + match and the F seed
         * H[t][0] - go (sw.cpp:47-49 and :22-26,38).  Constant without the INDEL boundary, a ramp
         * of -ge per row with it; kept as running values so a step costs two subtractions.
         * The value for t == 0 is never used: a lane sits out the steps of rows < 1 (step()).
         */
        bdec = indel ? T::rep(c.s_ge) : 0u;
        bh = T::rep(c.bias + c.s_match + (indel ? -c.s_go + c.s_ge : 0));
        bf = T::rep(c.bias - c.s_go + (indel ? -c.s_go + c.s_ge : 0));
    }

    /*
     * A lane only WORKS on rows 1..tl0 (tl0 >= tl1: the longer ref of a unit sits in the low half);
     * at the other steps of the warp's loop -- before its row 1 (the lane skew), after its last row
     * (a longer ref in another lane group) -- it sits out and its registers keep what they hold.  So
     * the state strip_begin() sets up is still intact when row 1 arrives, and after the loop the
     * registers still hold row tl0: the last row (sw.cpp:116-127) and the last-column maximum
     * (sw.cpp:100-104) are stored once, by all lanes together, in strip_end().  Only a high half
     * with a shorter ref has to be captured on the way, at its row tl1.
     */
    SW_HD bool works_at(u32 t) const { const int i = (int)t - (int)lane; return i >= 1 && i <= (int)tl0; }
    SW_HD bool hi_event_at(u32 t) const { return T::PAIRS == 2 && tl1 != tl0 && (int)t - (int)lane == (int)tl1; }
    /* true if this lane works at all four steps t0 .. t0 + 3 and has no high-half capture among them */
    SW_HD bool quiet4(u32 t0) const
    {
        const int i0 = (int)t0 - (int)lane;
        bool q = i0 >= 1 && i0 + 3 <= (int)tl0;
        if (T::PAIRS == 2 && tl1 != tl0 && (int)tl1 >= i0 && (int)tl1 <= i0 + 3) q = false;
        return q;
    }
    SW_HD void capture_hi()
    {
        if (!active) return;
#pragma unroll
        for (int k = 0; k < K; k++) lastrow[(size_t)(nslots + slot) * K + k] = hm[k];
        if (lane == G - 1 && strip + 1 == nstrips) { colbest[2] = best1 >> 16; colbest[3] = best1 & 0xffffu; }
    }
    SW_HD void strip_end()
    {
        if (!active) return;
        const bool owner = (lane == G - 1) && (strip + 1 == nstrips);
#pragma unroll
        for (int k = 0; k < K; k++) lastrow[(size_t)slot * K + k] = hm[k];
        if (T::PAIRS == 2 && tl1 == tl0) {
#pragma unroll
            for (int k = 0; k < K; k++) lastrow[(size_t)(nslots + slot) * K + k] = hm[k];
        }
        if (owner) {
            if (T::PAIRS == 2) {
                colbest[0] = best0 >> 16; colbest[1] = best0 & 0xffffu;
                if (tl1 == tl0) { colbest[2] = best1 >> 16; colbest[3] = best1 & 0xffffu; }
            } else { colbest[0] = best0; colbest[1] = best1; }
        }
    }

    /*
     * One wavefront step: this lane computes ref row i = t - lane over its K
     * registers.  h_sh / f_sh are the neighbour's previous-step outputs.
     * MASKED = false: the caller knows that every lane of the warp works at this step (quiet4());
     * MASKED = true: the lane checks for itself, and sits out or captures its high half.
     */
    template <bool MASKED = true>
    SW_HD void step(const CellConsts &c, u32 t, u32 h_sh, u32 f_sh)
    {
        const int i = (int)t - (int)lane;
        const u32 ra = T::PAIRS == 2 ? (rq0[0] | (rq1[0] << 16)) : rq0[0];
        fetch_ref(t + MGLSW_REF_AHEAD);

        u32 h_in = h_sh, f_in = f_sh;
        if (STRIPS) {
            const Carry cin = cin_next;
            cin_next = fetch_carry(i + 1);
            if (lane == 0) {
                h_in = strip == 0 ? bh : cin.h;
                f_in = strip == 0 ? bf : cin.f;
            }
        } else if (lane == 0) {
            h_in = bh;
            f_in = bf;
        }
        bh -= bdec;
        bf -= bdec;

        if (!MASKED || works_at(t)) {
            u32 diag = hleft;
            hleft = h_in;
            u32 f = f_in;
            u32 f_km2 = f_in, f_mid_ = f_in;
#pragma unroll
            for (int k = 0; k < K; k++) {
                const u32 old = hm[k];
                sw_cell<T>(c, ra, bq[k], diag, e[k], f, hm[k]);
                diag = old;
                if (k == K - 2) f_km2 = f;
                if (k == KM - 1) f_mid_ = f;
            }
            f_mid = f_mid_;
            if (n == (u32)K) { h_out = hm[K - 1]; f_out = f; }
            else             { h_out = K >= 2 ? hm[K >= 2 ? K - 2 : 0] : h_in; f_out = f_km2; }

            /* running maximum of this lane's last column; meaningful in the lane that owns column ql */
            if (T::PAIRS == 2) {
#if SW_ON_DEVICE
                const u32 k0 = __byte_perm((u32)i, h_out, 0x5410);       /* low half of h_out : low 16 bits of the row  */
                const u32 k1 = __byte_perm((u32)i, h_out, 0x7610);       /* high half of h_out : low 16 bits of the row */
#else
                const u32 k0 = (h_out << 16) | ((u32)i & 0xffffu);
                const u32 k1 = (h_out & 0xffff0000u) | ((u32)i & 0xffffu);
#endif
                best0 = k0 > best0 ? k0 : best0;
                best1 = k1 > best1 ? k1 : best1;
            } else if ((int)h_out >= (int)best0) { best0 = h_out; best1 = (u32)i; }

            if (STRIPS && active && lane == G - 1) {
                Carry *col = carry_cols + (size_t)(strip & 1u) * carry_rows;
#if SW_ON_DEVICE
                __stcg(reinterpret_cast<uint2 *>(col + i), make_uint2(h_out, f_out));
#else
                col[i].h = h_out; col[i].f = f_out;
#endif
            }
            if (MASKED && hi_event_at(t)) capture_hi();
        }

        /*
         * Checkpoints (sw_layout.h).  Every step: (H + match, F) leaving register KM - 1 and leaving the
         * lane, into this lane's slice of the region being assembled.  After the last step of a row
         * block: H + match and E of every register, then the region goes out to HBM.  On the device
         * the region is assembled in shared memory and leaves with one bulk asynchronous copy.
         */
#if SW_ON_DEVICE
        u32 *slice = stage + wlane * (LS16 * 4u);
        if ((t % MGLSW_RB) == 0) {                     /* the bulk copy of the previous region must have read the staging buffer */
            if (wlane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
        }
        *reinterpret_cast<uint4 *>(slice + (t % MGLSW_RB) * 4u) = make_uint4(hm[KM - 1], f_mid, h_out, f_out);
#else
        u32 *slice = ckpt_cur + (size_t)(t / MGLSW_RB) * (LS16 * 128u) + wlane * (LS16 * 4u);
        { u32 *p = slice + (t % MGLSW_RB) * 4u; p[0] = hm[KM - 1]; p[1] = f_mid; p[2] = h_out; p[3] = f_out; }
#endif
        if ((t % MGLSW_RB) == MGLSW_RB - 1) {
#pragma unroll
            for (int q = 0; q < Q2; q++) {
                /* chunk q: section q / SC = 0 H[0, KM)  1 E[0, KM)  2 H[KM, K)  3 E[KM, K); register = section base + 4 (q % SC) + x */
                u32 w[4];
#pragma unroll
                for (int x = 0; x < 4; x++) {
                    const int sec = q / SC, off = 4 * (q % SC) + x;
                    const int reg = (sec & 2) ? KM + off : off;
                    const bool ok = (sec & 2) ? reg < K : off < KM;
                    const int rs = ok ? reg : 0;
                    w[x] = ok ? ((sec & 1) ? e[rs] : hm[rs]) : 0u;
                }
                u32 *p = slice + MGLSW_RB * 4u + q * 4;
#if SW_ON_DEVICE
                *reinterpret_cast<uint4 *>(p) = make_uint4(w[0], w[1], w[2], w[3]);
#else
                p[0] = w[0]; p[1] = w[1]; p[2] = w[2]; p[3] = w[3];
#endif
            }
#if SW_ON_DEVICE
            /* one TMA bulk store (UBLKCP) moves the assembled region: shared memory -> HBM, off the issue slots */
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (wlane == 0) {
                u32 *dst = ckpt_cur + (size_t)(t / MGLSW_RB) * (LS16 * 128u);
                const u32 src = (u32)__cvta_generic_to_shared(stage);
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n\tcp.async.bulk.commit_group;"
                             :: "l"(dst), "r"(src), "n"(LS16 * 512) : "memory");
            }
#endif
        }
    }
};

}  // namespace mglsw
#endif
