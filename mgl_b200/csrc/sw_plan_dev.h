/*
 * sw_plan_dev.h -- the chunk planner as data-parallel stages: one body per stage, run by one thread per pair /
 * sorted position / unit.  sw_plan_dev.cu runs the bodies inside ONE kernel of one thread block per chunk (with a
 * block-wide radix sort and two scans between them); tests/emu runs the same bodies in loops on the CPU, so the
 * planning LOGIC is checked against the oracle in a container without a GPU.  The host planner (sw_plan.cpp) computes
 * the same plan on host threads and stays available ("host_plan" option; the 2-bit transport uses it).
 *
 * Why on the device: on a box whose host cores are shared by eight GPUs the host planner's ~140 ns of CPU per pair
 * (sort by shape, 112-byte unit descriptors, checkpoint layout) bound the end-to-end step.  Here the
 * host uploads the pair table (36 bytes per pair) and reads back a summary of a few kilobytes per chunk; the unit
 * descriptors are born in HBM.
 *
 * It plays the role of the reference's per-call setup (align_avx, sw_avx.cpp:6-60: widening, padding, band
 * geometry) for a whole chunk; the reference has no batch and nothing to sort.
 *
 * Stages (n pairs, U padded units):
 *   A  per pair      validity, running maximum of the 16-bit bias the chunk needs
 *   B  per pair      arithmetic (needs the bias), bucket key                      -> radix sort (key, pair)
 *   C  per position  which positions open a unit, launch heads                    -> sum-scan: unit ranks
 *   E  one thread    launch table: positions, units (padded to whole warps), warps
 *   F  per position  unit descriptors (and the inactive copies that pad a launch's last warp)
 *   G  per unit      warp geometry: rows the warp iterates, row blocks, checkpoint words   -> sum-scan: placements
 *   H  per unit      checkpoint / last-row / carry offsets; totals
 *   I  one thread    summary for the host
 */
#ifndef MGL_B200_SW_PLAN_DEV_H
#define MGL_B200_SW_PLAN_DEV_H
#include "sw_plan_rules.h"

namespace mglsw {

#define MGLSW_PLAN_MAX_LAUNCHES 192      /* kind (2) x lane group (2) x K (24) x pipelined (2) */

struct PlanDevIn {
    const u64 *ref_off, *alt_off;        /* the chunk's slice of the caller's pair table (device copies) */
    const u32 *tl, *ql;
    const int *strategy;
    const u64 *ops_end;                  /* per pair: one-past-the-end word of its CIGAR scratch (a running sum the host makes while staging) */
    u32 n;
    u64 arena_bytes;                     /* the caller's arena size: every pair must lie inside it */
    Scores s;
    int allow_s16;
    int score_only;                      /* checkpoint regions folded onto a sink that fits the L2 (sw_engine.cu plan_range) */
};

struct PlanLaunchDev {                   /* one fill launch = one (arithmetic, lane group, K, pipelined) bucket */
    u32 q0, q1;                          /* sorted positions [q0, q1) */
    u32 urank0, units_real;              /* units before q0 (unpadded), units of this launch before padding */
    u32 first_unit, n_units, n_warps;    /* in the chunk's unit array, padded to whole warps */
    u32 kind, G, K, pipe;
    u32 reserved;
};

struct PlanSummary {                     /* what the host reads back per chunk */
    u32 n_launches, n_units;
    u32 err_invalid, err_pair;           /* 0: none, else 0xffffffff - (first offending pair) */
    u32 s16_bias, pairs_s16, pairs_s32, overflow;
    u64 cells, ckpt_words, carry_entries, lastrow_words, scratch_words;
    PlanLaunchDev launch[MGLSW_PLAN_MAX_LAUNCHES];
};

struct PlanState {                       /* zeroed before stage A */
    u32 need_bias;
    u32 n_heads;
    u32 head_q[MGLSW_PLAN_MAX_LAUNCHES];
    unsigned long long maxw;             /* largest checkpoint footprint of one warp, words */
    unsigned long long cells;
    u32 pairs_s16, pairs_s32;
    u64 tot_ckpt, tot_lastrow, tot_carry;
    PlanSummary sum;
};

struct U3 { u64 a, b, c; };              /* per unit: checkpoint words (warp leaders only), last-row words, carry entries */
SW_HD U3 operator+(const U3 &x, const U3 &y) { U3 r; r.a = x.a + y.a; r.b = x.b + y.b; r.c = x.c + y.c; return r; }

#if defined(__CUDA_ARCH__)
#define PD_ATOMIC_MAX_U32(p, v) atomicMax((p), (v))
#define PD_ATOMIC_ADD_U32(p, v) atomicAdd((p), (v))
#define PD_ATOMIC_MAX_U64(p, v) atomicMax((p), (unsigned long long)(v))
#else
#define PD_ATOMIC_MAX_U32(p, v) ([](u32 *q_, u32 v_) { const u32 o_ = *q_; if (v_ > o_) *q_ = v_; return o_; }((p), (v)))
#define PD_ATOMIC_ADD_U32(p, v) ([](u32 *q_, u32 v_) { const u32 o_ = *q_; *q_ = o_ + v_; return o_; }((p), (v)))
#define PD_ATOMIC_MAX_U64(p, v) ([](unsigned long long *q_, unsigned long long v_) { const unsigned long long o_ = *q_; if (v_ > o_) *q_ = v_; return o_; }((p), (unsigned long long)(v)))
#endif

/* A: validity (an offset check that cannot wrap); returns the 16-bit bias this pair needs (0: not a u16x2 candidate),
   of which the caller keeps the maximum in st->need_bias.  need16[p] / hi16[p] keep the pair's bounds for stage B. */
SW_HD u32 pd_stage_a(const PlanDevIn &in, u32 p, PlanState *st, u32 *need16, u32 *hi16)
{
    need16[p] = 0u; hi16[p] = 0u;
    const u32 tl = in.tl[p], ql = in.ql[p];
    const u64 ro = in.ref_off[p], ao = in.alt_off[p], ab = in.arena_bytes;
    if (ro > ab || tl > ab - ro || ao > ab || ql > ab - ao) { PD_ATOMIC_MAX_U32(&st->sum.err_invalid, 0xffffffffu - p); return 0u; }
    if (!pair_ok(in.s, tl, ql)) { PD_ATOMIC_MAX_U32(&st->sum.err_pair, 0xffffffffu - p); return 0u; }
    long long lo, hi;
    if (!(in.allow_s16 != 0 && ql <= (u32)MGLSW_G_SHORT * MGLSW_KMAX && s16_bounds_hd(in.s, tl, ql, lo, hi))) return 0u;    /* sw_plan_rules.h s16_candidate */
    need16[p] = (u32)(-lo); hi16[p] = (u32)(hi < 0 ? 0 : hi);       /* hi - lo < 65000 and lo < 0: both fit easily */
    return (u32)(-lo);
}

/* B: arithmetic and bucket key; returns the arithmetic (the wrapper counts pairs and cells) */
SW_HD int pd_stage_b(const PlanDevIn &in, u32 p, const PlanState *st, const u32 *need16, const u32 *hi16, u32 *key, u32 *idx)
{
    const u32 tl = in.tl[p], ql = in.ql[p];
    const u32 bias16 = (u32)s16_bias_of((long long)st->need_bias);
    /* a candidate stays inside 16 bits under the chunk's bias (sw_plan_rules.h s16_fits, from the bounds stage A kept) */
    const int kind = (need16[p] != 0u && need16[p] <= bias16 && hi16[p] <= 65535u - bias16) ? KIND_S16X2 : KIND_S32;
    key[p] = pair_sort_key(kind, tl, ql, in.strategy[p]);        /* ascending sort == heavy work first */
    idx[p] = p;
    return kind;
}

/* C: which positions open a unit (sw_plan_rules.h opens_unit), and the heads of launches */
SW_HD void pd_stage_c(u32 q, const u32 *skey, PlanState *st, u32 *uhead)
{
    const u32 d = q == 0 ? ~0u : (skey[q] ^ skey[q - 1]);
    uhead[q] = opens_unit(key_kind(sort_key_to_key(skey[q])), q, (d >> MGLSW_KEY_PAIR_SHIFT) != 0) ? 1u : 0u;
    if ((d >> MGLSW_KEY_LAUNCH_SHIFT) != 0) {
        const u32 slot = PD_ATOMIC_ADD_U32(&st->n_heads, 1u);
        if (slot < MGLSW_PLAN_MAX_LAUNCHES) st->head_q[slot] = q;
    }
}

/* E: the launch table (one thread; at most MGLSW_PLAN_MAX_LAUNCHES entries) */
SW_HD void pd_stage_e(u32 n, const u32 *skey, const u32 *urank, const u32 *uhead, PlanState *st)
{
    PlanSummary &S = st->sum;
    u32 nh = st->n_heads;
    if (nh > MGLSW_PLAN_MAX_LAUNCHES) { nh = MGLSW_PLAN_MAX_LAUNCHES; S.overflow = 1; }
    for (u32 a = 1; a < nh; a++) {                               /* insertion sort of the heads by position */
        const u32 v = st->head_q[a];
        u32 b = a;
        for (; b > 0 && st->head_q[b - 1] > v; b--) st->head_q[b] = st->head_q[b - 1];
        st->head_q[b] = v;
    }
    const u32 total = n ? urank[n - 1] + uhead[n - 1] : 0u;
    u32 run = 0;
    for (u32 l = 0; l < nh; l++) {
        PlanLaunchDev &L = S.launch[l];
        L.q0 = st->head_q[l];
        L.q1 = l + 1 < nh ? st->head_q[l + 1] : n;
        L.urank0 = urank[L.q0];
        L.units_real = (L.q1 < n ? urank[L.q1] : total) - L.urank0;
        const u32 key = sort_key_to_key(skey[L.q0]);
        L.kind = (u32)key_kind(key); L.G = (u32)key_G(key); L.K = (u32)key_K(key); L.pipe = (u32)key_pipe(key);
        const u32 upw = 32u / L.G;
        L.n_units = ceil_div_u32(L.units_real, upw) * upw;
        L.first_unit = run;
        L.n_warps = L.n_units / upw;
        L.reserved = 0;
        run += L.n_units;
    }
    S.n_launches = nh;
    S.n_units = run;
}

/* the launch that holds sorted position q / unit u (binary search over at most 192 entries) */
SW_HD u32 pd_launch_of_q(const PlanSummary &S, u32 q)
{
    u32 lo = 0, hi = S.n_launches;
    while (hi - lo > 1) { const u32 mid = (lo + hi) / 2; if (S.launch[mid].q0 <= q) lo = mid; else hi = mid; }
    return lo;
}
SW_HD u32 pd_launch_of_unit(const PlanSummary &S, u32 u)
{
    u32 lo = 0, hi = S.n_launches;
    while (hi - lo > 1) { const u32 mid = (lo + hi) / 2; if (S.launch[mid].first_unit <= u) lo = mid; else hi = mid; }
    return lo;
}

/* F: the unit a position opens; the launch's last unit also writes the inactive copies that pad its last warp */
SW_HD void pd_stage_f(const PlanDevIn &in, u32 q, const u32 *order, const u32 *uhead, const u32 *urank,
                      const PlanState *st, UnitDesc *units)
{
    if (!uhead[q]) return;
    const PlanSummary &S = st->sum;
    const PlanLaunchDev &L = S.launch[pd_launch_of_q(S, q)];
    const u32 r = urank[q] - L.urank0;
    UnitDesc u;
    u.ckpt_off = 0; u.reserved0 = 0; u.carry_off = 0; u.lastrow_off = 0;
    u.tlmax = 0; u.rb_count = 0; u.group = 0;
    u32 p0 = order[q];
    /* the next position rides along when it opens no unit of its own; the longer ref sits in the low half
       (sw_fill.h) -- the key orders refs in steps of 8 rows only */
    const bool two = L.kind == (u32)KIND_S16X2 && q + 1 < in.n && !uhead[q + 1];
    u32 p1 = two ? order[q + 1] : NO_PAIR;
    if (two && in.tl[p1] > in.tl[p0]) { const u32 t = p0; p0 = p1; p1 = t; }
    u.pair[0] = p0; u.pair[1] = NO_PAIR;
    u.ref_off[0] = u.ref_off[1] = in.ref_off[p0];
    u.alt_off[0] = u.alt_off[1] = in.alt_off[p0];
    u.tl[0] = u.tl[1] = in.tl[p0];
    u.ql = in.ql[p0];
    if (two) {
        u.pair[1] = p1; u.ref_off[1] = in.ref_off[p1]; u.alt_off[1] = in.alt_off[p1]; u.tl[1] = in.tl[p1];
    }
    unit_geometry(u, L.G);
    u.flags = UNIT_ACTIVE | ((in.strategy[p0] & (OS_INDEL | OS_LEADING_INDEL)) ? UNIT_INDEL : 0u)
            | (L.G == (u32)MGLSW_G_LONG ? UNIT_G32 : 0u) | (L.kind == (u32)KIND_S32 ? UNIT_S32 : 0u);
    u.regs = L.K;
    units[L.first_unit + r] = u;
    if (r + 1 == L.units_real) {
        u.flags &= ~UNIT_ACTIVE;
        for (u32 x = L.units_real; x < L.n_units; x++) units[L.first_unit + x] = u;
    }
}

/* G: warp geometry, by the warp's first unit; every unit reports what the placement scan sums */
SW_HD void pd_stage_g(u32 u, PlanState *st, UnitDesc *units, U3 *u3)
{
    const PlanSummary &S = st->sum;
    const PlanLaunchDev &L = S.launch[pd_launch_of_unit(S, u)];
    const u32 upw = 32u / L.G, g = (u - L.first_unit) % upw;
    U3 v; v.a = 0; v.b = 0; v.c = 0;
    if (g == 0) {
        u32 tlmax = 0;
        for (u32 x = 0; x < upw; x++) { const u32 t = units[u + x].carry_rows - 1; tlmax = t > tlmax ? t : tlmax; }
        const u32 rb = row_blocks(tlmax, L.G);
        for (u32 x = 0; x < upw; x++) { units[u + x].tlmax = tlmax; units[u + x].rb_count = rb; units[u + x].group = x; }
        v.a = warp_ckpt_words(units[u].nstrips, rb, L.K);        /* G = 32: one unit per warp; G = 8: one strip */
        if (v.a > st->maxw) PD_ATOMIC_MAX_U64(&st->maxw, v.a);   /* the plain read only spares most of the atomics */
    }
    const UnitDesc &me = units[u];
    if (me.flags & UNIT_ACTIVE) {
        v.b = unit_lastrow_words(me, (int)L.kind, L.G, L.K);
        if (L.G == (u32)MGLSW_G_LONG) v.c = 2ull * me.carry_rows;     /* only strip-mined units hand columns over through HBM */
    }
    u3[u] = v;
}

/* H: placements from the exclusive scan of G's reports; the last unit leaves the totals */
SW_HD void pd_stage_h(const PlanDevIn &in, u32 u, PlanState *st, UnitDesc *units, const U3 *u3, const U3 *u3x)
{
    const PlanSummary &S = st->sum;
    const PlanLaunchDev &L = S.launch[pd_launch_of_unit(S, u)];
    const u32 upw = 32u / L.G, g = (u - L.first_unit) % upw, w = (u - L.first_unit) / upw;
    const u64 maxw = st->maxw;
    u64 sink = (64ull << 20) / (maxw * 4 + 1);
    sink = sink < 1 ? 1 : (sink > 64 ? 64 : sink);
    UnitDesc &me = units[u];
    me.ckpt_off = in.score_only ? (u64)(w % sink) * maxw : u3x[u - g].a;
    if (me.flags & UNIT_ACTIVE) {
        me.lastrow_off = u3x[u].b;
        if (L.G == (u32)MGLSW_G_LONG) me.carry_off = u3x[u].c;
    }
    if (u + 1 == S.n_units) {
        st->tot_ckpt = in.score_only ? sink * maxw : u3x[u].a + u3[u].a;
        st->tot_lastrow = u3x[u].b + u3[u].b;
        st->tot_carry = u3x[u].c + u3[u].c;
    }
}

/* I: the summary */
SW_HD void pd_stage_i(const PlanDevIn &in, PlanState *st)
{
    const u64 *ops_end = in.ops_end;
    PlanSummary &S = st->sum;
    S.s16_bias = (u32)s16_bias_of((long long)st->need_bias);
    S.pairs_s16 = st->pairs_s16; S.pairs_s32 = st->pairs_s32;
    S.cells = st->cells;
    S.ckpt_words = st->tot_ckpt; S.lastrow_words = st->tot_lastrow; S.carry_entries = st->tot_carry;
    S.scratch_words = in.n ? ops_end[in.n - 1] : 0;
}

}  // namespace mglsw
#endif
