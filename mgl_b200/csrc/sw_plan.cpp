/* sw_plan.cpp -- see sw_plan.h. */
#include "sw_plan.h"

#include <algorithm>

#include <cstring>

#include "sw_cell.h"
#include "sw_layout.h"
#include "sw_plan_rules.h"

namespace mglsw {

namespace {

long long absll(long long v) { return v < 0 ? -v : v; }

/* a coarse magnitude of one cell's score differences; only used to refuse absurd score sets */
long long diff_bound(const Scores &s)
{
    const long long g = std::max(s.go, s.ge);
    return 2 * ((long long)s.match + g) + absll(s.mismatch) + s.go + s.ge;
}

/*
 * Stable sort of (key, index) by the 64-bit key.  The keys of one chunk differ in few bits (config 2: 9 bits
 * of ref length, 4 of alt length, 2 of K), so the bits that vary at all are first gathered into a compact key
 * (order-preserving: the gather keeps bit order), and that is sorted by LSD radix passes of <= 11 bits --
 * one pass for a batch of short reads instead of three 16-bit passes over 64 Ki buckets.
 */
u64 gather_bits(u64 v, u64 mask)                     /* portable: one iteration per mask bit */
{
    u64 r = 0;
    for (u64 bit = 1; mask; mask &= mask - 1, bit <<= 1)
        if (v & mask & (0 - mask)) r |= bit;
    return r;
}

#if defined(__x86_64__) && defined(__GNUC__)
__attribute__((target("bmi2"))) void gather_all_bmi2(u64 *key, u32 n, u64 mask)
{
    for (u32 k = 0; k < n; k++) key[k] = __builtin_ia32_pext_di(key[k], mask);
}
#endif

void gather_all(u64 *key, u32 n, u64 mask)
{
#if defined(__x86_64__) && defined(__GNUC__)
    static const bool have_bmi2 = __builtin_cpu_supports("bmi2");
    if (have_bmi2) { gather_all_bmi2(key, n, mask); return; }
#endif
    for (u32 k = 0; k < n; k++) key[k] = gather_bits(key[k], mask);
}

/* returns the mask of the key bits that vary; sort_key then holds the sorted keys reduced to those bits */
u64 sort_by_key(ChunkPlan &w, u32 n)
{
    std::vector<u64> &key = w.sort_key;
    std::vector<u32> &idx = w.sort_idx;
    if (n < 2) return 0;
    u64 diff = 0;
    for (u32 k = 1; k < n; k++) diff |= key[k] ^ key[0];
    if (diff == 0) return 0;
    const int bits = __builtin_popcountll(diff);
    if (bits < 64) gather_all(key.data(), n, diff);
    const int passes = (bits + 10) / 11, per = (bits + passes - 1) / passes;       /* <= 11 bits per pass */
    w.sort_tmp.resize(n); w.sort_idx_tmp.resize(n); w.sort_hist.resize((size_t)1 << per);
    u64 *ks = key.data(), *kd = w.sort_tmp.data();
    u32 *is = idx.data(), *id = w.sort_idx_tmp.data();
    const u32 dm = (1u << per) - 1u;
    for (int pass = 0; pass < passes; pass++) {
        const int sh = pass * per;
        u32 *hist = w.sort_hist.data();
        std::fill(hist, hist + dm + 1, 0u);
        for (u32 k = 0; k < n; k++) hist[(ks[k] >> sh) & dm]++;
        u32 run = 0;
        for (u32 d = 0; d <= dm; d++) { const u32 c = hist[d]; hist[d] = run; run += c; }
        for (u32 k = 0; k < n; k++) { const u32 at = hist[(ks[k] >> sh) & dm]++; kd[at] = ks[k]; id[at] = is[k]; }
        std::swap(ks, kd); std::swap(is, id);
    }
    if (is != idx.data()) { std::memcpy(idx.data(), is, n * sizeof(u32)); std::memcpy(key.data(), ks, n * sizeof(u64)); }
    return diff;
}

}  // namespace

PlanStatus check_scores(const Scores &s)
{
    if (s.match <= 0 && s.mismatch == 0 && s.go == 0 && s.ge == 0) return PLAN_BAD_SCORES;
    /* keeps every intermediate of a cell far inside 32 bits whatever the lengths */
    if (2 * diff_bound(s) >= (1ll << 28)) return PLAN_BAD_SCORES;
    return PLAN_OK;
}

bool s16_bounds(const Scores &s, u32 tl, u32 ql, long long &lo, long long &hi) { return s16_bounds_hd(s, tl, ql, lo, hi); }

PlanStatus plan_chunk(const PairIn *pairs, u32 n, const Scores &s, bool allow_s16, ChunkPlan &out)
{
    out.reset();
    if (check_scores(s) != PLAN_OK) return PLAN_BAD_SCORES;
    out.pair_unit.assign(n, 0);
    out.ops_end.assign(n, 0);
    out.units.reserve(n + 64);

    /* pass 1: per-pair arithmetic class, and the shared 16-bit bias */
    std::vector<unsigned char> &cand = out.cand;
    cand.assign(n, 0);
    long long need_bias = 0;
    for (u32 p = 0; p < n; p++) {
        const PairIn &pr = pairs[p];
        if (!pair_ok(s, pr.tl, pr.ql)) return PLAN_BAD_PAIR;
        long long need;
        if (s16_candidate(s, allow_s16, pr.tl, pr.ql, need)) {
            cand[p] = 1;
            need_bias = std::max(need_bias, need);
        }
    }
    const long long bias16 = s16_bias_of(need_bias);
    out.s16_bias = (int)bias16;

    /* pass 2: sort key = (kind, G, K, nstrips, indel, ql, tl), all descending so heavy work launches first;
       one launch per (kind, G, K): the strip count is a per-warp run-time value, not a kernel */
    out.full_key.resize(n); out.sort_key.resize(n); out.sort_idx.resize(n);
    u64 *const fkey = out.full_key.data();
    for (u32 p = 0; p < n; p++) {
        const PairIn &pr = pairs[p];
        const int kind = (cand[p] && s16_fits(s, pr.tl, pr.ql, bias16)) ? KIND_S16X2 : KIND_S32;
        const u64 key = pair_sort_key(kind, pr.tl, pr.ql, pr.strategy);      /* sw_plan_rules.h: ascending sort == heavy first */
        fkey[p] = key;
        out.sort_key[p] = key;
        out.sort_idx[p] = p;
    }
    const u64 varying = sort_by_key(out, n);
    const u32 *const order = out.sort_idx.data();     /* pair ids in launch order */
    /* the sorted keys, reduced to their varying bits: two keys agree on a field iff their reduced keys agree on the
       field's reduced mask (read in order: no random access to the full keys) */
    const u64 *const skey = out.sort_key.data();
    const u64 kmask = (1ull << MGLSW_KEY_BITS) - 1;
    const u64 m_launch = gather_bits(kmask & (~0ull << MGLSW_KEY_LAUNCH_SHIFT), varying);     /* kind, G, K, pipelined */
    const u64 m_pair = gather_bits(kmask & (~0ull << MGLSW_KEY_PAIR_SHIFT), varying);         /* ... + indel, alt length */

    /* pass 3: units, warps, launches */
    u32 pos = 0;
    while (pos < n) {
        u32 end = pos;
        while (end < n && ((skey[end] ^ skey[pos]) & m_launch) == 0) end++;

        const PairIn &first = pairs[order[pos]];
        Launch L;
        const u32 nk = sort_key_to_key((u32)fkey[order[pos]]);
        L.kind = key_kind(nk);
        L.G = key_G(nk);
        L.pipe = key_pipe(nk);
        L.nstrips = L.G == MGLSW_G_SHORT ? 1u : ceil_div_u32(first.ql, (u32)MGLSW_G_LONG * MGLSW_KMAX);   /* of its first unit */
        L.K = (int)regs_per_lane(first.ql, L.nstrips * L.G);
        L.first_unit = (u32)out.units.size();
        L.cells = 0;
        const u32 upw = 32u / (u32)L.G;                          /* units per warp */

        u32 q = pos;
        while (q < end) {
            UnitDesc u;
            std::memset(&u, 0, sizeof u);
            const u32 p0 = order[q];
            const PairIn &a = pairs[p0];
            u.pair[0] = p0; u.pair[1] = NO_PAIR;
            u.ref_off[0] = u.ref_off[1] = a.ref_off;
            u.alt_off[0] = u.alt_off[1] = a.alt_off;
            u.tl[0] = u.tl[1] = a.tl;
            u.ql = a.ql;
            out.pair_unit[p0] = (u32)out.units.size() * 2;
            L.cells += (u64)a.tl * a.ql;
            (L.kind == KIND_S32 ? out.pairs_s32 : out.pairs_s16)++;
            q++;
            /* the next position rides along unless it opens a unit of its own (sw_plan_rules.h opens_unit) */
            if (L.kind == KIND_S16X2 && q < n && !opens_unit(L.kind, q, ((skey[q] ^ skey[q - 1]) & m_pair) != 0)) {
                u32 p1 = order[q];
                L.cells += (u64)pairs[p1].tl * pairs[p1].ql;
                if (pairs[p1].tl > a.tl) {                   /* the longer ref sits in the low half (sw_fill.h); the key orders refs in steps of 8 only */
                    u.pair[0] = p1; u.ref_off[0] = pairs[p1].ref_off; u.alt_off[0] = pairs[p1].alt_off; u.tl[0] = pairs[p1].tl;
                    out.pair_unit[p1] = (u32)out.units.size() * 2;
                    p1 = p0;
                }
                const PairIn &b = pairs[p1];
                u.pair[1] = p1; u.ref_off[1] = b.ref_off; u.alt_off[1] = b.alt_off; u.tl[1] = b.tl;
                out.pair_unit[p1] = (u32)out.units.size() * 2 + 1;
                out.pairs_s16++;
                q++;
            }
            u.nstrips = L.G == MGLSW_G_SHORT ? 1u : ceil_div_u32(u.ql, (u32)MGLSW_G_LONG * MGLSW_KMAX);
            const u32 slots = u.nstrips * (u32)L.G;
            u.col_base = u.ql / slots;
            u.col_rem = u.ql % slots;
            u.carry_rows = std::max(u.tl[0], u.tl[1]) + 1;
            u.flags = UNIT_ACTIVE | ((pairs[p0].strategy & (OS_INDEL | OS_LEADING_INDEL)) ? UNIT_INDEL : 0u)
                    | (L.G == MGLSW_G_LONG ? UNIT_G32 : 0u) | (L.kind == KIND_S32 ? UNIT_S32 : 0u);
            u.regs = (u32)L.K;
            out.units.push_back(u);
        }
        /* pad the last warp with inactive copies */
        while ((out.units.size() - L.first_unit) % upw) {
            UnitDesc u = out.units.back();
            u.flags &= ~UNIT_ACTIVE;
            out.units.push_back(u);
        }
        L.n_units = (u32)out.units.size() - L.first_unit;
        L.n_warps = L.n_units / upw;

        /* per-warp geometry and buffer placement */
        for (u32 w = 0; w < L.n_warps; w++) {
            UnitDesc *wu = &out.units[L.first_unit + w * upw];
            u32 tlmax = 0;
            for (u32 g = 0; g < upw; g++) tlmax = std::max(tlmax, wu[g].carry_rows - 1);
            const u32 rb = row_blocks(tlmax, (u32)L.G);
            for (u32 g = 0; g < upw; g++) {
                wu[g].tlmax = tlmax;
                wu[g].rb_count = rb;
                wu[g].group = g;
                wu[g].ckpt_off = out.ckpt_words;
                if (wu[g].flags & UNIT_ACTIVE) {
                    if (L.G == MGLSW_G_LONG) {            /* only strip-mined units hand columns over through HBM */
                        wu[g].carry_off = out.carry_entries;
                        out.carry_entries += 2ull * wu[g].carry_rows;
                    }
                    wu[g].lastrow_off = out.lastrow_words;
                    out.lastrow_words += (u64)(L.kind == KIND_S16X2 ? 2 : 1) * wu[g].nstrips * L.G * L.K;
                }
            }
            out.ckpt_words += warp_ckpt_words(wu[0].nstrips, rb, (u32)L.K);   /* G = 32: one unit per warp; G = 8: one strip */
        }
        out.cells += L.cells;
        out.launches.push_back(L);
        pos = end;
    }

    /* CIGAR scratch: tl + ql + 4 elements per pair, in chunk order */
    u64 acc = 0;
    for (u32 p = 0; p < n; p++) {
        acc += (u64)pairs[p].tl + pairs[p].ql + 4;
        out.ops_end[p] = acc;
    }
    out.scratch_words = acc;
    out.n_units = (u32)out.units.size();
    return PLAN_OK;
}

}  // namespace mglsw
