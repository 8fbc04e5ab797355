/* sw_plan.cpp -- see sw_plan.h. */
#include "sw_plan.h"

#include <algorithm>
#include <cstring>

#include "sw_cell.h"
#include "sw_layout.h"

namespace mglsw {

namespace {

long long absll(long long v) { return v < 0 ? -v : v; }

/* margin for the garbage rows a lane computes before its row 1 and after its last row */
long long ramp_margin(const Scores &s, int G)
{
    return (long long)(G + 4) * ((long long)s.match + absll(s.mismatch) + s.go + s.ge);
}

/* a coarse magnitude of one cell's score differences; only used to refuse absurd score sets */
long long diff_bound(const Scores &s)
{
    const long long g = std::max(s.go, s.ge);
    return 2 * ((long long)s.match + g) + absll(s.mismatch) + s.go + s.ge;
}

struct Keyed {
    u64 key;
    u32 idx;
};

/* stable LSD radix sort on the 64-bit key, 16 bits a pass; passes whose digit is constant are skipped */
void radix_sort(std::vector<Keyed> &v)
{
    const size_t n = v.size();
    if (n < 2) return;
    std::vector<Keyed> tmp(n);
    std::vector<u32> hist(65536);
    Keyed *src = v.data(), *dst = tmp.data();
    for (int pass = 0; pass < 4; pass++) {
        const int sh = 16 * pass;
        std::fill(hist.begin(), hist.end(), 0u);
        for (size_t k = 0; k < n; k++) hist[(src[k].key >> sh) & 0xffffu]++;
        if (hist[(src[0].key >> sh) & 0xffffu] == n) continue;
        u32 run = 0;
        for (u32 &h : hist) { const u32 c = h; h = run; run += c; }
        for (size_t k = 0; k < n; k++) dst[hist[(src[k].key >> sh) & 0xffffu]++] = src[k];
        std::swap(src, dst);
    }
    if (src != v.data()) std::memcpy(v.data(), src, n * sizeof(Keyed));
}

}  // namespace

PlanStatus check_scores(const Scores &s)
{
    if (s.match <= 0 && s.mismatch == 0 && s.go == 0 && s.ge == 0) return PLAN_BAD_SCORES;
    /* keeps every intermediate of a cell far inside 32 bits whatever the lengths */
    if (2 * diff_bound(s) >= (1ll << 28)) return PLAN_BAD_SCORES;
    return PLAN_OK;
}

static void value_bounds(const Scores &s, u32 tl, u32 ql, int G, long long &lo, long long &hi)
{
    const long long m = ramp_margin(s, G);
    lo = -(3ll * s.go + (long long)s.ge * ((long long)tl + ql + 2) + absll(s.mismatch)) - m;
    hi = (long long)s.match * ((long long)std::min(tl, ql) + 2) + m;
}

bool s16_bounds(const Scores &s, u32 tl, u32 ql, long long &lo, long long &hi)
{
    value_bounds(s, tl, ql, MGLSW_G_SHORT, lo, hi);
    if (s.match - s.mismatch >= 32768) return false;
    if (tl > 65535u) return false;                         /* the fused last-column maximum packs (value, row) in 32 bits */
    return hi - lo < 65000;
}

PlanStatus plan_chunk(const PairIn *pairs, u32 n, const Scores &s, bool allow_s16, ChunkPlan &out)
{
    out = ChunkPlan();
    if (check_scores(s) != PLAN_OK) return PLAN_BAD_SCORES;
    out.pair_unit.assign(n, 0);
    out.ops_end.assign(n, 0);
    out.units.reserve(n / 2 + 64);

    /* pass 1: per-pair arithmetic class, and the shared 16-bit bias */
    std::vector<unsigned char> cand(n, 0);
    long long need_bias = 0;
    for (u32 p = 0; p < n; p++) {
        const PairIn &pr = pairs[p];
        if (pr.tl == 0 || pr.ql == 0 || pr.tl > (1u << 24) || pr.ql > (1u << 24)) return PLAN_BAD_PAIR;
        long long lo, hi;
        value_bounds(s, pr.tl, pr.ql, MGLSW_G_LONG, lo, hi);
        if (-lo >= (1ll << 30) || hi >= (1ll << 30)) return PLAN_BAD_PAIR;
        if (allow_s16 && pr.ql <= (u32)MGLSW_G_SHORT * MGLSW_KMAX && s16_bounds(s, pr.tl, pr.ql, lo, hi)) {
            cand[p] = 1;
            need_bias = std::max(need_bias, -lo);
        }
    }
    const long long bias16 = std::min<long long>(need_bias, 49152);
    out.s16_bias = (int)bias16;

    /* pass 2: sort key = (kind, G, K, nstrips, indel, ql, tl), all descending so heavy work launches first;
       one launch per (kind, G, K): the strip count is a per-warp run-time value, not a kernel */
    std::vector<Keyed> order(n);
    for (u32 p = 0; p < n; p++) {
        const PairIn &pr = pairs[p];
        int kind = KIND_S32;
        if (cand[p]) {
            long long lo, hi;
            s16_bounds(s, pr.tl, pr.ql, lo, hi);
            if (-lo <= bias16 && hi <= 65535 - bias16) kind = KIND_S16X2;
        }
        u32 G, nstrips;
        if (pr.ql <= (u32)MGLSW_G_SHORT * MGLSW_KMAX) { G = MGLSW_G_SHORT; nstrips = 1; }
        else { G = MGLSW_G_LONG; nstrips = ceil_div_u32(pr.ql, (u32)MGLSW_G_LONG * MGLSW_KMAX); }
        const u32 K = regs_per_lane(pr.ql, nstrips * G);
        const u32 indel = (pr.strategy & (OS_INDEL | OS_LEADING_INDEL)) ? 1u : 0u;
        /* 1 + 1 + 12 + 5 + 1 bits of launch key, then 24 + 20 bits of shape (tl clipped for the key only) */
        u64 key = (u64)(kind == KIND_S32 ? 1 : 0);
        key = (key << 1) | (G == (u32)MGLSW_G_LONG ? 1u : 0u);
        key = (key << 5) | K;
        key = (key << 12) | std::min<u32>(nstrips, 4095);
        key = (key << 1) | indel;
        key = (key << 24) | pr.ql;
        key = (key << 20) | std::min<u32>(pr.tl, (1u << 20) - 1);
        order[p].key = ~key;            /* ascending sort of ~key == descending key */
        order[p].idx = p;
    }
    radix_sort(order);

    /* pass 3: units, warps, launches */
    auto launch_key = [](u64 k) { return (~k) >> 57; };          /* kind, G, K */
    auto pair_key = [](u64 k) { return (~k) >> 20; };            /* ... + indel + ql     */
    u32 pos = 0;
    while (pos < n) {
        const u64 lk = launch_key(order[pos].key);
        u32 end = pos;
        while (end < n && launch_key(order[end].key) == lk) end++;

        const PairIn &first = pairs[order[pos].idx];
        Launch L;
        const u64 nk = ~order[pos].key;
        L.kind = ((nk >> 63) & 1) ? KIND_S32 : KIND_S16X2;
        L.G = ((nk >> 62) & 1) ? MGLSW_G_LONG : MGLSW_G_SHORT;
        L.nstrips = L.G == MGLSW_G_SHORT ? 1u : ceil_div_u32(first.ql, (u32)MGLSW_G_LONG * MGLSW_KMAX);   /* of its first unit */
        L.K = (int)regs_per_lane(first.ql, L.nstrips * L.G);
        L.first_unit = (u32)out.units.size();
        L.cells = 0;
        const u32 upw = 32u / (u32)L.G;                          /* units per warp */

        u32 q = pos;
        while (q < end) {
            UnitDesc u;
            std::memset(&u, 0, sizeof u);
            const u32 p0 = order[q].idx;
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
            if (L.kind == KIND_S16X2 && q < end && pair_key(order[q].key) == pair_key(order[q - 1].key)) {
                const u32 p1 = order[q].idx;
                const PairIn &b = pairs[p1];
                u.pair[1] = p1; u.ref_off[1] = b.ref_off; u.alt_off[1] = b.alt_off; u.tl[1] = b.tl;
                out.pair_unit[p1] = (u32)out.units.size() * 2 + 1;
                L.cells += (u64)b.tl * b.ql;
                out.pairs_s16++;
                q++;
            }
            u.nstrips = L.G == MGLSW_G_SHORT ? 1u : ceil_div_u32(u.ql, (u32)MGLSW_G_LONG * MGLSW_KMAX);
            const u32 slots = u.nstrips * (u32)L.G;
            u.col_base = u.ql / slots;
            u.col_rem = u.ql % slots;
            u.carry_rows = std::max(u.tl[0], u.tl[1]) + 1;
            u.flags = UNIT_ACTIVE | ((a.strategy & (OS_INDEL | OS_LEADING_INDEL)) ? UNIT_INDEL : 0u)
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
    return PLAN_OK;
}

}  // namespace mglsw
