/*
 * tests/emu/sw_emu.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Lock-step CPU emulation of the GPU path: the real host planner
 * (mgl_b200/csrc/sw_plan.cpp) plus the very same SW_HD fill-lane and traceback
 * code the CUDA kernels run, with a warp emulated as 32 FillLane objects whose
 * shuffles are array reads.  It exists so that the kernel LOGIC (cell
 * predicates, lane skew, strip carry, trace layout, traceback) can be checked
 * against the oracle in a container without a GPU.  Nothing under mgl_b200/
 * links or loads it; the product has no CPU path.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <vector>

#include <numeric>

#include "sw_config.h"
#include "sw_fill.h"
#include "sw_plan.h"
#include "sw_plan_dev.h"
#include "sw_trace.h"

using namespace mglsw;

namespace {

int g_device_plan = 0;
int g_through_state = 0;      /* 1: every walk is saved and resumed between two rebuilt blocks (Walker::save / resume) */
int g_clean_shortcut = 1;      /* the whole-path shortcut of sw_start_kernel (sw_trace.h clean_path); 0: every pair is walked */

/*
 * The device planner (sw_plan_dev.h) on the CPU: the same stage bodies the CUDA kernels of sw_plan_dev.cu run, one
 * loop iteration per thread, with std::stable_sort / std::partial_sum where the device uses a radix sort and scans.
 * Produces the ChunkPlan the emulated kernels consume, so that planner + fill + traceback are checked end to end.
 */
PlanStatus plan_chunk_dev_emu(const PairIn *pairs, u32 n, u64 arena_bytes, const Scores &s, bool allow_s16, bool score_only, ChunkPlan &out)
{
    out.reset();
    if (check_scores(s) != PLAN_OK) return PLAN_BAD_SCORES;
    std::vector<u64> ro(n), ao(n);
    std::vector<u32> tl(n), ql(n);
    std::vector<int> strat(n);
    for (u32 p = 0; p < n; p++) { ro[p] = pairs[p].ref_off; ao[p] = pairs[p].alt_off; tl[p] = pairs[p].tl; ql[p] = pairs[p].ql; strat[p] = pairs[p].strategy; }
    PlanDevIn in;
    in.ref_off = ro.data(); in.alt_off = ao.data(); in.tl = tl.data(); in.ql = ql.data(); in.strategy = strat.data();
    in.n = n; in.arena_bytes = arena_bytes; in.s = s; in.allow_s16 = allow_s16 ? 1 : 0; in.score_only = score_only ? 1 : 0;
    PlanState st;
    std::memset(&st, 0, sizeof st);
    std::vector<u64> ops_end(n);
    { u64 run = 0; for (u32 p = 0; p < n; p++) { run += (u64)tl[p] + ql[p] + 4; ops_end[p] = run; } }     /* the host makes this while staging */
    in.ops_end = ops_end.data();
    std::vector<u32> key(n), skey(n), idx(n), order(n), uhead(n), urank(n);
    std::vector<u32> need16(n), hi16(n);
    for (u32 p = 0; p < n; p++) { const u32 need = pd_stage_a(in, p, &st, need16.data(), hi16.data()); st.need_bias = std::max(st.need_bias, need); }
    for (u32 p = 0; p < n; p++) {
        const int kind = pd_stage_b(in, p, &st, need16.data(), hi16.data(), key.data(), idx.data());
        (kind == KIND_S16X2 ? st.pairs_s16 : st.pairs_s32)++;
        st.cells += (u64)tl[p] * ql[p];
    }
    std::stable_sort(idx.begin(), idx.end(), [&](u32 a, u32 b) { return key[a] < key[b]; });
    for (u32 q = 0; q < n; q++) { order[q] = idx[q]; skey[q] = key[idx[q]]; }
    for (u32 q = n; q-- > 0;) pd_stage_c(q, skey.data(), &st, uhead.data());              /* any thread order: heads arrive unsorted */
    { u32 run = 0; for (u32 q = 0; q < n; q++) { urank[q] = run; run += uhead[q]; } }
    pd_stage_e(n, skey.data(), urank.data(), uhead.data(), &st);
    const u32 U = st.sum.n_units;
    std::vector<UnitDesc> units(U);
    std::memset(units.data(), 0xee, U * sizeof(UnitDesc));                                /* device memory is not zeroed either */
    for (u32 q = 0; q < n; q++) pd_stage_f(in, q, order.data(), uhead.data(), urank.data(), &st, units.data());
    std::vector<U3> u3(U), u3x(U);
    for (u32 u = 0; u < U; u++) pd_stage_g(u, &st, units.data(), u3.data());
    { U3 run; run.a = run.b = run.c = 0; for (u32 u = 0; u < U; u++) { u3x[u] = run; run = run + u3[u]; } }
    for (u32 u = 0; u < U; u++) pd_stage_h(in, u, &st, units.data(), u3.data(), u3x.data());
    pd_stage_i(in, &st);

    const PlanSummary &S = st.sum;
    if (S.err_invalid) return PLAN_BAD_PAIR;
    if (S.err_pair) return PLAN_BAD_PAIR;
    if (S.overflow) return PLAN_BAD_PAIR;
    out.units = units;
    out.ops_end = ops_end;
    out.pair_unit.assign(n, 0xffffffffu);
    for (u32 u = 0; u < U; u++) {
        if (!(units[u].flags & UNIT_ACTIVE)) continue;
        for (u32 h = 0; h < 2; h++) if (units[u].pair[h] != NO_PAIR && (h == 0 || !(units[u].flags & UNIT_S32))) {
            if (out.pair_unit[units[u].pair[h]] != 0xffffffffu) return PLAN_BAD_PAIR;     /* a pair placed twice */
            out.pair_unit[units[u].pair[h]] = u * 2 + h;
        }
    }
    for (u32 p = 0; p < n; p++) if (out.pair_unit[p] == 0xffffffffu) return PLAN_BAD_PAIR;   /* a pair not placed */
    for (u32 l = 0; l < S.n_launches; l++) {
        const PlanLaunchDev &D = S.launch[l];
        Launch L;
        L.kind = (int)D.kind; L.G = (int)D.G; L.K = (int)D.K; L.pipe = (int)D.pipe;
        L.nstrips = units[D.first_unit].nstrips; L.first_unit = D.first_unit; L.n_units = D.n_units; L.n_warps = D.n_warps; L.cells = 0;
        out.launches.push_back(L);
    }
    out.ckpt_words = S.ckpt_words; out.carry_entries = S.carry_entries; out.lastrow_words = S.lastrow_words;
    out.scratch_words = S.scratch_words; out.cells = S.cells; out.pairs_s16 = S.pairs_s16; out.pairs_s32 = S.pairs_s32;
    out.s16_bias = (int)S.s16_bias;
    return PLAN_OK;
}

template <class T, int G, int K>
void emu_launch(const FillArgs &a, const Launch &L)
{
    const u32 upw = 32 / G;
    std::vector<FillLane<T, G, K>> lanes(32);
    for (u32 w = 0; w < L.n_warps; w++) {
        const UnitDesc *wu = a.units + L.first_unit + (size_t)w * upw;
        for (u32 l = 0; l < 32; l++) lanes[l].setup(a, wu[l / G], l, L.first_unit + w * upw + l / G);
        const u32 steps = wu[0].rb_count * MGLSW_RB;
        for (u32 s = 0; s < wu[0].nstrips; s++) {
            for (u32 l = 0; l < 32; l++) lanes[l].strip_begin(a.c, s);
            for (u32 t = 0; t < steps; t++) {
                /* the kernel's choice between the unmasked and the masked step, per group of four steps (sw_fill_kernel.cuh) */
                bool quiet = true;
                for (u32 l = 0; l < 32; l++) quiet = quiet && lanes[l].quiet4(t & ~3u);
                u32 hs[32], fs[32];
                for (u32 l = 0; l < 32; l++) { hs[l] = lanes[l].h_out; fs[l] = lanes[l].f_out; }
                for (u32 l = 0; l < 32; l++) {
                    const u32 src = (l % G) ? l - 1 : l;      /* __shfl_up_sync(.., 1, G) */
                    if (quiet) lanes[l].template step<false>(a.c, t, hs[src], fs[src]);
                    else lanes[l].template step<true>(a.c, t, hs[src], fs[src]);
                }
            }
            for (u32 l = 0; l < 32; l++) lanes[l].strip_end();
        }
    }
}

#define EMU_CASE(TT, GG, KK) if (L.G == GG && L.K == KK) { emu_launch<TT, GG, KK>(a, L); return true; }

bool emu_dispatch(const FillArgs &a, const Launch &L)
{
    if (L.kind == KIND_S16X2) { MGLSW_FOR_K_1_24(EMU_CASE, S16x2, 8) }
    else {
        MGLSW_FOR_K_1_24(EMU_CASE, S32, 8)
        MGLSW_FOR_K_7_24(EMU_CASE, S32, 32)
    }
    return false;
}

}  // namespace

extern "C" {

/* 1: plan with the device planner's stage bodies (sw_plan_dev.h) instead of the host planner (sw_plan.cpp) */
void emu_use_device_plan(int on) { g_device_plan = on; }
void emu_use_clean_shortcut(int on) { g_clean_shortcut = on; }
void emu_use_through_state(int on) { g_through_state = on; }

/*
 * Both planners on one chunk: sizes and bucket structure must agree exactly (the unit arrays may differ in the
 * order of equal-key pairs only).  out[0..7] host, out[8..15] device: launches, units, s16 pairs, s32 pairs,
 * checkpoint words, last-row words, carry entries, scratch words.  Returns 0, or a planner's status.
 */
int emu_compare_planners(int n, const uint64_t *ref_off, const uint32_t *ref_len, const uint64_t *alt_off, const uint32_t *alt_len,
                         const int32_t *strategies, int match, int mismatch, int gap_open, int gap_extend, int allow_s16, uint64_t *out16)
{
    const Scores s = normalise_scores(match, mismatch, gap_open, gap_extend);
    std::vector<PairIn> pairs(n);
    u64 bytes = 0;
    for (int p = 0; p < n; p++) {
        pairs[p].ref_off = ref_off[p]; pairs[p].alt_off = alt_off[p]; pairs[p].tl = ref_len[p]; pairs[p].ql = alt_len[p]; pairs[p].strategy = strategies[p];
        bytes = std::max<u64>(bytes, std::max(ref_off[p] + ref_len[p], alt_off[p] + alt_len[p]));
    }
    ChunkPlan a, b;
    PlanStatus ps = plan_chunk(pairs.data(), (u32)n, s, allow_s16 != 0, a);
    if (ps != PLAN_OK) return (int)ps;
    ps = plan_chunk_dev_emu(pairs.data(), (u32)n, bytes, s, allow_s16 != 0, false, b);
    if (ps != PLAN_OK) return 100 + (int)ps;
    const ChunkPlan *pl[2] = {&a, &b};
    for (int k = 0; k < 2; k++) {
        uint64_t *o = out16 + 8 * k;
        o[0] = pl[k]->launches.size(); o[1] = pl[k]->units.size(); o[2] = pl[k]->pairs_s16; o[3] = pl[k]->pairs_s32;
        o[4] = pl[k]->ckpt_words; o[5] = pl[k]->lastrow_words; o[6] = pl[k]->carry_entries; o[7] = pl[k]->scratch_words;
    }
    /* launch tables agree field by field */
    if (a.launches.size() != b.launches.size()) return 200;
    for (size_t l = 0; l < a.launches.size(); l++) {
        const Launch &x = a.launches[l], &y = b.launches[l];
        if (x.kind != y.kind || x.G != y.G || x.K != y.K || x.pipe != y.pipe || x.first_unit != y.first_unit || x.n_units != y.n_units || x.n_warps != y.n_warps) return 201 + (int)l;
    }
    return 0;
}

/*
 * Align a batch through planner + emulated kernels.  ops_out: per pair a
 * region of ops_stride words, elements left to right.  Returns 0, or a
 * negative code.  stats (optional, 7 words): launches, s16 pairs, s32 pairs, units, blocks replayed, diagonal hops,
 * pairs written down by the whole-path shortcut.
 */
int emu_align_batch(int n, const uint8_t *arena, const uint64_t *ref_off, const uint32_t *ref_len,
                    const uint64_t *alt_off, const uint32_t *alt_len, const int32_t *strategies,
                    int match, int mismatch, int gap_open, int gap_extend, int allow_s16,
                    int32_t *offsets, uint32_t *n_ops, uint32_t *ops_out, int ops_stride, uint32_t *stats)
{
    const Scores s = normalise_scores(match, mismatch, gap_open, gap_extend);
    std::vector<PairIn> pairs(n);
    for (int p = 0; p < n; p++) {
        pairs[p].ref_off = ref_off[p]; pairs[p].alt_off = alt_off[p];
        pairs[p].tl = ref_len[p]; pairs[p].ql = alt_len[p];
        pairs[p].strategy = strategies[p];
    }
    ChunkPlan plan;
    u64 caller_bytes = 0;
    for (int p = 0; p < n; p++) caller_bytes = std::max<u64>(caller_bytes, std::max(ref_off[p] + ref_len[p], alt_off[p] + alt_len[p]));
    const PlanStatus ps = g_device_plan ? plan_chunk_dev_emu(pairs.data(), (u32)n, caller_bytes, s, allow_s16 != 0, false, plan)
                                        : plan_chunk(pairs.data(), (u32)n, s, allow_s16 != 0, plan);
    if (ps != PLAN_OK) return -(int)ps;

    std::vector<u32> ckpt(plan.ckpt_words + 4, 0xdeadbeefu);
    std::vector<Carry> carry(plan.carry_entries + 1);
    std::vector<u32> colbest(plan.units.size() * 4 + 4, 0);
    /* the fill reads a few bytes outside each ref (sw_fill.h fetch_ref): give the arena the padding the engine gives it */
    u64 arena_bytes = 0;
    u32 max_tl = 0;
    for (int p = 0; p < n; p++) {
        arena_bytes = std::max<u64>(arena_bytes, std::max(ref_off[p] + ref_len[p], alt_off[p] + alt_len[p]));
        max_tl = std::max(max_tl, ref_len[p]);
    }
    std::vector<uint8_t> padded(MGLSW_ARENA_FRONT_PAD + arena_bytes + arena_tail_pad(max_tl), 0x5a);
    std::memcpy(padded.data() + MGLSW_ARENA_FRONT_PAD, arena, arena_bytes);
    arena = padded.data() + MGLSW_ARENA_FRONT_PAD;
    std::vector<u32> lastrow(plan.lastrow_words + 1, 0);
    std::vector<u32> scratch(plan.scratch_words + 1, 0);

    u32 n16 = 0, n32 = 0;
    if (stats) stats[4] = stats[5] = stats[6] = 0;
    for (const Launch &L : plan.launches) {
        FillArgs a;
        a.arena = arena; a.units = plan.units.data(); a.ckpt = ckpt.data();
        a.carry = carry.data(); a.lastrow = lastrow.data(); a.colbest = colbest.data(); a.n_warps = L.n_warps;
        a.c = L.kind == KIND_S16X2 ? make_consts<S16x2>(s, plan.s16_bias) : make_consts<S32>(s, 0);
        if (!emu_dispatch(a, L)) { std::fprintf(stderr, "emu: no kernel for kind %d G %d K %d\n", L.kind, L.G, L.K); return -10; }
    }
    for (int p = 0; p < n; p++) {
        const UnitDesc &u = plan.units[plan.pair_unit[p] >> 1];
        const u32 half = plan.pair_unit[p] & 1u;
        const u32 G = (u.flags & UNIT_G32) ? MGLSW_G_LONG : MGLSW_G_SHORT;
        PairResult r;
        u64 nibbuf[MGLSW_TRACE_BUF];
        u32 *end = scratch.data() + plan.ops_end[p];
        const bool s32 = (u.flags & UNIT_S32) != 0;
        const CellConsts cc = s32 ? make_consts<S32>(s, 0) : make_consts<S16x2>(s, plan.s16_bias);
        const u32 *cb4 = colbest.data() + (size_t)(plan.pair_unit[p] >> 1) * 4;
        const StartCell sc = s32 ? select_start<S32>(u, half, G, u.regs, cb4, lastrow.data(), strategies[p])
                                 : select_start<S16x2>(u, half, G, u.regs, cb4, lastrow.data(), strategies[p]);
        if (s32) n32++; else n16++;
        /* what sw_start_kernel does on the device: a path without a gap is written down at once (sw_trace.h clean_path) */
        bool clean = false;
        if (g_clean_shortcut && clean_path_possible(sc, cc, (u.flags & UNIT_INDEL) ? 1 : 0)) {
            const int L = sc.i < sc.j ? sc.i : sc.j;
            const uint8_t *rf = arena + u.ref_off[half] + (sc.i - L), *al = arena + u.alt_off[half] + (sc.j - L);
            int nm = 0;
            for (int t = 0; t < L; t++) nm += rf[t] != al[t] ? 1 : 0;
            clean = clean_path(sc, strategies[p], cc, (u.flags & UNIT_INDEL) ? 1 : 0, nm, end, r);
            if (clean && stats) stats[6]++;
        }
        if (!clean) {
            r = traceback_pair(u, half, s32, G, u.regs, ckpt.data(), arena, cc, sc, strategies[p], end, nibbuf, 1u, g_through_state != 0);
        }
        if (stats) { stats[4] += r.builds; stats[5] += r.hops; }
        if (const char *dump = std::getenv("EMU_DUMP")) {        /* developer aid: per-pair replay / hop counts in thread order */
            static FILE *fp = nullptr;
            if (!fp) fp = std::fopen(dump, "w");
            if (fp) { std::fprintf(fp, "%u %u %u %d %d\n", plan.pair_unit[p], r.builds, r.hops, r.score, clean ? 1 : 0); std::fflush(fp); }
        }
        offsets[p] = r.offset;
        n_ops[p] = r.n_ops;
        if ((int)r.n_ops > ops_stride) return -11;
        std::memcpy(ops_out + (size_t)p * ops_stride, end - r.n_ops, r.n_ops * sizeof(u32));
    }
    if (stats) { stats[0] = (u32)plan.launches.size(); stats[1] = n16; stats[2] = n32; stats[3] = (u32)plan.units.size(); }
    return 0;
}

}  // extern "C"
