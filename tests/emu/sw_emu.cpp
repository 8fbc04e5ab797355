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

#include "sw_config.h"
#include "sw_fill.h"
#include "sw_plan.h"
#include "sw_trace.h"

using namespace mglsw;

namespace {

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

/*
 * Align a batch through planner + emulated kernels.  ops_out: per pair a
 * region of ops_stride words, elements left to right.  Returns 0, or a
 * negative code.  stats (optional, 6 words): launches, s16 pairs, s32 pairs, units, blocks replayed, diagonal hops.
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
    const PlanStatus ps = plan_chunk(pairs.data(), (u32)n, s, allow_s16 != 0, plan);
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
    if (stats) stats[4] = stats[5] = 0;
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
        if (u.flags & UNIT_S32) { r = traceback_pair<S32>(u, half, G, u.regs, ckpt.data(), arena, make_consts<S32>(s, 0), colbest.data() + (size_t)(plan.pair_unit[p] >> 1) * 4, lastrow.data(), strategies[p], end, nibbuf, 1u); n32++; }
        else { r = traceback_pair<S16x2>(u, half, G, u.regs, ckpt.data(), arena, make_consts<S16x2>(s, plan.s16_bias), colbest.data() + (size_t)(plan.pair_unit[p] >> 1) * 4, lastrow.data(), strategies[p], end, nibbuf, 1u); n16++; }
        if (stats) { stats[4] += r.builds; stats[5] += r.hops; }
        if (const char *dump = std::getenv("EMU_DUMP")) {        /* developer aid: per-pair replay / hop counts in thread order */
            static FILE *fp = nullptr;
            if (!fp) fp = std::fopen(dump, "w");
            if (fp) { std::fprintf(fp, "%u %u %u\n", plan.pair_unit[p], r.builds, r.hops); std::fflush(fp); }
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
