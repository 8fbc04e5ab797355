/*
 * tests/emu/ph_emu.cpp -- TEST INFRASTRUCTURE: a lock-step CPU emulation of a warp running the SAME PhLane code
 * the CUDA kernel compiles (mgl_b200/csrc/ph_lane.h); shuffles become array reads, the shared-memory carry row a
 * vector.  Lets the kernel logic be checked against the golden vectors in a container without a GPU.
 */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

#include "../../mgl_b200/csrc/ph_lane.h"
#include "../../mgl_b200/csrc/ph_tables.h"

using namespace mglph;

template <class N, int RPL, int G = 32>
static N run_pair(const Tables<N> &tb, const PairView &pv)
{
    const int nb = ph_bands(pv.rl, RPL, G);
    const N y_initial = tb.initial / (N)pv.hl;
    std::vector<Top<N>> carry((size_t)pv.hl + 2);
    std::vector<PhLane<N, RPL, G>> L(G);
    N result = (N)0;
    for (int b = 0; b < nb; b++) {
        for (int l = 0; l < G; l++) L[(size_t)l].band_begin(tb, pv, b, l, y_initial);
        for (int t = 0; t < pv.hl + G - 1; t++) {
            Top<N> bots[32];
            for (int l = 0; l < G; l++) bots[l] = L[(size_t)l].bottom;              /* what __shfl_up_sync delivers */
            for (int l = 0; l < G; l++) {
                Top<N> top = l > 0 ? bots[l - 1] : Top<N>{(N)0, (N)0, (N)0};
                if (l == 0) {
                    const int c = t + 1;
                    if (b == 0) { top.m = (N)0; top.x = (N)0; top.y = y_initial; }
                    else if (c <= pv.hl) top = carry[(size_t)c];
                }
                const bool active = L[(size_t)l].step(pv, t, top, ph_hap_byte(pv, t, l));
                if (active && l == G - 1 && b + 1 < nb) carry[(size_t)(t - (G - 1) + 1)] = L[(size_t)l].bottom;
            }
        }
    }
    const int ll = ph_last_lane(pv.rl, RPL, G);
    result = L[(size_t)ll].result;
    return result;
}

template <class N>
static N run_any(const Tables<N> &tb, const PairView &pv, int rpl)
{
    switch (rpl) {
    case 1: return run_pair<N, 1>(tb, pv);
    case 2: return run_pair<N, 2>(tb, pv);
    case 3: return run_pair<N, 3>(tb, pv);
    case 4: return run_pair<N, 4>(tb, pv);
    case 5: return run_pair<N, 5>(tb, pv);
    case 6: return run_pair<N, 6>(tb, pv);
    case 8: return run_pair<N, 8>(tb, pv);
    /* half-warp groups: rpl = 100 + rows per lane */
    case 105: return run_pair<N, 5, 16>(tb, pv);
    case 106: return run_pair<N, 6, 16>(tb, pv);
    case 107: return run_pair<N, 7, 16>(tb, pv);
    case 108: return run_pair<N, 8, 16>(tb, pv);
    case 110: return run_pair<N, 10, 16>(tb, pv);
    default: return run_pair<N, 8>(tb, pv);
    }
}

extern "C" int ph_emu(const int32_t *lengths, const uint8_t *reads, const uint8_t *haps, int use_double, int rpl, double *out)
{
    const HostTables<float> &hf = tables_f();
    const HostTables<double> &hd = tables_d();
    const Tables<float> tf{hf.ph2pr.data(), hf.mm.data(), hf.initial};
    const Tables<double> td{hd.ph2pr.data(), hd.mm.data(), hd.initial};
    const int32_t R = lengths[0];
    const int32_t *rl = lengths + 1;
    const int32_t H = lengths[1 + R];
    const int32_t *hl = lengths + 2 + R;
    const uint8_t *rp = reads;
    for (int32_t r = 0; r < R; r++) {
        const int n = rl[r];
        PairView pv;
        pv.rl = n; pv.rs = rp; pv.q = rp + n; pv.ins = rp + 2 * (size_t)n; pv.del = rp + 3 * (size_t)n; pv.gcp = rp + 4 * (size_t)n;
        rp += 5 * (size_t)n;
        const uint8_t *hp = haps;
        for (int32_t h = 0; h < H; h++) {
            /* the kernel's haplotype buffer is padded on both sides (ph_hap_byte): give the emulator a padded copy */
            std::vector<uint8_t> padded((size_t)hl[h] + 2 * MGLPH_HAP_PAD, 0);
            std::copy(hp, hp + hl[h], padded.begin() + MGLPH_HAP_PAD);
            pv.hap = padded.data() + MGLPH_HAP_PAD; pv.hl = hl[h];
            double res;
            bool redo = use_double != 0;
            if (!redo) {
                const float f = run_any<float>(tf, pv, rpl);
                if ((double)f < MGLPH_MIN_ACCEPTED) redo = true;
                else res = std::log10((double)f) - (double)hf.log10_initial;
            }
            if (redo) res = std::log10(run_any<double>(td, pv, rpl)) - hd.log10_initial;
            out[(size_t)r * H + h] = res;
            hp += hl[h];
        }
    }
    return 0;
}
