/*
 * sw_synth.cpp -- deterministic synthetic pairs for the benchmark configs of
 * BASELINE.json.  A pair is a pure function of
 * (config, seed, pair index) through a counter-based generator (splitmix64),
 * so any pair can be regenerated alone for an oracle spot check and a shard
 * [first, first + n) needs no other shard's data.
 *   config 2: ref U[300,500]; alt = 150-base window of it, mutated, 0-5 random bases each side
 *   config 3: ref U[300,1000]; alt = mutated ref re-cut to U[300,1000]
 *   config 4: ref U[5000,10000]; alt = mutated ref re-cut to U[5000,10000]
 *   config 5: 90 % shape 2, 9.9 % shape 3, 0.1 % shape 4 by hash of the index
 * Mutation: 1 % substitution, 0.2 % insertion, 0.2 % deletion per base.
 */
#include <cstdint>
#include <cstring>

#include "../../include/mgl_synth.h"

namespace {

struct Rng {
    uint64_t s;
    uint64_t next() { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
    uint32_t below(uint32_t n) { return (uint32_t)((next() >> 32) * (uint64_t)n >> 32); }
    uint32_t range(uint32_t lo, uint32_t hi) { return lo + below(hi - lo + 1); }
};

const char kBase[4] = {'A', 'C', 'G', 'T'};

inline Rng rng_for(uint64_t seed, uint64_t idx)
{
    Rng r;
    r.s = seed * 0xD1342543DE82EF95ull + idx * 0x9E3779B97F4A7C15ull + 0x2545F4914F6CDD1Dull;
    r.next();
    return r;
}

int shape_of(int config, uint64_t seed, uint64_t idx)
{
    if (config != 5) return config;
    Rng r = rng_for(seed ^ 0xC5C5C5C5ull, idx);
    const uint32_t v = r.below(1000);
    return v < 900 ? 2 : (v < 999 ? 3 : 4);
}

/* Generates pair idx.  ref/alt may be NULL (sizes only).  Returns lengths. */
void gen_pair(int config, uint64_t seed, uint64_t idx, uint8_t *ref, uint8_t *alt, uint32_t *tl_out, uint32_t *ql_out)
{
    const int shape = shape_of(config, seed, idx);
    Rng r = rng_for(seed, idx);
    uint32_t tl, target_ql, win_lo, win_len, lflank = 0, rflank = 0;
    if (shape == 2) { tl = r.range(300, 500); win_len = 150; win_lo = r.below(tl - win_len + 1); lflank = r.below(6); rflank = r.below(6); target_ql = 0; }
    else if (shape == 3) { tl = r.range(300, 1000); target_ql = r.range(300, 1000); win_lo = 0; win_len = tl; }
    else { tl = r.range(5000, 10000); target_ql = r.range(5000, 10000); win_lo = 0; win_len = tl; }

    /* ref: tl uniform bases, generated from its own stream so that it can be replayed for the window */
    Rng rr = rng_for(seed ^ 0x5EEDull, idx);
    if (ref) for (uint32_t i = 0; i < tl; i++) ref[i] = (uint8_t)kBase[rr.below(4)];

    /* alt = lflank random + mutate(ref[win_lo .. win_lo + win_len)) (+ re-cut) + rflank random */
    Rng rw = rng_for(seed ^ 0x5EEDull, idx);               /* replay of the ref stream */
    for (uint32_t i = 0; i < win_lo; i++) rw.below(4);
    uint32_t q = 0;
    auto put = [&](uint8_t c) { if (alt) alt[q] = c; q++; };
    for (uint32_t i = 0; i < lflank; i++) put((uint8_t)kBase[r.below(4)]);
    uint32_t body = 0;                                      /* mutated bases emitted so far */
    const uint32_t body_cap = shape == 2 ? 0xffffffffu : target_ql;
    for (uint32_t i = 0; i < win_len && body < body_cap; i++) {
        const uint8_t c = (uint8_t)kBase[rw.below(4)];
        const uint32_t x = r.below(1000);
        if (x < 2) continue;                                /* deletion */
        if (x < 4 && body < body_cap) { put((uint8_t)kBase[r.below(4)]); body++; }   /* insertion before the base */
        if (body >= body_cap) break;
        put(x >= 4 && x < 14 ? (uint8_t)kBase[r.below(4)] : c);                      /* 1 % substitution */
        body++;
    }
    if (shape != 2) while (body < target_ql) { put((uint8_t)kBase[r.below(4)]); body++; }   /* re-cut: extend */
    for (uint32_t i = 0; i < rflank; i++) put((uint8_t)kBase[r.below(4)]);
    if (q == 0) put('A');
    *tl_out = tl;
    *ql_out = q;
}

}  // namespace

extern "C" {

MGL_SYNTH_API int mgl_sw_synth_sizes(int config, uint64_t seed, uint64_t first, int64_t n, uint32_t *ref_len, uint32_t *alt_len)
{
    if (config < 2 || config > 5 || n < 0 || !ref_len || !alt_len) return -1;
#pragma omp parallel for schedule(static)
    for (int64_t k = 0; k < n; k++) gen_pair(config, seed, first + (uint64_t)k, nullptr, nullptr, &ref_len[k], &alt_len[k]);
    return 0;
}

MGL_SYNTH_API int mgl_sw_synth_fill(int config, uint64_t seed, uint64_t first, int64_t n, const uint64_t *ref_off,
                                 const uint64_t *alt_off, uint8_t *arena)
{
    if (config < 2 || config > 5 || n < 0 || !ref_off || !alt_off || !arena) return -1;
#pragma omp parallel for schedule(static)
    for (int64_t k = 0; k < n; k++) {
        uint32_t tl, ql;
        gen_pair(config, seed, first + (uint64_t)k, arena + ref_off[k], arena + alt_off[k], &tl, &ql);
    }
    return 0;
}

}  // extern "C"
