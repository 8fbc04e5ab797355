/*
 * mgl_synth.h -- C ABI of libmgl_synth.so: deterministic synthetic pairs of BASELINE.json configs 2..5
 * (mgl_b200/csrc/sw_synth.cpp).  Benchmark and test input only; its own small library (plain C++, no CUDA) so that
 * a process timing the REFERENCE never maps the product library.  No counterpart in the reference, which ships no
 * benchmark inputs; shapes and seeds are SURVEY.md 8d's.
 */
#ifndef MGL_SYNTH_H
#define MGL_SYNTH_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGL_SYNTH_API __attribute__((visibility("default")))

/* a pair is a pure function of (config, seed, pair index): sizes first, then bytes; 0 or -1 on a bad argument */
MGL_SYNTH_API int mgl_sw_synth_sizes(int config, uint64_t seed, uint64_t first, int64_t n, uint32_t *ref_len, uint32_t *alt_len);
MGL_SYNTH_API int mgl_sw_synth_fill(int config, uint64_t seed, uint64_t first, int64_t n, const uint64_t *ref_off,
                                    const uint64_t *alt_off, uint8_t *arena);

#ifdef __cplusplus
}
#endif
#endif
