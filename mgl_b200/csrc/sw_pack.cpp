/*
 * sw_pack.cpp -- host side of the 2-bit transport form (sw_pack.h).  Runs on the planner threads of
 * Engine::align_batch, one chunk's byte range at a time, into pinned staging.
 */
#include "sw_pack.h"

#include <cstring>
#include <immintrin.h>

namespace mglsw {

namespace {

struct CodeTable {
    uint8_t t[256];
    CodeTable() { std::memset(t, 0xff, sizeof t); t['A'] = 0; t['C'] = 1; t['T'] = 2; t['G'] = 3; }
};
const CodeTable g_codes;

/* bases [i, n) one at a time; dst bytes from i / 4 on (i is a multiple of 4) */
bool pack2_tail(const uint8_t *src, size_t i, size_t n, uint8_t *dst)
{
    const size_t out_end = pack2_bytes(n);
    size_t o = i / 4;
    unsigned bad = 0;
    for (; i + 4 <= n; i += 4, o++) {
        const unsigned a = g_codes.t[src[i]], b = g_codes.t[src[i + 1]], c = g_codes.t[src[i + 2]], d = g_codes.t[src[i + 3]];
        bad |= a | b | c | d;
        dst[o] = (uint8_t)((a & 3) | (b & 3) << 2 | (c & 3) << 4 | (d & 3) << 6);
    }
    if (i < n) {
        unsigned v = 0;
        for (unsigned k = 0; i < n; i++, k++) { const unsigned a = g_codes.t[src[i]]; bad |= a; v |= (a & 3) << (2 * k); }
        dst[o++] = (uint8_t)v;
    }
    for (; o < out_end; o++) dst[o] = 0;
    return (bad & 0x80) == 0;
}

/* whole 128-base groups of src[0, n); returns the bases consumed through `done` */
__attribute__((target("avx2")))
bool pack2_avx2(const uint8_t *src, size_t n, uint8_t *dst, size_t &done)
{
    const __m256i three = _mm256_set1_epi8(3);
    const __m256i bases = _mm256_setr_epi8('A', 'C', 'T', 'G', 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0,
                                           'A', 'C', 'T', 'G', 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0);
    const __m256i w1 = _mm256_set1_epi16(0x0401);        /* c0 + 4 c1 per byte pair                      */
    const __m256i w2 = _mm256_set1_epi32(0x00100001);    /* (c0 + 4 c1) + 16 (c2 + 4 c3): one byte value per 32-bit lane */
    const __m256i order = _mm256_setr_epi32(0, 4, 1, 5, 2, 6, 3, 7);   /* undoes the per-128-bit-lane interleave of the two packs */
    size_t i = 0;
    const size_t groups = n / 128;
    for (size_t g0 = 0; g0 < groups;) {
        const size_t g1 = groups - g0 < 32 ? groups : g0 + 32;          /* validity is looked at every 4 KiB */
        __m256i ok = _mm256_set1_epi8(-1);
        for (; g0 < g1; g0++, i += 128) {
            __m256i q[4];
            for (int k = 0; k < 4; k++) {
                const __m256i v = _mm256_loadu_si256((const __m256i *)(src + i + 32 * k));
                const __m256i code = _mm256_and_si256(_mm256_srli_epi16(v, 1), three);
                ok = _mm256_and_si256(ok, _mm256_cmpeq_epi8(_mm256_shuffle_epi8(bases, code), v));
                q[k] = _mm256_madd_epi16(_mm256_maddubs_epi16(code, w1), w2);
            }
            const __m256i b = _mm256_packus_epi16(_mm256_packus_epi32(q[0], q[1]), _mm256_packus_epi32(q[2], q[3]));
            _mm256_storeu_si256((__m256i *)(dst + i / 4), _mm256_permutevar8x32_epi32(b, order));
        }
        if (_mm256_movemask_epi8(ok) != -1) { done = i; return false; }
    }
    done = i;
    return true;
}

}  // namespace

bool pack2_scalar(const uint8_t *src, size_t n, uint8_t *dst) { return pack2_tail(src, 0, n, dst); }

bool pack2(const uint8_t *src, size_t n, uint8_t *dst)
{
    static const bool have_avx2 = __builtin_cpu_supports("avx2");
    size_t done = 0;
    if (have_avx2 && !pack2_avx2(src, n, dst, done)) return false;
    return pack2_tail(src, done, n, dst);
}

}  // namespace mglsw
