/*
 * oracle/ref/shim.h -- forced-include (-include) header used ONLY to compile
 * the reference's own sources, unmodified and in place, into oracle/_ref/.
 * TEST INFRASTRUCTURE.  It repairs two defects that make the reference's AVX2
 * path abort on this glibc without touching reference source (SURVEY.md
 * section 0 fact 3, Appendix C):
 *   - sw_avx.cpp:17-42 calls aligned_alloc(size, alignment), arguments swapped
 *     relative to C11;
 *   - sw_avx.cpp:161-183 uses aligned 256-bit loads/stores on addresses that
 *     advance 4 bytes per step.
 */
#ifndef MGL_ORACLE_REF_SHIM_H
#define MGL_ORACLE_REF_SHIM_H
#include <stdlib.h>
#include <x86intrin.h>
static inline void *mgl_ref_aligned_alloc(size_t size, size_t alignment)
{
    void *p = 0;
    size_t rounded = (size + alignment - 1) / alignment * alignment;
    if (posix_memalign(&p, alignment, rounded ? rounded : alignment)) return 0;
    return p;
}
#define aligned_alloc(sz, al) mgl_ref_aligned_alloc((sz), (al))
#define _mm256_load_si256(p) _mm256_loadu_si256(p)
#define _mm256_store_si256(p, v) _mm256_storeu_si256((p), (v))
#endif
