/*
 * oracle/ref/tbb_stub/tbb/tbb.h -- TEST INFRASTRUCTURE.
 * Intel TBB is not in this image.  The reference's PairHMM uses exactly two things from it
 * (com_microsoft_mgl_pairhmm_MicrosoftPairHmm.cc:131 tbb::parallel_for over a blocked_range of reads;
 * compute_prob_avx{f,d}.cc scalable_aligned_malloc / scalable_free).  This stand-in runs the range on
 * the calling thread (reads are independent, so results are unchanged; oracle/ref/pairhmm_driver.cpp
 * parallelises over reads outside) and maps the allocator to posix_memalign.
 */
#ifndef MGL_ORACLE_TBB_STUB_H
#define MGL_ORACLE_TBB_STUB_H
#include <cstddef>
#include <cstdio>    /* the real tbb.h brings these in; the reference relies on it */
#include <cstdlib>
#include <cstring>
namespace tbb {
template <class T>
class blocked_range {
public:
    blocked_range(T b, T e) : b_(b), e_(e) {}
    T begin() const { return b_; }
    T end() const { return e_; }
private:
    T b_, e_;
};
template <class Range, class Body>
void parallel_for(const Range &r, const Body &body) { body(r); }
}  // namespace tbb
static inline void *scalable_aligned_malloc(size_t size, size_t align)
{
    void *p = nullptr;
    if (align < sizeof(void *)) align = sizeof(void *);
    return posix_memalign(&p, align, size ? size : align) == 0 ? p : nullptr;
}
static inline void scalable_free(void *p) { free(p); }
#endif
