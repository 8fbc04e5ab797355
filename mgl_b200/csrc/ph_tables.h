/*
 * ph_tables.h -- the quality tables of the PairHMM in the reference's arithmetic (Context.h): built once on the
 * host in float and in double, uploaded by the engine.  Qualities reach the kernels masked to 0..127
 * (compute_prob_scalar.cc:73-83), so the triangular matchToMatchProb table stops at 127 (8 256 entries).
 */
#ifndef MGL_B200_PH_TABLES_H
#define MGL_B200_PH_TABLES_H
#include <vector>

namespace mglph {

#define MGLPH_MM_ENTRIES ((128 * 129) / 2)
#define MGLPH_MIN_ACCEPTED 1e-28f          /* pairhmm_common.h:31 */

template <class N>
struct HostTables {
    std::vector<N> ph2pr;      /* 128 */
    std::vector<N> mm;         /* MGLPH_MM_ENTRIES */
    N initial;                 /* INITIAL_CONSTANT */
    N log10_initial;           /* LOG10_INITIAL_CONSTANT */
};

const HostTables<float> &tables_f();
const HostTables<double> &tables_d();

}  // namespace mglph
#endif
