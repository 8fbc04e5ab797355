/*
 * ph_tables.cpp -- see ph_tables.h.  Follows Context.h: the "Jacobian log" table (:41-45), matchToMatchProb through
 * approximateLog10SumLog10 (:47-79), ph2pr and the initial constants of the two constructors (:95-110, :131-146).
 * The intermediate types matter for the last digits (e.g. -0.1f * i is a float product in both precisions), so they
 * are kept as the reference has them.
 */
#include "ph_tables.h"

#include <cmath>

namespace mglph {
namespace {

constexpr int MAX_QUAL = 254;
constexpr double JAC_TOL = 8.0, JAC_STEP = 0.0001, JAC_INV_STEP = 1.0 / JAC_STEP;
constexpr int JAC_SIZE = (int)(JAC_TOL / JAC_STEP) + 1;

template <class N>
struct Builder {
    std::vector<N> jac;
    static int fast_round(N d) { return d > (N)0.0 ? (int)(d + (N)0.5) : (int)(d - (N)0.5); }
    N approx_log10_sum(N small, N big) const
    {
        if (small > big) { N t = big; big = small; small = t; }
        if (std::isinf(small) || std::isinf(big)) return big;
        const N diff = big - small;
        if (diff >= (N)JAC_TOL) return big;
        return big + jac[(size_t)fast_round((N)(diff * (N)JAC_INV_STEP))];
    }
    void build(HostTables<N> &t)
    {
        jac.resize(JAC_SIZE);
        for (int k = 0; k < JAC_SIZE; k++) jac[(size_t)k] = (N)std::log10(1.0 + std::pow(10.0, -((double)k) * JAC_STEP));
        const double inv_ln10 = 1.0 / std::log(10);
        t.mm.resize(MGLPH_MM_ENTRIES);
        for (int i = 0, offset = 0; i <= MAX_QUAL && i < 128; offset += ++i)
            for (int j = 0; j <= i; j++) {
                const double log10_sum = approx_log10_sum(-0.1f * i, -0.1f * j);
                const double m2m_log10 = std::log1p(-std::fmin(1.0, std::pow(10, log10_sum))) * inv_ln10;
                t.mm[(size_t)(offset + j)] = (N)std::pow(10, m2m_log10);
            }
    }
};

}  // namespace

const HostTables<float> &tables_f()
{
    static const HostTables<float> t = [] {
        HostTables<float> x;
        Builder<float>().build(x);
        x.ph2pr.resize(128);
        for (int q = 0; q < 128; q++) x.ph2pr[(size_t)q] = powf(10.f, -((float)q) / 10.f);
        x.initial = ldexpf(1.f, 120);
        x.log10_initial = log10f(x.initial);
        return x;
    }();
    return t;
}

const HostTables<double> &tables_d()
{
    static const HostTables<double> t = [] {
        HostTables<double> x;
        Builder<double>().build(x);
        x.ph2pr.resize(128);
        for (int q = 0; q < 128; q++) x.ph2pr[(size_t)q] = std::pow(10.0, -((double)q) / 10.0);
        x.initial = std::ldexp(1.0, 1020);
        x.log10_initial = std::log10(x.initial);
        return x;
    }();
    return t;
}

}  // namespace mglph
