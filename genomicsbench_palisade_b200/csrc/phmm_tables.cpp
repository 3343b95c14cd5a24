// Host-side generation of the pair-HMM probability tables.
//
// The reference builds these once per process in Context<float> / Context<double>
// (tools/GKL/src/main/native/pairhmm/Context.h): a Jacobian-logarithm table (:42-47), the
// match-to-match transition table indexed by (max(ins,del), min(ins,del)) (:50-61) and the
// Phred -> probability table ph2pr (:105-107 double via pow, :145-147 float via powf).  The
// values depend on the host libm, and the float tables are NOT the rounded double tables (the
// Jacobian approximation runs in the table's own type, :67-90), so both are generated here with
// the host libm -- exactly what the CPU reference on the same machine would hold -- and uploaded;
// the GPU never recomputes them.  Must be compiled without FP contraction.
#include <cmath>

#include "phmm_host.h"

namespace phmm {
namespace {

const int kMaxQual = 254;                               // Context.h:7
const int kJacobianSize = 80001;                        // (int)(8.0 / 0.0001) + 1, Context.h:8-11
const double kJacobianStep = 0.0001;
const double kJacobianTol = 8.0;

template <typename T>
struct TableBuilder {
    std::vector<T> jac;

    TableBuilder() : jac(kJacobianSize)
    {
        for (int k = 0; k < kJacobianSize; ++k)
            jac[k] = (T)std::log10(1.0 + std::pow(10.0, -((double)k) * kJacobianStep));
    }

    // log10(10^x + 10^y) by table lookup on |x - y|, evaluated in T (Context.h:67-90).
    T log10_sum(T lo, T hi) const
    {
        if (lo > hi) std::swap(lo, hi);
        const T diff = hi - lo;
        if (diff >= (T)kJacobianTol) return hi;
        const T scaled = (T)(diff * (T)(1.0 / kJacobianStep));
        const int idx = (scaled > (T)0.0) ? (int)(scaled + (T)0.5) : (int)(scaled - (T)0.5);
        return hi + jac[idx];
    }

    // 1 - P(open a gap) for insertion/deletion open qualities (a, b), a >= b (Context.h:50-61).
    T match_to_match(int a, int b) const
    {
        const double inv_ln10 = 1.0 / std::log(10);
        const double s = log10_sum((T)(-0.1 * a), (T)(-0.1 * b));
        const double l = std::log1p(-std::min(1.0, std::pow(10, s))) * inv_ln10;
        return (T)std::pow(10, l);
    }
};

HostTables make_tables()
{
    HostTables t;
    TableBuilder<float> bf;
    TableBuilder<double> bd;
    t.mm_f.resize(MM_REACHABLE);
    t.mm_d.resize(MM_REACHABLE);
    // kernel-side qualities are masked with 127 (avx-pairhmm-template.h:110-112), so only the
    // first 128 rows of the triangular table are reachable
    for (int a = 0, off = 0; a <= 127 && a <= kMaxQual; off += ++a)
        for (int b = 0; b <= a; ++b) {
            t.mm_f[off + b] = bf.match_to_match(a, b);
            t.mm_d[off + b] = bd.match_to_match(a, b);
        }
    for (int x = 0; x < 128; ++x) {
        t.ph2pr_d[x] = std::pow(10.0, -((double)x) / 10.0);
        t.ph2pr_f[x] = powf(10.f, -((float)x) / 10.f);
    }
    t.init_d = std::ldexp(1.0, 1020);                   // Context.h:109
    t.log10_init_d = std::log10(t.init_d);
    t.init_f = ldexpf(1.f, 120);                        // Context.h:149
    t.log10_init_f = log10f(t.init_f);
    return t;
}

}  // namespace

const HostTables &host_tables()
{
    static const HostTables t = make_tables();
    return t;
}

}  // namespace phmm
