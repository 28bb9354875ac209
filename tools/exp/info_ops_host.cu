// Host-side check of InfoOps (scan_elems.cuh): composition in tree order against the step-by-step recursion, and the
// transitions it yields against the state-space model the naturals were built from.   nvcc -o /tmp/info_ops_host ...
#include <cstdio>
#include <cmath>
#include <cstdlib>
#include <vector>
#include "../../vi_diffusion_processes_b200/csrc/scan_elems.cuh"
using namespace vdp;
static double rnd() { return 2.0 * rand() / RAND_MAX - 1.0; }
template <int D> static double mdiff(const Mat<D>& a, const Mat<D>& b) { double m = 0; for (int i = 0; i < D * D; ++i) m = fmax(m, fabs(a.a[i] - b.a[i])); return m; }
template <int D> static double vdiff(const Vec<D>& a, const Vec<D>& b) { double m = 0; for (int i = 0; i < D; ++i) m = fmax(m, fabs(a.v[i] - b.v[i])); return m; }
template <int D> int run() {
    const int N = 37;
    std::vector<Mat<D>> A(N), cQ(N), Q(N), thd(N + 1), ths(N);
    std::vector<Vec<D>> b(N), th1(N + 1);
    Mat<D> cP = mat_zero<D>(); Vec<D> mu0;
    for (int i = 0; i < D; ++i) { mu0.v[i] = rnd(); for (int j = 0; j <= i; ++j) cP(i, j) = (i == j ? 0.7 : 0.0) + 0.1 * rnd(); }
    for (int k = 0; k < N; ++k) {
        cQ[k] = mat_zero<D>();
        for (int i = 0; i < D; ++i) { b[k].v[i] = 0.3 * rnd(); for (int j = 0; j < D; ++j) { A[k](i, j) = (i == j ? 0.6 : 0.0) + 0.1 * rnd(); if (j <= i) cQ[k](i, j) = (i == j ? 0.5 : 0.0) + 0.1 * rnd(); } }
        Q[k] = mmT(cQ[k], cQ[k]);
    }
    for (int k = 0; k <= N; ++k) { thd[k] = mat_zero<D>(); th1[k] = vec_zero<D>(); }
    { const Mat<D> Pi = chol_inverse(cP); thd[0] = -0.5 * Pi; th1[0] = mv(Pi, mu0); }
    for (int k = 0; k < N; ++k) {
        const Mat<D> Qi = chol_inverse(cQ[k]);
        thd[k + 1] = thd[k + 1] + (-0.5 * Qi);
        thd[k] = thd[k] + (-0.5 * Tmm(A[k], mm(Qi, A[k])));
        ths[k] = mm(Qi, A[k]);
        th1[k + 1] = th1[k + 1] + mv(Qi, b[k]);
        th1[k] = th1[k] - Tmv(A[k], mv(Qi, b[k]));
    }
    using Ops = InfoOps<D>;
    typename Ops::State s; s.J = mirror_lower(-2.0 * thd[N]); s.h = th1[N];
    const typename Ops::State s0 = s;
    double eA = 0, eb = 0, eQ = 0;
    bool ok = true;
    for (int k = N - 1; k >= 0; --k) {
        const Mat<D> cov = chol_inverse(chol(s.J, ok));
        eA = fmax(eA, mdiff(mm(cov, ths[k]), A[k])); eb = fmax(eb, vdiff(mv(cov, s.h), b[k])); eQ = fmax(eQ, mdiff(cov, Q[k]));
        s = Ops::apply(Ops::step(ths[k], mirror_lower(-2.0 * thd[k]), th1[k]), s);
    }
    const Mat<D> P0 = chol_inverse(chol(s.J, ok));
    printf("D=%d  A %.1e  b %.1e  Q %.1e  P0 %.1e  mu0 %.1e\n", D, eA, eb, eQ, mdiff(P0, mmT(cP, cP)), vdiff(mv(P0, s.h), mu0));
    // tree order with push for runs of 3
    std::vector<typename Ops::Elem> es;
    for (int k = N - 1; k >= 0;) {
        typename Ops::Elem e = Ops::identity();
        for (int t = 0; t < 3 && k >= 0; ++t, --k) Ops::push(e, ths[k], mirror_lower(-2.0 * thd[k]), th1[k]);
        es.push_back(e);
    }
    while (es.size() > 1) {
        std::vector<typename Ops::Elem> nx;
        for (size_t i = 0; i < es.size(); i += 2) nx.push_back(i + 1 < es.size() ? Ops::combine(es[i], es[i + 1]) : es[i]);
        es = nx;
    }
    const typename Ops::State s2 = Ops::apply(es[0], s0);
    const double d = fmax(mdiff(s2.J, s.J), vdiff(s2.h, s.h));
    printf("      tree vs sequential %.1e (ok=%d)\n", d, (int)ok);
    return (d < 1e-12 && eA < 1e-12 && eb < 1e-12 && eQ < 1e-12 && ok) ? 0 : 1;
}
int main() { srand(3); int rc = run<1>() | run<2>() | run<3>(); printf(rc ? "FAIL\n" : "PASS\n"); return rc; }
