// SDE drifts f(x) and their Jacobians as __device__ functors, selected at compile time.
// Reference: markovflow/sde/sde.py (OU :152-163, double-well :200-211, Benes :244-255, sine :288-299,
// sqrt :332-343, MLP 1-3-1 :394-416, Van der Pol :454-469).  The reference obtains Jacobians from a
// GradientTape (:75-90, :484-499); every drift is a closed form so the analytic Jacobian is the same
// function.  LORENZ63 and OU_ND are extensions (the reference has no 3-D SDE; BASELINE config 5).
#pragma once
#include "smallmat.cuh"

namespace vdp {

enum DriftId : int {
    DRIFT_OU = 0, DRIFT_DW = 1, DRIFT_BENES = 2, DRIFT_SINE = 3, DRIFT_SQRT = 4,
    DRIFT_VDP = 5, DRIFT_MLP131 = 6, DRIFT_LORENZ63 = 7, DRIFT_OU_ND = 8
};

constexpr int VDP_MAX_DRIFT_PARAMS = 16;
struct DriftParams { double p[VDP_MAX_DRIFT_PARAMS]; };

// f(x) and J_ij = d f_i / d x_j in one call (they share sub-expressions).
template <int ID, int D> struct Drift;

template <int D> struct Drift<DRIFT_OU, D> {
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<D>& x, Vec<D>& f, Mat<D>& J) {
        J = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) { f(i) = -P.p[0] * x(i); J(i, i) = -P.p[0]; }
    }
};
template <int D> struct Drift<DRIFT_OU_ND, D> : Drift<DRIFT_OU, D> {};

template <int D> struct Drift<DRIFT_DW, D> {
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<D>& x, Vec<D>& f, Mat<D>& J) {
        J = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) {
            const double x2 = x(i) * x(i);
            f(i) = (P.p[0] * x(i)) * (P.p[1] - x2);
            J(i, i) = P.p[0] * fma(-3.0, x2, P.p[1]);
        }
    }
};
template <int D> struct Drift<DRIFT_BENES, D> {
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<D>& x, Vec<D>& f, Mat<D>& J) {
        J = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) {
            const double t = tanh(x(i));
            f(i) = P.p[0] * t;
            J(i, i) = P.p[0] * (1.0 - t * t);
        }
    }
};
template <int D> struct Drift<DRIFT_SINE, D> {
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<D>& x, Vec<D>& f, Mat<D>& J) {
        J = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) {
            double s, c;
            sincos(x(i) - P.p[0], &s, &c);
            f(i) = s;
            J(i, i) = c;
        }
    }
};
template <int D> struct Drift<DRIFT_SQRT, D> {
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<D>& x, Vec<D>& f, Mat<D>& J) {
        J = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) {
            const double r = sqrt(P.p[0] * fabs(x(i)));
            const double sgn = (x(i) > 0.0) ? 1.0 : ((x(i) < 0.0) ? -1.0 : 0.0);
            f(i) = r;
            J(i, i) = 0.5 / r * P.p[0] * sgn;
        }
    }
};
// 1-3-1 relu MLP per coordinate; p = (w1[3], b1[3], w2[3], b2).  float64 (the reference's Keras layers
// run in float32 and cast back, sde.py:411 -- documented deviation).
template <int D> struct Drift<DRIFT_MLP131, D> {
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<D>& x, Vec<D>& f, Mat<D>& J) {
        J = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) {
            double acc = 0.0, g = 0.0;
#pragma unroll
            for (int h = 0; h < 3; ++h) {
                const double pre = x(i) * P.p[h] + P.p[3 + h];
                const bool on = pre > 0.0;
                acc += (on ? pre : 0.0) * P.p[6 + h];
                g += on ? (P.p[h] * P.p[6 + h]) : 0.0;
            }
            f(i) = acc + P.p[9];
            J(i, i) = g;
        }
    }
};
// Van der Pol: f = tau [a (x1 - x1^3/3 - x2), x1 / a];  p = (a, tau).  D must be 2.
template <> struct Drift<DRIFT_VDP, 2> {
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<2>& x, Vec<2>& f, Mat<2>& J) {
        const double a = P.p[0], tau = P.p[1];
        const double x1 = x(0), x2 = x(1);
        f(0) = tau * (a * (x1 - (1.0 / 3.0) * (x1 * x1 * x1) - x2));
        f(1) = tau * ((1.0 / a) * x1);
        J(0, 0) = tau * a * (1.0 - x1 * x1);
        J(0, 1) = -tau * a;
        J(1, 0) = tau / a;
        J(1, 1) = 0.0;
    }
};
// Lorenz-63 (extension): f = [s (y - x), x (r - z) - y, x y - b z];  p = (sigma, rho, beta).  D must be 3.
template <> struct Drift<DRIFT_LORENZ63, 3> {
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<3>& x, Vec<3>& f, Mat<3>& J) {
        const double s = P.p[0], r = P.p[1], b = P.p[2];
        const double X = x(0), Y = x(1), Z = x(2);
        f(0) = s * (Y - X);
        f(1) = X * (r - Z) - Y;
        f(2) = X * Y - b * Z;
        J(0, 0) = -s;    J(0, 1) = s;    J(0, 2) = 0.0;
        J(1, 0) = r - Z; J(1, 1) = -1.0; J(1, 2) = -X;
        J(2, 0) = Y;     J(2, 1) = X;    J(2, 2) = -b;
    }
};

// Polynomial drifts.  degree = p > 0: every component of f is a polynomial of total degree <= p in x.  The integrands
// of the KL term along the Gaussian path (sde_utils.py:317-354) and of its reverse-mode adjoint are then polynomials
// of degree <= 2 p in the quadrature variable, which a Gauss-Hermite rule with p + 1 points per dimension integrates
// exactly -- as does the reference's 20-point rule: the two agree to rounding, for the value and for the gradient
// (both are the same function of (m, S)).  (p + 1)^D nodes instead of 20^D: 16 instead of 400 for Van der Pol,
// 27 / 8 instead of 8000 for Lorenz-63 / a linear drift in D = 3.  degree = 0: not a polynomial, the 20-point rule.
template <int ID> struct DriftPoly { static constexpr int degree = 0; };
template <> struct DriftPoly<DRIFT_OU> { static constexpr int degree = 1; };
template <> struct DriftPoly<DRIFT_OU_ND> { static constexpr int degree = 1; };
template <> struct DriftPoly<DRIFT_DW> { static constexpr int degree = 3; };
template <> struct DriftPoly<DRIFT_VDP> { static constexpr int degree = 3; };
template <> struct DriftPoly<DRIFT_LORENZ63> { static constexpr int degree = 2; };
// points per dimension of the exact rule (20 where there is none)
template <int ID> struct ExactOrder { static constexpr int H = DriftPoly<ID>::degree > 0 ? DriftPoly<ID>::degree + 1 : 20; };

// D = 1 quadrature loops (kl_kernels.cuh): residual of one Euler step against a linear transition and its slope,
//   d(x) = x + dt f(x) - (A x + b),     j(x) = d'(x) = 1 + dt f'(x) - A,
// from per-transition constants prepared once outside the node loop.  The generic form calls the drift functor;
// polynomial drifts collapse to a Horner form (DW: d = x (k1 - a x^2) - b, j = k1 - 3 a x^2 with a = dt s,
// k1 = 1 + a c - A: 4 float64 instructions per node instead of 9).
// kCubic: d is a polynomial of degree <= 3 in x; taylor() returns its coefficients about m,
// d(m + e) = c0 + c1 e + c2 e^2 + c3 e^3, for the closed-form Gaussian expectations (CubicMoments).
template <int ID> struct DriftStep1 {
    static constexpr bool kCubic = false;
    struct K { double omA, nb, dt; };
    static __device__ __forceinline__ K prepare(const DriftParams&, double dt, double A, double b) {
        K k;
        k.omA = 1.0 - A; k.nb = -b; k.dt = dt;
        return k;
    }
    static __device__ __forceinline__ void eval(const DriftParams& P, const K& k, double X, double& d, double& j) {
        Vec<1> x, f;
        Mat<1> J;
        x(0) = X;
        Drift<ID, 1>::eval(P, x, f, J);
        d = fma(k.dt, f(0), fma(k.omA, X, k.nb));
        j = fma(k.dt, J(0, 0), k.omA);
    }
};
template <> struct DriftStep1<DRIFT_DW> {
    static constexpr bool kCubic = true;
    struct K { double k1, na, na3, nb; };
    static __device__ __forceinline__ void taylor(const K& k, double m, double& c0, double& c1, double& c2,
                                                  double& c3) {
        const double m2 = m * m;
        c0 = fma(m, fma(k.na, m2, k.k1), k.nb);
        c1 = fma(k.na3, m2, k.k1);
        c2 = k.na3 * m;
        c3 = k.na;
    }
    static __device__ __forceinline__ K prepare(const DriftParams& P, double dt, double A, double b) {
        K k;
        const double a = dt * P.p[0];
        k.k1 = fma(a, P.p[1], 1.0 - A); k.na = -a; k.na3 = -3.0 * a; k.nb = -b;
        return k;
    }
    static __device__ __forceinline__ void eval(const DriftParams&, const K& k, double X, double& d, double& j) {
        const double x2 = X * X;
        d = fma(X, fma(k.na, x2, k.k1), k.nb);
        j = fma(k.na3, x2, k.k1);
    }
};
template <> struct DriftStep1<DRIFT_OU> {
    static constexpr bool kCubic = true;
    struct K { double k1, nb; };
    static __device__ __forceinline__ void taylor(const K& k, double m, double& c0, double& c1, double& c2,
                                                  double& c3) {
        c0 = fma(k.k1, m, k.nb);
        c1 = k.k1;
        c2 = 0.0;
        c3 = 0.0;
    }
    static __device__ __forceinline__ K prepare(const DriftParams& P, double dt, double A, double b) {
        K k;
        k.k1 = fma(-dt, P.p[0], 1.0 - A); k.nb = -b;
        return k;
    }
    static __device__ __forceinline__ void eval(const DriftParams&, const K& k, double X, double& d, double& j) {
        d = fma(k.k1, X, k.nb);
        j = k.k1;
    }
};
template <> struct DriftStep1<DRIFT_OU_ND> : DriftStep1<DRIFT_OU> {};

// Expectations over x = m + e, e ~ N(0, S), of a cubic d(x) = c0 + c1 e + c2 e^2 + c3 e^3, its slope j = d' and their
// products, from the central moments E e^2 = S, E e^4 = 3 S^2, E e^6 = 15 S^3.  An H-point Gauss-Hermite rule
// integrates polynomials up to degree 2H - 1 exactly, so for a polynomial drift these are the values the reference's
// 20-point rule (sde_utils.py:317-330) returns, up to rounding; expanding about m keeps the cancellation of the
// quadrature sums (d is the residual of a linearisation about m) and no more.
struct CubicMoments {
    double Ed, Ed2, Ede, Ejd, Ejde;   // E d, E d^2, E d e, E j d, E j d e
    template <bool GRAD>
    __device__ __forceinline__ void eval(double c0, double c1, double c2, double c3, double S) {
        const double E4 = 3.0 * (S * S), E6 = 5.0 * (E4 * S);
        const double c11 = c1 * c1, c02 = c0 * c2, c13 = c1 * c3, c22 = c2 * c2, c33 = c3 * c3;
        Ed2 = fma(c0, c0, fma(fma(2.0, c02, c11), S, fma(fma(2.0, c13, c22), E4, c33 * E6)));
        if constexpr (GRAD) {
            Ed = fma(c2, S, c0);
            Ede = fma(c1, S, c3 * E4);
            Ejd = fma(c0, c1, fma(3.0 * fma(c1, c2, c0 * c3), S, 5.0 * ((c2 * c3) * E4)));
            Ejde = fma(fma(2.0, c02, c11), S, fma(2.0 * fma(2.0, c13, c22), E4, 3.0 * (c33 * E6)));
        }
    }
};

// Derivatives of f and J with respect to every entry of the parameter block: what a tape watching
// prior_sde.trainable_variables propagates through drift / gradient_drift (prior-parameter learning,
// variational_cvi_sde.py:495-518, vi_sde.py:457-470).  df[p](i) = d f_i / d P.p[p], dJ[p](i, j) = d J_ij / d P.p[p].
template <int ID, int D> struct DriftParamGrad;

template <int D> struct DriftParamGrad<DRIFT_OU, D> {
    static constexpr int NP = 1;
    static __device__ __forceinline__ void eval(const DriftParams&, const Vec<D>& x, Vec<D>* df, Mat<D>* dJ) {
        dJ[0] = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) { df[0](i) = -x(i); dJ[0](i, i) = -1.0; }
    }
};
template <int D> struct DriftParamGrad<DRIFT_OU_ND, D> : DriftParamGrad<DRIFT_OU, D> {};

template <int D> struct DriftParamGrad<DRIFT_DW, D> {
    static constexpr int NP = 2;
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<D>& x, Vec<D>* df, Mat<D>* dJ) {
        dJ[0] = mat_zero<D>();
        dJ[1] = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) {
            const double x2 = x(i) * x(i);
            df[0](i) = x(i) * (P.p[1] - x2);
            df[1](i) = P.p[0] * x(i);
            dJ[0](i, i) = fma(-3.0, x2, P.p[1]);
            dJ[1](i, i) = P.p[0];
        }
    }
};
template <int D> struct DriftParamGrad<DRIFT_BENES, D> {
    static constexpr int NP = 1;
    static __device__ __forceinline__ void eval(const DriftParams&, const Vec<D>& x, Vec<D>* df, Mat<D>* dJ) {
        dJ[0] = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) {
            const double t = tanh(x(i));
            df[0](i) = t;
            dJ[0](i, i) = 1.0 - t * t;
        }
    }
};
template <int D> struct DriftParamGrad<DRIFT_SINE, D> {
    static constexpr int NP = 1;
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<D>& x, Vec<D>* df, Mat<D>* dJ) {
        dJ[0] = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) {
            double s, c;
            sincos(x(i) - P.p[0], &s, &c);
            df[0](i) = -c;
            dJ[0](i, i) = s;
        }
    }
};
template <int D> struct DriftParamGrad<DRIFT_SQRT, D> {
    static constexpr int NP = 1;
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<D>& x, Vec<D>* df, Mat<D>* dJ) {
        dJ[0] = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) {
            const double r = sqrt(P.p[0] * fabs(x(i)));
            const double sgn = (x(i) > 0.0) ? 1.0 : ((x(i) < 0.0) ? -1.0 : 0.0);
            df[0](i) = 0.5 * fabs(x(i)) / r;
            dJ[0](i, i) = 0.25 * sgn / r;
        }
    }
};
// p = (w1[3], b1[3], w2[3], b2); relu'(0) = 0 as in TensorFlow
template <int D> struct DriftParamGrad<DRIFT_MLP131, D> {
    static constexpr int NP = 10;
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<D>& x, Vec<D>* df, Mat<D>* dJ) {
#pragma unroll
        for (int p = 0; p < NP; ++p) dJ[p] = mat_zero<D>();
#pragma unroll
        for (int i = 0; i < D; ++i) {
#pragma unroll
            for (int h = 0; h < 3; ++h) {
                const double pre = x(i) * P.p[h] + P.p[3 + h];
                const double on = (pre > 0.0) ? 1.0 : 0.0;
                df[h](i) = on * x(i) * P.p[6 + h];
                dJ[h](i, i) = on * P.p[6 + h];
                df[3 + h](i) = on * P.p[6 + h];
                df[6 + h](i) = on * pre;
                dJ[6 + h](i, i) = on * P.p[h];
            }
            df[9](i) = 1.0;
        }
    }
};
template <> struct DriftParamGrad<DRIFT_VDP, 2> {
    static constexpr int NP = 2;
    static __device__ __forceinline__ void eval(const DriftParams& P, const Vec<2>& x, Vec<2>* df, Mat<2>* dJ) {
        const double a = P.p[0], tau = P.p[1];
        const double x1 = x(0), x2 = x(1);
        const double g = x1 - (1.0 / 3.0) * (x1 * x1 * x1) - x2;
        const double ia = 1.0 / a;
        df[0](0) = tau * g;            df[0](1) = -tau * x1 * ia * ia;
        df[1](0) = a * g;              df[1](1) = x1 * ia;
        dJ[0](0, 0) = tau * (1.0 - x1 * x1); dJ[0](0, 1) = -tau; dJ[0](1, 0) = -tau * ia * ia; dJ[0](1, 1) = 0.0;
        dJ[1](0, 0) = a * (1.0 - x1 * x1);   dJ[1](0, 1) = -a;   dJ[1](1, 0) = ia;             dJ[1](1, 1) = 0.0;
    }
};
template <> struct DriftParamGrad<DRIFT_LORENZ63, 3> {
    static constexpr int NP = 3;
    static __device__ __forceinline__ void eval(const DriftParams&, const Vec<3>& x, Vec<3>* df, Mat<3>* dJ) {
#pragma unroll
        for (int p = 0; p < NP; ++p) { df[p] = vec_zero<3>(); dJ[p] = mat_zero<3>(); }
        df[0](0) = x(1) - x(0); dJ[0](0, 0) = -1.0; dJ[0](0, 1) = 1.0;
        df[1](1) = x(0);        dJ[1](1, 0) = 1.0;
        df[2](2) = -x(2);       dJ[2](2, 2) = -1.0;
    }
};
// number of entries of the parameter block a drift reads (0 for an unknown id)
__host__ __device__ inline int drift_num_params(int id) {
    switch (id) {
        case DRIFT_OU: case DRIFT_OU_ND: case DRIFT_BENES: case DRIFT_SINE: case DRIFT_SQRT: return 1;
        case DRIFT_DW: case DRIFT_VDP: return 2;
        case DRIFT_LORENZ63: return 3;
        case DRIFT_MLP131: return 10;
        default: return 0;
    }
}

// Which (drift, D) pairs exist: the element-wise drifts for any D, VDP only D=2, LORENZ63 only D=3.
template <int ID, int D> struct DriftValid { static constexpr bool value = true; };
template <int D> struct DriftValid<DRIFT_VDP, D> { static constexpr bool value = (D == 2); };
template <int D> struct DriftValid<DRIFT_LORENZ63, D> { static constexpr bool value = (D == 3); };

// True when the reference's gradient_drift returns the full [D, D] Jacobian (batch_jacobian subclasses);
// otherwise it returns d(sum_i f_i)/dx_j, a [D] vector (sde.py:86-90).
__host__ __device__ inline bool drift_full_jacobian(int id) {
    return id == DRIFT_VDP || id == DRIFT_LORENZ63 || id == DRIFT_OU_ND;
}

}  // namespace vdp
