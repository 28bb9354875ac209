// D x D (D <= 3) dense blocks held in registers.  Every routine is the scalar restatement of the
// tf.linalg.* call the reference makes on a [.., D, D] block (cholesky, triangular_solve,
// cholesky_solve, solve, matmul); loops are fully unrolled so nothing spills to local memory.
// The operation order inside each routine follows oracle/gauss_markov_np.py (tri_solve, chol_solve)
// so that the device path and the oracle round the same way wherever that is cheap.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#include "fastmath.cuh"

namespace vdp {

template <int D>
struct Mat {
    double a[D * D];
    __host__ __device__ __forceinline__ double& operator()(int i, int j) { return a[i * D + j]; }
    __host__ __device__ __forceinline__ const double& operator()(int i, int j) const { return a[i * D + j]; }
};

template <int D>
struct Vec {
    double v[D];
    __host__ __device__ __forceinline__ double& operator()(int i) { return v[i]; }
    __host__ __device__ __forceinline__ const double& operator()(int i) const { return v[i]; }
};

#define VDP_HD __host__ __device__ __forceinline__

template <int D> VDP_HD Mat<D> mat_load(const double* __restrict__ p) {
    Mat<D> m;
#pragma unroll
    for (int i = 0; i < D * D; ++i) m.a[i] = p[i];
    return m;
}
template <int D> VDP_HD void mat_store(double* __restrict__ p, const Mat<D>& m) {
#pragma unroll
    for (int i = 0; i < D * D; ++i) p[i] = m.a[i];
}
template <int D> VDP_HD Vec<D> vec_load(const double* __restrict__ p) {
    Vec<D> x;
#pragma unroll
    for (int i = 0; i < D; ++i) x.v[i] = p[i];
    return x;
}
template <int D> VDP_HD void vec_store(double* __restrict__ p, const Vec<D>& x) {
#pragma unroll
    for (int i = 0; i < D; ++i) p[i] = x.v[i];
}
template <int D> VDP_HD Mat<D> mat_zero() {
    Mat<D> m;
#pragma unroll
    for (int i = 0; i < D * D; ++i) m.a[i] = 0.0;
    return m;
}
template <int D> VDP_HD Vec<D> vec_zero() {
    Vec<D> x;
#pragma unroll
    for (int i = 0; i < D; ++i) x.v[i] = 0.0;
    return x;
}
template <int D> VDP_HD Mat<D> mat_eye() {
    Mat<D> m = mat_zero<D>();
#pragma unroll
    for (int i = 0; i < D; ++i) m(i, i) = 1.0;
    return m;
}
template <int D> VDP_HD Mat<D> transpose(const Mat<D>& A) {
    Mat<D> T;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j) T(i, j) = A(j, i);
    return T;
}
template <int D> VDP_HD Mat<D> operator+(const Mat<D>& A, const Mat<D>& B) {
    Mat<D> C;
#pragma unroll
    for (int i = 0; i < D * D; ++i) C.a[i] = A.a[i] + B.a[i];
    return C;
}
template <int D> VDP_HD Mat<D> operator-(const Mat<D>& A, const Mat<D>& B) {
    Mat<D> C;
#pragma unroll
    for (int i = 0; i < D * D; ++i) C.a[i] = A.a[i] - B.a[i];
    return C;
}
template <int D> VDP_HD Mat<D> operator*(double s, const Mat<D>& A) {
    Mat<D> C;
#pragma unroll
    for (int i = 0; i < D * D; ++i) C.a[i] = s * A.a[i];
    return C;
}
template <int D> VDP_HD Vec<D> operator+(const Vec<D>& x, const Vec<D>& y) {
    Vec<D> z;
#pragma unroll
    for (int i = 0; i < D; ++i) z.v[i] = x.v[i] + y.v[i];
    return z;
}
template <int D> VDP_HD Vec<D> operator-(const Vec<D>& x, const Vec<D>& y) {
    Vec<D> z;
#pragma unroll
    for (int i = 0; i < D; ++i) z.v[i] = x.v[i] - y.v[i];
    return z;
}
template <int D> VDP_HD Vec<D> operator*(double s, const Vec<D>& x) {
    Vec<D> z;
#pragma unroll
    for (int i = 0; i < D; ++i) z.v[i] = s * x.v[i];
    return z;
}
// C = A B
template <int D> VDP_HD Mat<D> mm(const Mat<D>& A, const Mat<D>& B) {
    Mat<D> C;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j) {
            double s = A(i, 0) * B(0, j);
#pragma unroll
            for (int k = 1; k < D; ++k) s += A(i, k) * B(k, j);
            C(i, j) = s;
        }
    return C;
}
// C = A B^T
template <int D> VDP_HD Mat<D> mmT(const Mat<D>& A, const Mat<D>& B) {
    Mat<D> C;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j) {
            double s = A(i, 0) * B(j, 0);
#pragma unroll
            for (int k = 1; k < D; ++k) s += A(i, k) * B(j, k);
            C(i, j) = s;
        }
    return C;
}
// C = A^T B
template <int D> VDP_HD Mat<D> Tmm(const Mat<D>& A, const Mat<D>& B) {
    Mat<D> C;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j) {
            double s = A(0, i) * B(0, j);
#pragma unroll
            for (int k = 1; k < D; ++k) s += A(k, i) * B(k, j);
            C(i, j) = s;
        }
    return C;
}
// y = A x
template <int D> VDP_HD Vec<D> mv(const Mat<D>& A, const Vec<D>& x) {
    Vec<D> y;
#pragma unroll
    for (int i = 0; i < D; ++i) {
        double s = A(i, 0) * x(0);
#pragma unroll
        for (int k = 1; k < D; ++k) s += A(i, k) * x(k);
        y(i) = s;
    }
    return y;
}
// y = A^T x
template <int D> VDP_HD Vec<D> Tmv(const Mat<D>& A, const Vec<D>& x) {
    Vec<D> y;
#pragma unroll
    for (int i = 0; i < D; ++i) {
        double s = A(0, i) * x(0);
#pragma unroll
        for (int k = 1; k < D; ++k) s += A(k, i) * x(k);
        y(i) = s;
    }
    return y;
}
template <int D> VDP_HD double dot(const Vec<D>& x, const Vec<D>& y) {
    double s = x(0) * y(0);
#pragma unroll
    for (int k = 1; k < D; ++k) s += x(k) * y(k);
    return s;
}
// outer product x y^T
template <int D> VDP_HD Mat<D> outer(const Vec<D>& x, const Vec<D>& y) {
    Mat<D> C;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j) C(i, j) = x(i) * y(j);
    return C;
}
template <int D> VDP_HD Mat<D> sym(const Mat<D>& A) {
    Mat<D> C;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j) C(i, j) = 0.5 * (A(i, j) + A(j, i));
    return C;
}
// band_to_block(symmetric=True): keep the lower triangle, mirror it
template <int D> VDP_HD Mat<D> mirror_lower(const Mat<D>& A) {
    Mat<D> C;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j) C(i, j) = (j <= i) ? A(i, j) : A(j, i);
    return C;
}
template <int D> VDP_HD Mat<D> tril(const Mat<D>& A) {
    Mat<D> C;
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j) C(i, j) = (j <= i) ? A(i, j) : 0.0;
    return C;
}
template <int D> VDP_HD double trace(const Mat<D>& A) {
    double s = A(0, 0);
#pragma unroll
    for (int k = 1; k < D; ++k) s += A(k, k);
    return s;
}

// tf.linalg.cholesky: lower factor from the lower triangle.  ok=false if a pivot is <= 0 or NaN.
template <int D> VDP_HD Mat<D> chol(const Mat<D>& X, bool& ok) {
    Mat<D> L = mat_zero<D>();
#pragma unroll
    for (int j = 0; j < D; ++j) {
        double d = X(j, j);
#pragma unroll
        for (int k = 0; k < j; ++k) d -= L(j, k) * L(j, k);
        if (!(d > 0.0)) ok = false;
        double ljj, rjj;
        sqrt_rsqrt_fast(d, ljj, rjj);
        L(j, j) = ljj;
#pragma unroll
        for (int i = j + 1; i < D; ++i) {
            double s = X(i, j);
#pragma unroll
            for (int k = 0; k < j; ++k) s -= L(i, k) * L(j, k);
            L(i, j) = div_r(s, ljj, rjj);
        }
    }
    return L;
}
// tf.linalg.triangular_solve(L, B, lower=True): X = L^-1 B
template <int D> VDP_HD Mat<D> trsm(const Mat<D>& L, const Mat<D>& B) {
    Mat<D> X;
    double rd[D];
#pragma unroll
    for (int i = 0; i < D; ++i) rd[i] = rcp_fast(L(i, i));
#pragma unroll
    for (int c = 0; c < D; ++c)
#pragma unroll
        for (int i = 0; i < D; ++i) {
            double acc = B(i, c);
#pragma unroll
            for (int j = 0; j < i; ++j) acc -= L(i, j) * X(j, c);
            X(i, c) = div_r(acc, L(i, i), rd[i]);
        }
    return X;
}
// adjoint=True: X = L^-T B
template <int D> VDP_HD Mat<D> trsmT(const Mat<D>& L, const Mat<D>& B) {
    Mat<D> X;
    double rd[D];
#pragma unroll
    for (int i = 0; i < D; ++i) rd[i] = rcp_fast(L(i, i));
#pragma unroll
    for (int c = 0; c < D; ++c)
#pragma unroll
        for (int i = D - 1; i >= 0; --i) {
            double acc = B(i, c);
#pragma unroll
            for (int j = i + 1; j < D; ++j) acc -= L(j, i) * X(j, c);
            X(i, c) = div_r(acc, L(i, i), rd[i]);
        }
    return X;
}
template <int D> VDP_HD Vec<D> trsv(const Mat<D>& L, const Vec<D>& b) {
    Vec<D> x;
#pragma unroll
    for (int i = 0; i < D; ++i) {
        double acc = b(i);
#pragma unroll
        for (int j = 0; j < i; ++j) acc -= L(i, j) * x(j);
        x(i) = div_fast(acc, L(i, i));
    }
    return x;
}
template <int D> VDP_HD Vec<D> trsvT(const Mat<D>& L, const Vec<D>& b) {
    Vec<D> x;
#pragma unroll
    for (int i = D - 1; i >= 0; --i) {
        double acc = b(i);
#pragma unroll
        for (int j = i + 1; j < D; ++j) acc -= L(j, i) * x(j);
        x(i) = div_fast(acc, L(i, i));
    }
    return x;
}
// tf.linalg.cholesky_solve(L, B) = (L L^T)^-1 B
template <int D> VDP_HD Mat<D> chol_solve(const Mat<D>& L, const Mat<D>& B) { return trsmT(L, trsm(L, B)); }
template <int D> VDP_HD Vec<D> chol_solve(const Mat<D>& L, const Vec<D>& b) { return trsvT(L, trsv(L, b)); }
template <int D> VDP_HD Mat<D> chol_inverse(const Mat<D>& L) { return chol_solve(L, mat_eye<D>()); }

// tf.linalg.solve(A, B): LU with partial pivoting (what LAPACK getrf/getrs does), X = A^-1 B
template <int D> VDP_HD Mat<D> lu_solve(Mat<D> A, Mat<D> B) {
#pragma unroll
    for (int k = 0; k < D; ++k) {
        int p = k;
        double best = fabs(A(k, k));
#pragma unroll
        for (int i = k + 1; i < D; ++i) {
            const double v = fabs(A(i, k));
            if (v > best) { best = v; p = i; }
        }
#pragma unroll
        for (int i = k + 1; i < D; ++i) {
            if (i == p) {
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    double t = A(k, j); A(k, j) = A(i, j); A(i, j) = t;
                    t = B(k, j); B(k, j) = B(i, j); B(i, j) = t;
                }
            }
        }
        const double piv = A(k, k);
        const double rpiv = rcp_fast(piv);
#pragma unroll
        for (int i = k + 1; i < D; ++i) {
            const double f = div_r(A(i, k), piv, rpiv);
            A(i, k) = f;
#pragma unroll
            for (int j = k + 1; j < D; ++j) A(i, j) -= f * A(k, j);
#pragma unroll
            for (int j = 0; j < D; ++j) B(i, j) -= f * B(k, j);
        }
    }
    Mat<D> X;
    double rd[D];
#pragma unroll
    for (int i = 0; i < D; ++i) rd[i] = rcp_fast(A(i, i));
#pragma unroll
    for (int c = 0; c < D; ++c)
#pragma unroll
        for (int i = D - 1; i >= 0; --i) {
            double acc = B(i, c);
#pragma unroll
            for (int j = i + 1; j < D; ++j) acc -= A(i, j) * X(j, c);
            X(i, c) = div_r(acc, A(i, i), rd[i]);
        }
    return X;
}

// 2 sum_i log L_ii
template <int D> VDP_HD double logdet_from_chol(const Mat<D>& L) {
    double s = log(L(0, 0));
#pragma unroll
    for (int k = 1; k < D; ++k) s += log(L(k, k));
    return 2.0 * s;
}

// Reverse mode of L = chol(S) as TensorFlow registers it (Murray 2016):
// Sbar = 0.5 (P + P^T), P = L^-T Phi(L^T Lbar) L^-1, Phi = lower triangle with halved diagonal.
template <int D> VDP_HD Mat<D> chol_adjoint(const Mat<D>& L, const Mat<D>& Lbar) {
    Mat<D> mid = Tmm(L, Lbar);
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
        for (int j = 0; j < D; ++j) mid(i, j) = (j < i) ? mid(i, j) : ((j == i) ? 0.5 * mid(i, j) : 0.0);
    Mat<D> P = trsmT(L, mid);                       // L^-T mid
    P = transpose(trsmT(L, transpose(P)));          // (.) L^-1
    return sym(P);
}

}  // namespace vdp
