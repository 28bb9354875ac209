// Associative elements of the parallel-in-time sweeps and the block-level scan over them.
//
// Each sequential recursion of the block-tridiagonal layer (sweeps.cuh) is a composition of per-step
// maps on a small carry.  A chunk of consecutive steps is summarised by one element of a family that is
// closed under composition; scanning the chunk elements gives every chunk its incoming carry, after
// which the chunk runs the reference-order recursion of sweeps.cuh from that carry.  Only the carries
// see a rounding order different from the sequential reference (measured effect on the site natural
// parameters: 1e-11 relative).
//
//   SchurElem   forward banded Cholesky (cholesky_band, block_tri_diag.py:428-440) in information form:
//               S_out = Lb - Lab^T (S_in + La)^-1 Lab.  Composition eliminates the shared node by a
//               Schur complement on positive-definite blocks (the pairwise-potential elements of the
//               parallel smoother of Sarkka & Garcia-Fernandez), i.e. it is a block Cholesky of the same
//               precision in nested-dissection order, backward stable like the sequential order.
//   InfoElem    the same Schur complements taken from the bottom with the linear term riding along: the forward
//               transitions of the Gaussian straight from its naturals (see InfoOps below).
//   AffSymElem  backward inverse band (inverse_from_cholesky_band, block_tri_diag.py:318-337):
//               S_out = C + G^T S_in G, all terms positive semi-definite.
//   AffVecElem  affine vector recursions (solve_triang_mat on the A^-1 block, state_space_model.py:251,
//               ssm_gaussian_transformations.py:502):  x_out = M x_in + v.
#pragma once
#include "smallmat.cuh"

namespace vdp {

// ---------------------------------------------------------------------------------------------------
template <int D>
struct SchurElem {
    Mat<D> La, Lab, Lb;
    int id;   // 1 = identity map
};

template <int D>
struct SchurOps {
    using Elem = SchurElem<D>;
    using State = Mat<D>;
    __device__ __forceinline__ static Elem identity() {
        Elem e;
        e.La = mat_zero<D>();
        e.Lab = mat_zero<D>();
        e.Lb = mat_zero<D>();
        e.id = 1;
        return e;
    }
    // one transition  S' = Kd_next - Ks S^-1 Ks^T
    __device__ __forceinline__ static Elem step(const Mat<D>& Ks, const Mat<D>& Kd_next) {
        Elem e;
        e.La = mat_zero<D>();
        e.Lab = transpose(Ks);
        e.Lb = Kd_next;
        e.id = 0;
        return e;
    }
    // e1 first, then e2
    __device__ __forceinline__ static Elem combine(const Elem& e1, const Elem& e2) {
        if (e1.id) return e2;
        if (e2.id) return e1;
        bool ok = true;
        const Mat<D> Lm = chol(e1.Lb + e2.La, ok);
        const Mat<D> WU = trsm(Lm, transpose(e1.Lab));
        const Mat<D> WV = trsm(Lm, e2.Lab);
        Elem r;
        r.La = e1.La - Tmm(WU, WU);
        r.Lab = Tmm(WU, WV);
        r.Lb = e2.Lb - Tmm(WV, WV);
        r.id = 0;
        return r;
    }
    // e followed by one transition (cheaper than combine(e, step()))
    __device__ __forceinline__ static void push(Elem& e, const Mat<D>& Ks, const Mat<D>& Kd_next) {
        if (e.id) { e = step(Ks, Kd_next); return; }
        bool ok = true;
        const Mat<D> Lm = chol(e.Lb, ok);
        const Mat<D> WU = trsm(Lm, transpose(e.Lab));
        const Mat<D> WV = trsm(Lm, transpose(Ks));
        e.La = e.La - Tmm(WU, WU);
        e.Lab = Tmm(WU, WV);
        e.Lb = Kd_next - Tmm(WV, WV);
    }
    __device__ __forceinline__ static State apply(const Elem& e, const State& S) {
        if (e.id) return S;
        bool ok = true;
        const Mat<D> Lx = chol(S + e.La, ok);
        const Mat<D> W = trsm(Lx, e.Lab);
        return e.Lb - Tmm(W, W);
    }
};

// D = 1: the same maps with one reciprocal per composition instead of chol + two triangular solves
template <>
struct SchurOps<1> {
    using Elem = SchurElem<1>;
    using State = Mat<1>;
    __device__ __forceinline__ static Elem identity() {
        Elem e;
        e.La.a[0] = 0.0; e.Lab.a[0] = 0.0; e.Lb.a[0] = 0.0;
        e.id = 1;
        return e;
    }
    __device__ __forceinline__ static Elem step(const Mat<1>& Ks, const Mat<1>& Kd_next) {
        Elem e;
        e.La.a[0] = 0.0; e.Lab = Ks; e.Lb = Kd_next;
        e.id = 0;
        return e;
    }
    __device__ __forceinline__ static Elem combine(const Elem& e1, const Elem& e2) {
        if (e1.id) return e2;
        if (e2.id) return e1;
        const double r = rcp_fast(e1.Lb.a[0] + e2.La.a[0]);
        const double ur = e1.Lab.a[0] * r, v = e2.Lab.a[0];
        Elem o;
        o.La.a[0] = e1.La.a[0] - ur * e1.Lab.a[0];
        o.Lab.a[0] = ur * v;
        o.Lb.a[0] = e2.Lb.a[0] - (v * r) * v;
        o.id = 0;
        return o;
    }
    __device__ __forceinline__ static void push(Elem& e, const Mat<1>& Ks, const Mat<1>& Kd_next) {
        if (e.id) { e = step(Ks, Kd_next); return; }
        const double r = rcp_fast(e.Lb.a[0]);
        const double ur = e.Lab.a[0] * r, v = Ks.a[0];
        e.La.a[0] = e.La.a[0] - ur * e.Lab.a[0];
        e.Lab.a[0] = ur * v;
        e.Lb.a[0] = Kd_next.a[0] - (v * r) * v;
    }
    __device__ __forceinline__ static State apply(const Elem& e, const State& S) {
        if (e.id) return S;
        State o;
        o.a[0] = e.Lb.a[0] - div_fast(e.Lab.a[0], S.a[0] + e.La.a[0]) * e.Lab.a[0];
        return o;
    }
};

// ---------------------------------------------------------------------------------------------------
// Information-form backward sweep: the precision K = (-2 theta_d, -theta_s) with its linear term theta_1 taken apart
// from the BOTTOM.  State at node k+1 = (J, h), the precision and linear term of x_{k+1} given x_k = 0 with everything
// after k+1 integrated out; one step down
//     J_k = Kd_k - Ts_k^T J_{k+1}^-1 Ts_k,     h_k = t1_k + Ts_k^T J_{k+1}^-1 h_{k+1}        (Ts_k = theta_s[k] = -K_{k+1,k})
// and on the way  A_k = J_{k+1}^-1 Ts_k,  b_k = J_{k+1}^-1 h_{k+1},  Q_k = J_{k+1}^-1: the forward transitions of the
// Gaussian the naturals describe (what naturals_to_ssm_params, ssm_gaussian_transformations.py:333-511, returns), without
// a Cholesky sweep from the top and without the inverse band.  Elements:
//     J' = Lb - Lab^T (J + La)^-1 Lab,     h' = vb + Lab^T (J + La)^-1 (h + va)
// closed under composition (the matrix part is SchurElem; the vector part rides on the same Schur complement).
template <int D>
struct InfoElem {
    Mat<D> La, Lab, Lb;
    Vec<D> va, vb;
    int id;   // 1 = identity map
};
template <int D>
struct InfoState {
    Mat<D> J;
    Vec<D> h;
};

template <int D>
struct InfoOps {
    using Elem = InfoElem<D>;
    using State = InfoState<D>;
    VDP_HD static Elem identity() {
        Elem e;
        e.La = mat_zero<D>();
        e.Lab = mat_zero<D>();
        e.Lb = mat_zero<D>();
        e.va = vec_zero<D>();
        e.vb = vec_zero<D>();
        e.id = 1;
        return e;
    }
    VDP_HD static Elem step(const Mat<D>& Ts, const Mat<D>& Kd, const Vec<D>& t1) {
        Elem e;
        e.La = mat_zero<D>();
        e.Lab = Ts;
        e.Lb = Kd;
        e.va = vec_zero<D>();
        e.vb = t1;
        e.id = 0;
        return e;
    }
    // e1 first (the upper stretch of the chain), then e2
    VDP_HD static Elem combine(const Elem& e1, const Elem& e2) {
        if (e1.id) return e2;
        if (e2.id) return e1;
        bool ok = true;
        const Mat<D> Lm = chol(e1.Lb + e2.La, ok);
        const Mat<D> WU = trsm(Lm, transpose(e1.Lab));
        const Mat<D> WV = trsm(Lm, e2.Lab);
        const Vec<D> wg = trsv(Lm, e1.vb + e2.va);
        Elem r;
        r.La = e1.La - Tmm(WU, WU);
        r.Lab = Tmm(WU, WV);
        r.Lb = e2.Lb - Tmm(WV, WV);
        r.va = e1.va + Tmv(WU, wg);
        r.vb = e2.vb + Tmv(WV, wg);
        r.id = 0;
        return r;
    }
    // e followed by one step (La = 0, va = 0 of the step folded in)
    VDP_HD static void push(Elem& e, const Mat<D>& Ts, const Mat<D>& Kd, const Vec<D>& t1) {
        if (e.id) { e = step(Ts, Kd, t1); return; }
        bool ok = true;
        const Mat<D> Lm = chol(e.Lb, ok);
        const Mat<D> WU = trsm(Lm, transpose(e.Lab));
        const Mat<D> WV = trsm(Lm, Ts);
        const Vec<D> wg = trsv(Lm, e.vb);
        e.La = e.La - Tmm(WU, WU);
        e.Lab = Tmm(WU, WV);
        e.Lb = Kd - Tmm(WV, WV);
        e.va = e.va + Tmv(WU, wg);
        e.vb = t1 + Tmv(WV, wg);
    }
    VDP_HD static State apply(const Elem& e, const State& s) {
        if (e.id) return s;
        bool ok = true;
        const Mat<D> Lx = chol(s.J + e.La, ok);
        const Mat<D> W = trsm(Lx, e.Lab);
        State r;
        r.J = e.Lb - Tmm(W, W);
        r.h = e.vb + Tmv(W, trsv(Lx, s.h + e.va));
        return r;
    }
};

// ---------------------------------------------------------------------------------------------------
template <int D>
struct AffSymElem {
    Mat<D> C, G;
};

template <int D>
struct AffSymOps {
    using Elem = AffSymElem<D>;
    using State = Mat<D>;
    __device__ __forceinline__ static Elem identity() {
        Elem e;
        e.C = mat_zero<D>();
        e.G = mat_eye<D>();
        return e;
    }
    __device__ __forceinline__ static Elem combine(const Elem& e1, const Elem& e2) {
        Elem r;
        r.C = e2.C + Tmm(e2.G, mm(e1.C, e2.G));
        r.G = mm(e1.G, e2.G);
        return r;
    }
    __device__ __forceinline__ static State apply(const Elem& e, const State& S) {
        return e.C + Tmm(e.G, mm(S, e.G));
    }
};

// ---------------------------------------------------------------------------------------------------
template <int D>
struct AffVecElem {
    Mat<D> M;
    Vec<D> v;
};

template <int D>
struct AffVecOps {
    using Elem = AffVecElem<D>;
    using State = Vec<D>;
    __device__ __forceinline__ static Elem identity() {
        Elem e;
        e.M = mat_eye<D>();
        e.v = vec_zero<D>();
        return e;
    }
    __device__ __forceinline__ static Elem combine(const Elem& e1, const Elem& e2) {
        Elem r;
        r.M = mm(e2.M, e1.M);
        r.v = e2.v + mv(e2.M, e1.v);
        return r;
    }
    __device__ __forceinline__ static State apply(const Elem& e, const State& x) { return e.v + mv(e.M, x); }
};

// ---------------------------------------------------------------------------------------------------
// warp shuffle of a register struct
template <typename T>
__device__ __forceinline__ T shfl_any(const T& v, int src_lane) {
    static_assert(sizeof(T) % 4 == 0, "struct must be made of 32-bit words");
    T r;
    const int* pi = reinterpret_cast<const int*>(&v);
    int* po = reinterpret_cast<int*>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 4); ++i) po[i] = __shfl_sync(0xffffffffu, pi[i], src_lane);
    return r;
}

// Exclusive scan of one element per thread over a block of T threads, thread 0 first.
// Returns the composition of the elements of all preceding threads (identity for thread 0) and, in
// `total`, the composition of all T elements (the same value in every thread).
// `warp_tot` is shared scratch of T/32 elements; the call contains block-wide barriers.
template <class Ops, int T>
__device__ __forceinline__ typename Ops::Elem block_exclusive_scan(const typename Ops::Elem& mine,
                                                                   typename Ops::Elem* warp_tot,
                                                                   typename Ops::Elem& total) {
    using E = typename Ops::Elem;
    constexpr int W = T / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    E inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const E other = shfl_any(inc, lane >= d ? lane - d : lane);
        if (lane >= d) inc = Ops::combine(other, inc);
    }
    E excl = shfl_any(inc, lane > 0 ? lane - 1 : 0);
    if (lane == 0) excl = Ops::identity();
    __syncthreads();   // previous use of warp_tot is over
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    E tot = warp_tot[0];
    E wp = Ops::identity();
#pragma unroll
    for (int w = 1; w < W; ++w) {
        if (w == warp) wp = tot;
        tot = Ops::combine(tot, warp_tot[w]);
    }
    total = tot;
    if (warp == 0) return excl;
    return Ops::combine(wp, excl);
}

}  // namespace vdp
