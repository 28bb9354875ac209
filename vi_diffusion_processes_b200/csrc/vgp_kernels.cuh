// VDP (Archambeau et al. 2007) fixed-point sweep: markovflow/models/vi_sde.py (VariationalMarkovGP).
//
//   forward_pass      vi_sde.py:171-204   LinearDrift(-A, b).to_ssm (drift.py:64-116) -> marginals
//   E_sde, its grads  vi_sde.py:206-240, 422-434 -> squared_drift_difference_along_Gaussian_path
//                     (sde_utils.py:182-249), H = 20, reverse mode through the quadrature and chol(S)
//   jump conditions   vi_sde.py:262-287   likelihood gradients at the observed grid points
//   update_lagrange   vi_sde.py:289-347   backward Euler sweep for (lambda, psi)
//   update_param      vi_sde.py:377-414   damped fixed point for (A, b), E[f], E[grad f] with H = 10
//   update_initial_statistics :242-260, KL_initial_state :416-420, elbo :436-455
//
// Layouts: A [B,N,D,D], b [B,N,D], lambda [B,N,D], psi [B,N,D,D], m [B,N+1,D], S [B,N+1,D,D].
#pragma once
#include "kl_kernels.cuh"
#include "tile_sweeps.cuh"

namespace vdp {

constexpr double kVgpClipMin = -5000.0, kVgpClipMax = 5000.0;   // vi_sde.py:59-60

// tf.where(is_nan(x), 1e-8, x) then clip_by_value
__device__ __forceinline__ double vgp_fix(double x, double lo, double hi) {
    if (x != x) x = 1e-8;
    return fmin(fmax(x, lo), hi);
}
template <int D> __device__ __forceinline__ Mat<D> vgp_fix(Mat<D> x, double lo, double hi) {
#pragma unroll
    for (int i = 0; i < D * D; ++i) x.a[i] = vgp_fix(x.a[i], lo, hi);
    return x;
}
template <int D> __device__ __forceinline__ Vec<D> vgp_fix(Vec<D> x, double lo, double hi) {
#pragma unroll
    for (int i = 0; i < D; ++i) x.v[i] = vgp_fix(x.v[i], lo, hi);
    return x;
}

// LinearDrift(a_sign * A, b).to_ssm (+ the stabilize_system clip of vi_sde.py:184-197); one thread per transition
template <int D>
__global__ void k_vgp_to_ssm(const double* __restrict__ A, const double* __restrict__ bv, double a_sign, Mat<D> q,
                             const double* __restrict__ time_grid, int stabilize, double* __restrict__ sA,
                             double* __restrict__ sb, double* __restrict__ scQ, int64_t B, int64_t N,
                             int64_t* status) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= B * N) return;
    constexpr int DD = D * D;
    const int64_t k = t % N;
    const double delta = time_grid[k + 1] - time_grid[k];
    Mat<D> At = delta * (a_sign * mat_load<D>(A + t * DD)) + mat_eye<D>();
    Vec<D> bt = delta * vec_load<D>(bv + t * D);
    bool ok = true;
    const Mat<D> cQ = chol(delta * q, ok);
    if (!ok) report_not_pd(status, t);
    if (stabilize) { At = vgp_fix(At, -1.0, 1.0); bt = vgp_fix(bt, -1.0, 1.0); }
    mat_store<D>(sA + t * DD, At);
    vec_store<D>(sb + t * D, bt);
    mat_store<D>(scQ + t * DD, cQ);
}

// VariationalMarkovGP.forward_pass (vi_sde.py:171-204) in ONE forward sweep: LinearDrift(-A, b).to_ssm (+ the
// stabilize_system clip, :184-197) evaluated in the staged tile, the marginal means m_{k+1} = A_k m_k + b_k
// (state_space_model.py:251) and the marginal covariances by the forward recursion
//     S_{k+1} = A_k S_k A_k^T + Q_k,   S_0 = chol P_0 chol P_0^T,   Q_k = chol(q delta_k) chol(q delta_k)^T,
// scanned together as PairOps<AffVecOps, AffSymOps> (element of S: C = Q, G = A^T).  The reference takes the
// covariances from the rebuilt precision (banded Cholesky + inverse band, state_space_model.py:262); the forward
// recursion is the same quantity by another rounding route (sums of positive semi-definite terms; pinned to the oracle
// at 1e-9 by the VDP parity tests, and to the reference's route -- VDP_VGP_REFACTOR_PRECISION -- at 1e-11).
// 32 instead of 112 bytes per grid step (D = 1) and one sweep instead of a pointwise pass + two sweeps.
template <int D>
using FwdOps = PairOps<AffVecOps<D>, AffSymOps<D>>;

// process noise of one step as the state-space model holds it, chol(delta q) chol(delta q)^T (drift.py:64-116 keeps the
// factor); a scalar needs no round trip through the square root
template <int D>
__device__ __forceinline__ Mat<D> step_noise(double delta, const Mat<D>& q, bool& ok) {
    if constexpr (D == 1) {
        Mat<1> r;
        r.a[0] = delta * q.a[0];
        if (!(r.a[0] > 0.0)) ok = false;
        return r;
    } else {
        const Mat<D> cQ = chol(delta * q, ok);
        return mmT(cQ, cQ);
    }
}

// E_sde term of one transition and its gradients for a polynomial scalar drift in closed form (the quantities of
// k_vgp_esde<.., CLOSED> below: term = 0.5 dt E || (-A) x + b - f(x) ||^2 / q at the marginal N(m, S) of the
// transition's first grid point; dEdm, dEdS divided by dt as vi_sde.py:206-225 does).  Scalar algebra: the adjoint
// of chol(S) collapses to a division by S, so no square root is needed (sinv = 1 / q, hoisted by the caller).
template <int DRIFT>
__device__ __forceinline__ void vgp_esde_closed1(const DriftParams& P, double sinv, double dt, double A0, double b0,
                                                 double m1, double Sk, double& term, double& dm, double& dS, bool& ok) {
    if (!(Sk > 0.0)) ok = false;
    const double a1 = -A0;
    const typename DriftStep1<DRIFT>::K kst = DriftStep1<DRIFT>::prepare(P, -1.0, 1.0 - a1, -b0);
    double c0, c1, c2, c3;
    DriftStep1<DRIFT>::taylor(kst, m1, c0, c1, c2, c3);
    CubicMoments cm;
    cm.template eval<true>(c0, c1, c2, c3, Sk);
    term = (0.5 * dt) * (sinv * cm.Ed2);
    dm = sinv * cm.Ejd;
    dS = (0.5 * sinv) * (cm.Ejde * rcp_fast(Sk));
}

// DRIFT >= 0 (D = 1, polynomial drift, FULL mode): the sweep also evaluates the E_sde terms and their gradients of
// every transition from the marginals it has just computed -- per-tile partial sums of the terms and dEdm / dEdS leave
// with the tile -- so the separate pass over A, b, m, S (k_vgp_esde, 56 bytes per grid step) is not needed.
struct VgpEsdeArgs {
    DriftParams P;
    double q = 1.0, dt = 0.0;
    double *dEdm = nullptr, *dEdS = nullptr;   // [B, N]
    double* partials = nullptr;                // [B, nseg * tiles_per_seg]
    int tiles_per_seg = 0;
};

template <int D, int T, int L, int MODE, int NSTAGE = 1, int DRIFT = -1>
__global__ void __launch_bounds__(T)
k_vgp_forward_tiled(const double* __restrict__ A, const double* __restrict__ bv, Mat<D> q,
                    const double* __restrict__ time_grid, int stabilize, const double* __restrict__ mu0,
                    const double* __restrict__ cholP0, double* __restrict__ m, double* __restrict__ S, int64_t B,
                    int64_t N, int nseg, int64_t seg_len, const typename FwdOps<D>::State* __restrict__ carry_in,
                    typename FwdOps<D>::State* __restrict__ init_out, typename FwdOps<D>::Elem* __restrict__ seg_elem,
                    int64_t* status, VgpEsdeArgs es = VgpEsdeArgs()) {
    constexpr int DD = D * D, TL = T * L;
    constexpr bool ESDE = (DRIFT >= 0) && (MODE == SWEEP_FULL);
    static_assert(DRIFT < 0 || (D == 1 && NSTAGE == 1), "the fused E_sde terms are scalar, single staging buffer");
    using Ops = FwdOps<D>;
    constexpr int SM = tile_slots(TL * DD), SV = tile_slots(TL * D);
    constexpr int kStage = SM + SV;
    extern __shared__ __align__(16) double smem[];
    double* stage0 = smem + kRingHeaderDoubles;
    double* rawGm = stage0 + NSTAGE * kStage;                 // ESDE: dEdm / dEdS of the tile, then the reduction
    double* rawGS = rawGm + (ESDE ? SM : 0);
    double* xcarry = rawGS + (ESDE ? SM : 0);                 // ESDE: stored marginal of the tile's last grid point
    typename Ops::Elem* wtot = reinterpret_cast<typename Ops::Elem*>(xcarry + (ESDE ? 4 : 0));
    TileRing<NSTAGE> ring;
    ring.init(smem);
    const int64_t b = blockIdx.x / nseg;
    const int seg = (int)(blockIdx.x % nseg);
    const int64_t t0 = seg * seg_len, t1 = min(N, t0 + seg_len);
    const int ntiles = (int)((t1 - t0 + TL - 1) / TL);
    auto issue = [&](int ti) {
        const int64_t tk = t0 + (int64_t)ti * TL;
        const int cnt = (int)min((int64_t)TL, t1 - tk);
        double* base = stage0 + (ti % NSTAGE) * kStage;
        uint64_t* bar = ring.bar + ti % NSTAGE;
        uint32_t bytes = tile_fetch(base, A + (b * N + tk) * DD, cnt * DD, bar);
        bytes += tile_fetch(base + SM, bv + (b * N + tk) * D, cnt * D, bar);
        mbar_arrive_expect_tx(bar, bytes);
    };
    if (threadIdx.x == 0 && ntiles > 0) issue(0);
    typename Ops::State st;
    st.a = vec_zero<D>();
    st.b = mat_zero<D>();
    if (MODE == SWEEP_FULL && seg > 0) st = carry_in[b * nseg + seg];
    typename Ops::Elem seg_tot = Ops::identity();
    const double es_sinv = ESDE ? 1.0 / es.q : 0.0;
    for (int ti = 0; ti < ntiles; ++ti) {
        const int64_t tk = t0 + (int64_t)ti * TL;
        const int cnt = (int)min((int64_t)TL, t1 - tk);
        if (NSTAGE == 1 && ti > 0) {
            __syncthreads();
            if (threadIdx.x == 0) { bulk_wait_read(); issue(ti); }
        }
        ring.wait(ti);
        double* base = stage0 + (ti % NSTAGE) * kStage;
        double* sA = base + tile_off(A + (b * N + tk) * DD);       // drift A -> SSM A_k -> S_{k+1}
        double* sb = base + SM + tile_off(bv + (b * N + tk) * D);  // drift b -> SSM b_k -> m_{k+1}
        if (seg == 0 && tk == 0) {
            st.a = vec_load<D>(mu0 + b * D);
            const Mat<D> c0 = mat_load<D>(cholP0 + b * DD);
            st.b = mmT(c0, c0);
            if (MODE != SWEEP_FULL && threadIdx.x == 0) init_out[b] = st;
            if (MODE == SWEEP_FULL && threadIdx.x == 0) {
                vec_store<D>(m + b * (N + 1) * D, st.a);
                mat_store<D>(S + b * (N + 1) * DD, st.b);
            }
        }
        const int k0 = threadIdx.x * L;
        const int len = min(L, cnt - k0);
        // LinearDrift(-A, b).to_ssm on this thread's chunk, in place; chunk element on the way
        typename Ops::Elem e = Ops::identity();
        bool ok = true;
        for (int t = 0; t < len; ++t) {
            const double delta = time_grid[tk + k0 + t + 1] - time_grid[tk + k0 + t];
            Mat<D> At = delta * (-1.0 * mat_load<D>(sA + (k0 + t) * DD)) + mat_eye<D>();
            Vec<D> bt = delta * vec_load<D>(sb + (k0 + t) * D);
            if (stabilize) { At = vgp_fix(At, -1.0, 1.0); bt = vgp_fix(bt, -1.0, 1.0); }
            if (!ESDE) {   // (ESDE keeps the drift's own A, b staged: the replay needs them for the E_sde terms)
                mat_store<D>(sA + (k0 + t) * DD, At);
                vec_store<D>(sb + (k0 + t) * D, bt);
            }
            typename Ops::Elem s;
            s.a.M = At;
            s.a.v = bt;
            s.b.C = step_noise<D>(delta, q, ok);
            s.b.G = transpose(At);
            e = Ops::combine(e, s);
        }
        if (!ok) report_not_pd(status, b * (N + 1) + tk + k0);
        if (NSTAGE > 1 && threadIdx.x == 0 && ti + 1 < ntiles) {
            bulk_wait_read();
            issue(ti + 1);
        }
        typename Ops::Elem total;
        const typename Ops::Elem excl = block_exclusive_scan<Ops, T>(e, wtot, total);
        if (MODE != SWEEP_FULL) {
            seg_tot = Ops::combine(seg_tot, total);
            continue;
        }
        typename Ops::State cur = Ops::apply(excl, st);
        const typename Ops::State tile_in = st;
        st = Ops::apply(total, st);
        double* sGm = rawGm;
        double* sGS = rawGS;
        double keepA = 0.0, keepb = 0.0, acc = 0.0;
        if constexpr (ESDE) {   // outputs placed with the parity of their destination (bulk stores)
            sGm += tile_off(es.dEdm + (b * N + tk));
            sGS += tile_off(es.dEdS + (b * N + tk));
        }
        for (int t = 0; t < len; ++t) {
            const double delta = time_grid[tk + k0 + t + 1] - time_grid[tk + k0 + t];
            Mat<D> At;
            Vec<D> bt;
            if constexpr (ESDE) {
                const Mat<D> A0 = mat_load<D>(sA + (k0 + t) * DD);
                const Vec<D> b0 = vec_load<D>(sb + (k0 + t) * D);
                if (t == 0) {
                    keepA = A0.a[0];       // the marginal of the chunk's first grid point is the neighbour's to store
                    keepb = b0.v[0];
                } else {
                    double term, dm, dS;
                    vgp_esde_closed1<(DRIFT < 0 ? 0 : DRIFT)>(es.P, es_sinv, es.dt, A0.a[0], b0.v[0], cur.a.v[0],
                                                              cur.b.a[0], term, dm, dS, ok);
                    acc += term;
                    sGm[k0 + t] = dm;
                    sGS[k0 + t] = dS;
                }
                At = delta * (-1.0 * A0) + mat_eye<D>();
                bt = delta * b0;
                if (stabilize) { At = vgp_fix(At, -1.0, 1.0); bt = vgp_fix(bt, -1.0, 1.0); }
            } else {
                At = mat_load<D>(sA + (k0 + t) * DD);
                bt = vec_load<D>(sb + (k0 + t) * D);
            }
            cur.a = bt + mv(At, cur.a);
            cur.b = step_noise<D>(delta, q, ok) + mm(At, mmT(cur.b, At));
            vec_store<D>(sb + (k0 + t) * D, cur.a);
            mat_store<D>(sA + (k0 + t) * DD, cur.b);
        }
        if constexpr (ESDE) {
            if (len > 0 && k0 + len == cnt) {   // the tile's last stored marginal, for the next tile's first transition
                xcarry[2 * (ti & 1)] = cur.a.v[0];
                xcarry[2 * (ti & 1) + 1] = cur.b.a[0];
            }
            __syncthreads();
            if (len > 0) {
                double m1, S1;
                if (threadIdx.x > 0) { m1 = sb[k0 - 1]; S1 = sA[k0 - 1]; }
                else if (ti > 0) { m1 = xcarry[2 * ((ti - 1) & 1)]; S1 = xcarry[2 * ((ti - 1) & 1) + 1]; }
                else { m1 = tile_in.a.v[0]; S1 = tile_in.b.a[0]; }   // sequence start (or the segment's incoming carry)
                double term, dm, dS;
                vgp_esde_closed1<(DRIFT < 0 ? 0 : DRIFT)>(es.P, es_sinv, es.dt, keepA, keepb, m1, S1, term, dm, dS, ok);
                acc += term;
                sGm[k0] = dm;
                sGS[k0] = dS;
            }
            if (!ok) report_not_pd(status, b * (N + 1) + tk + k0);
        }
        fence_async_smem();
        __syncthreads();
        tile_store<T>(m + (b * (N + 1) + tk + 1) * D, sb, cnt * D);
        tile_store<T>(S + (b * (N + 1) + tk + 1) * DD, sA, cnt * DD);
        if constexpr (ESDE) {
            tile_store<T>(es.dEdm + (b * N + tk), sGm, cnt);
            tile_store<T>(es.dEdS + (b * N + tk), sGS, cnt);
        }
        if (threadIdx.x == 0) bulk_commit();
        if constexpr (ESDE) {
            // sum of the tile's terms: fixed tree over the block (scratch behind the carry slots' scan area is busy:
            // the staged b array has been handed to the bulk store, so the reduction uses its own slots)
            double* red = reinterpret_cast<double*>(wtot + T / 32);
            red[threadIdx.x] = acc;
            __syncthreads();
            for (int s2 = T / 2; s2 > 0; s2 >>= 1) {
                if ((int)threadIdx.x < s2) red[threadIdx.x] += red[threadIdx.x + s2];
                __syncthreads();
            }
            if (threadIdx.x == 0)
                es.partials[((int64_t)b * nseg + seg) * es.tiles_per_seg + ti] = red[0];
        }
    }
    if (threadIdx.x == 0) {
        if (MODE != SWEEP_FULL) seg_elem[b * nseg + seg] = seg_tot;
        bulk_wait_all();
    }
}

template <int D, int T, int L, int NSTAGE = 1, bool ESDE = false>
constexpr size_t vgp_forward_tiled_smem() {
    return ((size_t)kRingHeaderDoubles + (size_t)NSTAGE * (tile_slots(T * L * D * D) + tile_slots(T * L * D)) +
            (ESDE ? 2 * (size_t)tile_slots(T * L * D * D) + 4 + T : 0)) * sizeof(double) +
           (T / 32) * sizeof(typename FwdOps<D>::Elem);
}

// E_sde terms and (optionally) gradients: one group of G lanes per transition (b, k), marginal of node k.
//   term_k = 0.5 dt E_{N(m_k,S_k)} || (-A_k) x + b_k - f(x) ||^2_{q^-1}
//   dEdm = d term / d m_k / dt, dEdS = d term / d S_k / dt       (vi_sde.py:206-225 divides by dt)
// CLOSED (polynomial scalar drifts): the Gauss-Hermite sums in closed form (CubicMoments, drifts.cuh)
template <int DRIFT, int D, int H, int G, bool CLOSED = false>
__global__ void k_vgp_esde(DriftParams P, Mat<D> q, double dt, double pi_factor, const double* __restrict__ A,
                           const double* __restrict__ bv, const double* __restrict__ m,
                           const double* __restrict__ S, double* __restrict__ terms, double* __restrict__ dEdm,
                           double* __restrict__ dEdS, int64_t B, int64_t N, int64_t* status) {
    const int64_t t = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / G;
    const int lane = threadIdx.x % G;
    if (t >= B * N) return;
    constexpr int DD = D * D;
    const int64_t b = t / N, k = t % N;
    const int64_t node = b * (N + 1) + k;
    bool ok = true;
    const Mat<D> Al = -1.0 * mat_load<D>(A + t * DD);
    const Vec<D> bl = vec_load<D>(bv + t * D);
    const Vec<D> mk = vec_load<D>(m + node * D);
    const Mat<D> LS = chol(mat_load<D>(S + node * DD), ok);
    const Mat<D> cq = chol(q, ok);
    const Mat<D> sinv = chol_inverse(cq);
    const Mat<D> sinv_s = sym(sinv);
    double quad = 0.0;
    Vec<D> gm = vec_zero<D>();
    Mat<D> Lb = mat_zero<D>();
    constexpr int NP = NodeCount<H, D>::value;
    constexpr bool kScalar = (D == 1 && G == 1);
    if constexpr (kScalar) {
        // one thread per transition: nodes unrolled (tables as immediates), q^-1 and pi^(-1/2) factored out
        const double sL = 1.4142135623730951 * LS(0, 0), m1 = mk(0), a1 = Al(0, 0), b1 = bl(0);
        double q_ = 0.0, g_ = 0.0, l_ = 0.0;
        if constexpr (CLOSED) {
            // d(x) = a1 x + b1 - f(x) = x - f(x) - ((1 - a1) x - b1): DriftStep1 with dt = -1
            const typename DriftStep1<DRIFT>::K kst = DriftStep1<DRIFT>::prepare(P, -1.0, 1.0 - a1, -b1);
            double c0, c1, c2, c3;
            DriftStep1<DRIFT>::taylor(kst, m1, c0, c1, c2, c3);
            const double Sk = LS(0, 0) * LS(0, 0);
            CubicMoments cm;
            if (dEdm) cm.template eval<true>(c0, c1, c2, c3, Sk); else cm.template eval<false>(c0, c1, c2, c3, Sk);
            quad = sinv(0, 0) * cm.Ed2;
            if (dEdm) {
                gm(0) = (dt * sinv(0, 0)) * cm.Ejd;
                Lb(0, 0) = (dt * sinv(0, 0)) * cm.Ejde / LS(0, 0);
            }
        } else {
        if (dEdm) {
#pragma unroll
            for (int p = 0; p < H; ++p) {
                Vec<1> X, f;
                Mat<1> J;
                X(0) = fma(sL, gh_x<H>(p), m1);
                Drift<DRIFT, 1>::eval(P, X, f, J);
                const double d = fma(a1, X(0), b1) - f(0);
                const double wd = gh_w<H>(p) * d;
                q_ = fma(wd, d, q_);
                const double s1 = (a1 - J(0, 0)) * wd;
                g_ += s1;
                l_ = fma(s1, gh_x<H>(p), l_);
            }
        } else {
#pragma unroll
            for (int p = 0; p < H; ++p) {
                Vec<1> X, f;
                Mat<1> J;
                X(0) = fma(sL, gh_x<H>(p), m1);
                Drift<DRIFT, 1>::eval(P, X, f, J);
                const double d = fma(a1, X(0), b1) - f(0);
                q_ = fma(gh_w<H>(p) * d, d, q_);
            }
        }
        const double sc = pi_factor * sinv(0, 0);
        quad = sc * q_;
        gm(0) = (dt * sc) * g_;
        Lb(0, 0) = (1.4142135623730951 * dt * sc) * l_;
        }
    }
    for (int p = lane; p < (kScalar ? 0 : NP); p += G) {
        Vec<D> xi, f;
        Mat<D> J;
        double w;
        gh_node<H, D>(p, xi, w);
        w *= pi_factor;
        const Vec<D> X = gh_point(LS, mk, xi);
        Drift<DRIFT, D>::eval(P, X, f, J);
        const Vec<D> AX = mv(Al, X);
        Vec<D> d;
#pragma unroll
        for (int i = 0; i < D; ++i) d(i) = (AX(i) + bl(i)) - f(i);
        quad += w * dot(d, mv(sinv, d));
        if (dEdm) {
            const Vec<D> rb = (dt * w) * mv(sinv_s, d);
#pragma unroll
            for (int i = 0; i < D; ++i) {
                double s = 0.0;
#pragma unroll
                for (int j = 0; j < D; ++j) s += (Al(j, i) - J(j, i)) * rb(j);
                gm(i) += s;
#pragma unroll
                for (int j = 0; j <= i; ++j) Lb(i, j) += 1.4142135623730951 * (s * xi(j));
            }
        }
    }
    quad = group_sum<G>(quad);
    if (dEdm) { group_sum<G>(gm); group_sum<G>(Lb); }
    if (lane != 0) return;
    if (!ok) report_not_pd(status, node);
    terms[t] = 0.5 * dt * quad;
    if (dEdm) {
        const Mat<D> gS = chol_adjoint(LS, Lb);
        const double idt = 1.0 / dt;
        vec_store<D>(dEdm + t * D, idt * gm);
        mat_store<D>(dEdS + t * DD, idt * gS);
    }
}

// jump conditions: likelihood gradients scattered on the grid (buffers zeroed by the caller)
template <int D>
__global__ void k_vgp_jump(const double* __restrict__ m, const int64_t* __restrict__ obs_idx,
                           const double* __restrict__ obs_y, Mat<D> lik_chol, double* __restrict__ dobs_m,
                           double* __restrict__ dobs_S, int64_t B, int64_t N, int64_t n_obs) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * n_obs) return;
    constexpr int DD = D * D;
    const int64_t b = i / n_obs;
    const int64_t k = obs_idx[i];
    if (k < 0 || k > N) return;
    const int64_t node = b * (N + 1) + k;
    const Mat<D> inv = chol_inverse(lik_chol);
    const Vec<D> gmu = mv(sym(inv), vec_load<D>(obs_y + i * D) - vec_load<D>(m + node * D));
    const Mat<D> gS = -0.5 * inv;
#pragma unroll
    for (int e = 0; e < D; ++e) atomicAdd(dobs_m + node * D + e, gmu(e));
#pragma unroll
    for (int e = 0; e < DD; ++e) atomicAdd(dobs_S + node * DD + e, gS.a[e]);
}

// the entries k_vgp_jump is about to accumulate into, set to zero (consecutive sweeps: the rest of the two buffers
// is still zero from the sweep before, so the dense memsets are not repeated)
template <int D>
__global__ void k_vgp_jump_clear(const int64_t* __restrict__ obs_idx, double* __restrict__ dobs_m,
                                 double* __restrict__ dobs_S, int64_t B, int64_t N, int64_t n_obs) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * n_obs) return;
    constexpr int DD = D * D;
    const int64_t k = obs_idx[i];
    if (k < 0 || k > N) return;
    const int64_t node = (i / n_obs) * (N + 1) + k;
#pragma unroll
    for (int e = 0; e < D; ++e) dobs_m[node * D + e] = 0.0;
#pragma unroll
    for (int e = 0; e < DD; ++e) dobs_S[node * DD + e] = 0.0;
}

// update_param: one group per transition (b, k), marginal of node k, H = 10
template <int DRIFT, int D, int H, int G, bool CLOSED = false>
__global__ void k_vgp_update_param(DriftParams P, Mat<D> q, double pi_factor, int stabilize, double lr,
                                   const double* __restrict__ m, const double* __restrict__ S,
                                   double* __restrict__ lam, double* __restrict__ psi, double* __restrict__ A,
                                   double* __restrict__ bv, int64_t B, int64_t N, int64_t* status) {
    const int64_t t = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / G;
    const int lane = threadIdx.x % G;
    if (t >= B * N) return;
    constexpr int DD = D * D;
    const int64_t b = t / N, k = t % N;
    const int64_t node = b * (N + 1) + k;
    bool ok = true;
    Vec<D> ef;
    Mat<D> ej;
    const Vec<D> mk = vec_load<D>(m + node * D);
    if constexpr (CLOSED) {
        // E f = c0 + c2 S, E f' = c1 + 3 c3 S with f(m + e) = c0 + c1 e + c2 e^2 + c3 e^3 (DriftStep1 with dt = 1,
        // A = 1, b = 0 is d = f)
        const typename DriftStep1<DRIFT>::K kst = DriftStep1<DRIFT>::prepare(P, 1.0, 1.0, 0.0);
        double c0, c1, c2, c3;
        DriftStep1<DRIFT>::taylor(kst, mk(0), c0, c1, c2, c3);
        const double Sk = S[node];
        if (!(Sk > 0.0)) ok = false;
        ef(0) = fma(c2, Sk, c0);
        ej(0, 0) = fma(3.0 * c3, Sk, c1);
    } else {
        gh_expect_point<DRIFT, D, H, G>(P, mk, mat_load<D>(S + node * DD), lane, pi_factor, ef, ej, ok);
    }
    if (lane != 0) return;
    if (!ok) report_not_pd(status, node);
    Mat<D> ps = mat_load<D>(psi + t * DD);
    Vec<D> la = vec_load<D>(lam + t * D);
    if (stabilize) {
        ps = vgp_fix(ps, kVgpClipMin, kVgpClipMax);
        la = vgp_fix(la, kVgpClipMin, kVgpClipMax);
        mat_store<D>(psi + t * DD, ps);
        vec_store<D>(lam + t * D, la);
    }
    const Mat<D> At = (-1.0 * ej) + 2.0 * mm(q, ps);
    const Vec<D> bt = (ef + mv(At, mk)) - mv(q, la);
    mat_store<D>(A + t * DD, (1.0 - lr) * mat_load<D>(A + t * DD) + lr * At);
    vec_store<D>(bv + t * D, (1.0 - lr) * vec_load<D>(bv + t * D) + lr * bt);
}


// ---------------------------------------------------------------------------------------------------
// update_lagrange (vi_sde.py:337-347): for t = N-1 .. 1
//     d_psi = psi_t A_t + psi_t A_t - dEdS_t ;  d_lam = A_t lam_t - dEdm_t
//     psi_{t-1} = psi_t - dt d_psi - dObsS_t ;  lam_{t-1} = lam_t - dt d_lam - dObsm_t
// with lam_{N-1} = 0, psi_{N-1} = 1e-10 I.  "Transition" j = t - 1 maps state j+1 to state j; the maps are
// affine (lam: left-affine, psi: right-affine), scanned together.
template <int D>
struct RightAffElem { Mat<D> R, V; };
template <int D>
struct RightAffOps {
    using Elem = RightAffElem<D>;
    using State = Mat<D>;
    __device__ __forceinline__ static Elem identity() {
        Elem e;
        e.R = mat_eye<D>();
        e.V = mat_zero<D>();
        return e;
    }
    __device__ __forceinline__ static Elem combine(const Elem& e1, const Elem& e2) {
        Elem r;
        r.R = mm(e1.R, e2.R);
        r.V = mm(e1.V, e2.R) + e2.V;
        return r;
    }
    __device__ __forceinline__ static State apply(const Elem& e, const State& X) { return mm(X, e.R) + e.V; }
};
template <int D>
using LagrangeOps = PairOps<AffVecOps<D>, RightAffOps<D>>;

template <int D, int T, int L, int MODE>
__global__ void __launch_bounds__(T)
k_vgp_lagrange_tiled(const double* __restrict__ A, const double* __restrict__ dEdm, const double* __restrict__ dEdS,
                     const double* __restrict__ dobs_m, const double* __restrict__ dobs_S, double dt, int stabilize,
                     double* __restrict__ lam, double* __restrict__ psi, int64_t B, int64_t N, int nseg,
                     int64_t seg_len, const typename LagrangeOps<D>::State* __restrict__ carry_in,
                     typename LagrangeOps<D>::State* __restrict__ init_out,
                     typename LagrangeOps<D>::Elem* __restrict__ seg_elem) {
    constexpr int DD = D * D, TL = T * L;
    using Ops = LagrangeOps<D>;
    extern __shared__ __align__(16) double smem[];
    double* sA = smem;               // A_t                      index i <-> t = lo + 1 + i
    double* sES = sA + TL * DD;      // dEdS_t  -> psi_{t-1}
    double* sOS = sES + TL * DD;     // dObsS_t
    double* sEm = sOS + TL * DD;     // dEdm_t  -> lam_{t-1}
    double* sOm = sEm + TL * D;      // dObsm_t
    typename Ops::Elem* wtot = reinterpret_cast<typename Ops::Elem*>(sOm + TL * D);
    const int64_t b = blockIdx.x / nseg;
    const int seg = (int)(blockIdx.x % nseg);
    const int64_t NT = N - 1;        // transitions j = 0 .. N-2
    const int64_t t0 = seg * seg_len, t1 = min(NT, t0 + seg_len);
    typename Ops::State st;
    st.a = vec_zero<D>();
    st.b = 1e-10 * mat_eye<D>();
    if (t1 == NT || NT <= 0) {
        if (threadIdx.x == 0 && seg == nseg - 1) {
            if (MODE == SWEEP_FULL) {
                vec_store<D>(lam + (b * N + N - 1) * D, st.a);
                mat_store<D>(psi + (b * N + N - 1) * DD, st.b);
            } else {
                init_out[b] = st;
            }
        }
    } else if (MODE == SWEEP_FULL) {
        st = carry_in[b * nseg + seg];
    }
    typename Ops::Elem seg_tot = Ops::identity();
    for (int64_t hi = t1; hi > t0; hi -= TL) {
        const int64_t lo = max(t0, hi - TL);
        const int cnt = (int)(hi - lo);
        __syncthreads();
        // transition j uses the inputs of step t = j + 1
        stage_in<T>(sA, A + (b * N + lo + 1) * DD, (int64_t)cnt * DD);
        stage_in<T>(sES, dEdS + (b * N + lo + 1) * DD, (int64_t)cnt * DD);
        stage_in<T>(sEm, dEdm + (b * N + lo + 1) * D, (int64_t)cnt * D);
        stage_in<T>(sOS, dobs_S + (b * (N + 1) + lo + 1) * DD, (int64_t)cnt * DD);
        stage_in<T>(sOm, dobs_m + (b * (N + 1) + lo + 1) * D, (int64_t)cnt * D);
        stage_wait();
        if (stabilize) {
            for (int i = threadIdx.x; i < cnt * DD; i += T) {
                sES[i] = vgp_fix(sES[i], kVgpClipMin, kVgpClipMax);
                sOS[i] = vgp_fix(sOS[i], kVgpClipMin, kVgpClipMax);
            }
            for (int i = threadIdx.x; i < cnt * D; i += T) {
                sEm[i] = vgp_fix(sEm[i], kVgpClipMin, kVgpClipMax);
                sOm[i] = vgp_fix(sOm[i], kVgpClipMin, kVgpClipMax);
            }
        }
        __syncthreads();
        const BwdChunk ch = bwd_chunk<L>(cnt);
        typename Ops::Elem e = Ops::identity();
        for (int k = ch.top - 1; k >= ch.bot; --k) {
            const Mat<D> At = mat_load<D>(sA + k * DD);
            typename Ops::Elem s;
            s.a.M = mat_eye<D>() - dt * At;
            s.a.v = dt * vec_load<D>(sEm + k * D) - vec_load<D>(sOm + k * D);
            s.b.R = mat_eye<D>() - (2.0 * dt) * At;
            s.b.V = dt * mat_load<D>(sES + k * DD) - mat_load<D>(sOS + k * DD);
            e = Ops::combine(e, s);
        }
        typename Ops::Elem total;
        const typename Ops::Elem excl = block_exclusive_scan<Ops, T>(e, wtot, total);
        if (MODE != SWEEP_FULL) {
            seg_tot = Ops::combine(seg_tot, total);
            continue;
        }
        typename Ops::State mine = Ops::apply(excl, st);
        st = Ops::apply(total, st);
        for (int k = ch.top - 1; k >= ch.bot; --k) {
            const Mat<D> At = mat_load<D>(sA + k * DD);
            const Mat<D> pa = mm(mine.b, At);
            const Mat<D> dpsi = (pa + pa) - mat_load<D>(sES + k * DD);
            const Vec<D> dlam = mv(At, mine.a) - vec_load<D>(sEm + k * D);
            mine.b = (mine.b - dt * dpsi) - mat_load<D>(sOS + k * DD);
            mine.a = (mine.a - dt * dlam) - vec_load<D>(sOm + k * D);
            mat_store<D>(sES + k * DD, mine.b);
            vec_store<D>(sEm + k * D, mine.a);
        }
        __syncthreads();
        stage_out<T>(psi + (b * N + lo) * DD, sES, (int64_t)cnt * DD);
        stage_out<T>(lam + (b * N + lo) * D, sEm, (int64_t)cnt * D);
    }
    if (MODE != SWEEP_FULL && threadIdx.x == 0) seg_elem[b * nseg + seg] = seg_tot;
}

template <int D, int T, int L>
constexpr size_t lagrange_tiled_smem() {
    return (3 * (size_t)T * L * D * D + 2 * (size_t)T * L * D) * sizeof(double) +
           (T / 32) * sizeof(typename LagrangeOps<D>::Elem);
}

// update_initial_statistics (vi_sde.py:242-260), one thread per sequence
template <int D>
__global__ void k_vgp_update_initial(Vec<D> p_mu, Mat<D> p_cov, double lr, const double* __restrict__ lam,
                                     const double* __restrict__ psi, double* __restrict__ q_mean,
                                     double* __restrict__ q_chol, int64_t B, int64_t N, int64_t* status) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    constexpr int DD = D * D;
    bool ok = true;
    const Mat<D> Lp = chol(p_cov, ok);
    const Vec<D> qm = p_mu - mv(mmT(Lp, Lp), vec_load<D>(lam + b * N * D));
    const Mat<D> pci = chol_inverse(Lp);
    const Mat<D> prior_cov_inv = mmT(pci, pci);   // as written in the reference (:250-251)
    const Mat<D> qc_new = lu_solve(prior_cov_inv + 2.0 * mat_load<D>(psi + b * N * DD), mat_eye<D>());
    const Mat<D> qch = mat_load<D>(q_chol + b * DD);
    const Mat<D> qc = (1.0 - lr) * mmT(qch, qch) + lr * qc_new;
    vec_store<D>(q_mean + b * D, (1.0 - lr) * vec_load<D>(q_mean + b * D) + lr * qm);
    mat_store<D>(q_chol + b * DD, chol(qc, ok));
    if (!ok) report_not_pd(status, b * (N + 1));
}

// elbo = E_obs - E_sde - KL[q(x0) || p(x0)]  (vi_sde.py:436-455); ve and esde are per-sequence sums
template <int D>
__global__ void k_vgp_elbo_combine(const double* __restrict__ ve, const double* __restrict__ esde, Vec<D> p_mu,
                                   Mat<D> p_cov, const double* __restrict__ q_mean,
                                   const double* __restrict__ q_chol, double* __restrict__ elbo, int64_t B) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    constexpr int DD = D * D;
    bool ok = true;
    const Mat<D> qch = mat_load<D>(q_chol + b * DD);
    const Mat<D> Lp = chol(p_cov, ok), Lq = chol(mmT(qch, qch), ok);
    const Mat<D> Mq = trsm(Lp, Lq);
    const Vec<D> diff = trsv(Lp, p_mu - vec_load<D>(q_mean + b * D));
    double fro = 0.0;
#pragma unroll
    for (int e = 0; e < DD; ++e) fro += Mq.a[e] * Mq.a[e];
    const double kl0 = 0.5 * (fro + dot(diff, diff) - (double)D + logdet_from_chol(Lp) - logdet_from_chol(Lq));
    elbo[b] = (ve[b] - esde[b]) - kl0;
}

}  // namespace vdp
