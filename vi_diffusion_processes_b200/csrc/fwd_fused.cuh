// One forward pass per site update for what k_mean_tiled -> k_kl_grad -> k_chol_tiled do in three
// (D = 1, polynomial drift, naturals carried with the sites held in theta, one segment per sequence):
//
//   marginal means of the posterior rebuilt after the data-site update      state_space_model.py:251
//   gradient of KL[q || p_SDE] in expectation parameters, per transition    sde_utils.py:473-547 (closed-form moments)
//   girsanov site update, applied to theta = nat(p) + g + data sites        variational_cvi_sde.py:279-299
//   banded Cholesky of the precision of the posterior that follows          block_tri_diag.py:428-440
//
// A tile (T chunks of L grid steps of one sequence) is staged once: A, b, S of the rebuilt posterior and the three
// arrays of theta (48 B per grid step).  The means never leave the CTA; theta' goes back to HBM (it is the state the
// next update starts from) together with the factor of -2 theta_d', -theta_s' (40 B per step).  The three separate
// kernels move 24 + 72 + 32 B per step.
//
// Node ownership.  Transition term k gives a "left" gradient part to grid point k and a "right" part to k + 1, so
// grid point k is final once terms k - 1 and k are known.  Thread c computes the terms of its chunk and finishes the
// points k0 + 1 .. k0 + L - 1 on the way; the right part of its last term goes to the next thread (shared memory) or,
// from the tile's last term, to the next tile (the point tile_start + cnt is the next tile's point 0 and is neither
// updated nor stored by this tile).  The Cholesky recursion therefore lags one step: tile t factorises through
// transitions [tile_start - 1, tile_start + cnt - 1), i.e. up to the last point it has finished, and the one
// transition left over at the end of the sequence is done by thread 0 after the last tile.
#pragma once
#include "tile_sweeps.cuh"

#ifndef VDP_FUSED_MINBLOCKS
#define VDP_FUSED_MINBLOCKS 4
#endif
#ifndef VDP_FUSED_UNROLL
#define VDP_FUSED_UNROLL 3   /* unroll of the term loop of a full chunk */
#endif

#ifndef VDP_FUSED_PREFETCH
#define VDP_FUSED_PREFETCH 1   /* measured: 0.83 -> 0.75 ms on the C4 shape */
#endif
// tile shape of the fused pass (defaults: the shape of the D = 1 mean sweep, so the means round the same way)
#ifndef VDP_TF1
#define VDP_TF1 VDP_T1
#endif
#ifndef VDP_LF1
#define VDP_LF1 VDP_LM1
#endif

namespace vdp {

constexpr int kFusedUnroll = VDP_FUSED_UNROLL;
constexpr int kFusedT = VDP_TF1, kFusedL = VDP_LF1;

struct FwdFusedArgs {
    DriftParams P;
    double q = 0.0, dt = 0.0, lr = 0.0;          // diffusion, drift step, girsanov step size
    double x0_mu = 0.0, x0_cov = 1.0;            // p(x0)
    const double* time_grid = nullptr;           // [N+1]
    const double *A = nullptr, *bo = nullptr, *mu0 = nullptr, *Sd = nullptr;   // rebuilt q: [B,N], [B,N], [B], [B,N+1]
    double *th1 = nullptr, *thd = nullptr, *ths = nullptr;                     // theta, updated in place
    double *Ld = nullptr, *Ls = nullptr;                                       // factor of the updated precision
    const double* lr_dev = nullptr;              // per-sequence step size on the device (the trainer loop)
    const int32_t* active = nullptr;
    int64_t* status = nullptr;
};

// doubles in front of the staged arrays: mbarrier, tile-to-tile carries, scan scratch (one 32-byte element per warp)
template <int T>
__host__ __device__ constexpr int fused_header() { return 8 + (T / 32) * 4; }
template <int T, int L>
__host__ __device__ constexpr int fused_region() { return tile_slots(T * L + 1) + 2; }   // 2 margin slots in front
template <int T, int L>
constexpr size_t fwd_fused_smem() {
    return ((size_t)fused_header<T>() + 6 * (size_t)fused_region<T, L>()) * sizeof(double);
}

// The KL-gradient terms of one chunk (transitions k0 .. k0 + len - 1 of the staged tile) and the site update of the
// grid points they finish.  FULL: len == L at compile time (no per-step guards).  m: mean at the chunk's first point
// on entry, at its last point + 1 on exit;  g0m / g0S: left part of the first term;  prm / prS: right part of the last.
template <int DRIFT, int L, bool FULL, int UNROLL>
__device__ __forceinline__ void fused_chunk_terms(const FwdFusedArgs& a, const double* __restrict__ tgp, int len,
                                                  int k0, const double* sA, const double* sB, const double* sS,
                                                  double* s1, double* sD, double* sT, double nlr, double& m,
                                                  double& g0m, double& g0S, double& prm, double& prS, int& bad) {
    double Sk = sS[k0];
    double t_lo = __ldg(tgp);
    bool ok = true;
    const int cnt = FULL ? L : len;
#pragma unroll UNROLL
    for (int t = 0; t < cnt; ++t) {
        const int j = k0 + t;
        const double t_hi = __ldg(tgp + t + 1);
        const double Ak = sA[j];
        const double mn = sB[j] + Ak * m;
        const double Sn = sS[j + 1];
        const double Qp = (t_hi - t_lo) * a.q;
        if (!(Sk > 0.0)) ok = false;
        double Lgm, LgS, Rgm, RgS, gC;
        kl_grad_closed_core<DRIFT>(a.P, Qp, a.dt, Ak, rcp_fast(Sk), m, mn, Sk, Sn, Lgm, LgS, Rgm, RgS, gC, ok);
        if (!ok) { bad = t + 1; ok = true; }
        if (t == 0) {
            g0m = Lgm;
            g0S = LgS;
        } else {
            // point j: left part of term j + right part of term j - 1;  S = eta_d - m m^T
            const double gSj = LgS + prS;
            const double gmj = (Lgm + prm) - 2.0 * (gSj * m);
            s1[j] = fma(nlr, gmj, s1[j]);
            sD[j] = fma(nlr, gSj, sD[j]);
        }
        sT[j] = fma(nlr, gC, sT[j]);
        prm = Rgm;
        prS = RgS;
        m = mn;
        Sk = Sn;
        t_lo = t_hi;
    }
}

// The lagged Cholesky recursion of one chunk from its incoming Schur complement s: factor entries into oLd / oLs.
template <int L, bool FULL>
__device__ __forceinline__ int fused_chunk_chol(const double* cKd, const double* cKs, int k0, int len, double s,
                                                double* oLd, double* oLs) {
    int badc = -1;
    if constexpr (FULL) {
        double ks[L], kd[L];
#pragma unroll
        for (int t = 0; t < L; ++t) {
            ks[t] = -cKs[k0 + t];
            kd[t] = (t < L - 1) ? -2.0 * cKd[k0 + t + 1] : 1.0;
        }
#pragma unroll
        for (int t = 0; t < L; ++t) {
            if (!(s > 0.0) && badc < 0) badc = t;
            const double r = rcp_fast(s);
            double Lk, rL;
            sqrt_rsqrt_fast(s, Lk, rL);
            oLd[k0 + t] = Lk;
            oLs[k0 + t] = div_r(ks[t], Lk, rL);
            if (t < L - 1) s = fma(-(ks[t] * ks[t]), r, kd[t]);
        }
    } else {
#pragma unroll 1
        for (int t = 0; t < len; ++t) {
            if (!(s > 0.0) && badc < 0) badc = t;
            const double ks = -cKs[k0 + t];
            const double r = rcp_fast(s);
            double Lk, rL;
            sqrt_rsqrt_fast(s, Lk, rL);
            oLd[k0 + t] = Lk;
            oLs[k0 + t] = div_r(ks, Lk, rL);
            if (t < len - 1) s = fma(-(ks * ks), r, -2.0 * cKd[k0 + t + 1]);
        }
    }
    return badc;
}

template <int DRIFT, int T, int L>
__global__ void __launch_bounds__(T, VDP_FUSED_MINBLOCKS) k_fwd_fused(FwdFusedArgs a, int64_t B, int64_t N) {
    constexpr int TL = T * L, SR = fused_region<T, L>();
    static_assert(sizeof(SchurElem<1>) <= 4 * sizeof(double) && sizeof(AffVecElem<1>) <= 4 * sizeof(double), "scan scratch");
    extern __shared__ __align__(16) double smem[];
    // header: [0] mbarrier, [2..5] right gradient part handed from tile to tile (two parities), [6..7] theta_s' of
    // the tile's last transition (two parities), [8..) scan scratch
    double* carryR = smem + 2;
    double* carryTs = smem + 6;
    void* wtot = smem + 8;
    double* rawA = smem + fused_header<T>() + 2;       // A_k            -> L_{k+1,k} (lagged)
    double* rawB = rawA + SR;                     // b_k
    double* rawS = rawB + SR;                     // S_k            -> L_kk (lagged)
    double* raw1 = rawS + SR;                     // theta_1
    double* rawD = raw1 + SR;                     // theta_d
    double* rawT = rawD + SR;                     // theta_s
    TileRing<1> ring;
    ring.init(smem);
    const int64_t b = blockIdx.x;
    const int64_t n = N + 1;
    const int ntiles = (int)((N + TL - 1) / TL);
    const double* gA = a.A + b * N;
    const double* gB = a.bo + b * N;
    const double* gS = a.Sd + b * n;
    double* g1 = a.th1 + b * n;
    double* gD = a.thd + b * n;
    double* gT = a.ths + b * N;
    auto issue = [&](int ti) {   // thread 0: start the copies of tile ti
        const int64_t tk = (int64_t)ti * TL;
        const int cnt = (int)min((int64_t)TL, N - tk);
        uint32_t bytes = tile_fetch(rawA, gA + tk, cnt, ring.bar);
        bytes += tile_fetch(rawB, gB + tk, cnt, ring.bar);
        bytes += tile_fetch(rawS, gS + tk, cnt + 1, ring.bar);
        bytes += tile_fetch(raw1, g1 + tk, cnt + 1, ring.bar);
        bytes += tile_fetch(rawD, gD + tk, cnt + 1, ring.bar);
        bytes += tile_fetch(rawT, gT + tk, cnt, ring.bar);
        mbar_arrive_expect_tx(ring.bar, bytes);
    };
    if (threadIdx.x == 0 && ntiles > 0) issue(0);
    double lr = a.lr;
    if (a.lr_dev) lr = (!a.active || a.active[b]) ? a.lr_dev[b] : 0.0;
    const double nlr = -lr;
    Vec<1> st;                 // mean at the tile's first grid point
    st.v[0] = a.mu0[b];
    Mat<1> S_tile;             // Schur complement at the grid point before the tile's first (lagged recursion)
    S_tile.a[0] = 1.0;
    const int k0 = threadIdx.x * L;
    for (int ti = 0; ti < ntiles; ++ti) {
        const int64_t tk = (int64_t)ti * TL;
        const int cnt = (int)min((int64_t)TL, N - tk);
        const bool last_tile = (tk + cnt == N);
        const int len = min(L, cnt - k0);
        if (ti > 0) {          // single buffer: refill once everybody has left the previous tile
            __syncthreads();
            if (threadIdx.x == 0) { bulk_wait_read(); issue(ti); }
        }
        ring.wait(ti);
#if VDP_FUSED_PREFETCH
        if (threadIdx.x == 0 && ti + 1 < ntiles) {   // the next tile's lines travel to L2 while this one is worked on
            const int64_t tn = tk + TL;
            const int cn2 = (int)min((int64_t)TL, N - tn);
            bulk_prefetch_l2(gA + tn, cn2);
            bulk_prefetch_l2(gB + tn, cn2);
            bulk_prefetch_l2(gS + tn + 1, cn2);
            bulk_prefetch_l2(g1 + tn + 1, cn2);
            bulk_prefetch_l2(gD + tn + 1, cn2);
            bulk_prefetch_l2(gT + tn, cn2);
        }
#endif
        double* sA = rawA + tile_off(gA + tk);
        double* sB = rawB + tile_off(gB + tk);
        double* sS = rawS + tile_off(gS + tk);
        double* s1 = raw1 + tile_off(g1 + tk);
        double* sD = rawD + tile_off(gD + tk);
        double* sT = rawT + tile_off(gT + tk);
        // ---- means: chunk element, block scan, replay (state_space_model.py:251)
        using VOps = AffVecOps<1>;
        VOps::Elem e = VOps::identity();
        if (len == L) {
#pragma unroll
            for (int t = 0; t < L; ++t) {
                AffVecElem<1> s;
                s.M.a[0] = sA[k0 + t];
                s.v.v[0] = sB[k0 + t];
                e = VOps::combine(e, s);
            }
        } else {
#pragma unroll 1
            for (int t = 0; t < len; ++t) {
                AffVecElem<1> s;
                s.M.a[0] = sA[k0 + t];
                s.v.v[0] = sB[k0 + t];
                e = VOps::combine(e, s);
            }
        }
        VOps::Elem totalV;
        const VOps::Elem exclV = block_exclusive_scan<VOps, T>(e, reinterpret_cast<VOps::Elem*>(wtot), totalV);
        const double m0 = VOps::apply(exclV, st).v[0];
        st = VOps::apply(totalV, st);
        // ---- KL gradient per transition, site update of the finished grid points
        double m = m0;
        double g0m = 0.0, g0S = 0.0;          // left part of the chunk's first term (its point waits for the neighbour)
        double prm = 0.0, prS = 0.0;          // right part of the chunk's last term
        int bad = 0;
        // (the grid times are L2 / L1 resident and shared by every sequence: read per term)
        if (len == L)
            fused_chunk_terms<DRIFT, L, true, kFusedUnroll>(a, a.time_grid + tk + k0, L, k0, sA, sB, sS, s1, sD, sT, nlr, m,
                                                            g0m, g0S, prm, prS, bad);
        else if (len > 0)
            fused_chunk_terms<DRIFT, L, false, 1>(a, a.time_grid + tk + k0, len, k0, sA, sB, sS, s1, sD, sT, nlr, m, g0m,
                                                  g0S, prm, prS, bad);
        if (bad) report_not_pd(a.status, b * n + tk + k0 + bad - 1);
        // hand the right part of the chunk's last term on: through the (consumed) slots of the chunk's last transition
        // to the next thread, or to the next tile; the sequence's last grid point is finished here
        if (len > 0) {
            if (k0 + len == cnt) {
                carryR[2 * (ti & 1)] = prm;
                carryR[2 * (ti & 1) + 1] = prS;
                carryTs[ti & 1] = sT[cnt - 1];
                if (last_tile) {
                    s1[cnt] = fma(nlr, prm - 2.0 * (prS * m), s1[cnt]);
                    sD[cnt] = fma(nlr, prS, sD[cnt]);
                }
            } else {
                sB[k0 + len - 1] = prm;
                sA[k0 + len - 1] = prS;
            }
        }
        __syncthreads();
        if (len > 0) {
            double inm, inS;
            if (threadIdx.x > 0) {
                inm = sB[k0 - 1];
                inS = sA[k0 - 1];
            } else if (ti > 0) {
                inm = carryR[2 * ((ti - 1) & 1)];
                inS = carryR[2 * ((ti - 1) & 1) + 1];
            } else {
                // KL[q(x0) || p(x0)]  (sde_utils.py:536-541)
                const double S0 = sS[0];
                if (!(S0 > 0.0) || !(a.x0_cov > 0.0)) report_not_pd(a.status, b * n);
                const double Pinv = 1.0 / a.x0_cov;
                inm = Pinv * (m0 - a.x0_mu);
                inS = 0.5 * (Pinv - 1.0 / S0);
            }
            const double gSj = g0S + inS;
            const double gmj = (g0m + inm) - 2.0 * (gSj * m0);
            s1[k0] = fma(nlr, gmj, s1[k0]);
            sD[k0] = fma(nlr, gSj, sD[k0]);
        }
        if (threadIdx.x == 0) sT[-1] = (ti > 0) ? carryTs[(ti - 1) & 1] : 0.0;
        __syncthreads();
        // ---- Cholesky of K = (-2 theta_d', -theta_s'), one step behind: local transition j' is the sequence's
        // transition tk - 1 + j' (j' = 0 of the first tile is void: K_s = 0 makes S_0 = K_d[0] whatever the carry)
        const double* cKd = sD - 1;
        const double* cKs = sT - 1;
        const SchurElem<1> eC = chol_tile_summary<1, L, true>(cKd, cKs, k0, len, -2.0, -1.0);
        SchurElem<1> totalC;
        const SchurElem<1> exclC = block_exclusive_scan<SchurOps<1>, T>(eC, reinterpret_cast<SchurElem<1>*>(wtot), totalC);
        double s = SchurOps<1>::apply(exclC, S_tile).a[0];
        S_tile = SchurOps<1>::apply(totalC, S_tile);
        // factor entries go where A and S were: slot j' of oLd / oLs <-> grid point / transition tk - 1 + j',
        // placed with the parity of the destination so that they leave as bulk stores
        double* gLd = a.Ld + b * n + tk - 1;
        double* gLs = a.Ls + b * N + tk - 1;
        double* oLd = rawS + (int)((reinterpret_cast<uintptr_t>(gLd) >> 3) & 1u);
        double* oLs = rawA + (int)((reinterpret_cast<uintptr_t>(gLs) >> 3) & 1u);
        {
            int badc = -1;
            if (len == L) badc = fused_chunk_chol<L, true>(cKd, cKs, k0, L, s, oLd, oLs);
            else if (len > 0) badc = fused_chunk_chol<L, false>(cKd, cKs, k0, len, s, oLd, oLs);
            if (badc >= 0 && (ti > 0 || k0 + badc > 0)) report_not_pd(a.status, b * n + tk - 1 + k0 + badc);
        }
        fence_async_smem();
        __syncthreads();
        const int skip = (ti == 0) ? 1 : 0;
        const int cn = cnt + (last_tile ? 1 : 0);
        tile_store<T>(g1 + tk, s1, cn);
        tile_store<T>(gD + tk, sD, cn);
        tile_store<T>(gT + tk, sT, cnt);
        tile_store<T>(gLd + skip, oLd + skip, cnt - skip);
        tile_store<T>(gLs + skip, oLs + skip, cnt - skip);
        if (threadIdx.x == 0) bulk_commit();
        if (last_tile && threadIdx.x == 0) {
            // the transition the lagged recursion has not reached: N - 1, and the last diagonal entry
            double sN = S_tile.a[0];
            const double ks = -sT[cnt - 1], kdN = -2.0 * sD[cnt];
            if (!(sN > 0.0)) report_not_pd(a.status, b * n + N - 1);
            const double Lk = sqrt(sN);
            a.Ld[b * n + N - 1] = Lk;
            a.Ls[b * N + N - 1] = ks / Lk;
            sN = kdN - (ks * ks) / sN;
            if (!(sN > 0.0)) report_not_pd(a.status, b * n + N);
            a.Ld[b * n + N] = sqrt(sN);
        }
    }
    if (threadIdx.x == 0) bulk_wait_all();
}

}  // namespace vdp
