// Fused-level entry points of the C ABI (included at the end of vdp_all.inl; one translation unit so the
// constant-memory quadrature tables exist once).

namespace vdp {

// launch configuration for a quadrature kernel: G lanes per grid point
template <int D> struct QuadGroup { static constexpr int value = (D == 1) ? 1 : 32; };

template <int D>
static int launch_chol_nat(const double* thd, const double* ths, double* Ld, double* Ls, int64_t B, int64_t n,
                           int64_t* status, cudaStream_t st) {
    k_chol_fwd<D><<<grid_for(B, kSeqBlock), kSeqBlock, 0, st>>>(thd, ths, -2.0, -1.0, Ld, Ls, B, n, status);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// naturals -> SSM (F1 + B1)
template <int D>
static int run_nat2ssm(const double* th1, const double* thd, const double* ths, double* A, double* b, double* cholQ,
                       double* mu0, double* cholP0, double* Ld, double* Ls, int64_t B, int64_t N, int64_t* status,
                       cudaStream_t st, SegScratch sc = SegScratch(nullptr, 0), double* Sd_tiled = nullptr,
                       bool factor_ready = false) {
    int rc;
    if (use_tiled_sweeps(B, N)) {
        // factor_ready: (Ld, Ls) already hold the factor of (-2 thd, -ths) (the fused forward pass of the site update)
        if (!factor_ready) {
            prof_mark(VDP_PHASE_CHOL_FWD, st);
            if ((rc = run_chol_tiled<D>(thd, ths, -2.0, -1.0, Ld, Ls, B, N, status, sc, st))) return rc;
        }
        prof_mark(VDP_PHASE_NAT2SSM_BWD, st);
        // Sd_tiled (tiled route only): marginal covariances of the Gaussian the naturals describe, written on the way
        return run_nat2ssm_bwd_tiled<D>(Ld, Ls, th1, thd, ths, A, b, cholQ, mu0, cholP0, B, N, status, sc, st, Sd_tiled);
    }
    prof_mark(VDP_PHASE_CHOL_FWD, st);
    rc = launch_chol_nat<D>(thd, ths, Ld, Ls, B, N + 1, status, st);
    if (rc) return rc;
    prof_mark(VDP_PHASE_NAT2SSM_BWD, st);
    k_nat2ssm_bwd<D><<<grid_for(B, kSeqBlock), kSeqBlock, 0, st>>>(Ld, Ls, th1, thd, ths, A, b, cholQ, mu0, cholP0,
                                                                   B, N, status);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// SSM -> marginals (F2 + B2)
template <int D>
static int run_marginals(const double* A, const double* b, const double* cholQ, const double* mu0,
                         const double* cholP0, double* m, double* Sd, double* Ld, double* Ls, int64_t B, int64_t N,
                         int64_t* status, cudaStream_t st, SegScratch sc = SegScratch(nullptr, 0)) {
    if (use_tiled_sweeps(B, N)) {
        int rc;
        prof_mark(VDP_PHASE_MARG_FWD, st);
        if ((rc = run_marg_fwd_tiled<D>(A, b, cholQ, mu0, cholP0, m, Ld, Ls, B, N, status, sc, st))) return rc;
        prof_mark(VDP_PHASE_INVBAND_BWD, st);
        if ((rc = run_invband_tiled<D>(Ld, Ls, Sd, nullptr, B, N, sc, st))) return rc;
        prof_mark(VDP_PHASE_END, st);
        return VDP_OK;
    }
    prof_mark(VDP_PHASE_MARG_FWD, st);
    k_marg_fwd<D><<<grid_for(B, kSeqBlock), kSeqBlock, 0, st>>>(A, b, cholQ, mu0, cholP0, m, Ld, Ls, B, N, status);
    VDP_CHECK_LAUNCH();
    prof_mark(VDP_PHASE_INVBAND_BWD, st);
    k_invband_bwd<D><<<grid_for(B, kSeqBlock), kSeqBlock, 0, st>>>(Ld, Ls, Sd, nullptr, B, N + 1);
    VDP_CHECK_LAUNCH();
    prof_mark(VDP_PHASE_END, st);
    return VDP_OK;
}

}  // namespace vdp

#if VDP_EMIT_CORE
extern "C" {

int vdp_gh_expect(int drift_id, const double* params_host, const double* m, const double* S, int H, double* Ef,
                  double* EJ, int64_t count, int D, int64_t* status, void* stream) {
    VDP_REQUIRE(m && S && count >= 0, "m, S, count");
    VDP_REQUIRE(H == 10 || H == 20, "H must be 10 or 20 (the two rules the reference uses)");
    if (count == 0) return VDP_OK;
    const DriftParams P = make_params(params_host);
    const int fj = drift_full_jacobian(drift_id) ? 1 : 0;
    VDP_DISPATCH_D(D, VDP_DISPATCH_DRIFT(drift_id, {
        constexpr int G = QuadGroup<kD>::value;
        const unsigned grid = grid_for(count * G, kPtBlock);
        if (H == 10)
            k_gh_expect<kDrift, kD, 10, G><<<grid, kPtBlock, 0, as_stream(stream)>>>(P, m, S, pi_factor(kD), fj, Ef, EJ, count, status);
        else
            k_gh_expect<kDrift, kD, 20, G><<<grid, kPtBlock, 0, as_stream(stream)>>>(P, m, S, pi_factor(kD), fj, Ef, EJ, count, status);
    }));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_linearize_to_ssm(int drift_id, const double* params_host, const double* q_host, const double* time_grid,
                         const double* m, const double* S, double* A, double* b, double* cholQ, double* A_lin,
                         double* b_lin, int clip, double clip_lo, double clip_hi, int64_t B, int64_t N, int D,
                         int64_t* status, void* stream) {
    VDP_REQUIRE(q_host && time_grid && m && S && A && b && cholQ && B >= 0 && N >= 1, "pointers, B, N");
    if (B == 0) return VDP_OK;
    const DriftParams P = make_params(params_host);
    const int fj = drift_full_jacobian(drift_id) ? 1 : 0;
    VDP_DISPATCH_D(D, VDP_DISPATCH_DRIFT(drift_id, {
        constexpr int G = QuadGroup<kD>::value;
        bool done = false;
        // polynomial drift in D >= 2: E f and E J through the exact low-order rule (DriftPoly / ExactOrder in
        // drifts.cuh), one thread per grid point; the GH_QUADRATURE switch keeps the reference's 10 points
        if constexpr (kD >= 2 && DriftPoly<kDrift>::degree > 0) {
            if (!g_cfg.gh_quadrature) {
                k_linearize<kDrift, kD, ExactOrder<kDrift>::H, 1><<<grid_for(B * N, kPtBlock), kPtBlock, 0,
                                                                   as_stream(stream)>>>(
                    P, host_mat<kD>(q_host), time_grid, m, S, pi_factor(kD), fj, A, b, cholQ, A_lin, b_lin, clip,
                    clip_lo, clip_hi, B, N, status);
                done = true;
            }
        }
        if (!done)
            k_linearize<kDrift, kD, 10, G><<<grid_for(B * N * G, kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                P, host_mat<kD>(q_host), time_grid, m, S, pi_factor(kD), fj, A, b, cholQ, A_lin, b_lin, clip, clip_lo,
                clip_hi, B, N, status);
    }));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_ssm_to_naturals(const double* A, const double* b, const double* cholQ, const double* mu0,
                        const double* cholP0, double* th1, double* thd, double* ths, int64_t B, int64_t N, int D,
                        void* stream) {
    VDP_REQUIRE(A && b && cholQ && mu0 && cholP0 && th1 && thd && ths && B >= 0 && N >= 1, "pointers, B, N");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_ssm_to_naturals<kD><<<grid_for(B * (N + 1), kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                          A, b, cholQ, mu0, cholP0, nullptr, nullptr, nullptr, th1, thd, ths, B, N)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

size_t vdp_naturals_to_ssm_workspace(int64_t B, int64_t N, int D) {
    return padded((size_t)B * (N + 1) * D * D) + padded((size_t)B * N * D * D) + kSegScratchBytes;
}

int vdp_naturals_to_ssm(const double* th1, const double* thd, const double* ths, double* A, double* b,
                        double* cholQ, double* mu0, double* cholP0, void* workspace, size_t workspace_bytes,
                        int64_t B, int64_t N, int D, int64_t* status, void* stream) {
    VDP_REQUIRE(th1 && thd && ths && A && b && cholQ && mu0 && cholP0 && B >= 0 && N >= 1, "pointers, B, N");
    if (B == 0) return VDP_OK;
    Arena ar(workspace, workspace_bytes);
    double* Ld = ar.take((size_t)B * (N + 1) * D * D);
    double* Ls = ar.take((size_t)B * N * D * D);
    if (!workspace || !Ld || !Ls) { set_error("workspace too small"); return VDP_E_WORKSPACE; }
    SegScratch sc(ar.take_bytes(kSegScratchBytes), kSegScratchBytes);
    VDP_DISPATCH_D(D, { int rc = run_nat2ssm<kD>(th1, thd, ths, A, b, cholQ, mu0, cholP0, Ld, Ls, B, N, status,
                                                 as_stream(stream), sc); if (rc) return rc; });
    return VDP_OK;
}

size_t vdp_ssm_marginals_workspace(int64_t B, int64_t N, int D) { return vdp_naturals_to_ssm_workspace(B, N, D); }

int vdp_ssm_marginals(const double* A, const double* b, const double* cholQ, const double* mu0,
                      const double* cholP0, double* m, double* Sd, void* workspace, size_t workspace_bytes,
                      int64_t B, int64_t N, int D, int64_t* status, void* stream) {
    VDP_REQUIRE(A && b && cholQ && mu0 && cholP0 && m && Sd && B >= 0 && N >= 1, "pointers, B, N");
    if (B == 0) return VDP_OK;
    Arena ar(workspace, workspace_bytes);
    double* Ld = ar.take((size_t)B * (N + 1) * D * D);
    double* Ls = ar.take((size_t)B * N * D * D);
    if (!workspace || !Ld || !Ls) { set_error("workspace too small"); return VDP_E_WORKSPACE; }
    SegScratch sc(ar.take_bytes(kSegScratchBytes), kSegScratchBytes);
    VDP_DISPATCH_D(D, { int rc = run_marginals<kD>(A, b, cholQ, mu0, cholP0, m, Sd, Ld, Ls, B, N, status,
                                                   as_stream(stream), sc); if (rc) return rc; });
    return VDP_OK;
}

int vdp_ssm_to_expectations(const double* A, const double* m, const double* Sd, double* e1, double* ed, double* es,
                            int64_t B, int64_t N, int D, void* stream) {
    VDP_REQUIRE(A && m && Sd && e1 && ed && es && B >= 0 && N >= 1, "pointers, B, N");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_ssm_to_expectations<kD><<<grid_for(B * (N + 1), kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                          A, m, Sd, e1, ed, es, B, N)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_expectations_to_ssm(const double* e1, const double* ed, const double* es, double* A, double* b,
                            double* cholQ, double* mu0, double* cholP0, int64_t B, int64_t N, int D,
                            int64_t* status, void* stream) {
    VDP_REQUIRE(e1 && ed && es && A && b && cholQ && mu0 && cholP0 && B >= 0 && N >= 1, "pointers, B, N");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_expectations_to_ssm<kD><<<grid_for(B * N, kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                          e1, ed, es, A, b, cholQ, mu0, cholP0, B, N, status)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}


int vdp_expectations_to_ssm_vjp(const double* e1, const double* ed, const double* es, const double* gA,
                                const double* gb, const double* gcholP0, const double* gcholQ, const double* gmu0,
                                double* g1, double* gd, double* gs, int64_t B, int64_t N, int D, int64_t* status,
                                void* stream) {
    VDP_REQUIRE(e1 && ed && es && gA && gb && gcholP0 && gcholQ && gmu0 && g1 && gd && gs && B >= 0 && N >= 1,
                "pointers, B, N");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_expectations_to_ssm_vjp<kD><<<grid_for(B * (N + 1), kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                          e1, ed, es, gA, gb, gcholP0, gcholQ, gmu0, nullptr, nullptr, nullptr, 0.0, g1, gd, gs, B, N,
                          status)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

size_t vdp_ssm_natgrad_workspace(int64_t B, int64_t N, int D) {
    const size_t n = (size_t)B * (N + 1), t = (size_t)B * N, DD = (size_t)D * D;
    return 2 * (padded(n * D) + padded(n * DD) + padded(t * DD)) + padded(n * D) + padded(n * DD) + padded(n * DD) +
           padded(t * DD) + kSegScratchBytes;
}

// SSMNaturalGradient._natgrad_step without momentum (ssm_natgrad.py:121-218):
//   eta = ssm_to_expectations(ssm); dL/d eta = VJP of expectations_to_ssm_params at eta with dL/d(ssm params);
//   theta = ssm_to_naturals(ssm); theta <- theta - gamma dL/d eta; ssm <- naturals_to_ssm_params(theta)
int vdp_ssm_natgrad_step(double* A, double* b, double* cholP0, double* cholQ, double* mu0, const double* dL_dA,
                         const double* dL_db, const double* dL_dcholP0, const double* dL_dcholQ,
                         const double* dL_dmu0, double gamma, void* workspace, size_t workspace_bytes, int64_t B,
                         int64_t N, int D, int64_t* status, void* stream) {
    VDP_REQUIRE(A && b && cholP0 && cholQ && mu0 && dL_dA && dL_db && dL_dcholP0 && dL_dcholQ && dL_dmu0 && B >= 0 &&
                    N >= 1, "pointers, B, N");
    if (B == 0) return VDP_OK;
    Arena ar(workspace, workspace_bytes);
    const size_t n = (size_t)B * (N + 1), t = (size_t)B * N, DD = (size_t)D * D;
    double* e1 = ar.take(n * D);
    double* edd = ar.take(n * DD);
    double* ess = ar.take(t * DD);
    double* th1 = ar.take(n * D);
    double* thd = ar.take(n * DD);
    double* ths = ar.take(t * DD);
    double* m = ar.take(n * D);
    double* Sd = ar.take(n * DD);
    double* Ld = ar.take(n * DD);
    double* Ls = ar.take(t * DD);
    void* seg = ar.take_bytes(kSegScratchBytes);
    if (!workspace || !seg) { set_error("workspace too small"); return VDP_E_WORKSPACE; }
    const SegScratch sc(seg, kSegScratchBytes);
    cudaStream_t st = as_stream(stream);
    VDP_DISPATCH_D(D, {
        int rc = run_marginals<kD>(A, b, cholQ, mu0, cholP0, m, Sd, Ld, Ls, B, N, status, st, sc);
        if (rc) return rc;
        k_ssm_to_expectations<kD><<<grid_for(B * (N + 1), kPtBlock), kPtBlock, 0, st>>>(A, m, Sd, e1, edd, ess, B, N);
        k_ssm_to_naturals<kD><<<grid_for(B * (N + 1), kPtBlock), kPtBlock, 0, st>>>(A, b, cholQ, mu0, cholP0, nullptr,
                                                                                  nullptr, nullptr, th1, thd, ths, B, N);
        k_expectations_to_ssm_vjp<kD><<<grid_for(B * (N + 1), kPtBlock), kPtBlock, 0, st>>>(
            e1, edd, ess, dL_dA, dL_db, dL_dcholP0, dL_dcholQ, dL_dmu0, th1, thd, ths, gamma, th1, thd, ths, B, N, status);
        VDP_CHECK_LAUNCH();
        rc = run_nat2ssm<kD>(th1, thd, ths, A, b, cholQ, mu0, cholP0, Ld, Ls, B, N, status, st, sc);
        if (rc) return rc;
    });
    return VDP_OK;
}

}  // extern "C"
#endif  // VDP_EMIT_CORE
