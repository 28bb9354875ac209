// VDP entry points (VariationalMarkovGP, vi_sde.py): forward pass, Lagrange sweep, parameter update,
// initial-state update, ELBO and the fused trainer iteration.  Included at the end of vdp_all.inl.

namespace vdp {

template <int D> struct LagrangeCfg;
template <> struct LagrangeCfg<1> { static constexpr int T = 128, L = 9; };
template <> struct LagrangeCfg<2> { static constexpr int T = 64, L = 5; };
template <> struct LagrangeCfg<3> { static constexpr int T = 64, L = 3; };

struct VgpWs {
    double *sA, *sb, *scQ, *Ld, *Ls, *terms, *dEdm, *dEdS, *dobs_m, *dobs_S, *ve, *esde;
    void* seg;
};
static inline size_t vgp_ws_bytes(int64_t B, int64_t N, int D) {
    const size_t n = (size_t)B * (N + 1), t = (size_t)B * N, DD = (size_t)D * D;
    return padded(t * DD) + padded(t * D) + padded(t * DD) + padded(n * DD) + padded(t * DD) + padded(t) +
           padded(t * D) + padded(t * DD) + padded(n * D) + padded(n * DD) + 2 * padded((size_t)B) + kSegScratchBytes;
}
static inline bool vgp_ws_take(void* ws, size_t bytes, int64_t B, int64_t N, int D, VgpWs& w) {
    if (!ws) return false;
    Arena ar(ws, bytes);
    const size_t n = (size_t)B * (N + 1), t = (size_t)B * N, DD = (size_t)D * D;
    w.sA = ar.take(t * DD);
    w.sb = ar.take(t * D);
    w.scQ = ar.take(t * DD);
    w.Ld = ar.take(n * DD);
    w.Ls = ar.take(t * DD);
    w.terms = ar.take(t);
    w.dEdm = ar.take(t * D);
    w.dEdS = ar.take(t * DD);
    w.dobs_m = ar.take(n * D);
    w.dobs_S = ar.take(n * DD);
    w.ve = ar.take((size_t)B);
    w.esde = ar.take((size_t)B);
    w.seg = ar.take_bytes(kSegScratchBytes);
    return w.seg != nullptr;
}

static int check_vgp(const vdp_vgp_problem* p, VgpWs& w) {
    VDP_REQUIRE(p != nullptr, "problem");
    VDP_REQUIRE(p->B >= 0 && p->N >= 1 && p->n_obs >= 0, "B, N, n_obs");
    VDP_REQUIRE(p->time_grid && p->A && p->b && p->lam && p->psi && p->q_mean && p->q_chol && p->m && p->S && p->elbo,
                "pointers");
    VDP_REQUIRE(p->n_obs == 0 || (p->obs_idx && p->obs_y), "observation pointers");
    if (p->B > 0 && !vgp_ws_take(p->workspace, p->workspace_bytes, p->B, p->N, p->D, w)) {
        set_error("workspace too small: need %zu bytes", vgp_ws_bytes(p->B, p->N, p->D));
        return VDP_E_WORKSPACE;
    }
    return VDP_OK;
}

// forward_pass as one fused sweep (k_vgp_forward_tiled): to_ssm, means and the covariance recursion
// DRIFT >= 0 (D = 1, polynomial drift) with esde_parts != NULL: the FULL pass also evaluates the E_sde terms and their
// gradients (w.dEdm, w.dEdS) and leaves *esde_parts partial sums of the terms per sequence in w.terms.
template <int D, int DRIFT = -1>
static int run_vgp_forward_tiled(const vdp_vgp_problem& p, const VgpWs& w, cudaStream_t st,
                                 int64_t* esde_parts = nullptr) {
    using C = TileCfg<D>;
    using Ops = FwdOps<D>;
    constexpr int L = C::Lm;
    constexpr int NST = 1;
    constexpr bool ES = DRIFT >= 0;
    constexpr size_t smem = vgp_forward_tiled_smem<D, C::T, L, NST>();
    constexpr size_t smem_full = vgp_forward_tiled_smem<D, C::T, L, NST, ES>();
    const int64_t B = p.B, N = p.N;
    SegScratch sc(w.seg, kSegScratchBytes);
    const SegPlan pl = plan_segments(B, N, C::T * L, true);
    const unsigned grid = (unsigned)(B * pl.nseg);
    const int stab = (p.flags & VDP_VGP_STABILIZE) ? 1 : 0;
    VgpEsdeArgs es;
    if (ES) {
        es.P = make_params(p.drift_params);
        es.q = p.q[0]; es.dt = p.dt;
        es.dEdm = w.dEdm; es.dEdS = w.dEdS; es.partials = w.terms;
        es.tiles_per_seg = (int)((pl.seg_len + C::T * L - 1) / (C::T * L));
        *esde_parts = (int64_t)pl.nseg * es.tiles_per_seg;
    }
    int rc;
    if ((rc = allow_smem(k_vgp_forward_tiled<D, C::T, L, SWEEP_FULL, NST, DRIFT>, smem_full))) return rc;
    if (pl.nseg == 1) {
        k_vgp_forward_tiled<D, C::T, L, SWEEP_FULL, NST, DRIFT><<<grid, C::T, smem_full, st>>>(
            p.A, p.b, host_mat<D>(p.q), p.time_grid, stab, p.q_mean, p.q_chol, p.m, p.S, B, N, 1, pl.seg_len, nullptr,
            nullptr, nullptr, p.status, es);
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    typename Ops::State* carry = sc.take<typename Ops::State>((size_t)B * pl.nseg);
    typename Ops::State* init = sc.take<typename Ops::State>((size_t)B);
    typename Ops::Elem* el = sc.take<typename Ops::Elem>((size_t)B * pl.nseg);
    if (!el) { set_error("segment scratch too small"); return VDP_E_WORKSPACE; }
    if ((rc = allow_smem(k_vgp_forward_tiled<D, C::T, L, SWEEP_SUMMARY, NST>, smem))) return rc;
    k_vgp_forward_tiled<D, C::T, L, SWEEP_SUMMARY, NST><<<grid, C::T, smem, st>>>(
        p.A, p.b, host_mat<D>(p.q), p.time_grid, stab, p.q_mean, p.q_chol, p.m, p.S, B, N, pl.nseg, pl.seg_len, nullptr,
        init, el, p.status);
    VDP_CHECK_LAUNCH();
    if ((rc = launch_seg_scan<Ops>(el, init, carry, B, pl.nseg, 0, st))) return rc;
    k_vgp_forward_tiled<D, C::T, L, SWEEP_FULL, NST, DRIFT><<<grid, C::T, smem_full, st>>>(
        p.A, p.b, host_mat<D>(p.q), p.time_grid, stab, p.q_mean, p.q_chol, p.m, p.S, B, N, pl.nseg, pl.seg_len, carry,
        nullptr, nullptr, p.status, es);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// vi_sde.py:171-204
template <int D>
static int vgp_forward(const vdp_vgp_problem& p, const VgpWs& w, cudaStream_t st) {
    const int64_t B = p.B, N = p.N;
    // tiled route: one fused sweep (covariances by the forward recursion); VDP_VGP_REFACTOR_PRECISION / the
    // REFACTOR_PRECISION switch keep the reference's route through the rebuilt precision
    if (use_tiled_sweeps(B, N) && !(p.flags & VDP_VGP_REFACTOR_PRECISION) && !g_cfg.refactor_precision)
        return run_vgp_forward_tiled<D>(p, w, st);
    k_vgp_to_ssm<D><<<grid_for(B * N, kPtBlock), kPtBlock, 0, st>>>(p.A, p.b, -1.0, host_mat<D>(p.q), p.time_grid,
                                                                    (p.flags & VDP_VGP_STABILIZE) ? 1 : 0, w.sA, w.sb,
                                                                    w.scQ, B, N, p.status);
    VDP_CHECK_LAUNCH();
    return run_marginals<D>(w.sA, w.sb, w.scQ, p.q_mean, p.q_chol, p.m, p.S, w.Ld, w.Ls, B, N, p.status, st,
                            SegScratch(w.seg, kSegScratchBytes));
}

template <int DRIFT, int D>
static int vgp_esde(const vdp_vgp_problem& p, const VgpWs& w, bool grads, cudaStream_t st) {
    constexpr int G = QuadGroup<D>::value;
    const DriftParams P = make_params(p.drift_params);
    if constexpr (D == 1 && G == 1 && DriftStep1<DRIFT>::kCubic) {
        if (!(p.flags & VDP_VGP_GH_QUADRATURE) && !g_cfg.gh_quadrature) {
            k_vgp_esde<DRIFT, D, 20, G, true><<<grid_for(p.B * p.N * G, kPtBlock), kPtBlock, 0, st>>>(
                P, host_mat<D>(p.q), p.dt, pi_factor(D), p.A, p.b, p.m, p.S, w.terms, grads ? w.dEdm : nullptr,
                grads ? w.dEdS : nullptr, p.B, p.N, p.status);
            VDP_CHECK_LAUNCH();
            return VDP_OK;
        }
    }
    // polynomial drift in D >= 2: the exact low-order rule, one thread per grid point (drifts.cuh DriftPoly)
    if constexpr (D >= 2 && DriftPoly<DRIFT>::degree > 0) {
        if (!(p.flags & VDP_VGP_GH_QUADRATURE) && !g_cfg.gh_quadrature) {
            k_vgp_esde<DRIFT, D, ExactOrder<DRIFT>::H, 1><<<grid_for(p.B * p.N, kPtBlock), kPtBlock, 0, st>>>(
                P, host_mat<D>(p.q), p.dt, pi_factor(D), p.A, p.b, p.m, p.S, w.terms, grads ? w.dEdm : nullptr,
                grads ? w.dEdS : nullptr, p.B, p.N, p.status);
            VDP_CHECK_LAUNCH();
            return VDP_OK;
        }
    }
    k_vgp_esde<DRIFT, D, 20, G><<<grid_for(p.B * p.N * G, kPtBlock), kPtBlock, 0, st>>>(
        P, host_mat<D>(p.q), p.dt, pi_factor(D), p.A, p.b, p.m, p.S, w.terms, grads ? w.dEdm : nullptr,
        grads ? w.dEdS : nullptr, p.B, p.N, p.status);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// vi_sde.py:289-347
// forward pass + E_sde terms and gradients.  *parts > 0: the forward sweep did both (partial sums of the terms in
// w.terms); *parts = 0: the two separate passes ran and w.terms holds one term per transition.
template <int DRIFT, int D>
static int vgp_forward_esde(const vdp_vgp_problem& p, const VgpWs& w, cudaStream_t st, int64_t* parts) {
    *parts = 0;
    int rc;
    if constexpr (D == 1 && DriftStep1<DRIFT>::kCubic) {
        if (use_tiled_sweeps(p.B, p.N) && !(p.flags & (VDP_VGP_REFACTOR_PRECISION | VDP_VGP_GH_QUADRATURE)) &&
            !g_cfg.refactor_precision && !g_cfg.gh_quadrature && !g_cfg.no_fused_forward)
            return run_vgp_forward_tiled<D, DRIFT>(p, w, st, parts);
    }
    if ((rc = vgp_forward<D>(p, w, st))) return rc;
    return vgp_esde<DRIFT, D>(p, w, true, st);
}

// have_grads: w.dEdm / w.dEdS already hold dE_sde/dm, dE_sde/dS of the current (A, b, m, S)
template <int DRIFT, int D>
// jumps_sparse: the jump buffers are zero away from the observed grid points (left so by the previous sweep)
static int vgp_lagrange(const vdp_vgp_problem& p, const VgpWs& w, cudaStream_t st, bool have_grads = false,
                        bool jumps_sparse = false) {
    using C = LagrangeCfg<D>;
    using Ops = LagrangeOps<D>;
    const int64_t B = p.B, N = p.N;
    const int stab = (p.flags & VDP_VGP_STABILIZE) ? 1 : 0;
    int rc;
    if (!have_grads && (rc = vgp_esde<DRIFT, D>(p, w, true, st))) return rc;
    const size_t n = (size_t)B * (N + 1);
    if (!jumps_sparse) {
        VDP_CHECK_CUDA(cudaMemsetAsync(w.dobs_m, 0, n * D * sizeof(double), st));
        VDP_CHECK_CUDA(cudaMemsetAsync(w.dobs_S, 0, n * D * D * sizeof(double), st));
    } else if (p.n_obs > 0) {
        k_vgp_jump_clear<D><<<grid_for(B * p.n_obs, kPtBlock), kPtBlock, 0, st>>>(p.obs_idx, w.dobs_m, w.dobs_S, B, N,
                                                                                  p.n_obs);
        VDP_CHECK_LAUNCH();
    }
    if (p.n_obs > 0) {
        k_vgp_jump<D><<<grid_for(B * p.n_obs, kPtBlock), kPtBlock, 0, st>>>(p.m, p.obs_idx, p.obs_y,
                                                                            host_mat<D>(p.lik_chol), w.dobs_m,
                                                                            w.dobs_S, B, N, p.n_obs);
        VDP_CHECK_LAUNCH();
    }
    constexpr size_t smem = lagrange_tiled_smem<D, C::T, C::L>();
    SegPlan pl;
    pl.nseg = 1;
    pl.seg_len = std::max<int64_t>(N - 1, 1);
    if (N - 1 >= 2) pl = plan_segments(B, N - 1, C::T * C::L, true);
    const unsigned grid = (unsigned)(B * pl.nseg);
    if ((rc = allow_smem(k_vgp_lagrange_tiled<D, C::T, C::L, SWEEP_FULL>, smem))) return rc;
    if (pl.nseg == 1) {
        k_vgp_lagrange_tiled<D, C::T, C::L, SWEEP_FULL><<<grid, C::T, smem, st>>>(
            p.A, w.dEdm, w.dEdS, w.dobs_m, w.dobs_S, p.dt, stab, p.lam, p.psi, B, N, 1, pl.seg_len, nullptr, nullptr,
            nullptr);
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    SegScratch sc(w.seg, kSegScratchBytes);
    typename Ops::State* carry = sc.take<typename Ops::State>((size_t)B * pl.nseg);
    typename Ops::State* init = sc.take<typename Ops::State>((size_t)B);
    typename Ops::Elem* el = sc.take<typename Ops::Elem>((size_t)B * pl.nseg);
    if (!el) { set_error("segment scratch too small"); return VDP_E_WORKSPACE; }
    if ((rc = allow_smem(k_vgp_lagrange_tiled<D, C::T, C::L, SWEEP_SUMMARY>, smem))) return rc;
    k_vgp_lagrange_tiled<D, C::T, C::L, SWEEP_SUMMARY><<<grid, C::T, smem, st>>>(
        p.A, w.dEdm, w.dEdS, w.dobs_m, w.dobs_S, p.dt, stab, p.lam, p.psi, B, N, pl.nseg, pl.seg_len, nullptr, init, el);
    VDP_CHECK_LAUNCH();
    if ((rc = launch_seg_scan<Ops>(el, init, carry, B, pl.nseg, 1, st))) return rc;
    k_vgp_lagrange_tiled<D, C::T, C::L, SWEEP_FULL><<<grid, C::T, smem, st>>>(
        p.A, w.dEdm, w.dEdS, w.dobs_m, w.dobs_S, p.dt, stab, p.lam, p.psi, B, N, pl.nseg, pl.seg_len, carry, nullptr,
        nullptr);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// vi_sde.py:377-414
template <int DRIFT, int D>
static int vgp_update_param(const vdp_vgp_problem& p, cudaStream_t st) {
    constexpr int G = QuadGroup<D>::value;
    const DriftParams P = make_params(p.drift_params);
    if constexpr (D == 1 && G == 1 && DriftStep1<DRIFT>::kCubic) {
        if (!(p.flags & VDP_VGP_GH_QUADRATURE) && !g_cfg.gh_quadrature) {
            k_vgp_update_param<DRIFT, D, 10, G, true><<<grid_for(p.B * p.N * G, kPtBlock), kPtBlock, 0, st>>>(
                P, host_mat<D>(p.q), pi_factor(D), (p.flags & VDP_VGP_STABILIZE) ? 1 : 0, p.lr, p.m, p.S, p.lam, p.psi,
                p.A, p.b, p.B, p.N, p.status);
            VDP_CHECK_LAUNCH();
            return VDP_OK;
        }
    }
    if constexpr (D >= 2 && DriftPoly<DRIFT>::degree > 0) {
        if (!(p.flags & VDP_VGP_GH_QUADRATURE) && !g_cfg.gh_quadrature) {
            k_vgp_update_param<DRIFT, D, ExactOrder<DRIFT>::H, 1><<<grid_for(p.B * p.N, kPtBlock), kPtBlock, 0, st>>>(
                P, host_mat<D>(p.q), pi_factor(D), (p.flags & VDP_VGP_STABILIZE) ? 1 : 0, p.lr, p.m, p.S, p.lam, p.psi,
                p.A, p.b, p.B, p.N, p.status);
            VDP_CHECK_LAUNCH();
            return VDP_OK;
        }
    }
    k_vgp_update_param<DRIFT, D, 10, G><<<grid_for(p.B * p.N * G, kPtBlock), kPtBlock, 0, st>>>(
        P, host_mat<D>(p.q), pi_factor(D), (p.flags & VDP_VGP_STABILIZE) ? 1 : 0, p.lr, p.m, p.S, p.lam, p.psi, p.A,
        p.b, p.B, p.N, p.status);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// vi_sde.py:242-260
template <int D>
static int vgp_update_initial(const vdp_vgp_problem& p, cudaStream_t st) {
    k_vgp_update_initial<D><<<grid_for(p.B, kPtBlock), kPtBlock, 0, st>>>(host_vec<D>(p.x0_mu), host_mat<D>(p.x0_cov),
                                                                          p.x0_lr, p.lam, p.psi, p.q_mean, p.q_chol,
                                                                          p.B, p.N, p.status);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// vi_sde.py:436-455 (m, S of the problem must be the current forward pass)
template <int DRIFT, int D>
// terms_ready >= 0: w.terms already holds the E_sde terms (terms_ready = 0: one per transition; > 0: that many partial
// sums per sequence, from the fused forward sweep)
static int vgp_elbo(const vdp_vgp_problem& p, const VgpWs& w, cudaStream_t st, int64_t terms_ready = -1) {
    int rc;
    if (terms_ready < 0 && (rc = vgp_esde<DRIFT, D>(p, w, false, st))) return rc;
    k_reduce_terms<<<(unsigned)p.B, 256, 0, st>>>(w.terms, terms_ready > 0 ? terms_ready : p.N, w.esde);
    VDP_CHECK_LAUNCH();
    if (p.n_obs > 0) {
        k_variational_expectation<D><<<grid_for(p.B * 32, kPtBlock), kPtBlock, 0, st>>>(
            p.m, p.S, p.obs_idx, p.obs_y, host_mat<D>(p.lik_chol), w.ve, p.B, p.N, p.n_obs);
        VDP_CHECK_LAUNCH();
    } else {
        VDP_CHECK_CUDA(cudaMemsetAsync(w.ve, 0, (size_t)p.B * sizeof(double), st));
    }
    k_vgp_elbo_combine<D><<<grid_for(p.B, kPtBlock), kPtBlock, 0, st>>>(w.ve, w.esde, host_vec<D>(p.x0_mu),
                                                                        host_mat<D>(p.x0_cov), p.q_mean, p.q_chol,
                                                                        p.elbo, p.B);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

}  // namespace vdp

#define VDP_VGP_ENTRY(NAME, BODY)                                                                   \
    int NAME(const vdp_vgp_problem* p, void* stream) {                                              \
        VgpWs w;                                                                                    \
        int rc = check_vgp(p, w);                                                                   \
        if (rc) return rc;                                                                          \
        if (p->B == 0) return VDP_OK;                                                               \
        cudaStream_t st = as_stream(stream);                                                        \
        (void)st;                                                                                   \
        VDP_DISPATCH_D(p->D, VDP_DISPATCH_DRIFT(p->drift_id, BODY));                                \
        return VDP_OK;                                                                              \
    }

extern "C" {

int vdp_linear_drift_to_ssm(const double* A, const double* b, const double* q_host, const double* time_grid,
                            double* ssm_A, double* ssm_b, double* ssm_cholQ, int64_t B, int64_t N, int D,
                            int64_t* status, void* stream) {
    VDP_REQUIRE(A && b && q_host && time_grid && ssm_A && ssm_b && ssm_cholQ && B >= 0 && N >= 1, "pointers, B, N");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_vgp_to_ssm<kD><<<grid_for(B * N, kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                          A, b, 1.0, host_mat<kD>(q_host), time_grid, 0, ssm_A, ssm_b, ssm_cholQ, B, N, status)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_sq_drift_difference(int drift_id, const double* params_host, const double* q_host, double dt,
                            const double* A, const double* b, const double* m, const double* S, double* value,
                            double* dEdm, double* dEdS, void* workspace, size_t workspace_bytes, int64_t B,
                            int64_t N, int D, int64_t* status, void* stream) {
    VDP_REQUIRE(q_host && A && b && m && S && value && B >= 0 && N >= 1, "pointers, B, N");
    VDP_REQUIRE((dEdm == nullptr) == (dEdS == nullptr), "dEdm and dEdS must both be given or both be NULL");
    if (B == 0) return VDP_OK;
    Arena ar(workspace, workspace_bytes);
    double* terms = workspace ? ar.take((size_t)B * N) : nullptr;
    if (!terms) { set_error("workspace too small"); return VDP_E_WORKSPACE; }
    const DriftParams P = make_params(params_host);
    cudaStream_t st = as_stream(stream);
    VDP_DISPATCH_D(D, VDP_DISPATCH_DRIFT(drift_id, {
        constexpr int G = QuadGroup<kD>::value;
        k_vgp_esde<kDrift, kD, 20, G><<<grid_for(B * N * G, kPtBlock), kPtBlock, 0, st>>>(
            P, host_mat<kD>(q_host), dt, pi_factor(kD), A, b, m, S, terms, dEdm, dEdS, B, N, status);
    }));
    VDP_CHECK_LAUNCH();
    k_reduce_terms<<<(unsigned)B, 256, 0, st>>>(terms, N, value);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

size_t vdp_sq_drift_difference_workspace(int64_t B, int64_t N, int D) { (void)D; return padded((size_t)B * N); }

size_t vdp_vgp_workspace(int64_t B, int64_t N, int D, int64_t n_obs) { (void)n_obs; return vgp_ws_bytes(B, N, D); }

VDP_VGP_ENTRY(vdp_vgp_forward_pass, { if ((rc = vgp_forward<kD>(*p, w, st))) return rc; })
VDP_VGP_ENTRY(vdp_vgp_update_lagrange, { if ((rc = vgp_lagrange<kDrift, kD>(*p, w, st))) return rc; })
VDP_VGP_ENTRY(vdp_vgp_update_param, { if ((rc = vgp_update_param<kDrift, kD>(*p, st))) return rc; })
VDP_VGP_ENTRY(vdp_vgp_update_initial_statistics, { if ((rc = vgp_update_initial<kD>(*p, st))) return rc; })
VDP_VGP_ENTRY(vdp_vgp_elbo, { if ((rc = vgp_elbo<kDrift, kD>(*p, w, st))) return rc; })
// one iteration of VIMarkovGPTrainer.perform_inference (vi_markov_gp_trainer.py:62-73)
// (the tail -- forward pass + E_sde quadrature with its gradients -- is what the next sweep starts from:
// VDP_VGP_STATE_CARRIED skips the head)
VDP_VGP_ENTRY(vdp_vgp_sweep, {
    const bool carried = (p->flags & VDP_VGP_STATE_CARRIED) != 0;
    int64_t parts = 0;
    if (!carried) { if ((rc = vgp_forward_esde<kDrift, kD>(*p, w, st, &parts))) return rc; }
    if ((rc = vgp_lagrange<kDrift, kD>(*p, w, st, true, carried))) return rc;
    if ((rc = vgp_update_param<kDrift, kD>(*p, st))) return rc;
    if (p->flags & VDP_VGP_UPDATE_X0) { if ((rc = vgp_update_initial<kD>(*p, st))) return rc; }
    if ((rc = vgp_forward_esde<kDrift, kD>(*p, w, st, &parts))) return rc;
    if ((rc = vgp_elbo<kDrift, kD>(*p, w, st, parts))) return rc;
})

}  // extern "C"
