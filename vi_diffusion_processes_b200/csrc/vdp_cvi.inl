// CVI-DP entry points: KL + gradient, the fused site update and classic_elbo
// (included at the end of vdp_all.inl after vdp_fused.inl).

namespace vdp {

template <int D> struct KlLaunch {
    static constexpr int G = QuadGroup<D>::value;
    static constexpr int groups = (G == 1) ? 128 : 8;   // first group of a block is the halo
    static constexpr int threads = groups * G;
};

template <int DRIFT, int D>
static int launch_kl_grad(const DriftParams& P, const double* q_host, double dt, const double* time_grid,
                          const double* x0_mu, const double* x0_cov, const double* Aq, const double* m,
                          const double* Sd, double* terms, double* g1, double* gd, double* gs, int upd, double lr,
                          const double* sc1, const double* scd, int64_t B, int64_t N, int64_t* status,
                          cudaStream_t st, int64_t node0 = 0, int64_t node1 = -1,
                          const KlNaturalsOut& nat = KlNaturalsOut(), bool gh_loops = false) {
    using KL = KlLaunch<D>;
    const int64_t total = ((node1 < 0) ? B * (N + 1) : node1) - node0;
    if (total <= 0) return VDP_OK;
    const unsigned grid = (unsigned)((total + KL::groups - 2) / (KL::groups - 1));
    const size_t smem = KL::groups * sizeof(KlHandoff<D>);
    // polynomial scalar drifts: the 20-point Gauss-Hermite sums are taken in closed form (CubicMoments, the same
    // values up to rounding); VDP_FLAG_GH_QUADRATURE / the GH_QUADRATURE switch run the quadrature loops instead
    if constexpr (D == 1 && KL::G == 1 && DriftStep1<DRIFT>::kCubic) {
        if (!gh_loops && !g_cfg.gh_quadrature) {
#define VDP_KL_CLOSED(U, NV)                                                                                          \
    k_kl_grad<DRIFT, D, 20, KL::G, true, U, NV><<<grid, KL::threads, smem, st>>>(                                         \
        P, host_mat<D>(q_host), dt, pi_factor(D), time_grid, host_vec<D>(x0_mu), host_mat<D>(x0_cov), Aq, m, Sd, terms, \
        g1, gd, gs, upd, lr, sc1, scd, B, N, status, node0, node1, nat)
            if (upd == 2 && !terms) VDP_KL_CLOSED(2, true);
            else if (upd == 2) VDP_KL_CLOSED(2, false);
            else if (upd == 1) VDP_KL_CLOSED(1, false);
            else VDP_KL_CLOSED(0, false);
#undef VDP_KL_CLOSED
            VDP_CHECK_LAUNCH();
            return VDP_OK;
        }
    }
    // D >= 2, polynomial drift: the exact low-order rule (DriftPoly / ExactOrder in drifts.cuh), a few lanes per grid
    // point; VDP_FLAG_GH_QUADRATURE / the GH_QUADRATURE switch keep the reference's 20 points per dimension
    if constexpr (D >= 2 && DriftPoly<DRIFT>::degree > 0) {
        if (!gh_loops && !g_cfg.gh_quadrature) {
            constexpr int HX = ExactOrder<DRIFT>::H, GX = KlExactGroup<D>::value;
            constexpr int threads = KlShape<D, HX, GX>::threads, groups = threads / GX;
            const unsigned gridx = (unsigned)((total + groups - 2) / (groups - 1));
            k_kl_grad<DRIFT, D, HX, GX><<<gridx, threads, groups * sizeof(KlHandoff<D>), st>>>(
                P, host_mat<D>(q_host), dt, pi_factor(D), time_grid, host_vec<D>(x0_mu), host_mat<D>(x0_cov), Aq, m, Sd,
                terms, g1, gd, gs, upd, lr, sc1, scd, B, N, status, node0, node1, nat);
            VDP_CHECK_LAUNCH();
            return VDP_OK;
        }
    }
    k_kl_grad<DRIFT, D, 20, KL::G><<<grid, KL::threads, smem, st>>>(
        P, host_mat<D>(q_host), dt, pi_factor(D), time_grid, host_vec<D>(x0_mu), host_mat<D>(x0_cov), Aq, m, Sd,
        terms, g1, gd, gs, upd, lr, sc1, scd, B, N, status, node0, node1, nat);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

static __global__ void k_elbo_combine(const double* __restrict__ ve, const double* __restrict__ kl,
                               double* __restrict__ elbo, int64_t B) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b < B) elbo[b] = ve[b] - kl[b];
}

// workspace carve-up shared by site_update and elbo
struct CviWs {
    double *th1, *thd, *ths, *Ld, *Ls, *terms, *sc1, *scd;
    void* seg;   // kSegScratchBytes of segment scratch for the tiled sweeps
};
static inline size_t cvi_ws_bytes(int64_t B, int64_t N, int D) {
    const size_t n = (size_t)B * (N + 1), t = (size_t)B * N, DD = (size_t)D * D;
    return padded(n * D) + padded(n * DD) + padded(t * DD) + padded(n * DD) + padded(t * DD) + padded(n) +
           padded(n * D) + padded(n * DD) + kSegScratchBytes;
}
static inline bool cvi_ws_take(void* ws, size_t bytes, int64_t B, int64_t N, int D, CviWs& w) {
    if (!ws) return false;
    Arena ar(ws, bytes);
    const size_t n = (size_t)B * (N + 1), t = (size_t)B * N, DD = (size_t)D * D;
    w.th1 = ar.take(n * D);
    w.thd = ar.take(n * DD);
    w.ths = ar.take(t * DD);
    w.Ld = ar.take(n * DD);
    w.Ls = ar.take(t * DD);
    w.terms = ar.take(n);
    w.sc1 = ar.take(n * D);
    w.scd = ar.take(n * DD);
    w.seg = ar.take_bytes(kSegScratchBytes);
    return w.seg != nullptr;
}

// (F): sites -> q as an SSM -> marginals.   variational_cvi_sde.py:161-192 + gauss_markov.py:107-117
// naturals_ready: the site-update kernel has already written the natural parameters into the workspace
// DRIFT >= 0 with kl_partials != NULL: classic_elbo follows on the whole chain; where the mean sweep can add up the KL
// terms on its way (tiled route, D = 1, polynomial drift in closed form) it does, and *kl_partials is the number of
// partial sums per sequence it left in w.terms (0: run_elbo makes its own pass).
// skip_cholQ (tiled route only): the caller does not read the process-noise factors of this rebuild
// stop_before_means (tiled route): q as an SSM and the marginal covariances only (the fused forward pass computes the
// means itself);  factor_ready (tiled route): w.Ld / w.Ls already hold the factor of the naturals' precision
template <int D, int DRIFT = -1>
static int run_posterior(const vdp_cvi_problem& p, const CviWs& w, cudaStream_t st, bool naturals_ready = false,
                         int64_t* kl_partials = nullptr, bool skip_cholQ = false, bool stop_before_means = false,
                         bool factor_ready = false) {
    const int64_t B = p.B, N = p.N;
    if (kl_partials) *kl_partials = 0;
    if (!naturals_ready) {
        prof_mark(VDP_PHASE_SITES_TO_NAT, st);
        k_ssm_to_naturals<D><<<grid_for(B * (N + 1), kPtBlock), kPtBlock, 0, st>>>(
            p.pA, p.pb, p.pcholQ, p.pmu0, p.pcholP0, p.g1, p.gd, p.gs, w.th1, w.thd, w.ths, B, N);
        VDP_CHECK_LAUNCH();
        if (p.n_obs > 0) {
            k_scatter_data<D><<<grid_for(B * p.n_obs, kPtBlock), kPtBlock, 0, st>>>(p.d1, p.d2, p.obs_idx, w.th1,
                                                                                w.thd, B, N, p.n_obs);
            VDP_CHECK_LAUNCH();
        }
    }
    const SegScratch sc(w.seg, kSegScratchBytes);
    // marginals (gauss_markov.py:107-117): the reference rebuilds the precision from the state-space model,
    // factorises it again and takes the diagonal of the inverse band.  That matrix is the one naturals_to_ssm started
    // from, and naturals_to_ssm itself walks through the inverse band of its factor (ssm_gaussian_transformations.py:
    // 443-504): on the tiled route the backward sweep writes those diagonal blocks out as the marginal covariances
    // and the forward sweep that follows computes the means alone.  VDP_REFACTOR_PRECISION=1 keeps the second
    // factorisation + inverse band (equal to rounding, tests/test_gpu_parity.py::test_marginals_from_first_factor).
    const bool ssm_only = (p.flags & VDP_FLAG_POSTERIOR_SSM_ONLY) != 0;
    const bool gh_loops = (p.flags & VDP_FLAG_GH_QUADRATURE) || g_cfg.gh_quadrature;
    const bool lite = use_tiled_sweeps(B, N) && !(p.flags & VDP_FLAG_REFACTOR_PRECISION) && !g_cfg.refactor_precision;
    int rc = run_nat2ssm<D>(w.th1, w.thd, w.ths, p.qA, p.qb, (skip_cholQ && lite && !ssm_only) ? nullptr : p.qcholQ,
                            p.qmu0, p.qcholP0, w.Ld, w.Ls, B, N, p.status, st, sc,
                            (lite && !ssm_only) ? p.fx_covs : nullptr, factor_ready);
    if (rc || ssm_only) return rc;
    if (stop_before_means && lite) return VDP_OK;
    if (lite) {
        prof_mark(VDP_PHASE_MARG_FWD, st);
        if constexpr (DRIFT >= 0 && D == 1) {
            if constexpr (DriftStep1<(DRIFT < 0 ? 0 : DRIFT)>::kCubic) {
                if (kl_partials && !gh_loops && !(p.flags & VDP_FLAG_SEPARATE_KL_VALUE) && !g_cfg.separate_kl_value) {
                    MeanKlArgs ka;
                    ka.P = make_params(p.drift_params);
                    ka.Qp = p.dt * p.q[0];
                    ka.dt = p.dt;
                    ka.Sd = p.fx_covs; ka.cholQ = p.qcholQ; ka.qcholP0 = p.qcholP0; ka.pmu0 = p.pmu0;
                    ka.pcholP0 = p.pcholP0; ka.partials = w.terms; ka.status = p.status;
                    if ((rc = run_mean_fwd_tiled<D, DRIFT>(p.qA, p.qb, p.qmu0, p.fx_mus, B, N, sc, st, &ka, kl_partials)))
                        return rc;
                    prof_mark(VDP_PHASE_END, st);
                    return VDP_OK;
                }
            }
        }
        if ((rc = run_mean_fwd_tiled<D>(p.qA, p.qb, p.qmu0, p.fx_mus, B, N, sc, st))) return rc;
        prof_mark(VDP_PHASE_END, st);
        return VDP_OK;
    }
    return run_marginals<D>(p.qA, p.qb, p.qcholQ, p.qmu0, p.qcholP0, p.fx_mus, p.fx_covs, w.Ld, w.Ls, B, N, p.status,
                            st, sc);
}

template <int DRIFT, int D>
static int run_elbo(const vdp_cvi_problem& p, const CviWs& w, const DriftParams& P, cudaStream_t st,
                    int64_t node0 = 0, int64_t node1 = -1, int64_t kl_partials = 0) {
    const int64_t B = p.B, N = p.N;
    constexpr int G = QuadGroup<D>::value;
    prof_mark(VDP_PHASE_ELBO, st);
    const bool ranged = node1 >= 0;
    const int64_t cnt = (ranged ? node1 : B * (N + 1)) - node0;
    if (ranged) VDP_CHECK_CUDA(cudaMemsetAsync(w.terms, 0, (size_t)B * (N + 1) * sizeof(double), st));
    if (B <= 64 && p.n_obs >= 1024)
        k_variational_expectation_block<D><<<(unsigned)B, 256, 0, st>>>(
            p.fx_mus, p.fx_covs, p.obs_idx, p.obs_y, host_mat<D>(p.lik_chol), p.ve, B, N, p.n_obs);
    else
        k_variational_expectation<D><<<grid_for(B * 32, kPtBlock), kPtBlock, 0, st>>>(
            p.fx_mus, p.fx_covs, p.obs_idx, p.obs_y, host_mat<D>(p.lik_chol), p.ve, B, N, p.n_obs);
    VDP_CHECK_LAUNCH();
    bool closed = false;
    if constexpr (D == 1 && G == 1 && DriftStep1<DRIFT>::kCubic)
        closed = !(p.flags & VDP_FLAG_GH_QUADRATURE) && !g_cfg.gh_quadrature;
    if (kl_partials > 0) {
        // the mean sweep of the rebuild has left kl_partials partial sums per sequence in w.terms
        k_reduce_terms<<<(unsigned)B, 256, 0, st>>>(w.terms, kl_partials, p.kl);
        VDP_CHECK_LAUNCH();
        k_elbo_combine<<<grid_for(B, kPtBlock), kPtBlock, 0, st>>>(p.ve, p.kl, p.elbo, B);
        VDP_CHECK_LAUNCH();
        prof_mark(VDP_PHASE_END, st);
        return VDP_OK;
    }
    if (cnt > 0 && closed) {
        if constexpr (D == 1 && G == 1 && DriftStep1<DRIFT>::kCubic) {
            k_kl_value<DRIFT, D, 20, G, true><<<grid_for(cnt * G, kPtBlock), kPtBlock, 0, st>>>(
                P, host_mat<D>(p.q), p.dt, pi_factor(D), p.qA, p.qb, p.qcholQ, p.fx_mus, p.fx_covs, p.qmu0,
                p.qcholP0, p.pmu0, p.pcholP0, w.terms, B, N, p.status, node0, node1);
            VDP_CHECK_LAUNCH();
        }
    } else if (cnt > 0 && D >= 2 && DriftPoly<DRIFT>::degree > 0 && !(p.flags & VDP_FLAG_GH_QUADRATURE) &&
               !g_cfg.gh_quadrature) {
        // exact low-order rule of a polynomial drift, one thread per grid point (value only: little state per thread)
        constexpr int HX = (D >= 2) ? ExactOrder<DRIFT>::H : 20;
        k_kl_value<DRIFT, D, HX, 1><<<grid_for(cnt, kPtBlock), kPtBlock, 0, st>>>(
            P, host_mat<D>(p.q), p.dt, pi_factor(D), p.qA, p.qb, p.qcholQ, p.fx_mus, p.fx_covs, p.qmu0, p.qcholP0,
            p.pmu0, p.pcholP0, w.terms, B, N, p.status, node0, node1);
        VDP_CHECK_LAUNCH();
    } else if (cnt > 0) {
        k_kl_value<DRIFT, D, 20, G><<<grid_for(cnt * G, kPtBlock), kPtBlock, 0, st>>>(
            P, host_mat<D>(p.q), p.dt, pi_factor(D), p.qA, p.qb, p.qcholQ, p.fx_mus, p.fx_covs, p.qmu0, p.qcholP0,
            p.pmu0, p.pcholP0, w.terms, B, N, p.status, node0, node1);
        VDP_CHECK_LAUNCH();
    }
    launch_reduce_terms(w.terms, N + 1, p.kl, B, st);
    VDP_CHECK_LAUNCH();
    k_elbo_combine<<<grid_for(B, kPtBlock), kPtBlock, 0, st>>>(p.ve, p.kl, p.elbo, B);
    VDP_CHECK_LAUNCH();
    prof_mark(VDP_PHASE_END, st);
    return VDP_OK;
}

// node0 / node1 >= 0: sharded quadrature (vdp_cvi_site_update_sharded): the rebuilds run on the whole chain, the
// quadrature and the site update on the flat node range [node0, node1) only; stop_after_sites ends phase 0.
template <int DRIFT, int D>
static int run_site_update(const vdp_cvi_problem& p, const CviWs& w, cudaStream_t st, int64_t node0 = 0,
                           int64_t node1 = -1, bool stop_after_sites = false) {
    const int64_t B = p.B, N = p.N;
    const DriftParams P = make_params(p.drift_params);
    int rc;
    // the naturals written by the previous update's girsanov kernel are still valid away from the observations
    const bool carried = (p.flags & VDP_FLAG_NATURALS_CARRIED) && p.n_obs <= 128 && node1 < 0 &&
                         !g_cfg.no_carried_naturals;
    // sites held in theta (VDP_FLAG_SITES_IN_NATURALS, only together with carried naturals): the girsanov update moves
    // theta itself, g1 / gd / gs are neither read nor written (vdp_cvi_materialize_sites forms them on request)
    const bool lazy = carried && (p.flags & VDP_FLAG_SITES_IN_NATURALS) && !stop_after_sites &&
                      !(p.flags & VDP_FLAG_SKIP_GIRSANOV);
    // (A) update_data_sites  variational_cvi_sde.py:301-317
    if (lazy && p.n_obs > 0) {
        // ... and the patch of theta at the observed grid points in the same launch (one block per sequence)
        prof_mark(VDP_PHASE_SITES_TO_NAT, st);
        k_obs_update_patch<D><<<(unsigned)B, kObsBlock, 0, st>>>(
            p.fx_mus, p.obs_idx, p.obs_y, host_mat<D>(p.lik_chol), p.lr_data, p.lr_data_dev, p.active_dev,
            (p.flags & VDP_FLAG_SKIP_DATA_UPDATE) ? 1 : 0, p.d1, p.d2, w.th1, w.thd, w.sc1, w.scd, N, p.n_obs);
        VDP_CHECK_LAUNCH();
    } else {
        if (!(p.flags & VDP_FLAG_SKIP_DATA_UPDATE) && p.n_obs > 0) {
            prof_mark(VDP_PHASE_DATA_SITES, st);
            k_data_site_update<D><<<grid_for(B * p.n_obs, kPtBlock), kPtBlock, 0, st>>>(
                p.fx_mus, p.obs_idx, p.obs_y, host_mat<D>(p.lik_chol), p.lr_data, p.d1, p.d2, B, N, p.n_obs,
                p.lr_data_dev, p.active_dev);
            VDP_CHECK_LAUNCH();
        }
        if (carried && p.n_obs > 0) {
            prof_mark(VDP_PHASE_SITES_TO_NAT, st);
            k_patch_naturals<D><<<grid_for(B * p.n_obs, kPtBlock), kPtBlock, 0, st>>>(
                p.pA, p.pb, p.pcholQ, p.pmu0, p.pcholP0, p.g1, p.gd, p.d1, p.d2, p.obs_idx, w.th1, w.thd, w.sc1, w.scd,
                B, N, p.n_obs);
            VDP_CHECK_LAUNCH();
        }
    }
    // the rebuild between the two site updates feeds k_kl_grad (A, means, marginal covariances): no chol Q needed
    const bool girsanov_follows = !(p.flags & VDP_FLAG_SKIP_GIRSANOV);
    // One forward pass for mean sweep -> KL gradient / site update -> factorisation of the next rebuild
    // (fwd_fused.cuh): scalar chains, polynomial drift in closed form, sites held in theta, one segment per sequence
    bool fused_fwd = false;
    if constexpr (D == 1 && DriftStep1<DRIFT>::kCubic) {
        fused_fwd = lazy && !(p.flags & (VDP_FLAG_GH_QUADRATURE | VDP_FLAG_REFACTOR_PRECISION | VDP_FLAG_NO_FUSED_FORWARD)) &&
                    !g_cfg.gh_quadrature && !g_cfg.refactor_precision && !g_cfg.no_fused_forward &&
                    use_tiled_sweeps(B, N) && plan_segments(B, N, kFusedT * kFusedL, true).nseg == 1;
    }
    if ((rc = run_posterior<D>(p, w, st, carried, nullptr, girsanov_follows, fused_fwd))) return rc;
    if constexpr (D == 1 && DriftStep1<DRIFT>::kCubic) {
        if (fused_fwd) {
            prof_mark(VDP_PHASE_KL_GRAD, st);
            if (p.n_obs > 0) {
                k_add_data_step_seq<1><<<(unsigned)B, kObsBlock, 0, st>>>(
                    p.obs_idx, w.sc1, w.scd, w.th1, w.thd, p.lr_girsanov, p.lr_girsanov_dev, p.active_dev, N, p.n_obs);
                VDP_CHECK_LAUNCH();
            }
            FwdFusedArgs fa;
            fa.P = P;
            fa.q = p.q[0]; fa.dt = p.dt; fa.lr = p.lr_girsanov;
            fa.x0_mu = p.x0_mu[0]; fa.x0_cov = p.x0_cov[0];
            fa.time_grid = p.time_grid;
            fa.A = p.qA; fa.bo = p.qb; fa.mu0 = p.qmu0; fa.Sd = p.fx_covs;
            fa.th1 = w.th1; fa.thd = w.thd; fa.ths = w.ths;
            fa.Ld = w.Ld; fa.Ls = w.Ls;
            fa.lr_dev = p.lr_girsanov_dev; fa.active = p.active_dev;
            fa.status = p.status;
            constexpr size_t smem = fwd_fused_smem<kFusedT, kFusedL>();
            if ((rc = allow_smem(k_fwd_fused<DRIFT, kFusedT, kFusedL>, smem))) return rc;
            k_fwd_fused<DRIFT, kFusedT, kFusedL><<<(unsigned)B, kFusedT, smem, st>>>(fa, B, N);
            VDP_CHECK_LAUNCH();
            const bool want = !(p.flags & VDP_FLAG_SKIP_ELBO);
            int64_t parts = 0;
            if ((rc = run_posterior<D, DRIFT>(p, w, st, true, want ? &parts : nullptr, false, false, true))) return rc;
            if (want) return run_elbo<DRIFT, D>(p, w, P, st, 0, -1, parts);
            return VDP_OK;
        }
    }
    if (p.flags & VDP_FLAG_SKIP_GIRSANOV) {
        if (!(p.flags & VDP_FLAG_SKIP_ELBO)) return run_elbo<DRIFT, D>(p, w, P, st);
        return VDP_OK;
    }
    // (B) update_girsanov_sites  variational_cvi_sde.py:279-299
    const size_t n = (size_t)B * (N + 1);
    prof_mark(VDP_PHASE_KL_GRAD, st);
    // data sites on the dense grid (variational_cvi_sde.py:286-288).  With carried naturals the dense copy is carried
    // too: k_patch_naturals has rewritten its observed entries.  (A per-block gather of the observations into shared
    // memory inside the kernel, which would save the dense reads, measured slower: the kernel is bound by the FP64
    // pipe and instruction issue, and the gather's extra instructions cost more than the reads.)
    if (!carried) {
        VDP_CHECK_CUDA(cudaMemsetAsync(w.sc1, 0, n * D * sizeof(double), st));
        VDP_CHECK_CUDA(cudaMemsetAsync(w.scd, 0, n * D * D * sizeof(double), st));
        if (p.n_obs > 0) {
            k_scatter_data<D><<<grid_for(B * p.n_obs, kPtBlock), kPtBlock, 0, st>>>(p.d1, p.d2, p.obs_idx, w.sc1, w.scd,
                                                                                B, N, p.n_obs);
            VDP_CHECK_LAUNCH();
        }
    }
    // whole-chain update: the kernel also writes the naturals of the posterior that follows (no second pass over
    // the sites); the sharded quadrature exchanges site ranges first and rebuilds from the sites
    const bool fuse_nat = lazy || ((node1 < 0) && !stop_after_sites && !(p.flags & VDP_FLAG_NO_FUSED_NATURALS) &&
                                   !g_cfg.no_fused_naturals);
    KlNaturalsOut nat;
    if (fuse_nat) {
        if (!lazy) { nat.pA = p.pA; nat.pb = p.pb; nat.pcholQ = p.pcholQ; nat.pmu0 = p.pmu0; nat.pcholP0 = p.pcholP0; }
        nat.th1 = w.th1; nat.thd = w.thd; nat.ths = w.ths;
    }
    nat.lr_dev = p.lr_girsanov_dev;
    nat.active = p.active_dev;
    if ((rc = launch_kl_grad<DRIFT, D>(P, p.q, p.dt, p.time_grid, p.x0_mu, p.x0_cov, p.qA, p.fx_mus, p.fx_covs,
                                       nullptr, p.g1, p.gd, p.gs, lazy ? 2 : 1, p.lr_girsanov, w.sc1, w.scd, B, N,
                                       p.status, st, node0, node1, nat, (p.flags & VDP_FLAG_GH_QUADRATURE) != 0)))
        return rc;
    if (stop_after_sites) { prof_mark(VDP_PHASE_END, st); return VDP_OK; }
    const bool want_elbo = !(p.flags & VDP_FLAG_SKIP_ELBO);
    int64_t kl_partials = 0;
    if ((rc = run_posterior<D, DRIFT>(p, w, st, fuse_nat, (want_elbo && node1 < 0) ? &kl_partials : nullptr))) return rc;
    // (C) classic_elbo  variational_cvi_sde.py:339-352
    if (want_elbo) return run_elbo<DRIFT, D>(p, w, P, st, 0, -1, kl_partials);
    return VDP_OK;
}

static int check_problem(const vdp_cvi_problem* p) {
    VDP_REQUIRE(p != nullptr, "problem");
    VDP_REQUIRE(p->B >= 0 && p->N >= 1 && p->n_obs >= 0, "B, N, n_obs");
    VDP_REQUIRE(p->time_grid && p->pA && p->pb && p->pcholQ && p->pmu0 && p->pcholP0, "prior SSM pointers");
    VDP_REQUIRE(p->g1 && p->gd && p->gs, "girsanov site pointers");
    VDP_REQUIRE(p->n_obs == 0 || (p->d1 && p->d2 && p->obs_idx && p->obs_y), "data site pointers");
    VDP_REQUIRE(p->fx_mus && p->fx_covs && p->qA && p->qb && p->qcholQ && p->qmu0 && p->qcholP0, "output pointers");
    VDP_REQUIRE(p->elbo && p->ve && p->kl, "elbo pointers");
    return VDP_OK;
}

}  // namespace vdp

#if VDP_EMIT_CVI
extern "C" {

size_t vdp_cvi_kl_workspace(int64_t B, int64_t N, int D) { (void)D; return padded((size_t)B * (N + 1)); }

int vdp_cvi_kl_and_grad(int drift_id, const double* params_host, const double* q_host, double dt,
                        const double* time_grid, const double* x0_mu_host, const double* x0_cov_host,
                        const double* A, const double* m, const double* Sd, double* kl, double* g1, double* gd,
                        double* gs, void* workspace, size_t workspace_bytes, int64_t B, int64_t N, int D,
                        int64_t* status, void* stream) {
    VDP_REQUIRE(q_host && time_grid && x0_mu_host && x0_cov_host && A && m && Sd && kl && g1 && gd && gs,
                "pointers");
    VDP_REQUIRE(B >= 0 && N >= 1, "B, N");
    if (B == 0) return VDP_OK;
    Arena ar(workspace, workspace_bytes);
    double* terms = workspace ? ar.take((size_t)B * (N + 1)) : nullptr;
    if (!terms) { set_error("workspace too small"); return VDP_E_WORKSPACE; }
    const DriftParams P = make_params(params_host);
    cudaStream_t st = as_stream(stream);
    VDP_DISPATCH_D(D, VDP_DISPATCH_DRIFT(drift_id, {
        int rc = launch_kl_grad<kDrift, kD>(P, q_host, dt, time_grid, x0_mu_host, x0_cov_host, A, m, Sd, terms, g1,
                                            gd, gs, 0, 0.0, nullptr, nullptr, B, N, status, st);
        if (rc) return rc;
    }));
    launch_reduce_terms(terms, N + 1, kl, B, st);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

size_t vdp_cvi_workspace(int64_t B, int64_t N, int D, int64_t n_obs) { (void)n_obs; return cvi_ws_bytes(B, N, D); }

int vdp_cvi_site_update(const vdp_cvi_problem* p, void* stream) {
    int rc = check_problem(p);
    if (rc) return rc;
    if (p->B == 0) return VDP_OK;
    CviWs w;
    if (!cvi_ws_take(p->workspace, p->workspace_bytes, p->B, p->N, p->D, w)) {
        set_error("workspace too small: need %zu bytes", cvi_ws_bytes(p->B, p->N, p->D));
        return VDP_E_WORKSPACE;
    }
    VDP_DISPATCH_D(p->D, VDP_DISPATCH_DRIFT(p->drift_id, {
        rc = run_site_update<kDrift, kD>(*p, w, as_stream(stream));
        if (rc) return rc;
    }));
    return VDP_OK;
}

int vdp_cvi_materialize_sites(const vdp_cvi_problem* p, void* stream) {
    int rc = check_problem(p);
    if (rc) return rc;
    if (p->B == 0) return VDP_OK;
    CviWs w;
    if (!cvi_ws_take(p->workspace, p->workspace_bytes, p->B, p->N, p->D, w)) {
        set_error("workspace too small: need %zu bytes", cvi_ws_bytes(p->B, p->N, p->D));
        return VDP_E_WORKSPACE;
    }
    VDP_DISPATCH_D(p->D, (k_sites_from_naturals<kD><<<grid_for(p->B * (p->N + 1), kPtBlock), kPtBlock, 0,
                                                     as_stream(stream)>>>(
                              p->pA, p->pb, p->pcholQ, p->pmu0, p->pcholP0, w.th1, w.thd, w.ths, w.sc1, w.scd, p->g1,
                              p->gd, p->gs, p->B, p->N)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// ---- the steps of a site update one at a time, without the posterior rebuild between them: what a caller that
// rebuilds the posterior itself (time-sharded chains, vdp_dist_sweep_*) strings together
int vdp_cvi_update_data_sites(const vdp_cvi_problem* p, void* stream) {
    int rc = check_problem(p);
    if (rc) return rc;
    if (p->B == 0 || p->n_obs == 0) return VDP_OK;
    VDP_DISPATCH_D(p->D, (k_data_site_update<kD><<<grid_for(p->B * p->n_obs, kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                             p->fx_mus, p->obs_idx, p->obs_y, host_mat<kD>(p->lik_chol), p->lr_data, p->d1, p->d2, p->B,
                             p->N, p->n_obs, p->lr_data_dev, p->active_dev)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_cvi_full_sites(const vdp_cvi_problem* p, double* th1, double* thd, double* ths, void* stream) {
    int rc = check_problem(p);
    if (rc) return rc;
    VDP_REQUIRE(th1 && thd && ths, "th1, thd, ths");
    if (p->B == 0) return VDP_OK;
    cudaStream_t st = as_stream(stream);
    VDP_DISPATCH_D(p->D, {
        k_ssm_to_naturals<kD><<<grid_for(p->B * (p->N + 1), kPtBlock), kPtBlock, 0, st>>>(
            p->pA, p->pb, p->pcholQ, p->pmu0, p->pcholP0, p->g1, p->gd, p->gs, th1, thd, ths, p->B, p->N);
        if (p->n_obs > 0)
            k_scatter_data<kD><<<grid_for(p->B * p->n_obs, kPtBlock), kPtBlock, 0, st>>>(p->d1, p->d2, p->obs_idx, th1,
                                                                                     thd, p->B, p->N, p->n_obs);
    });
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_cvi_update_girsanov_sites_range(const vdp_cvi_problem* p, int64_t node_begin, int64_t node_end, void* stream) {
    int rc = check_problem(p);
    if (rc) return rc;
    VDP_REQUIRE(node_begin >= 0 && node_end >= node_begin && node_end <= p->B * (p->N + 1), "node range");
    if (p->B == 0 || node_end == node_begin) return VDP_OK;
    CviWs w;
    if (!cvi_ws_take(p->workspace, p->workspace_bytes, p->B, p->N, p->D, w)) {
        set_error("workspace too small: need %zu bytes", cvi_ws_bytes(p->B, p->N, p->D));
        return VDP_E_WORKSPACE;
    }
    const DriftParams P = make_params(p->drift_params);
    cudaStream_t st = as_stream(stream);
    const size_t n = (size_t)p->B * (p->N + 1);
    VDP_DISPATCH_D(p->D, VDP_DISPATCH_DRIFT(p->drift_id, {
        VDP_CHECK_CUDA(cudaMemsetAsync(w.sc1, 0, n * kD * sizeof(double), st));
        VDP_CHECK_CUDA(cudaMemsetAsync(w.scd, 0, n * kD * kD * sizeof(double), st));
        if (p->n_obs > 0) {
            k_scatter_data<kD><<<grid_for(p->B * p->n_obs, kPtBlock), kPtBlock, 0, st>>>(p->d1, p->d2, p->obs_idx, w.sc1,
                                                                                     w.scd, p->B, p->N, p->n_obs);
            VDP_CHECK_LAUNCH();
        }
        rc = launch_kl_grad<kDrift, kD>(P, p->q, p->dt, p->time_grid, p->x0_mu, p->x0_cov, p->qA, p->fx_mus, p->fx_covs,
                                        nullptr, p->g1, p->gd, p->gs, 1, p->lr_girsanov, w.sc1, w.scd, p->B, p->N,
                                        p->status, st, node_begin, node_end, [&] {
                                            KlNaturalsOut nat;
                                            nat.lr_dev = p->lr_girsanov_dev;
                                            nat.active = p->active_dev;
                                            return nat;
                                        }(),
                                        (p->flags & VDP_FLAG_GH_QUADRATURE) != 0);
        if (rc) return rc;
    }));
    return VDP_OK;
}

int vdp_cvi_elbo_terms_range(const vdp_cvi_problem* p, int64_t node_begin, int64_t node_end, void* stream) {
    int rc = check_problem(p);
    if (rc) return rc;
    VDP_REQUIRE(node_begin >= 0 && node_end >= node_begin && node_end <= p->B * (p->N + 1), "node range");
    if (p->B == 0) return VDP_OK;
    CviWs w;
    if (!cvi_ws_take(p->workspace, p->workspace_bytes, p->B, p->N, p->D, w)) {
        set_error("workspace too small: need %zu bytes", cvi_ws_bytes(p->B, p->N, p->D));
        return VDP_E_WORKSPACE;
    }
    const DriftParams P = make_params(p->drift_params);
    VDP_DISPATCH_D(p->D, VDP_DISPATCH_DRIFT(p->drift_id, {
        if ((rc = run_elbo<kDrift, kD>(*p, w, P, as_stream(stream), node_begin, node_end))) return rc;
    }));
    return VDP_OK;
}

int vdp_cvi_site_update_sharded(const vdp_cvi_problem* p, int phase, int64_t node_begin, int64_t node_end,
                                void* stream) {
    int rc = check_problem(p);
    if (rc) return rc;
    VDP_REQUIRE(phase == 0 || phase == 1, "phase must be 0 or 1");
    VDP_REQUIRE(node_begin >= 0 && node_end >= node_begin && node_end <= p->B * (p->N + 1), "node range");
    if (p->B == 0) return VDP_OK;
    CviWs w;
    if (!cvi_ws_take(p->workspace, p->workspace_bytes, p->B, p->N, p->D, w)) {
        set_error("workspace too small: need %zu bytes", cvi_ws_bytes(p->B, p->N, p->D));
        return VDP_E_WORKSPACE;
    }
    const DriftParams P = make_params(p->drift_params);
    VDP_DISPATCH_D(p->D, VDP_DISPATCH_DRIFT(p->drift_id, {
        if (phase == 0) {
            rc = run_site_update<kDrift, kD>(*p, w, as_stream(stream), node_begin, node_end, true);
        } else {
            rc = run_posterior<kD>(*p, w, as_stream(stream));
            if (!rc) rc = run_elbo<kDrift, kD>(*p, w, P, as_stream(stream), node_begin, node_end);
        }
        if (rc) return rc;
    }));
    return VDP_OK;
}

int vdp_cvi_elbo(const vdp_cvi_problem* p, void* stream) {
    int rc = check_problem(p);
    if (rc) return rc;
    if (p->B == 0) return VDP_OK;
    CviWs w;
    if (!cvi_ws_take(p->workspace, p->workspace_bytes, p->B, p->N, p->D, w)) {
        set_error("workspace too small: need %zu bytes", cvi_ws_bytes(p->B, p->N, p->D));
        return VDP_E_WORKSPACE;
    }
    const DriftParams P = make_params(p->drift_params);
    VDP_DISPATCH_D(p->D, VDP_DISPATCH_DRIFT(p->drift_id, {
        if ((rc = run_posterior<kD>(*p, w, as_stream(stream)))) return rc;
        if (p->flags & VDP_FLAG_POSTERIOR_SSM_ONLY) return VDP_OK;
        if ((rc = run_elbo<kDrift, kD>(*p, w, P, as_stream(stream)))) return rc;
    }));
    return VDP_OK;
}

}  // extern "C"
#endif  // VDP_EMIT_CVI
