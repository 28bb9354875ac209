// Prior-parameter learning entry points: gradients with respect to the drift parameters of
// KL_q_p, of -VE through the re-linearised prior, and of E_sde.  Included at the end of vdp_all.inl.

namespace vdp {

template <int D> struct ParamGradLaunch {
    static constexpr int G = QuadGroup<D>::value;
    static constexpr int groups = (G == 1) ? 128 : 8;
    static constexpr int threads = groups * G;
};
template <int D> static inline int64_t param_grad_blocks(int64_t N) {
    return (N + ParamGradLaunch<D>::groups - 1) / ParamGradLaunch<D>::groups;
}
static inline int64_t param_grad_blocks_any(int64_t N, int D) {
    return D == 1 ? param_grad_blocks<1>(N) : param_grad_blocks<2>(N);
}

template <int DRIFT, int D>
static int launch_drift_param_grad(const DriftParams& P, const Mat<D>& W, double c_x, double c_f, double a_sign,
                                   double out_scale, const double* A, const double* b, const double* m,
                                   const double* S, int64_t pts, double* partials, double* grads, int64_t B,
                                   int64_t N, int64_t* status, cudaStream_t st) {
    using L = ParamGradLaunch<D>;
    constexpr int NP = DriftParamGrad<DRIFT, D>::NP;
    const int64_t nbx = param_grad_blocks<D>(N);
    const dim3 grid((unsigned)nbx, (unsigned)B);
    k_drift_param_grad<DRIFT, D, 20, L::G><<<grid, L::threads, L::groups * NP * sizeof(double), st>>>(
        P, W, c_x, c_f, a_sign, pi_factor(D), A, b, m, S, pts, partials, N, status);
    VDP_CHECK_LAUNCH();
    k_reduce_partials<<<(unsigned)B, 256, 0, st>>>(partials, nbx, NP, out_scale, grads);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// partial sums [B, blocks per sequence, NP]: the segment scratch when they fit, else the Ls area (N D^2 doubles per
// sequence, enough once N >= 19)
static inline double* cvi_partials_area(const CviWs& w, int64_t B, int64_t N, int D) {
    const size_t need = (size_t)B * param_grad_blocks_any(N, D) * kMaxParams * sizeof(double);
    if (need <= kSegScratchBytes) return (double*)w.seg;
    if (need <= (size_t)B * N * D * D * sizeof(double)) return w.Ls;
    return nullptr;
}

static inline int64_t sens_chunk_len(int64_t N) {
    int64_t L = (int64_t)std::ceil(std::sqrt((double)(N + 1)));
    return std::min<int64_t>(std::max<int64_t>(L, 16), 2048);
}

template <int DRIFT, int D>
static int run_ve_prior_grad(const vdp_cvi_problem& p, const CviWs& w, const double* path_m, const double* path_S,
                             int clip, double lo, double hi, double* grads, cudaStream_t st) {
    using L = ParamGradLaunch<D>;
    constexpr int NP = DriftParamGrad<DRIFT, D>::NP;
    constexpr int DD = D * D;
    const int64_t B = p.B, N = p.N;
    const size_t n = (size_t)B * (N + 1);
    int rc;
    // dist_q.marginals under the current (re-linearised) prior: q as an SSM, posterior means and covariances
    if ((rc = run_posterior<D>(p, w, st))) return rc;
    // cotangents of -VE at the observed points, scattered onto the grid
    VDP_CHECK_CUDA(cudaMemsetAsync(w.sc1, 0, n * D * sizeof(double), st));
    VDP_CHECK_CUDA(cudaMemsetAsync(w.scd, 0, n * DD * sizeof(double), st));
    if (p.n_obs > 0) {
        k_ve_cotangents<D><<<grid_for(B * p.n_obs, kPtBlock), kPtBlock, 0, st>>>(
            p.fx_mus, p.obs_idx, p.obs_y, host_mat<D>(p.lik_chol), w.sc1, w.scd, B, N, p.n_obs);
        VDP_CHECK_LAUNCH();
    }
    // sensitivity sweeps: u -> th1, V_kk -> thd, V_{k+1,k} -> ths; chunk summaries and carries in the Ld area
    const int64_t Lc = sens_chunk_len(N);
    const int64_t nc = (N + 1 + Lc - 1) / Lc;
    const size_t elem_bytes = (size_t)B * nc * sizeof(SensElem<D>), carry_bytes = (size_t)B * nc * sizeof(SensState<D>);
    // Ld and Ls are adjacent in the workspace and both free here
    if (elem_bytes + carry_bytes > (size_t)(reinterpret_cast<char*>(w.terms) - reinterpret_cast<char*>(w.Ld))) {
        set_error("workspace too small for the sensitivity sweeps");
        return VDP_E_WORKSPACE;
    }
    SensElem<D>* elems = reinterpret_cast<SensElem<D>*>(w.Ld);
    SensState<D>* carry = reinterpret_cast<SensState<D>*>(reinterpret_cast<char*>(w.Ld) + elem_bytes);
    double *u = w.th1, *Vd = w.thd, *Vs = w.ths;
    const unsigned gchunks = grid_for(B * nc, 64);
    k_sens_summary<D, SENS_FWD><<<gchunks, 64, 0, st>>>(p.qA, p.fx_covs, w.sc1, w.scd, elems, B, N, Lc, nc);
    VDP_CHECK_LAUNCH();
    k_sens_carry<D><<<grid_for(B, kSeqBlock), kSeqBlock, 0, st>>>(elems, carry, B, nc);
    VDP_CHECK_LAUNCH();
    k_sens_apply<D, SENS_FWD><<<gchunks, 64, 0, st>>>(p.qA, p.fx_covs, w.sc1, w.scd, carry, u, Vd, Vs, B, N, Lc, nc);
    VDP_CHECK_LAUNCH();
    const int64_t ncb = (N + Lc - 1) / Lc;
    const unsigned gchunks_b = grid_for(B * ncb, 64);
    k_sens_summary<D, SENS_BWD><<<gchunks_b, 64, 0, st>>>(p.qA, p.fx_covs, w.sc1, w.scd, elems, B, N, Lc, ncb);
    VDP_CHECK_LAUNCH();
    k_sens_carry<D><<<grid_for(B, kSeqBlock), kSeqBlock, 0, st>>>(elems, carry, B, ncb);
    VDP_CHECK_LAUNCH();
    k_sens_apply<D, SENS_BWD><<<gchunks_b, 64, 0, st>>>(p.qA, p.fx_covs, w.sc1, w.scd, carry, u, Vd, Vs, B, N, Lc,
                                                        ncb);
    VDP_CHECK_LAUNCH();
    // H = 10 expectations of df/dp, dJ/dp on the linearisation path and the contraction with the sensitivities
    double* partials = reinterpret_cast<double*>(w.seg);
    const int64_t nbx = param_grad_blocks<D>(N);
    if ((size_t)B * nbx * NP * sizeof(double) > kSegScratchBytes) {
        partials = w.terms;   // B (N+1) doubles
        if ((size_t)B * nbx * NP > n) {
            set_error("workspace too small for the parameter-gradient partial sums");
            return VDP_E_WORKSPACE;
        }
    }
    const DriftParams P = make_params(p.drift_params);
    const dim3 grid((unsigned)nbx, (unsigned)B);
    k_ve_param_grad<DRIFT, D, 10, L::G><<<grid, L::threads, L::groups * NP * sizeof(double), st>>>(
        P, pi_factor(D), p.time_grid, path_m, path_S, clip, lo, hi, p.pA, p.pb, p.pcholQ, p.fx_mus, u, Vd, Vs, partials,
        N, p.status);
    VDP_CHECK_LAUNCH();
    k_reduce_partials<<<(unsigned)B, 256, 0, st>>>(partials, nbx, NP, 1.0, grads);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

}  // namespace vdp

extern "C" {

int vdp_drift_num_params(int drift_id) { return drift_num_params(drift_id); }

size_t vdp_drift_param_grad_workspace(int64_t B, int64_t N, int D) {
    return padded((size_t)B * param_grad_blocks_any(N, D) * kMaxParams);
}

int vdp_drift_param_grad(int drift_id, const double* params_host, const double* q_host, double q_scale, double c_x,
                         double c_f, double a_sign, double out_scale, const double* A, const double* b,
                         const double* m, const double* S, int64_t points_per_sequence, double* grads, void* workspace,
                         size_t workspace_bytes, int64_t B, int64_t N, int D, int64_t* status, void* stream) {
    VDP_REQUIRE(q_host && A && b && m && S && grads, "pointers");
    VDP_REQUIRE(B >= 0 && B <= 65535 && N >= 1 && points_per_sequence >= N, "B (<= 65535), N, points_per_sequence");
    if (B == 0) return VDP_OK;
    if (!workspace || workspace_bytes < vdp_drift_param_grad_workspace(B, N, D)) {
        set_error("workspace too small: need %zu bytes", vdp_drift_param_grad_workspace(B, N, D));
        return VDP_E_WORKSPACE;
    }
    const DriftParams P = make_params(params_host);
    VDP_DISPATCH_D(D, VDP_DISPATCH_DRIFT(drift_id, {
        bool ok = true;
        const Mat<kD> cq = chol(q_scale * host_mat<kD>(q_host), ok);
        if (!ok) { set_error("q_scale * q is not positive definite"); return VDP_E_BADARG; }
        const Mat<kD> W = sym(chol_inverse(cq));
        int rc = launch_drift_param_grad<kDrift, kD>(P, W, c_x, c_f, a_sign, out_scale, A, b, m, S,
                                                     points_per_sequence, (double*)workspace, grads, B, N, status,
                                                     as_stream(stream));
        if (rc) return rc;
    }));
    return VDP_OK;
}

int vdp_cvi_grad_kl_prior_params(const vdp_cvi_problem* p, double* grads, void* stream) {
    int rc = check_problem(p);
    if (rc) return rc;
    VDP_REQUIRE(grads != nullptr && p->B <= 65535, "grads, B <= 65535");
    if (p->B == 0) return VDP_OK;
    CviWs w;
    if (!cvi_ws_take(p->workspace, p->workspace_bytes, p->B, p->N, p->D, w)) {
        set_error("workspace too small: need %zu bytes", cvi_ws_bytes(p->B, p->N, p->D));
        return VDP_E_WORKSPACE;
    }
    double* partials = cvi_partials_area(w, p->B, p->N, p->D);
    if (!partials) { set_error("workspace too small for the partial sums"); return VDP_E_WORKSPACE; }
    const DriftParams P = make_params(p->drift_params);
    cudaStream_t st = as_stream(stream);
    VDP_DISPATCH_D(p->D, VDP_DISPATCH_DRIFT(p->drift_id, {
        if ((rc = run_posterior<kD>(*p, w, st))) return rc;   // KL_q_p rebuilds dist_q (variational_cvi_sde.py:450-462)
        bool ok = true;
        const Mat<kD> cq = chol(p->dt * host_mat<kD>(p->q), ok);
        if (!ok) { set_error("dt * q is not positive definite"); return VDP_E_BADARG; }
        const Mat<kD> W = sym(chol_inverse(cq));
        rc = launch_drift_param_grad<kDrift, kD>(P, W, 1.0, p->dt, 1.0, p->dt, p->qA, p->qb, p->fx_mus, p->fx_covs,
                                                 p->N + 1, partials, grads, p->B, p->N, p->status, st);
        if (rc) return rc;
    }));
    return VDP_OK;
}

int vdp_cvi_grad_ve_prior_params(const vdp_cvi_problem* p, const double* path_m, const double* path_S, int clip,
                                 double clip_lo, double clip_hi, double* grads, void* stream) {
    int rc = check_problem(p);
    if (rc) return rc;
    VDP_REQUIRE(grads != nullptr && path_m != nullptr && path_S != nullptr && p->B <= 65535,
                "grads, path_m, path_S, B <= 65535");
    VDP_REQUIRE(path_m != p->fx_mus && path_S != p->fx_covs, "fx_mus / fx_covs must be scratch buffers, not the path");
    if (p->B == 0) return VDP_OK;
    CviWs w;
    if (!cvi_ws_take(p->workspace, p->workspace_bytes, p->B, p->N, p->D, w)) {
        set_error("workspace too small: need %zu bytes", cvi_ws_bytes(p->B, p->N, p->D));
        return VDP_E_WORKSPACE;
    }
    VDP_DISPATCH_D(p->D, VDP_DISPATCH_DRIFT(p->drift_id, {
        rc = run_ve_prior_grad<kDrift, kD>(*p, w, path_m, path_S, clip, clip_lo, clip_hi, grads, as_stream(stream));
        if (rc) return rc;
    }));
    return VDP_OK;
}

}  // extern "C"
