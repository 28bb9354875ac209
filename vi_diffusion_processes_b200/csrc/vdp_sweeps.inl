// Host-side launch logic of the four time sweeps: choice between the thread-per-sequence kernels
// (btd_kernels.cuh; many short sequences) and the parallel-in-time tiled kernels (tile_sweeps.cuh), and
// the segment plan of the latter.  Included by vdp_all.inl before vdp_fused.inl.

namespace vdp {

// tile shapes: T chunks of L steps; L odd (conflict-free chunk stride in shared memory)
template <int D> struct TileCfg;
// (the D = 1 shape can be overridden at build time for tuning runs: -DVDP_T1=.. -DVDP_LC1=.. ...)
#ifndef VDP_T1
#define VDP_T1 128
#endif
#ifndef VDP_LC1
#define VDP_LC1 9
#endif
#ifndef VDP_LI1
#define VDP_LI1 17
#endif
#ifndef VDP_LM1
#define VDP_LM1 9
#endif
#ifndef VDP_LN1
#define VDP_LN1 9
#endif
template <> struct TileCfg<1> { static constexpr int T = VDP_T1, Lc = VDP_LC1, Li = VDP_LI1, Lm = VDP_LM1, Ln = VDP_LN1; };
template <> struct TileCfg<2> { static constexpr int T = 64, Lc = 9, Li = 9, Lm = 5, Ln = 5; };
#ifndef VDP_T3
#define VDP_T3 64
#endif
#ifndef VDP_LC3
#define VDP_LC3 5
#endif
#ifndef VDP_LM3
#define VDP_LM3 3
#endif
#ifndef VDP_LN3
#define VDP_LN3 3
#endif
template <> struct TileCfg<3> { static constexpr int T = VDP_T3, Lc = VDP_LC3, Li = 5, Lm = VDP_LM3, Ln = VDP_LN3; };

// staging buffers per sweep kernel: 2 = the copies of the next tile land while the current one is processed
#ifndef VDP_SWEEP_STAGES
#define VDP_SWEEP_STAGES 2
#endif
#ifndef VDP_STAGES_C
#define VDP_STAGES_C VDP_SWEEP_STAGES
#endif
// (measured on B200, C4 shape, T = 128: Cholesky L = 9 with 2 stages 0.228 ms (L = 17: 0.257; 1 stage: 0.232);
// naturals_to_ssm L = 9 with 1 stage 0.407 ms (2 stages halve the CTAs per SM: 0.538; L = 5 with 2 stages 0.427);
// means L = 9 with 1 stage 0.216 ms (2 stages 0.226): the ring only pays where it does not cost occupancy)
#ifndef VDP_STAGES_N
#define VDP_STAGES_N 1
#endif
#ifndef VDP_STAGES_M
#define VDP_STAGES_M 1
#endif
constexpr int kSweepStages = VDP_STAGES_C;   // Cholesky sweep
constexpr int kStagesN = VDP_STAGES_N;       // naturals_to_ssm backward sweep
constexpr int kStagesM = VDP_STAGES_M;       // mean sweep

#ifndef VDP_TARGET_BLOCKS
#define VDP_TARGET_BLOCKS (148 * 4)
#endif
constexpr int64_t kTargetBlocks = VDP_TARGET_BLOCKS;   // CTAs wanted in flight before sequences are split in time
constexpr int64_t kMaxSegTotal = 4096;         // bound on B * nseg (sizes the segment scratch)
constexpr size_t kSegScratchBytes = 8u << 20;  // part of every fused-level workspace

struct SegScratch {
    char* base;
    size_t size, used;
    SegScratch(void* p, size_t n) : base((char*)p), size(p ? n : 0), used(0) {}
    template <class X> X* take(size_t count) {
        const size_t bytes = (count * sizeof(X) + 255) & ~size_t(255);
        if (used + bytes > size) return nullptr;
        X* r = reinterpret_cast<X*>(base + used);
        used += bytes;
        return r;
    }
};

struct SegPlan {
    int nseg;
    int64_t seg_len;
};

// vdp_config_set("SWEEPS", 1 | 2) overrides the heuristic (tests run both mappings)
static inline bool use_tiled_sweeps(int64_t B, int64_t N) {
    if (g_cfg.sweeps == 1) return false;
    if (g_cfg.sweeps == 2) return N >= 1;
    return N >= 512 || (B < 148 * 64 && N >= 32);
}

static inline SegPlan plan_segments(int64_t B, int64_t N, int tile_steps, bool have_scratch,
                                    int64_t target_blocks = kTargetBlocks) {
    int64_t want = 1;
    const int forced = g_cfg.nseg;
    if (forced > 0) {
        want = forced;
    } else if (B < target_blocks) {
        want = (target_blocks + B - 1) / B;
        want = std::min<int64_t>(want, std::max<int64_t>(1, N / tile_steps));   // >= one full tile per segment
    }
    want = std::min<int64_t>(want, std::max<int64_t>(1, kMaxSegTotal / std::max<int64_t>(B, 1)));
    want = std::min<int64_t>(want, N);
    if (!have_scratch) want = 1;
    SegPlan p;
    p.seg_len = (N + want - 1) / want;
    p.nseg = (int)((N + p.seg_len - 1) / p.seg_len);
    return p;
}

template <class K>
static inline int allow_smem(K kernel, size_t bytes) {
    if (bytes > 48 * 1024)
        VDP_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return VDP_OK;
}

template <class Ops>
static inline int launch_seg_scan(const typename Ops::Elem* el, const typename Ops::State* init,
                                  typename Ops::State* carry, int64_t B, int nseg, int reverse, cudaStream_t st) {
    // few sequences, many segments: scan the segments of a sequence in parallel
    if (B <= 16 && nseg >= 16) {
        k_seg_scan_par<Ops><<<(unsigned)B, kSegScanThreads, 0, st>>>(el, init, carry, nseg, reverse);
    } else {
        k_seg_scan<Ops><<<grid_for(B, 32), 32, 0, st>>>(el, init, carry, B, nseg, reverse);
    }
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// ---- cholesky_band ---------------------------------------------------------------------------------
template <int D>
static int run_chol_tiled(const double* Xd, const double* Xs, double scaleD, double scaleS, double* Ld, double* Ls,
                          int64_t B, int64_t N, int64_t* status, SegScratch sc, cudaStream_t st) {
    using C = TileCfg<D>;
    constexpr size_t smem = chol_tiled_smem<D, C::T, C::Lc, kSweepStages>();
    const SegPlan pl = plan_segments(B, N, C::T * C::Lc, sc.size > 0);
    const unsigned grid = (unsigned)(B * pl.nseg);
    int rc;
    if ((rc = allow_smem(k_chol_tiled<D, C::T, C::Lc, SWEEP_FULL, kSweepStages>, smem))) return rc;
    if (pl.nseg == 1) {
        k_chol_tiled<D, C::T, C::Lc, SWEEP_FULL, kSweepStages><<<grid, C::T, smem, st>>>(Xd, Xs, scaleD, scaleS, Ld, Ls, B, N, 1,
                                                                           pl.seg_len, nullptr, nullptr, nullptr, status);
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    Mat<D>* carry = sc.take<Mat<D>>((size_t)B * pl.nseg);
    Mat<D>* init = sc.take<Mat<D>>((size_t)B);
    SchurElem<D>* el = sc.take<SchurElem<D>>((size_t)B * pl.nseg);
    if (!el) { set_error("segment scratch too small"); return VDP_E_WORKSPACE; }
    if ((rc = allow_smem(k_chol_tiled<D, C::T, C::Lc, SWEEP_SUMMARY, kSweepStages>, smem))) return rc;
    k_chol_tiled<D, C::T, C::Lc, SWEEP_SUMMARY, kSweepStages><<<grid, C::T, smem, st>>>(Xd, Xs, scaleD, scaleS, Ld, Ls, B, N, pl.nseg,
                                                                          pl.seg_len, nullptr, init, el, status);
    VDP_CHECK_LAUNCH();
    if ((rc = launch_seg_scan<SchurOps<D>>(el, init, carry, B, pl.nseg, 0, st))) return rc;
    k_chol_tiled<D, C::T, C::Lc, SWEEP_FULL, kSweepStages><<<grid, C::T, smem, st>>>(Xd, Xs, scaleD, scaleS, Ld, Ls, B, N, pl.nseg,
                                                                       pl.seg_len, carry, nullptr, nullptr, status);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// ---- inverse_from_cholesky_band --------------------------------------------------------------------
template <int D>
static int run_invband_tiled(const double* Ld, const double* Ls, double* Sd, double* Ss, int64_t B, int64_t N,
                             SegScratch sc, cudaStream_t st) {
    using C = TileCfg<D>;
    constexpr size_t smem = invband_tiled_smem<D, C::T, C::Li>();
    const SegPlan pl = plan_segments(B, N, C::T * C::Li, sc.size > 0);
    const unsigned grid = (unsigned)(B * pl.nseg);
    int rc;
    if ((rc = allow_smem(k_invband_tiled<D, C::T, C::Li, SWEEP_FULL>, smem))) return rc;
    if (pl.nseg == 1) {
        k_invband_tiled<D, C::T, C::Li, SWEEP_FULL><<<grid, C::T, smem, st>>>(Ld, Ls, Sd, Ss, B, N, 1, pl.seg_len,
                                                                              nullptr, nullptr, nullptr);
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    Mat<D>* carry = sc.take<Mat<D>>((size_t)B * pl.nseg);
    Mat<D>* init = sc.take<Mat<D>>((size_t)B);
    AffSymElem<D>* el = sc.take<AffSymElem<D>>((size_t)B * pl.nseg);
    if (!el) { set_error("segment scratch too small"); return VDP_E_WORKSPACE; }
    if ((rc = allow_smem(k_invband_tiled<D, C::T, C::Li, SWEEP_SUMMARY>, smem))) return rc;
    k_invband_tiled<D, C::T, C::Li, SWEEP_SUMMARY><<<grid, C::T, smem, st>>>(Ld, Ls, Sd, Ss, B, N, pl.nseg, pl.seg_len,
                                                                             nullptr, init, el);
    VDP_CHECK_LAUNCH();
    if ((rc = launch_seg_scan<AffSymOps<D>>(el, init, carry, B, pl.nseg, 1, st))) return rc;
    k_invband_tiled<D, C::T, C::Li, SWEEP_FULL><<<grid, C::T, smem, st>>>(Ld, Ls, Sd, Ss, B, N, pl.nseg, pl.seg_len,
                                                                          carry, nullptr, nullptr);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// ---- naturals_to_ssm_params, backward half ---------------------------------------------------------
template <int D, bool COV>
static int run_nat2ssm_bwd_tiled_impl(const double* Ld, const double* Ls, const double* th1, const double* thd,
                                      const double* ths, double* A, double* b, double* cholQ, double* mu0,
                                      double* cholP0, int64_t B, int64_t N, int64_t* status, SegScratch sc,
                                      cudaStream_t st, double* Sd) {
    using C = TileCfg<D>;
    constexpr size_t smem = nat2ssm_tiled_smem<D, C::T, C::Ln, true, false, kStagesN>();
    constexpr size_t smem_full = nat2ssm_tiled_smem<D, C::T, C::Ln, true, COV, kStagesN>();
    const SegPlan pl = plan_segments(B, N, C::T * C::Ln, sc.size > 0);
    const unsigned grid = (unsigned)(B * pl.nseg);
    int rc;
    if ((rc = allow_smem(k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_FULL, true, COV, kStagesN>, smem_full))) return rc;
    if (pl.nseg == 1) {
        k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_FULL, true, COV, kStagesN><<<grid, C::T, smem_full, st>>>(
            Ld, Ls, th1, thd, ths, A, b, cholQ, mu0, cholP0, B, N, 1, pl.seg_len, nullptr, nullptr, nullptr, nullptr,
            nullptr, nullptr, status, 1, 1, Sd);
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    Mat<D>* carryS = sc.take<Mat<D>>((size_t)B * pl.nseg);
    Vec<D>* carryY = sc.take<Vec<D>>((size_t)B * pl.nseg);
    Mat<D>* initS = sc.take<Mat<D>>((size_t)B);
    Vec<D>* initY = sc.take<Vec<D>>((size_t)B);
    AffSymElem<D>* elS = sc.take<AffSymElem<D>>((size_t)B * pl.nseg);
    AffVecElem<D>* elY = sc.take<AffVecElem<D>>((size_t)B * pl.nseg);
    if (!elY) { set_error("segment scratch too small"); return VDP_E_WORKSPACE; }
    if ((rc = allow_smem(k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_SUMMARY, true, false, kStagesN>, smem))) return rc;
    if ((rc = allow_smem(k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_SUMMARY2, true, false, kStagesN>, smem))) return rc;
    k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_SUMMARY, true, false, kStagesN><<<grid, C::T, smem, st>>>(
        Ld, Ls, th1, thd, ths, A, b, cholQ, mu0, cholP0, B, N, pl.nseg, pl.seg_len, nullptr, nullptr, initS, initY, elS,
        elY, status);
    VDP_CHECK_LAUNCH();
    if ((rc = launch_seg_scan<AffSymOps<D>>(elS, initS, carryS, B, pl.nseg, 1, st))) return rc;
    k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_SUMMARY2, true, false, kStagesN><<<grid, C::T, smem, st>>>(
        Ld, Ls, th1, thd, ths, A, b, cholQ, mu0, cholP0, B, N, pl.nseg, pl.seg_len, carryS, nullptr, initS, initY, elS,
        elY, status);
    VDP_CHECK_LAUNCH();
    if ((rc = launch_seg_scan<AffVecOps<D>>(elY, initY, carryY, B, pl.nseg, 1, st))) return rc;
    k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_FULL, true, COV, kStagesN><<<grid, C::T, smem_full, st>>>(
        Ld, Ls, th1, thd, ths, A, b, cholQ, mu0, cholP0, B, N, pl.nseg, pl.seg_len, carryS, carryY, nullptr, nullptr,
        nullptr, nullptr, status, 1, 1, Sd);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// Sd != NULL: also write the marginal covariances (the diagonal of the inverse band the sweep computes on its way)
template <int D>
static int run_nat2ssm_bwd_tiled(const double* Ld, const double* Ls, const double* th1, const double* thd,
                                 const double* ths, double* A, double* b, double* cholQ, double* mu0, double* cholP0,
                                 int64_t B, int64_t N, int64_t* status, SegScratch sc, cudaStream_t st,
                                 double* Sd = nullptr) {
    if (Sd) return run_nat2ssm_bwd_tiled_impl<D, true>(Ld, Ls, th1, thd, ths, A, b, cholQ, mu0, cholP0, B, N, status, sc,
                                                       st, Sd);
    return run_nat2ssm_bwd_tiled_impl<D, false>(Ld, Ls, th1, thd, ths, A, b, cholQ, mu0, cholP0, B, N, status, sc, st,
                                                nullptr);
}


// ---- marginal means + Cholesky of the rebuilt precision --------------------------------------------
template <int D>
static int run_marg_fwd_tiled(const double* A, const double* b, const double* cholQ, const double* mu0,
                              const double* cholP0, double* m, double* Ld, double* Ls, int64_t B, int64_t N,
                              int64_t* status, SegScratch sc, cudaStream_t st) {
    using C = TileCfg<D>;
    using Ops = MargOps<D>;
    constexpr size_t smem = marg_tiled_smem<D, C::T, C::Lm>();
    const SegPlan pl = plan_segments(B, N, C::T * C::Lm, sc.size > 0);
    const unsigned grid = (unsigned)(B * pl.nseg);
    int rc;
    if ((rc = allow_smem(k_marg_tiled<D, C::T, C::Lm, SWEEP_FULL>, smem))) return rc;
    if (pl.nseg == 1) {
        k_marg_tiled<D, C::T, C::Lm, SWEEP_FULL><<<grid, C::T, smem, st>>>(A, b, cholQ, mu0, cholP0, m, Ld, Ls, B, N, 1,
                                                                           pl.seg_len, nullptr, nullptr, nullptr, status);
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    typename Ops::State* carry = sc.take<typename Ops::State>((size_t)B * pl.nseg);
    typename Ops::State* init = sc.take<typename Ops::State>((size_t)B);
    typename Ops::Elem* el = sc.take<typename Ops::Elem>((size_t)B * pl.nseg);
    if (!el) { set_error("segment scratch too small"); return VDP_E_WORKSPACE; }
    if ((rc = allow_smem(k_marg_tiled<D, C::T, C::Lm, SWEEP_SUMMARY>, smem))) return rc;
    k_marg_tiled<D, C::T, C::Lm, SWEEP_SUMMARY><<<grid, C::T, smem, st>>>(A, b, cholQ, mu0, cholP0, m, Ld, Ls, B, N,
                                                                          pl.nseg, pl.seg_len, nullptr, init, el, status);
    VDP_CHECK_LAUNCH();
    if ((rc = launch_seg_scan<Ops>(el, init, carry, B, pl.nseg, 0, st))) return rc;
    k_marg_tiled<D, C::T, C::Lm, SWEEP_FULL><<<grid, C::T, smem, st>>>(A, b, cholQ, mu0, cholP0, m, Ld, Ls, B, N, pl.nseg,
                                                                       pl.seg_len, carry, nullptr, nullptr, status);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// ---- marginal means alone (the posterior rebuild of the site update: the precision factor is already there) ----
// DRIFT >= 0 (D = 1, polynomial drift): the sweep also adds up the KL terms of classic_elbo (MeanKlArgs); kl->partials
// receives *n_partials sums per sequence, to be added by k_reduce_terms.
template <int D, int DRIFT = -1>
static int run_mean_fwd_tiled(const double* A, const double* b, const double* mu0, double* m, int64_t B, int64_t N,
                              SegScratch sc, cudaStream_t st, const MeanKlArgs* kl = nullptr,
                              int64_t* n_partials = nullptr) {
    using C = TileCfg<D>;
    using Ops = AffVecOps<D>;
    constexpr bool KL = DRIFT >= 0;
    constexpr int L = (D == 1) ? C::Lm : 2 * C::Lm;   // the same tile (and rounding order of the means) with and without KL
    constexpr size_t smem = mean_tiled_smem<D, C::T, L, false, kStagesM>();
    constexpr size_t smem_full = mean_tiled_smem<D, C::T, L, KL, kStagesM>();
    const SegPlan pl = plan_segments(B, N, C::T * L, sc.size > 0);
    const unsigned grid = (unsigned)(B * pl.nseg);
    MeanKlArgs ka;
    if (KL) {
        ka = *kl;
        ka.tiles_per_seg = (int)((pl.seg_len + C::T * L - 1) / (C::T * L));
        *n_partials = (int64_t)pl.nseg * ka.tiles_per_seg;
        VDP_CHECK_CUDA(cudaMemsetAsync(ka.partials, 0, (size_t)B * *n_partials * sizeof(double), st));
    }
    int rc;
    if ((rc = allow_smem(k_mean_tiled<D, C::T, L, SWEEP_FULL, DRIFT, kStagesM>, smem_full))) return rc;
    if (pl.nseg == 1) {
        k_mean_tiled<D, C::T, L, SWEEP_FULL, DRIFT, kStagesM><<<grid, C::T, smem_full, st>>>(A, b, mu0, m, B, N, 1, pl.seg_len,
                                                                                 nullptr, nullptr, nullptr, ka);
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    typename Ops::State* carry = sc.take<typename Ops::State>((size_t)B * pl.nseg);
    typename Ops::State* init = sc.take<typename Ops::State>((size_t)B);
    typename Ops::Elem* el = sc.take<typename Ops::Elem>((size_t)B * pl.nseg);
    if (!el) { set_error("segment scratch too small"); return VDP_E_WORKSPACE; }
    if ((rc = allow_smem(k_mean_tiled<D, C::T, L, SWEEP_SUMMARY, -1, kStagesM>, smem))) return rc;
    k_mean_tiled<D, C::T, L, SWEEP_SUMMARY, -1, kStagesM><<<grid, C::T, smem, st>>>(A, b, mu0, m, B, N, pl.nseg, pl.seg_len, nullptr,
                                                                    init, el);
    VDP_CHECK_LAUNCH();
    if ((rc = launch_seg_scan<Ops>(el, init, carry, B, pl.nseg, 0, st))) return rc;
    k_mean_tiled<D, C::T, L, SWEEP_FULL, DRIFT, kStagesM><<<grid, C::T, smem_full, st>>>(A, b, mu0, m, B, N, pl.nseg, pl.seg_len,
                                                                             carry, nullptr, nullptr, ka);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

}  // namespace vdp
