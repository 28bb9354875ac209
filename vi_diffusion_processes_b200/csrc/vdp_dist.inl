// Time-segment sharding of ONE long sequence over ranks (BASELINE config 5): every rank owns a contiguous range
// of grid points and runs the tiled sweeps on it; what crosses ranks is one associative element per rank and
// sweep (plus the natural initial state of the first / last rank), exchanged by the caller (all-gather over
// NCCL through torch.distributed -- a few hundred bytes).  Per sweep:
//
//     vdp_dist_sweep_summary   local segment elements -> this rank's total element (+ natural init state)
//     [caller: all-gather of the (total, init) blobs of all ranks]
//     vdp_dist_sweep_finish    this rank's incoming carry from the other ranks' totals -> segment carries ->
//                              the reference-order recursion over the local range
//
// The backward half of naturals_to_ssm_params has two dependent recursions (Sigma, then y through A(Sigma)):
// stage 0 summarises Sigma, stage 1 (given the gathered stage-0 blobs) summarises y, finish needs both.
// Included at the end of vdp_all.inl.  B = 1 (one sequence); local arrays carry the halo slots documented in
// include/vdp_b200.h.

namespace vdp {

// staging buffers of the time-sharded naturals_to_ssm sweep (its D = 3 tile is 60 KB: a second buffer leaves one CTA
// of 64 threads per SM)
#ifndef VDP_DIST_STAGES_N
#define VDP_DIST_STAGES_N 1
#endif
constexpr int kDistStagesN = VDP_DIST_STAGES_N;
// segments a rank's stretch is split into.  The D = 3 tiles are large (Cholesky 92 KB, naturals_to_ssm 60-74 KB of shared
// memory per CTA: 2 and 3 resident CTAs per SM), so 4 CTAs per SM worth of segments run as 1.3-2 waves; 6 per SM is a
// whole number of waves for both (measured on the 10^7-step chain: 592 segments 20.3 ms, 444: 19.5, 888: 19.3)
#ifndef VDP_DIST_TARGET3
#define VDP_DIST_TARGET3 (148 * 6)
#endif
template <int D>
constexpr int64_t kDistTarget = (D == 3) ? VDP_DIST_TARGET3 : kTargetBlocks;
// grid steps per thread of the information-form sweep and of the moment sweep (three staged arrays each: 27 KB per
// CTA of 32 threads in D = 3 at 5 steps per thread).  Measured on the 10^7-step chain, ms per update (64 threads, 888
// segments unless noted): L = 3 / 3: 12.6; 5 / 3: 11.4; 3 / 5: 12.1; 5 / 5 with 592 segments: 9.9; 32 threads, 1184
// segments, 5 / 5: 9.5 (3 / 3: 10.7; 7 / 7: 10.8; 5 / 9: 9.7; 888 segments: 10.7; 1776: 10.4; 2368: 9.7)
// (profiles/r02b_tune_c5_tile_shapes.log)
#ifndef VDP_LINFO3
#define VDP_LINFO3 5
#endif
#ifndef VDP_LMOM3
#define VDP_LMOM3 5
#endif
#ifndef VDP_TINFO3
#define VDP_TINFO3 32
#endif
template <int D>
constexpr int kInfoL = (D == 3) ? VDP_LINFO3 : TileCfg<D>::Ln;
template <int D>
constexpr int kMomentsL = (D == 3) ? VDP_LMOM3 : TileCfg<D>::Lm;
// threads per CTA of the two sweeps: one warp in D = 3 (the block scan is then a warp scan with no barrier and no
// shared-memory round; 27 KB of shared memory per CTA leave 8 of them per SM)
template <int D>
constexpr int kInfoT = (D == 3) ? VDP_TINFO3 : TileCfg<D>::T;
// ... and the segments their stretch is split into: 8 per SM
#ifndef VDP_INFO_TARGET3
#define VDP_INFO_TARGET3 (148 * 8)
#endif
template <int D>
constexpr int64_t kInfoTarget = (D == 3) ? VDP_INFO_TARGET3 : kTargetBlocks;

template <class Ops>
__global__ void k_seg_total(const typename Ops::Elem* __restrict__ el, int nseg, int reverse,
                            typename Ops::Elem* __restrict__ total) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    typename Ops::Elem t = Ops::identity();
    for (int i = 0; i < nseg; ++i) t = Ops::combine(t, el[reverse ? nseg - 1 - i : i]);
    *total = t;
}

// composition of all nseg segment elements of the rank's stretch (block-wide scan when there are many)
template <class Ops>
static inline void launch_seg_total(const typename Ops::Elem* el, int nseg, int reverse, typename Ops::Elem* total,
                                    cudaStream_t st) {
    if (nseg >= 16) k_seg_total_par<Ops><<<1, kSegScanThreads, 0, st>>>(el, nseg, reverse, total);
    else k_seg_total<Ops><<<1, 32, 0, st>>>(el, nseg, reverse, total);
}

// incoming carry of `rank`: the natural initial state (held by rank 0 forward / rank G-1 backward) pushed
// through the totals of the ranks that come before it in sweep direction
template <class Ops>
__global__ void k_rank_init(const char* __restrict__ totals, size_t total_stride, const char* __restrict__ inits,
                            size_t init_stride, int rank, int nranks, int reverse,
                            typename Ops::State* __restrict__ out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int src = reverse ? nranks - 1 : 0;
    typename Ops::State st = *reinterpret_cast<const typename Ops::State*>(inits + (size_t)src * init_stride);
    if (!reverse) {
        for (int r = 0; r < rank; ++r)
            st = Ops::apply(*reinterpret_cast<const typename Ops::Elem*>(totals + (size_t)r * total_stride), st);
    } else {
        for (int r = nranks - 1; r > rank; --r)
            st = Ops::apply(*reinterpret_cast<const typename Ops::Elem*>(totals + (size_t)r * total_stride), st);
    }
    *out = st;
}

template <class Ops>
struct DistBufs {
    typename Ops::State* carry;
    typename Ops::State* init;
    typename Ops::Elem* el;
};
template <class Ops>
static bool dist_take(SegScratch& sc, int nseg, DistBufs<Ops>& b) {
    b.carry = sc.take<typename Ops::State>((size_t)nseg);
    b.init = sc.take<typename Ops::State>(1);
    b.el = sc.take<typename Ops::Elem>((size_t)nseg);
    return b.el != nullptr;
}

// local carries from the gathered blobs: rank-level incoming carry, then the local segment scan
template <class Ops>
static int dist_carries(const DistBufs<Ops>& b, const void* totals, const void* inits, int rank, int nranks,
                        int reverse, int nseg, cudaStream_t st) {
    k_rank_init<Ops><<<1, 32, 0, st>>>((const char*)totals, sizeof(typename Ops::Elem), (const char*)inits,
                                       sizeof(typename Ops::State), rank, nranks, reverse, b.init);
    VDP_CHECK_LAUNCH();
    return launch_seg_scan<Ops>(b.el, b.init, b.carry, 1, nseg, reverse, st);
}

template <int D>
static int dist_run(const vdp_dist_sweep& p, int phase, void* total_blob, void* init_blob, const void* totals,
                    const void* inits, cudaStream_t st) {
    // phase 0 = summary (stage p.stage), phase 1 = finish
    using C = TileCfg<D>;
    const int64_t N = p.N_local;
    const int first = p.rank == 0, last = p.rank == p.nranks - 1;
    SegScratch sc(p.workspace, p.workspace_bytes);
    int rc;
    switch (p.kind) {
    case VDP_SWEEP_CHOL: {
        using Ops = SchurOps<D>;
        constexpr size_t smem = chol_tiled_smem<D, C::T, C::Lc, kSweepStages>();
        const SegPlan pl = plan_segments(1, N, C::T * C::Lc, true, kDistTarget<D>);
        DistBufs<Ops> b;
        if (!dist_take<Ops>(sc, pl.nseg, b)) break;
        if (phase == 0) {
            if ((rc = allow_smem(k_chol_tiled<D, C::T, C::Lc, SWEEP_SUMMARY, kSweepStages>, smem))) return rc;
            k_chol_tiled<D, C::T, C::Lc, SWEEP_SUMMARY, kSweepStages><<<pl.nseg, C::T, smem, st>>>(
                p.in[0], p.in[1], p.scaleD, p.scaleS, p.out[0], p.out[1], 1, N, pl.nseg, pl.seg_len, nullptr,
                (Mat<D>*)init_blob, b.el, p.status, first);
            VDP_CHECK_LAUNCH();
            launch_seg_total<Ops>(b.el, pl.nseg, 0, (typename Ops::Elem*)total_blob, st);
        } else {
            if ((rc = dist_carries<Ops>(b, totals, inits, p.rank, p.nranks, 0, pl.nseg, st))) return rc;
            if ((rc = allow_smem(k_chol_tiled<D, C::T, C::Lc, SWEEP_FULL, kSweepStages>, smem))) return rc;
            k_chol_tiled<D, C::T, C::Lc, SWEEP_FULL, kSweepStages><<<pl.nseg, C::T, smem, st>>>(
                p.in[0], p.in[1], p.scaleD, p.scaleS, p.out[0], p.out[1], 1, N, pl.nseg, pl.seg_len, b.carry, nullptr,
                nullptr, p.status, first);
        }
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    case VDP_SWEEP_MARG: {
        using Ops = MargOps<D>;
        constexpr size_t smem = marg_tiled_smem<D, C::T, C::Lm>();
        const SegPlan pl = plan_segments(1, N, C::T * C::Lm, true, kDistTarget<D>);
        DistBufs<Ops> b;
        if (!dist_take<Ops>(sc, pl.nseg, b)) break;
        if (phase == 0) {
            if ((rc = allow_smem(k_marg_tiled<D, C::T, C::Lm, SWEEP_SUMMARY>, smem))) return rc;
            k_marg_tiled<D, C::T, C::Lm, SWEEP_SUMMARY><<<pl.nseg, C::T, smem, st>>>(
                p.in[0], p.in[1], p.in[2], p.in[3], p.in[4], p.out[0], p.out[1], p.out[2], 1, N, pl.nseg, pl.seg_len,
                nullptr, (typename Ops::State*)init_blob, b.el, p.status, first, last);
            VDP_CHECK_LAUNCH();
            launch_seg_total<Ops>(b.el, pl.nseg, 0, (typename Ops::Elem*)total_blob, st);
        } else {
            if ((rc = dist_carries<Ops>(b, totals, inits, p.rank, p.nranks, 0, pl.nseg, st))) return rc;
            if ((rc = allow_smem(k_marg_tiled<D, C::T, C::Lm, SWEEP_FULL>, smem))) return rc;
            k_marg_tiled<D, C::T, C::Lm, SWEEP_FULL><<<pl.nseg, C::T, smem, st>>>(
                p.in[0], p.in[1], p.in[2], p.in[3], p.in[4], p.out[0], p.out[1], p.out[2], 1, N, pl.nseg, pl.seg_len,
                b.carry, nullptr, nullptr, p.status, first, last);
        }
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    case VDP_SWEEP_INVBAND: {
        using Ops = AffSymOps<D>;
        constexpr size_t smem = invband_tiled_smem<D, C::T, C::Li>();
        const SegPlan pl = plan_segments(1, N, C::T * C::Li, true, kDistTarget<D>);
        DistBufs<Ops> b;
        if (!dist_take<Ops>(sc, pl.nseg, b)) break;
        if (phase == 0) {
            if ((rc = allow_smem(k_invband_tiled<D, C::T, C::Li, SWEEP_SUMMARY>, smem))) return rc;
            k_invband_tiled<D, C::T, C::Li, SWEEP_SUMMARY><<<pl.nseg, C::T, smem, st>>>(
                p.in[0], p.in[1], p.out[0], p.out[1], 1, N, pl.nseg, pl.seg_len, nullptr, (Mat<D>*)init_blob, b.el,
                last);
            VDP_CHECK_LAUNCH();
            launch_seg_total<Ops>(b.el, pl.nseg, 1, (typename Ops::Elem*)total_blob, st);
        } else {
            if ((rc = dist_carries<Ops>(b, totals, inits, p.rank, p.nranks, 1, pl.nseg, st))) return rc;
            if ((rc = allow_smem(k_invband_tiled<D, C::T, C::Li, SWEEP_FULL>, smem))) return rc;
            k_invband_tiled<D, C::T, C::Li, SWEEP_FULL><<<pl.nseg, C::T, smem, st>>>(
                p.in[0], p.in[1], p.out[0], p.out[1], 1, N, pl.nseg, pl.seg_len, b.carry, nullptr, nullptr, last);
        }
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    case VDP_SWEEP_NAT2SSM: {
        using OpsS = AffSymOps<D>;
        using OpsY = AffVecOps<D>;
        constexpr size_t smem = nat2ssm_tiled_smem<D, C::T, C::Ln, false, false, kDistStagesN>();
        const SegPlan pl = plan_segments(1, N, C::T * C::Ln, true, kDistTarget<D>);
        DistBufs<OpsS> bs;
        DistBufs<OpsY> by;
        if (!dist_take<OpsS>(sc, pl.nseg, bs) || !dist_take<OpsY>(sc, pl.nseg, by)) break;
#define VDP_N2S_ARGS(MODE_)                                                                                        \
    k_nat2ssm_tiled<D, C::T, C::Ln, MODE_, false, false, kDistStagesN><<<pl.nseg, C::T, smem, st>>>(                                             \
        p.in[0], p.in[1], p.in[2], p.in[3], p.in[4], p.out[0], p.out[1], p.out[2], p.out[3], p.out[4], 1, N, pl.nseg,  \
        pl.seg_len, bs.carry, by.carry, (Mat<D>*)init_blob, (Vec<D>*)init_blob, bs.el, by.el, p.status, first, last)
        if (phase == 0 && p.stage == 0) {
            if ((rc = allow_smem(k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_SUMMARY, false, false, kDistStagesN>, smem))) return rc;
            VDP_N2S_ARGS(SWEEP_SUMMARY);
            VDP_CHECK_LAUNCH();
            launch_seg_total<OpsS>(bs.el, pl.nseg, 1, (typename OpsS::Elem*)total_blob, st);
        } else if (phase == 0) {
            if ((rc = dist_carries<OpsS>(bs, p.prev_totals, p.prev_inits, p.rank, p.nranks, 1, pl.nseg, st))) return rc;
            if ((rc = allow_smem(k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_SUMMARY2, false, false, kDistStagesN>, smem))) return rc;
            VDP_N2S_ARGS(SWEEP_SUMMARY2);
            VDP_CHECK_LAUNCH();
            launch_seg_total<OpsY>(by.el, pl.nseg, 1, (typename OpsY::Elem*)total_blob, st);
        } else {
            // the Sigma carries were left in the workspace by the stage-1 summary
            if ((rc = dist_carries<OpsY>(by, totals, inits, p.rank, p.nranks, 1, pl.nseg, st))) return rc;
            if (p.out[5]) {   // marginal covariances on the way
                constexpr size_t smem_cov = nat2ssm_tiled_smem<D, C::T, C::Ln, false, true, kDistStagesN>();
                if ((rc = allow_smem(k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_FULL, false, true, kDistStagesN>, smem_cov)))
                    return rc;
                k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_FULL, false, true, kDistStagesN><<<pl.nseg, C::T, smem_cov, st>>>(
                    p.in[0], p.in[1], p.in[2], p.in[3], p.in[4], p.out[0], p.out[1], p.out[2], p.out[3], p.out[4], 1, N,
                    pl.nseg, pl.seg_len, bs.carry, by.carry, (Mat<D>*)init_blob, (Vec<D>*)init_blob, bs.el, by.el,
                    p.status, first, last, p.out[5]);
            } else {
                if ((rc = allow_smem(k_nat2ssm_tiled<D, C::T, C::Ln, SWEEP_FULL, false, false, kDistStagesN>, smem)))
                    return rc;
                VDP_N2S_ARGS(SWEEP_FULL);
            }
        }
#undef VDP_N2S_ARGS
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    case VDP_SWEEP_MEAN: {
        using Ops = AffVecOps<D>;
        constexpr int L = (D == 1) ? C::Lm : 2 * C::Lm;
        constexpr size_t smem = mean_tiled_smem<D, C::T, L, false, kStagesM>();
        const SegPlan pl = plan_segments(1, N, C::T * L, true, kDistTarget<D>);
        DistBufs<Ops> b;
        if (!dist_take<Ops>(sc, pl.nseg, b)) break;
        if (phase == 0) {
            if ((rc = allow_smem(k_mean_tiled<D, C::T, L, SWEEP_SUMMARY, -1, kStagesM>, smem))) return rc;
            k_mean_tiled<D, C::T, L, SWEEP_SUMMARY, -1, kStagesM><<<pl.nseg, C::T, smem, st>>>(
                p.in[0], p.in[1], p.in[2], p.out[0], 1, N, pl.nseg, pl.seg_len, nullptr, (Vec<D>*)init_blob, b.el,
                MeanKlArgs(), first);
            VDP_CHECK_LAUNCH();
            launch_seg_total<Ops>(b.el, pl.nseg, 0, (typename Ops::Elem*)total_blob, st);
        } else {
            if ((rc = dist_carries<Ops>(b, totals, inits, p.rank, p.nranks, 0, pl.nseg, st))) return rc;
            if ((rc = allow_smem(k_mean_tiled<D, C::T, L, SWEEP_FULL, -1, kStagesM>, smem))) return rc;
            k_mean_tiled<D, C::T, L, SWEEP_FULL, -1, kStagesM><<<pl.nseg, C::T, smem, st>>>(
                p.in[0], p.in[1], p.in[2], p.out[0], 1, N, pl.nseg, pl.seg_len, b.carry, nullptr, nullptr, MeanKlArgs(),
                first);
        }
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    case VDP_SWEEP_INFO: {
        using Ops = InfoOps<D>;
        constexpr int TI = kInfoT<D>;
        constexpr int L = kInfoL<D>;
        constexpr size_t smem = info_tiled_smem<D, TI, L, 1>();
        const SegPlan pl = plan_segments(1, N, TI * L, true, kInfoTarget<D>);
        DistBufs<Ops> b;
        if (!dist_take<Ops>(sc, pl.nseg, b)) break;
        if (phase == 0) {
            if ((rc = allow_smem(k_info_tiled<D, TI, L, SWEEP_SUMMARY, 1>, smem))) return rc;
            k_info_tiled<D, TI, L, SWEEP_SUMMARY, 1><<<pl.nseg, TI, smem, st>>>(
                p.in[0], p.in[1], p.in[2], p.out[0], p.out[1], p.out[2], p.out[3], p.out[4], 1, N, pl.nseg, pl.seg_len,
                nullptr, (typename Ops::State*)init_blob, b.el, p.status, first, last);
            VDP_CHECK_LAUNCH();
            launch_seg_total<Ops>(b.el, pl.nseg, 1, (typename Ops::Elem*)total_blob, st);
        } else {
            if ((rc = dist_carries<Ops>(b, totals, inits, p.rank, p.nranks, 1, pl.nseg, st))) return rc;
            if ((rc = allow_smem(k_info_tiled<D, TI, L, SWEEP_FULL, 1>, smem))) return rc;
            k_info_tiled<D, TI, L, SWEEP_FULL, 1><<<pl.nseg, TI, smem, st>>>(
                p.in[0], p.in[1], p.in[2], p.out[0], p.out[1], p.out[2], p.out[3], p.out[4], 1, N, pl.nseg, pl.seg_len,
                b.carry, nullptr, nullptr, p.status, first, last);
        }
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    case VDP_SWEEP_MOMENTS: {
        using Ops = MomOps<D>;
        constexpr int TI = kInfoT<D>;
        constexpr int L = kMomentsL<D>;
        constexpr size_t smem = moments_tiled_smem<D, TI, L, 1>();
        const SegPlan pl = plan_segments(1, N, TI * L, true, kInfoTarget<D>);
        DistBufs<Ops> b;
        if (!dist_take<Ops>(sc, pl.nseg, b)) break;
        if (phase == 0) {
            if ((rc = allow_smem(k_moments_tiled<D, TI, L, SWEEP_SUMMARY, 1>, smem))) return rc;
            k_moments_tiled<D, TI, L, SWEEP_SUMMARY, 1><<<pl.nseg, TI, smem, st>>>(
                p.in[0], p.in[1], p.in[2], p.in[3], p.in[4], p.out[0], p.out[1], 1, N, pl.nseg, pl.seg_len, nullptr,
                (typename Ops::State*)init_blob, b.el, first);
            VDP_CHECK_LAUNCH();
            launch_seg_total<Ops>(b.el, pl.nseg, 0, (typename Ops::Elem*)total_blob, st);
        } else {
            if ((rc = dist_carries<Ops>(b, totals, inits, p.rank, p.nranks, 0, pl.nseg, st))) return rc;
            if ((rc = allow_smem(k_moments_tiled<D, TI, L, SWEEP_FULL, 1>, smem))) return rc;
            k_moments_tiled<D, TI, L, SWEEP_FULL, 1><<<pl.nseg, TI, smem, st>>>(
                p.in[0], p.in[1], p.in[2], p.in[3], p.in[4], p.out[0], p.out[1], 1, N, pl.nseg, pl.seg_len, b.carry,
                nullptr, nullptr, first);
        }
        VDP_CHECK_LAUNCH();
        return VDP_OK;
    }
    default:
        set_error("unknown sweep kind %d", p.kind);
        return VDP_E_BADARG;
    }
    set_error("workspace too small for the distributed sweep");
    return VDP_E_WORKSPACE;
}

static int check_dist(const vdp_dist_sweep* p) {
    VDP_REQUIRE(p != nullptr, "sweep");
    VDP_REQUIRE(p->nranks >= 1 && p->rank >= 0 && p->rank < p->nranks, "rank, nranks");
    VDP_REQUIRE(p->N_local >= 1, "N_local");
    VDP_REQUIRE(p->workspace != nullptr, "workspace");
    return VDP_OK;
}

}  // namespace vdp

extern "C" {

size_t vdp_dist_workspace(int D) { (void)D; return kSegScratchBytes; }

int vdp_dist_blob_bytes(int kind, int stage, int D, size_t* total_bytes, size_t* init_bytes) {
    VDP_REQUIRE(total_bytes && init_bytes, "outputs");
    VDP_DISPATCH_D(D, {
        switch (kind) {
        case VDP_SWEEP_CHOL: *total_bytes = sizeof(SchurElem<kD>); *init_bytes = sizeof(Mat<kD>); break;
        case VDP_SWEEP_MARG: *total_bytes = sizeof(typename MargOps<kD>::Elem); *init_bytes = sizeof(typename MargOps<kD>::State); break;
        case VDP_SWEEP_INVBAND: *total_bytes = sizeof(AffSymElem<kD>); *init_bytes = sizeof(Mat<kD>); break;
        case VDP_SWEEP_MEAN: *total_bytes = sizeof(AffVecElem<kD>); *init_bytes = sizeof(Vec<kD>); break;
        case VDP_SWEEP_INFO: *total_bytes = sizeof(InfoElem<kD>); *init_bytes = sizeof(InfoState<kD>); break;
        case VDP_SWEEP_MOMENTS: *total_bytes = sizeof(typename MomOps<kD>::Elem); *init_bytes = sizeof(typename MomOps<kD>::State); break;
        case VDP_SWEEP_NAT2SSM:
            if (stage == 0) { *total_bytes = sizeof(AffSymElem<kD>); *init_bytes = sizeof(Mat<kD>); }
            else { *total_bytes = sizeof(AffVecElem<kD>); *init_bytes = sizeof(Vec<kD>); }
            break;
        default: set_error("unknown sweep kind %d", kind); return VDP_E_BADARG;
        }
    });
    return VDP_OK;
}

int vdp_dist_sweep_summary(const vdp_dist_sweep* p, void* total_blob, void* init_blob, void* stream) {
    int rc = check_dist(p);
    if (rc) return rc;
    VDP_REQUIRE(total_blob && init_blob, "blobs");
    VDP_DISPATCH_D(p->D, { if ((rc = dist_run<kD>(*p, 0, total_blob, init_blob, nullptr, nullptr, as_stream(stream)))) return rc; });
    return VDP_OK;
}

int vdp_dist_sweep_finish(const vdp_dist_sweep* p, const void* all_totals, const void* all_inits, void* stream) {
    int rc = check_dist(p);
    if (rc) return rc;
    VDP_REQUIRE(all_totals && all_inits, "gathered blobs");
    VDP_DISPATCH_D(p->D, { if ((rc = dist_run<kD>(*p, 1, nullptr, nullptr, all_totals, all_inits, as_stream(stream)))) return rc; });
    return VDP_OK;
}

}  // extern "C"
