// Parallel-in-time ("tiled") versions of the four sequential sweeps of the block-tridiagonal layer.
//
// Mapping: one CTA walks one SEGMENT of one sequence, tile by tile in sweep direction.  A tile is T
// chunks of L consecutive grid steps; it is staged in shared memory with coalesced loads, thread c owns
// chunk c.  Per tile: (A) every thread composes the associative element of its chunk (scan_elems.cuh),
// (scan) a block-wide exclusive scan turns the elements and the tile's incoming carry into one carry
// per chunk, (C) every thread runs the reference-order recursion of sweeps.cuh over its chunk, writing
// into the staged tile, which is then stored with coalesced writes.  L is odd so that the chunk stride
// in shared memory (L * D * D doubles) is conflict-free without padding.
//
// With one segment per sequence (many sequences, BASELINE config 4) the inputs are read from HBM once
// and the sweep is a single launch.  With several segments per sequence (few long sequences, config 5,
// and time-segment sharding over GPUs) the launch sequence is
//     SUMMARY (segment elements)  ->  k_seg_scan (segment carries)  ->  FULL.
//
// Reference operations: see sweeps.cuh (same step functions; only the carries differ in rounding order).
#pragma once
#include "common.cuh"
#include "scan_elems.cuh"
#include "sweeps.cuh"
#include "kl_kernels.cuh"
#include "tma_tile.cuh"

// L2 prefetch of the next tile in the single-buffer sweeps (naturals_to_ssm, means)
#ifndef VDP_PREFETCH_N
#define VDP_PREFETCH_N 0
#endif
#ifndef VDP_PREFETCH_M
#define VDP_PREFETCH_M 1
#endif

namespace vdp {

enum { SWEEP_FULL = 0, SWEEP_SUMMARY = 1, SWEEP_SUMMARY2 = 2 };

// Tile staging: global -> shared with 8-byte cp.async (LDGSTS), so that every load of a tile is in flight
// at once and no register staging limits the memory-level parallelism; any 8-byte alignment works.
__device__ __forceinline__ void cp_async8(double* s, const double* g) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(s);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(sa), "l"(g));
}
template <int T>
__device__ __forceinline__ void stage_in(double* __restrict__ s, const double* __restrict__ g, int64_t count) {
    for (int64_t i = threadIdx.x; i < count; i += T) cp_async8(s + i, g + i);
}
// wait for this thread's copies (a __syncthreads() must follow before other threads' elements are read)
__device__ __forceinline__ void stage_wait() {
    asm volatile("cp.async.commit_group;\n" ::);
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}
// scale the elements this thread staged itself (call between stage_wait and the barrier)
template <int T>
__device__ __forceinline__ void stage_scale(double* __restrict__ s, int64_t count, double scale) {
    if (scale == 1.0) return;
    for (int64_t i = threadIdx.x; i < count; i += T) s[i] = scale * s[i];
}
// start the lines of [g, g + count) on their way to L2 (the next tile of this CTA: its cp.async loads then find them
// there instead of in HBM); one 128-byte line per call
template <int T>
__device__ __forceinline__ void prefetch_range(const double* __restrict__ g, int64_t count) {
    for (int64_t i = (int64_t)threadIdx.x * 16; i < count; i += (int64_t)T * 16)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(g + i));
}
template <int T>
__device__ __forceinline__ void stage_out(double* __restrict__ g, const double* __restrict__ s, int64_t count) {
    for (int64_t i = threadIdx.x; i < count; i += T) g[i] = s[i];
}

// two recursions scanned together
template <class OpsA, class OpsB>
struct PairOps {
    struct Elem { typename OpsA::Elem a; typename OpsB::Elem b; };
    struct State { typename OpsA::State a; typename OpsB::State b; };
    __device__ __forceinline__ static Elem identity() {
        Elem e;
        e.a = OpsA::identity();
        e.b = OpsB::identity();
        return e;
    }
    __device__ __forceinline__ static Elem combine(const Elem& e1, const Elem& e2) {
        Elem r;
        r.a = OpsA::combine(e1.a, e2.a);
        r.b = OpsB::combine(e1.b, e2.b);
        return r;
    }
    __device__ __forceinline__ static State apply(const Elem& e, const State& s) {
        State r;
        r.a = OpsA::apply(e.a, s.a);
        r.b = OpsB::apply(e.b, s.b);
        return r;
    }
};

// carries of the segments of every sequence: state at the start (forward) / top (reverse) of each segment
template <class Ops>
__global__ void k_seg_scan(const typename Ops::Elem* __restrict__ elems, const typename Ops::State* __restrict__ init,
                           typename Ops::State* __restrict__ carry, int64_t B, int nseg, int reverse) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    typename Ops::State st = init[b];
    for (int i = 0; i < nseg; ++i) {
        const int s = reverse ? nseg - 1 - i : i;
        carry[b * nseg + s] = st;
        st = Ops::apply(elems[b * nseg + s], st);
    }
}

// The same carries for FEW sequences with MANY segments (one long chain split in time): one block per sequence, every
// thread takes a run of consecutive segments (in sweep order), composes them, the block scans the run totals -- log
// depth instead of nseg sequential compositions -- and each thread then replays its run from its incoming state with
// `apply`, which is cheaper than `combine`.  256 threads keep the D = 3 elements (up to 40 doubles) in registers.
// (592 segments of a D = 3 chain: 1 ms -> 0.03 ms per scan.)
constexpr int kSegScanThreads = 256;
template <class Ops>
__global__ void __launch_bounds__(kSegScanThreads) k_seg_scan_par(const typename Ops::Elem* __restrict__ elems,
                                                                 const typename Ops::State* __restrict__ init,
                                                                 typename Ops::State* __restrict__ carry, int nseg,
                                                                 int reverse) {
    constexpr int T = kSegScanThreads;
    __shared__ typename Ops::Elem wtot[T / 32];
    const int64_t b = blockIdx.x;
    const int per = (nseg + T - 1) / T;
    const int i0 = threadIdx.x * per, i1 = min(nseg, i0 + per);      // positions in sweep order
    typename Ops::Elem e = Ops::identity();
    for (int i = i0; i < i1; ++i) e = Ops::combine(e, elems[b * nseg + (reverse ? nseg - 1 - i : i)]);
    typename Ops::Elem total;
    const typename Ops::Elem excl = block_exclusive_scan<Ops, T>(e, wtot, total);
    typename Ops::State st = Ops::apply(excl, init[b]);
    for (int i = i0; i < i1; ++i) {
        const int s = reverse ? nseg - 1 - i : i;
        carry[b * nseg + s] = st;
        st = Ops::apply(elems[b * nseg + s], st);
    }
}
// composition of all segment elements of one sequence (the rank total of the time-sharded sweeps)
template <class Ops>
__global__ void __launch_bounds__(kSegScanThreads) k_seg_total_par(const typename Ops::Elem* __restrict__ elems, int nseg,
                                                                  int reverse, typename Ops::Elem* __restrict__ total_out) {
    constexpr int T = kSegScanThreads;
    __shared__ typename Ops::Elem wtot[T / 32];
    const int per = (nseg + T - 1) / T;
    const int i0 = threadIdx.x * per, i1 = min(nseg, i0 + per);
    typename Ops::Elem e = Ops::identity();
    for (int i = i0; i < i1; ++i) e = Ops::combine(e, elems[reverse ? nseg - 1 - i : i]);
    typename Ops::Elem total;
    block_exclusive_scan<Ops, T>(e, wtot, total);
    if (threadIdx.x == 0) *total_out = total;
}

// ---------------------------------------------------------------------------------------------------
// forward Cholesky over one staged tile: sKd holds X_kk for nodes [0, cnt] of the tile (in: X, out: L_kk),
// sKs holds X_{k+1,k} for transitions [0, cnt) (in: X, out: L_{k+1,k}); the factorised matrix is
// K = (scaleD X_d, scaleS X_s) (SCALED; the scales are applied as the operands are read, -2 / -1 for naturals).
template <int D, bool SCALED>
__device__ __forceinline__ Mat<D> tile_block(const double* p, double scale) {
    Mat<D> m = mat_load<D>(p);
    if (SCALED) {
#pragma unroll
        for (int e = 0; e < D * D; ++e) m.a[e] *= scale;
    }
    return m;
}
// Phase A: chunk element.
template <int D, int L, bool SCALED = false>
__device__ __forceinline__ SchurElem<D> chol_tile_summary(const double* sKd, const double* sKs, int k0, int len,
                                                          double scaleD = 1.0, double scaleS = 1.0) {
    constexpr int DD = D * D;
    SchurElem<D> e = SchurOps<D>::identity();
    if constexpr (D == 1) {
        if (len == L) {   // full chunk: preloaded operands, unrolled; Lb carries the only serial dependency
            double ks[L], kd[L];
#pragma unroll
            for (int t = 0; t < L; ++t) {
                ks[t] = SCALED ? scaleS * sKs[k0 + t] : sKs[k0 + t];
                kd[t] = SCALED ? scaleD * sKd[k0 + t + 1] : sKd[k0 + t + 1];
            }
            double La = 0.0, Lab = ks[0], Lb = kd[0];
#pragma unroll
            for (int t = 1; t < L; ++t) {
                const double r = rcp_fast(Lb);
                const double ur = Lab * r;
                La = fma(-ur, Lab, La);
                Lab = ur * ks[t];
                Lb = fma(-(ks[t] * ks[t]), r, kd[t]);
            }
            e.La.a[0] = La; e.Lab.a[0] = Lab; e.Lb.a[0] = Lb; e.id = 0;
            return e;
        }
    }
    for (int t = 0; t < len; ++t)
        SchurOps<D>::push(e, tile_block<D, SCALED>(sKs + (k0 + t) * DD, scaleS),
                          tile_block<D, SCALED>(sKd + (k0 + t + 1) * DD, scaleD));
    return e;
}
// Phase C: reference-order recursion from the chunk's carry S (the Schur complement at node k0).
// `last_node` = local index of the sequence's final node if it is in this tile, else -1.
template <int D, int L, bool SCALED = false>
__device__ __forceinline__ void chol_tile_recursion(double* sKd, double* sKs, int k0, int len, const Mat<D>& S,
                                                    int last_node, int64_t flat_node0, int64_t* status,
                                                    double scaleD = 1.0, double scaleS = 1.0) {
    constexpr int DD = D * D;
    if (len <= 0) return;
    if constexpr (D == 1) {
        // Scalar.  The only step-to-step dependency is the Schur complement S_{t+1} = Kd_{t+1} - Ks_t^2 / S_t
        // (one reciprocal + one fma on the critical path); the factor entries L_t = sqrt(S_t) and
        // L_{t+1,t} = Ks_t / L_t hang off it and overlap with the following steps.
        if (len == L && last_node < 0) {   // full chunk: operands preloaded, loops unrolled
            double ks[L], kd[L];
#pragma unroll
            for (int t = 0; t < L; ++t) {
                ks[t] = SCALED ? scaleS * sKs[k0 + t] : sKs[k0 + t];
                kd[t] = (t < L - 1) ? (SCALED ? scaleD * sKd[k0 + t + 1] : sKd[k0 + t + 1]) : 0.0;
            }
            double s = S.a[0];   // (node k0 + L belongs to the next chunk: never read here, its owner rewrites it)
            int bad = -1;
#pragma unroll
            for (int t = 0; t < L; ++t) {
                if (!(s > 0.0) && bad < 0) bad = t;
                const double r = rcp_fast(s);
                double Lk, rL;
                sqrt_rsqrt_fast(s, Lk, rL);
                sKd[k0 + t] = Lk;
                sKs[k0 + t] = div_r(ks[t], Lk, rL);
                if (t < L - 1) s = fma(-(ks[t] * ks[t]), r, kd[t]);
            }
            if (bad >= 0) report_not_pd(status, flat_node0 + k0 + bad);
            return;
        }
        double s = S.a[0];
        for (int t = 0; t < len; ++t) {
            if (!(s > 0.0)) report_not_pd(status, flat_node0 + k0 + t);
            const double ks = SCALED ? scaleS * sKs[k0 + t] : sKs[k0 + t];
            const double r = rcp_fast(s);
            double Lk, rL;
            sqrt_rsqrt_fast(s, Lk, rL);
            sKd[k0 + t] = Lk;
            sKs[k0 + t] = div_r(ks, Lk, rL);
            if (t < len - 1 || k0 + len == last_node)
                s = fma(-(ks * ks), r, SCALED ? scaleD * sKd[k0 + t + 1] : sKd[k0 + t + 1]);
        }
        if (k0 + len == last_node) {
            if (!(s > 0.0)) report_not_pd(status, flat_node0 + last_node);
            sKd[last_node] = sqrt_fast(s);
        }
        return;
    }
    bool ok = true;
    Mat<D> Lk = chol(S, ok);
    if (!ok) { report_not_pd(status, flat_node0 + k0); ok = true; }
    mat_store<D>(sKd + k0 * DD, Lk);
    for (int t = 0; t < len; ++t) {
        const Mat<D> Ksub = tile_block<D, SCALED>(sKs + (k0 + t) * DD, scaleS);
        const Mat<D> Lsub = transpose(trsm(Lk, transpose(Ksub)));
        mat_store<D>(sKs + (k0 + t) * DD, Lsub);
        if (t < len - 1 || k0 + t + 1 == last_node) {
            const Mat<D> Sn = tile_block<D, SCALED>(sKd + (k0 + t + 1) * DD, scaleD) - mmT(Lsub, Lsub);
            Lk = chol(Sn, ok);
            if (!ok) { report_not_pd(status, flat_node0 + k0 + t + 1); ok = true; }
            mat_store<D>(sKd + (k0 + t + 1) * DD, Lk);
        }
    }
}

// shared-memory layout of a sweep kernel with NSTAGE staged tiles: [mbarriers][stage 0 arrays][stage 1 arrays]...
template <int NSTAGE>
struct TileRing {
    uint64_t* bar;
    __device__ __forceinline__ void init(double* smem) {
        bar = reinterpret_cast<uint64_t*>(smem);
        if (threadIdx.x == 0) {
#pragma unroll
            for (int s = 0; s < NSTAGE; ++s) mbar_init(bar + s, 1);
            mbar_fence_init();
        }
        __syncthreads();
    }
    // all threads: tile `ti` (the ti-th this CTA staged) has landed in its stage
    __device__ __forceinline__ void wait(int ti) {
        mbar_wait(bar + ti % NSTAGE, (uint32_t)((ti / NSTAGE) & 1));
        if (threadIdx.x == 0) ldgsts_wait_all();   // odd tail elements of the tile (tile_fetch)
        __syncthreads();
    }
};
constexpr int kRingHeaderDoubles = 2;   // room for up to two mbarriers, keeps the stage buffers 16-byte aligned

// cholesky_band of K = (scaleD * Xd, scaleS * Xs)    block_tri_diag.py:428-440
// Tiles move by TMA bulk copies through a ring of NSTAGE staging buffers (tma_tile.cuh): the copies of tile t + 1
// are issued as soon as every thread has left tile t - 1 and land while tile t is scanned and replayed.
template <int D, int T, int L, int MODE, int NSTAGE = 2>
__global__ void __launch_bounds__(T)
k_chol_tiled(const double* __restrict__ Xd, const double* __restrict__ Xs, double scaleD, double scaleS,
             double* __restrict__ Ld, double* __restrict__ Ls, int64_t B, int64_t N, int nseg, int64_t seg_len,
             const Mat<D>* __restrict__ carry_in, Mat<D>* __restrict__ init_out,
             SchurElem<D>* __restrict__ seg_elem, int64_t* status, int first = 1) {
    constexpr int DD = D * D, TL = T * L;
    constexpr int SD = tile_slots((TL + 1) * DD), SS = tile_slots(TL * DD);
    extern __shared__ __align__(16) double smem[];
    double* stage0 = smem + kRingHeaderDoubles;
    SchurElem<D>* wtot = reinterpret_cast<SchurElem<D>*>(stage0 + NSTAGE * (SD + SS));
    TileRing<NSTAGE> ring;
    ring.init(smem);
    const int64_t b = blockIdx.x / nseg;
    const int seg = (int)(blockIdx.x % nseg);
    const int64_t t0 = seg * seg_len, t1 = min(N, t0 + seg_len);
    const double* xd = Xd + b * (N + 1) * DD;
    const double* xs = Xs + b * N * DD;
    const int ntiles = (int)((t1 - t0 + TL - 1) / TL);
    auto issue = [&](int ti) {   // thread 0: start the copies of tile ti
        const int64_t tk = t0 + (int64_t)ti * TL;
        const int cnt = (int)min((int64_t)TL, t1 - tk);
        double* base = stage0 + (ti % NSTAGE) * (SD + SS);
        uint64_t* bar = ring.bar + ti % NSTAGE;
        uint32_t bytes = tile_fetch(base, xd + tk * DD, (cnt + 1) * DD, bar);
        bytes += tile_fetch(base + SD, xs + tk * DD, cnt * DD, bar);
        mbar_arrive_expect_tx(bar, bytes);
    };
    if (threadIdx.x == 0 && ntiles > 0) issue(0);
    Mat<D> S_tile = mat_zero<D>();
    if (MODE == SWEEP_FULL && (seg > 0 || !first)) S_tile = carry_in[b * nseg + seg];
    SchurElem<D> seg_tot = SchurOps<D>::identity();
    for (int ti = 0; ti < ntiles; ++ti) {
        const int64_t tk = t0 + (int64_t)ti * TL;
        const int cnt = (int)min((int64_t)TL, t1 - tk);
        if (NSTAGE == 1 && ti > 0) {   // single buffer: refill once everybody has left the previous tile
            __syncthreads();
            if (threadIdx.x == 0) { bulk_wait_read(); issue(ti); }
        }
        ring.wait(ti);
        if (NSTAGE > 1 && threadIdx.x == 0 && ti + 1 < ntiles) {
            bulk_wait_read();          // the bulk stores of tile ti - 1 have read the buffer tile ti + 1 goes into
            issue(ti + 1);
        }
        double* base = stage0 + (ti % NSTAGE) * (SD + SS);
        double* sKd = base + tile_off(xd + tk * DD);
        double* sKs = base + SD + tile_off(xs + tk * DD);
        if (seg == 0 && tk == 0 && first) {
            S_tile = tile_block<D, true>(sKd, scaleD);
            if (MODE != SWEEP_FULL && threadIdx.x == 0) init_out[b] = S_tile;
        }
        const int k0 = threadIdx.x * L;
        const int len = min(L, cnt - k0);
        const SchurElem<D> e = chol_tile_summary<D, L, true>(sKd, sKs, k0, len, scaleD, scaleS);
        SchurElem<D> total;
        const SchurElem<D> excl = block_exclusive_scan<SchurOps<D>, T>(e, wtot, total);
        if (MODE != SWEEP_FULL) {
            seg_tot = SchurOps<D>::combine(seg_tot, total);
            continue;
        }
        const Mat<D> S = SchurOps<D>::apply(excl, S_tile);
        S_tile = SchurOps<D>::apply(total, S_tile);
        const int last_node = (tk + cnt == N) ? cnt : -1;
        chol_tile_recursion<D, L, true>(sKd, sKs, k0, len, S, last_node, b * (N + 1) + tk, status, scaleD, scaleS);
        fence_async_smem();
        __syncthreads();
        double* gd = Ld + (b * (N + 1) + tk) * DD;
        double* gs = Ls + (b * N + tk) * DD;
        const int nd = (cnt + (last_node >= 0 ? 1 : 0)) * DD, ns = cnt * DD;
        const bool bulk_d = tile_off(gd) == tile_off(xd + tk * DD), bulk_s = tile_off(gs) == tile_off(xs + tk * DD);
        if (threadIdx.x == 0) {
            if (bulk_d) tile_store_bulk(gd, sKd, nd);
            if (bulk_s) tile_store_bulk(gs, sKs, ns);
            bulk_commit();
        }
        if (!bulk_d) stage_out<T>(gd, sKd, nd);
        if (!bulk_s) stage_out<T>(gs, sKs, ns);
    }
    if (MODE != SWEEP_FULL && threadIdx.x == 0) seg_elem[b * nseg + seg] = seg_tot;
    if (threadIdx.x == 0) bulk_wait_all();
}

template <int D, int T, int L, int NSTAGE = 2>
constexpr size_t chol_tiled_smem() {
    return ((size_t)kRingHeaderDoubles + (size_t)NSTAGE * (tile_slots((T * L + 1) * D * D) + tile_slots(T * L * D * D))) *
               sizeof(double) +
           (T / 32) * sizeof(SchurElem<D>);
}


// ---------------------------------------------------------------------------------------------------
// StateSpaceModel.marginal_means (state_space_model.py:251) + _build_precision (:431-483) + cholesky_band:
// forward sweep giving m [B,N+1,D] and the Cholesky factor (Ld, Ls) of the rebuilt precision.
template <int D>
using MargOps = PairOps<SchurOps<D>, AffVecOps<D>>;

template <int D, int T, int L, int MODE>
__global__ void __launch_bounds__(T)
k_marg_tiled(const double* __restrict__ A, const double* __restrict__ bo, const double* __restrict__ cholQ,
             const double* __restrict__ mu0, const double* __restrict__ cholP0, double* __restrict__ m,
             double* __restrict__ Ld, double* __restrict__ Ls, int64_t B, int64_t N, int nseg, int64_t seg_len,
             const typename MargOps<D>::State* __restrict__ carry_in,
             typename MargOps<D>::State* __restrict__ init_out,
             typename MargOps<D>::Elem* __restrict__ seg_elem, int64_t* status, int first = 1, int last = 1) {
    constexpr int DD = D * D, TL = T * L;
    using Ops = MargOps<D>;
    extern __shared__ __align__(16) double smem[];
    double* sC = smem;                        // chol factor of node i           (TL + 2)
    double* sKd = sC;                         // ... overwritten by: precision diagonal -> L_kk  (TL + 1)
    double* sA = sC + (TL + 2) * DD;          // A of transition i               (TL + 1)
    double* sKs = sA + (TL + 1) * DD;         // precision sub-diagonal -> L_sub (TL)
    double* sb = sKs + TL * DD;               // offsets -> m_{k+1}              (TL)
    typename Ops::Elem* wtot = reinterpret_cast<typename Ops::Elem*>(sb + TL * D);
    const int64_t b = blockIdx.x / nseg;
    const int seg = (int)(blockIdx.x % nseg);
    const int64_t t0 = seg * seg_len, t1 = min(N, t0 + seg_len);
    typename Ops::State st;
    st.a = mat_zero<D>();
    st.b = vec_zero<D>();
    if (MODE == SWEEP_FULL && (seg > 0 || !first)) st = carry_in[b * nseg + seg];
    typename Ops::Elem seg_tot = Ops::identity();
    for (int64_t tk = t0; tk < t1; tk += TL) {
        const int cnt = (int)min((int64_t)TL, t1 - tk);
        // time-sharded ranks: the arrays carry one halo transition past N (from the next rank) and cholQ one
        // before index 0 (the factor of this rank's first node)
        const bool halo_has_next = (tk + cnt < N) || !last;
        __syncthreads();
        // node factors: node k -> cholQ[k-1], node 0 -> cholP0
        if (tk == 0 && first) {
            stage_in<T>(sC, cholP0 + b * DD, DD);
            stage_in<T>(sC + DD, cholQ + b * N * DD, (int64_t)(cnt + (halo_has_next ? 1 : 0)) * DD);
        } else {
            stage_in<T>(sC, cholQ + (b * N + tk - 1) * DD, (int64_t)(cnt + 1 + (halo_has_next ? 1 : 0)) * DD);
        }
        stage_in<T>(sA, A + (b * N + tk) * DD, (int64_t)(cnt + (halo_has_next ? 1 : 0)) * DD);
        stage_in<T>(sb, bo + (b * N + tk) * D, (int64_t)cnt * D);
        stage_wait();
        __syncthreads();
        // phase 0: precision blocks of nodes [0, cnt], written over the staged node factors (sKd is sC: one staged
        // array less, 6 instead of 4 CTAs per SM at D = 1).  Every thread walks its own chunk in order and holds the
        // factor of the node after its chunk -- the first one of the next thread's chunk -- in registers; thread 0
        // also does node cnt, whose factor nobody else overwrites.
        const int k0 = threadIdx.x * L;
        const int len = min(L, cnt - k0);
        {
            Mat<D> c_after = mat_eye<D>(), c_last = mat_eye<D>(), c_halo = mat_eye<D>(), A_halo = mat_zero<D>();
            if (len > 0) c_after = mat_load<D>(sC + (k0 + len) * DD);
            if (threadIdx.x == 0) {
                c_last = mat_load<D>(sC + cnt * DD);
                if (halo_has_next) { A_halo = mat_load<D>(sA + cnt * DD); c_halo = mat_load<D>(sC + (cnt + 1) * DD); }
            }
            __syncthreads();
            Mat<D> ci = mat_eye<D>();
            if (len > 0) ci = mat_load<D>(sC + k0 * DD);
            for (int t = 0; t < len; ++t) {
                const int i = k0 + t;
                const Mat<D> cnext = (t == len - 1) ? c_after : mat_load<D>(sC + (i + 1) * DD);
                Mat<D> diag, sub;
                precision_blocks(ci, true, mat_load<D>(sA + i * DD), cnext, diag, sub);
                mat_store<D>(sKd + i * DD, diag);
                mat_store<D>(sKs + i * DD, sub);
                ci = cnext;
            }
            if (threadIdx.x == 0) {
                Mat<D> diag, sub = mat_zero<D>();
                precision_blocks(c_last, halo_has_next, A_halo, c_halo, diag, sub);
                mat_store<D>(sKd + cnt * DD, diag);
            }
        }
        __syncthreads();
        if (seg == 0 && tk == 0) {
            if (first) {
                st.a = mat_load<D>(sKd);
                st.b = vec_load<D>(mu0 + b * D);
                if (MODE != SWEEP_FULL && threadIdx.x == 0) init_out[b] = st;
            }
            if (MODE == SWEEP_FULL && threadIdx.x == 0) vec_store<D>(m + b * (N + 1) * D, st.b);
        }
        typename Ops::Elem e;
        e.a = chol_tile_summary<D, L>(sKd, sKs, k0, len);
        e.b = AffVecOps<D>::identity();
        for (int t = 0; t < len; ++t) {
            AffVecElem<D> s;
            s.M = mat_load<D>(sA + (k0 + t) * DD);
            s.v = vec_load<D>(sb + (k0 + t) * D);
            e.b = AffVecOps<D>::combine(e.b, s);
        }
        typename Ops::Elem total;
        const typename Ops::Elem excl = block_exclusive_scan<Ops, T>(e, wtot, total);
        if (MODE != SWEEP_FULL) {
            seg_tot = Ops::combine(seg_tot, total);
            continue;
        }
        const typename Ops::State mine = Ops::apply(excl, st);
        st = Ops::apply(total, st);
        const int last_node = (tk + cnt == N) ? cnt : -1;
        chol_tile_recursion<D, L>(sKd, sKs, k0, len, mine.a, last_node, b * (N + 1) + tk, status);
        Vec<D> mk = mine.b;
        for (int t = 0; t < len; ++t) {
            mk = vec_load<D>(sb + (k0 + t) * D) + mv(mat_load<D>(sA + (k0 + t) * DD), mk);
            vec_store<D>(sb + (k0 + t) * D, mk);
        }
        __syncthreads();
        stage_out<T>(Ld + (b * (N + 1) + tk) * DD, sKd, (int64_t)(cnt + (last_node >= 0 ? 1 : 0)) * DD);
        stage_out<T>(Ls + (b * N + tk) * DD, sKs, (int64_t)cnt * DD);
        stage_out<T>(m + (b * (N + 1) + tk + 1) * D, sb, (int64_t)cnt * D);
    }
    if (MODE != SWEEP_FULL && threadIdx.x == 0) seg_elem[b * nseg + seg] = seg_tot;
}

template <int D, int T, int L>
constexpr size_t marg_tiled_smem() {
    return ((size_t)(T * L + 2) * D * D + (size_t)(T * L + 1) * D * D + (size_t)T * L * D * D + (size_t)T * L * D) *
               sizeof(double) +
           (T / 32) * sizeof(typename MargOps<D>::Elem);
}


// ---------------------------------------------------------------------------------------------------
// Marginal means alone, m_{k+1} = A_k m_k + b_k (state_space_model.py:251): the forward sweep of the posterior
// rebuild when the Cholesky factor of the precision is already at hand (the one naturals_to_ssm was computed from is
// the factor of the same matrix the reference rebuilds from the state-space model).
// DRIFT >= 0 (D = 1, polynomial drift, FULL mode): the sweep also has everything the KL term of classic_elbo
// (variational_cvi_sde.py:446-486) needs once S_k and chol Q_k are staged beside A_k and b_k, so it adds the terms up
// on its way -- per thread over its chunk, then a fixed tree over the CTA -- into partials[b, seg, tile]; the
// separate pass over A, b, m again (k_kl_value) and the per-grid-point terms array are not needed then.
struct MeanKlArgs {
    DriftParams P;
    double Qp = 0.0, dt = 0.0;              // q dt, dt
    const double *Sd = nullptr, *cholQ = nullptr, *qcholP0 = nullptr, *pmu0 = nullptr, *pcholP0 = nullptr;
    double* partials = nullptr;             // [B, nseg, tiles_per_seg], zeroed by the caller
    int64_t* status = nullptr;
    int tiles_per_seg = 0;
};

template <int D, int T, int L, int MODE, int DRIFT = -1, int NSTAGE = 2>
__global__ void __launch_bounds__(T)
k_mean_tiled(const double* __restrict__ A, const double* __restrict__ bo, const double* __restrict__ mu0,
             double* __restrict__ m, int64_t B, int64_t N, int nseg, int64_t seg_len,
             const Vec<D>* __restrict__ carry_in, Vec<D>* __restrict__ init_out,
             AffVecElem<D>* __restrict__ seg_elem, MeanKlArgs kl = MeanKlArgs(), int first = 1) {
    constexpr int DD = D * D, TL = T * L;
    constexpr bool KL = (DRIFT >= 0) && (MODE == SWEEP_FULL);
    static_assert(DRIFT < 0 || D == 1, "the fused KL terms are scalar");
    using Ops = AffVecOps<D>;
    constexpr int SM = tile_slots(TL * DD), SV = tile_slots(TL * D), S1 = tile_slots(TL);
    constexpr int kStage = SM + SV + (KL ? 2 * S1 : 0);                   // A, b, [S, chol Q]
    extern __shared__ __align__(16) double smem[];
    double* stage0 = smem + kRingHeaderDoubles;
    typename Ops::Elem* wtot = reinterpret_cast<typename Ops::Elem*>(stage0 + NSTAGE * kStage);
    TileRing<NSTAGE> ring;
    ring.init(smem);
    const int64_t b = blockIdx.x / nseg;
    const int seg = (int)(blockIdx.x % nseg);
    const int64_t t0 = seg * seg_len, t1 = min(N, t0 + seg_len);
    const int ntiles = (int)((t1 - t0 + TL - 1) / TL);
    auto issue = [&](int ti) {   // thread 0: start the copies of tile ti
        const int64_t tk = t0 + (int64_t)ti * TL;
        const int cnt = (int)min((int64_t)TL, t1 - tk);
        double* base = stage0 + (ti % NSTAGE) * kStage;
        uint64_t* bar = ring.bar + ti % NSTAGE;
        uint32_t bytes = tile_fetch(base, A + (b * N + tk) * DD, cnt * DD, bar);
        bytes += tile_fetch(base + SM, bo + (b * N + tk) * D, cnt * D, bar);
        if (KL) {
            bytes += tile_fetch(base + SM + SV, kl.Sd + (b * (N + 1) + tk), cnt, bar);
            bytes += tile_fetch(base + SM + SV + S1, kl.cholQ + (b * N + tk), cnt, bar);
        }
        mbar_arrive_expect_tx(bar, bytes);
    };
    if (threadIdx.x == 0 && ntiles > 0) issue(0);
    // first = 0 (time-sharded ranks after the first): the chain starts from the carry of the ranks before it
    Vec<D> st = vec_zero<D>();
    if (MODE == SWEEP_FULL && (seg > 0 || !first)) st = carry_in[b * nseg + seg];
    typename Ops::Elem seg_tot = Ops::identity();
    for (int ti = 0; ti < ntiles; ++ti) {
        const int64_t tk = t0 + (int64_t)ti * TL;
        const int cnt = (int)min((int64_t)TL, t1 - tk);
        if (NSTAGE == 1 && ti > 0) {
            __syncthreads();
            if (threadIdx.x == 0) { bulk_wait_read(); issue(ti); }
        }
        ring.wait(ti);
        if (VDP_PREFETCH_M && NSTAGE == 1 && threadIdx.x == 0 && ti + 1 < ntiles) {
            const int64_t tn = tk + TL;
            const int nc = (int)min((int64_t)TL, t1 - tn);
            bulk_prefetch_l2(A + (b * N + tn) * DD, nc * DD);
            bulk_prefetch_l2(bo + (b * N + tn) * D, nc * D);
            if (KL) {
                bulk_prefetch_l2(kl.Sd + (b * (N + 1) + tn), nc);
                bulk_prefetch_l2(kl.cholQ + (b * N + tn), nc);
            }
        }
        double* base = stage0 + (ti % NSTAGE) * kStage;
        double* sA = base + tile_off(A + (b * N + tk) * DD);                  // A of transition i
        double* sb = base + SM + tile_off(bo + (b * N + tk) * D);             // offsets -> m_{k+1}
        double* sS = base + SM + SV + (KL ? tile_off(kl.Sd + (b * (N + 1) + tk)) : 0);       // S of node i (KL only)
        double* sQ = base + SM + SV + S1 + (KL ? tile_off(kl.cholQ + (b * N + tk)) : 0);     // chol Q of transition i
        if (seg == 0 && tk == 0) {
            if (first) {
                st = vec_load<D>(mu0 + b * D);
                if (MODE != SWEEP_FULL && threadIdx.x == 0) init_out[b] = st;
            }
            if (MODE == SWEEP_FULL && threadIdx.x == 0) vec_store<D>(m + b * (N + 1) * D, st);
        }
        const int k0 = threadIdx.x * L;
        const int len = min(L, cnt - k0);
        typename Ops::Elem e = Ops::identity();
        for (int t = 0; t < len; ++t) {
            AffVecElem<D> s;
            s.M = mat_load<D>(sA + (k0 + t) * DD);
            s.v = vec_load<D>(sb + (k0 + t) * D);
            e = Ops::combine(e, s);
        }
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
        Vec<D> mk = Ops::apply(excl, st);
        st = Ops::apply(total, st);
        double kl_acc = 0.0, kl_ratio = 1.0;   // (L <= 9 ratios of O(1) each: the product stays far from the range ends)
        for (int t = 0; t < len; ++t) {
            const Vec<D> bk = vec_load<D>(sb + (k0 + t) * D);
            const Mat<D> Ak = mat_load<D>(sA + (k0 + t) * DD);
            if constexpr (KL) {
                bool ok = true;
                double r;
                kl_acc += kl_value_scalar_closed<(DRIFT < 0 ? 0 : DRIFT)>(kl.P, kl.Qp, kl.dt, Ak.a[0], bk.v[0],
                                                                          sQ[k0 + t], mk.v[0], sS[k0 + t], ok, r);
                kl_ratio *= r;
                if (!ok) { report_not_pd(kl.status, b * (N + 1) + tk + k0 + t); kl_ratio = 1.0; }
            }
            mk = bk + mv(Ak, mk);
            vec_store<D>(sb + (k0 + t) * D, mk);
        }
        if (KL && len > 0) kl_acc -= log(kl_ratio);      // sum of 0.5 (-2 log ratio_t)
        fence_async_smem();
        __syncthreads();
        tile_store<T>(m + (b * (N + 1) + tk + 1) * D, sb, cnt * D);
        if (threadIdx.x == 0) bulk_commit();
        if constexpr (KL) {
            if (seg == 0 && tk == 0 && threadIdx.x == 0) {
                bool ok = true;
                kl_acc += kl_initial_state<D>(mu0, kl.qcholP0, kl.pmu0, kl.pcholP0, b, ok);
                if (!ok) report_not_pd(kl.status, b * (N + 1));
            }
            sS[threadIdx.x] = kl_acc;          // (the staged S values have been consumed: barrier above)
            __syncthreads();
            for (int s2 = T / 2; s2 > 0; s2 >>= 1) {
                if ((int)threadIdx.x < s2) sS[threadIdx.x] += sS[threadIdx.x + s2];
                __syncthreads();
            }
            if (threadIdx.x == 0)
                kl.partials[((int64_t)b * nseg + seg) * kl.tiles_per_seg + (tk - t0) / TL] = sS[0];
        }
    }
    if (threadIdx.x == 0) {
        if (MODE != SWEEP_FULL) seg_elem[b * nseg + seg] = seg_tot;
        bulk_wait_all();
    }
}

template <int D, int T, int L, bool KL = false, int NSTAGE = 2>
constexpr size_t mean_tiled_smem() {
    return ((size_t)kRingHeaderDoubles +
            (size_t)NSTAGE * (tile_slots(T * L * D * D) + tile_slots(T * L * D) + (KL ? 2 * tile_slots(T * L) : 0))) *
               sizeof(double) +
           (T / 32) * sizeof(AffVecElem<D>);
}


// ---------------------------------------------------------------------------------------------------
// Backward sweeps.  A segment's tiles are walked from its top; thread c owns the c-th chunk from the top
// of the tile, local transitions [bot, top), processed downwards; its carry is the state at node `top`.
struct BwdChunk { int top, bot, len; };
template <int L>
__device__ __forceinline__ BwdChunk bwd_chunk(int cnt) {
    BwdChunk c;
    c.top = cnt - (int)threadIdx.x * L;
    c.bot = max(0, c.top - L);
    c.len = c.top - c.bot;
    return c;
}

// element of the inverse-band recursion over a chunk (top-down)
template <int D>
__device__ __forceinline__ AffSymElem<D> invband_tile_summary(const double* sLd, const double* sLs, const BwdChunk& ch) {
    constexpr int DD = D * D;
    AffSymElem<D> e = AffSymOps<D>::identity();
    for (int k = ch.top - 1; k >= ch.bot; --k) {
        const Mat<D> U = trsm(mat_load<D>(sLd + k * DD), mat_eye<D>());
        AffSymElem<D> s;
        s.G = mm(mat_load<D>(sLs + k * DD), U);
        s.C = Tmm(U, U);
        e = AffSymOps<D>::combine(e, s);
    }
    return e;
}

// inverse_from_cholesky_band (block_tri_diag.py:318-337): Sd [B,n,D,D] and optionally Ss [B,n-1,D,D]
template <int D, int T, int L, int MODE>
__global__ void __launch_bounds__(T)
k_invband_tiled(const double* __restrict__ Ld, const double* __restrict__ Ls, double* __restrict__ Sd,
                double* __restrict__ Ss, int64_t B, int64_t N, int nseg, int64_t seg_len,
                const Mat<D>* __restrict__ carry_in, Mat<D>* __restrict__ init_out,
                AffSymElem<D>* __restrict__ seg_elem, int last = 1) {
    constexpr int DD = D * D, TL = T * L;
    using Ops = AffSymOps<D>;
    extern __shared__ __align__(16) double smem[];
    double* sLd = smem;             // L_kk -> Sigma_kk       (TL)
    double* sLs = sLd + TL * DD;    // L_sub -> Sigma_{k+1,k} (TL)
    AffSymElem<D>* wtot = reinterpret_cast<AffSymElem<D>*>(sLs + TL * DD);
    const int64_t b = blockIdx.x / nseg;
    const int seg = (int)(blockIdx.x % nseg);
    const int64_t t0 = seg * seg_len, t1 = min(N, t0 + seg_len);
    Mat<D> S_tile;
    if (t1 == N && last) {
        S_tile = invband_last(mat_load<D>(Ld + (b * (N + 1) + N) * DD));
        if (threadIdx.x == 0) {
            if (MODE == SWEEP_FULL) mat_store<D>(Sd + (b * (N + 1) + N) * DD, S_tile);
            else init_out[b] = S_tile;
        }
    } else {
        S_tile = (MODE == SWEEP_FULL) ? carry_in[b * nseg + seg] : mat_zero<D>();
    }
    AffSymElem<D> seg_tot = Ops::identity();
    for (int64_t hi = t1; hi > t0; hi -= TL) {
        const int64_t lo = max(t0, hi - TL);
        const int cnt = (int)(hi - lo);
        __syncthreads();
        stage_in<T>(sLd, Ld + (b * (N + 1) + lo) * DD, (int64_t)cnt * DD);
        stage_in<T>(sLs, Ls + (b * N + lo) * DD, (int64_t)cnt * DD);
        stage_wait();
        __syncthreads();
        if (lo > t0) {
            const int64_t nlo = max(t0, lo - TL);
            prefetch_range<T>(Ld + (b * (N + 1) + nlo) * DD, (lo - nlo) * DD);
            prefetch_range<T>(Ls + (b * N + nlo) * DD, (lo - nlo) * DD);
        }
        const BwdChunk ch = bwd_chunk<L>(cnt);
        const AffSymElem<D> e = invband_tile_summary<D>(sLd, sLs, ch);
        AffSymElem<D> total;
        const AffSymElem<D> excl = block_exclusive_scan<Ops, T>(e, wtot, total);
        if (MODE != SWEEP_FULL) {
            seg_tot = Ops::combine(seg_tot, total);
            continue;
        }
        Mat<D> S = Ops::apply(excl, S_tile);
        S_tile = Ops::apply(total, S_tile);
        for (int k = ch.top - 1; k >= ch.bot; --k) {
            Mat<D> Ssub, Skk;
            invband_step(mat_load<D>(sLd + k * DD), mat_load<D>(sLs + k * DD), S, Ssub, Skk);
            S = Skk;
            mat_store<D>(sLd + k * DD, Skk);
            mat_store<D>(sLs + k * DD, Ssub);
        }
        __syncthreads();
        stage_out<T>(Sd + (b * (N + 1) + lo) * DD, sLd, (int64_t)cnt * DD);
        if (Ss) stage_out<T>(Ss + (b * N + lo) * DD, sLs, (int64_t)cnt * DD);
    }
    if (MODE != SWEEP_FULL && threadIdx.x == 0) seg_elem[b * nseg + seg] = seg_tot;
}

template <int D, int T, int L>
constexpr size_t invband_tiled_smem() {
    return 2 * (size_t)T * L * D * D * sizeof(double) + (T / 32) * sizeof(AffSymElem<D>);
}


// ---------------------------------------------------------------------------------------------------
// naturals_to_ssm_params, backward half (ssm_gaussian_transformations.py:443-504): inverse band ->
// A_k -> conditional precision -> covariance and its Cholesky factor -> (A^-T)-solve of theta1 -> offsets.
// Two dependent recursions: Sigma (AffSym) and, through A_k(Sigma), the vector y (AffVec); the tile runs
// scan / recursion for Sigma first, composes the y elements on the way, then scans and runs y.
// Multi-segment launch order: SUMMARY (Sigma elements) -> k_seg_scan -> SUMMARY2 (y elements, needs the
// Sigma carries) -> k_seg_scan -> FULL.
// FROML: the precision blocks the conditional precisions need, K_kk = -2 theta_d and K_{k+1,k} = -theta_s, are
// rebuilt from the Cholesky factor already staged (K_kk = L_kk L_kk^T + L_{k,k-1} L_{k,k-1}^T,
// K_{k+1,k} = L_{k+1,k} L_kk^T) instead of being read again: 16 of 64 B per grid step less traffic (D = 1) and
// 3 instead of 5 staged arrays (twice the CTAs per SM).  Single-rank launches only (needs L_{k,k-1} below the range).
// COV: the diagonal blocks of the inverse band the Sigma recursion walks through are the marginal covariances of the
// Gaussian the naturals describe; they are staged out as well (Sd [B, N+1, D, D]), which makes a separate
// inverse-band sweep over the same factor unnecessary in the posterior rebuild.
template <int D, int T, int L, int MODE, bool FROML = false, bool COV = false, int NSTAGE = 2>
__global__ void __launch_bounds__(T)
k_nat2ssm_tiled(const double* __restrict__ Ld, const double* __restrict__ Ls, const double* __restrict__ th1,
                const double* __restrict__ thd, const double* __restrict__ ths, double* __restrict__ A,
                double* __restrict__ bo, double* __restrict__ cholQ, double* __restrict__ mu0,
                double* __restrict__ cholP0, int64_t B, int64_t N, int nseg, int64_t seg_len,
                const Mat<D>* __restrict__ carryS, const Vec<D>* __restrict__ carryY, Mat<D>* __restrict__ initS,
                Vec<D>* __restrict__ initY, AffSymElem<D>* __restrict__ seg_elemS,
                AffVecElem<D>* __restrict__ seg_elemY, int64_t* status, int first = 1, int last = 1,
                double* __restrict__ Sd = nullptr) {
    constexpr int DD = D * D, TL = T * L;
    constexpr int SM = tile_slots(TL * DD), SV = tile_slots(TL * D);      // slots of a staged matrix / vector array
    constexpr int kStage = (FROML ? 2 : 4) * SM + SV;                     // Ld, Ls, [thd, ths,] th1
    extern __shared__ __align__(16) double smem[];
    double* stage0 = smem + kRingHeaderDoubles;
    double* rawSd = stage0 + NSTAGE * kStage;                              // marginal covariances (COV; output only)
    AffSymElem<D>* wtotS = reinterpret_cast<AffSymElem<D>*>(rawSd + (COV ? SM : 0));
    AffVecElem<D>* wtotY = reinterpret_cast<AffVecElem<D>*>(wtotS);   // used after the Sigma scan is over
    TileRing<NSTAGE> ring;
    ring.init(smem);
    const int64_t b = blockIdx.x / nseg;
    const int seg = (int)(blockIdx.x % nseg);
    const int64_t t0 = seg * seg_len, t1 = min(N, t0 + seg_len);
    const int64_t n = N + 1;
    const int ntiles = (int)((t1 - t0 + TL - 1) / TL);
    auto issue = [&](int ti) {   // thread 0: start the copies of the ti-th tile from the top
        const int64_t hi = t1 - (int64_t)ti * TL, lo = max(t0, hi - TL);
        const int cnt = (int)(hi - lo);
        double* base = stage0 + (ti % NSTAGE) * kStage;
        uint64_t* bar = ring.bar + ti % NSTAGE;
        uint32_t bytes = tile_fetch(base, Ld + (b * n + lo) * DD, cnt * DD, bar);
        bytes += tile_fetch(base + SM, Ls + (b * N + lo) * DD, cnt * DD, bar);
        if (MODE != SWEEP_SUMMARY) {
            if (!FROML) {
                bytes += tile_fetch(base + 2 * SM, thd + (b * n + lo) * DD, cnt * DD, bar);
                bytes += tile_fetch(base + 3 * SM, ths + (b * N + lo) * DD, cnt * DD, bar);
            }
            bytes += tile_fetch(base + (FROML ? 2 : 4) * SM, th1 + (b * n + lo) * D, cnt * D, bar);
        }
        mbar_arrive_expect_tx(bar, bytes);
    };
    if (threadIdx.x == 0 && ntiles > 0) issue(0);
    Mat<D> S_tile = mat_zero<D>();
    Vec<D> y_tile = vec_zero<D>();
    if (t1 == N && last) {
        S_tile = invband_last(mat_load<D>(Ld + (b * n + N) * DD));
        y_tile = vec_load<D>(th1 + (b * n + N) * D);
        if (threadIdx.x == 0) {
            if (MODE == SWEEP_SUMMARY) initS[b] = S_tile;
            if (MODE == SWEEP_SUMMARY2) initY[b] = y_tile;
            if (MODE == SWEEP_FULL) {   // outputs of the last node
                bool ok = true;
                Mat<D> cov, cc, KN;
                if (FROML) {
                    const Mat<D> Ln = mat_load<D>(Ld + (b * n + N) * DD), Lm = mat_load<D>(Ls + (b * N + N - 1) * DD);
                    KN = mmT(Ln, Ln) + mmT(Lm, Lm);
                } else {
                    KN = -2.0 * mat_load<D>(thd + (b * n + N) * DD);
                }
                cov_from_precision(mirror_lower(KN), cov, cc, ok);
                if (!ok) report_not_pd(status, b * n + N);
                vec_store<D>(bo + (b * N + N - 1) * D, mv(cov, y_tile));
                if (cholQ) mat_store<D>(cholQ + (b * N + N - 1) * DD, cc);
                if (COV) mat_store<D>(Sd + (b * n + N) * DD, S_tile);
            }
        }
    } else {
        if (MODE != SWEEP_SUMMARY) S_tile = carryS[b * nseg + seg];
        if (MODE == SWEEP_FULL) y_tile = carryY[b * nseg + seg];
    }
    AffSymElem<D> seg_totS = AffSymOps<D>::identity();
    AffVecElem<D> seg_totY = AffVecOps<D>::identity();
    for (int ti = 0; ti < ntiles; ++ti) {
        const int64_t hi = t1 - (int64_t)ti * TL, lo = max(t0, hi - TL);
        const int cnt = (int)(hi - lo);
        if (NSTAGE == 1 && ti > 0) {
            __syncthreads();
            if (threadIdx.x == 0) { bulk_wait_read(); issue(ti); }
        }
        ring.wait(ti);
        if (VDP_PREFETCH_N && NSTAGE == 1 && threadIdx.x == 0 && ti + 1 < ntiles) {
            // single buffer: the next tile's lines travel to L2 while this one is worked on
            const int64_t nhi = lo, nlo = max(t0, nhi - TL);
            const int nc = (int)(nhi - nlo);
            bulk_prefetch_l2(Ld + (b * n + nlo) * DD, nc * DD);
            bulk_prefetch_l2(Ls + (b * N + nlo) * DD, nc * DD);
            if (MODE != SWEEP_SUMMARY) {
                if (!FROML) {
                    bulk_prefetch_l2(thd + (b * n + nlo) * DD, nc * DD);
                    bulk_prefetch_l2(ths + (b * N + nlo) * DD, nc * DD);
                }
                bulk_prefetch_l2(th1 + (b * n + nlo) * D, nc * D);
            }
        }
        double* base = stage0 + (ti % NSTAGE) * kStage;
        double* sLd = base + tile_off(Ld + (b * n + lo) * DD);                 // L_kk  -> chol of the conditional covariance
        double* sLs = base + SM + tile_off(Ls + (b * N + lo) * DD);            // L_sub -> A_k
        double* sTd = base + 2 * SM + tile_off(thd + (b * n + lo) * DD);       // theta_diag -> conditional covariance (not FROML)
        double* sTs = base + 3 * SM + tile_off(ths + (b * N + lo) * DD);       // theta_sub                         (not FROML)
        double* sT1 = base + (FROML ? 2 : 4) * SM + tile_off(th1 + (b * n + lo) * D);   // theta_1 -> offset of node k
        double* sSd = rawSd + (COV ? tile_off(Sd + (b * n + lo) * DD) : 0);    // placed with the parity of its destination
        const BwdChunk ch = bwd_chunk<L>(cnt);
        // FROML: L_{bot,bot-1}, the block below this thread's chunk (the neighbour overwrites it in its recursion)
        Mat<D> Lbelow = mat_zero<D>();
        if (FROML && MODE != SWEEP_SUMMARY && ch.len > 0) {
            if (ch.bot >= 1) Lbelow = mat_load<D>(sLs + (ch.bot - 1) * DD);
            else if (lo >= 1) Lbelow = mat_load<D>(Ls + (b * N + lo - 1) * DD);
        }
        const AffSymElem<D> eS = invband_tile_summary<D>(sLd, sLs, ch);
        if (NSTAGE > 1 && threadIdx.x == 0 && ti + 1 < ntiles) {
            // (here rather than at the top of the tile: the bulk stores of the previous tile, whose buffer the next
            // tile goes into, have had this thread's summary phase to read their source)
            bulk_wait_read();
            issue(ti + 1);
        } else if (COV && threadIdx.x == 0) {
            bulk_wait_read();   // sSd is one buffer: its previous store must have read it before the recursion below
        }
        AffSymElem<D> totalS;
        const AffSymElem<D> exclS = block_exclusive_scan<AffSymOps<D>, T>(eS, wtotS, totalS);
        if (MODE == SWEEP_SUMMARY) {
            seg_totS = AffSymOps<D>::combine(seg_totS, totalS);
            continue;
        }
        Mat<D> S = AffSymOps<D>::apply(exclS, S_tile);
        S_tile = AffSymOps<D>::apply(totalS, S_tile);
        AffVecElem<D> eY = AffVecOps<D>::identity();
        bool ok = true;
        for (int k = ch.top - 1; k >= ch.bot; --k) {
            Mat<D> Ssub, Skk;
            const Mat<D> Lkk = mat_load<D>(sLd + k * DD), Lsub = mat_load<D>(sLs + k * DD);
            invband_step(Lkk, Lsub, S, Ssub, Skk);
            S = Skk;
            if (COV && MODE == SWEEP_FULL) mat_store<D>(sSd + k * DD, Skk);
            const Mat<D> Ak = transition_from_cov(Skk, Ssub);
            Mat<D> prec;
            if (FROML) {
                const Mat<D> Lkm1 = (k > ch.bot) ? mat_load<D>(sLs + (k - 1) * DD) : Lbelow;
                prec = (mmT(Lkk, Lkk) + mmT(Lkm1, Lkm1)) + Tmm(Ak, mmT(Lsub, Lkk));
            } else {
                prec = -2.0 * mat_load<D>(sTd + k * DD);
                prec = prec + Tmm(Ak, -1.0 * mat_load<D>(sTs + k * DD));
            }
            mat_store<D>(sLs + k * DD, Ak);
            prec = mirror_lower(prec);
            Mat<D> cov, cc;
            cov_from_precision(prec, cov, cc, ok);
            if (!ok) { report_not_pd(status, b * n + lo + k); ok = true; }
            if (!FROML) mat_store<D>(sTd + k * DD, cov);
            mat_store<D>(sLd + k * DD, cc);
            AffVecElem<D> s;
            s.M = transpose(Ak);
            s.v = vec_load<D>(sT1 + k * D);
            eY = AffVecOps<D>::combine(eY, s);
        }
        AffVecElem<D> totalY;
        const AffVecElem<D> exclY = block_exclusive_scan<AffVecOps<D>, T>(eY, wtotY, totalY);
        if (MODE == SWEEP_SUMMARY2) {
            seg_totY = AffVecOps<D>::combine(seg_totY, totalY);
            continue;
        }
        Vec<D> y = AffVecOps<D>::apply(exclY, y_tile);
        y_tile = AffVecOps<D>::apply(totalY, y_tile);
        for (int k = ch.top - 1; k >= ch.bot; --k) {
            y = vec_load<D>(sT1 + k * D) + Tmv(mat_load<D>(sLs + k * DD), y);
            Mat<D> cov;
            if (FROML) { const Mat<D> cc = mat_load<D>(sLd + k * DD); cov = mmT(cc, cc); }
            else cov = mat_load<D>(sTd + k * DD);
            vec_store<D>(sT1 + k * D, mv(cov, y));
        }
        fence_async_smem();
        __syncthreads();
        // (cholQ == NULL: the caller has no use for the process-noise factors, e.g. the rebuild between the data-site
        // and the girsanov-site update, whose consumer reads A, the means and the marginal covariances only)
        tile_store<T>(A + (b * N + lo) * DD, sLs, cnt * DD);
        if (COV) tile_store<T>(Sd + (b * n + lo) * DD, sSd, cnt * DD);
        if (lo >= 1 || !first) {   // time-sharded ranks own one slot before index 0 of bo / cholQ
            tile_store<T>(bo + (b * N + lo - 1) * D, sT1, cnt * D);
            if (cholQ) tile_store<T>(cholQ + (b * N + lo - 1) * DD, sLd, cnt * DD);
        } else {
            stage_out<T>(mu0 + b * D, sT1, D);
            stage_out<T>(cholP0 + b * DD, sLd, DD);
            tile_store<T>(bo + b * N * D, sT1 + D, (cnt - 1) * D);
            if (cholQ) tile_store<T>(cholQ + b * N * DD, sLd + DD, (cnt - 1) * DD);
        }
        if (threadIdx.x == 0) bulk_commit();
    }
    if (threadIdx.x == 0) {
        if (MODE == SWEEP_SUMMARY) seg_elemS[b * nseg + seg] = seg_totS;
        if (MODE == SWEEP_SUMMARY2) seg_elemY[b * nseg + seg] = seg_totY;
        bulk_wait_all();
    }
}

template <int D, int T, int L, bool FROML = false, bool COV = false, int NSTAGE = 2>
constexpr size_t nat2ssm_tiled_smem() {
    return ((size_t)kRingHeaderDoubles +
            (size_t)NSTAGE * ((FROML ? 2 : 4) * tile_slots(T * L * D * D) + tile_slots(T * L * D)) +
            (COV ? tile_slots(T * L * D * D) : 0)) * sizeof(double) +
           (T / 32) * (sizeof(AffSymElem<D>) > sizeof(AffVecElem<D>) ? sizeof(AffSymElem<D>) : sizeof(AffVecElem<D>));
}

// ---------------------------------------------------------------------------------------------------
// naturals_to_ssm_params (ssm_gaussian_transformations.py:333-511) as ONE backward sweep in information form
// (InfoOps, scan_elems.cuh): no Cholesky sweep from the top, no inverse band, one recursion instead of two dependent
// ones.  Transition k is produced from the state (J, h) at node k+1:
//     A_k = J^-1 theta_s[k],  b_k = J^-1 h,  chol Q_k = chol(J^-1);   node 0 of the chain: mu0 = J_0^-1 h_0, chol P0.
// Same values as the reference route in exact arithmetic (A from the conditional precision instead of from the
// covariances of the inverse band); a route of its own, held to the oracle like the others.
// in: th1 [B,N+1,D], thd [B,N+1,D,D], ths [B,N,D,D];  out: A, bo, cholQ [B,N,...] (all at the index of their transition),
// mu0, cholP0.  Time-sharded: `last` = the stretch ends with the chain's final node (its naturals are the start state);
// otherwise the top segment starts from carry_in; `first` = node 0 of the stretch is node 0 of the chain.
template <int D, int T, int L, int MODE, int NSTAGE = 1>
__global__ void __launch_bounds__(T)
k_info_tiled(const double* __restrict__ th1, const double* __restrict__ thd, const double* __restrict__ ths,
             double* __restrict__ A, double* __restrict__ bo, double* __restrict__ cholQ, double* __restrict__ mu0,
             double* __restrict__ cholP0, int64_t B, int64_t N, int nseg, int64_t seg_len,
             const InfoState<D>* __restrict__ carry_in, InfoState<D>* __restrict__ init_out,
             InfoElem<D>* __restrict__ seg_elem, int64_t* status, int first = 1, int last = 1) {
    constexpr int DD = D * D, TL = T * L;
    using Ops = InfoOps<D>;
    constexpr int SM = tile_slots(TL * DD), SV = tile_slots(TL * D);
    constexpr int kStage = 2 * SM + SV;                                    // thd, ths, th1
    extern __shared__ __align__(16) double smem[];
    double* stage0 = smem + kRingHeaderDoubles;
    InfoElem<D>* wtot = reinterpret_cast<InfoElem<D>*>(stage0 + NSTAGE * kStage);
    TileRing<NSTAGE> ring;
    ring.init(smem);
    const int64_t b = blockIdx.x / nseg;
    const int seg = (int)(blockIdx.x % nseg);
    const int64_t t0 = seg * seg_len, t1 = min(N, t0 + seg_len);
    const int64_t n = N + 1;
    const int ntiles = (int)((t1 - t0 + TL - 1) / TL);
    auto issue = [&](int ti) {   // thread 0: start the copies of the ti-th tile from the top
        const int64_t hi = t1 - (int64_t)ti * TL, lo = max(t0, hi - TL);
        const int cnt = (int)(hi - lo);
        double* base = stage0 + (ti % NSTAGE) * kStage;
        uint64_t* bar = ring.bar + ti % NSTAGE;
        uint32_t bytes = tile_fetch(base, thd + (b * n + lo) * DD, cnt * DD, bar);
        bytes += tile_fetch(base + SM, ths + (b * N + lo) * DD, cnt * DD, bar);
        bytes += tile_fetch(base + 2 * SM, th1 + (b * n + lo) * D, cnt * D, bar);
        mbar_arrive_expect_tx(bar, bytes);
    };
    if (threadIdx.x == 0 && ntiles > 0) issue(0);
    InfoState<D> st;
    st.J = mat_eye<D>();
    st.h = vec_zero<D>();
    if (t1 == N && last) {
        st.J = mirror_lower(-2.0 * mat_load<D>(thd + (b * n + N) * DD));
        st.h = vec_load<D>(th1 + (b * n + N) * D);
        if (MODE != SWEEP_FULL && threadIdx.x == 0) init_out[b] = st;
    } else if (MODE == SWEEP_FULL) {
        st = carry_in[b * nseg + seg];
    }
    InfoElem<D> seg_tot = Ops::identity();
    for (int ti = 0; ti < ntiles; ++ti) {
        const int64_t hi = t1 - (int64_t)ti * TL, lo = max(t0, hi - TL);
        const int cnt = (int)(hi - lo);
        if (NSTAGE == 1 && ti > 0) {
            __syncthreads();
            if (threadIdx.x == 0) { bulk_wait_read(); issue(ti); }
        }
        ring.wait(ti);
        double* base = stage0 + (ti % NSTAGE) * kStage;
        double* sTd = base + tile_off(thd + (b * n + lo) * DD);               // theta_d[k] -> chol Q_k
        double* sTs = base + SM + tile_off(ths + (b * N + lo) * DD);          // theta_s[k] -> A_k
        double* sT1 = base + 2 * SM + tile_off(th1 + (b * n + lo) * D);       // theta_1[k] -> b_k
        const BwdChunk ch = bwd_chunk<L>(cnt);
        InfoElem<D> e = Ops::identity();
        for (int k = ch.top - 1; k >= ch.bot; --k)
            Ops::push(e, mat_load<D>(sTs + k * DD), mirror_lower(-2.0 * mat_load<D>(sTd + k * DD)),
                      vec_load<D>(sT1 + k * D));
        if (NSTAGE > 1 && threadIdx.x == 0 && ti + 1 < ntiles) {
            bulk_wait_read();
            issue(ti + 1);
        }
        InfoElem<D> total;
        const InfoElem<D> excl = block_exclusive_scan<Ops, T>(e, wtot, total);
        if (MODE != SWEEP_FULL) {
            seg_tot = Ops::combine(seg_tot, total);
            continue;
        }
        InfoState<D> s = Ops::apply(excl, st);
        st = Ops::apply(total, st);
        bool ok = true;
        for (int k = ch.top - 1; k >= ch.bot; --k) {
            const Mat<D> Ts = mat_load<D>(sTs + k * DD);
            const Mat<D> Kd = -2.0 * mat_load<D>(sTd + k * DD);
            const Vec<D> t1k = vec_load<D>(sT1 + k * D);
            Mat<D> cov, cc;
            cov_from_precision(s.J, cov, cc, ok);
            if (!ok) { report_not_pd(status, b * n + lo + k + 1); ok = true; }
            const Mat<D> Ak = mm(cov, Ts);
            const Vec<D> bk = mv(cov, s.h);
            mat_store<D>(sTs + k * DD, Ak);
            mat_store<D>(sTd + k * DD, cc);
            vec_store<D>(sT1 + k * D, bk);
            s.J = mirror_lower(Kd - Tmm(Ts, Ak));
            s.h = t1k + Tmv(Ts, bk);
        }
        if (first && lo == 0 && ch.len > 0 && ch.bot == 0) {   // node 0 of the chain: the initial state
            Mat<D> cov, cc;
            cov_from_precision(s.J, cov, cc, ok);
            if (!ok) report_not_pd(status, b * n);
            vec_store<D>(mu0 + b * D, mv(cov, s.h));
            mat_store<D>(cholP0 + b * DD, cc);
        }
        fence_async_smem();
        __syncthreads();
        tile_store<T>(A + (b * N + lo) * DD, sTs, cnt * DD);
        tile_store<T>(cholQ + (b * N + lo) * DD, sTd, cnt * DD);
        tile_store<T>(bo + (b * N + lo) * D, sT1, cnt * D);
        if (threadIdx.x == 0) bulk_commit();
    }
    if (threadIdx.x == 0) {
        if (MODE != SWEEP_FULL) seg_elem[b * nseg + seg] = seg_tot;
        bulk_wait_all();
    }
}

template <int D, int T, int L, int NSTAGE = 1>
constexpr size_t info_tiled_smem() {
    return ((size_t)kRingHeaderDoubles + (size_t)NSTAGE * (2 * tile_slots(T * L * D * D) + tile_slots(T * L * D))) *
               sizeof(double) +
           (T / 32) * sizeof(InfoElem<D>);
}

// ---------------------------------------------------------------------------------------------------
// Marginal means and covariances of a state-space model by the moment recursion
//     m_{k+1} = A_k m_k + b_k,     S_{k+1} = A_k S_k A_k^T + Q_k,     Q_k = chol Q_k chol Q_k^T
// (what StateSpaceModel.marginals, state_space_model.py:231-262, returns; the reference goes through the precision and
// its inverse band).  Forward sweep; the two recursions share A and are scanned together.
// out: m [B,N+1,D], Sd [B,N+1,D,D].  Time-sharded: `first` = the stretch starts at node 0 of the chain (state from
// mu0 / cholP0), otherwise from carry_in; the last row written is the state of the node after the stretch.
template <int D>
using MomOps = PairOps<AffSymOps<D>, AffVecOps<D>>;

template <int D, int T, int L, int MODE, int NSTAGE = 1>
__global__ void __launch_bounds__(T)
k_moments_tiled(const double* __restrict__ A, const double* __restrict__ bo, const double* __restrict__ cholQ,
                const double* __restrict__ mu0, const double* __restrict__ cholP0, double* __restrict__ m,
                double* __restrict__ Sd, int64_t B, int64_t N, int nseg, int64_t seg_len,
                const typename MomOps<D>::State* __restrict__ carry_in,
                typename MomOps<D>::State* __restrict__ init_out, typename MomOps<D>::Elem* __restrict__ seg_elem,
                int first = 1) {
    constexpr int DD = D * D, TL = T * L;
    using Ops = MomOps<D>;
    constexpr int SM = tile_slots(TL * DD), SV = tile_slots(TL * D);
    constexpr int kStage = 2 * SM + SV;                                    // A, chol Q, b
    extern __shared__ __align__(16) double smem[];
    double* stage0 = smem + kRingHeaderDoubles;
    typename Ops::Elem* wtot = reinterpret_cast<typename Ops::Elem*>(stage0 + NSTAGE * kStage);
    TileRing<NSTAGE> ring;
    ring.init(smem);
    const int64_t b = blockIdx.x / nseg;
    const int seg = (int)(blockIdx.x % nseg);
    const int64_t t0 = seg * seg_len, t1 = min(N, t0 + seg_len);
    const int64_t n = N + 1;
    const int ntiles = (int)((t1 - t0 + TL - 1) / TL);
    auto issue = [&](int ti) {
        const int64_t tk = t0 + (int64_t)ti * TL;
        const int cnt = (int)min((int64_t)TL, t1 - tk);
        double* base = stage0 + (ti % NSTAGE) * kStage;
        uint64_t* bar = ring.bar + ti % NSTAGE;
        uint32_t bytes = tile_fetch(base, A + (b * N + tk) * DD, cnt * DD, bar);
        bytes += tile_fetch(base + SM, cholQ + (b * N + tk) * DD, cnt * DD, bar);
        bytes += tile_fetch(base + 2 * SM, bo + (b * N + tk) * D, cnt * D, bar);
        mbar_arrive_expect_tx(bar, bytes);
    };
    if (threadIdx.x == 0 && ntiles > 0) issue(0);
    typename Ops::State st;
    st.a = mat_zero<D>();
    st.b = vec_zero<D>();
    if (MODE == SWEEP_FULL && (seg > 0 || !first)) st = carry_in[b * nseg + seg];
    typename Ops::Elem seg_tot = Ops::identity();
    for (int ti = 0; ti < ntiles; ++ti) {
        const int64_t tk = t0 + (int64_t)ti * TL;
        const int cnt = (int)min((int64_t)TL, t1 - tk);
        if (NSTAGE == 1 && ti > 0) {
            __syncthreads();
            if (threadIdx.x == 0) { bulk_wait_read(); issue(ti); }
        }
        ring.wait(ti);
        double* base = stage0 + (ti % NSTAGE) * kStage;
        double* sA = base + tile_off(A + (b * N + tk) * DD);                  // A of transition i
        double* sQ = base + SM + tile_off(cholQ + (b * N + tk) * DD);         // chol Q of transition i -> S of node i + 1
        double* sb = base + 2 * SM + tile_off(bo + (b * N + tk) * D);         // offset of transition i -> m of node i + 1
        if (seg == 0 && tk == 0) {
            if (first) {
                const Mat<D> c0 = mat_load<D>(cholP0 + b * DD);
                st.a = mmT(c0, c0);
                st.b = vec_load<D>(mu0 + b * D);
                if (MODE != SWEEP_FULL && threadIdx.x == 0) init_out[b] = st;
            }
            if (MODE == SWEEP_FULL && threadIdx.x == 0) {
                vec_store<D>(m + b * n * D, st.b);
                mat_store<D>(Sd + b * n * DD, st.a);
            }
        }
        const int k0 = threadIdx.x * L;
        const int len = min(L, cnt - k0);
        typename Ops::Elem e = Ops::identity();
        for (int t = 0; t < len; ++t) {
            typename Ops::Elem s;
            const Mat<D> Ak = mat_load<D>(sA + (k0 + t) * DD), cc = mat_load<D>(sQ + (k0 + t) * DD);
            s.a.G = transpose(Ak);
            s.a.C = mmT(cc, cc);
            s.b.M = Ak;
            s.b.v = vec_load<D>(sb + (k0 + t) * D);
            e = Ops::combine(e, s);
        }
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
        typename Ops::State s = Ops::apply(excl, st);
        st = Ops::apply(total, st);
        for (int t = 0; t < len; ++t) {
            const Mat<D> Ak = mat_load<D>(sA + (k0 + t) * DD), cc = mat_load<D>(sQ + (k0 + t) * DD);
            s.b = vec_load<D>(sb + (k0 + t) * D) + mv(Ak, s.b);
            s.a = sym(mmT(cc, cc) + mmT(mm(Ak, s.a), Ak));
            vec_store<D>(sb + (k0 + t) * D, s.b);
            mat_store<D>(sQ + (k0 + t) * DD, s.a);
        }
        fence_async_smem();
        __syncthreads();
        tile_store<T>(m + (b * n + tk + 1) * D, sb, cnt * D);
        tile_store<T>(Sd + (b * n + tk + 1) * DD, sQ, cnt * DD);
        if (threadIdx.x == 0) bulk_commit();
    }
    if (threadIdx.x == 0) {
        if (MODE != SWEEP_FULL) seg_elem[b * nseg + seg] = seg_tot;
        bulk_wait_all();
    }
}

template <int D, int T, int L, int NSTAGE = 1>
constexpr size_t moments_tiled_smem() {
    return ((size_t)kRingHeaderDoubles + (size_t)NSTAGE * (2 * tile_slots(T * L * D * D) + tile_slots(T * L * D))) *
               sizeof(double) +
           (T / 32) * sizeof(typename MomOps<D>::Elem);
}


}  // namespace vdp
