// CVISitesTrainer._optimize_sites_under_stable_prior on the device (docs/diffusion_processes/cvi_dp_trainer.py:63-101):
// the site-update loop as a CUDA-graph WHILE node.  The body is one vdp_cvi_site_update (captured from the same launch
// sequence a direct call makes) followed by two small kernels: k_loop_record gathers the marginals at the test grid
// points (NLPD / RMSE are computed from them afterwards), k_loop_check applies the reference's per-iteration rules to
// every sequence -- ELBO dropped => step sizes / 10 (:84-87), |ELBO change| < tol => stop (:89-91) -- and sets the
// condition of the WHILE node.  The step sizes live in device memory (vdp_cvi_problem.lr_*_dev), so the captured
// launches pick up a decayed value without re-capture, and a stopped sequence is updated with step size 0, which
// leaves its state as it is.  Included by vdp_all.inl (translation unit "cvi").
#include <map>
#include <mutex>
#include <string>

namespace vdp {

template <int D>
__global__ void k_loop_record(const double* __restrict__ fx_mus, const double* __restrict__ fx_covs,
                              const int64_t* __restrict__ test_idx, double* __restrict__ test_mus,
                              double* __restrict__ test_covs, const int32_t* __restrict__ counters, int64_t B,
                              int64_t N, int n_test) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * n_test) return;
    constexpr int DD = D * D;
    const int64_t b = i / n_test;
    const int64_t k = test_idx[i];
    const int64_t it = counters[0];                 // updates finished before this one
    const int64_t node = b * (N + 1) + (k < 0 ? 0 : (k > N ? N : k));
    const int64_t o = it * B * n_test + i;
    vec_store<D>(test_mus + o * D, vec_load<D>(fx_mus + node * D));
    mat_store<D>(test_covs + o * DD, mat_load<D>(fx_covs + node * DD));
}

// one block; use_handle: set the WHILE condition
static __global__ void k_loop_check(const double* __restrict__ elbo, double* __restrict__ trace, double* __restrict__ lr_data,
                             double* __restrict__ lr_girsanov, int32_t* __restrict__ active, int32_t* __restrict__ n_itr,
                             int32_t* __restrict__ counters, int64_t B, int max_itr, double tol,
                             cudaGraphConditionalHandle handle, int use_handle) {
    __shared__ int n_active;
    if (threadIdx.x == 0) n_active = 0;
    __syncthreads();
    const int it = counters[0] + 1;                 // this update's number, 1-based
    int mine = 0;
    for (int64_t b = threadIdx.x; b < B; b += blockDim.x) {
        const double prev = trace[(int64_t)(it - 1) * B + b];
        if (!active[b]) { trace[(int64_t)it * B + b] = prev; continue; }
        const double e = elbo[b];
        trace[(int64_t)it * B + b] = e;
        n_itr[b] = it;
        // len(tmp_elbo_vals) > 2  <=>  it >= 2   (the list starts with the ELBO before the loop)
        if (it >= 2 && prev > e) { lr_girsanov[b] = lr_girsanov[b] / 10; lr_data[b] = lr_data[b] / 10; }
        if (it >= 2 && fabs(prev - e) < tol) active[b] = 0;
        else ++mine;
    }
    if (mine) atomicAdd(&n_active, mine);
    __syncthreads();
    if (threadIdx.x == 0) {
        counters[0] = it;
        counters[1] = n_active;
        if (use_handle) cudaGraphSetConditional(handle, (n_active > 0 && it < max_itr) ? 1u : 0u);
    }
}

static __global__ void k_loop_init(int32_t* __restrict__ active, int32_t* __restrict__ n_itr,
                                   int32_t* __restrict__ counters, int64_t B) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b < B) { active[b] = 1; n_itr[b] = 0; }
    if (b == 0) { counters[0] = 0; counters[1] = (int32_t)B; }
}

struct LoopGraph {
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
};
static std::map<std::string, LoopGraph> g_loop_graphs;
static std::mutex g_loop_mutex;

template <int DRIFT, int D>
static int loop_body(const vdp_cvi_problem& p, const vdp_cvi_loop& lp, const CviWs& w, cudaStream_t st,
                     cudaGraphConditionalHandle handle, int use_handle) {
    int rc = run_site_update<DRIFT, D>(p, w, st);
    if (rc) return rc;
    if (lp.n_test > 0 && lp.test_idx && lp.test_mus && lp.test_covs) {
        k_loop_record<D><<<grid_for(p.B * lp.n_test, kPtBlock), kPtBlock, 0, st>>>(
            p.fx_mus, p.fx_covs, lp.test_idx, lp.test_mus, lp.test_covs, lp.counters, p.B, p.N, lp.n_test);
        VDP_CHECK_LAUNCH();
    }
    k_loop_check<<<1, 256, 0, st>>>(p.elbo, lp.elbo_trace, lp.lr_data, lp.lr_girsanov, lp.active, lp.n_itr, lp.counters,
                                    p.B, lp.max_itr, lp.optim_tol, handle, use_handle);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

}  // namespace vdp

extern "C" int vdp_cvi_optimize_sites(const vdp_cvi_problem* hp, const vdp_cvi_loop* lp, void* stream) {
    int rc = check_problem(hp);
    if (rc) return rc;
    VDP_REQUIRE(lp && lp->max_itr >= 1 && lp->lr_data && lp->lr_girsanov && lp->active && lp->n_itr && lp->elbo_trace &&
                    lp->counters, "loop arrays");
    if (hp->B == 0) return VDP_OK;
    CviWs w;
    if (!cvi_ws_take(hp->workspace, hp->workspace_bytes, hp->B, hp->N, hp->D, w)) {
        set_error("workspace too small: need %zu bytes", cvi_ws_bytes(hp->B, hp->N, hp->D));
        return VDP_E_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    vdp_cvi_problem p = *hp;                       // every update reads its step sizes from the device
    p.lr_data_dev = lp->lr_data;
    p.lr_girsanov_dev = lp->lr_girsanov;
    p.active_dev = lp->active;
    p.flags &= ~(VDP_FLAG_SKIP_ELBO | VDP_FLAG_SKIP_GIRSANOV | VDP_FLAG_SKIP_DATA_UPDATE);
    k_loop_init<<<grid_for(p.B, kPtBlock), kPtBlock, 0, st>>>(lp->active, lp->n_itr, lp->counters, p.B);
    VDP_CHECK_LAUNCH();
    // first update: the caller's flags (the naturals may have to be rebuilt from the sites)
    VDP_DISPATCH_D(p.D, VDP_DISPATCH_DRIFT(p.drift_id, {
        if ((rc = loop_body<kDrift, kD>(p, *lp, w, st, 0, 0))) return rc;
    }));
    if (lp->max_itr < 2) return VDP_OK;
    // updates 2 .. max_itr: WHILE node; the graph is kept for the next call with the same problem
    vdp_cvi_problem q = p;
    q.flags = lp->later_flags & ~(VDP_FLAG_SKIP_ELBO | VDP_FLAG_SKIP_GIRSANOV | VDP_FLAG_SKIP_DATA_UPDATE);
    std::string key(reinterpret_cast<const char*>(&q), sizeof(q));
    key.append(reinterpret_cast<const char*>(lp), sizeof(*lp));
    for (const char* k : {"SWEEPS", "NSEG", "GH_QUADRATURE", "REFACTOR_PRECISION", "NO_CARRIED_NATURALS",
                          "NO_FUSED_NATURALS", "SEPARATE_KL_VALUE", "NO_FUSED_FORWARD"}) {
        int v = 0;
        vdp_config_get(k, &v);
        key.append(reinterpret_cast<const char*>(&v), sizeof(v));
    }
    std::lock_guard<std::mutex> lock(g_loop_mutex);
    auto it = g_loop_graphs.find(key);
    if (it == g_loop_graphs.end()) {
        if (g_loop_graphs.size() >= 8) {
            for (auto& kv : g_loop_graphs) {
                if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
                if (kv.second.graph) cudaGraphDestroy(kv.second.graph);
            }
            g_loop_graphs.clear();
        }
        LoopGraph lg;
        VDP_CHECK_CUDA(cudaGraphCreate(&lg.graph, 0));
        cudaGraphConditionalHandle handle;
        // default value 1 at every launch: the first update has left at least its own sequence(s) active
        // (k_loop_check never stops a sequence at it = 1) and max_itr >= 2
        VDP_CHECK_CUDA(cudaGraphConditionalHandleCreate(&handle, lg.graph, 1, cudaGraphCondAssignDefault));
        cudaGraphNodeParams np = {};
        np.type = cudaGraphNodeTypeConditional;
        np.conditional.handle = handle;
        np.conditional.type = cudaGraphCondTypeWhile;
        np.conditional.size = 1;
        cudaGraphNode_t node;
        VDP_CHECK_CUDA(cudaGraphAddNode(&node, lg.graph, nullptr, 0, &np));
        cudaGraph_t body = np.conditional.phGraph_out[0];
        cudaStream_t cs;
        VDP_CHECK_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
        VDP_CHECK_CUDA(cudaStreamBeginCaptureToGraph(cs, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
        int brc = VDP_OK;
        auto capture = [&]() -> int {
            VDP_DISPATCH_D(q.D, VDP_DISPATCH_DRIFT(q.drift_id, {
                int r2 = loop_body<kDrift, kD>(q, *lp, w, cs, handle, 1);
                if (r2) return r2;
            }));
            return VDP_OK;
        };
        brc = capture();
        cudaGraph_t captured = nullptr;
        const cudaError_t ce = cudaStreamEndCapture(cs, &captured);
        cudaStreamDestroy(cs);
        if (brc) { cudaGraphDestroy(lg.graph); return brc; }
        if (ce != cudaSuccess) {
            set_error("capture of the site-update loop failed: %s", cudaGetErrorString(ce));
            cudaGraphDestroy(lg.graph);
            return VDP_E_CUDA;
        }
        VDP_CHECK_CUDA(cudaGraphInstantiate(&lg.exec, lg.graph, 0));
        it = g_loop_graphs.emplace(key, lg).first;
    }
    VDP_CHECK_CUDA(cudaGraphLaunch(it->second.exec, st));
    return VDP_OK;
}
