// Host-buffer variant of the CVI-DP site update: every array of the problem lives in HOST memory (pinned
// for full PCIe speed).  The B independent sequences are cut in chunks; chunk c runs on stream c % slots with
// its own device slot:  H2D copies -> vdp_cvi_site_update -> D2H copies, all stream-ordered.  With >= 3 slots
// the upload of chunk c+1, the kernels of chunk c and the download of chunk c-1 overlap (two copy engines +
// SMs), so the call costs max(H2D, D2H, compute) instead of their sum.  Chunking over sequences is exact:
// no kernel on the path mixes sequences.
namespace vdp {

struct HostSlot {
    cudaStream_t st = nullptr;
    double *pA, *pb, *pcholQ, *pmu0, *pcholP0, *g1, *gd, *gs, *d1, *d2, *obs_y, *fx_mus, *fx_covs;
    double *qA, *qb, *qcholQ, *qmu0, *qcholP0, *elbo, *ve, *kl;
    int64_t* obs_idx;
    void* ws;
    // posterior means at the observed grid points: gathered on the host (pinned staging), scattered on the device
    double *mu_obs_dev = nullptr, *mu_obs_host = nullptr;
    cudaEvent_t staged = nullptr;
};

template <int D>
__global__ void k_set_at_obs(const double* __restrict__ mu_obs, const int64_t* __restrict__ obs_idx,
                             double* __restrict__ fx_mus, int64_t B, int64_t N, int64_t n_obs) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= B * n_obs) return;
    const int64_t b = i / n_obs, k = obs_idx[i];
    if (k < 0 || k > N) return;
#pragma unroll
    for (int e = 0; e < D; ++e) fx_mus[(b * (N + 1) + k) * D + e] = mu_obs[i * D + e];
}

}  // namespace vdp

struct vdp_cvi_host_ctx {
    int64_t chunk, N, n_obs;
    int D, device, nslots;
    size_t ws_bytes;
    char* pool = nullptr;            // one device allocation for all slots
    double* time_grid = nullptr;     // [N+1]
    int64_t* status = nullptr;       // device int64[2]
    vdp::HostSlot slot[4];
    // optional device-resident copy of the linearised prior of all B sequences (vdp_cvi_host_set_prior)
    int64_t prior_B = 0;
    char* prior_pool = nullptr;
    double *rA = nullptr, *rb = nullptr, *rQ = nullptr, *rmu0 = nullptr, *rP0 = nullptr;
};

namespace vdp {

struct HostSizes {
    size_t vN1, mN1, vN, mN, v1, m1, vo, mo, io, s;
    HostSizes(int64_t B, int64_t N, int D, int64_t n_obs) {
        const size_t d = D, dd = (size_t)D * D, b = (size_t)B;
        vN1 = padded(b * (N + 1) * d);  mN1 = padded(b * (N + 1) * dd);
        vN = padded(b * N * d);         mN = padded(b * N * dd);
        v1 = padded(b * d);             m1 = padded(b * dd);
        vo = padded(b * n_obs * d + 1); mo = padded(b * n_obs * dd + 1);
        io = padded(b * n_obs + 1);     s = padded(b);
    }
    size_t slot_bytes(size_t ws) const {
        return (mN + vN + mN + v1 + m1)            // prior
               + (vN1 + mN1 + mN) + (vo + mo)      // sites
               + io + 2 * vo                       // observations, gathered means
               + (vN1 + mN1)                       // marginals
               + (mN + vN + mN + v1 + m1)          // posterior SSM
               + 3 * s + ((ws + 255) & ~size_t(255));
    }
};

static inline cudaError_t h2d(void* dst, const void* src, size_t bytes, cudaStream_t st) {
    return bytes ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st) : cudaSuccess;
}
static inline cudaError_t d2h(void* dst, const void* src, size_t bytes, cudaStream_t st) {
    return (bytes && dst) ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess;
}

}  // namespace vdp

extern "C" {

int vdp_cvi_host_destroy(vdp_cvi_host_ctx* c) {
    if (!c) return VDP_OK;
    cudaSetDevice(c->device);
    for (int i = 0; i < c->nslots; ++i)
        if (c->slot[i].st) {
            cudaStreamSynchronize(c->slot[i].st);
            cudaStreamDestroy(c->slot[i].st);
            if (c->slot[i].staged) cudaEventDestroy(c->slot[i].staged);
            if (c->slot[i].mu_obs_host) cudaFreeHost(c->slot[i].mu_obs_host);
        }
    if (c->pool) cudaFree(c->pool);
    if (c->prior_pool) cudaFree(c->prior_pool);
    delete c;
    return VDP_OK;
}

int vdp_cvi_host_create(vdp_cvi_host_ctx** out, int64_t chunk, int64_t N, int D, int64_t n_obs, int slots,
                        int device) {
    VDP_REQUIRE(out && chunk >= 1 && N >= 1 && D >= 1 && D <= 3 && n_obs >= 0, "out, chunk, N, D, n_obs");
    VDP_REQUIRE(slots >= 1 && slots <= 4, "slots must be 1..4");
    VDP_CHECK_CUDA(cudaSetDevice(device));
    vdp_cvi_host_ctx* c = new vdp_cvi_host_ctx();
    c->chunk = chunk; c->N = N; c->n_obs = n_obs; c->D = D; c->device = device; c->nslots = slots;
    c->ws_bytes = vdp_cvi_workspace(chunk, N, D, n_obs);
    const HostSizes z(chunk, N, D, n_obs);
    const size_t per_slot = z.slot_bytes(c->ws_bytes);
    const size_t total = per_slot * slots + padded(N + 1) + 256;
    if (cudaMalloc((void**)&c->pool, total) != cudaSuccess) {
        cudaGetLastError();
        set_error("vdp_cvi_host_create: cannot allocate %zu device bytes", total);
        delete c;
        return VDP_E_CUDA;
    }
    Arena ar(c->pool, total);
    c->time_grid = ar.take(N + 1);
    c->status = (int64_t*)ar.take_bytes(256);
    for (int i = 0; i < slots; ++i) {
        HostSlot& s = c->slot[i];
        const size_t d = D, dd = (size_t)D * D, b = (size_t)chunk;
        s.pA = ar.take(b * N * dd); s.pb = ar.take(b * N * d); s.pcholQ = ar.take(b * N * dd);
        s.pmu0 = ar.take(b * d); s.pcholP0 = ar.take(b * dd);
        s.g1 = ar.take(b * (N + 1) * d); s.gd = ar.take(b * (N + 1) * dd); s.gs = ar.take(b * N * dd);
        s.d1 = ar.take(b * n_obs * d + 1); s.d2 = ar.take(b * n_obs * dd + 1);
        s.obs_idx = (int64_t*)ar.take(b * n_obs + 1); s.obs_y = ar.take(b * n_obs * d + 1);
        s.fx_mus = ar.take(b * (N + 1) * d); s.fx_covs = ar.take(b * (N + 1) * dd);
        s.qA = ar.take(b * N * dd); s.qb = ar.take(b * N * d); s.qcholQ = ar.take(b * N * dd);
        s.qmu0 = ar.take(b * d); s.qcholP0 = ar.take(b * dd);
        s.elbo = ar.take(b); s.ve = ar.take(b); s.kl = ar.take(b);
        s.mu_obs_dev = ar.take(b * n_obs * d + 1);
        s.ws = ar.take_bytes(c->ws_bytes);
        if (!s.ws || cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&s.staged, cudaEventDisableTiming) != cudaSuccess ||
            cudaHostAlloc((void**)&s.mu_obs_host, (b * n_obs * d + 1) * sizeof(double), cudaHostAllocDefault) !=
                cudaSuccess) {
            set_error("vdp_cvi_host_create: slot %d setup failed", i);
            vdp_cvi_host_destroy(c);
            return VDP_E_CUDA;
        }
    }
    *out = c;
    return VDP_OK;
}

int vdp_cvi_host_set_prior(vdp_cvi_host_ctx* c, int64_t B, const double* pA, const double* pb, const double* pcholQ,
                           const double* pmu0, const double* pcholP0) {
    VDP_REQUIRE(c && B >= 1 && pA && pb && pcholQ && pmu0 && pcholP0, "ctx, B, prior pointers (host)");
    VDP_CHECK_CUDA(cudaSetDevice(c->device));
    if (c->prior_pool && c->prior_B != B) { cudaFree(c->prior_pool); c->prior_pool = nullptr; }
    const size_t d = c->D, dd = (size_t)c->D * c->D, b = (size_t)B, N = (size_t)c->N;
    const size_t total = 2 * padded(b * N * dd) + padded(b * N * d) + padded(b * d) + padded(b * dd);
    if (!c->prior_pool && cudaMalloc((void**)&c->prior_pool, total) != cudaSuccess) {
        cudaGetLastError();
        set_error("vdp_cvi_host_set_prior: cannot allocate %zu device bytes", total);
        return VDP_E_CUDA;
    }
    c->prior_B = B;
    Arena ar(c->prior_pool, total);
    c->rA = ar.take(b * N * dd); c->rb = ar.take(b * N * d); c->rQ = ar.take(b * N * dd);
    c->rmu0 = ar.take(b * d); c->rP0 = ar.take(b * dd);
    cudaStream_t st = c->slot[0].st;
    VDP_CHECK_CUDA(h2d(c->rA, pA, b * N * dd * 8, st));
    VDP_CHECK_CUDA(h2d(c->rb, pb, b * N * d * 8, st));
    VDP_CHECK_CUDA(h2d(c->rQ, pcholQ, b * N * dd * 8, st));
    VDP_CHECK_CUDA(h2d(c->rmu0, pmu0, b * d * 8, st));
    VDP_CHECK_CUDA(h2d(c->rP0, pcholP0, b * dd * 8, st));
    VDP_CHECK_CUDA(cudaStreamSynchronize(st));
    return VDP_OK;
}

int vdp_cvi_host_site_update(vdp_cvi_host_ctx* c, const vdp_cvi_problem* hp, int64_t* status_host2) {
    VDP_REQUIRE(c && hp, "ctx, problem");
    VDP_REQUIRE(hp->N == c->N && hp->D == c->D && hp->n_obs == c->n_obs, "problem shape differs from the context");
    VDP_REQUIRE(hp->time_grid && hp->g1 && hp->gd && hp->gs && hp->fx_mus && hp->fx_covs, "host state pointers");
    VDP_REQUIRE(hp->n_obs == 0 || (hp->d1 && hp->d2 && hp->obs_idx && hp->obs_y), "host data-site pointers");
    const bool resident = (hp->pA == nullptr);
    if (resident) {
        VDP_REQUIRE(c->prior_pool && c->prior_B == hp->B, "pA == NULL needs vdp_cvi_host_set_prior for the same B");
    } else {
        VDP_REQUIRE(hp->pb && hp->pcholQ && hp->pmu0 && hp->pcholP0, "host prior pointers");
    }
    VDP_CHECK_CUDA(cudaSetDevice(c->device));
    const int64_t B = hp->B, N = c->N, no = c->n_obs;
    const size_t d = c->D, dd = d * d, S = sizeof(double);
    cudaStream_t s0 = c->slot[0].st;
    VDP_CHECK_CUDA(cudaMemsetAsync(c->status, 0, 2 * sizeof(int64_t), s0));
    VDP_CHECK_CUDA(h2d(c->time_grid, hp->time_grid, (N + 1) * S, s0));
    VDP_CHECK_CUDA(cudaStreamSynchronize(s0));
    int rc = VDP_OK;
    int64_t ci = 0;
    for (int64_t b0 = 0; b0 < B && rc == VDP_OK; b0 += c->chunk, ++ci) {
        const size_t nb = (size_t)std::min<int64_t>(c->chunk, B - b0), o = (size_t)b0;
        HostSlot& s = c->slot[ci % c->nslots];
        cudaStream_t st = s.st;
        vdp_cvi_problem p = *hp;
        p.B = (int64_t)nb;
        p.time_grid = c->time_grid;
        cudaError_t e = cudaSuccess;
        auto up = [&](void* dst, const void* src, size_t per_seq) {
            if (e == cudaSuccess) e = h2d(dst, (const char*)src + o * per_seq, nb * per_seq, st);
        };
        auto down = [&](void* dst, const void* src, size_t per_seq) {
            if (e == cudaSuccess && dst) e = d2h((char*)dst + o * per_seq, src, nb * per_seq, st);
        };
        if (resident) {
            p.pA = c->rA + o * N * dd; p.pb = c->rb + o * N * d; p.pcholQ = c->rQ + o * N * dd;
            p.pmu0 = c->rmu0 + o * d; p.pcholP0 = c->rP0 + o * dd;
        } else {
            up(s.pA, hp->pA, N * dd * S); up(s.pb, hp->pb, N * d * S); up(s.pcholQ, hp->pcholQ, N * dd * S);
            up(s.pmu0, hp->pmu0, d * S); up(s.pcholP0, hp->pcholP0, dd * S);
            p.pA = s.pA; p.pb = s.pb; p.pcholQ = s.pcholQ; p.pmu0 = s.pmu0; p.pcholP0 = s.pcholP0;
        }
        up(s.g1, hp->g1, (N + 1) * d * S); up(s.gd, hp->gd, (N + 1) * dd * S); up(s.gs, hp->gs, N * dd * S);
        // fx_covs is an output only, and of fx_mus the data-site update reads the observed grid points only:
        // gather those on the host, upload n_obs values per sequence instead of N + 1
        const bool need_mu = no > 0 && !(hp->flags & VDP_FLAG_SKIP_DATA_UPDATE);
        if (no > 0) {
            up(s.d1, hp->d1, no * d * S); up(s.d2, hp->d2, no * dd * S);
            up(s.obs_idx, hp->obs_idx, no * sizeof(int64_t)); up(s.obs_y, hp->obs_y, no * d * S);
        }
        if (need_mu && e == cudaSuccess) {
            e = cudaEventSynchronize(s.staged);      // the previous upload from this staging buffer is over
            for (size_t bb = 0; bb < nb; ++bb)
                for (int64_t j = 0; j < no; ++j) {
                    int64_t k = hp->obs_idx[(o + bb) * no + j];
                    k = k < 0 ? 0 : (k > N ? N : k);
                    for (size_t q = 0; q < d; ++q)
                        s.mu_obs_host[(bb * no + j) * d + q] = hp->fx_mus[((o + bb) * (N + 1) + k) * d + q];
                }
            if (e == cudaSuccess) e = h2d(s.mu_obs_dev, s.mu_obs_host, nb * no * d * S, st);
            if (e == cudaSuccess) e = cudaEventRecord(s.staged, st);
            if (e == cudaSuccess) {
                VDP_DISPATCH_D((int)d, (k_set_at_obs<kD><<<grid_for((int64_t)nb * no, kPtBlock), kPtBlock, 0, st>>>(
                                          s.mu_obs_dev, s.obs_idx, s.fx_mus, (int64_t)nb, N, no)));
                e = cudaGetLastError();
            }
        }
        if (e != cudaSuccess) { set_error("H2D copy: %s", cudaGetErrorString(e)); rc = VDP_E_CUDA; break; }
        p.g1 = s.g1; p.gd = s.gd; p.gs = s.gs; p.d1 = s.d1; p.d2 = s.d2; p.obs_idx = s.obs_idx; p.obs_y = s.obs_y;
        p.fx_mus = s.fx_mus; p.fx_covs = s.fx_covs;
        p.qA = s.qA; p.qb = s.qb; p.qcholQ = s.qcholQ; p.qmu0 = s.qmu0; p.qcholP0 = s.qcholP0;
        p.elbo = s.elbo; p.ve = s.ve; p.kl = s.kl;
        p.workspace = s.ws; p.workspace_bytes = c->ws_bytes; p.status = c->status;
        if ((rc = vdp_cvi_site_update(&p, st))) break;
        down(hp->g1, s.g1, (N + 1) * d * S); down(hp->gd, s.gd, (N + 1) * dd * S); down(hp->gs, s.gs, N * dd * S);
        down(hp->fx_mus, s.fx_mus, (N + 1) * d * S); down(hp->fx_covs, s.fx_covs, (N + 1) * dd * S);
        if (no > 0) { down(hp->d1, s.d1, no * d * S); down(hp->d2, s.d2, no * dd * S); }
        down(hp->qA, s.qA, N * dd * S); down(hp->qb, s.qb, N * d * S); down(hp->qcholQ, s.qcholQ, N * dd * S);
        down(hp->qmu0, s.qmu0, d * S); down(hp->qcholP0, s.qcholP0, dd * S);
        down(hp->elbo, s.elbo, S); down(hp->ve, s.ve, S); down(hp->kl, s.kl, S);
        if (e != cudaSuccess) { set_error("D2H copy: %s", cudaGetErrorString(e)); rc = VDP_E_CUDA; }
    }
    for (int i = 0; i < c->nslots; ++i)
        if (cudaStreamSynchronize(c->slot[i].st) != cudaSuccess && rc == VDP_OK) {
            set_error("stream synchronize: %s", cudaGetErrorString(cudaGetLastError()));
            rc = VDP_E_CUDA;
        }
    if (rc == VDP_OK && status_host2)
        VDP_CHECK_CUDA(cudaMemcpy(status_host2, c->status, 2 * sizeof(int64_t), cudaMemcpyDeviceToHost));
    return rc;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// Resident-state session: the state of the B sequences (girsanov and data sites, linearised prior, posterior,
// marginals, carried naturals) stays on the device between calls, as the reference's model object keeps its
// tf.Variables between trainer iterations; per site update the host supplies that update's inputs -- the observations
// -- and receives its results: ELBO terms and the posterior marginals (what cvi_dp_trainer.py:75-77 reads every
// iteration for ELBO / NLPD / RMSE).  Sequences are processed in chunks on rotating streams so that the download of
// one chunk's marginals overlaps the kernels of the next.
struct vdp_cvi_session {
    vdp_cvi_problem dev;          // device-resident problem of all B sequences (pointers into `pool`)
    int64_t chunk = 0;
    int device = 0, nstreams = 0, updates = 0;
    char* pool = nullptr;
    cudaStream_t st[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t ws_chunk = 0;          // workspace bytes of one chunk (each chunk keeps its own: carried naturals)
};

extern "C" {

int vdp_cvi_session_destroy(vdp_cvi_session* s) {
    if (!s) return VDP_OK;
    cudaSetDevice(s->device);
    for (int i = 0; i < s->nstreams; ++i)
        if (s->st[i]) { cudaStreamSynchronize(s->st[i]); cudaStreamDestroy(s->st[i]); }
    if (s->pool) cudaFree(s->pool);
    delete s;
    return VDP_OK;
}

int vdp_cvi_session_create(vdp_cvi_session** out, const vdp_cvi_problem* hp, int64_t chunk, int streams, int device) {
    VDP_REQUIRE(out && hp && chunk >= 1 && streams >= 1 && streams <= 4, "out, problem, chunk, streams (1..4)");
    VDP_REQUIRE(hp->B >= 1 && hp->N >= 1 && hp->D >= 1 && hp->D <= 3 && hp->n_obs >= 0, "B, N, D, n_obs");
    VDP_REQUIRE(hp->time_grid && hp->pA && hp->pb && hp->pcholQ && hp->pmu0 && hp->pcholP0 && hp->g1 && hp->gd &&
                    hp->gs && hp->fx_mus && hp->fx_covs, "host state pointers");
    VDP_REQUIRE(hp->n_obs == 0 || (hp->d1 && hp->d2 && hp->obs_idx && hp->obs_y), "host data-site pointers");
    VDP_CHECK_CUDA(cudaSetDevice(device));
    vdp_cvi_session* s = new vdp_cvi_session();
    s->device = device;
    s->chunk = std::min<int64_t>(chunk, hp->B);
    s->nstreams = streams;
    s->dev = *hp;
    const int64_t B = hp->B, N = hp->N, no = hp->n_obs;
    const size_t d = hp->D, dd = d * d, b = (size_t)B, S = sizeof(double);
    const int64_t nchunks = (B + s->chunk - 1) / s->chunk;
    s->ws_chunk = (vdp_cvi_workspace(s->chunk, N, hp->D, no) + 255) & ~size_t(255);
    const size_t total = padded(N + 1) + 256 + 2 * (2 * padded(b * N * dd) + padded(b * N * d) + padded(b * d) +
                                                    padded(b * dd))                          // prior, posterior
                         + padded(b * (N + 1) * d) + padded(b * (N + 1) * dd) + padded(b * N * dd)   // sites
                         + padded(b * no * d + 1) + padded(b * no * dd + 1) + padded(b * no + 1) + padded(b * no * d + 1)
                         + padded(b * (N + 1) * d) + padded(b * (N + 1) * dd) + 3 * padded(b) + (size_t)nchunks * s->ws_chunk;
    if (cudaMalloc((void**)&s->pool, total) != cudaSuccess) {
        cudaGetLastError();
        set_error("vdp_cvi_session_create: cannot allocate %zu device bytes", total);
        delete s;
        return VDP_E_CUDA;
    }
    Arena ar(s->pool, total);
    vdp_cvi_problem& p = s->dev;
    double* tg = ar.take(N + 1);
    p.time_grid = tg;
    p.status = (int64_t*)ar.take_bytes(256);
    double *pA = ar.take(b * N * dd), *pb = ar.take(b * N * d), *pQ = ar.take(b * N * dd), *pm = ar.take(b * d),
           *pP = ar.take(b * dd);
    p.pA = pA; p.pb = pb; p.pcholQ = pQ; p.pmu0 = pm; p.pcholP0 = pP;
    p.qA = ar.take(b * N * dd); p.qb = ar.take(b * N * d); p.qcholQ = ar.take(b * N * dd); p.qmu0 = ar.take(b * d);
    p.qcholP0 = ar.take(b * dd);
    p.g1 = ar.take(b * (N + 1) * d); p.gd = ar.take(b * (N + 1) * dd); p.gs = ar.take(b * N * dd);
    p.d1 = ar.take(b * no * d + 1); p.d2 = ar.take(b * no * dd + 1);
    int64_t* oi = (int64_t*)ar.take(b * no + 1);
    double* oy = ar.take(b * no * d + 1);
    p.obs_idx = oi; p.obs_y = oy;
    p.fx_mus = ar.take(b * (N + 1) * d); p.fx_covs = ar.take(b * (N + 1) * dd);
    p.elbo = ar.take(b); p.ve = ar.take(b); p.kl = ar.take(b);
    p.workspace = ar.take_bytes((size_t)nchunks * s->ws_chunk);
    p.workspace_bytes = s->ws_chunk;
    p.lr_data_dev = p.lr_girsanov_dev = nullptr;
    p.active_dev = nullptr;
    bool ok = p.workspace != nullptr;
    for (int i = 0; i < streams && ok; ++i) ok = cudaStreamCreateWithFlags(&s->st[i], cudaStreamNonBlocking) == cudaSuccess;
    if (!ok) { set_error("vdp_cvi_session_create: setup failed"); vdp_cvi_session_destroy(s); return VDP_E_CUDA; }
    cudaStream_t st = s->st[0];
    cudaError_t e = cudaMemsetAsync(p.status, 0, 2 * sizeof(int64_t), st);
    auto up = [&](void* dst, const void* src, size_t bytes) { if (e == cudaSuccess) e = h2d(dst, src, bytes, st); };
    up(tg, hp->time_grid, (N + 1) * S);
    up(pA, hp->pA, b * N * dd * S); up(pb, hp->pb, b * N * d * S); up(pQ, hp->pcholQ, b * N * dd * S);
    up(pm, hp->pmu0, b * d * S); up(pP, hp->pcholP0, b * dd * S);
    up(p.g1, hp->g1, b * (N + 1) * d * S); up(p.gd, hp->gd, b * (N + 1) * dd * S); up(p.gs, hp->gs, b * N * dd * S);
    up(p.fx_mus, hp->fx_mus, b * (N + 1) * d * S); up(p.fx_covs, hp->fx_covs, b * (N + 1) * dd * S);
    if (no > 0) {
        up(p.d1, hp->d1, b * no * d * S); up(p.d2, hp->d2, b * no * dd * S);
        up(oi, hp->obs_idx, b * no * sizeof(int64_t)); up(oy, hp->obs_y, b * no * d * S);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
        set_error("vdp_cvi_session_create: upload failed: %s", cudaGetErrorString(e));
        vdp_cvi_session_destroy(s);
        return VDP_E_CUDA;
    }
    *out = s;
    return VDP_OK;
}

// the view of sequences [b0, b0 + nb) of the resident problem
static vdp_cvi_problem session_chunk(const vdp_cvi_session* s, int64_t b0, int64_t nb, int64_t ci) {
    vdp_cvi_problem p = s->dev;
    const size_t N = (size_t)p.N, d = (size_t)p.D, dd = d * d, no = (size_t)p.n_obs, o = (size_t)b0;
    p.B = nb;
    p.pA += o * N * dd; p.pb += o * N * d; p.pcholQ += o * N * dd; p.pmu0 += o * d; p.pcholP0 += o * dd;
    p.qA += o * N * dd; p.qb += o * N * d; p.qcholQ += o * N * dd; p.qmu0 += o * d; p.qcholP0 += o * dd;
    p.g1 += o * (N + 1) * d; p.gd += o * (N + 1) * dd; p.gs += o * N * dd;
    p.d1 += o * no * d; p.d2 += o * no * dd; p.obs_idx += o * no; p.obs_y += o * no * d;
    p.fx_mus += o * (N + 1) * d; p.fx_covs += o * (N + 1) * dd;
    p.elbo += o; p.ve += o; p.kl += o;
    p.workspace = (char*)s->dev.workspace + (size_t)ci * s->ws_chunk;
    p.workspace_bytes = s->ws_chunk;
    return p;
}

int vdp_cvi_session_site_update(vdp_cvi_session* s, const double* obs_y_host, double lr_data, double lr_girsanov,
                                int flags, double* elbo_host, double* ve_host, double* kl_host, double* fx_mus_host,
                                double* fx_covs_host, int64_t* status_host2) {
    VDP_REQUIRE(s != nullptr, "session");
    VDP_CHECK_CUDA(cudaSetDevice(s->device));
    const int64_t B = s->dev.B, N = s->dev.N, no = s->dev.n_obs;
    const size_t d = (size_t)s->dev.D, dd = d * d, S = sizeof(double);
    // from the second update on every chunk's workspace holds the naturals of its previous update: carried, with the
    // girsanov sites held in theta (vdp_cvi_session_download_state materialises them)
    int fl = flags & (VDP_FLAG_GH_QUADRATURE | VDP_FLAG_REFACTOR_PRECISION | VDP_FLAG_SEPARATE_KL_VALUE |
                      VDP_FLAG_NO_FUSED_FORWARD);
    if (s->updates > 0) fl |= VDP_FLAG_NATURALS_CARRIED | VDP_FLAG_SITES_IN_NATURALS;
    int rc = VDP_OK;
    int64_t ci = 0;
    for (int64_t b0 = 0; b0 < B && rc == VDP_OK; b0 += s->chunk, ++ci) {
        const int64_t nb = std::min<int64_t>(s->chunk, B - b0);
        cudaStream_t st = s->st[ci % s->nstreams];
        vdp_cvi_problem p = session_chunk(s, b0, nb, ci);
        p.lr_data = lr_data;
        p.lr_girsanov = lr_girsanov;
        p.flags = fl;
        cudaError_t e = cudaSuccess;
        if (obs_y_host && no > 0)
            e = h2d((void*)p.obs_y, obs_y_host + (size_t)b0 * no * d, (size_t)nb * no * d * S, st);
        if (e != cudaSuccess) { set_error("H2D copy: %s", cudaGetErrorString(e)); rc = VDP_E_CUDA; break; }
        if ((rc = vdp_cvi_site_update(&p, st))) break;
        const size_t o = (size_t)b0;
        if (e == cudaSuccess && elbo_host) e = d2h(elbo_host + o, p.elbo, nb * S, st);
        if (e == cudaSuccess && ve_host) e = d2h(ve_host + o, p.ve, nb * S, st);
        if (e == cudaSuccess && kl_host) e = d2h(kl_host + o, p.kl, nb * S, st);
        if (e == cudaSuccess && fx_mus_host) e = d2h(fx_mus_host + o * (N + 1) * d, p.fx_mus, nb * (N + 1) * d * S, st);
        if (e == cudaSuccess && fx_covs_host) e = d2h(fx_covs_host + o * (N + 1) * dd, p.fx_covs, nb * (N + 1) * dd * S, st);
        if (e != cudaSuccess) { set_error("D2H copy: %s", cudaGetErrorString(e)); rc = VDP_E_CUDA; }
    }
    for (int i = 0; i < s->nstreams; ++i)
        if (cudaStreamSynchronize(s->st[i]) != cudaSuccess && rc == VDP_OK) {
            set_error("stream synchronize: %s", cudaGetErrorString(cudaGetLastError()));
            rc = VDP_E_CUDA;
        }
    if (rc == VDP_OK) s->updates += 1;
    if (rc == VDP_OK && status_host2)
        VDP_CHECK_CUDA(cudaMemcpy(status_host2, s->dev.status, 2 * sizeof(int64_t), cudaMemcpyDeviceToHost));
    return rc;
}

int vdp_cvi_session_download_state(vdp_cvi_session* s, const vdp_cvi_problem* hp) {
    VDP_REQUIRE(s && hp && hp->B == s->dev.B && hp->N == s->dev.N && hp->D == s->dev.D && hp->n_obs == s->dev.n_obs,
                "session, problem of the session's shape (host pointers)");
    VDP_CHECK_CUDA(cudaSetDevice(s->device));
    const int64_t B = s->dev.B, N = s->dev.N, no = s->dev.n_obs;
    const size_t d = (size_t)s->dev.D, dd = d * d, b = (size_t)B, S = sizeof(double);
    cudaStream_t st = s->st[0];
    int rc = VDP_OK;
    if (s->updates > 1) {      // the sites live in theta: g = theta - nat(p) - scatter(data sites), chunk by chunk
        int64_t ci = 0;
        for (int64_t b0 = 0; b0 < B && rc == VDP_OK; b0 += s->chunk, ++ci) {
            vdp_cvi_problem p = session_chunk(s, b0, std::min<int64_t>(s->chunk, B - b0), ci);
            rc = vdp_cvi_materialize_sites(&p, st);
        }
        if (rc) return rc;
    }
    const vdp_cvi_problem& p = s->dev;
    cudaError_t e = cudaSuccess;
    auto down = [&](void* dst, const void* src, size_t bytes) { if (e == cudaSuccess && dst) e = d2h(dst, src, bytes, st); };
    down(hp->g1, p.g1, b * (N + 1) * d * S); down(hp->gd, p.gd, b * (N + 1) * dd * S); down(hp->gs, p.gs, b * N * dd * S);
    if (no > 0) { down(hp->d1, p.d1, b * no * d * S); down(hp->d2, p.d2, b * no * dd * S); }
    down(hp->fx_mus, p.fx_mus, b * (N + 1) * d * S); down(hp->fx_covs, p.fx_covs, b * (N + 1) * dd * S);
    down(hp->qA, p.qA, b * N * dd * S); down(hp->qb, p.qb, b * N * d * S); down(hp->qcholQ, p.qcholQ, b * N * dd * S);
    down(hp->qmu0, p.qmu0, b * d * S); down(hp->qcholP0, p.qcholP0, b * dd * S);
    down(hp->elbo, p.elbo, b * S); down(hp->ve, p.ve, b * S); down(hp->kl, p.kl, b * S);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { set_error("D2H copy: %s", cudaGetErrorString(e)); return VDP_E_CUDA; }
    return VDP_OK;
}

}  // extern "C"
