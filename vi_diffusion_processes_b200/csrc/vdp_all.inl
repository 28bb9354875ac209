// C ABI of libvdp_b200 (see include/vdp_b200.h).  Host-side launch logic only; kernels live in the
// .cuh files next to this one.  Compiled for sm_100a as several translation units (tu_*.cu), each of which defines
// one VDP_EMIT_* macro and includes this file: the macro selects which extern "C" entry points (and hence which
// kernel instantiations) the unit emits, so that the units build in parallel.
#ifndef VDP_EMIT_CORE
#define VDP_EMIT_CORE 0
#endif
#ifndef VDP_EMIT_CVI
#define VDP_EMIT_CVI 0
#endif
#ifndef VDP_EMIT_PRIOR
#define VDP_EMIT_PRIOR 0
#endif
#ifndef VDP_EMIT_VGP
#define VDP_EMIT_VGP 0
#endif
#ifndef VDP_EMIT_DIST
#define VDP_EMIT_DIST 0
#endif
#include <stdarg.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

#include "btd_kernels.cuh"
#include "common.cuh"
#include "kl_kernels.cuh"
#include "pointwise_kernels.cuh"
#include "prior_grad_kernels.cuh"
#include "quad_core.cuh"
#include "tile_sweeps.cuh"
#include "vgp_kernels.cuh"
#include "sim_kernels.cuh"

namespace vdp {

// library-wide state lives in the core translation unit (the library is compiled as several units, see tu_*.cu)
#if VDP_EMIT_CORE
thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}
#else
extern thread_local char g_error[512];
#endif

constexpr int kSeqBlock = 32;   // threads per block for one-thread-per-sequence sweeps (spread over SMs)
constexpr int kPtBlock = 128;   // threads per block for pointwise kernels

static inline double pi_factor(int D) { return std::pow(M_PI, -D * 0.5); }

// bump allocator over the caller's workspace
struct Arena {
    char* base;
    size_t size, used;
    Arena(void* p, size_t n) : base((char*)p), size(n), used(0) {}
    void* take_bytes(size_t bytes) {
        bytes = (bytes + 255) & ~size_t(255);
        if (!base || used + bytes > size) return nullptr;
        void* r = base + used;
        used += bytes;
        return r;
    }
    double* take(size_t doubles) {
        size_t bytes = (doubles * sizeof(double) + 255) & ~size_t(255);
        if (used + bytes > size) return nullptr;
        double* r = (double*)(base + used);
        used += bytes;
        return r;
    }
};
static inline size_t padded(size_t doubles) { return (doubles * sizeof(double) + 255) & ~size_t(255); }

// ---- optional per-phase device timing (vdp_profile_*): CUDA events recorded on the launching stream at
// phase boundaries of the fused pipelines; off by default, costs one cudaEventRecord per phase when on.
struct Profiler {
    bool on = false;
    std::vector<cudaEvent_t> pool;
    std::vector<int> phase;   // phase[i] = phase that starts at event i
    size_t used = 0;
};
#if VDP_EMIT_CORE
Profiler g_prof;

static int env_flag(const char* name) {
    const char* e = getenv(name);
    return (e && *e) ? atoi(e) : 0;
}
static Config config_from_env() {
    Config c;
    const char* e = getenv("VDP_SWEEPS");
    if (e && !strcmp(e, "seq")) c.sweeps = 1;
    if (e && !strcmp(e, "tiled")) c.sweeps = 2;
    c.nseg = env_flag("VDP_NSEG");
    c.gh_quadrature = env_flag("VDP_GH_QUADRATURE");
    c.refactor_precision = env_flag("VDP_REFACTOR_PRECISION");
    c.no_carried_naturals = env_flag("VDP_NO_CARRIED_NATURALS");
    c.no_fused_naturals = env_flag("VDP_NO_FUSED_NATURALS");
    c.separate_kl_value = env_flag("VDP_SEPARATE_KL_VALUE");
    c.no_fused_forward = env_flag("VDP_NO_FUSED_FORWARD");
    return c;
}
Config g_cfg = config_from_env();   // the environment is read here, once, at load time
static int* config_field(const char* key) {
    if (!key) return nullptr;
    if (!strncmp(key, "VDP_", 4)) key += 4;
    if (!strcmp(key, "SWEEPS")) return &g_cfg.sweeps;
    if (!strcmp(key, "NSEG")) return &g_cfg.nseg;
    if (!strcmp(key, "GH_QUADRATURE")) return &g_cfg.gh_quadrature;
    if (!strcmp(key, "REFACTOR_PRECISION")) return &g_cfg.refactor_precision;
    if (!strcmp(key, "NO_CARRIED_NATURALS")) return &g_cfg.no_carried_naturals;
    if (!strcmp(key, "NO_FUSED_NATURALS")) return &g_cfg.no_fused_naturals;
    if (!strcmp(key, "SEPARATE_KL_VALUE")) return &g_cfg.separate_kl_value;
    if (!strcmp(key, "NO_FUSED_FORWARD")) return &g_cfg.no_fused_forward;
    return nullptr;
}
#else
extern Profiler g_prof;
#endif
static inline void prof_mark(int phase, cudaStream_t st) {
    if (!g_prof.on) return;
    if (g_prof.used == g_prof.pool.size()) {
        if (g_prof.pool.size() >= 65536) return;
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        g_prof.pool.push_back(e);
        g_prof.phase.push_back(0);
    }
    g_prof.phase[g_prof.used] = phase;
    cudaEventRecord(g_prof.pool[g_prof.used], st);
    ++g_prof.used;
}

}  // namespace vdp

#include "vdp_sweeps.inl"
#include "fwd_fused.cuh"

using namespace vdp;

#if VDP_EMIT_CORE
extern "C" {

int vdp_version(void) { return 200; }

int vdp_config_set(const char* key, int value) {
    int* f = config_field(key);
    if (!f) { set_error("unknown configuration key %s", key ? key : "(null)"); return VDP_E_BADARG; }
    *f = value;
    return VDP_OK;
}
int vdp_config_get(const char* key, int* value) {
    int* f = config_field(key);
    if (!f || !value) { set_error("unknown configuration key %s", key ? key : "(null)"); return VDP_E_BADARG; }
    *value = *f;
    return VDP_OK;
}
const char* vdp_last_error(void) { return g_error; }

int64_t vdp_read_status(const int64_t* status, int64_t* out_host2, void* stream) {
    int64_t h[2] = {0, 0};
    if (status) {
        if (cudaStreamSynchronize(as_stream(stream)) != cudaSuccess) return VDP_E_CUDA;
        if (cudaMemcpy(h, status, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) return VDP_E_CUDA;
    }
    if (out_host2) { out_host2[0] = h[0]; out_host2[1] = h[1]; }
    return h[0];
}

int vdp_profile_enable(int on) {
    g_prof.on = on != 0;
    g_prof.used = 0;
    return VDP_OK;
}

int vdp_profile_read(double* ms_by_phase, int64_t* spans_by_phase, int n_phases) {
    for (int i = 0; i < n_phases; ++i) { ms_by_phase[i] = 0.0; spans_by_phase[i] = 0; }
    if (g_prof.used == 0) return VDP_OK;
    VDP_CHECK_CUDA(cudaEventSynchronize(g_prof.pool[g_prof.used - 1]));
    for (size_t i = 0; i + 1 < g_prof.used; ++i) {
        const int ph = g_prof.phase[i];
        if (ph < 0 || ph >= n_phases) continue;   // VDP_PHASE_END marks close a span, they do not open one
        float ms = 0.f;
        VDP_CHECK_CUDA(cudaEventElapsedTime(&ms, g_prof.pool[i], g_prof.pool[i + 1]));
        ms_by_phase[ph] += ms;
        spans_by_phase[ph] += 1;
    }
    g_prof.used = 0;
    return VDP_OK;
}

// diagnostic: the fast float64 primitives of fastmath.cuh, element-wise (tests measure their error in ulps)
__global__ void k_test_fastmath(const double* __restrict__ x, const double* __restrict__ y, double* __restrict__ rcp,
                                double* __restrict__ sq, double* __restrict__ rsq, double* __restrict__ quo,
                                int64_t n) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    double s, rs;
    sqrt_rsqrt_fast(x[i], s, rs);
    rcp[i] = rcp_fast(x[i]);
    sq[i] = s;
    rsq[i] = rs;
    quo[i] = div_fast(y[i], x[i]);
}

int vdp_test_fastmath(const double* x, const double* y, double* rcp, double* sq, double* rsq, double* quo, int64_t n,
                      void* stream) {
    VDP_REQUIRE(x && y && rcp && sq && rsq && quo && n >= 0, "pointers, n");
    if (n == 0) return VDP_OK;
    k_test_fastmath<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(x, y, rcp, sq, rsq, quo, n);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// diagnostic: float64 FMA throughput of the device (the roofline denominator of the quadrature kernels): every thread
// runs `iters` rounds of 16 independent DFMA chains; 2 * 16 * iters * threads flops in all
__global__ void k_test_fp64_fma(double* __restrict__ out, int64_t iters, double a, double b) {
    double x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = (double)(threadIdx.x + j);
    for (int64_t i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = fma(x[j], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += x[j];
    out[blockIdx.x * (int64_t)blockDim.x + threadIdx.x] = s;
}

int vdp_test_fp64_fma(double* out, int64_t blocks, int threads, int64_t iters, void* stream) {
    VDP_REQUIRE(out && blocks > 0 && threads > 0 && threads <= 1024 && iters >= 0, "out, blocks, threads, iters");
    k_test_fp64_fma<<<(unsigned)blocks, threads, 0, as_stream(stream)>>>(out, iters, 0.999999, 1e-9);
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

// ---------------------------------------------------------------------------------------------------
// op level
// ---------------------------------------------------------------------------------------------------

int vdp_btd_cholesky(const double* Kd, const double* Ks, double* Ld, double* Ls, int64_t B, int64_t n, int D,
                     int64_t* status, void* stream) {
    VDP_REQUIRE(Kd && Ld && B >= 0 && n >= 1, "Kd, Ld, B, n");
    VDP_REQUIRE((Ks == nullptr) == (Ls == nullptr), "Ks and Ls must both be given or both be NULL");
    if (B == 0) return VDP_OK;
    if (Ks && n >= 2 && use_tiled_sweeps(B, n - 1)) {
        VDP_DISPATCH_D(D, { int rc = run_chol_tiled<kD>(Kd, Ks, 1.0, 1.0, Ld, Ls, B, n - 1, status,
                                                        SegScratch(nullptr, 0), as_stream(stream)); if (rc) return rc; });
        return VDP_OK;
    }
    VDP_DISPATCH_D(D, (k_chol_fwd<kD><<<grid_for(B, kSeqBlock), kSeqBlock, 0, as_stream(stream)>>>(
                          Kd, Ks, 1.0, 1.0, Ld, Ls, B, n, status)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_btd_inverse_band(const double* Ld, const double* Ls, double* Sd, double* Ss, int64_t B, int64_t n, int D,
                         void* stream) {
    VDP_REQUIRE(Ld && Sd && B >= 0 && n >= 1, "Ld, Sd, B, n");
    if (B == 0) return VDP_OK;
    if (Ls && n >= 2 && use_tiled_sweeps(B, n - 1)) {
        VDP_DISPATCH_D(D, { int rc = run_invband_tiled<kD>(Ld, Ls, Sd, Ss, B, n - 1, SegScratch(nullptr, 0),
                                                           as_stream(stream)); if (rc) return rc; });
        return VDP_OK;
    }
    VDP_DISPATCH_D(D, (k_invband_bwd<kD><<<grid_for(B, kSeqBlock), kSeqBlock, 0, as_stream(stream)>>>(
                          Ld, Ls, Sd, Ls ? Ss : nullptr, B, n)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_btd_inverse_band_jvp(const double* Ld, const double* Ls, const double* dKd, const double* dKs, double* dSd,
                             double* dSs, int64_t B, int64_t n, int D, void* stream) {
    VDP_REQUIRE(Ld && Ls && dKd && dKs && dSd && dSs && B >= 0 && n >= 2, "pointers, B, n >= 2");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_invband_jvp<kD><<<grid_for(B, kSeqBlock), kSeqBlock, 0, as_stream(stream)>>>(
                          Ld, Ls, dKd, dKs, dSd, dSs, B, n)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_btd_solve(const double* Ld, const double* Ls, const double* rhs, double* x, int transpose, int64_t B,
                  int64_t n, int D, void* stream) {
    VDP_REQUIRE(Ld && rhs && x && B >= 0 && n >= 1, "Ld, rhs, x, B, n");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_btd_solve<kD><<<grid_for(B, kSeqBlock), kSeqBlock, 0, as_stream(stream)>>>(
                          Ld, Ls, rhs, x, transpose, B, n)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_btd_matvec(const double* Md, const double* Ms, const double* x, double* y, int symmetric, int transpose,
                   int64_t B, int64_t n, int D, void* stream) {
    VDP_REQUIRE(Md && x && y && B >= 0 && n >= 1, "Md, x, y, B, n");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_btd_matvec<kD><<<grid_for(B * n, kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                          Md, Ms, x, y, symmetric, transpose, B, n)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_btd_logdet(const double* Ld, double* out, int64_t B, int64_t n, int D, void* stream) {
    VDP_REQUIRE(Ld && out && B >= 0 && n >= 1, "Ld, out, B, n");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_btd_logdet<kD><<<grid_for(B, kSeqBlock), kSeqBlock, 0, as_stream(stream)>>>(Ld, out, B, n)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_btd_block_to_band(const double* Md, const double* Ms, double* band, int64_t B, int64_t n, int D,
                          void* stream) {
    VDP_REQUIRE(Md && band && B >= 0 && n >= 1, "Md, band, B, n");
    if (B == 0) return VDP_OK;
    const int64_t total = B * (Ms ? 2 * D : D) * n * D;
    VDP_DISPATCH_D(D, (k_block_to_band<kD><<<grid_for(total, kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                          Md, Ms, band, B, n)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_btd_band_to_block(const double* band, double* Md, double* Ms, int symmetric, int64_t B, int64_t n, int D,
                          void* stream) {
    VDP_REQUIRE(band && Md && B >= 0 && n >= 1, "band, Md, B, n");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_band_to_block<kD><<<grid_for(B * n, kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                          band, Md, Ms, symmetric, B, n)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

int vdp_btd_conditional_precisions(const double* As, const double* Kd, const double* Ks, double* out, int64_t B,
                                   int64_t n, int D, void* stream) {
    VDP_REQUIRE(Kd && out && B >= 0 && n >= 1 && (n == 1 || (As && Ks)), "As, Kd, Ks, out, B, n");
    if (B == 0) return VDP_OK;
    VDP_DISPATCH_D(D, (k_conditional_precisions<kD><<<grid_for(B * n, kPtBlock), kPtBlock, 0, as_stream(stream)>>>(
                          As, Kd, Ks, out, B, n)));
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}

}  // extern "C"
#endif  // VDP_EMIT_CORE

#include "vdp_fused.inl"

#if VDP_EMIT_CORE
extern "C" int vdp_euler_maruyama(int drift_id, const double* params_host, const double* q_host, const double* x0,
                                  const double* time_grid, const double* noise, unsigned long long seed, double* out,
                                  int64_t B, int64_t n, int D, int64_t* status, void* stream) {
    VDP_REQUIRE(q_host && x0 && time_grid && out && B >= 0 && n >= 1, "pointers, B, n");
    if (B == 0) return VDP_OK;
    const DriftParams P = make_params(params_host);
    VDP_DISPATCH_D(D, VDP_DISPATCH_DRIFT(drift_id, {
        bool ok = true;
        const Mat<kD> cq = chol(host_mat<kD>(q_host), ok);
        if (!ok) { set_error("q is not positive definite"); return VDP_E_BADARG; }
        k_euler_maruyama<kDrift, kD><<<grid_for(B, 64), 64, 0, as_stream(stream)>>>(P, cq, x0, time_grid, noise, seed,
                                                                                 out, B, n);
    }));
    (void)status;
    VDP_CHECK_LAUNCH();
    return VDP_OK;
}
#endif  // VDP_EMIT_CORE

#include "vdp_cvi.inl"
#if VDP_EMIT_PRIOR
#include "vdp_prior.inl"
#endif
#if VDP_EMIT_CVI
#include "vdp_host.inl"
#include "vdp_trainer.inl"
#endif
#if VDP_EMIT_VGP
#include "vdp_vgp.inl"
#endif
#if VDP_EMIT_DIST
#include "vdp_dist.inl"
#include "vdp_noalloc.inl"
#include "vdp_ssm_density.inl"
#include "vdp_comm.inl"
#endif
