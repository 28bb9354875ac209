// Shared host/device plumbing for the vdp_b200 library: error reporting, the device status word,
// dispatch over the state dimension D and the drift id.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/vdp_b200.h"
#include "drifts.cuh"
#include "smallmat.cuh"

namespace vdp {

void set_error(const char* fmt, ...);

// Process-wide tuning / route switches (vdp_config_set).  Defaults come from the environment ONCE, when the library is
// loaded (VDP_SWEEPS=seq|tiled, VDP_NSEG, VDP_GH_QUADRATURE, ... -- test and tuning hooks); no entry point reads the
// environment afterwards.  The numerics routes can also be chosen per call through the problem's flags
// (VDP_FLAG_GH_QUADRATURE, ...), which is what a caller that needs re-entrancy per stream uses.
struct Config {
    int sweeps = 0;                 // 0 = heuristic, 1 = thread per sequence, 2 = parallel-in-time tiles
    int nseg = 0;                   // > 0: time segments per sequence of the tiled sweeps
    int gh_quadrature = 0;          // 20-node Gauss-Hermite loops instead of the closed-form moments (cubic drifts)
    int refactor_precision = 0;     // marginals through a second factorisation, as the reference (four sweeps)
    int no_carried_naturals = 0;    // ignore VDP_FLAG_NATURALS_CARRIED
    int no_fused_naturals = 0;      // rebuild the naturals from the sites after the girsanov update
    int no_fused_forward = 0;       // mean sweep, KL gradient and the next factorisation as three kernels instead of one
    int separate_kl_value = 0;      // KL terms of classic_elbo in their own pass instead of inside the mean sweep
};
extern Config g_cfg;

#define VDP_CHECK_CUDA(expr)                                                                   \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            ::vdp::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,             \
                             cudaGetErrorString(_e));                                          \
            return VDP_E_CUDA;                                                                 \
        }                                                                                      \
    } while (0)

#define VDP_CHECK_LAUNCH()  VDP_CHECK_CUDA(cudaGetLastError())

#define VDP_REQUIRE(cond, msg)                                                                 \
    do {                                                                                       \
        if (!(cond)) {                                                                         \
            ::vdp::set_error("bad argument: %s (%s)", msg, #cond);                            \
            return VDP_E_BADARG;                                                               \
        }                                                                                      \
    } while (0)

// first failing block wins; status[0] = code, status[1] = flat index
__device__ __forceinline__ void report_not_pd(int64_t* status, int64_t flat_index) {
    if (status != nullptr) {
        if (atomicCAS(reinterpret_cast<unsigned long long*>(status), 0ULL,
                      static_cast<unsigned long long>(VDP_STATUS_NOT_PD)) == 0ULL) {
            status[1] = flat_index;
        }
    }
}

// flat node index -> (sequence, grid point): 32-bit division when the index fits (a 64-bit division by a run-time
// divisor is ~40 instructions, as much as the closed-form quadrature of a grid point)
__device__ __forceinline__ void split_index(int64_t i, int64_t n, int64_t& b, int64_t& k) {
    if ((uint64_t)i < 0x80000000ull && (uint64_t)n < 0x80000000ull) {
        const uint32_t bb = (uint32_t)i / (uint32_t)n;
        b = bb;
        k = (uint32_t)i - bb * (uint32_t)n;
    } else {
        b = i / n;
        k = i - b * n;
    }
}

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

inline unsigned grid_for(int64_t work, int block) { return (unsigned)((work + block - 1) / block); }

template <int V> struct IntTag { static constexpr int value = V; };

// Run `body` with `constexpr int kD` = D for D in 1..3.  The body sits in a generic lambda so that
// `if constexpr` on (drift, D) validity really discards the invalid instantiations; a `return x;`
// inside the body leaves the lambda, and a non-zero value is propagated to the caller.
#define VDP_D_CASE(V, ...)                                                                     \
    case V: {                                                                                  \
        auto _vdp_f = [&](auto _vdp_dt) -> int {                                               \
            constexpr int kD = decltype(_vdp_dt)::value;                                       \
            (void)kD;                                                                          \
            __VA_ARGS__;                                                                       \
            return VDP_OK;                                                                     \
        };                                                                                     \
        const int _vdp_rc = _vdp_f(::vdp::IntTag<V>{});                                        \
        if (_vdp_rc) return _vdp_rc;                                                           \
    } break;

#define VDP_DISPATCH_D(D, ...)                                                                 \
    switch (D) {                                                                               \
        VDP_D_CASE(1, __VA_ARGS__)                                                             \
        VDP_D_CASE(2, __VA_ARGS__)                                                             \
        VDP_D_CASE(3, __VA_ARGS__)                                                             \
        default: ::vdp::set_error("state dimension %d not supported (1..3)", D); return VDP_E_BADARG; \
    }

// inside VDP_DISPATCH_D: run `body` with `constexpr int kDrift` for every (drift, D) pair that exists
#define VDP_DRIFT_CASE(ID, ...)                                                                \
    case ID: {                                                                                 \
        auto _vdp_g = [&](auto _vdp_it) -> int {                                               \
            constexpr int kDrift = decltype(_vdp_it)::value;                                   \
            if constexpr (::vdp::DriftValid<kDrift, kD>::value) {                              \
                __VA_ARGS__;                                                                   \
                return VDP_OK;                                                                 \
            } else {                                                                           \
                ::vdp::set_error("drift %d is not defined for D=%d", kDrift, kD);              \
                return VDP_E_BADARG;                                                           \
            }                                                                                  \
        };                                                                                     \
        const int _vdp_rc2 = _vdp_g(::vdp::IntTag<ID>{});                                      \
        if (_vdp_rc2) return _vdp_rc2;                                                         \
    } break;

#define VDP_DISPATCH_DRIFT(id, ...)                                                            \
    switch (id) {                                                                              \
        VDP_DRIFT_CASE(::vdp::DRIFT_OU, __VA_ARGS__)                                           \
        VDP_DRIFT_CASE(::vdp::DRIFT_DW, __VA_ARGS__)                                           \
        VDP_DRIFT_CASE(::vdp::DRIFT_BENES, __VA_ARGS__)                                        \
        VDP_DRIFT_CASE(::vdp::DRIFT_SINE, __VA_ARGS__)                                         \
        VDP_DRIFT_CASE(::vdp::DRIFT_SQRT, __VA_ARGS__)                                         \
        VDP_DRIFT_CASE(::vdp::DRIFT_VDP, __VA_ARGS__)                                          \
        VDP_DRIFT_CASE(::vdp::DRIFT_MLP131, __VA_ARGS__)                                       \
        VDP_DRIFT_CASE(::vdp::DRIFT_LORENZ63, __VA_ARGS__)                                     \
        VDP_DRIFT_CASE(::vdp::DRIFT_OU_ND, __VA_ARGS__)                                        \
        default: ::vdp::set_error("unknown drift id %d", id); return VDP_E_BADARG;             \
    }

inline DriftParams make_params(const double* host) {
    DriftParams P;
    for (int i = 0; i < VDP_MAX_DRIFT_PARAMS; ++i) P.p[i] = host ? host[i] : 0.0;
    return P;
}

template <int D> struct HostMat { Mat<D> m; };
template <int D> inline Mat<D> host_mat(const double* p) {
    Mat<D> m;
    for (int i = 0; i < D * D; ++i) m.a[i] = p[i];
    return m;
}
template <int D> inline Vec<D> host_vec(const double* p) {
    Vec<D> v;
    for (int i = 0; i < D; ++i) v.v[i] = p[i];
    return v;
}

}  // namespace vdp
