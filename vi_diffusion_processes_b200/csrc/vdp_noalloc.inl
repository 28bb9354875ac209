// Allocation-free variant of vdp_ssm_marginals (included at the end of vdp_all.inl).
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch the sweeps need
// (Cholesky factor of the rebuilt precision + segment carries) lives in a file-scope __device__ array of fixed
// size instead of a caller-provided workspace.  Every GPU that loads the library holds the array until exit.
//
// In host code the array's name is only a stand-in for the symbol, not a device address: the address a kernel
// can use comes from cudaGetSymbolAddress, asked once per device with that device current.
#include <mutex>

namespace vdp {

constexpr size_t kStaticSegBytes = kSegScratchBytes;                    // segment carries (as every fused workspace)
constexpr size_t kStaticFactorBytes = VDP_STATIC_FACTOR_BYTES;          // Ld + Ls of the rebuilt precision
constexpr size_t kStaticScratchBytes = kStaticSegBytes + kStaticFactorBytes;
static_assert(kStaticScratchBytes == VDP_STATIC_SCRATCH_BYTES, "header and library disagree on the scratch size");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
static void* g_static_scratch_ptr[kMaxDevices] = {};
static std::mutex g_static_scratch_mutex;

// device address of g_static_scratch on the CURRENT device (resolved once per device)
static int static_scratch_address(void** out) {
    int dev = -1;
    VDP_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { set_error("device index %d out of range", dev); return VDP_E_BADARG; }
    std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
    if (!g_static_scratch_ptr[dev]) {
        void* p = nullptr;
        VDP_CHECK_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) { set_error("cudaGetSymbolAddress returned a null address"); return VDP_E_CUDA; }
        g_static_scratch_ptr[dev] = p;
    }
    *out = g_static_scratch_ptr[dev];
    return VDP_OK;
}

}  // namespace vdp

extern "C" {

size_t vdp_static_scratch_bytes(void) { return kStaticScratchBytes; }

int vdp_ssm_marginals_static(const double* A, const double* b, const double* cholQ, const double* mu0,
                             const double* cholP0, double* m, double* Sd, int64_t B, int64_t N, int D,
                             int64_t* status, void* stream) {
    VDP_REQUIRE(A && b && cholQ && mu0 && cholP0 && m && Sd && B >= 0 && N >= 1, "pointers, B, N");
    VDP_REQUIRE(D >= 1 && D <= 3, "D in 1..3");
    if (B == 0) return VDP_OK;
    const size_t need = padded((size_t)B * (N + 1) * D * D) + padded((size_t)B * N * D * D);
    if (need > kStaticFactorBytes) {
        set_error("vdp_ssm_marginals_static: B=%lld N=%lld D=%d needs %zu bytes of factor scratch, the static "
                  "array holds %zu (limit: 8 (2 N + 1) B D^2 <= %zu); use vdp_ssm_marginals with a workspace",
                  (long long)B, (long long)N, D, need, kStaticFactorBytes, kStaticFactorBytes);
        return VDP_E_WORKSPACE;
    }
    void* base = nullptr;
    int rc = static_scratch_address(&base);
    if (rc) return rc;
    // same layout as the caller-provided workspace of vdp_ssm_marginals: Ld, Ls, segment scratch.  Every word of
    // it that a kernel reads has been written earlier in the same call (the factor by the forward sweep, the
    // segment elements by the SUMMARY launches), so nothing has to start at zero and the previous call's contents
    // are never seen.
    Arena ar(base, kStaticScratchBytes);
    double* Ld = ar.take((size_t)B * (N + 1) * D * D);
    double* Ls = ar.take((size_t)B * N * D * D);
    if (!Ld || !Ls) { set_error("static scratch too small"); return VDP_E_WORKSPACE; }
    SegScratch sc((char*)base + kStaticFactorBytes, kStaticSegBytes);
    VDP_DISPATCH_D(D, { int rc2 = run_marginals<kD>(A, b, cholQ, mu0, cholP0, m, Sd, Ld, Ls, B, N, status,
                                                    as_stream(stream), sc); if (rc2) return rc2; });
    return VDP_OK;
}

}  // extern "C"
