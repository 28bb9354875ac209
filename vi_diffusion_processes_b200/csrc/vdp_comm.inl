// Rank-to-rank exchange of the time-sharded sweeps (BASELINE config 5) issued from C on the compute stream:
// a thin communicator over NCCL (NVLink 5 / NVSwitch on one box).  What crosses ranks is tiny -- one scan element
// and one initial state per rank and sweep, a few D x D halo blocks, four doubles of ELBO terms -- so the exchange is
// an all-gather of a few hundred bytes per rank: latency-bound, no fused compute + collective kernel is warranted,
// but every exchange is enqueued on the caller's stream between the kernels it separates, so a whole site update is
// one stream-ordered sequence with no host synchronisation.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: inside a PyTorch process that is the library torch already
// loaded), so the shared object has no link-time dependency on it and loads on boxes without NCCL.
// Included by vdp_all.inl (translation unit "dist").
#include <dlfcn.h>

struct vdp_comm {
    void* nccl = nullptr;   // ncclComm_t
    int rank = 0, nranks = 1, device = 0;
};

namespace vdp {

struct NcclId { char internal[128]; };
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;

static int nccl_load() {
    if (g_nccl.handle) return VDP_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { set_error("NCCL not found (dlopen libnccl.so.2): %s", dlerror()); return VDP_E_CUDA; }
    g_nccl.GetUniqueId = (int (*)(NcclId*))dlsym(h, "ncclGetUniqueId");
    g_nccl.CommInitRank = (int (*)(void**, int, NcclId, int))dlsym(h, "ncclCommInitRank");
    g_nccl.AllGather = (int (*)(const void*, void*, size_t, int, void*, cudaStream_t))dlsym(h, "ncclAllGather");
    g_nccl.CommDestroy = (int (*)(void*))dlsym(h, "ncclCommDestroy");
    g_nccl.GetErrorString = (const char* (*)(int))dlsym(h, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllGather || !g_nccl.CommDestroy) {
        set_error("libnccl.so.2 lacks an expected symbol");
        return VDP_E_CUDA;
    }
    g_nccl.handle = h;
    return VDP_OK;
}

#define VDP_CHECK_NCCL(expr)                                                                              \
    do {                                                                                                  \
        const int _r = (expr);                                                                            \
        if (_r != 0) {                                                                                    \
            ::vdp::set_error("%s failed: %s", #expr,                                                      \
                             ::vdp::g_nccl.GetErrorString ? ::vdp::g_nccl.GetErrorString(_r) : "NCCL error"); \
            return VDP_E_CUDA;                                                                            \
        }                                                                                                 \
    } while (0)

}  // namespace vdp

extern "C" {

int vdp_comm_unique_id(void* id128) {
    VDP_REQUIRE(id128 != nullptr, "id128");
    int rc = nccl_load();
    if (rc) return rc;
    VDP_CHECK_NCCL(g_nccl.GetUniqueId(reinterpret_cast<NcclId*>(id128)));
    return VDP_OK;
}

int vdp_comm_create(vdp_comm** out, const void* id128, int rank, int nranks) {
    VDP_REQUIRE(out && id128 && nranks >= 1 && rank >= 0 && rank < nranks, "out, id128, rank, nranks");
    int rc = nccl_load();
    if (rc) return rc;
    vdp_comm* c = new vdp_comm();
    c->rank = rank;
    c->nranks = nranks;
    VDP_CHECK_CUDA(cudaGetDevice(&c->device));
    NcclId id;
    memcpy(&id, id128, sizeof(id));
    const int r = g_nccl.CommInitRank(&c->nccl, nranks, id, rank);
    if (r != 0) {
        set_error("ncclCommInitRank failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "NCCL error");
        delete c;
        return VDP_E_CUDA;
    }
    *out = c;
    return VDP_OK;
}

int vdp_comm_all_gather(vdp_comm* c, const void* send, void* recv, size_t bytes_per_rank, void* stream) {
    VDP_REQUIRE(c && c->nccl && send && recv, "comm, send, recv");
    if (bytes_per_rank == 0) return VDP_OK;
    VDP_CHECK_NCCL(g_nccl.AllGather(send, recv, bytes_per_rank, /*ncclChar*/ 0, c->nccl, as_stream(stream)));
    return VDP_OK;
}

int vdp_comm_destroy(vdp_comm* c) {
    if (!c) return VDP_OK;
    if (c->nccl && g_nccl.CommDestroy) g_nccl.CommDestroy(c->nccl);
    delete c;
    return VDP_OK;
}

}  // extern "C"
