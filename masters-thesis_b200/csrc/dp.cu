// Data-parallel gradient exchange of the training step: NCCL called from inside the library.
//
// Replaces what Lightning's implicit DDP does to the reference when it sees several GPUs
// (train.py:169 strategy='auto' -> torch DistributedDataParallel -> Reducer -> ncclAllReduce per bucket): a
// sum all-reduce of a slice of the flat gradient vector, enqueued on the caller's (communication) stream.
// The kernels pre-scale gradients by 1/world_size, so the sum is the DDP mean.
//
// libnccl is bound at run time (dlopen): inside a PyTorch process the already loaded libnccl.so.2 is reused, so
// the process never holds two copies of NCCL; a process that never calls b200_dp_* needs no NCCL at all.  Only the
// stable core of the NCCL API is used (unique id, comm init/destroy, all-reduce, error string).
#include <dlfcn.h>

#include <mutex>

#include "common.cuh"

namespace b200 {

// the part of nccl.h this file needs (NCCL 2.x ABI)
struct NcclUniqueId { char internal[128]; };
typedef struct ncclComm* NcclComm;
constexpr int kNcclSuccess = 0, kNcclFloat32 = 7, kNcclSum = 0;

struct NcclApi {
    int (*GetUniqueId)(NcclUniqueId*);
    int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int);
    int (*CommDestroy)(NcclComm);
    int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
    const char* (*GetErrorString)(int);
    int (*GetVersion)(int*);
    bool ok;
};

static NcclApi g_nccl = {};
static std::once_flag g_nccl_once;

static void nccl_bind() {
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);       // the copy the host framework already loaded
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return;
    auto sym = [&](const char* n) { return dlsym(h, n); };
    g_nccl.GetUniqueId = (int (*)(NcclUniqueId*))sym("ncclGetUniqueId");
    g_nccl.CommInitRank = (int (*)(NcclComm*, int, NcclUniqueId, int))sym("ncclCommInitRank");
    g_nccl.CommDestroy = (int (*)(NcclComm))sym("ncclCommDestroy");
    g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t))sym("ncclAllReduce");
    g_nccl.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
    g_nccl.GetVersion = (int (*)(int*))sym("ncclGetVersion");
    g_nccl.ok = g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.CommDestroy && g_nccl.AllReduce && g_nccl.GetErrorString;
}

static int nccl_ready(const char* who) {
    std::call_once(g_nccl_once, nccl_bind);
    if (!g_nccl.ok) { set_error("%s: libnccl.so.2 could not be loaded (%s)", who, dlerror() ? dlerror() : "missing symbols"); return B200_ERR_UNSUPPORTED; }
    return B200_OK;
}

static int nccl_fail(int rc, const char* what) {
    set_error("NCCL error %d (%s) in %s", rc, g_nccl.GetErrorString(rc), what);
    return -2000 - rc;
}

struct DpComm {
    NcclComm comm;
    int world, rank, device;
};

}  // namespace b200

using namespace b200;

extern "C" int b200_dp_nccl_version(void) {
    if (nccl_ready("dp_nccl_version")) return 0;
    int v = 0;
    if (g_nccl.GetVersion) g_nccl.GetVersion(&v);
    return v;
}

extern "C" int b200_dp_unique_id(void* id_out_host) {
    B200_CHECK_ARG(id_out_host != nullptr, "dp_unique_id: null pointer");
    int rc = nccl_ready("dp_unique_id");
    if (rc) return rc;
    NcclUniqueId id;
    const int n = g_nccl.GetUniqueId(&id);
    if (n != kNcclSuccess) return nccl_fail(n, "ncclGetUniqueId");
    memcpy(id_out_host, &id, sizeof(id));
    return B200_OK;
}

extern "C" int b200_dp_init(const void* id_host, int world_size, int rank, void** comm_out) {
    B200_CHECK_ARG(id_host && comm_out, "dp_init: null pointer");
    B200_CHECK_ARG(world_size >= 1 && rank >= 0 && rank < world_size, "dp_init: bad rank %d of %d", rank, world_size);
    int rc = nccl_ready("dp_init");
    if (rc) return rc;
    NcclUniqueId id;
    memcpy(&id, id_host, sizeof(id));
    DpComm* c = new DpComm{nullptr, world_size, rank, 0};
    cudaError_t ce = cudaGetDevice(&c->device);
    if (ce != cudaSuccess) { delete c; return cuda_fail(ce, "cudaGetDevice"); }
    const int n = g_nccl.CommInitRank(&c->comm, world_size, id, rank);
    if (n != kNcclSuccess) { delete c; return nccl_fail(n, "ncclCommInitRank"); }
    *comm_out = c;
    return B200_OK;
}

extern "C" int b200_dp_allreduce(void* comm, float* buf, int64_t count, void* stream) {
    B200_CHECK_ARG(comm && buf && count > 0, "dp_allreduce: bad arguments");
    DpComm* c = (DpComm*)comm;
    TimedLaunch _t("ncclAllReduce", (cudaStream_t)stream);
    const int n = g_nccl.AllReduce(buf, buf, (size_t)count, kNcclFloat32, kNcclSum, c->comm, (cudaStream_t)stream);
    if (n != kNcclSuccess) return nccl_fail(n, "ncclAllReduce");
    return B200_OK;
}

extern "C" int b200_dp_destroy(void* comm) {
    if (!comm) return B200_OK;
    DpComm* c = (DpComm*)comm;
    const int n = g_nccl.CommDestroy(c->comm);
    delete c;
    if (n != kNcclSuccess) return nccl_fail(n, "ncclCommDestroy");
    return B200_OK;
}
