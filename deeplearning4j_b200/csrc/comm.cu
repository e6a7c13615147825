// Gradient exchange behind the C ABI: NCCL all-reduce of the flat [1,P] vectors on the caller's CUDA stream.
//
// This is what a GradientsAccumulator plugged into ParallelWrapper (TrainingMode.CUSTOM,
// deeplearning4j-scaleout/.../parallelism/ParallelWrapper.java:838-843) calls from storeUpdate / applyUpdate
// (nn/optimize/solvers/accumulation/GradientsAccumulator.java:27-41) instead of the reference's host-side queues of
// threshold-encoded messages (EncodedGradientsAccumulator.java:244-315, 440-520) or Nd4j.averageAndPropagate
// (ParallelWrapper.java:316-376). NCCL is bound at run time (dlopen of libnccl.so.2 -- in a PyTorch process that is
// the copy torch already loaded), so the library itself carries no link-time dependency and still loads on a box
// without NCCL; only dl4jb200_comm_* then fail, loudly.
#include <dlfcn.h>
#include <nccl.h>

#include "common.cuh"

namespace {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
};
NcclApi g_nccl;

template <typename F>
bool bind(F &fn, const char *name) {
    fn = reinterpret_cast<F>(dlsym(g_nccl.handle, name));
    return fn != nullptr;
}

int nccl_load() {
    if (g_nccl.handle) return DL4JB200_OK;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        dl4j_set_error("comm: NCCL is not available (dlopen libnccl.so.2: %s)", dlerror());
        return DL4JB200_NOT_SUPPORTED;
    }
    g_nccl.handle = h;
    if (!(bind(g_nccl.GetUniqueId, "ncclGetUniqueId") && bind(g_nccl.CommInitRank, "ncclCommInitRank") &&
          bind(g_nccl.CommDestroy, "ncclCommDestroy") && bind(g_nccl.AllReduce, "ncclAllReduce") &&
          bind(g_nccl.Broadcast, "ncclBroadcast") && bind(g_nccl.GroupStart, "ncclGroupStart") &&
          bind(g_nccl.GroupEnd, "ncclGroupEnd") && bind(g_nccl.GetErrorString, "ncclGetErrorString") &&
          bind(g_nccl.GetVersion, "ncclGetVersion"))) {
        dl4j_set_error("comm: libnccl lacks an expected symbol");
        g_nccl.handle = nullptr;
        return DL4JB200_NOT_SUPPORTED;
    }
    return DL4JB200_OK;
}

}  // namespace

struct dl4jb200_comm {
    ncclComm_t comm;
    dl4jb200_ctx *ctx;
    int nranks, rank;
};

#define DL4J_NCCL_TRY(expr)                                                                          \
    do {                                                                                             \
        ncclResult_t _r = (expr);                                                                    \
        if (_r != ncclSuccess) {                                                                     \
            dl4j_set_error("%s failed: %s (%s:%d)", #expr, g_nccl.GetErrorString(_r), __FILE__, __LINE__); \
            return DL4JB200_ERR_CUDA;                                                                \
        }                                                                                            \
    } while (0)

extern "C" {

int dl4jb200_comm_unique_id(void *id_out, size_t bytes) {
    DL4J_REQUIRE(id_out && bytes >= sizeof(ncclUniqueId), "comm_unique_id: need a buffer of at least %zu bytes", sizeof(ncclUniqueId));
    int rc = nccl_load();
    if (rc) return rc;
    ncclUniqueId id;
    DL4J_NCCL_TRY(g_nccl.GetUniqueId(&id));
    memcpy(id_out, &id, sizeof(id));
    return DL4JB200_OK;
}

int dl4jb200_comm_init(dl4jb200_ctx *ctx, int nranks, int rank, const void *id, dl4jb200_comm **out) {
    DL4J_REQUIRE(ctx && id && out, "comm_init: NULL argument");
    DL4J_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "comm_init: rank %d of %d", rank, nranks);
    int rc = nccl_load();
    if (rc) return rc;
    DL4J_CUDA_TRY(cudaSetDevice(ctx->device));
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    ncclComm_t c;
    DL4J_NCCL_TRY(g_nccl.CommInitRank(&c, nranks, uid, rank));
    *out = new dl4jb200_comm{c, ctx, nranks, rank};
    return DL4JB200_OK;
}

int dl4jb200_comm_destroy(dl4jb200_comm *comm) {
    if (!comm) return DL4JB200_OK;
    if (g_nccl.handle) g_nccl.CommDestroy(comm->comm);
    delete comm;
    return DL4JB200_OK;
}

int dl4jb200_comm_nranks(dl4jb200_comm *comm) { return comm ? comm->nranks : 0; }

int dl4jb200_comm_nccl_version(void) {
    int v = 0;
    if (nccl_load() == DL4JB200_OK) g_nccl.GetVersion(&v);
    return v;
}

static int allreduce(dl4jb200_comm *comm, float *buf, size_t n, ncclRedOp_t op, void *stream, const char *who) {
    DL4J_REQUIRE(comm && buf, "%s: NULL argument", who);
    if (n == 0) return DL4JB200_OK;
    cudaStream_t s = stream ? (cudaStream_t)stream : comm->ctx->stream;
    DL4J_NCCL_TRY(g_nccl.AllReduce(buf, buf, n, ncclFloat32, op, comm->comm, s));
    return DL4JB200_OK;
}

int dl4jb200_comm_allreduce_sum_f32(dl4jb200_comm *comm, float *buf, size_t n, void *stream) {
    return allreduce(comm, buf, n, ncclSum, stream, "comm_allreduce_sum_f32");
}

int dl4jb200_comm_allreduce_avg_f32(dl4jb200_comm *comm, float *buf, size_t n, void *stream) {
    return allreduce(comm, buf, n, ncclAvg, stream, "comm_allreduce_avg_f32");
}

// Several disjoint spans of one buffer in ONE NCCL group (the scattered BatchNormalization running statistics inside
// the flat parameter vector): offsets / counts in floats.
int dl4jb200_comm_allreduce_avg_spans_f32(dl4jb200_comm *comm, float *base, const long long *offsets, const long long *counts,
                                          int nspans, void *stream) {
    DL4J_REQUIRE(comm && base && (nspans == 0 || (offsets && counts)), "comm_allreduce_avg_spans_f32: NULL argument");
    if (nspans == 0) return DL4JB200_OK;
    cudaStream_t s = stream ? (cudaStream_t)stream : comm->ctx->stream;
    DL4J_NCCL_TRY(g_nccl.GroupStart());
    for (int i = 0; i < nspans; i++) {
        ncclResult_t r = g_nccl.AllReduce(base + offsets[i], base + offsets[i], (size_t)counts[i], ncclFloat32, ncclAvg, comm->comm, s);
        if (r != ncclSuccess) {
            g_nccl.GroupEnd();
            dl4j_set_error("comm_allreduce_avg_spans_f32: ncclAllReduce failed: %s", g_nccl.GetErrorString(r));
            return DL4JB200_ERR_CUDA;
        }
    }
    DL4J_NCCL_TRY(g_nccl.GroupEnd());
    return DL4JB200_OK;
}

int dl4jb200_comm_broadcast_f32(dl4jb200_comm *comm, float *buf, size_t n, int root, void *stream) {
    DL4J_REQUIRE(comm && buf && root >= 0 && root < comm->nranks, "comm_broadcast_f32: bad argument");
    if (n == 0) return DL4JB200_OK;
    cudaStream_t s = stream ? (cudaStream_t)stream : comm->ctx->stream;
    DL4J_NCCL_TRY(g_nccl.Broadcast(buf, buf, n, ncclFloat32, root, comm->comm, s));
    return DL4JB200_OK;
}

}  // extern "C"
