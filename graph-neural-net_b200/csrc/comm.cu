// comm.cu -- NCCL gradient all-reduce for the data-parallel path.  The reference is serial
// (README.md:13); its batch sum (src/SoftmaxCrossEntropyNeuralNet.java:305-322) becomes a sum
// over GPUs here.  NCCL symbols are resolved with dlopen/dlsym (see comm.hpp).
#include "comm.hpp"

#include <dlfcn.h>

namespace b2nn {

namespace {
struct NcclUniqueId { char internal[128]; };
typedef void *NcclComm;
typedef int (*PFN_GetUniqueId)(NcclUniqueId *);
typedef int (*PFN_CommInitRank)(NcclComm *, int, NcclUniqueId, int);
typedef int (*PFN_CommDestroy)(NcclComm);
typedef int (*PFN_AllReduce)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t);
typedef const char *(*PFN_GetErrorString)(int);

struct NcclApi {
    void *handle = nullptr;
    PFN_GetUniqueId GetUniqueId = nullptr;
    PFN_CommInitRank CommInitRank = nullptr;
    PFN_CommDestroy CommDestroy = nullptr;
    PFN_AllReduce AllReduce = nullptr;
    PFN_GetErrorString GetErrorString = nullptr;
};

NcclApi *nccl() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.handle ? &api : nullptr;
    tried = true;
    // prefer a copy that is already mapped (torch's bundled NCCL), then the usual sonames
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *n : names) {
        h = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) {
        const char *env = getenv("B2NN_NCCL_LIB");
        if (env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
    }
    for (const char *n : names) {
        if (h) break;
        h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    }
    if (!h) return nullptr;
    api.handle = h;
    api.GetUniqueId = (PFN_GetUniqueId)dlsym(h, "ncclGetUniqueId");
    api.CommInitRank = (PFN_CommInitRank)dlsym(h, "ncclCommInitRank");
    api.CommDestroy = (PFN_CommDestroy)dlsym(h, "ncclCommDestroy");
    api.AllReduce = (PFN_AllReduce)dlsym(h, "ncclAllReduce");
    api.GetErrorString = (PFN_GetErrorString)dlsym(h, "ncclGetErrorString");
    if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllReduce) {
        api.handle = nullptr;
        return nullptr;
    }
    return &api;
}

int nccl_fail(NcclApi *a, int rc, const char *what) {
    set_error("%s failed: %s (%d)", what, (a && a->GetErrorString) ? a->GetErrorString(rc) : "?", rc);
    return B2NN_ERR_NCCL;
}
}  // namespace

int comm_unique_id(void *id128) {
    NcclApi *a = nccl();
    B2_REQUIRE(a != nullptr, B2NN_ERR_NCCL, "NCCL library not found (set B2NN_NCCL_LIB)");
    NcclUniqueId id;
    int rc = a->GetUniqueId(&id);
    if (rc != 0) return nccl_fail(a, rc, "ncclGetUniqueId");
    memcpy(id128, id.internal, 128);
    return B2NN_OK;
}

int comm_init(b2nn_net *net, const void *id128, int rank, int world) {
    B2_REQUIRE(world >= 1 && rank >= 0 && rank < world, B2NN_ERR_INVALID_ARG, "bad rank/world %d/%d", rank, world);
    B2_REQUIRE(net->precision == B2NN_PREC_BF16, B2NN_ERR_UNSUPPORTED,
               "data-parallel training is implemented for the BF16 path only (tiny FP64 nets are replicas only)");
    if (world == 1) { net->rank = 0; net->world = 1; return B2NN_OK; }
    NcclApi *a = nccl();
    B2_REQUIRE(a != nullptr, B2NN_ERR_NCCL, "NCCL library not found (set B2NN_NCCL_LIB)");
    NcclUniqueId id;
    memcpy(id.internal, id128, 128);
    NcclComm c = nullptr;
    int rc = a->CommInitRank(&c, world, id, rank);
    if (rc != 0) return nccl_fail(a, rc, "ncclCommInitRank");
    net->nccl_comm = c;
    net->rank = rank;
    net->world = world;
    const int L = net->n_dims - 1;
    for (int l = 0; l < L; l++) {
        auto &y = net->wide[l];
        if (y.g == nullptr) {
            B2_CHECK_CUDA(cudaMalloc(&y.g, (size_t)y.in * y.ld_n * sizeof(float)));
            B2_CHECK_CUDA(cudaMemset(y.g, 0, (size_t)y.in * y.ld_n * sizeof(float)));
        }
    }
    return B2NN_OK;
}

void comm_destroy(b2nn_net *net) {
    if (net->nccl_comm) {
        NcclApi *a = nccl();
        if (a) a->CommDestroy((NcclComm)net->nccl_comm);
        net->nccl_comm = nullptr;
    }
}

int comm_allreduce_sum_f32(b2nn_net *net, float *buf, size_t count, cudaStream_t st) {
    NcclApi *a = nccl();
    B2_REQUIRE(a != nullptr && net->nccl_comm != nullptr, B2NN_ERR_NCCL, "communicator not initialised");
    int rc = a->AllReduce(buf, buf, count, /*ncclFloat32*/ 7, /*ncclSum*/ 0, (NcclComm)net->nccl_comm, st);
    if (rc != 0) return nccl_fail(a, rc, "ncclAllReduce");
    return B2NN_OK;
}

}  // namespace b2nn
