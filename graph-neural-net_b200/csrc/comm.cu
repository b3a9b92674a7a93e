// comm.cu -- NCCL gradient all-reduce for the data-parallel path.  The reference is serial
// (README.md:13); its batch sum (src/SoftmaxCrossEntropyNeuralNet.java:305-322) becomes a sum
// over GPUs here.  NCCL symbols are resolved with dlopen/dlsym (see comm.hpp).
#include "comm.hpp"

#include <dlfcn.h>

#include <cstring>


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
    B2_REQUIRE(is_wide_precision(net->precision), B2NN_ERR_UNSUPPORTED,
               "data-parallel training is implemented for the tensor-core path only (tiny FP64 nets are replicas only)");
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

// ---- peer memory (CUDA IPC): lets the fused reduce/update kernel (dp_kernels.cuh) address every rank's arena ----
int comm_ipc_export(b2nn_net *net, void *handle64) {
    B2_REQUIRE(net->arena != nullptr && is_wide_precision(net->precision), B2NN_ERR_UNSUPPORTED,
               "the peer-memory exchange is for the tensor-core (BF16 / TF32) nets");
    B2_REQUIRE(net->n_dims - 1 <= DP_MAX_LAYERS, B2NN_ERR_UNSUPPORTED,
               "the peer-memory exchange supports up to %d weight matrices (flag block layout)", DP_MAX_LAYERS);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    B2_CHECK_CUDA(cudaIpcGetMemHandle(&h, net->arena));
    memcpy(handle64, &h, 64);
    return B2NN_OK;
}

int comm_ipc_import(b2nn_net *net, const void *handles, int world) {
    B2_REQUIRE(world == net->world && world <= DP_MAX_WORLD, B2NN_ERR_INVALID_ARG,
               "world %d does not match the communicator (%d) or exceeds %d", world, net->world, DP_MAX_WORLD);
    B2_REQUIRE(net->n_dims - 1 <= DP_MAX_LAYERS, B2NN_ERR_UNSUPPORTED,
               "the peer-memory exchange supports up to %d weight matrices (flag block layout)", DP_MAX_LAYERS);
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    net->peer_arena.assign(world, nullptr);
    for (int p = 0; p < world; p++) {
        if (p == net->rank) { net->peer_arena[p] = net->arena; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char *)handles + 64 * (size_t)p, 64);
        void *ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            set_error("cudaIpcOpenMemHandle(rank %d) -> %s", p, cudaGetErrorString(e));
            for (int q = 0; q < p; q++)
                if (q != net->rank && net->peer_arena[q]) cudaIpcCloseMemHandle(net->peer_arena[q]);
            net->peer_arena.clear();
            return B2NN_ERR_CUDA;
        }
        net->peer_arena[p] = (char *)ptr;
    }
    return B2NN_OK;
}

void comm_destroy(b2nn_net *net) {
    if (net->vmm_arena) comm_vmm_destroy(net);  // also forgets the (non-IPC) peer mappings
    for (size_t p = 0; p < net->peer_arena.size(); p++)
        if ((int)p != net->rank && net->peer_arena[p]) cudaIpcCloseMemHandle(net->peer_arena[p]);
    net->peer_arena.clear();
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
