// comm_vmm.cu -- NVSwitch multicast (NVLS) plumbing for the data-parallel exchange.
//
// multimem.ld_reduce / multimem.st address a MULTICAST object: one virtual address that stands for the same
// offset in every GPU's memory, with the reduction / replication done inside the NVSwitch.  Memory bound to a
// multicast object must come from the virtual-memory-management API (cuMemCreate), and processes share such
// allocations through POSIX file descriptors.  So, for a one-process-per-GPU group:
//   1. every rank moves its parameter arena into a cuMemCreate allocation and exports its fd
//      (comm_vmm_export); the launcher all-gathers (pid, fd); every rank duplicates the peers' fds with
//      pidfd_getfd, imports and maps them (comm_vmm_import) -> unicast peer pointers (flags, shard gather);
//   2. rank 0 creates the multicast object and exports its fd (comm_mc_create), the others import it
//      (comm_mc_import); all add their device (comm_mc_add_device), [barrier], bind their arena and map the
//      object (comm_mc_bind) -> net->mc_base.
// Driver entry points are looked up at run time (no link-time dependency on libcuda).
#include <sys/syscall.h>
#include <unistd.h>

#include <cerrno>
#include <cstring>

#include "comm.hpp"
#include "wide_path.hpp"

namespace b2nn {

namespace {

template <typename F>
bool drv(const char *name, F *out) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
        p == nullptr)
        return false;
    *out = reinterpret_cast<F>(p);
    return true;
}

struct Drv {
    CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long);
    CUresult (*MemRelease)(CUmemGenericAllocationHandle);
    CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long);
    CUresult (*MemAddressFree)(CUdeviceptr, size_t);
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
    CUresult (*MemUnmap)(CUdeviceptr, size_t);
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t);
    CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags);
    CUresult (*MemExportToShareableHandle)(void *, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                           unsigned long long);
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle *, void *, CUmemAllocationHandleType);
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *);
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                                 unsigned long long);
    CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags);
    CUresult (*DeviceGet)(CUdevice *, int);
    CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice);
    bool ok = false;
};

Drv *driver() {
    static Drv d;
    static bool tried = false;
    if (tried) return d.ok ? &d : nullptr;
    tried = true;
    d.ok = drv("cuMemCreate", &d.MemCreate) && drv("cuMemRelease", &d.MemRelease) &&
           drv("cuMemAddressReserve", &d.MemAddressReserve) && drv("cuMemAddressFree", &d.MemAddressFree) &&
           drv("cuMemMap", &d.MemMap) && drv("cuMemUnmap", &d.MemUnmap) && drv("cuMemSetAccess", &d.MemSetAccess) &&
           drv("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity) &&
           drv("cuMemExportToShareableHandle", &d.MemExportToShareableHandle) &&
           drv("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle) &&
           drv("cuMulticastCreate", &d.MulticastCreate) && drv("cuMulticastAddDevice", &d.MulticastAddDevice) &&
           drv("cuMulticastBindMem", &d.MulticastBindMem) &&
           drv("cuMulticastGetGranularity", &d.MulticastGetGranularity) && drv("cuDeviceGet", &d.DeviceGet) &&
           drv("cuDeviceGetAttribute", &d.DeviceGetAttribute);
    return d.ok ? &d : nullptr;
}

#define B2_CHECK_DRV(expr)                                                          \
    do {                                                                            \
        CUresult _r = (expr);                                                       \
        if (_r != CUDA_SUCCESS) {                                                   \
            b2nn::set_error("%s:%d: %s -> CUresult %d", __FILE__, __LINE__, #expr, (int)_r); \
            return B2NN_ERR_CUDA;                                                   \
        }                                                                           \
    } while (0)

CUmemAllocationProp arena_prop(int device) {
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof prop);
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = device;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

int map_rw(Drv *d, CUdeviceptr *va, size_t bytes, size_t align, CUmemGenericAllocationHandle h, int device) {
    B2_CHECK_DRV(d->MemAddressReserve(va, bytes, align, 0, 0));
    B2_CHECK_DRV(d->MemMap(*va, bytes, 0, h, 0));
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof acc);
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = device;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    B2_CHECK_DRV(d->MemSetAccess(*va, bytes, &acc, 1));
    return B2NN_OK;
}

// duplicate file descriptor `fd` of process `pid` into this process (Linux >= 5.6)
int steal_fd(int pid, int fd) {
    if (pid == (int)getpid()) return dup(fd);
    const int pidfd = (int)syscall(434 /*pidfd_open*/, pid, 0);
    if (pidfd < 0) return -1;
    const int mine = (int)syscall(438 /*pidfd_getfd*/, pidfd, fd, 0);
    close(pidfd);
    return mine;
}

}  // namespace

// 1a. move the arena into an exportable cuMemCreate allocation; returns this process's pid and the arena's fd
int comm_vmm_export(b2nn_net *net, int32_t *pid, int32_t *fd) {
    Drv *d = driver();
    B2_REQUIRE(d != nullptr, B2NN_ERR_UNSUPPORTED, "CUDA virtual-memory / multicast driver entry points not available");
    B2_REQUIRE(net->arena != nullptr && is_wide_precision(net->precision) && net->world > 1, B2NN_ERR_UNSUPPORTED,
               "needs a tensor-core (BF16 / TF32) net with a communicator");
    // before the arena moves: the flag block at its head holds DP_MAX_LAYERS rows per handshake array
    B2_REQUIRE(net->n_dims - 1 <= DP_MAX_LAYERS, B2NN_ERR_UNSUPPORTED,
               "the in-switch exchange supports up to %d weight matrices (flag block layout)", DP_MAX_LAYERS);
    B2_REQUIRE(!net->vmm_arena, B2NN_ERR_INVALID_ARG, "arena already exported");
    B2_REQUIRE(net->peer_arena.empty(), B2NN_ERR_INVALID_ARG, "the arena is already shared through CUDA IPC");
    {
        CUdevice dev;
        int mc = 0, fdh = 0;
        B2_CHECK_DRV(d->DeviceGet(&dev, net->device));
        B2_CHECK_DRV(d->DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
        B2_CHECK_DRV(d->DeviceGetAttribute(&fdh, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
        B2_REQUIRE(mc != 0 && fdh != 0, B2NN_ERR_UNSUPPORTED, "device %d: multicast %d, fd handles %d", net->device, mc, fdh);
    }
    B2_CHECK_CUDA(cudaDeviceSynchronize());
    const CUmemAllocationProp prop = arena_prop(net->device);
    size_t gran = 0, mc_gran = 0;
    B2_CHECK_DRV(d->MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof mp);
    mp.numDevices = (unsigned)net->world;
    mp.size = net->arena_bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    B2_CHECK_DRV(d->MulticastGetGranularity(&mc_gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    if (mc_gran > gran) gran = mc_gran;
    const size_t bytes = (net->arena_bytes + gran - 1) / gran * gran;
    CUmemGenericAllocationHandle h;
    B2_CHECK_DRV(d->MemCreate(&h, bytes, &prop, 0));
    CUdeviceptr va = 0;
    B2_TRY(map_rw(d, &va, bytes, gran, h, net->device));
    B2_CHECK_CUDA(cudaMemset((void *)va, 0, bytes));
    B2_CHECK_CUDA(cudaMemcpy((void *)va, net->arena, net->arena_bytes, cudaMemcpyDeviceToDevice));
    int efd = -1;
    B2_CHECK_DRV(d->MemExportToShareableHandle(&efd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    // swap: every pointer into the arena is an offset from its base
    char *old_arena = net->arena;
    B2_TRY(wide_rebase_arena(net, (char *)va));
    B2_CHECK_CUDA(cudaFree(old_arena));
    net->vmm_arena = true;
    net->vmm_handle = h;
    net->vmm_bytes = bytes;
    net->vmm_gran = gran;
    net->vmm_fd = efd;
    *pid = (int32_t)getpid();
    *fd = efd;
    return B2NN_OK;
}

// 1b. map every peer's arena (unicast)
int comm_vmm_import(b2nn_net *net, const int32_t *pids, const int32_t *fds, int world) {
    Drv *d = driver();
    B2_REQUIRE(d != nullptr && net->vmm_arena, B2NN_ERR_INVALID_ARG, "call comm_vmm_export first");
    B2_REQUIRE(world == net->world && world <= DP_MAX_WORLD, B2NN_ERR_INVALID_ARG, "bad world %d", world);
    net->peer_arena.assign(world, nullptr);
    net->peer_handles.assign(world, 0);
    for (int p = 0; p < world; p++) {
        if (p == net->rank) { net->peer_arena[p] = net->arena; continue; }
        const int fd = steal_fd(pids[p], fds[p]);
        B2_REQUIRE(fd >= 0, B2NN_ERR_CUDA, "pidfd_getfd(pid %d, fd %d) failed (errno %d)", pids[p], fds[p], errno);
        CUmemGenericAllocationHandle h;
        CUresult r = d->MemImportFromShareableHandle(&h, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
        close(fd);
        B2_REQUIRE(r == CUDA_SUCCESS, B2NN_ERR_CUDA, "cuMemImportFromShareableHandle(rank %d) -> %d", p, (int)r);
        CUdeviceptr va = 0;
        B2_TRY(map_rw(d, &va, net->vmm_bytes, net->vmm_gran, h, net->device));
        net->peer_handles[p] = h;
        net->peer_arena[p] = (char *)va;
    }
    return B2NN_OK;
}

// 2a. rank 0: create the multicast object, export its fd
int comm_mc_create(b2nn_net *net, int32_t *pid, int32_t *fd) {
    Drv *d = driver();
    B2_REQUIRE(d != nullptr && net->vmm_arena, B2NN_ERR_INVALID_ARG, "call comm_vmm_export first");
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof mp);
    mp.numDevices = (unsigned)net->world;
    mp.size = net->vmm_bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    B2_CHECK_DRV(d->MulticastCreate(&net->mc_handle, &mp));
    int efd = -1;
    B2_CHECK_DRV(d->MemExportToShareableHandle(&efd, net->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    net->mc_fd = efd;
    *pid = (int32_t)getpid();
    *fd = efd;
    return B2NN_OK;
}

// 2b. other ranks: import rank 0's multicast object
int comm_mc_import(b2nn_net *net, int32_t pid, int32_t fd) {
    Drv *d = driver();
    B2_REQUIRE(d != nullptr && net->vmm_arena, B2NN_ERR_INVALID_ARG, "call comm_vmm_export first");
    const int mine = steal_fd(pid, fd);
    B2_REQUIRE(mine >= 0, B2NN_ERR_CUDA, "pidfd_getfd(pid %d, fd %d) failed (errno %d)", pid, fd, errno);
    CUresult r = d->MemImportFromShareableHandle(&net->mc_handle, (void *)(uintptr_t)mine,
                                                 CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    close(mine);
    B2_REQUIRE(r == CUDA_SUCCESS, B2NN_ERR_CUDA, "cuMemImportFromShareableHandle(multicast) -> %d", (int)r);
    return B2NN_OK;
}

// 2c. every rank: join the team (all must have joined before anyone binds)
int comm_mc_add_device(b2nn_net *net) {
    Drv *d = driver();
    B2_REQUIRE(d != nullptr && net->mc_handle != 0, B2NN_ERR_INVALID_ARG, "no multicast object");
    CUdevice dev;
    B2_CHECK_DRV(d->DeviceGet(&dev, net->device));
    B2_CHECK_DRV(d->MulticastAddDevice(net->mc_handle, dev));
    return B2NN_OK;
}

// 2d. every rank: bind its arena at offset 0 and map the multicast object
int comm_mc_bind(b2nn_net *net) {
    Drv *d = driver();
    B2_REQUIRE(d != nullptr && net->mc_handle != 0, B2NN_ERR_INVALID_ARG, "no multicast object");
    B2_CHECK_DRV(d->MulticastBindMem(net->mc_handle, 0, net->vmm_handle, 0, net->vmm_bytes, 0));
    CUdeviceptr va = 0;
    B2_TRY(map_rw(d, &va, net->vmm_bytes, net->vmm_gran, net->mc_handle, net->device));
    net->mc_base = (char *)va;
    net->use_multimem = true;
    return B2NN_OK;
}

void comm_vmm_destroy(b2nn_net *net) {
    Drv *d = driver();
    if (!d || !net->vmm_arena) return;
    cudaDeviceSynchronize();
    if (net->mc_base) { d->MemUnmap((CUdeviceptr)net->mc_base, net->vmm_bytes); d->MemAddressFree((CUdeviceptr)net->mc_base, net->vmm_bytes); }
    if (net->mc_handle) d->MemRelease(net->mc_handle);
    for (size_t p = 0; p < net->peer_handles.size(); p++)
        if ((int)p != net->rank && net->peer_handles[p]) {
            d->MemUnmap((CUdeviceptr)net->peer_arena[p], net->vmm_bytes);
            d->MemAddressFree((CUdeviceptr)net->peer_arena[p], net->vmm_bytes);
            d->MemRelease(net->peer_handles[p]);
        }
    net->peer_arena.clear();
    net->peer_handles.clear();
    d->MemUnmap((CUdeviceptr)net->arena, net->vmm_bytes);
    d->MemAddressFree((CUdeviceptr)net->arena, net->vmm_bytes);
    d->MemRelease(net->vmm_handle);
    if (net->vmm_fd >= 0) close(net->vmm_fd);
    if (net->mc_fd >= 0) close(net->mc_fd);
    net->arena = nullptr;
    net->mc_base = nullptr;
    net->use_multimem = false;
    net->mc_handle = 0;
    net->vmm_arena = false;
}

}  // namespace b2nn
