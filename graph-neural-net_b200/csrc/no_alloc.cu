// no_alloc.cu -- b2nn_propagate_device_noalloc: the FP64 forward pass for callers that may not allocate.
//
// No counterpart in the reference.  Code under CUDA graph capture, tight
// serving loops and harnesses that forbid cudaMalloc cannot use b2nn_propagate, whose work buffers (the activated input
// rows and every layer's z / a rows) are cudaMalloc'ed on demand and whose rows cross PCIe.  This variant runs the same
// kernels (f64::gather_rows, f64::gemm_strict<EPI_FWD>, f64::output_rows: S:164-198 / G:183-221) with those work
// buffers carved out of ONE file-scope __device__ array of fixed size; inputs and outputs are the caller's DEVICE
// buffers.  It calls no cudaMalloc / cudaFree / allocating library routine and does not synchronise.
//
// The array is shared by every call on a device and keeps the last call's contents.  Nothing in it has to start at
// zero: every row range a kernel reads was fully written by an earlier kernel of the same call.  Calls that overlap on
// the GPU (two streams, two nets on one device) RACE on it; calls on one stream are ordered by the stream.
#include "common.hpp"
#include "f64_kernels.cuh"

namespace b2nn {

// 4 MB of FP64 work rows; every GPU that loads the library holds it until the process exits
constexpr size_t kNoAllocMaxDoubles = (size_t)B2NN_NOALLOC_MAX_DOUBLES;
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocMaxDoubles];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, once per device, with that device current.
static double *noalloc_scratch_of_device(int device, int *status) {
    static double *ptr[kMaxDevices] = {};
    *status = B2NN_OK;
    if (device < 0 || device >= kMaxDevices) { *status = B2NN_ERR_INVALID_ARG; return nullptr; }
    if (!ptr[device]) {
        double *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr) {
            set_error("cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", device, cudaGetErrorString(e));
            *status = B2NN_ERR_CUDA;
            return nullptr;
        }
        ptr[device] = p;   // the same value whichever thread stores it
    }
    return ptr[device];
}

static inline size_t round32(size_t n) { return (n + 31) & ~(size_t)31; }   // 256-byte aligned sub-buffers

}  // namespace b2nn

using namespace b2nn;

extern "C" int b2nn_noalloc_max_doubles(int64_t *n) {
    B2_REQUIRE(n != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    *n = (int64_t)kNoAllocMaxDoubles;
    return B2NN_OK;
}

extern "C" int b2nn_propagate_device_noalloc(b2nn_net *net, const double *x_dev, int32_t n, double *y_dev) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "net is null");
    B2_REQUIRE(x_dev != nullptr && y_dev != nullptr && n > 0, B2NN_ERR_INVALID_ARG, "null buffer or n <= 0");
    B2_REQUIRE(net->precision == B2NN_PREC_F64, B2NN_ERR_UNSUPPORTED,
               "b2nn_propagate_device_noalloc serves FP64 nets only (tensor-core nets keep their rows in the net's arena)");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    const int L = net->n_dims - 1, d0 = net->dims[0], dL = net->dims[L];
    // work rows: a0 = f(x), then z_l for l = 1..L and a_l for l = 1..L-1
    size_t need = round32((size_t)n * d0);
    for (int l = 1; l <= L; l++) need += round32((size_t)n * net->dims[l]) * (l < L ? 2 : 1);
    B2_REQUIRE(need <= kNoAllocMaxDoubles, B2NN_ERR_INVALID_ARG,
               "b2nn_propagate_device_noalloc: %d rows of this net need %zu doubles of work rows, the fixed scratch array "
               "holds %zu (B2NN_NOALLOC_MAX_DOUBLES); split the rows or use b2nn_propagate",
               (int)n, need, kNoAllocMaxDoubles);
    int st = B2NN_OK;
    double *scratch = noalloc_scratch_of_device(net->device, &st);
    if (st != B2NN_OK) return st;

    double *a_prev = scratch;
    size_t off = round32((size_t)n * d0);
    f64::gather_rows<<<n, 128, 0, net->stream>>>(x_dev, nullptr, nullptr, nullptr, n, d0, dL, net->inner_act,
                                                  net->inner_param, a_prev, nullptr);
    net->launches++;
    double *zL = nullptr;
    for (int l = 1; l <= L; l++) {
        const int dl = net->dims[l], dk = net->dims[l - 1];
        double *z = scratch + off;
        off += round32((size_t)n * dl);
        double *a = nullptr;
        if (l < L) { a = scratch + off; off += round32((size_t)n * dl); }
        f64::GemmArgs g{};
        g.M = n; g.N = dl; g.K = dk;
        g.A = a_prev; g.sam = dk; g.sak = 1;
        g.B = net->f64[l - 1].w; g.sbk = dl; g.sbn = 1;
        g.C = z; g.ldc = dl;
        g.C2 = a; g.ldc2 = dl;
        g.act = net->inner_act; g.act_param = net->inner_param;
        const bool small_m = n <= 32;   // the tile shapes f64_path.cu picks: the sums keep the reference's ascending-k order either way
        dim3 grid((dl + 63) / 64, (n + (small_m ? 15 : 63)) / (small_m ? 16 : 64));
        if (small_m) f64::gemm_strict<1, f64::EPI_FWD><<<grid, 256, 0, net->stream>>>(g);
        else f64::gemm_strict<4, f64::EPI_FWD><<<grid, 256, 0, net->stream>>>(g);
        net->launches++;
        a_prev = a;
        zL = z;
    }
    f64::output_rows<<<(n + 127) / 128, 128, 0, net->stream>>>(zL, nullptr, n, dL, net->kind, net->last_act, net->last_param,
                                                               net->loss, (net->flags & B2NN_FLAG_STABLE_SOFTMAX) ? 1 : 0,
                                                               y_dev, nullptr, nullptr, nullptr);
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}
