// api.cu -- C-ABI of libb2nn.so, part 1: lifecycle, weights, runtime (include/b2nn.h).
#include <cmath>
#include <cstring>

#include "comm.hpp"
#include "common.hpp"
#include "f64_path.hpp"
#include "wide_path.hpp"

namespace b2nn {
static thread_local char g_err[1024] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char *get_error() { return g_err; }
}  // namespace b2nn

using namespace b2nn;

extern "C" {

const char *b2nn_last_error(void) { return get_error(); }
int b2nn_version(void) { return B2NN_VERSION; }

int b2nn_debug_tile_list(int32_t M, int32_t N, int32_t pairs, int32_t *m0, int32_t *n0, int32_t *bn, int32_t cap,
                         int32_t *total) {
    return wide_tile_list(M, N, pairs, m0, n0, bn, cap, total);
}

int b2nn_device_count(int32_t *n) {
    B2_REQUIRE(n != nullptr, B2NN_ERR_INVALID_ARG, "n is null");
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess) { cudaGetLastError(); c = 0; }
    *n = c;
    return B2NN_OK;
}

int b2nn_create(const int32_t *dims, int32_t n_dims, int32_t net_kind, int32_t inner_act, double inner_param,
                int32_t last_act, double last_param, int32_t loss, int32_t precision, int32_t flags, int64_t seed,
                int32_t device, b2nn_net **out) {
    B2_REQUIRE(out != nullptr, B2NN_ERR_INVALID_ARG, "out is null");
    *out = nullptr;
    B2_REQUIRE(dims != nullptr && n_dims >= 2, B2NN_ERR_INVALID_ARG,
               "layerDims must hold at least 2 entries (SoftmaxCrossEntropyNeuralNet.java:105)");
    for (int i = 0; i < n_dims; i++)
        B2_REQUIRE(dims[i] > 0, B2NN_ERR_INVALID_ARG, "layerDims[%d] = %d must be positive", i, dims[i]);
    B2_REQUIRE(net_kind == B2NN_NET_SOFTMAX_CE || net_kind == B2NN_NET_GENERAL, B2NN_ERR_INVALID_ARG,
               "unknown net kind %d", net_kind);
    B2_REQUIRE(inner_act >= 0 && inner_act <= B2NN_ACT_IDENTITY, B2NN_ERR_INVALID_ARG,
               "unknown inner activation %d: only the built-in ActivationFunction singletons map to kernels",
               inner_act);
    if (net_kind == B2NN_NET_GENERAL) {
        B2_REQUIRE(last_act >= 0 && last_act <= B2NN_ACT_IDENTITY, B2NN_ERR_INVALID_ARG,
                   "unknown last activation %d", last_act);
        B2_REQUIRE(loss == B2NN_LOSS_HALF_SQUARED_ERROR, B2NN_ERR_INVALID_ARG, "unknown loss %d", loss);
    }
    B2_REQUIRE(precision == B2NN_PREC_F64 || precision == B2NN_PREC_BF16 || precision == B2NN_PREC_TF32,
               B2NN_ERR_INVALID_ARG, "unknown precision %d", precision);

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device: libb2nn has no CPU fallback");
        return B2NN_ERR_NO_DEVICE;
    }
    if (device < 0) B2_CHECK_CUDA(cudaGetDevice(&device));
    B2_REQUIRE(device < ndev, B2NN_ERR_INVALID_ARG, "device %d out of range (%d devices)", device, ndev);
    B2_CHECK_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    B2_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
    if (is_wide_precision(precision))
        B2_REQUIRE(prop.major == 10, B2NN_ERR_UNSUPPORTED,
                   "the BF16 / TF32 path is tcgen05 code for sm_100a; device is sm_%d%d", prop.major, prop.minor);

    b2nn_net *net = new b2nn_net();
    net->device = device;
    net->n_dims = n_dims;
    net->dims.assign(dims, dims + n_dims);
    net->kind = net_kind; net->inner_act = inner_act; net->inner_param = inner_param;
    net->last_act = last_act; net->last_param = last_param; net->loss = loss;
    net->precision = precision; net->flags = flags;
    net->num_sms = prop.multiProcessorCount;
    {
        const char *fu = getenv("B2NN_FUSED_UPDATE");
        net->fused_update = !(fu != nullptr && fu[0] == '0');   // single-GPU nets; B2NN_FUSED_UPDATE=0: separate update kernel
        net->trace = getenv("B2NN_TRACE") != nullptr;
        { const char *pd = getenv("B2NN_PDL"); net->pdl = !(pd != nullptr && pd[0] == '0'); }
        net->dp_ctas = prop.multiProcessorCount;  // one 128-thread CTA per SM beside the GEMM (measured best at 8 GPUs)
        if (const char *dc = getenv("B2NN_DP_CTAS")) { if (atoi(dc) > 0) net->dp_ctas = atoi(dc); }
        if (const char *to = getenv("B2NN_DP_TIMEOUT_MS")) { if (atof(to) > 0) net->dp_timeout_ns = (unsigned long long)(atof(to) * 1e6); }
        const char *dt = getenv("B2NN_DYNAMIC_TILES");
        net->dynamic_tiles = dt != nullptr && dt[0] == '1';
        net->dynamic_tiles_forced = dt != nullptr;
        const char *br = getenv("B2NN_BACKWARD_REORDER");
        net->backward_reorder = !(br != nullptr && br[0] == '0');
        const char *ou = getenv("B2NN_OVERLAP_UPDATE");
        net->overlap_update = !(ou != nullptr && ou[0] == '0');
        const char *ff = getenv("B2NN_F64_FUSED");
        net->fused_f64_disabled = ff != nullptr && ff[0] == '0';
        net->coop_launch = prop.cooperativeLaunch != 0;
        net->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
    }
    net->w_off.resize(n_dims);
    int64_t tot = 0;
    for (int l = 0; l + 1 < n_dims; l++) { net->w_off[l] = tot; tot += (int64_t)dims[l] * dims[l + 1]; }
    net->w_off[n_dims - 1] = tot;
    net->n_weights = tot;

    // appendLayer (S:139-156): layer-major, row i, col j, W = nextDouble() - 0.5
    std::vector<double> w((size_t)tot);
    JavaRandom rng(seed);
    for (int l = 0; l + 1 < n_dims; l++) {
        const double sc = (flags & B2NN_FLAG_SCALED_INIT) ? 2.0 * std::sqrt(6.0 / (double)dims[l]) : 1.0;
        double *W = w.data() + net->w_off[l];
        const size_t n = (size_t)dims[l] * dims[l + 1];
        if (flags & B2NN_FLAG_SCALED_INIT)
            for (size_t q = 0; q < n; q++) W[q] = (rng.next_double() - 0.5) * sc;
        else
            for (size_t q = 0; q < n; q++) W[q] = rng.next_double() - 0.5;
    }

    int st = B2NN_OK;
    cudaError_t ce = cudaStreamCreateWithFlags(&net->stream, cudaStreamNonBlocking);
    if (ce != cudaSuccess) {
        set_error("cudaStreamCreate: %s", cudaGetErrorString(ce));
        delete net;
        return B2NN_ERR_CUDA;
    }
    net->own_stream = true;
    st = (precision == B2NN_PREC_F64) ? f64_init(net, w) : wide_init(net, w);
    if (st != B2NN_OK) {
        b2nn_destroy(net);
        return st;
    }
    *out = net;
    return B2NN_OK;
}

int b2nn_destroy(b2nn_net *net) {
    if (!net) return B2NN_OK;
    cudaSetDevice(net->device);
    cudaDeviceSynchronize();
    comm_destroy(net);
    f64_fused_free(net);
    f64_free(net);
    wide_free(net);
    for (auto e : net->prof_events) cudaEventDestroy(e);
    if (net->prof_stamps) cudaFree(net->prof_stamps);
    if (net->pinned) cudaFreeHost(net->pinned);
    if (net->dscratch) cudaFree(net->dscratch);
    if (net->own_stream && net->stream) cudaStreamDestroy(net->stream);
    delete net;
    return B2NN_OK;
}

int b2nn_dims(const b2nn_net *net, int32_t *input_dim, int32_t *output_dim) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "net is null");
    if (input_dim) *input_dim = net->dims[0];                 // getInputDim S:383
    if (output_dim) *output_dim = net->dims[net->n_dims - 1]; // getOutputDim S:392
    return B2NN_OK;
}

int b2nn_num_weights(const b2nn_net *net, int64_t *n) {
    B2_REQUIRE(net != nullptr && n != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    *n = net->n_weights;
    return B2NN_OK;
}

static int xfer_weights(b2nn_net *net, double *w, bool to_host, bool velocity) {
    B2_REQUIRE(net != nullptr && w != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    if (net->precision == B2NN_PREC_F64) {
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        for (int l = 0; l + 1 < net->n_dims; l++) {
            const size_t n = (size_t)net->dims[l] * net->dims[l + 1] * sizeof(double);
            double *dev = velocity ? net->f64[l].v : net->f64[l].w;
            if (to_host) B2_CHECK_CUDA(cudaMemcpy(w + net->w_off[l], dev, n, cudaMemcpyDeviceToHost));
            else B2_CHECK_CUDA(cudaMemcpy(dev, w + net->w_off[l], n, cudaMemcpyHostToDevice));
        }
        return B2NN_OK;
    }
    return to_host ? wide_download_weights(net, w, velocity) : wide_upload_weights(net, w, velocity);
}

int b2nn_get_weights(b2nn_net *net, double *w) { return xfer_weights(net, w, true, false); }
int b2nn_set_weights(b2nn_net *net, const double *w) { return xfer_weights(net, const_cast<double *>(w), false, false); }
int b2nn_get_velocity(b2nn_net *net, double *v) { return xfer_weights(net, v, true, true); }
int b2nn_set_velocity(b2nn_net *net, const double *v) { return xfer_weights(net, const_cast<double *>(v), false, true); }

int b2nn_sync(b2nn_net *net) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "net is null");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    if (net->aux_stream) B2_CHECK_CUDA(cudaStreamSynchronize(net->aux_stream));
    return check_device_error_words(net);
}

int b2nn_set_stream(b2nn_net *net, void *cuda_stream) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "net is null");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    if (net->own_stream && net->stream) cudaStreamDestroy(net->stream);
    net->stream = reinterpret_cast<cudaStream_t>(cuda_stream);
    net->own_stream = false;
    return B2NN_OK;
}

int b2nn_kernel_launches(const b2nn_net *net, int64_t *n) {
    B2_REQUIRE(net != nullptr && n != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    *n = net->launches;
    return B2NN_OK;
}

// slots back to (start = max, end = 0), behind whatever the stream still holds
static int stamps_reset(b2nn_net *net) {
    const size_t n = (size_t)net->prof_stamp_cap;
    std::vector<unsigned long long> init(4 * n, 0ull);
    for (size_t i = 0; i < n; i++) init[4 * i] = ~0ull;
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    B2_CHECK_CUDA(cudaMemcpy(net->prof_stamps, init.data(), init.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    net->prof_stamp_used = 0;
    net->prof_stamp_flops = 0;
    return B2NN_OK;
}

int b2nn_profile_enable(b2nn_net *net, int32_t on) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "net is null");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    if (on && net->prof_events.empty()) {
        net->prof_events.resize(4096);
        net->prof_cls.assign(2048, 0);
        for (auto &e : net->prof_events) B2_CHECK_CUDA(cudaEventCreate(&e));
    }
    B2_TRY(profile_flush(net));
    if (on == 3 && net->prof_stamps == nullptr) {
        net->prof_stamp_cap = 16384;
        B2_CHECK_CUDA(cudaMalloc(&net->prof_stamps, 4 * sizeof(unsigned long long) * (size_t)net->prof_stamp_cap));
    }
    net->prof_stamping = on == 3;
    if (on == 3) B2_TRY(stamps_reset(net));
    net->profiling = on == 1 || on == 2;   // mode 3 records no events and serialises nothing
    net->prof_grouped = on == 2;   // one event pair per run of consecutive contraction launches instead of one per launch
    net->prof_flops = 0; net->prof_ms_acc = 0; net->prof_launch_acc = 0;
    for (int c = 0; c < 4; c++) { net->prof_ms_cls[c] = 0; net->prof_n_cls[c] = 0; }
    return B2NN_OK;
}

int b2nn_profile_read(b2nn_net *net, double *gemm_ms, int64_t *gemm_launches, double *gemm_flops) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "net is null");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    B2_TRY(profile_flush(net));
    if (gemm_ms) *gemm_ms = net->prof_ms_acc;
    if (gemm_launches) *gemm_launches = net->prof_launch_acc;
    if (gemm_flops) *gemm_flops = net->prof_flops;
    return B2NN_OK;
}

int b2nn_profile_read_classes(b2nn_net *net, double *ms4, int64_t *launches4) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "net is null");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    B2_TRY(profile_flush(net));
    for (int c = 0; c < 4; c++) {
        if (ms4) ms4[c] = net->prof_ms_cls[c];
        if (launches4) launches4[c] = net->prof_n_cls[c];
    }
    return B2NN_OK;
}

int b2nn_profile_read_stamps(b2nn_net *net, double *busy_ms, double *sum_ms, int64_t *launches, double *flops) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "net is null");
    B2_REQUIRE(net->prof_stamps != nullptr, B2NN_ERR_INVALID_ARG, "b2nn_profile_enable(net, 3) first");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    const int n = net->prof_stamp_used;
    std::vector<unsigned long long> h(4 * (size_t)n);
    if (n) B2_CHECK_CUDA(cudaMemcpy(h.data(), net->prof_stamps, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    // launches of one stream start in order: fold [start, end] intervals into their union (consecutive launches overlap
    // under programmatic dependent launch) and their plain sum
    double busy = 0, sum = 0, clk = 0, clk_ns = 0;
    unsigned long long cur0 = 0, cur1 = 0;
    int64_t valid = 0;
    for (int i = 0; i < n; i++) {
        const unsigned long long t0 = h[4 * (size_t)i], t1 = h[4 * (size_t)i + 1];
        if (t1 <= t0) continue;   // a slot whose kernel never ran
        valid++;
        clk += (double)h[4 * (size_t)i + 2]; clk_ns += (double)h[4 * (size_t)i + 3];
        sum += (double)(t1 - t0);
        if (cur1 == 0) { cur0 = t0; cur1 = t1; }
        else if (t0 <= cur1) { if (t1 > cur1) cur1 = t1; }
        else { busy += (double)(cur1 - cur0); cur0 = t0; cur1 = t1; }
    }
    if (cur1) busy += (double)(cur1 - cur0);
    if (busy_ms) *busy_ms = busy * 1e-6;
    if (sum_ms) *sum_ms = sum * 1e-6;
    if (launches) *launches = valid;
    if (flops) *flops = net->prof_stamp_flops;
    net->prof_stamp_mhz = clk_ns > 0 ? clk / clk_ns * 1e3 : 0.0;
    return stamps_reset(net);
}

int b2nn_profile_stamp_clock_mhz(const b2nn_net *net, double *mhz) {
    B2_REQUIRE(net != nullptr && mhz != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    *mhz = net->prof_stamp_mhz;
    return B2NN_OK;
}

int b2nn_k1_profile(b2nn_net *net, int32_t on, double *cycles16_or_null, int32_t *steps_or_null) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "net is null");
    if (cycles16_or_null) memcpy(cycles16_or_null, net->k1_prof_cycles, sizeof net->k1_prof_cycles);
    if (steps_or_null) *steps_or_null = net->k1_prof_iters;
    net->k1_prof = on != 0;
    return B2NN_OK;
}

int b2nn_comm_unique_id(void *id128) {
    B2_REQUIRE(id128 != nullptr, B2NN_ERR_INVALID_ARG, "id is null");
    return comm_unique_id(id128);
}

int b2nn_comm_init(b2nn_net *net, const void *id128, int32_t rank, int32_t world) {
    B2_REQUIRE(net != nullptr && id128 != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    return comm_init(net, id128, rank, world);
}

int b2nn_comm_ipc_export(b2nn_net *net, void *handle64) {
    B2_REQUIRE(net != nullptr && handle64 != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    return comm_ipc_export(net, handle64);
}

int b2nn_comm_ipc_import(b2nn_net *net, const void *handles, int32_t world) {
    B2_REQUIRE(net != nullptr && handles != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    return comm_ipc_import(net, handles, world);
}

int b2nn_comm_nvls_export(b2nn_net *net, int32_t *pid_fd) {
    B2_REQUIRE(net != nullptr && pid_fd != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    return comm_vmm_export(net, &pid_fd[0], &pid_fd[1]);
}

int b2nn_comm_nvls_import(b2nn_net *net, const int32_t *pid_fd, int32_t world) {
    B2_REQUIRE(net != nullptr && pid_fd != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    B2_REQUIRE(world >= 1 && world <= DP_MAX_WORLD, B2NN_ERR_INVALID_ARG, "bad world %d", world);
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    int32_t pids[DP_MAX_WORLD], fds[DP_MAX_WORLD];
    for (int p = 0; p < world; p++) { pids[p] = pid_fd[2 * p]; fds[p] = pid_fd[2 * p + 1]; }
    return comm_vmm_import(net, pids, fds, world);
}

int b2nn_comm_nvls_team(b2nn_net *net, int32_t *pid_fd) {
    B2_REQUIRE(net != nullptr && pid_fd != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    if (net->rank == 0) B2_TRY(comm_mc_create(net, &pid_fd[0], &pid_fd[1]));
    else B2_TRY(comm_mc_import(net, pid_fd[0], pid_fd[1]));
    return comm_mc_add_device(net);
}

int b2nn_comm_mode(const b2nn_net *net, int32_t *mode) {
    B2_REQUIRE(net != nullptr && mode != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    *mode = net->world <= 1 ? 0 : net->peer_arena.empty() ? 1 : net->use_multimem ? 3 : 2;
    return B2NN_OK;
}

int b2nn_comm_nvls_bind(b2nn_net *net) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    return comm_mc_bind(net);
}

}  // extern "C"
