// f64_path.cu -- host orchestration of the FP64 parity path (general shapes).
// S = src/SoftmaxCrossEntropyNeuralNet.java, G = src/GeneralNeuralNet.java (reference).
#include "f64_path.hpp"

#include "f64_kernels.cuh"

namespace b2nn {

static int launch_gemm(b2nn_net *net, int epi, const f64::GemmArgs &g) {
    const bool small_m = g.M <= 32;
    dim3 grid((g.N + 63) / 64, (g.M + (small_m ? 15 : 63)) / (small_m ? 16 : 64));
    dim3 block(256);
#define B2_F64_LAUNCH(TMT, EPI) f64::gemm_strict<TMT, EPI><<<grid, block, 0, net->stream>>>(g)
    if (small_m) {
        switch (epi) {
        case f64::EPI_FWD: B2_F64_LAUNCH(1, f64::EPI_FWD); break;
        case f64::EPI_DX: B2_F64_LAUNCH(1, f64::EPI_DX); break;
        case f64::EPI_DW_UPDATE: B2_F64_LAUNCH(1, f64::EPI_DW_UPDATE); break;
        default: B2_F64_LAUNCH(1, f64::EPI_STORE); break;
        }
    } else {
        switch (epi) {
        case f64::EPI_FWD: B2_F64_LAUNCH(4, f64::EPI_FWD); break;
        case f64::EPI_DX: B2_F64_LAUNCH(4, f64::EPI_DX); break;
        case f64::EPI_DW_UPDATE: B2_F64_LAUNCH(4, f64::EPI_DW_UPDATE); break;
        default: B2_F64_LAUNCH(4, f64::EPI_STORE); break;
        }
    }
#undef B2_F64_LAUNCH
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

int f64_init(b2nn_net *net, const std::vector<double> &w_host) {
    const int L = net->n_dims - 1;
    net->f64.resize(L);
    for (int l = 0; l < L; l++) {
        const size_t n = (size_t)net->dims[l] * net->dims[l + 1];
        B2_CHECK_CUDA(cudaMalloc(&net->f64[l].w, n * sizeof(double)));
        B2_CHECK_CUDA(cudaMalloc(&net->f64[l].v, n * sizeof(double)));
        B2_CHECK_CUDA(cudaMemcpy(net->f64[l].w, w_host.data() + net->w_off[l], n * sizeof(double),
                                 cudaMemcpyHostToDevice));
        B2_CHECK_CUDA(cudaMemset(net->f64[l].v, 0, n * sizeof(double)));  // previousUpdate zeros, S:152
    }
    B2_CHECK_CUDA(cudaMalloc(&net->f64_grad, (size_t)net->n_weights * sizeof(double)));
    return B2NN_OK;
}

void f64_free(b2nn_net *net) {
    for (auto &l : net->f64) { cudaFree(l.w); cudaFree(l.v); }
    net->f64.clear();
    auto &fw = net->fw;
    for (auto p : fw.z) cudaFree(p);
    for (auto p : fw.a) cudaFree(p);
    for (auto p : fw.d) cudaFree(p);
    fw.z.clear(); fw.a.clear(); fw.d.clear();
    cudaFree(fw.t); cudaFree(fw.row_loss); cudaFree(fw.pred); cudaFree(fw.idx); cudaFree(fw.xin);
    fw.t = nullptr; fw.row_loss = nullptr; fw.pred = nullptr; fw.idx = nullptr; fw.xin = nullptr;
    fw.cap_rows = 0;
    cudaFree(net->f64_grad);
    net->f64_grad = nullptr;
}

int f64_ensure_work(b2nn_net *net, int rows) {
    auto &fw = net->fw;
    if (rows <= fw.cap_rows) return B2NN_OK;
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    for (auto p : fw.z) cudaFree(p);
    for (auto p : fw.a) cudaFree(p);
    for (auto p : fw.d) cudaFree(p);
    cudaFree(fw.t); cudaFree(fw.row_loss); cudaFree(fw.pred); cudaFree(fw.idx); cudaFree(fw.xin);
    const int nd = net->n_dims;
    fw.z.assign(nd, nullptr); fw.a.assign(nd, nullptr); fw.d.assign(nd, nullptr);
    fw.cap_rows = 0;
    fw.iota_rows = 0;
    for (int l = 0; l < nd; l++) {
        const size_t n = (size_t)rows * net->dims[l] * sizeof(double);
        B2_CHECK_CUDA(cudaMalloc(&fw.z[l], n));
        B2_CHECK_CUDA(cudaMalloc(&fw.a[l], n));
        B2_CHECK_CUDA(cudaMalloc(&fw.d[l], n));
    }
    B2_CHECK_CUDA(cudaMalloc(&fw.t, (size_t)rows * net->dims[nd - 1] * sizeof(double)));
    B2_CHECK_CUDA(cudaMalloc(&fw.row_loss, (size_t)rows * sizeof(double)));
    B2_CHECK_CUDA(cudaMalloc(&fw.pred, (size_t)rows * sizeof(int32_t)));
    B2_CHECK_CUDA(cudaMalloc(&fw.idx, (size_t)rows * sizeof(int32_t)));
    B2_CHECK_CUDA(cudaMalloc(&fw.xin, (size_t)rows * net->dims[0] * sizeof(double)));
    fw.cap_rows = rows;
    return B2NN_OK;
}

// a0 = f(X[idx]) and the target rows, from device-resident X / T / labels
int f64_gather(b2nn_net *net, const double *X, const double *T, const int32_t *labels, const int32_t *idx_dev,
               int rows, bool want_targets) {
    auto &fw = net->fw;
    f64::gather_rows<<<rows, 128, 0, net->stream>>>(X, T, labels, idx_dev, rows, net->dims[0],
                                                     net->dims[net->n_dims - 1], net->inner_act, net->inner_param,
                                                     fw.a[0], want_targets ? fw.t : nullptr);
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

// propagate S:169-193: z_l = a_{l-1} W_{l-1}; a_l = f(z_l)
int f64_forward(b2nn_net *net, int rows) {
    auto &fw = net->fw;
    const int L = net->n_dims - 1;
    for (int l = 1; l <= L; l++) {
        f64::GemmArgs g{};
        g.M = rows; g.N = net->dims[l]; g.K = net->dims[l - 1];
        g.A = fw.a[l - 1]; g.sam = net->dims[l - 1]; g.sak = 1;
        g.B = net->f64[l - 1].w; g.sbk = net->dims[l]; g.sbn = 1;
        g.C = fw.z[l]; g.ldc = net->dims[l];
        g.C2 = (l < L) ? fw.a[l] : nullptr; g.ldc2 = net->dims[l];
        g.act = net->inner_act; g.act_param = net->inner_param;
        B2_TRY(launch_gemm(net, f64::EPI_FWD, g));
    }
    return B2NN_OK;
}

int f64_output(b2nn_net *net, int rows, bool have_targets, double *P, bool want_loss, bool want_delta,
               bool want_pred) {
    auto &fw = net->fw;
    const int L = net->n_dims - 1;
    f64::output_rows<<<(rows + 127) / 128, 128, 0, net->stream>>>(
        fw.z[L], have_targets ? fw.t : nullptr, rows, net->dims[L], net->kind, net->last_act, net->last_param,
        net->loss, (net->flags & B2NN_FLAG_STABLE_SOFTMAX) ? 1 : 0, P, want_loss ? fw.row_loss : nullptr,
        want_delta ? fw.d[L] : nullptr, want_pred ? fw.pred : nullptr);
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

// inner-layer deltas S:260-278: d_j = (W_j d_{j+1}) (.) f'(z_j), j = L-1 .. 1
int f64_deltas(b2nn_net *net, int rows) {
    auto &fw = net->fw;
    const int L = net->n_dims - 1;
    for (int jl = L - 1; jl >= 1; jl--) {
        f64::GemmArgs g{};
        g.M = rows; g.N = net->dims[jl]; g.K = net->dims[jl + 1];
        g.A = fw.d[jl + 1]; g.sam = net->dims[jl + 1]; g.sak = 1;
        g.B = net->f64[jl].w; g.sbk = 1; g.sbn = net->dims[jl + 1];  // B(k,n) = W_j[n][k]
        g.C = fw.d[jl]; g.ldc = net->dims[jl];
        g.Z = fw.z[jl]; g.ldz = net->dims[jl];
        g.act = net->inner_act; g.act_param = net->inner_param;
        B2_TRY(launch_gemm(net, f64::EPI_DX, g));
    }
    return B2NN_OK;
}

// batch gradient + momentum update S:305-344, or plain store of the gradient (grad_out != null)
int f64_gradients(b2nn_net *net, int rows, double step, double momentum, double *grad_out) {
    auto &fw = net->fw;
    const int L = net->n_dims - 1;
    for (int l = 0; l < L; l++) {
        f64::GemmArgs g{};
        g.M = net->dims[l]; g.N = net->dims[l + 1]; g.K = rows;
        g.A = fw.a[l]; g.sam = 1; g.sak = net->dims[l];           // A(m,k) = a_l[k][m]
        g.B = fw.d[l + 1]; g.sbk = net->dims[l + 1]; g.sbn = 1;
        g.ldc = net->dims[l + 1];
        if (grad_out) {
            g.C = grad_out + net->w_off[l];
            B2_TRY(launch_gemm(net, f64::EPI_STORE, g));
        } else {
            g.C = net->f64[l].w; g.V = net->f64[l].v;
            g.step = step; g.momentum = momentum; g.batch = rows;   // batch.size(), S:325
            B2_TRY(launch_gemm(net, f64::EPI_DW_UPDATE, g));
        }
    }
    return B2NN_OK;
}

int f64_iota(b2nn_net *net, int rows) {
    if (net->fw.iota_rows >= rows) return B2NN_OK;
    net->fw.iota_rows = rows;
    f64::iota_rows<<<(rows + 127) / 128, 128, 0, net->stream>>>(net->fw.idx, rows);
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

int f64_reduce(b2nn_net *net, int rows, const int32_t *labels_dev, double *loss_sum_dev, long long *hits_dev) {
    auto &fw = net->fw;
    f64::reduce_rows<<<1, 256, 0, net->stream>>>(loss_sum_dev ? fw.row_loss : nullptr, hits_dev ? fw.pred : nullptr,
                                                labels_dev, rows, loss_sum_dev, hits_dev);
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

}  // namespace b2nn
